#!/bin/bash
# Round-2 GPU pass B: fused-sampler parity, fused bench (conv chains), ncu full capture of the splice-conv kernel.
# Usage under gpurun: bash scripts/gpu_r2b.sh <tag> [ncu]
set -u
TAG=${1:-r02b}
OUT=gpurun_out
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_fused.py tests/test_gpu_aug_parity.py -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/${TAG}_pytest.log
tail -15 $OUT/${TAG}_pytest.log
timeout 600 python scripts/bench_fused.py --batches 128,1024,4096 --cases augcl_splice_conv,svea_conv,augcl_splice,svea_overlay > $OUT/${TAG}_fused.jsonl 2> $OUT/${TAG}_fused.err; echo "bench_fused rc=$?"
cut -c1-200 $OUT/${TAG}_fused.jsonl
if [ "${2:-}" = "ncu" ]; then
  CMD="python scripts/bench_fused.py --batches 1024 --cases augcl_splice_conv"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:ring_sample_fused -s 30 -c 1 -f -o $OUT/${TAG}_fused_conv $CMD > $OUT/${TAG}_ncu.log 2>&1
  echo "ncu rc=$?"
fi
