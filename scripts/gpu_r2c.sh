#!/bin/bash
# Round-2 GPU pass C (one GPU): full parity suite, smoke, bench at the driver's settings, reference arm, kernel sweeps,
# ncu launch list of the bench command and --set full captures of the headline and the conv-chain kernels.
set -u
TAG=${1:-r02c}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/${TAG}_pytest.log
tail -4 $OUT/${TAG}_pytest.log
timeout 300 python __graft_entry__.py --smoke > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/${TAG}_smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > $OUT/${TAG}_bench20.json 2> $OUT/${TAG}_bench20.err; echo "bench20 rc=$?"
tail -c 600 $OUT/${TAG}_bench20.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $OUT/${TAG}_benchref.json 2> $OUT/${TAG}_benchref.err; echo "benchref rc=$?"
timeout 600 python scripts/bench_fused.py --batches 128,256,1024,4096 > $OUT/${TAG}_fused.jsonl 2> $OUT/${TAG}_fused.err; echo "bench_fused rc=$?"
timeout 600 python scripts/bench_augs.py --batches 128,2048 > $OUT/${TAG}_augs.jsonl 2> $OUT/${TAG}_augs.err; echo "bench_augs rc=$?"
NCU_CMD="python bench.py --steps 20 --warmup 3 --no-sweep --no-cpu-baseline"
timeout 300 $NCU_CMD > $OUT/${TAG}_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/${TAG}_launches.csv $NCU_CMD > $OUT/${TAG}_ncu1.log 2>&1
echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ring_sample_band -s 5 -c 2 -f -o $OUT/${TAG}_ring_sample $NCU_CMD > $OUT/${TAG}_ncu2.log 2>&1
echo "ncu full (band) rc=$?"
for CB in "augcl_splice_conv 1024" "augcl_splice_conv 128" "svea_overlay 256" "drq_shift_both 16384"; do
  set -- $CB
  CMD="python scripts/bench_fused.py --batches $2 --cases $1"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:ring_sample_fused -s 30 -c 1 -f -o $OUT/${TAG}_fused_$1_$2 $CMD > $OUT/${TAG}_ncu_$1_$2.log 2>&1
  echo "ncu full ($1 $2) rc=$?"
done
