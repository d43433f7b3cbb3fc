#!/bin/bash
# One GPU-box pass: parity tests, smoke, bench line, ncu launch list + one full capture of the sample kernel.
# Usage (from the repo root, under gpurun):  bash scripts/gpu_check.sh [tag]
set -u
TAG=${1:-r01}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/${TAG}_pytest.log
timeout 300 python __graft_entry__.py --smoke > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/${TAG}_smoke.log
timeout 900 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"
tail -c 3000 $OUT/${TAG}_bench.json
NCU_CMD="python bench.py --steps 20 --warmup 3 --no-sweep --no-cpu-baseline"
timeout 300 $NCU_CMD > $OUT/${TAG}_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv $NCU_CMD > $OUT/${TAG}_ncu1.log 2>&1
echo "ncu launches rc=$?"
timeout 300 $NCU_CMD > $OUT/${TAG}_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ring_sample -s 5 -c 3 -f -o $OUT/${TAG}_ring_sample $NCU_CMD > $OUT/${TAG}_ncu2.log 2>&1
echo "ncu full rc=$?"
