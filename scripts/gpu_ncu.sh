#!/bin/bash
# ncu evidence for the hot kernels (one GPU).  Usage under gpurun: bash scripts/gpu_ncu.sh <tag>
set -u
TAG=${1:-r01}
OUT=gpurun_out
mkdir -p $OUT
CMD="python scripts/ncu_kernels.py"
timeout 300 $CMD > $OUT/${TAG}_ncu_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'ring_sample_band|random_conv_tma|overlay_vec|splice_vec|window_copy_vec' -s 5 -c 10 -f -o $OUT/${TAG}_kernels $CMD > $OUT/${TAG}_ncu_full.log 2>&1
echo "ncu full rc=$?"
BCMD="python bench.py --steps 20 --warmup 3 --no-sweep --no-cpu-baseline"
timeout 300 $BCMD > $OUT/${TAG}_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv $BCMD > $OUT/${TAG}_ncu1.log 2>&1
echo "ncu launches rc=$?"
