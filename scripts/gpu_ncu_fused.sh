#!/bin/bash
# ncu --set full capture of one fused-sampler launch.  Usage under gpurun: bash scripts/gpu_ncu_fused.sh <tag> <case> <batch> [opts]
set -u
TAG=$1; CASE=$2; B=$3; OPTS=${4:-}
OUT=gpurun_out; mkdir -p $OUT
CMD="python scripts/bench_fused.py --batches $B --cases $CASE"
[ -n "$OPTS" ] && CMD="$CMD --opts $OPTS"
timeout 300 $CMD > $OUT/${TAG}_plain.jsonl 2>&1 || { echo "plain run failed"; tail -5 $OUT/${TAG}_plain.jsonl; exit 1; }
cat $OUT/${TAG}_plain.jsonl | cut -c1-160
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ring_sample_fused -s 30 -c 1 -f -o $OUT/${TAG} $CMD > $OUT/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"
