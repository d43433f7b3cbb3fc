#!/bin/bash
# Option sweep of the fused sampler.  Usage under gpurun: bash scripts/gpu_sweep_fused.sh <tag> "<cases>" "<batches>" opt1 opt2 ...
set -u
TAG=$1; CASES=$2; BATCHES=$3; shift 3
OUT=gpurun_out; mkdir -p $OUT
: > $OUT/${TAG}.jsonl
for O in "$@"; do
  OPT=""; [ "$O" != "default" ] && OPT="--opts $O"
  timeout 300 python scripts/bench_fused.py --batches $BATCHES --cases $CASES $OPT 2>&1 | grep '^{' | while read -r L; do echo "{\"opts\": \"$O\", ${L:1}"; done >> $OUT/${TAG}.jsonl
done
python - <<PY
import json
for l in open("$OUT/${TAG}.jsonl"):
    d=json.loads(l); print(d["opts"].ljust(52), d["case"].ljust(18), str(d["batch"]).rjust(5), str(d["fused_us"]).rjust(8), d["frac_of_measured_hbm"])
PY
