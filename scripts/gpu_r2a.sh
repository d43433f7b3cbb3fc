#!/bin/bash
# Round-2 GPU pass A: parity tests, smoke, host-side profile of the drop-in call, bench at the driver's settings.
set -u
TAG=${1:-r02a}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_smi.txt 2>&1
lscpu | head -20 > $OUT/${TAG}_lscpu.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/${TAG}_pytest.log
tail -5 $OUT/${TAG}_pytest.log
timeout 300 python __graft_entry__.py --smoke > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/${TAG}_smoke.log
timeout 300 python scripts/profile_e2e.py > $OUT/${TAG}_profile_e2e.log 2>&1; echo "profile rc=$?"
head -40 $OUT/${TAG}_profile_e2e.log
timeout 900 python bench.py --steps 20 --warmup 5 > $OUT/${TAG}_bench20.json 2> $OUT/${TAG}_bench20.err; echo "bench20 rc=$?"
tail -c 1500 $OUT/${TAG}_bench20.err
python - <<PY
import json
d=json.loads(open("$OUT/${TAG}_bench20.json").read().strip().splitlines()[-1])
for k in ("value","ms_per_step","roofline","clocks","timing","cpu_baseline","torch_cuda_reference"):
    v=d.get(k)
    if k=="roofline": v={a:v[a] for a in ("achieved","frac")}
    print(k, v)
e=d["e2e"]; print({k:v for k,v in e.items() if not k.endswith("note") and k!="api"})
print(d.get("fused_configs")); print(d.get("dedup_ring")); print(d.get("batch_sweep"))
PY
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $OUT/${TAG}_benchref.json 2> $OUT/${TAG}_benchref.err; echo "benchref rc=$?"
cut -c1-300 $OUT/${TAG}_benchref.json
