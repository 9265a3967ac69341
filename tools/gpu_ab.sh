#!/usr/bin/env bash
# A/B runs of bench.py under environment switches on one GPU; usage (on the box): bash tools/gpu_ab.sh <tag> "ENV=..;ENV=.." ...
TAG="$1"; shift
O=gpurun_out; mkdir -p $O
i=0
for cfg in "$@"; do
  i=$((i+1))
  env $(echo "$cfg" | tr ';' ' ') python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/${TAG}_ab$i.json 2> $O/${TAG}_ab$i.err || tail -5 $O/${TAG}_ab$i.err
  python - <<PY
import json
try:
    d=json.load(open("$O/${TAG}_ab$i.json"))
    print("$cfg", "| ms/step", round(d["ms_per_step"],4), "stage", {k:round(v,4) for k,v in d["config"]["stage_ms"].items()}, "e2e", round(d["e2e"]["ms_per_step"],3), "launch", d["config"]["launch"][:40])
except Exception as e: print("$cfg", "failed", e)
PY
done
