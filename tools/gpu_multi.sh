#!/usr/bin/env bash
# one multi-GPU bench run on the box: bash tools/gpu_multi.sh <tag> <ngpus> [ENV=..;ENV=..]
TAG="$1"; N="$2"; CFG="${3:-}"
O=gpurun_out; mkdir -p $O
env $(echo "$CFG" | tr ';' ' ') python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline > $O/${TAG}_n$N.json 2> $O/${TAG}_n$N.err || tail -c 600 $O/${TAG}_n$N.err
python - <<PY
import json
try:
    d=json.load(open("$O/${TAG}_n$N.json"))
    print("N", d["n_gpus"], "$CFG", "ms/step", round(d["ms_per_step"],4), "value", round(d["value"]/1e9,3), "G el/s  e2e", round(d["e2e"]["ms_per_step"],3), d["config"]["stage_ms"])
    print(" per rank", d["config"].get("per_rank"), d["config"].get("numa"))
except Exception as e: print("parse failed", e)
PY
