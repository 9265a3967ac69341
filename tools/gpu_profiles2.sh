#!/usr/bin/env bash
# second profiling job: ghost-penalty stage + ncu of k_facetpatch_c / k_rows_packed, operators on deformed meshes + RHS (tools/bench_deformed_ops.py),
# one rank's share of an 8-GPU run on one GPU (NGSX_BENCH_EMULATE).  usage (on the box): bash tools/gpu_profiles2.sh <tag>
TAG="${1:-prof2}"
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests -q -m gpu -k "order3 or spacetime or facet or ghost" --maxfail=5 > $O/${TAG}_tests.log 2>&1; echo "tests rc=$?"; tail -n 2 $O/${TAG}_tests.log
python tools/bench_ghostpen.py 119 10 > $O/${TAG}_ghostpen.json 2> $O/${TAG}_ghostpen.err; echo "ghostpen rc=$?"; tail -c 300 $O/${TAG}_ghostpen.err
python tools/bench_deformed_ops.py 94 10 > $O/${TAG}_defops.json 2> $O/${TAG}_defops.err; echo "defops rc=$?"; tail -c 300 $O/${TAG}_defops.err; cat $O/${TAG}_defops.json
ncu --set full --clock-control none --import-source on -k regex:"k_facetpatch_c|k_rows_packed" -s 8 -c 2 -f -o $O/${TAG}_prof_ghostpen python tools/bench_ghostpen.py 119 2 > $O/${TAG}_ncu_ghostpen.log 2>&1; tail -n 1 $O/${TAG}_ncu_ghostpen.log
for r in 0 3; do
  NGSX_BENCH_EMULATE=$r/8 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/${TAG}_emu$r.json 2> $O/${TAG}_emu$r.err; echo "emu$r rc=$?"
  python - <<PY
import json
try:
    d=json.load(open("$O/${TAG}_emu$r.json")); c=d["config"]
    print("emu $r/8 ms/step", round(d["ms_per_step"],4), "stage", {k:round(v,4) for k,v in c["stage_ms"].items()}, "serial", c["stage_ms_one_stream"])
except Exception as e: print("parse failed", e)
PY
done
