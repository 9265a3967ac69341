#!/usr/bin/env bash
# one comprehensive GPU-box job (the pod's slots are scarce): GPU tests, bench (both arms), launch list, ncu captures, stage benches.
# Everything lands in gpurun_out/.  usage (on the box): bash tools/gpu_job.sh <tag> [quick]
TAG="${1:-job}"; MODE="${2:-full}"
O=gpurun_out
mkdir -p $O
python __graft_entry__.py > $O/${TAG}_build.log 2>&1 || { tail -20 $O/${TAG}_build.log; exit 1; }
timeout 1200 python -m pytest tests -q -m gpu --maxfail=12 > $O/${TAG}_tests.log 2>&1; echo "tests rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed" $O/${TAG}_tests.log | tail -n 16
python bench.py --steps 20 --warmup 5 > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"; tail -c 300 $O/${TAG}_bench.err
python - <<PY
import json
try:
    d=json.load(open("$O/${TAG}_bench.json"))
    print("ms/step", d["ms_per_step"], "stage", d["config"]["stage_ms_one_stream"], "e2e", d["e2e"]["ms_per_step"], "parity", (d.get("parity") or {}).get("ok"))
    print("roofline", {k: d["roofline"][k] for k in ("achieved","frac")}, d["roofline"]["assemble_stage"])
except Exception as e: print("bench parse failed", e)
PY
if [ "$MODE" = "quick" ]; then exit 0; fi
python bench.py --impl reference --steps 3 --warmup 1 > $O/${TAG}_ref.json 2> $O/${TAG}_ref.err; echo "ref rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/${TAG}_ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_rows -s 2 -c 1 -f -o $O/${TAG}_prof_rows python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/${TAG}_ncu_rows.log 2>&1
python tools/bench_deformed.py 119 10 2 > $O/${TAG}_deformed.json 2> $O/${TAG}_deformed.err; echo "deformed rc=$?"; tail -c 300 $O/${TAG}_deformed.err
ncu --set full --clock-control none --import-source on -k regex:k_rows_packed -s 2 -c 1 -f -o $O/${TAG}_prof_packed python tools/bench_deformed.py 119 2 2 > $O/${TAG}_ncu_packed.log 2>&1
ls -la $O | tail -20
