#!/usr/bin/env bash
# order-3 job on the box: the order-3 parity tests, the stage timing (tools/bench_order3.py) and an ncu capture of k_elmat3.
# usage: bash tools/gpu_order3.sh <tag> [nx]
TAG="${1:-o3}"; NX="${2:-80}"
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -q -m gpu -k "order3" --maxfail=8 > $O/${TAG}_tests.log 2>&1; echo "tests rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed|Error" $O/${TAG}_tests.log | tail -n 12
python tools/bench_order3.py $NX 10 > $O/${TAG}_order3.json 2> $O/${TAG}_order3.err; echo "bench rc=$?"; tail -c 400 $O/${TAG}_order3.err; cat $O/${TAG}_order3.json
ncu --set full --clock-control none --import-source on -k regex:k_elmat3 -s 6 -c 2 -f -o $O/${TAG}_prof_elmat3 python tools/bench_order3.py $NX 2 > $O/${TAG}_ncu.log 2>&1; tail -n 3 $O/${TAG}_ncu.log
