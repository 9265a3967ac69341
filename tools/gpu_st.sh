#!/usr/bin/env bash
# space-time job on the box: f-3 parity tests, stage timing (tools/bench_spacetime.py), ncu of the space-time kernels.  usage: bash tools/gpu_st.sh <tag>
TAG="${1:-st}"
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -q -m gpu -k "spacetime" --maxfail=8 > $O/${TAG}_tests.log 2>&1; echo "tests rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed|Error|assert" $O/${TAG}_tests.log | tail -n 20
python tools/bench_spacetime.py 1000 10 > $O/${TAG}_spacetime.json 2> $O/${TAG}_spacetime.err; echo "bench rc=$?"; tail -c 400 $O/${TAG}_spacetime.err; cat $O/${TAG}_spacetime.json
ncu --set full --clock-control none --import-source on -k regex:"k_st_(integrate|elmat|classify)" -s 6 -c 6 -f -o $O/${TAG}_prof_st python tools/bench_spacetime.py 1000 2 > $O/${TAG}_ncu.log 2>&1; tail -n 2 $O/${TAG}_ncu.log
