#!/usr/bin/env bash
# symmetric storage + fast element vectors: parity tests, stage tools.  usage (on the box): bash tools/gpu_sym.sh <tag>
TAG="${1:-sym}"
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -q -m gpu -k "symmetric or elvec or linear or rhs or xfem or nxfem" --maxfail=8 > $O/${TAG}_tests.log 2>&1; echo "tests rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed|Error" $O/${TAG}_tests.log | tail -n 12
python tools/bench_symmetric.py 119 10 > $O/${TAG}_symmetric.json 2> $O/${TAG}_symmetric.err; echo "sym rc=$?"; tail -c 400 $O/${TAG}_symmetric.err; cat $O/${TAG}_symmetric.json
python tools/bench_deformed_ops.py 94 10 > $O/${TAG}_defops.json 2> $O/${TAG}_defops.err; echo "defops rc=$?"; tail -c 300 $O/${TAG}_defops.err; python -c "
import json; d=json.load(open('$O/${TAG}_defops.json')); print({k:v for k,v in d['ms'].items() if 'Linear' in k}, d.get('rhs_elements_per_s_affine'))"
