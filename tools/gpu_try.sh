#!/usr/bin/env bash
# build (abort on failure) -> CPU tests -> one gpurun call: GPU tests, bench, optional ncu capture of kernels matching $1
set -eo pipefail
cd "$(dirname "$0")/.."
python __graft_entry__.py 2> /tmp/build.err || { grep -v deprecated /tmp/build.err | head -30; exit 1; }
python -m pytest tests -x -q -m "not gpu" 2>&1 | tail -1
TAG="${2:-try}"
NCU=""
if [ -n "$1" ]; then
  NCU="&& ncu --set full --clock-control none --import-source on -k regex:\"$1\" -s 1 -c ${3:-1} -o gpurun_out/prof_$TAG python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_$TAG.log 2>&1"
fi
/usr/local/graft/bin/gpurun --timeout 1200 -- "(timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 8); python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err $NCU; python -c \"
import json;d=json.load(open('gpurun_out/bench_$TAG.json'));print(d['ms_per_step'],d['config']['stage_ms'],d['roofline']['kernel_ms'],d['e2e']['ms_per_step'])\"; tail -n 3 gpurun_out/bench_$TAG.err"
