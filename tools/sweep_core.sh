#!/usr/bin/env bash
# build variants of ngsx_core.cu (tuning knobs) into build/variants/, then time each on the GPU box
# usage: tools/sweep_core.sh "name:-DNGSX_INT_MINB=2" ...
set -eo pipefail
cd "$(dirname "$0")/.."
python __graft_entry__.py 2>/dev/null
mkdir -p build/variants
C=ngsxfem_b200/csrc
names=()
for v in "$@"; do
  n="${v%%:*}"; f="${v#*:}"; names+=("$n")
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC $f -c $C/ngsx_core.cu -o build/variants/core_$n.o
  nvcc -shared -o build/variants/lib_$n.so build/variants/core_$n.o $C/ngsx_assemble.o $C/ngsx_rows.o $C/ngsx_rowsmat.o $C/ngsx_graph.o $C/ngsx_facetpatch.o $C/ngsx_xfem.o $C/ngsx_spacetime.o 2>/dev/null
done
cmd="cp ngsxfem_b200/libngsx_b200.so build/variants/lib_base.so; for n in base ${names[*]}; do cp build/variants/lib_\$n.so ngsxfem_b200/libngsx_b200.so; touch ngsxfem_b200/libngsx_b200.so; python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c \"import json,sys;d=json.loads(sys.stdin.read());print('\$n',d['ms_per_step'],d['config']['stage_ms'])\"; done"
/usr/local/graft/bin/gpurun --timeout 1500 -- "$cmd" 2>&1 | grep -v "^\[gpurun\] sending"
