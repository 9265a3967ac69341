#!/usr/bin/env bash
# profiling record of the kernels outside the headline step: stage benches (with clock records) + ncu --set full of their kernels.
# usage (on the box): bash tools/gpu_profiles.sh <tag>
TAG="${1:-prof}"
O=gpurun_out; mkdir -p $O
python tools/bench_ghostpen.py 119 10 > $O/${TAG}_ghostpen.json 2> $O/${TAG}_ghostpen.err; echo "ghostpen rc=$?"; tail -c 300 $O/${TAG}_ghostpen.err
python tools/bench_nxfem.py 94 10 > $O/${TAG}_nxfem.json 2> $O/${TAG}_nxfem.err; echo "nxfem rc=$?"; tail -c 300 $O/${TAG}_nxfem.err
python tools/bench_tracefem.py 94 10 > $O/${TAG}_tracefem.json 2> $O/${TAG}_tracefem.err; echo "tracefem rc=$?"; tail -c 300 $O/${TAG}_tracefem.err
ncu --set full --clock-control none --import-source on -k regex:"k_facetpatch_c|k_rows_packed" -s 8 -c 2 -f -o $O/${TAG}_prof_ghostpen python tools/bench_ghostpen.py 119 2 > $O/${TAG}_ncu_ghostpen.log 2>&1; tail -n 1 $O/${TAG}_ncu_ghostpen.log
ncu --set full --clock-control none --import-source on -k regex:"k_xfem_mom" -s 4 -c 1 -f -o $O/${TAG}_prof_nxfem python tools/bench_nxfem.py 94 2 > $O/${TAG}_ncu_nxfem.log 2>&1; tail -n 1 $O/${TAG}_ncu_nxfem.log
ncu --set full --clock-control none --import-source on -k regex:"k_cutmat_mom|k_fused_sweep|k_geom|k_classify_cut|k_int_cut" -s 12 -c 6 -f -o $O/${TAG}_prof_step python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/${TAG}_ncu_step.log 2>&1; tail -n 1 $O/${TAG}_ncu_step.log
python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"
ls -la $O | grep ${TAG}
