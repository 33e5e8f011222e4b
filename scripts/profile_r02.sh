#!/bin/bash
# Round-2 profile pass (run under gpurun, ONE GPU).  Each ncu pass follows a plain run of the same command that exited 0.
#   1. launch list of the bench step           -> gpurun_out/r02_launches.csv
#   2. ncu --set full: the three training kernels (perf_train.py --only step)   -> gpurun_out/r02_train.ncu-rep
#   3. ncu --set full: the render kernel (800x800x64)                           -> gpurun_out/r02_render.ncu-rep
#   4. ncu --set full: the HEAD tcgen05 kernel (800x800x64 fenceposts)           -> gpurun_out/r02_head.ncu-rep
set -u
mkdir -p gpurun_out
NCU="ncu --clock-control none"
python scripts/perf_train.py --iters 3 --only step > gpurun_out/r02_perf_train.json 2>&1 || { echo "perf_train failed"; exit 1; }
$NCU --metrics gpu__time_duration.sum -c 300 --csv --log-file gpurun_out/r02_launches.csv python scripts/perf_train.py --iters 2 --only step > gpurun_out/r02_launches.log 2>&1
echo "launch list rc=$?"
$NCU --set full --import-source on -k regex:'g1_fused_tc_kernel|g1_bwd_tc_kernel|g1_wgrad_tc_kernel' -s 6 -c 3 -o gpurun_out/r02_train -f python scripts/perf_train.py --iters 2 --only step > gpurun_out/r02_ncu_train.log 2>&1
echo "ncu train rc=$?"
python scripts/perf_render.py --iters 2 > gpurun_out/r02_perf_render.json 2>&1 || { echo "perf_render failed"; exit 1; }
$NCU --set full --import-source on -k regex:'g1_fused_tc_kernel' -s 1 -c 1 -o gpurun_out/r02_render -f python scripts/perf_render.py --iters 1 > gpurun_out/r02_ncu_render.log 2>&1
echo "ncu render rc=$?"
python scripts/perf_head.py --hw 800 --prec bf16 --iters 2 > gpurun_out/r02_perf_head.json 2>&1 || { echo "perf_head failed"; exit 1; }
$NCU --set full --import-source on -k regex:'head_tc_kernel|head_composite_kernel' -s 2 -c 2 -o gpurun_out/r02_head -f python scripts/perf_head.py --hw 800 --prec bf16 --iters 1 > gpurun_out/r02_ncu_head.log 2>&1
echo "ncu head rc=$?"
ls -la gpurun_out/r02_*
