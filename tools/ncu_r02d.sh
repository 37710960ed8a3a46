#!/bin/bash
# Round-2 late ncu evidence (one gpurun call, 1 GPU): staircase GEMM, triangular-B panel solve, PCA projection.
set -u
mkdir -p gpurun_out
python tools/misc2_one.py > gpurun_out/r02d_misc2_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/r02d_misc2_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"dgemm_kernel" -s 5 -c 5 -f -o gpurun_out/r02d_gemm_forms python tools/misc2_one.py > gpurun_out/r02d_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"pca_project" -s 2 -c 2 -f -o gpurun_out/r02d_pca_project python tools/misc2_one.py > gpurun_out/r02d_ncu2.log 2>&1
ls -la gpurun_out/*.ncu-rep; tail -2 gpurun_out/r02d_ncu1.log gpurun_out/r02d_ncu2.log
