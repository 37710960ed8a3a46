#!/bin/bash
# Round-2 ncu evidence (one gpurun call, 1 GPU): launch list of one Learn at N = 32768, full captures of the new kernels.
set -u
mkdir -p gpurun_out
python tools/chol_one.py 32768 > gpurun_out/r02_one_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/r02_one_plain.log; exit 1; }
L=$(awk '{print $2}' gpurun_out/r02_one_plain.log | tail -1)
ncu --metrics gpu__time_duration.sum --clock-control none -s $((L + 2)) -c $((L + 2)) --csv --log-file gpurun_out/r02_launches.csv \
    python tools/chol_one.py 32768 > gpurun_out/r02_one_ncu.log 2>&1
python tools/chol_one.py 2048 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:potrf_tile3 -s 20 -c 1 -f -o gpurun_out/r02_tile3 python tools/chol_one.py 2048 > gpurun_out/r02_tile3_ncu.log 2>&1
python tools/misc_one.py > gpurun_out/r02_misc_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"trmv_lower|gram_mma_transform|centre_norm" -c 12 -f -o gpurun_out/r02_misc python tools/misc_one.py > gpurun_out/r02_misc_ncu.log 2>&1
ls -la gpurun_out/*.ncu-rep
