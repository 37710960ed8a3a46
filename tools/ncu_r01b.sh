set -x
CMD="python bench.py --steps 1 --warmup 3 --skip-cpu --skip-checks --predict-steps 3"
$CMD > gpurun_out/r01b_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 7000 --csv --log-file gpurun_out/r01b_launches.csv $CMD > gpurun_out/r01b_ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:dgemm_kernel -s 15 -c 3 -f -o gpurun_out/r01b_dgemm $CMD --skip-predict > gpurun_out/r01b_ncu_dgemm.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemv_t_kernel -s 3 -c 2 -f -o gpurun_out/r01b_gemv $CMD > gpurun_out/r01b_ncu_gemv.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gram_kernel -c 1 -f -o gpurun_out/r01b_gram $CMD --skip-predict > gpurun_out/r01b_ncu_gram.log 2>&1
tail -2 gpurun_out/r01b_plain.log | cut -c1-600
ls -la gpurun_out
