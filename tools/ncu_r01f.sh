# launch list after the Gram / derivative-contraction rewrite: one ncu invocation per gpurun call
set -x
CMD="python bench.py --steps 1 --warmup 3 --skip-cpu --skip-checks --skip-predict --skip-loglik --skip-pca --skip-sparse"
$CMD > gpurun_out/r01f_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/r01f_launches.csv $CMD > gpurun_out/r01f_ncu_launches.log 2>&1
tail -c 600 gpurun_out/r01f_plain.log
