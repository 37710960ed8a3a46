set -x
CMD="python tools/pca_probe.py"
$CMD > gpurun_out/r01c_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:pca_ -c 24 -f -o gpurun_out/r01c_pca $CMD > gpurun_out/r01c_ncu.log 2>&1
tail -3 gpurun_out/r01c_plain.log
tail -3 gpurun_out/r01c_ncu.log
