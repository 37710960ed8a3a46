# ncu --set full of the value Gram kernel (one capture per kernel family)
set -x
CMD="python tools/gram_one.py 16384 2"
$CMD > gpurun_out/gram_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gram_rows_kernel -c 4 -f -o gpurun_out/s9_gram_rows $CMD > gpurun_out/gram_ncu.log 2>&1
tail -n 3 gpurun_out/gram_plain.log
