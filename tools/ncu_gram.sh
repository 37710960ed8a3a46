# ncu --set full of the value Gram kernel (one capture per kernel family)
set -x
CMD="python tools/gram_one.py 16384 2"
$CMD > gpurun_out/gram_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:gram_rows_kernel|gram_dcontract_kernel" -c 6 -f -o gpurun_out/r01f_gram $CMD > gpurun_out/gram_ncu.log 2>&1
tail -n 3 gpurun_out/gram_plain.log
