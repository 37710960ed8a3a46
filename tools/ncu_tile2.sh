PYTHONPATH=. python tools/tile_probe.py > gpurun_out/tile_plain.log 2>&1 &&
PYTHONPATH=. ncu --set full --clock-control none --import-source on -k regex:potrf_tile2 -s 4 -c 1 -f -o gpurun_out/r01d_tile2 python tools/tile_probe.py > gpurun_out/ncu_tile2d.log 2>&1
cat gpurun_out/tile_plain.log
