set -x
timeout 120 python tools/exp_stage.py 256 textured > gpurun_out/r2c5_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"fast_cells" -s 3 -c 1 -o gpurun_out/r2c5_fastcells -f python tools/exp_stage.py 256 textured > gpurun_out/r2c5_ncu.log 2>&1
tail -3 gpurun_out/r2c5_ncu.log
