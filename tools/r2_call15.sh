timeout 600 ncu --set full --clock-control none --import-source on -k regex:"resize_level_words" -c 2 -o gpurun_out/r02_resize_v3 -f python tools/exp_stage.py 1024 textured > gpurun_out/r02_resize_v3.log 2>&1
tail -2 gpurun_out/r02_resize_v3.log
