timeout 600 ncu --set full --clock-control none --import-source on -k regex:"fast_cells|orient_describe|blur_kernel|quadtree_select" -c 4 -o gpurun_out/r02_b -f python tools/exp_stage.py 1024 textured > gpurun_out/r02_b.log 2>&1
tail -2 gpurun_out/r02_b.log
