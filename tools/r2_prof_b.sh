timeout 900 ncu --set full --clock-control none --import-source on -k regex:"orient_describe|blur_kernel|quadtree_select|quadtree_pyramid|fast_cells|resize_level_words" -c 12 -o gpurun_out/r02_full -f python bench.py --steps 1 --warmup 0 --no-cpu --no-configs --no-match > gpurun_out/r02_ncu_full.log 2>&1
tail -2 gpurun_out/r02_ncu_full.log
