timeout 900 ncu --set full --clock-control none --import-source on -k regex:"top2_scan" -c 1 -o gpurun_out/r02_full_match -f python bench.py --steps 1 --warmup 0 --no-cpu --no-configs --frames 64 > gpurun_out/r02_ncu_full_match.log 2>&1
tail -2 gpurun_out/r02_ncu_full_match.log
