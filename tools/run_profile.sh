set -x
python -m pytest tests -m gpu -x -q > gpurun_out/final_gpu_tests.log 2>&1; tail -2 gpurun_out/final_gpu_tests.log
python bench.py --impl reference > gpurun_out/final_bench_ref.json 2> gpurun_out/final_bench_ref.err
python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err
tail -c 300 gpurun_out/final_bench.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/final_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/final_ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"orient_describe|blur_kernel|quadtree_select|fast_cells|resize_level_words|top2_scan" -c 12 -o gpurun_out/final_full -f python bench.py --steps 1 --warmup 0 --no-cpu > gpurun_out/final_ncu_full.log 2>&1
tail -2 gpurun_out/final_ncu_full.log
