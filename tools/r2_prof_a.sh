timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu --no-configs > gpurun_out/r02_plain.json 2>/dev/null; tail -c 150 gpurun_out/r02_plain.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches_1024frames.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-configs > gpurun_out/r02_ncu_launches.log 2>&1
tail -2 gpurun_out/r02_launches_1024frames.csv | cut -c1-200
