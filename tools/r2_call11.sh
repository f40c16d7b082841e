set -x
# item 10: TMA 2-D tile probe, exact output
( echo "# tools/probe_tma_tile.cu on $(nvidia-smi --query-gpu=name,driver_version --format=csv,noheader): uint8 640x480 image, pitch 640, box 48x43,"
  echo "# element strides {1,1}, CU_TENSOR_MAP_INTERLEAVE_NONE, SWIZZLE_NONE, L2_PROMOTION_NONE, OOB_FILL_NONE; tile origin (x, 450): rows 480.. are out of bounds"
  echo "# variant 0 = libcu++ cde::cp_async_bulk_tensor_2d_global_to_shared; 1 = raw PTX, descriptor as __grid_constant__ param; 2 = raw PTX, descriptor in global memory"
  for v in 0 1 2; do for x in 0 16 13 27 600 608; do timeout 30 tools/_probe/tma_v${v}_x${x} 2>&1 || echo "exit $?"; done; done ) > gpurun_out/r02_tma_probe.txt
cat gpurun_out/r02_tma_probe.txt
# launch list of the bench command
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu --no-configs > gpurun_out/r02_plain.json 2>/dev/null; tail -c 200 gpurun_out/r02_plain.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches_1024frames.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-configs > gpurun_out/r02_ncu_launches.log 2>&1
tail -3 gpurun_out/r02_launches_1024frames.csv
