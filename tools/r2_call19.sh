ORBX_E2E_TRACE=3 timeout 300 python tools/exp_e2e.py 2>&1 | grep -E "chunk 1[3-9]|stream end|pipelined|compute"
for s in "96,96,96,96,96,96,64,64,64,64,48,48,32,32,16,16" "64,64,64,64,64,64,64,64,64,64,64,64,64,64,48,32,24,16,8"; do echo "sched=$s"; ORBX_CHUNK_SCHEDULE=$s timeout 300 python tools/exp_e2e.py 2>&1 | grep -E "pipelined chunk 0"; done
