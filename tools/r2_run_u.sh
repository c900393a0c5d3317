python tools/graph_profile.py 2>/dev/null | grep -v "^initialize" > gpurun_out/r2u_graph_profile.txt
MT_CONCURRENT_D=0 MT_WGRAD_SIDE=0 MT_PHASE_FORK=0 MT_PREPACK=0 python tools/graph_profile.py 2>/dev/null | grep -v "^initialize" > gpurun_out/r2u_graph_profile_serial.txt
head -45 gpurun_out/r2u_graph_profile_serial.txt
grep -A30 "^span" gpurun_out/r2u_graph_profile.txt
