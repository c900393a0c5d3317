MT_PDL=0 MT_CONCURRENT_D=0 MT_WGRAD_SIDE=0 MT_PHASE_FORK=0 MT_PREPACK=0 python tools/graph_profile.py 2>/dev/null | grep -v "^initialize" > gpurun_out/r2_graph_profile_insitu_serial_new.txt
head -48 gpurun_out/r2_graph_profile_insitu_serial_new.txt
python tools/graph_profile.py 2>/dev/null | grep -v "^initialize" > gpurun_out/r2_graph_profile_insitu_new.txt
grep "^span" gpurun_out/r2_graph_profile_insitu_new.txt
