# Regenerates the files under profiles/ that describe the CURRENT kernels (run on a B200 box from the repo
# root; outputs land in gpurun_out/ and are copied into profiles/ by hand after a look).
set -x
python bench.py > gpurun_out/r1_bench_1gpu.json 2> gpurun_out/bench.err
python tools/graph_profile.py 2>/dev/null | grep -v "^initialize" > gpurun_out/graph_profile.txt
MT_CONCURRENT_D=0 MT_WGRAD_SIDE=0 MT_PHASE_FORK=0 MT_PREPACK=0 python tools/graph_profile.py 2>/dev/null | grep -v "^initialize" > gpurun_out/graph_profile_serial.txt
if [ "$1" = "configs" ]; then
{
MT_B=32 python tools/run_config.py
MT_B=16 MT_SIZE=512 python tools/run_config.py
MT_MODEL=BaseModel python tools/run_config.py
MT_MODEL=BaseModel MT_FLAGS=concat,reparam python tools/run_config.py
MT_FLAGS=use_dis_content,ms_dis python tools/run_config.py
MT_FLAGS=dis_sn python tools/run_config.py
MT_FLAGS=use_dropout python tools/run_config.py
MT_FLAGS=up_type=nearest python tools/run_config.py
MT_FLAGS=enc_norm=layer,dec_norm=instance,dis_norm=instance python tools/run_config.py
} 2>/dev/null | grep -v "^initialize" > gpurun_out/other_configs.txt
cat gpurun_out/other_configs.txt
fi
E="MT_CUDA_GRAPH=0 MT_CONCURRENT_D=0 MT_WGRAD_SIDE=0 MT_PHASE_FORK=0 MT_PREPACK=0"
env $E python tools/profile_step.py > gpurun_out/plain_step.log 2>&1 && \
env $E ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_step_b8.csv python tools/profile_step.py > gpurun_out/ncu_step.log 2>&1
tail -2 gpurun_out/ncu_step.log
python tools/summarize_launches.py gpurun_out/launches_step_b8.csv > gpurun_out/launches_summary.txt; head -12 gpurun_out/launches_summary.txt
