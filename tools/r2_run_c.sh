# round 2, GPU call C: full GPU suite + smoke + bench + launch list of one iteration
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r2c_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2c_pytest.log
tail -25 gpurun_out/r2c_pytest.log | cut -c1-600
timeout 600 python __graft_entry__.py smoke > gpurun_out/r2c_smoke.log 2>&1; tail -2 gpurun_out/r2c_smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2c_bench.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], d['roofline']['frac'], d['roofline_hbm']['frac'])"
E="MT_CUDA_GRAPH=0 MT_CONCURRENT_D=0 MT_WGRAD_SIDE=0 MT_PHASE_FORK=0 MT_PREPACK=0"
env $E python tools/profile_step.py > gpurun_out/r2c_plain_step.log 2>&1 && \
env $E ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2c_launches_step_b8.csv python tools/profile_step.py > gpurun_out/r2c_ncu_step.log 2>&1
tail -2 gpurun_out/r2c_ncu_step.log
python tools/summarize_launches.py gpurun_out/r2c_launches_step_b8.csv > gpurun_out/r2c_launches_summary.txt; head -40 gpurun_out/r2c_launches_summary.txt
