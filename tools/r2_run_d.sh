# round 2, GPU call D: small-kernel rewrite + D steps off the critical path: op probes, determinism, bench, launch list
set -x
mkdir -p gpurun_out
timeout 600 python tools/gpu_probe.py linear_relu linear_plain_n3 linear_wide style_mlp style_mlp_base dis_head dis_head_s8 misc_ops ln_relu_n32 determinism_bench step_dim64_256_B8_light graph_vs_eager_dc > gpurun_out/r2d_probe.log 2>&1
cut -c1-700 gpurun_out/r2d_probe.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-baselines > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2d_bench.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], d['roofline']['frac'], d['roofline_hbm']['frac'])"
MT_FAST_LOSS=0 timeout 600 python bench.py --steps 10 --warmup 3 --no-baselines > gpurun_out/r2d_bench_slowloss.json 2> /dev/null
python -c "
import json; d=json.load(open('gpurun_out/r2d_bench_slowloss.json')); print('MT_FAST_LOSS=0', d['value'], d['ms_per_step'])"
E="MT_CUDA_GRAPH=0 MT_CONCURRENT_D=0 MT_WGRAD_SIDE=0 MT_PHASE_FORK=0 MT_PREPACK=0"
env $E python tools/profile_step.py > gpurun_out/r2d_plain_step.log 2>&1 && \
env $E ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2d_launches_step_b8.csv python tools/profile_step.py > gpurun_out/r2d_ncu_step.log 2>&1
python tools/summarize_launches.py gpurun_out/r2d_launches_step_b8.csv > gpurun_out/r2d_launches_summary.txt; head -45 gpurun_out/r2d_launches_summary.txt
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r2d_pytest.log 2>&1; tail -5 gpurun_out/r2d_pytest.log | cut -c1-500
