# round 2, GPU call B: new op cases (linear / style MLP / D heads / losses / modules), full GPU suite, probes, bench
set -x
mkdir -p gpurun_out
timeout 600 python tools/gpu_probe.py linear_relu linear_plain_n3 linear_wide style_mlp style_mlp_base dis_head dis_head_s8 losses module_forwards misc_ops in_big_mean adam_flat > gpurun_out/r2b_probe_ops.log 2>&1
cat gpurun_out/r2b_probe_ops.log
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2b_pytest.log
tail -30 gpurun_out/r2b_pytest.log
timeout 900 python tools/gpu_probe.py step_dim64_256_B8_light determinism_bench step_dim8_64 > gpurun_out/r2b_probe.log 2>&1
cat gpurun_out/r2b_probe.log | cut -c1-1500
timeout 600 python tests/tools/precision_experiment.py 64 256 8 > gpurun_out/r2b_precision.log 2>&1; tail -2 gpurun_out/r2b_precision.log | cut -c1-3000
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2b_bench.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], d['roofline']['frac'], d['roofline_hbm']['frac'])"
