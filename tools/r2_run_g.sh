set -x
mkdir -p gpurun_out
timeout 600 python tools/gpu_probe.py misc_ops misc_ops_c64 misc_ops_c32 misc_ops_c128 misc_ops_c256 dgrad_accumulate determinism_bench step_dim64_256_B8_light > gpurun_out/r2g_probe.log 2>&1
cut -c1-900 gpurun_out/r2g_probe.log
timeout 300 python tools/host_gap.py 2>/dev/null | tail -8
timeout 600 python bench.py --steps 20 --warmup 5 --no-baselines > gpurun_out/r2g_bench.json 2> gpurun_out/r2g_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2g_bench.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], d['roofline']['frac'], d['roofline_hbm']['frac'])"
MT_FUSE_RES_GRAD=0 timeout 600 python bench.py --steps 20 --warmup 5 --no-baselines 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('MT_FUSE_RES_GRAD=0', d['value'], d['ms_per_step'])"
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r2g_pytest.log 2>&1; tail -5 gpurun_out/r2g_pytest.log | cut -c1-500
