set -x
mkdir -p gpurun_out
for P in 0 1 0 1; do
MT_PDL=$P timeout 600 python bench.py --steps 20 --warmup 5 --no-baselines 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('MT_PDL=$P', d['value'], d['ms_per_step'], d['e2e']['value'])"
done
MT_PDL=1 timeout 900 python tools/gpu_probe.py determinism_bench step_dim64_256_B8_light hot_n16 in_relu_n16 ln_relu_n32 graph_vs_eager 2>&1 | cut -c1-400
