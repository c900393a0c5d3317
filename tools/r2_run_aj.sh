timeout 900 python -m pytest tests/test_kernels_gpu.py -x -q 2>&1 | tail -3
for C in 1 0 1 0; do
MT_STARVED_BN=$C timeout 600 python bench.py --steps 20 --warmup 5 --no-baselines 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('starved_bn=$C', d['value'], d['ms_per_step'], d['e2e']['value'], d['clocks']['sm_mhz'], d['roofline']['achieved'])"; done
