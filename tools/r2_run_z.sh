timeout 900 python -m pytest tests/test_kernels_gpu.py -x -q 2>&1 | tail -3
for C in 0 1; do echo "MT_WGRAD_TAPS=$C"; MT_WGRAD_TAPS=$C timeout 600 python tools/ncu_conv_classes.py 2>&1 | grep "wgrad" | grep stem; done
for C in 0 1 0 1; do
MT_WGRAD_TAPS=$C timeout 600 python bench.py --steps 20 --warmup 5 --no-baselines 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('wgrad_taps=$C', d['value'], d['ms_per_step'], d['e2e']['value'], d['clocks']['sm_mhz'], d['roofline']['achieved'])"; done
