timeout 900 python -m pytest tests/test_kernels_gpu.py -x -q 2>&1 | tail -5
for S in 256,256,3,1,66 64,128,3,1,130 64,64,3,1,130 128,128,3,1,66; do
for B in 1 0; do MT_WGRAD_BOX=$B MT_SHAPE=$S timeout 300 python tools/wgrad_time.py 2>/dev/null | tail -1 | sed "s/^/box=$B /"; done; done
for C in 1 0 1 0; do
MT_WGRAD_BOX=$C timeout 600 python bench.py --steps 20 --warmup 5 --no-baselines 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('wgrad_box=$C', d['value'], d['ms_per_step'], d['e2e']['value'], d['clocks']['sm_mhz'], d['roofline']['achieved'])"; done
