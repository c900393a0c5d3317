timeout 900 python -m pytest tests/test_kernels_gpu.py -x -q 2>&1 | tail -4
for H in 1 0; do echo "MT_HALO=$H"; MT_HALO=$H MT_SHAPE=3,64,7,1,262,stem MT_STATS=1 timeout 300 python tools/conv_timeline.py 2>&1 | grep "back-to-back"; done
for C in 1 0 1 0; do
MT_HALO=$C timeout 600 python bench.py --steps 20 --warmup 5 --no-baselines 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('halo=$C', d['value'], d['ms_per_step'], d['e2e']['value'], d['clocks']['sm_mhz'], d['roofline']['achieved'])"; done
