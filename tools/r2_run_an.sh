timeout 900 python -m pytest tests/test_kernels_gpu.py -x -q 2>&1 | tail -4
for C in 1 0; do echo "MT_HALO_S2=$C"; MT_HALO_S2=$C timeout 600 python tools/ncu_conv_classes.py 2>/dev/null | grep -v wgrad | grep "down 3x3 s2 64\|convT 3x3 s2 128"; done
for C in 1 0 1 0; do
MT_HALO_S2=$C timeout 600 python bench.py --steps 20 --warmup 5 --no-baselines 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('halo_s2=$C', d['value'], d['ms_per_step'], d['e2e']['value'], d['clocks']['sm_mhz'], d['roofline']['achieved'])"; done
