timeout 900 python -m pytest tests/test_kernels_gpu.py -x -q -k "conv_z or linear or misc" 2>&1 | tail -4
timeout 900 python -m pytest tests/test_model_gpu.py -x -q -k "base" 2>&1 | tail -4
for C in 1 0 1 0; do
MT_FOLD_Z=$C timeout 900 python bench.py --config 5-base --steps 6 --warmup 3 --no-baselines 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('fold_z=$C', d['value'], d['ms_per_step'], d['e2e']['value'], d['clocks']['sm_mhz'])"; done
