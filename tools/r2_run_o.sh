for D in 0 64 0 64 0 64; do
MT_IGEMM_DBG=$D timeout 600 python bench.py --steps 20 --warmup 5 --no-baselines 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('dbg=$D', d['value'], d['ms_per_step'], d['e2e']['value'], d['clocks']['sm_mhz'], d['roofline']['achieved'])"; done
