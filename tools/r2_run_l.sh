for P in 1 0 1 0 1; do
MT_PDL=$P timeout 600 python bench.py --steps 20 --warmup 5 --no-baselines 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('MT_PDL=$P', d['value'], d['ms_per_step'], d['e2e']['value'], d['clocks']['sm_mhz'])"
done
