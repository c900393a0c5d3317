set -x
mkdir -p gpurun_out
for P in 1 0 1; do
MT_PDL=$P timeout 600 python bench.py --steps 20 --warmup 5 --no-baselines 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('MT_PDL=$P', d['value'], d['ms_per_step'], d['e2e']['value'])"
done
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/r2k_pytest.log 2>&1; tail -4 gpurun_out/r2k_pytest.log | cut -c1-600
timeout 600 python __graft_entry__.py smoke 2>&1 | tail -1
