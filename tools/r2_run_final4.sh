timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
for c in 3 4 5-adain; do
timeout 900 python bench.py --config $c --steps 10 --warmup 3 --no-baselines 2>/dev/null > gpurun_out/r2h_bench_cfg$c.json; python -c "
import json; d=json.load(open('gpurun_out/r2h_bench_cfg$c.json')); print('cfg $c', d['value'], d['ms_per_step'], d['e2e']['value'], d['clocks'])"
done
