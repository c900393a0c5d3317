timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py 2>/dev/null > gpurun_out/r2g_bench_default.json; python -c "
import json; d=json.load(open('gpurun_out/r2g_bench_default.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['steps'], d['warmup'], d['roofline']['frac'], d['clocks'])"
timeout 900 python bench.py --config 5-base --steps 6 --warmup 3 --no-baselines 2>/dev/null > gpurun_out/r2_bench_cfg5-base_1gpu.json; python -c "
import json; d=json.load(open('gpurun_out/r2_bench_cfg5-base_1gpu.json')); print('5-base', d['value'], d['ms_per_step'], d['e2e']['value'], d['clocks'])"
