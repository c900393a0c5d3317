# final evidence, 1 GPU: the other BASELINE configs
for c in 4 5-adain 5-base; do
  timeout 900 python bench.py --config $c --steps 10 --warmup 3 --no-baselines > gpurun_out/r2_bench_cfg${c}_1gpu.json 2> gpurun_out/r2_bench_cfg${c}_1gpu.err; echo "cfg $c rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/r2_bench_cfg${c}_1gpu.json')); print('cfg $c N=1', d['value'], d['ms_per_step'], d['e2e']['value'], d['clocks'])"
done
