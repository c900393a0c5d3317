set -x
mkdir -p gpurun_out
python tools/graph_profile.py 2>/dev/null | grep -v "^initialize" > gpurun_out/r2h_graph_profile.txt
MT_CONCURRENT_D=0 MT_WGRAD_SIDE=0 MT_PHASE_FORK=0 MT_PREPACK=0 python tools/graph_profile.py 2>/dev/null | grep -v "^initialize" > gpurun_out/r2h_graph_profile_serial.txt
head -50 gpurun_out/r2h_graph_profile_serial.txt
grep -A32 "^span" gpurun_out/r2h_graph_profile.txt
for c in 4 5-adain 5-base; do
  timeout 900 python bench.py --config $c --steps 6 --warmup 3 --no-baselines > gpurun_out/r2h_bench_cfg$c.json 2> gpurun_out/r2h_bench_cfg$c.err; echo "cfg $c rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/r2h_bench_cfg$c.json')); print('$c', d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['all_convs']['frac'], d['roofline_hbm']['frac'])"
done
