set -x
mkdir -p gpurun_out
timeout 600 python tools/gpu_probe.py misc_ops misc_ops_c64 misc_ops_c32 misc_ops_c128 linear_relu linear_plain_n3 linear_wide style_mlp style_mlp_base determinism_bench > gpurun_out/r2i_probe.log 2>&1
cut -c1-600 gpurun_out/r2i_probe.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-baselines > gpurun_out/r2i_bench.json 2> gpurun_out/r2i_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2i_bench.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], d['roofline']['frac'], d['roofline_hbm']['frac'])"
python tools/graph_profile.py 2>/dev/null | grep -v "^initialize" > gpurun_out/r2i_graph_profile.txt
grep -E "rgb_head|linear_|colsum|dis_head|^span|^sum" gpurun_out/r2i_graph_profile.txt
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r2i_pytest.log 2>&1; tail -3 gpurun_out/r2i_pytest.log | cut -c1-500
