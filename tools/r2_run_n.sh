bash tools/r2_run_m.sh > gpurun_out/r2n_timeline.log 2>&1
grep -A1 "===\|epilogue of\|back-to-back" gpurun_out/r2n_timeline.log | grep -v "^tile\|^--\|^shape\|ring"
timeout 900 python -m pytest tests/test_kernels_gpu.py -x -q 2>&1 | tail -3
for i in 1 2; do timeout 600 python bench.py --steps 20 --warmup 5 --no-baselines 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['clocks']['sm_mhz'], d['roofline']['achieved'])"; done
