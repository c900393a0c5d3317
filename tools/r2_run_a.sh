# round 2, GPU call A: full GPU test suite with the new op cases, the bench-geometry model cases, the
# determinism checks and a bench line (outputs under gpurun_out/)
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/r2a_smi.txt; nproc >> gpurun_out/r2a_smi.txt; free -g >> gpurun_out/r2a_smi.txt
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2a_pytest.log
tail -5 gpurun_out/r2a_pytest.log
timeout 1500 python tools/gpu_probe.py step_dim8_64 step_dim64_256_B8 step_dim64_256_B8_light graph_vs_eager_bench sampling_256_n8 determinism_small determinism_bench step_hinge step_basec > gpurun_out/r2a_probe.log 2>&1
cat gpurun_out/r2a_probe.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench rc=$?"
cat gpurun_out/r2a_bench.json | head -c 3000
