# final evidence, call 1 (1 GPU): full GPU test suite, full bench line (with baselines), launch list, conv-class ncu table
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r2f_pytest.log 2>&1; tail -3 gpurun_out/r2f_pytest.log | cut -c1-300
timeout 1200 python bench.py --steps 20 --warmup 5 > gpurun_out/r2f_bench_1gpu.json 2> gpurun_out/r2f_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2f_bench_1gpu.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], d['roofline']['frac'], d['roofline_hbm']['frac'], d['clocks'], d['cpu_baseline']['value'], {k: v.get('value') for k, v in d['gpu_baseline'].items() if isinstance(v, dict)}, d['sampling']['value'])"
E="MT_CUDA_GRAPH=0 MT_CONCURRENT_D=0 MT_WGRAD_SIDE=0 MT_PHASE_FORK=0 MT_PREPACK=0"
env $E python tools/profile_step.py > gpurun_out/r2f_plain_step.log 2>&1 && \
env $E ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2f_launches_step_b8.csv python tools/profile_step.py > gpurun_out/r2f_ncu_step.log 2>&1
python tools/summarize_launches.py gpurun_out/r2f_launches_step_b8.csv > gpurun_out/r2f_launches_summary.txt; head -30 gpurun_out/r2f_launches_summary.txt | cut -c1-120
M=gpu__time_duration.sum,sm__inst_executed_pipe_tensor.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,smsp__inst_executed_pipe_uniform.sum,l1tex__m_xbar2l1tex_read_bytes.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__cycles_elapsed.max
python tools/ncu_conv_classes.py > gpurun_out/r2f_classes_plain.txt 2>&1 && \
ncu --profile-from-start off --metrics $M --clock-control none -k regex:igemm --csv --log-file gpurun_out/r2f_ncu_conv_classes.csv python tools/ncu_conv_classes.py > gpurun_out/r2f_ncu_classes.log 2>&1
python tools/ncu_conv_classes.py --join gpurun_out/r2f_ncu_conv_classes.csv > gpurun_out/r2f_ncu_conv_classes_table.txt 2>&1; head -50 gpurun_out/r2f_ncu_conv_classes_table.txt | cut -c1-140
