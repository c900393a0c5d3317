# round 2, GPU call E: ncu evidence (one ncu session: per-class tensor-pipe counters, --set full of the hot conv and
# of the norm kernels).  Every command first runs without ncu.
set -x
mkdir -p gpurun_out
M="gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_tensor.sum,l1tex__m_xbar2l1tex_read_bytes.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__cycles_elapsed.avg,sm__throughput.avg.pct_of_peak_sustained_elapsed"
python tools/ncu_conv_classes.py > gpurun_out/r2e_conv_classes_plain.txt 2>&1 && \
ncu --profile-from-start off --metrics $M --clock-control none -k regex:igemm --csv --log-file gpurun_out/r2e_conv_classes.csv python tools/ncu_conv_classes.py > gpurun_out/r2e_conv_classes_ncu.log 2>&1
python tools/ncu_conv_classes.py --join gpurun_out/r2e_conv_classes.csv > gpurun_out/r2e_conv_classes_table.txt 2>&1; cat gpurun_out/r2e_conv_classes_table.txt
cat gpurun_out/r2e_conv_classes_plain.txt | tail -30
# --set full of the hot conv: first measured launch of the class list (res 3x3 256->256 fwd, N = 16)
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:igemm_conv_kernel -c 1 -o gpurun_out/r2e_conv_hot python tools/ncu_conv_classes.py > gpurun_out/r2e_conv_hot_ncu.log 2>&1
ncu -i gpurun_out/r2e_conv_hot.ncu-rep --page raw --csv > gpurun_out/r2e_ncu_conv_hot_raw.csv 2>/dev/null
# norm kernels
python tools/ncu_norm.py > gpurun_out/r2e_norm_events.txt 2>&1 && \
ncu --set full --clock-control none -k regex:"affine_act_fwd|norm_bwd_apply|plane_reduce" -o gpurun_out/r2e_norm MT_REPS=1 python tools/ncu_norm.py > gpurun_out/r2e_norm_ncu.log 2>&1
MT_REPS=1 ncu --set full --clock-control none -k regex:"affine_act_fwd|norm_bwd_apply|plane_reduce" -f -o gpurun_out/r2e_norm python tools/ncu_norm.py > gpurun_out/r2e_norm_ncu.log 2>&1
ncu -i gpurun_out/r2e_norm.ncu-rep --page raw --csv > gpurun_out/r2e_ncu_norm_raw.csv 2>/dev/null
cat gpurun_out/r2e_norm_events.txt
ls -la gpurun_out/r2e_*
