# --set full captures of the round-2 kernels (after a plain run of the same command): halo conv (3x3 64->128) and box-mode wgrad (hot)
python tools/ncu_conv_classes.py > /dev/null 2>&1 || exit 1
ncu --profile-from-start off --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"igemm_halo_kernel<.int.2, .int.128, .int.3, .int.3, .int.1>" -c 1 -f -o gpurun_out/r2_halo python tools/ncu_conv_classes.py > gpurun_out/r2_halo_ncu.log 2>&1
ncu -i gpurun_out/r2_halo.ncu-rep --page raw --csv > gpurun_out/r2_ncu_halo_raw.csv 2>/dev/null
ncu --profile-from-start off --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"igemm_wgrad_kernel<.int.2, .int.128, .int.3, .bool.1" -c 1 -f -o gpurun_out/r2_wgbox python tools/ncu_conv_classes.py > gpurun_out/r2_wgbox_ncu.log 2>&1
ncu -i gpurun_out/r2_wgbox.ncu-rep --page raw --csv > gpurun_out/r2_ncu_wgrad_box_raw.csv 2>/dev/null
ls -la gpurun_out/r2_ncu_halo_raw.csv gpurun_out/r2_ncu_wgrad_box_raw.csv
rm -f gpurun_out/r2_halo.ncu-rep gpurun_out/r2_wgbox.ncu-rep
