# ncu evidence of round 2, second call: full captures of the two dominant digit-GEMM roles and of the HBM-side kernels
mkdir -p gpurun_out
CMD2="python bench.py --steps 1 --warmup 1 --no-aux --no-cpu-baseline --no-native --particles 128"
$CMD2 > gpurun_out/r02_ncu_plain3.log 2>&1 &&
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:OzInvuTiles -s 36 -c 2 -o gpurun_out/r02_prof_invu $CMD2 > gpurun_out/r02_ncu_invu.log 2>&1
echo "invu exit $?"
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:OzSyrkTiles -s 44 -c 2 -o gpurun_out/r02_prof_syrk $CMD2 > gpurun_out/r02_ncu_syrk.log 2>&1
echo "syrk exit $?"
ncu --set full --clock-control none --import-source on -k 'regex:k_cov_lower|k_colnorm_alpha|k_wy' -c 6 -o gpurun_out/r02_prof_hbm $CMD2 > gpurun_out/r02_ncu_hbm.log 2>&1
echo "hbm exit $?"
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:OzInvpTiles|OzTrsmTiles' -s 80 -c 2 -o gpurun_out/r02_prof_panel $CMD2 > gpurun_out/r02_ncu_panel.log 2>&1
echo "panel exit $?"
ls -la gpurun_out/*.ncu-rep
