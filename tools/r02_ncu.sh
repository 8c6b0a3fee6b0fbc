# ncu evidence of round 2 (one GPU call; every ncu run follows a plain run of the same command that exited 0)
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-aux --no-cpu-baseline --no-native"
$CMD > gpurun_out/r02_ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 1300 -c 1500 --csv --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/r02_ncu_list.log 2>&1
echo "launch list exit $?"
CMD2="python bench.py --steps 1 --warmup 1 --no-aux --no-cpu-baseline --no-native --particles 128"
$CMD2 > gpurun_out/r02_ncu_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:OzInvuTiles -s 40 -c 2 -o gpurun_out/r02_prof_invu $CMD2 > gpurun_out/r02_ncu_invu.log 2>&1
echo "invu exit $?"
ncu --set full --clock-control none --import-source on -k regex:OzSyrkTiles -s 40 -c 2 -o gpurun_out/r02_prof_syrk $CMD2 > gpurun_out/r02_ncu_syrk.log 2>&1
echo "syrk exit $?"
ncu --set full --clock-control none --import-source on -k 'regex:k_cov_lower|k_colnorm_alpha|k_wy|k_diag_block|k_oz_slice_mt|OzInvpTiles|OzTrsmTiles' -s 30 -c 14 -o gpurun_out/r02_prof_misc $CMD2 > gpurun_out/r02_ncu_misc.log 2>&1
echo "misc exit $?"
ls -la gpurun_out/*.ncu-rep
