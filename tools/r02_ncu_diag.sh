# source-level ncu capture of the diagonal-block kernel (C2-sized run of tools/ab_diag.py)
mkdir -p gpurun_out
CMD="python tools/ab_diag.py"
$CMD > gpurun_out/r02_ncu_diag_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_diag_block -s 20 -c 2 -o gpurun_out/r02_prof_diag $CMD > gpurun_out/r02_ncu_diag.log 2>&1
echo "ncu exit $?"; ls -la gpurun_out/*.ncu-rep
