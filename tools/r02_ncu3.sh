# ncu launch list with DRAM bytes per launch (the roofline's `traffic`), final round-2 code, default settings
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-aux --no-cpu-baseline --no-native"
$CMD > gpurun_out/r02_ncu_plain4.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 700 -c 1000 --csv --log-file gpurun_out/r02_launches_dram.csv $CMD > gpurun_out/r02_ncu_list2.log 2>&1
echo "launch list exit $?"
python tools/launch_summary.py gpurun_out/r02_launches_dram.csv | head -20
