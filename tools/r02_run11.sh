mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/r02_pytest_gpu_v11.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02_pytest_gpu_v11.log; tail -3 gpurun_out/r02_pytest_gpu_v11.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke_v11.log 2>&1; echo "smoke exit $?"
timeout 900 python bench.py --steps 8 --warmup 3 > gpurun_out/r02_bench_v11.json 2> gpurun_out/r02_bench_v11.err; echo "bench exit $?"
python -c "
import json;d=json.load(open('gpurun_out/r02_bench_v11.json'));print(d['value'],d['e2e']['value'],d['loss_checksum'],d['clocks']['sm_mhz']);print(d['roofline']['kernel_share_of_step']);a=d['aux'];print(a['strong_scaling_slices_1gpu']);print(a['config_c2']['loss']);print(a['len_scale_opt_default']['seconds'], a['c4_experiments']['seconds'], a['c4_experiments']['rank0_phase_seconds'])"
