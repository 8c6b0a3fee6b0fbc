# the driver's end-of-round sequence on one B200: GPU tests, smoke, reference arm, our arm
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/r02_final_pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02_final_pytest_gpu.log; tail -3 gpurun_out/r02_final_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_final_smoke.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/r02_final_smoke.log
timeout 900 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/r02_final_bench_ref.json 2> gpurun_out/r02_final_bench_ref.err; echo "ref exit $?"
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r02_final_bench.json 2> gpurun_out/r02_final_bench.err; echo "bench exit $?"
python -c "
import json
r=json.load(open('gpurun_out/r02_final_bench_ref.json')); d=json.load(open('gpurun_out/r02_final_bench.json'))
print('reference', round(r['value'],3), r['cpu_baseline']['cores'], 'cores;', 'ours', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'checksum', d['loss_checksum'], 'sm', d['clocks']['sm_mhz'], d['clocks']['reasons'])
print('same config', r['config']==d['config'], 'launches', d['gpu_launches'], 'roofline frac', round(d['roofline']['frac'],3), 'whole step', round(d['roofline']['whole_step_frac_of_peak'],3), 'cpu_baseline', round(d['cpu_baseline']['value'],2))
a=d['aux']; print('c2', round(a['config_c2']['loss']['value']), 'lso', round(a['len_scale_opt_default']['seconds'],4), 'c4', round(a['c4_experiments']['seconds'],3), 'c5', round(a['c5_surface']['value']), 'predict', round(a['grid_predict']['value']))"
