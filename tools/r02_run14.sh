mkdir -p gpurun_out
python tools/kmeans_profile.py 2>&1 | grep -E "alone|fit_predict:|N=200000" | tee gpurun_out/r02_kmeans_profile.log
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/r02_pytest_gpu_v14.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02_pytest_gpu_v14.log; tail -3 gpurun_out/r02_pytest_gpu_v14.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke_v14.log 2>&1; echo "smoke exit $?"
timeout 900 python bench.py --steps 8 --warmup 3 > gpurun_out/r02_bench_v14.json 2> gpurun_out/r02_bench_v14.err; echo "bench exit $?"
python -c "
import json;d=json.load(open('gpurun_out/r02_bench_v14.json'));print(d['value'],d['e2e']['value'],d['loss_checksum'],d['clocks']['sm_mhz'],d['cpu_baseline']['value']);a=d['aux'];print(a['config_c2']['loss']['value']);print(a['len_scale_opt_default']['seconds'], a['len_scale_opt_default']['best_length_scales'], a['c4_experiments']['seconds'], a['c4_experiments']['rank0_phase_seconds'], a['c4_experiments']['rank0_len_scale_opt_split_seconds'])"
