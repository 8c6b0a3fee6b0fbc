mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/r02_pytest_gpu_v5.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02_pytest_gpu_v5.log
tail -4 gpurun_out/r02_pytest_gpu_v5.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke_v5.log 2>&1; echo "smoke exit $?"
timeout 900 python bench.py --steps 8 --warmup 3 > gpurun_out/r02_bench_v5.json 2> gpurun_out/r02_bench_v5.err; echo "bench exit $?"
python -c "
import json;d=json.load(open('gpurun_out/r02_bench_v5.json'));print(d['value'],d['e2e']['value'],d['loss_checksum'],d['roofline']['kernel_share_of_step']);print(d['aux']['strong_scaling_slices_1gpu']);print(d['aux']['config_c2']);print(d['aux']['c4_experiments'])"
GPK_DIAG_VARIANT=0 timeout 300 python tools/ab_diag.py
