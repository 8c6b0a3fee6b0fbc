mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_memory_hygiene.py tests/test_gpu_large_golden.py -m gpu -q -p no:cacheprovider 2>&1 | tail -15 > gpurun_out/r02_pytest_hygiene.log; tail -8 gpurun_out/r02_pytest_hygiene.log
timeout 900 python bench.py --steps 8 --warmup 3 > gpurun_out/r02_bench_v4.json 2> gpurun_out/r02_bench_v4.err; echo "bench exit $?"
python -c "
import json;d=json.load(open('gpurun_out/r02_bench_v4.json'));print(d['value'],d['e2e']['value'],d['loss_checksum'],d['roofline']['kernel_share_of_step']);print(d['aux']['strong_scaling_slices_1gpu']);print(d['aux']['config_c2'])"
