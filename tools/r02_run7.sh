mkdir -p gpurun_out
# the new test alone in a fresh process, then the whole GPU suite (it is collected last)
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/r02_pytest_no_alloc_alone.log 2>&1; echo "alone exit $?"; tail -2 gpurun_out/r02_pytest_no_alloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q -s -p no:cacheprovider > gpurun_out/r02_pytest_gpu_v7.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02_pytest_gpu_v7.log
tail -3 gpurun_out/r02_pytest_gpu_v7.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke_v7.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/r02_smoke_v7.log
timeout 900 python bench.py --steps 8 --warmup 3 > gpurun_out/r02_bench_v7.json 2> gpurun_out/r02_bench_v7.err; echo "bench exit $?"
python -c "
import json;d=json.load(open('gpurun_out/r02_bench_v7.json'));print(d['value'],d['e2e']['value'],d['loss_checksum'],d['clocks']);print(d['roofline']['traffic'],d['roofline']['algorithmic_bytes_per_launch'],d['roofline']['frac'],d['roofline']['whole_step_frac_of_peak'])"
timeout 600 python bench.py --impl reference --steps 8 --warmup 3 > gpurun_out/r02_bench_v7_ref.json 2> gpurun_out/r02_bench_v7_ref.err; echo "ref exit $?"; cut -c1-160 gpurun_out/r02_bench_v7_ref.json
