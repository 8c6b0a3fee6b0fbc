mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/r02_pytest_gpu_v2.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02_pytest_gpu_v2.log
tail -4 gpurun_out/r02_pytest_gpu_v2.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke_v2.log 2>&1; echo "smoke exit $?"
timeout 900 python bench.py --steps 8 --warmup 3 > gpurun_out/r02_bench_v2.json 2> gpurun_out/r02_bench_v2.err; echo "bench exit $?"
tail -3 gpurun_out/r02_bench_v2.err
bash tools/sanitize.sh memcheck small fp64 emu surface
