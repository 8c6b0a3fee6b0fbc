mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/r02_gpuinfo.txt 2>&1
nproc >> gpurun_out/r02_gpuinfo.txt
timeout 1500 python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/r02_pytest_gpu_v1.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02_pytest_gpu_v1.log
tail -5 gpurun_out/r02_pytest_gpu_v1.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke_v1.log 2>&1; echo "smoke exit $?"
timeout 900 python bench.py --steps 8 --warmup 3 > gpurun_out/r02_bench_v1.json 2> gpurun_out/r02_bench_v1.err; echo "bench exit $?"
timeout 300 python tools/pso_traj_report.py > gpurun_out/r02_pso_trajectories.log 2>&1; echo "traj exit $?"
tail -3 gpurun_out/r02_bench_v1.err
