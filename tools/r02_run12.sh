mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_device_pso.py -m gpu -x -q -p no:cacheprovider 2>&1 | tail -3
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29641 bench.py --gpus 2 --steps 8 --warmup 3 > gpurun_out/r02_bench_2gpu_v2.json 2> gpurun_out/r02_bench_2gpu_v2.err; echo "bench 2gpu exit $?"
python -c "
import json;d=json.load(open('gpurun_out/r02_bench_2gpu_v2.json'));print(d['value'],d['e2e']['value'],d['multirank_identical_to_1rank'],d['loss_checksum'],d['cpu_baseline']['value']);a=d['aux'];print(a['c4_experiments']['seconds'],a['c4_experiments']['rank0_phase_seconds'],a['c5_surface']['value'],a['weak']['value'])"
timeout 600 python -m pytest tests/test_gpu_multirank.py tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider 2>&1 | tail -2
