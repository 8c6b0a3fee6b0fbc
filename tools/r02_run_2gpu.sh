mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 8 --warmup 3 > gpurun_out/r02_bench_2gpu.json 2> gpurun_out/r02_bench_2gpu.err; echo "bench 2gpu exit $?"
tail -5 gpurun_out/r02_bench_2gpu.err
python -c "
import json;d=json.load(open('gpurun_out/r02_bench_2gpu.json'));print(d['value'],d['e2e'],d['multirank_identical_to_1rank'],d['checked'],d['loss_checksum']);print(d['aux'])"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --impl reference --gpus 2 --steps 4 --warmup 1 > gpurun_out/r02_bench_2gpu_ref.json 2> gpurun_out/r02_bench_2gpu_ref.err; echo "ref arm exit $?"; cut -c1-300 gpurun_out/r02_bench_2gpu_ref.json
timeout 600 python -m pytest tests/test_gpu_multirank.py -m gpu -q -s -p no:cacheprovider 2>&1 | tail -4
