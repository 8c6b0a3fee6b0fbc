mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus 8 --steps 8 --warmup 3 > gpurun_out/r02_bench_8gpu.json 2> gpurun_out/r02_bench_8gpu.err; echo "bench 8gpu exit $?"
tail -5 gpurun_out/r02_bench_8gpu.err
python -c "
import json;d=json.load(open('gpurun_out/r02_bench_8gpu.json'));print(d['value'],d['e2e'],d['multirank_identical_to_1rank'],d['checked'],d['loss_checksum']);print(d['aux'])"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29622 bench.py --gpus 4 --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_4gpu.json 2> gpurun_out/r02_bench_4gpu.err; echo "bench 4gpu exit $?"
python -c "
import json;d=json.load(open('gpurun_out/r02_bench_4gpu.json'));print(d['value'],d['e2e'],d['multirank_identical_to_1rank'],d['loss_checksum']);print(d['aux'])"
