mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29631 bench.py --impl reference --gpus 8 --steps 8 --warmup 3 > gpurun_out/r02_bench_8gpu_v2_ref.json 2> gpurun_out/r02_bench_8gpu_v2_ref.err; echo "ref exit $?"; cut -c1-140 gpurun_out/r02_bench_8gpu_v2_ref.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29632 bench.py --gpus 8 --steps 8 --warmup 3 > gpurun_out/r02_bench_8gpu_v2.json 2> gpurun_out/r02_bench_8gpu_v2.err; echo "bench 8gpu exit $?"
python -c "
import json;d=json.load(open('gpurun_out/r02_bench_8gpu_v2.json'));print(d['value'],d['e2e']['value'],d['multirank_identical_to_1rank'],d['loss_checksum'],d['cpu_baseline']['value']);a=d['aux'];print(a['c4_experiments']['seconds'],a['c5_surface']['value'],a['weak']['value'])"
