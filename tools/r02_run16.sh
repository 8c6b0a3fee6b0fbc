mkdir -p gpurun_out
GPK_DIAG_VARIANT=3 timeout 300 python tools/ab_diag.py > gpurun_out/r02_ab_diag_hoist_v3.json 2> gpurun_out/r02_ab_diag_hoist_v3.err; echo "exit $?"
python -c "
import json;d=json.load(open('gpurun_out/r02_ab_diag_hoist_v3.json'));print(d['variant'],d['info'],d['digest'][:16],'c2 ms',round(d['c2_ms_per_swarm'],3),'diag avg us',round(d['c2_diag_us_per_launch'],1))"
bash tools/r02_small_prof.sh 2>&1 | grep -E "^N=100|^N=21" -A3 | grep -E "^N=|chol_inv"
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_device_pso.py -m gpu -x -q -p no:cacheprovider 2>&1 | tail -2
python tools/pso_host_profile.py 2>&1 | grep -E "per swarm" | head -2
