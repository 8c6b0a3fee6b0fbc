mkdir -p gpurun_out
for v in 0 3 4; do
  GPK_DIAG_VARIANT=$v timeout 300 python tools/ab_diag.py > gpurun_out/r02_ab_diag_la_v$v.json 2> gpurun_out/r02_ab_diag_la_v$v.err; echo "variant $v exit $?"
  python -c "
import json;d=json.load(open('gpurun_out/r02_ab_diag_la_v$v.json'));print(d['variant'],d['info'],d['digest'][:16],'c2 ms',round(d['c2_ms_per_swarm'],3),'diag avg us',round(d['c2_diag_us_per_launch'],1))"
done
python - <<'PY'
import numpy as np
a=np.load('gpurun_out/ab_diag_v0.npz')
for v in (3,4):
    b=np.load(f'gpurun_out/ab_diag_v{v}.npz')
    print(v, {k: bool(np.array_equal(a[k], b[k])) for k in a.files})
PY
echo "== small kernel, look-ahead schedule"
GPK_SMALL_VARIANT=1 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_workflow.py tests/test_gpu_multirank.py tests/test_gpu_memory_hygiene.py -m gpu -x -q -p no:cacheprovider 2>&1 | tail -2
sed -i 's|^GPK_LIB=\$PWD/gpurun_out/libgpk_prof.so python tools/small_profile.py \(.*\)$|GPK_LIB=$PWD/gpurun_out/libgpk_prof.so python tools/small_profile.py \1|' tools/r02_small_prof.sh
GPK_SMALL_VARIANT=1 bash tools/r02_small_prof.sh 2>&1 | grep -E "^N=|chol_inv" | tee gpurun_out/r02_small_profile_la.log
GPK_SMALL_VARIANT=1 python tools/pso_host_profile.py 2>&1 | grep -E "per swarm" 
