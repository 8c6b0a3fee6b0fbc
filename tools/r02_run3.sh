mkdir -p gpurun_out
for v in 0 1 2; do
  GPK_DIAG_VARIANT=$v timeout 300 python tools/ab_diag.py > gpurun_out/r02_ab_diag_v$v.json 2> gpurun_out/r02_ab_diag_v$v.err; echo "variant $v exit $?"
  cat gpurun_out/r02_ab_diag_v$v.json
done
python - <<'PY'
import numpy as np
a=np.load('gpurun_out/ab_diag_v0.npz'); b=np.load('gpurun_out/ab_diag_v2.npz')
for n in (640,1024,300):
    r0,r2,s0,s2=a[f'r_{n}'],b[f'r_{n}'],a[f'sd_{n}'],b[f'sd_{n}']
    print(n,'counts equal',np.array_equal(a[f'c_{n}'],b[f'c_{n}']),'resid rel',np.abs(r0-r2).max()/np.abs(r0).max(),'sd rel',np.abs(s2/s0-1).max())
PY
timeout 600 python bench.py --steps 8 --warmup 3 --no-aux --no-cpu-baseline > gpurun_out/r02_bench_v3_noaux.json 2> gpurun_out/r02_bench_v3.err; echo "bench exit $?"
python -c "
import json;d=json.load(open('gpurun_out/r02_bench_v3_noaux.json'));print(d['value'],d['loss_checksum'],d['checked'],d['roofline']['kernel_share_of_step'])"
timeout 900 python -m pytest tests -m gpu -q -x -p no:cacheprovider 2>&1 | tail -5
