"""cProfile of the default-workflow len_scale_opt (N -> 100 points, 40 particles, 500 iterations): where the host time of
a PSO iteration goes (the loss kernel itself is ~130 us of it)."""
import cProfile, contextlib, io, os, pstats, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from gaussian_process_b200 import pso, synthetic
x, y, e = synthetic.problem(1024, 2, 0)
with contextlib.redirect_stdout(io.StringIO()):
    np.random.seed(0); pso.len_scale_opt(x, y, e, False, lhs_random_state=0, max_iter=20)
    np.random.seed(0)
    t0 = time.perf_counter(); pso.len_scale_opt(x, y, e, False, lhs_random_state=0); dt = time.perf_counter() - t0
    pr = cProfile.Profile(); np.random.seed(0); pr.enable(); pso.len_scale_opt(x, y, e, False, lhs_random_state=0); pr.disable()
print(f"len_scale_opt: {dt:.3f} s")
st = io.StringIO(); pstats.Stats(pr, stream=st).sort_stats("cumulative").print_stats(28); print(st.getvalue()[:6000])
