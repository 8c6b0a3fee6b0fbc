"""Where the MMA thread of the digit GEMM waits (instrumented build, tools/build_prof.sh):
    GPK_LIB=tools/libgpk_prof.so python tools/oz_stall.py [particles] [groups]
Per kernel kind (SUB = SYRK, STORE = TRSM / phase-2 panel / generic, STORE_NEG = inverse update): share of the MMA
thread's loop spent waiting for operands (full barriers), for the epilogue to drain TMEM (freebar), for the tile queue,
and issuing; average cycles per tile and per K chunk."""
import ctypes as C
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import bench
from gaussian_process_b200 import engine, pso, _lib

P = int(sys.argv[1]) if len(sys.argv) > 1 else 256
G = int(sys.argv[2]) if len(sys.argv) > 2 else 4
lib = _lib.load()
assert hasattr(lib, "gpk_debug_oz_prof"), "not the instrumented build: set GPK_LIB=tools/libgpk_prof.so"
lib.gpk_debug_oz_prof.restype = C.c_int
lib.gpk_debug_oz_prof.argtypes = [C.POINTER(C.c_ulonglong), C.c_int]
_lib.check(lib.gpk_set_groups(G, 1), "gpk_set_groups")
x, y, e = bench.make_problem(0)
pos, lo, hi = bench.make_particles(x, P, 0)
s, ex = pso.sigma_grid()
ctx = engine.LossContext(x, y, e, s)
dpos = engine.to_dev(pos); dact = torch.ones(P, dtype=torch.uint8, device="cuda")
for _ in range(2):
    ctx.counts_device(dpos, dact, P)
buf = (C.c_ulonglong * 32)()
assert lib.gpk_debug_oz_prof(buf, 1) == 0
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
ctx.counts_device(dpos, dact, P)
e1.record()
torch.cuda.synchronize()
assert lib.gpk_debug_oz_prof(buf, 0) == 0
v = np.array(list(buf), dtype=np.float64).reshape(4, 8)
print(f"{P} particles, {G} groups: {e0.elapsed_time(e1):.1f} ms per swarm evaluation (instrumented)")
names = {0: "SUB (SYRK)", 1: "STORE (TRSM, phase-2 panel)", 2: "STORE_NEG (inverse update)", 3: "SUMSQ (predict)"}
for k in range(4):
    tot, full, free, q, issue, tiles, chunks, ctas = v[k]
    if tot == 0:
        continue
    ideal = chunks * 1088.0
    print(f"{names[k]:30s}: CTAs {int(ctas)}, tiles {int(tiles)}, chunks/tile {chunks / tiles:.1f}; loop cycles per tile {tot / tiles:.0f}"
          f" (ideal MMA {ideal / tiles:.0f} = {100 * ideal / tot:.1f} %); wait operands {100 * full / tot:.1f} %, wait epilogue {100 * free / tot:.1f} %,"
          f" wait queue+make {100 * q / tot:.1f} %, issue {100 * issue / tot:.1f} %; per tile: operands {full / tiles:.0f}, epilogue {free / tiles:.0f}, queue {q / tiles:.0f} clk")
