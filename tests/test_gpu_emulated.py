"""Emulated-FP64 primitives (int8 tcgen05 digit GEMM, include/gpk.h gpk_oz_*) through the C ABI:
digit records bit-exact against a NumPy restatement of the split, int32 accumulators exact against NumPy
integer sums, FP64 result against a long-double reference.  Run on the B200 box: pytest -m gpu."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

NS, DIGIT = 7, 4096
CHUNK = NS * DIGIT
BIAS = 0x0000808080808080


def host_digits(x, rowinv):
    """(R, K) float64 -> (R, K, 7) int8 balanced digits: bytes of rint(u 2^55) + 0x..8080, low six re-centred."""
    v = np.rint(x * rowinv[:, None] * 2.0 ** 55).astype(np.int64)
    w = (v + BIAS) ^ BIAS
    return np.stack([((w >> (8 * (6 - s))) & 0xFF).astype(np.uint8).view(np.int8) for s in range(NS)], axis=-1)


def record_offsets(rows, k):
    """Byte offset of digit 0 of element (r, c) in the record layout [rb][kc][digit][rg16][khalf][8][16]."""
    r = np.arange(rows)[:, None]
    c = np.arange(k)[None, :]
    kch = k // 32
    return (((r // 128) * kch + c // 32) * CHUNK + ((r % 128) // 8) * 256 + ((c % 32) // 16) * 128 + (r % 8) * 16
            + (c % 16))


def scales(x):
    mx = np.abs(x).max(axis=1)
    sig = 2.0 ** (np.floor(np.log2(mx)) + 1)
    sig = np.where(mx > 0.98 * sig, 2 * sig, sig)
    return sig


@pytest.fixture(scope="module")
def env():
    import torch
    from gaussian_process_b200 import _lib, engine
    return torch, _lib.load(), _lib, engine


def _operands(m, n, k, seed):
    rng = np.random.default_rng(seed)
    a = rng.uniform(-1, 1, (m, k)) * 2.0 ** rng.integers(-4, 5, (m, 1))
    b = rng.uniform(-1, 1, (n, k)) * 2.0 ** rng.integers(-4, 5, (n, 1))
    a[:, ::7] *= 1e-6                      # wide dynamic range inside a row
    b[:, 3::11] *= 1e-9
    return a, b


@pytest.mark.parametrize("groups", [5, 6, 7, 8])
def test_digit_gemm_exact(env, groups):
    torch, lib, _lib, engine = env
    m, n, k = 256, 128, 384
    a, b = _operands(m, n, k, 3)
    sa, sb = scales(a), scales(b)
    da, db = engine.to_dev(a), engine.to_dev(b)
    dia, dib = engine.to_dev(1.0 / sa), engine.to_dev(1.0 / sb)
    dsa, dsb = engine.to_dev(sa), engine.to_dev(sb)
    ra = torch.empty(lib.gpk_oz_slice_bytes(m, k), dtype=torch.uint8, device="cuda")
    rb = torch.empty(lib.gpk_oz_slice_bytes(n, k), dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    _lib.check(lib.gpk_oz_slice(da.data_ptr(), k, m, k, dia.data_ptr(), ra.data_ptr(), st), "gpk_oz_slice")
    _lib.check(lib.gpk_oz_slice(db.data_ptr(), k, n, k, dib.data_ptr(), rb.data_ptr(), st), "gpk_oz_slice")
    # 1. digit records, byte for byte
    dig_a, dig_b = host_digits(a, 1.0 / sa), host_digits(b, 1.0 / sb)
    for rec, dig, rows in ((ra, dig_a, m), (rb, dig_b, n)):
        h = rec.cpu().numpy().view(np.int8)
        off = record_offsets(rows, k)
        for s in range(NS):
            np.testing.assert_array_equal(h[off + s * DIGIT], dig[:, :, s])
    # the split is error free to 2^-56 of the row scale
    u = sum(dig_a[:, :, s].astype(np.longdouble) * np.longdouble(2.0) ** (-7 - 8 * s) for s in range(NS))
    assert np.max(np.abs(u * sa[:, None] - a) / sa[:, None]) <= 2.0 ** -56
    # 2. GEMM: raw int32 accumulators and the FP64 result
    c = torch.full((m, n), float("nan"), dtype=torch.float64, device="cuda")
    tiles = (m // 128) * (n // 64)
    dbg = torch.zeros((tiles, groups, 128, 64), dtype=torch.int32, device="cuda")
    _lib.check(lib.gpk_oz_gemm_tn(ra.data_ptr(), rb.data_ptr(), m, n, k, dsa.data_ptr(), dsb.data_ptr(), c.data_ptr(),
                                  n, dbg.data_ptr(), groups, st), "gpk_oz_gemm_tn")
    torch.cuda.synchronize()
    acc = np.zeros((groups, m, n), dtype=np.int64)
    for s in range(NS):
        for t in range(NS):
            if s + t < groups:
                acc[s + t] += dig_a[:, :, s].astype(np.int64) @ dig_b[:, :, t].astype(np.int64).T
    got = dbg.cpu().numpy().reshape(m // 128, n // 64, groups, 128, 64).transpose(2, 0, 3, 1, 4).reshape(groups, m, n)
    np.testing.assert_array_equal(got, acc)                                   # integer work: exact
    ref = (a.astype(np.longdouble) @ b.astype(np.longdouble).T).astype(np.float64)
    err = np.abs(c.cpu().numpy() - ref) / (sa[:, None] * sb[None, :])
    # groups = 8: FP64 rounding of the result only; fewer groups: + the dropped digit products, each of weight
    # <= 2^(-8 g) of the scale product per contraction index (worst case, k = 384)
    assert err.max() < {8: 4e-15, 7: 4e-14, 6: 1e-11, 5: 3e-9}[groups]


def test_emulated_loss_matches_native_and_oracle(env):
    """Whole blocked path, both modes, N = 640 (5 block rows): counts equal, mean/sd at the FP64 rounding level."""
    torch, lib, _lib, engine = env
    from oracle import gp_oracle as orc
    x, y, e = orc.synthetic_problem(640, 2, 11)
    pos, lo, hi = orc.synthetic_particles(x, 6, 11)
    s, ex = orc.sigma_grid()
    out = {}
    for mode in ("fp64", "emu", "emu28"):
        engine.set_precision(mode)
        out[mode] = engine.LossContext(x, y, e, s).counts(pos, want_detail=True)
        assert not out[mode][1].any()
    engine.set_precision("auto")
    mu_r, sd_r = orc.gp(x, y, e, x, pos[0])
    for mode in ("fp64", "emu"):
        np.testing.assert_allclose(out[mode][2][0] + y, mu_r, rtol=1e-9, atol=1e-12)
        np.testing.assert_allclose(out[mode][3][0], sd_r, rtol=1e-9)
    np.testing.assert_array_equal(out["emu"][0], out["fp64"][0])
    np.testing.assert_allclose(out["emu28"][3], out["fp64"][3], rtol=1e-7)


def test_emulated_mode_rejects_zero_uncertainty(env):
    torch, lib, _lib, engine = env
    x = np.random.default_rng(0).random((2, 200))
    y, e = np.zeros(200), np.zeros(200)
    s = np.linspace(0.001, 3, 1000)
    engine.set_precision("emu")
    try:
        with pytest.raises(ValueError):
            engine.LossContext(x, y, e, s).counts(np.array([[0.3, 0.3]]))
    finally:
        engine.set_precision("auto")
    assert engine._resolve_precision(e) == 0          # auto falls back to the native FP64 kernels


@pytest.mark.parametrize("mode", ["fp64", "emu"])
def test_graph_replay_equals_direct_launches(env, mode, monkeypatch):
    """128 < N <= 2048 replays the evaluation from a CUDA graph: same counts as direct launches, for changing
    positions, active masks and swarm sizes on one context."""
    torch, lib, _lib, engine = env
    from oracle import gp_oracle as orc
    x, y, e = orc.synthetic_problem(300, 2, 5)
    s, ex = orc.sigma_grid()
    engine.set_precision(mode)
    try:
        ctx_g = engine.LossContext(x, y, e, s)
        ctx_d = engine.LossContext(x, y, e, s)
        for seed, p in ((1, 6), (2, 6), (3, 9), (4, 6)):
            pos, lo, hi = orc.synthetic_particles(x, p, seed)
            act = (np.arange(p) % 3 != seed % 3).astype(np.uint8)
            monkeypatch.setenv("GPK_GRAPH", "1")
            cg, ig = ctx_g.counts(pos, active=act)
            assert len(ctx_g._graphs) == 1
            monkeypatch.setenv("GPK_GRAPH", "0")
            cd, idd = ctx_d.counts(pos, active=act)
            live = act.astype(bool)
            np.testing.assert_array_equal(cg[live], cd[live])
            np.testing.assert_array_equal(ig[live], idd[live])
        assert len(ctx_d._graphs) == 0
    finally:
        engine.set_precision("auto")


def test_emulated_mode_size_guard(env):
    """Exact int32 accumulation needs contraction lengths below 2^14: N <= 16256 for every emulated mode."""
    torch, lib, _lib, engine = env
    prev = engine.get_precision()
    engine.set_precision("emu")
    try:
        with pytest.raises(ValueError):
            engine._resolve_precision(np.full(4, 0.1), 20000)
    finally:
        engine.set_precision(prev)
    assert engine._resolve_precision(np.full(4, 0.1), 20000) in (0, 1) and (prev != "auto" or engine._resolve_precision(np.full(4, 0.1), 20000) == 0)
    assert lib.gpk_loss_workspace_bytes(1, 20000, 2, 0) > 0
    dummy = torch.zeros(64, dtype=torch.float64, device="cuda")
    st = lib.gpk_loss_batched(dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(), None, dummy.data_ptr(),
                              1, 20000, 2, 10, 1, 0, 0.0, dummy.data_ptr(), dummy.data_ptr(), None, None, None, None,
                              dummy.data_ptr(), 64, None)
    assert st == 2 and b"16256" in lib.gpk_last_error_string()
