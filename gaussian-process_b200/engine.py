"""Host-side engine: NumPy in / NumPy out wrappers around libgpk.  PyTorch supplies device
memory, the current CUDA stream and pinned staging only -- every computation is a libgpk kernel.

Reference functions served (file:line in rferguson22/Gaussian-Process):
  cov_batched      kernel_func            GP_func.py:49-65 (+ :21 diagonal)
  LossContext      wass_loss -> counts    find_len_scales.py:154-163
  GPFit / gp       GP                     GP_func.py:12-45
  prob_surface     generate_prob_surf     calc_prob_surf.py:67-81
"""
import os

import numpy as np
import torch

from . import _lib

NB = 128
MAX_D = 8                     # include/gpk.h GPK_MAX_D
MAX_S = 4096                  # include/gpk.h GPK_MAX_S
DEFAULT_CHUNK = 8192          # grid points per predict pass (2048 measured 17 % slower than 4096: too few CTAs per launch)


_cuda_ok = False
_dev_cache = {}


def device():
    global _cuda_ok
    if not _cuda_ok:                                         # checked until it succeeds once (the check itself is ~10 us)
        if not torch.cuda.is_available():
            raise RuntimeError("gaussian-process_b200 needs a CUDA device (B200); there is no CPU fallback")
        _cuda_ok = True
    idx = torch.cuda.current_device()
    dev = _dev_cache.get(idx)
    if dev is None:
        dev = _dev_cache[idx] = torch.device("cuda", idx)
    return dev


def _stream():
    return torch.cuda.current_stream(device().index).cuda_stream      # an int index skips torch's device-resolution path


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def to_dev(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.to(dtype)
    return t.to(device(), non_blocking=False)


# Arithmetic of the two dense contractions of the blocked path (N > 128), include/gpk.h GPK_PREC_*.  It is a per-call argument
# of the C ABI (the library keeps no precision state); what lives here is the caller's *policy*:
#   "fp64"   native FP64 tensor instructions (DMMA)
#   "emu"    FP64 emulated on the int8 tcgen05 tensor cores (7-digit Ozaki split, 34 digit products): same rounding
#            level as an FP64 dot product, 1.7x the loss-evaluation throughput at N = 4096; needs e > 0
#   "emu28"  28 digit products (~2^8 larger truncation error)
#   "emu21"  6 digits, 21 products;  "emu15"  5 digits, 15 products -- for use under the a-posteriori check below
#   "auto"   "emu" when every e_i is finite and > 0 and N >= 1024 (below that the native kernels are faster:
#            profiles/r01_mode_crossover.log), else "fp64"                        (default; env GPK_PRECISION)
_MODES = {"fp64": 0, "emu": 1, "emu28": 2, "emu21": 3, "emu15": 4, "emu28_34": 5, "emu34_28": 6}
MODE_NAMES = {v: k for k, v in _MODES.items()}
EMU_MIN_N = 1024
EMU_MAX_N = 16256             # contraction length below 2^14: the int32 sums of 8 digit products of 2^14 each cannot overflow
_precision_pref = os.environ.get("GPK_PRECISION", "auto")

# Checked mixed precision: every emulated factorisation is probed a posteriori, rho = |W^T W K v - v|_inf / |v|_inf for one
# fixed pseudo-random v (include/gpk.h).  rho estimates the relative error the computed inverse puts on alpha = K^-1 y, i.e. on
# mu - y; a particle (or fit) whose rho exceeds the tolerance is evaluated again by the native FP64 kernels in the same call.
# 1e-10 leaves a factor 10 to the 1e-9 parity bar.  0 switches the check off.                      (env GPK_CHECK_TOL)
_check_tol = float(os.environ.get("GPK_CHECK_TOL", "1e-10"))

# Which algebraic form the loss of the one-CTA path (N <= 128) uses: "reduced" (mu - y = -e^2 alpha, var from diag K^-1) or
# "reference" (GP_func.py:36-40 as written: K_s^T alpha - y, 1 - sum v^2).                         (env GPK_LOSS_FORM)
_FORMS = {"reduced": 0, "reference": 1}
_loss_form = os.environ.get("GPK_LOSS_FORM", "reduced")


def set_precision(mode):
    global _precision_pref
    mode = MODE_NAMES.get(mode, mode)
    if mode != "auto" and mode not in _MODES:
        raise ValueError(f"precision must be one of auto, {', '.join(_MODES)} (got {mode!r})")
    _precision_pref = mode


def get_precision():
    return _precision_pref


def set_check_tol(tol):
    global _check_tol
    _check_tol = float(tol)


def get_check_tol():
    return _check_tol


def set_loss_form(form):
    global _loss_form
    if form not in _FORMS:
        raise ValueError(f"loss form must be one of {', '.join(_FORMS)} (got {form!r})")
    _loss_form = form


def get_loss_form():
    return _loss_form


def _resolve_precision(e_host, n=None):
    """Mode number for a problem with uncertainties e_host (host array) and n known points."""
    pref = _precision_pref
    ok = bool(np.all(np.isfinite(e_host)) and np.all(np.abs(e_host) > 1e-150)) if e_host.size else False
    if pref == "auto":
        return 1 if ok and (n is None or EMU_MIN_N <= n <= EMU_MAX_N) else 0
    if pref not in _MODES:
        raise ValueError(f"GPK_PRECISION={pref!r}: expected auto, {', '.join(_MODES)}")
    if pref != "fp64" and not ok:
        raise ValueError("emulated-FP64 mode needs finite, non-zero e_known (use GPK_PRECISION=fp64)")
    if pref != "fp64" and n is not None and n > EMU_MAX_N:
        raise ValueError(f"emulated-FP64 mode supports N <= {EMU_MAX_N} (exact int32 accumulation); use GPK_PRECISION=fp64")
    return _MODES[pref]


# CUDA-graph replay of the blocked loss evaluation where launch latency shows (128 < N <= 2048); GPK_GRAPH=0 disables
GRAPH_MIN_N, GRAPH_MAX_N = 128, 2048


def _graphs_enabled():
    return os.environ.get("GPK_GRAPH", "1") != "0" and not _lib.timing_active()


_ws_cache = {}


def workspace(nbytes, tag="main"):
    """Grow-only scratch buffer per (device, tag)."""
    key = (torch.cuda.current_device(), tag)
    buf = _ws_cache.get(key)
    if buf is None or buf.numel() < nbytes:
        _ws_cache.pop(key, None)
        buf = None
        torch.cuda.empty_cache()
        buf = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device())
        _ws_cache[key] = buf
    return buf


def release_workspaces():
    _ws_cache.clear()
    torch.cuda.empty_cache()


def free_bytes():
    free, _total = torch.cuda.mem_get_info()
    return int(free)


# ------------------------------------------------------------------------------------ (a)
def cov_batched(x1, x2, ls, e=None):
    """K[p] = kernel_func(x1, x2, ls[p]) (+ diag(e^2) if e is given).  x1 (d,N1), x2 (d,N2), ls (P,d)."""
    lib = _lib.load()
    x1, x2 = _f64(x1), _f64(x2)
    ls = np.atleast_2d(_f64(ls))
    d, n1 = x1.shape
    n2 = x2.shape[1]
    p = ls.shape[0]
    dx1, dx2, dls = to_dev(x1), to_dev(x2), to_dev(ls)
    de = to_dev(_f64(e)) if e is not None else None
    out = torch.empty((p, n1, n2), dtype=torch.float64, device=device())
    st = lib.gpk_cov_batched(dx1.data_ptr(), dx2.data_ptr(), de.data_ptr() if de is not None else None,
                             dls.data_ptr(), p, n1, n2, d, out.data_ptr(), n2, n1 * n2, _stream())
    _lib.check(st, "gpk_cov_batched")
    return out.cpu().numpy()


def potrf_batched(k):
    """Lower Cholesky factors of a batch (P,Np,Np), Np a multiple of 128.  Returns (L, info)."""
    lib = _lib.load()
    k = _f64(k)
    p, n, _ = k.shape
    dk = to_dev(k)
    info = torch.zeros(p, dtype=torch.int32, device=device())
    ws = workspace(lib.gpk_potrf_workspace_bytes(p, n), "potrf")
    st = lib.gpk_potrf_batched(dk.data_ptr(), p, n, info.data_ptr(), ws.data_ptr(), ws.numel(), _stream())
    _lib.check(st, "gpk_potrf_batched")
    return np.tril(dk.cpu().numpy()), info.cpu().numpy()


def _pad_factors(L, n):
    """(P,n,n) host factors -> (P,Np,Np) device batch with identity in the padding (what gpk_potrf_batched leaves)."""
    L = _f64(L)
    p = L.shape[0]
    n_pad = int(_lib.load().gpk_padded_n(n))
    full = np.zeros((p, n_pad, n_pad))
    full[:, :n, :n] = np.tril(L[:, :n, :n])
    idx = np.arange(n, n_pad)
    full[:, idx, idx] = 1.0
    return to_dev(full), n_pad


def fit_alpha_batched(L, y):
    """Staged fit (SURVEY 8b `gpk_fit_alpha_batched`): alpha_p = L_p^-T (L_p^-1 y) for a batch of Cholesky factors
    (P,N,N), GP_func.py:24.  Returns (P,N)."""
    lib = _lib.load()
    y = _f64(y)
    n = y.shape[0]
    dl, n_pad = _pad_factors(L, n)
    p = dl.shape[0]
    alpha = torch.zeros((p, n_pad), dtype=torch.float64, device=device())
    ws = workspace(lib.gpk_staged_workspace_bytes(p, n_pad), "staged")
    dy = to_dev(y)
    st = lib.gpk_fit_alpha_batched(dl.data_ptr(), dy.data_ptr(), p, n, n_pad, alpha.data_ptr(), ws.data_ptr(),
                                   ws.numel(), _stream())
    _lib.check(st, "gpk_fit_alpha_batched")
    return alpha[:, :n].cpu().numpy()


def selfpredict_counts(L, alpha, y, e, sigma_vals, active=None):
    """Staged calibration counts (SURVEY 8b `gpk_selfpredict_counts`) from factors (P,N,N) and alpha (P,N):
    returns (counts (P,S) int32, mu - y (P,N), sd (P,N)), find_len_scales.py:159-163 before the mean."""
    lib = _lib.load()
    y, e, sigma_vals, alpha = _f64(y), _f64(e), _f64(sigma_vals), _f64(alpha)
    n, s = y.shape[0], sigma_vals.shape[0]
    dl, n_pad = _pad_factors(L, n)
    p = dl.shape[0]
    dev = device()
    da = torch.zeros((p, n_pad), dtype=torch.float64, device=dev)
    da[:, :n] = to_dev(alpha)
    dact = to_dev(np.asarray(active, dtype=np.uint8)) if active is not None else None
    counts = torch.empty((p, s), dtype=torch.int32, device=dev)
    resid = torch.empty((p, n), dtype=torch.float64, device=dev)
    sd = torch.empty((p, n), dtype=torch.float64, device=dev)
    ws = workspace(lib.gpk_staged_workspace_bytes(p, n_pad), "staged")
    dy, de, dsig = to_dev(y), to_dev(e), to_dev(sigma_vals)           # named: the buffers must outlive the call
    st = lib.gpk_selfpredict_counts(dl.data_ptr(), da.data_ptr(), dy.data_ptr(), de.data_ptr(),
                                    dsig.data_ptr(), dact.data_ptr() if dact is not None else None, p, n, n_pad,
                                    s, counts.data_ptr(), resid.data_ptr(), sd.data_ptr(), ws.data_ptr(), ws.numel(), _stream())
    _lib.check(st, "gpk_selfpredict_counts")
    return counts.cpu().numpy(), resid.cpu().numpy(), sd.cpu().numpy()


# ------------------------------------------------------------------------------------ (b)
class LossContext:
    """Device-resident training set + sigma grid for repeated swarm evaluations (one per PSO
    iteration).  Only the (P,d) positions go up and the (P,S) int32 counts come down per call.

    After every evaluation `last_fallback` ((P,) uint8 device tensor) marks the particles the a-posteriori check sent to the
    native FP64 kernels and `last_probe` ((P,) float64 device tensor) holds their probe values; `fallbacks` counts them over
    the life of the context when the host-returning calls (counts / counts_view) are used."""

    def __init__(self, x, y, e, sigma_vals, max_particles=None, mem_fraction=0.85):
        self.lib = _lib.load()
        x, y, e, sigma_vals = _f64(x), _f64(y), _f64(e), _f64(sigma_vals)
        self.d, self.n = x.shape
        if self.d > MAX_D:
            raise ValueError(f"x_known has {self.d} dimensions; the kernels support at most {MAX_D} (include/gpk.h GPK_MAX_D)")
        self.s = sigma_vals.shape[0]
        if self.s > MAX_S:
            raise ValueError(f"{self.s} sigma scales; the count kernels support at most {MAX_S} (include/gpk.h GPK_MAX_S)")
        # the count is found by bisection over an ascending grid (the predicate is monotone in s); the reference compares
        # element-wise and accepts any order (find_len_scales.py:161-163): other grids are sorted here, un-permuted on return
        self._sig_unsort = None
        if self.s > 1 and np.any(np.diff(sigma_vals) < 0):
            order = np.argsort(sigma_vals, kind="stable")
            sigma_vals = np.ascontiguousarray(sigma_vals[order])
            unsort = np.empty_like(order)
            unsort[order] = np.arange(self.s)
            self._sig_unsort = to_dev(unsort.astype(np.int64))
        self.x, self.y, self.e, self.sig = to_dev(x), to_dev(y), to_dev(e), to_dev(sigma_vals)
        self.mem_fraction = mem_fraction
        self._e_ok = e
        self._pin_pos = None
        self._dev_pos = None
        self._pin_counts = None
        self._packed = None
        self._mode_key = None
        self._graphs = {}
        self.fallbacks = 0
        self.evaluated = 0
        self.last_fallback = None
        self.last_probe = None

    def _ws(self, p):
        if self._mode_key != _precision_pref:                 # the e-scan is O(N): once per policy, not per call
            self.mode = _resolve_precision(self._e_ok, self.n)
            self._mode_key = _precision_pref
        need = self.lib.gpk_loss_workspace_bytes(p, self.n, self.d, self.mode)
        have = _ws_cache.get((torch.cuda.current_device(), "loss"))
        if have is not None and have.numel() >= need:
            return have
        one = self.lib.gpk_loss_workspace_bytes(1, self.n, self.d, self.mode)
        budget = int((free_bytes() + (have.numel() if have is not None else 0)) * self.mem_fraction)
        if need > budget:                      # libgpk runs the swarm in waves inside this much memory
            need = max(one, budget // one * one)
        return workspace(need, "loss")

    def _form(self):
        return _FORMS[_loss_form] if self.n <= NB else 0

    def _launch(self, positions_dev, active_dev, p, counts, info, resid, sd, probe, fallback, ws, stream=None, mode=None,
                flags=None, tol=None):
        mode = self.mode if mode is None else mode
        st = self.lib.gpk_loss_batched(
            self.x.data_ptr(), self.y.data_ptr(), self.e.data_ptr(), positions_dev.data_ptr(),
            active_dev.data_ptr() if active_dev is not None else None, self.sig.data_ptr(), p, self.n, self.d,
            self.s, mode, self._form() if flags is None else flags,
            (_check_tol if mode != 0 else 0.0) if tol is None else tol,
            counts.data_ptr(), info.data_ptr(), resid.data_ptr() if resid is not None else None,
            sd.data_ptr() if sd is not None else None, probe.data_ptr() if probe is not None else None,
            fallback.data_ptr(), ws.data_ptr(), ws.numel(),
            stream if stream is not None else _stream())
        _lib.check(st, "gpk_loss_batched")

    def counts_device(self, positions_dev, active_dev, p, want_detail=False, stream=None, want_probe=False):
        """All inputs already on the device; returns device tensors (counts, info[, resid, sd]).  In the launch-bound
        range of the blocked path the evaluation is replayed from a CUDA graph; the tensors returned then belong to the
        context and are overwritten by its next evaluation."""
        dev = device()
        ws = self._ws(p)
        if not want_detail and not want_probe and self._sig_unsort is None and GRAPH_MIN_N < self.n <= GRAPH_MAX_N and _graphs_enabled():
            return self._counts_graphed(positions_dev, active_dev, p, ws)
        counts = torch.empty((p, self.s), dtype=torch.int32, device=dev)
        info = torch.empty(p, dtype=torch.int32, device=dev)
        resid = torch.empty((p, self.n), dtype=torch.float64, device=dev) if want_detail else None
        sd = torch.empty((p, self.n), dtype=torch.float64, device=dev) if want_detail else None
        self.last_fallback = torch.empty(p, dtype=torch.uint8, device=dev)
        self.last_probe = torch.empty(p, dtype=torch.float64, device=dev) if want_probe else None
        self._launch(positions_dev, active_dev, p, counts, info, resid, sd, self.last_probe, self.last_fallback, ws, stream)
        if self._sig_unsort is not None:
            counts = counts.index_select(1, self._sig_unsort)
        return (counts, info, resid, sd) if want_detail else (counts, info)

    def _counts_graphed(self, positions_dev, active_dev, p, ws):
        """One swarm evaluation of the blocked path is ~16 launches per block column on up to four streams; for
        N <= 2048 their launch latency is a visible share of the step.  Captured once per (swarm size, mode, check
        tolerance, workspace) into a CUDA graph with static input / output tensors, replayed afterwards."""
        key = (p, self.mode, _check_tol, ws.data_ptr(), ws.numel())
        g = self._graphs.get(key)
        if g is None:
            dev = device()
            st = {"pos": torch.empty((p, self.d), dtype=torch.float64, device=dev),
                  "act": torch.ones(p, dtype=torch.uint8, device=dev),
                  "counts": torch.empty((p, self.s), dtype=torch.int32, device=dev),
                  "info": torch.empty(p, dtype=torch.int32, device=dev),
                  "fb": torch.zeros(p, dtype=torch.uint8, device=dev),
                  "fb2": torch.zeros(p, dtype=torch.uint8, device=dev)}
            st["pos"].copy_(positions_dev)
            cur = torch.cuda.current_stream()
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(cur)
            # the a-posteriori check runs inside the graph in its flags-only form (negative tolerance): the ~6 T launches of
            # an in-call second pass, empty when nothing failed, would be a visible share of these launch-bound sizes
            gtol = -_check_tol if self.mode != 0 else 0.0
            with torch.cuda.stream(side):                  # warm-up outside the capture (attribute calls, streams)
                self._launch(st["pos"], st["act"], p, st["counts"], st["info"], None, None, None, st["fb"], ws, tol=gtol)
            cur.wait_stream(side)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                self._launch(st["pos"], st["act"], p, st["counts"], st["info"], None, None, None, st["fb"], ws, tol=gtol)
            self._graphs.clear()                           # one live graph per context (its workspace is shared)
            g = self._graphs[key] = (graph, st)
        graph, st = g
        st["pos"].copy_(positions_dev)
        if active_dev is not None:
            st["act"].copy_(active_dev)
        else:
            st["act"].fill_(1)
        graph.replay()
        if self.mode != 0 and _check_tol > 0 and bool(st["fb"].any().item()):
            # checked mixed precision: the failed particles again, by the native FP64 kernels, in place
            self._launch(st["pos"], st["fb"], p, st["counts"], st["info"], None, None, None, st["fb2"], ws, mode=0,
                         flags=2, tol=0.0)
        self.last_fallback = st["fb"]
        self.last_probe = None
        return st["counts"], st["info"]

    def counts(self, positions, active=None, want_detail=False, want_probe=False):
        """positions (P,d) host -> counts (P,S) int32 host (+ info, optional resid/sd[, probe])."""
        c, i, extra = self._counts_pinned(positions, active, want_detail, want_probe)
        res = [c.copy(), i.copy()]
        if want_detail:
            res += [extra[0].cpu().numpy(), extra[1].cpu().numpy()]
        if want_probe:
            res.append(self.last_probe.cpu().numpy())
        return tuple(res)

    def counts_view(self, positions, active=None):
        """As counts(), but the arrays returned are views of the context's page-locked staging buffers: valid until
        its next evaluation.  The PSO loop finishes the scores from them at once (one 160 KB copy less per iteration)."""
        c, i, _ = self._counts_pinned(positions, active, False, False)
        return c, i

    def _counts_pinned(self, positions, active, want_detail, want_probe):
        positions = np.atleast_2d(_f64(positions))
        p = positions.shape[0]
        dev = device()
        plain = not want_detail and not want_probe and self._sig_unsort is None
        if plain and not (GRAPH_MIN_N < self.n <= GRAPH_MAX_N and _graphs_enabled()):
            return self._counts_packed(positions, active, p, dev)
        if self._pin_pos is None or self._pin_pos.shape[0] != p or self._dev_pos.device != dev:
            self._pin_pos = torch.empty((p, self.d), dtype=torch.float64).pin_memory()
            self._pin_act = torch.empty(p, dtype=torch.uint8).pin_memory()
            self._pin_counts = torch.empty((p, self.s), dtype=torch.int32).pin_memory()
            self._pin_info = torch.empty(p, dtype=torch.int32).pin_memory()
            self._pin_fb = torch.empty(p, dtype=torch.uint8).pin_memory()
            self._np_pos, self._np_act = self._pin_pos.numpy(), self._pin_act.numpy()
            self._np_counts, self._np_info = self._pin_counts.numpy(), self._pin_info.numpy()
            self._np_fb = self._pin_fb.numpy()
            self._dev_pos = torch.empty((p, self.d), dtype=torch.float64, device=dev)
            self._dev_act = torch.empty(p, dtype=torch.uint8, device=dev)
        stream = torch.cuda.current_stream(dev.index)
        self._np_pos[...] = positions              # (the previous evaluation ended with a synchronize: staging is free)
        self._np_act[...] = 1 if active is None else np.asarray(active, dtype=np.uint8)
        self._dev_pos.copy_(self._pin_pos, non_blocking=True)
        self._dev_act.copy_(self._pin_act, non_blocking=True)
        out = self.counts_device(self._dev_pos, self._dev_act, p, want_detail, stream.cuda_stream, want_probe)
        self._pin_counts.copy_(out[0], non_blocking=True)
        self._pin_info.copy_(out[1], non_blocking=True)
        self._pin_fb.copy_(self.last_fallback, non_blocking=True)
        stream.synchronize()
        self.evaluated += int(self._np_act.sum())
        self.fallbacks += int(self._np_fb.sum())
        return self._np_counts, self._np_info, (out[2:] if want_detail else None)

    def _counts_packed(self, positions, active, p, dev):
        """The PSO loop's evaluation (counts and status only, direct launches): the inputs of a call travel as ONE block
        [positions | active] and its outputs as ONE block [counts | info | fallback flags], both preallocated with the
        context: one copy up, one launch, one copy down, one synchronise per swarm evaluation (was two copies up, three
        allocations and three copies down: 25 us of a 300 us iteration at the reference's PSO size,
        profiles/r02_pso_host_profile.log)."""
        pk = self._packed
        if pk is None or pk["p"] != p or pk["dev"] != dev:
            n_in = p * self.d * 8 + p
            o_info = p * self.s * 4
            o_fb = o_info + p * 4
            n_out = o_fb + p
            pin_in = torch.empty(n_in, dtype=torch.uint8).pin_memory()
            pin_out = torch.empty(n_out, dtype=torch.uint8).pin_memory()
            dev_in = torch.empty(n_in, dtype=torch.uint8, device=dev)
            dev_out = torch.empty(n_out, dtype=torch.uint8, device=dev)
            pk = self._packed = {
                "p": p, "dev": dev, "pin_in": pin_in, "pin_out": pin_out, "dev_in": dev_in, "dev_out": dev_out,
                "np_pos": pin_in[:p * self.d * 8].view(torch.float64).view(p, self.d).numpy(),
                "np_act": pin_in[p * self.d * 8:].numpy(),
                "np_counts": pin_out[:o_info].view(torch.int32).view(p, self.s).numpy(),
                "np_info": pin_out[o_info:o_fb].view(torch.int32).numpy(),
                "np_fb": pin_out[o_fb:].numpy(),
                "d_pos": dev_in.data_ptr(), "d_act": dev_in.data_ptr() + p * self.d * 8,
                "d_counts": dev_out.data_ptr(), "d_info": dev_out.data_ptr() + o_info, "d_fb": dev_out.data_ptr() + o_fb,
                "fb_view": dev_out[o_fb:]}
        stream = torch.cuda.current_stream(dev.index)
        pk["np_pos"][...] = positions              # (the previous evaluation ended with a synchronize: staging is free)
        pk["np_act"][...] = 1 if active is None else np.asarray(active, dtype=np.uint8)
        pk["dev_in"].copy_(pk["pin_in"], non_blocking=True)
        ws = self._ws(p)
        st = self.lib.gpk_loss_batched(
            self.x.data_ptr(), self.y.data_ptr(), self.e.data_ptr(), pk["d_pos"], pk["d_act"], self.sig.data_ptr(), p, self.n,
            self.d, self.s, self.mode, self._form(), _check_tol if self.mode != 0 else 0.0, pk["d_counts"], pk["d_info"],
            None, None, None, pk["d_fb"], ws.data_ptr(), ws.numel(), stream.cuda_stream)
        _lib.check(st, "gpk_loss_batched")
        pk["pin_out"].copy_(pk["dev_out"], non_blocking=True)
        stream.synchronize()
        self.last_fallback, self.last_probe = pk["fb_view"], None
        self.evaluated += int(pk["np_act"].sum())
        self.fallbacks += int(pk["np_fb"].sum())
        return pk["np_counts"], pk["np_info"], None


# ------------------------------------------------------------------------------------ (b)+(c)
class GPFit:
    """Cached fit state (W = L^-1, alpha, scaled points) on the device.  GP_func.py:21-24."""

    def __init__(self, x, y, e, ls):
        self.lib = lib = _lib.load()
        x, y, e, ls = _f64(x), _f64(y), _f64(e), _f64(ls).reshape(-1)
        self.d, self.n = x.shape
        if ls.shape[0] != self.d:
            raise ValueError(f"lengths has {ls.shape[0]} entries, x_known has {self.d} dimensions")
        self.np_ = lib.gpk_padded_n(self.n)
        dev = device()
        dx, dy, de = to_dev(x), to_dev(y), to_dev(e)
        self.ls = to_dev(ls)
        self.W = torch.empty((self.np_, self.np_), dtype=torch.float64, device=dev)
        self.alpha = torch.empty(self.np_, dtype=torch.float64, device=dev)
        self.xs = torch.empty((self.d, self.np_), dtype=torch.float64, device=dev)
        self.sq = torch.empty(self.np_, dtype=torch.float64, device=dev)
        if self.d > MAX_D:
            raise ValueError(f"x_known has {self.d} dimensions; the kernels support at most {MAX_D} (include/gpk.h GPK_MAX_D)")
        status = torch.zeros(2, dtype=torch.float64, device=dev)           # [probe, info] -> one device-to-host read
        info = torch.zeros(1, dtype=torch.int32, device=dev)
        self.mode = _resolve_precision(e, self.n)
        self.fell_back = False

        def run(mode):
            ws = workspace(lib.gpk_fit_workspace_bytes(self.n, self.d, mode), "fit")
            st = lib.gpk_fit(dx.data_ptr(), dy.data_ptr(), de.data_ptr(), self.ls.data_ptr(), self.n, self.d, mode,
                             self.W.data_ptr(), self.alpha.data_ptr(), self.xs.data_ptr(), self.sq.data_ptr(),
                             info.data_ptr(), status.data_ptr(), ws.data_ptr(), ws.numel(), _stream())
            _lib.check(st, "gpk_fit")
            status[1:2].copy_(info)
            rho, bad = status.tolist()
            return rho, int(bad)

        self.probe, bad = run(self.mode)
        if self.mode != 0 and _check_tol > 0 and not (self.probe <= _check_tol):
            # checked mixed precision: the emulated factor failed its a-posteriori residual probe -> native FP64 (DMMA)
            self.mode, self.fell_back = 0, True
            self.probe, bad = run(0)
        if bad != 0:
            # numpy.linalg.cholesky's behaviour at GP_func.py:22
            raise np.linalg.LinAlgError("Matrix is not positive definite")
        self.state = None
        if self.mode != 0 and self.n > NB:
            # digits of W for the int8 tensor-core prediction GEMM (emulated FP64), sliced once per fit
            self.state = torch.empty(lib.gpk_predict_state_bytes(self.n), dtype=torch.uint8, device=dev)
            st = lib.gpk_predict_prepare(self.W.data_ptr(), de.data_ptr(), self.n, self.state.data_ptr(),
                                         self.state.numel(), _stream())
            _lib.check(st, "gpk_predict_prepare")

    def predict_device(self, xfit_dev, m, chunk=DEFAULT_CHUNK):
        dev = device()
        mu = torch.empty(m, dtype=torch.float64, device=dev)
        sd = torch.empty(m, dtype=torch.float64, device=dev)
        chunk = int(max(NB, min(chunk, -(-m // NB) * NB)))
        if self.state is not None:
            ws = workspace(self.lib.gpk_predict_emu_workspace_bytes(self.n, chunk), "predict")
            st = self.lib.gpk_predict_emu(self.state.data_ptr(), self.alpha.data_ptr(), self.xs.data_ptr(),
                                          self.sq.data_ptr(), self.ls.data_ptr(), self.n, self.d, self.mode, xfit_dev.data_ptr(), m,
                                          mu.data_ptr(), sd.data_ptr(), chunk, ws.data_ptr(), ws.numel(), _stream())
            _lib.check(st, "gpk_predict_emu")
            return mu, sd
        ws = workspace(self.lib.gpk_predict_workspace_bytes(self.n, chunk), "predict")
        st = self.lib.gpk_predict(self.W.data_ptr(), self.alpha.data_ptr(), self.xs.data_ptr(), self.sq.data_ptr(),
                                  self.ls.data_ptr(), self.n, self.d, xfit_dev.data_ptr(), m, mu.data_ptr(),
                                  sd.data_ptr(), chunk, ws.data_ptr(), ws.numel(), _stream())
        _lib.check(st, "gpk_predict")
        return mu, sd

    def predict(self, x_fit, chunk=DEFAULT_CHUNK):
        """x_fit (d,M) host -> (mu (M,), sd (M,)) host float64.  GP_func.py:26-45."""
        x_fit = _f64(x_fit)
        if x_fit.shape[0] != self.d:
            raise ValueError(f"x_fit has {x_fit.shape[0]} dimensions, expected {self.d}")
        m = x_fit.shape[1]
        if m == 0:
            return np.zeros(0), np.zeros(0)
        mu, sd = self.predict_device(to_dev(x_fit), m, chunk)
        return mu.cpu().numpy(), sd.cpu().numpy()


def gp(x_known, y_known, e_known, x_fit, lengths, batch_size=10000):
    """GP(x_known, y_known, e_known, x_fit, lengths, batch_size).  `batch_size` only bounded host memory
    in the reference (GP_func.py:28-31); results do not depend on it and it is ignored here."""
    return GPFit(x_known, y_known, e_known, lengths).predict(x_fit)


# ------------------------------------------------------------------------------------ (d)
_pinned = {}


def _pinned_buffer(tag, nbytes):
    """Grow-only page-locked staging buffer (uint8) per tag."""
    buf = _pinned.get(tag)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(int(nbytes), dtype=torch.uint8).pin_memory()
        _pinned[tag] = buf
    return buf


def prob_surface_passes(table, points=100, rows_per_pass=None):
    """Generator over the passes of the surface reduction, in row order: yields (row_index[k] int64 (global),
    y[k,points], p[k,points]) of the emitting rows of one pass.  The arrays are views of page-locked staging buffers and
    stay valid only until the generator is advanced again.  Passes of `rows_per_pass` rows are pipelined over two
    streams: while the consumer works on pass k (copy out, CSV text), pass k+1 uploads, computes and drains."""
    lib = _lib.load()
    table = _f64(table)
    r, ncols = table.shape
    if r == 0 or ncols == 0:
        return
    if rows_per_pass is None:
        rows_per_pass = max(1, min(r, (1 << 27) // (points * 8)))        # 128 MB of y (and of p) per pass
    dev = device()
    streams = [torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)]
    main = torch.cuda.current_stream()
    for st_ in streams:
        st_.wait_stream(main)
    slot_bytes = rows_per_pass * points * 8
    pending = []                        # passes in flight, oldest first: (rows kept, staging views, first row, done event)

    def finished(item):
        k, ys, ps, ids, r0, done = item
        done.synchronize()
        idx = ids[:k].numpy()
        idx += r0
        return idx, ys[:k].numpy(), ps[:k].numpy()

    for i, r0 in enumerate(range(0, r, rows_per_pass)):
        slot = i & 1
        if len(pending) == 2:                                             # this slot's staging buffers are about to be reused
            yield finished(pending.pop(0))
        rc = min(rows_per_pass, r - r0)
        stage_in = _pinned_buffer(("surf_in", slot), rows_per_pass * ncols * 8)[:rc * ncols * 8].view(torch.float64).view(rc, ncols)
        stage_in.copy_(torch.from_numpy(table[r0:r0 + rc]))
        ys = _pinned_buffer(("surf_y", slot), slot_bytes).view(torch.float64)[:rc * points].view(rc, points)
        ps = _pinned_buffer(("surf_p", slot), slot_bytes).view(torch.float64)[:rc * points].view(rc, points)
        ids = _pinned_buffer(("surf_i", slot), rows_per_pass * 8).view(torch.int64)[:rc]
        with torch.cuda.stream(streams[slot]):
            dt = stage_in.to(dev, non_blocking=True)
            y = torch.empty((rc, points), dtype=torch.float64, device=dev)
            p = torch.empty((rc, points), dtype=torch.float64, device=dev)
            valid = torch.empty(rc, dtype=torch.int32, device=dev)
            st = lib.gpk_prob_surface(dt.data_ptr(), rc, ncols, points, y.data_ptr(), p.data_ptr(), valid.data_ptr(),
                                      streams[slot].cuda_stream)
            _lib.check(st, "gpk_prob_surface")
            # rows that do not emit are rare: ship everything when all rows are valid, else compact on the device first
            # (exclusive prefix sum of the valid flags + row gather, libgpk kernels)
            dst = torch.empty(rc, dtype=torch.int64, device=dev)
            cnt = torch.empty(1, dtype=torch.int64, device=dev)
            st = lib.gpk_surface_scan(valid.data_ptr(), rc, dst.data_ptr(), cnt.data_ptr(), streams[slot].cuda_stream)
            _lib.check(st, "gpk_surface_scan")
            k = int(cnt.item())                                           # (synchronises this stream only)
            if k == rc:
                ys.copy_(y, non_blocking=True)
                ps.copy_(p, non_blocking=True)
                keep = torch.arange(rc, dtype=torch.int64, device=dev)
            else:
                yc = torch.empty((max(k, 1), points), dtype=torch.float64, device=dev)
                pc = torch.empty((max(k, 1), points), dtype=torch.float64, device=dev)
                keep = torch.empty(max(k, 1), dtype=torch.int64, device=dev)
                st = lib.gpk_surface_gather(valid.data_ptr(), dst.data_ptr(), rc, points, y.data_ptr(), p.data_ptr(),
                                            yc.data_ptr(), pc.data_ptr(), keep.data_ptr(), streams[slot].cuda_stream)
                _lib.check(st, "gpk_surface_gather")
                ys[:k].copy_(yc[:k], non_blocking=True)
                ps[:k].copy_(pc[:k], non_blocking=True)
            ids[:k].copy_(keep[:k], non_blocking=True)
            done = torch.cuda.Event()
            done.record(streams[slot])
        pending.append((k, ys, ps, ids, r0, done))
    while pending:
        yield finished(pending.pop(0))
    main.wait_stream(streams[0])
    main.wait_stream(streams[1])


def prob_surface(table, points=100, rows_per_pass=None):
    """table (R, 2E) host payload -> (row_index[Rv], y[Rv,points], p[Rv,points]) of the emitting rows (the output is
    1600 B per row, so this path is bound by the host side: PCIe and the copy out of the staging buffers)."""
    table = _f64(table)
    r, ncols = table.shape
    if r == 0 or ncols == 0:
        return np.zeros(0, dtype=np.int64), np.zeros((0, points)), np.zeros((0, points))
    y_out = np.empty((r, points))
    p_out = np.empty((r, points))
    idx_out = np.empty(r, dtype=np.int64)
    y_t, p_t = torch.from_numpy(y_out), torch.from_numpy(p_out)
    n_out = 0
    for idx, ys, ps in prob_surface_passes(table, points, rows_per_pass):
        k = idx.shape[0]
        y_t[n_out:n_out + k].copy_(torch.from_numpy(ys))                  # threaded host copy out of the staging buffer
        p_t[n_out:n_out + k].copy_(torch.from_numpy(ps))
        idx_out[n_out:n_out + k] = idx
        n_out += k
    if n_out == r:
        return idx_out, y_out, p_out
    return idx_out[:n_out].copy(), y_out[:n_out].copy(), p_out[:n_out].copy()


def sum_gaussians(y, gaus, half_dy):
    lib = _lib.load()
    y, gaus = _f64(y), _f64(gaus)
    dy_, dg = to_dev(y), to_dev(gaus)
    out = torch.empty(y.shape[0], dtype=torch.float64, device=device())
    st = lib.gpk_sum_gaussians(dy_.data_ptr(), y.shape[0], dg.data_ptr(), gaus.shape[0], float(half_dy),
                               out.data_ptr(), _stream())
    _lib.check(st, "gpk_sum_gaussians")
    return out.cpu().numpy()
