"""Imports the UNMODIFIED Python reference from /root/reference -- TEST INFRASTRUCTURE.

Only usable in the build container (the GPU box has no /root/reference).  Used by
``oracle/make_golden.py`` to generate ``tests/golden/*`` and by the CPU tests that
cross-check the oracle restatement against the live reference.

The reference imports ``smt.sampling_methods.LHS`` at module top (find_len_scales.py:11);
smt is not installable here, so a stub module with a *seedable* centre-LHS is registered
first.  Nothing else is patched; for tracing, ``serial_pool()`` swaps the fork
pool for an in-process one so each swarm evaluation can be recorded."""
import contextlib
import os
import sys
import types
import warnings

import numpy as np

REFERENCE_DIR = os.environ.get("GP_REFERENCE_DIR", "/root/reference")

_lhs_state = {"seed": 0}


def set_lhs_seed(seed):
    _lhs_state["seed"] = seed


class _StubLHS:
    """Stand-in for smt.sampling_methods.LHS(xlimits=..., criterion="center")."""

    def __init__(self, xlimits, criterion="center", random_state=None):
        assert criterion == "center"
        self.xlimits = np.asarray(xlimits, dtype=np.float64)
        self.random_state = random_state

    def __call__(self, n):
        from oracle.gp_oracle import center_lhs
        seed = self.random_state if self.random_state is not None else _lhs_state["seed"]
        return center_lhs(self.xlimits, n, np.random.RandomState(seed))


def install_smt_stub():
    if "smt.sampling_methods" in sys.modules and hasattr(sys.modules["smt.sampling_methods"], "_gp_stub"):
        return
    smt = types.ModuleType("smt")
    sm = types.ModuleType("smt.sampling_methods")
    sm.LHS = _StubLHS
    sm._gp_stub = True
    smt.sampling_methods = sm
    sys.modules["smt"] = smt
    sys.modules["smt.sampling_methods"] = sm


def available():
    return os.path.isfile(os.path.join(REFERENCE_DIR, "GP_func.py"))


def load():
    """Returns a namespace with the reference modules (GP_func, find_len_scales, calc_prob_surf,
    convex_hull, read_in).  The modules are loaded under private names so they never shadow
    the drop-in modules of the same name."""
    import importlib.util
    if not available():
        raise RuntimeError(f"reference not found at {REFERENCE_DIR}")
    install_smt_stub()
    warnings.filterwarnings("ignore", category=DeprecationWarning)
    ns = types.SimpleNamespace()
    saved = {k: sys.modules.get(k) for k in ("GP_func", "find_len_scales", "calc_prob_surf", "convex_hull", "read_in")}
    sys.path.insert(0, REFERENCE_DIR)
    try:
        for name in ("GP_func", "find_len_scales", "calc_prob_surf", "convex_hull", "read_in"):
            sys.modules.pop(name, None)
        for name in ("GP_func", "find_len_scales", "calc_prob_surf", "convex_hull", "read_in"):
            spec = importlib.util.spec_from_file_location(name, os.path.join(REFERENCE_DIR, name + ".py"))
            mod = importlib.util.module_from_spec(spec)
            sys.modules[name] = mod          # find_len_scales does `from GP_func import GP`
            spec.loader.exec_module(mod)
            setattr(ns, name, mod)
    finally:
        sys.path.remove(REFERENCE_DIR)
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    return ns


class _SerialPool:
    def __init__(self, record):
        self._record = record

    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False

    def map(self, fn, iterable):
        args = list(iterable)
        out = [fn(a) for a in args]
        if self._record is not None:
            self._record.append((np.array([a[0] for a in args]), np.array(out)))
        return out


class _SerialCtx:
    def __init__(self, record):
        self._record = record

    def Pool(self, processes=None):
        return _SerialPool(self._record)


@contextlib.contextmanager
def serial_pool(ref_find_len_scales, record=None):
    """Within the block, the reference's `multiprocessing.get_context('fork').Pool` evaluates
    in-process, in order, and appends (positions, scores) of each map() call to `record`."""
    mp = ref_find_len_scales.multiprocessing
    orig = mp.get_context
    mp.get_context = lambda kind=None: _SerialCtx(record)
    try:
        yield
    finally:
        mp.get_context = orig
