"""Centre Latin-hypercube sampler with the call shape of ``smt.sampling_methods.LHS`` (the one
third-party call of find_len_scales.py:69-70).  Used when SMT is not installed; like SMT it draws
from a private RandomState, unseeded unless `random_state` is given."""
import numpy as np


class LHS:
    def __init__(self, xlimits, criterion="center", random_state=None):
        if criterion not in ("center", "c"):
            raise ValueError("only the 'center' criterion used by the reference is implemented")
        self.xlimits = np.asarray(xlimits, dtype=np.float64)
        if isinstance(random_state, np.random.RandomState):
            self._rng = random_state
        else:
            self._rng = np.random.RandomState(random_state)

    def __call__(self, n):
        lo, hi = self.xlimits[:, 0], self.xlimits[:, 1]
        edges = np.linspace(0, 1, n + 1)
        mid = (edges[:n] + edges[1:n + 1]) / 2
        unit = np.empty((n, self.xlimits.shape[0]))
        for j in range(unit.shape[1]):
            unit[:, j] = self._rng.permutation(mid)
        return lo + unit * (hi - lo)


def get_lhs_class():
    """SMT's LHS if importable (what the reference uses), else the built-in one."""
    try:
        from smt.sampling_methods import LHS as SmtLHS
        return SmtLHS
    except Exception:
        return LHS
