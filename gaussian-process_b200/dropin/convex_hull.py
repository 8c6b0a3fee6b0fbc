"""Drop-in for the reference's convex_hull module: `fill_convex_hull(points, step)` returns the same
sorted lattice, value for value, as convex_hull.py:203-224, built with array operations (gaussian_process_b200.lattice)
instead of per-point loops and tuple sets.  `round_to_res` and `sort_array` keep their reference meaning."""
import _bootstrap  # noqa: F401
import numpy as np

from gaussian_process_b200 import lattice


def fill_convex_hull(points, step):
    return lattice.fill_convex_hull(points, step)


def round_to_res(value, res):
    """Nearest multiple of `res` (convex_hull.py:13-24)."""
    return lattice._snap(np.float64(value), res)


def sort_array(array):
    """Rows de-duplicated and sorted (convex_hull.py:29-35)."""
    return lattice._sort_unique(array)
