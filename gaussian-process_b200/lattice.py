"""Prediction lattice inside the convex hull of the measured points: the same point set, in the same
order and with the same float64 values, as the reference's convex_hull.fill_convex_hull (convex_hull.py:203-224), built
with array operations instead of per-point Python loops and tuple sets.

What the reference does, restated: qhull facets -> every facet's edges rasterised at `step` along their longest axis
(:38-72) -> each facet scan-filled along its last non-degenerate axis (:122-150, :154-171) -> union, sorted (:29-35) ->
d times: scan-fill along the last axis, rotate the axes by one (:188-199).  Every generated coordinate goes through
`round_to_res` (:13-24): index k = round-half-even(value / res), value = k * res, then (res < 1) rounded to the number
of decimals res is written with.  Segment end points keep their raw values.

Each fill only ever adds lattice indices k0, k0 + 1, ... up to the first one whose value reaches the segment's end, so a
fill is: find the qualifying neighbour pairs, compute (k0, k_last) per pair, expand with repeat / arange.  The expansion
assumes that stepping by `res` from lattice value k lands on k + 1 (it does unless res is pathological); that is checked
on every generated point and the sequential restatement is used if it ever fails."""
import numpy as np
from scipy.spatial import ConvexHull


def _decimals(res):
    return str(res)[::-1].find('.')                       # convex_hull.py:22 (digits after the point in str(res))


def _value(k, res):
    """Lattice value of index k: second half of round_to_res (convex_hull.py:19-22)."""
    v = k * res
    if res < 1:
        v = np.around(v, _decimals(res))
    return v


def _index(value, res):
    return np.rint(value / res)                           # Python round() of a float64: half to even


def _snap(value, res):
    return _value(_index(value, res), res)


def _sort_unique(a):
    """sorted(set(map(tuple, a))) (convex_hull.py:29-35): rows de-duplicated, lexicographic order."""
    a = np.asarray(a, dtype=np.float64)
    if a.ndim != 2 or a.shape[0] < 2:
        return np.unique(a, axis=0)
    # stable lexicographic sort + neighbour comparison: the rows np.unique(a, axis=0) returns, without its
    # structured-dtype sort (4x faster at 4e4 rows); of rows that compare equal the first one inserted is kept,
    # as a Python set does
    s = a[np.lexsort(a.T[::-1])]
    keep = np.ones(s.shape[0], dtype=bool)
    keep[1:] = np.any(s[1:] != s[:-1], axis=1)
    return np.ascontiguousarray(s[keep])


def _run_lengths(k0, cur0, end, res):
    """Number of fill steps of each segment: indices k0, k0+1, ... are appended until the value reaches `end`
    (the loop tests the previous value, starting with the raw start `cur0`)."""
    n = np.zeros(len(k0), dtype=np.int64)
    go = cur0 < end
    if not go.any():
        return n
    kl = np.maximum(k0, np.ceil(end / res) - 1.0)
    for _ in range(4):                                    # settle on the first k >= k0 with value(k) >= end
        low = go & (_value(kl, res) < end)
        kl = np.where(low, kl + 1.0, kl)
    for _ in range(4):
        high = go & (kl > k0) & (_value(kl - 1.0, res) >= end)
        kl = np.where(high, kl - 1.0, kl)
    ok = (_value(kl, res) >= end) & ((kl == k0) | (_value(kl - 1.0, res) < end))
    if not ok[go].all():
        raise _Fallback
    n[go] = (kl[go] - k0[go]).astype(np.int64) + 1
    return n


class _Fallback(Exception):
    pass


def _expand(k0, n):
    """Concatenation of arange(k0[i], k0[i] + n[i]) and the segment number of every element."""
    total = int(n.sum())
    seg = np.repeat(np.arange(len(n)), n)
    first = np.cumsum(n) - n
    k = k0[seg] + (np.arange(total) - first[seg])
    return k, seg


def _check_unit_steps(k, res):
    if len(k) and not np.array_equal(_index(_value(k, res) + res, res), k + 1.0):
        raise _Fallback


def _straight_line(p1, p2, step):
    """Points of the rasterised segment p1-p2 (convex_hull.py:38-72), raw start first."""
    d = int(np.argmax(np.abs((p1 - p2) / step)))
    start, end = (p1, p2) if p1[d] < p2[d] else (p2, p1)
    diff = end - start
    res = step[d]
    k0 = _index(start[d] + res, res)
    n = _run_lengths(np.array([k0]), np.array([start[d]]), np.array([end[d]]), res)[0]
    if n == 0:
        return start[None, :].copy()
    k = k0 + np.arange(n, dtype=np.float64)
    _check_unit_steps(k, res)
    line = np.empty((n + 1, len(start)))
    line[0] = start
    line[1:, d] = _value(k, res)
    t = (line[1:, d] - start[d]) / diff[d]
    for i in range(len(start)):
        if i != d:
            line[1:, i] = _snap(t * diff[i] + start[i], step[i])
    return line


def _boundary_edges(vertices, step):
    m = len(vertices)
    lines = [_straight_line(vertices[0], vertices[1], step)]
    for i in range(1, m):
        lines.append(_straight_line(vertices[i], vertices[0] if i == m - 1 else vertices[i + 1], step))
    return _sort_unique(np.concatenate(lines))


def _boundary_fill(inc_dim, edges, step):
    """Scan fill along inc_dim between neighbouring rows that agree in every other coordinate (convex_hull.py:122-150)."""
    nd = edges.shape[1]
    if len(edges) < 2:
        return _sort_unique(edges)
    others = [j for j in range(nd) if j != inc_dim]
    a, b = edges[:-1], edges[1:]
    pair = np.all(a[:, others] == b[:, others], axis=1) if others else np.ones(len(a), dtype=bool)
    s, e = a[pair], b[pair]
    res = step[inc_dim]
    k0 = _index(s[:, inc_dim] + res, res)
    n = _run_lengths(k0, s[:, inc_dim], e[:, inc_dim], res)
    k, seg = _expand(k0, n)
    _check_unit_steps(k, res)
    new = np.empty((len(k), nd))
    new[:, inc_dim] = _value(k, res)
    for j in others:
        new[:, j] = _snap(s[:, j], step[j])[seg]         # every coordinate is re-snapped after the first step (:138-139)
    return _sort_unique(np.concatenate((new, edges)))


def _hull_boundary(vertices, step):
    edges = _boundary_edges(vertices, step)
    inc_dim = len(vertices) - 1
    while np.max(vertices[:, inc_dim]) - np.min(vertices[:, inc_dim]) == 0:
        inc_dim -= 1
    return _boundary_fill(inc_dim, edges, step)


def _fill(points, step):
    points = np.asarray(points, dtype=np.float64)
    hull = ConvexHull(points)
    step = np.asarray(step, dtype=np.float64)
    parts = [_hull_boundary(points[simplex], step) for simplex in hull.simplices]
    pts = _sort_unique(np.concatenate(parts))
    for _ in range(pts.shape[1]):
        pts = _boundary_fill(pts.shape[1] - 1, pts, step)
        pts = _sort_unique(np.concatenate((pts[:, 1:], pts[:, :1]), axis=1))      # rotate the axes (:194-196)
        step = np.concatenate((step[1:], step[:1]))
    return pts


def fill_convex_hull(points, step):
    """points (n, d), step (d,) -> (m, d) lattice points of the filled hull, sorted (convex_hull.py:203-224)."""
    try:
        return _fill(points, step)
    except _Fallback:
        from . import lattice_seq
        return lattice_seq.fill_convex_hull(points, step)
