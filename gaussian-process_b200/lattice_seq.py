"""Point-at-a-time form of the hull lattice (convex_hull.py:13-224 restated): the fallback of lattice.py for resolutions
where stepping by `res` does not advance the lattice index by exactly one, and the small-case cross-check of its tests.
Quadratic-ish in the grid size, like the reference."""
import numpy as np
from scipy.spatial import ConvexHull


def snap(value, res):
    """Nearest multiple of res, then rounded to the decimals res is written with when res < 1 (:13-24)."""
    v = round(value / res) * res
    if res < 1:
        v = np.around(v, str(res)[::-1].find('.'))
    return v


def unique_sorted(rows):
    return np.array(sorted(set(tuple(r) for r in rows)))              # :29-35


def raster(p, q, step):
    """Segment p-q sampled along its longest axis in units of step (:38-72)."""
    axis = int(np.argmax(abs((p - q) / step)))
    lo, hi = (p, q) if p[axis] < q[axis] else (q, p)
    delta = hi - lo
    cur = np.array(lo, dtype=float)
    out = [lo]
    while cur[axis] < hi[axis]:
        cur[axis] = snap(cur[axis] + step[axis], step[axis])
        t = (cur[axis] - lo[axis]) / delta[axis]
        for j in range(len(cur)):
            if j != axis:
                cur[j] = snap(t * delta[j] + lo[j], step[j])
        out.append(cur.copy())
    return np.array(out)


def scan_fill(axis, rows, step):
    """Fill between neighbouring rows that differ only along `axis` (:122-150)."""
    rest = [j for j in range(rows.shape[1]) if j != axis]
    out = []
    for a, b in zip(rows[:-1], rows[1:]):
        if (a[rest] == b[rest]).all():
            cur = a.copy()
            out.append(cur.copy())
            while cur[axis] < b[axis]:
                cur[axis] += step[axis]
                for j in range(len(cur)):
                    cur[j] = snap(cur[j], step[j])
                out.append(cur.copy())
            out.append(b.copy())
    out.append(rows[-1])
    return unique_sorted(np.concatenate((np.array(out), rows)))


def facet_boundary(verts, step):
    m = len(verts)
    lines = [raster(verts[0], verts[1], step)]                        # :80-97 (the first edge appears twice in 2-D)
    for i in range(1, m):
        lines.append(raster(verts[i], verts[0] if i == m - 1 else verts[i + 1], step))
    edges = unique_sorted(np.concatenate(lines))
    axis = m - 1
    while max(verts[:, axis]) - min(verts[:, axis]) == 0:             # :163-166
        axis -= 1
    return scan_fill(axis, edges, step)


def fill_convex_hull(points, step):
    points = np.asarray(points, dtype=float)
    step = np.asarray(step, dtype=float)
    hull = ConvexHull(points)
    pts = unique_sorted(np.concatenate([facet_boundary(points[s], step) for s in hull.simplices]))
    for _ in range(pts.shape[1]):                                     # :219-222
        pts = scan_fill(pts.shape[1] - 1, pts, step)
        pts = unique_sorted(np.concatenate((pts[:, 1:], pts[:, :1]), axis=1))
        step = np.concatenate((step[1:], step[:1]))
    return pts
