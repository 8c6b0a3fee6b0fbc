"""Generates tests/golden/hull_cases.npz: inputs and outputs of the REFERENCE's convex_hull.fill_convex_hull
(/root/reference/convex_hull.py, imported as is) on seeded random point sets: 2-D and 3-D, on-lattice and raw
coordinates, resolutions below and above 1, plus the shipped Test_file1 points.  Run in the build container:

    python tests/golden/make_hull_golden.py
"""
import os, sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def cases():
    rng = np.random.default_rng(2024)
    out = []
    for trial in range(36):
        d = 2 + trial % 2
        n = int(rng.integers(d + 2, 30))
        step = [float(rng.choice([0.01, 0.05, 0.1, 0.25, 0.5, 1.0, 2.0])) for _ in range(d)]
        if trial % 3 == 0:
            pts = np.stack([np.around(rng.integers(-10, 11, n) * s, 2) for s in step], 1)
        else:
            pts = rng.normal(size=(n, d)) * np.array(step) * rng.uniform(3, 9 if d == 2 else 6)
        out.append((pts, step))
    return out


def main():
    sys.path.insert(0, "/root/reference")
    import convex_hull as ref
    import pandas as pd
    store = {}
    k = 0
    for pts, step in cases():
        try:
            want = ref.fill_convex_hull(pts.copy(), list(step))
        except Exception:                       # degenerate input for qhull: not a case
            continue
        store[f"c{k}__points"], store[f"c{k}__step"], store[f"c{k}__lattice"] = pts, np.array(step), want
        k += 1
    df = pd.read_csv("/root/reference/Test_file1.txt")
    pts = df[["energy", "angle"]].to_numpy()
    for step in ([0.05, 0.02], [0.01, 0.01]):
        store[f"c{k}__points"], store[f"c{k}__step"] = pts, np.array(step)
        store[f"c{k}__lattice"] = ref.fill_convex_hull(pts.copy(), list(step))
        k += 1
    store["n_cases"] = np.array(k)
    np.savez_compressed(os.path.join(HERE, "hull_cases.npz"), **store)
    print("cases", k, "points", sum(len(store[f"c{i}__lattice"]) for i in range(k)))


if __name__ == "__main__":
    main()
