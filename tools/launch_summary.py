"""Summarise an ncu launch list (--metrics gpu__time_duration.sum[,dram__bytes_read.sum,dram__bytes_write.sum] --csv):
per kernel class the launch count, total time, share and mean DRAM traffic per launch.
usage: python tools/launch_summary.py gpurun_out/launches.csv > profiles/rNN_launch_summary.csv"""
import csv, re, sys
from collections import defaultdict


ROLES = {"OzSyrkTiles": "syrk", "OzTrsmTiles": "trsm", "OzInvpTiles": "invp", "OzInvuTiles": "invu", "OzPredTiles": "predict",
         "OzGemmTiles": "gemm"}


def kernel_class(name):
    """ncu prints the digit GEMM either with its functor ('void k_oz<0, 8, 64, OzSyrkTiles>(T4, int, int *)') or, in older
    lists, as a bare 'gpk::k_oz<...>' whose role is inferred from its neighbours on the stream (below)."""
    m = re.search(r"k_oz<.*?(Oz\w+Tiles)", name)
    if m:
        return f"k_oz[{ROLES.get(m.group(1), m.group(1))}]"
    if "k_oz<" in name:
        return "k_oz"
    m = re.match(r"(?:void\s+)?(?:gpk::)?(k_\w+)", name)
    return m.group(1) if m else "other:" + name.split("(")[0][-40:]


def main(path):
    lines = [l for l in open(path) if l.startswith('"')]
    rows = list(csv.DictReader(lines))
    per = defaultdict(dict)
    for r in rows:
        per[int(r["ID"])]["name"] = r["Kernel Name"]
        per[int(r["ID"])]["stream"] = r["Stream"]
        per[int(r["ID"])]["grid"] = r["Grid Size"]
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1.0)
        per[int(r["ID"])][r["Metric Name"]] = v * scale
    # the digit GEMM launches share one base name: label them by their neighbours on the same stream
    streams = defaultdict(list)
    for i in sorted(per):
        per[i]["cls"] = kernel_class(per[i]["name"])
        if not per[i]["cls"].startswith("other:"):
            streams[per[i]["stream"]].append(i)
    for ids in streams.values():
        particles = "?"
        for j, i in enumerate(ids):
            if per[i]["cls"] == "k_cov_lower":          # grid (tiles, P): the particle group every following launch works on
                particles = per[i]["grid"].strip("()").split(",")[1].strip()
            per[i]["P"] = particles
            if per[i]["cls"] != "k_oz":
                continue
            prev = per[ids[j - 1]]["cls"] if j else ""
            nxt = per[ids[j + 1]]["cls"] if j + 1 < len(ids) else ""
            role = ("invu" if prev == "k_oz_slice_mt" else "trsm" if prev == "k_oz_slice_d" else
                    "invp" if nxt == "k_oz_mt_scale" else "syrk" if nxt == "k_diag_block" else "?")
            per[i]["cls"] = f"k_oz[{role}]"
    agg = defaultdict(lambda: [0, 0.0, 0.0, 0.0])
    for k in per.values():
        if k["cls"].startswith("other:"):
            continue
        a = agg[f'{k["cls"]} P={k.get("P", "?")}']
        a[0] += 1
        a[1] += k.get("gpu__time_duration.sum", 0.0)
        a[2] += k.get("dram__bytes_read.sum", 0.0)
        a[3] += k.get("dram__bytes_write.sum", 0.0)
    total = sum(a[1] for c, a in agg.items() if not c.startswith("other:"))
    print(f"# {len(per)} launches, {total:.1f} ms in gpk kernels (cold-cache, serialised under ncu: compare SHARES)")
    print("kernel,launches,total_ms,share,mean_ms,dram_read_MB_per_launch,dram_write_MB_per_launch,dram_GBps")
    for c, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        if c.startswith("other:"):
            continue
        print(f"{c},{a[0]},{a[1]:.3f},{a[1] / total:.4f},{a[1] / a[0]:.4f},{a[2] / a[0] / 1e6:.2f},{a[3] / a[0] / 1e6:.2f},"
              f"{(a[2] + a[3]) / (a[1] * 1e-3) / 1e9 if a[1] else 0:.0f}")


if __name__ == "__main__":
    main(sys.argv[1])
