"""Summarise an .ncu-rep (read here, no GPU): one row per captured launch with the counters the roofline
needs.  usage: python tools/ncu_summary.py gpurun_out/x.ncu-rep > profiles/x.md"""
import csv, io, subprocess, sys

COLS = [
    ("gpu__time_duration.sum", "dur"),
    ("dram__bytes_read.sum", "dram_rd"),
    ("dram__bytes_write.sum", "dram_wr"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
    ("sm__inst_executed_pipe_tensor_op_dmma.avg.pct_of_peak_sustained_active", "dmma%act"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor%el"),
    ("TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed", "tc%el(realtime)"),
    ("sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tc_mem%el"),
    ("l1tex__m_xbar2l1tex_read_bytes.sum", "l2->sm"),
    ("sm__cycles_elapsed.max", "cycles"),
    ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "fp64%act"),
    ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "fp64inst%"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
    ("lts__t_sector_hit_rate.pct", "l2hit%"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem_conf"),
]

def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    names = [c for c in COLS if c[0] in idx]
    print("| kernel | " + " | ".join(s for _, s in names) + " |")
    print("|---|" + "---|" * len(names))
    for r in data:
        cells = []
        for m, s in names:
            v, u = r[idx[m]], units[idx[m]]
            try:
                f = float(v.replace(",", ""))
                if u in ("byte", "Kbyte", "Mbyte", "Gbyte"):
                    f *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
                    v = f"{f/1e6:.1f} MB"
                elif u in ("ns", "us", "ms", "s", "usecond", "msecond", "nsecond", "second"):
                    f *= {"ns": 1e-6, "nsecond": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1, "msecond": 1, "s": 1e3, "second": 1e3}[u]
                    v = f"{f:.3f} ms"
                else:
                    v = f"{f:.4g}"
            except ValueError:
                pass
            cells.append(v)
        print(f"| {r[idx['Kernel Name']]} | " + " | ".join(cells) + " |")

if __name__ == "__main__":
    main(sys.argv[1])
