"""Summarise `ncu -i X.ncu-rep --page raw --csv` (one row per profiled launch) into a markdown table.

    python tools/summarise_ncu_raw.py raw.csv [--last]      # --last: keep the last launch of every (kernel, grid)
"""
import csv
import re
import sys

COLS = [
    ("time ms", "gpu__time_duration.sum", lambda v, u: v / (1e6 if u in ("ns", "nsecond") else 1e3 if u in ("us", "usecond") else 1.0)),
    ("regs", "launch__registers_per_thread", None),
    ("DRAM rd MB", "dram__bytes_read.sum", lambda v, u: v / {"byte": 1e6, "Kbyte": 1e3, "Mbyte": 1.0, "Gbyte": 1e-3}.get(u, 1e6)),
    ("DRAM wr MB", "dram__bytes_write.sum", lambda v, u: v / {"byte": 1e6, "Kbyte": 1e3, "Mbyte": 1.0, "Gbyte": 1e-3}.get(u, 1e6)),
    ("DRAM thr %", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", None),
    ("L2 thr %", "lts__throughput.avg.pct_of_peak_sustained_elapsed", None),
    ("L1 thr %", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", None),
    ("warps act %", "sm__warps_active.avg.pct_of_peak_sustained_active", None),
    ("issue act %", "sm__inst_issued.avg.pct_of_peak_sustained_active", None),
    ("tensor pipe %", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", None),
    ("fma pipe %", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", None),
    ("L1 hit %", "l1tex__t_sector_hit_rate.pct", None),
    ("L2 hit %", "lts__t_sector_hit_rate.pct", None),
]


def num(s):
    try:
        return float(s.replace(",", ""))
    except Exception:
        return None


def main():
    path = sys.argv[1]
    last = "--last" in sys.argv
    with open(path, newline="") as f:
        lines = [l for l in f if not l.startswith("==")]
    rd = csv.reader(lines)
    header = next(rd)
    units = next(rd)
    idx = {h: i for i, h in enumerate(header)}
    rows = []
    for r in rd:
        if len(r) < len(header):
            continue
        name = r[idx["Kernel Name"]]
        name = re.sub(r"\(.*$", "", name).replace("void ", "").strip()
        grid = r[idx["Grid Size"]] if "Grid Size" in idx else ""
        vals = []
        for label, key, conv in COLS:
            cands = [h for h in header if h == key or h.startswith(key)]
            v = None
            if cands:
                i = idx[cands[0]]
                v = num(r[i])
                if v is not None and conv:
                    v = conv(v, units[i])
            vals.append(v)
        rows.append((name, grid, vals))
    if last:
        keep = {}
        for row in rows:
            keep[(row[0], row[1])] = row
        rows = list(keep.values())
    print("| kernel | grid | " + " | ".join(c[0] for c in COLS) + " |")
    print("|---|---|" + "---:|" * len(COLS))
    for name, grid, vals in rows:
        cells = ["-" if v is None else (f"{v:.4f}" if c[0] == "time ms" else f"{v:.1f}") for v, c in zip(vals, COLS)]
        print(f"| `{name[:70]}` | {grid} | " + " | ".join(cells) + " |")


if __name__ == "__main__":
    main()
