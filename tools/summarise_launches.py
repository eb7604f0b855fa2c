"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list into a per-kernel markdown table.

    python tools/summarise_launches.py gpurun_out/r1_bench_launches.csv "title" > profiles/r1_bench_launches.md

Own kernels (namespace roma::) are listed individually; library kernels are grouped by family.  The times are
ncu's serialised, cold-cache per-launch durations: use the SHARES, not the absolute numbers (B200_PROFILING.md).
"""
import csv
import re
import sys
from collections import defaultdict

FAMILIES = [("nvjet", "cuBLAS nvjet GEMM"), ("cutlass", "cutlass/cudnn conv"), ("xmma", "cudnn xmma conv"), ("cudnn", "cudnn (attention/conv)"),
            ("sdpa", "cudnn (attention/conv)"), ("batch_norm", "ATen batch_norm"), ("layer_norm", "ATen layer_norm"),
            ("max_pool", "ATen max_pool"), ("CatArray", "ATen cat"), ("elementwise", "ATen elementwise"),
            ("upsample", "ATen upsample"), ("multinomial", "ATen multinomial/topk"), ("topk", "ATen multinomial/topk"),
            ("sort", "ATen multinomial/topk"), ("gatherTopK", "ATen multinomial/topk"), ("reduce", "ATen reduce"),
            ("index", "ATen index/gather")]


def own_kernel_names():
    """Names of every __global__ function under roma_b200/csrc (ncu prints them without the outer namespace)."""
    from pathlib import Path
    names = set()
    for src in (Path(__file__).resolve().parent.parent / "roma_b200" / "csrc").glob("*.cu*"):
        text = src.read_text()
        if "__global__" in text:
            names.update(re.findall(r"\b(\w+_kernel)\b", text))
    return names


OWN = own_kernel_names()


def family(name: str) -> str:
    m = re.match(r"(?:void )?(?:\w+::)*(\w+)", name)
    if m and m.group(1) in OWN:
        tmpl = re.search(r"<([^>(]*)>", name)
        return "roma::" + m.group(1) + (f"<{tmpl.group(1)[:28]}>" if tmpl else "")
    for key, fam in FAMILIES:
        if key in name:
            return fam
    return "other library"


def main():
    path, title = sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else sys.argv[1]
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if l.startswith('"')]
    rd = csv.DictReader(lines)
    for r in rd:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        us = v / 1e3 if unit in ("ns", "nsecond") else v * 1e3 if unit in ("ms", "msecond") else v
        rows.append((r["Kernel Name"], us))
    tot = sum(us for _, us in rows)
    agg = defaultdict(lambda: [0, 0.0])
    for name, us in rows:
        a = agg[family(name)]
        a[0] += 1
        a[1] += us
    own = sum(v[1] for k, v in agg.items() if k.startswith("roma::"))
    print(f"# {title}\n")
    print(f"{len(rows)} launches, {tot / 1e3:.2f} ms summed kernel time (serialised, cold cache); own kernels "
          f"{own / 1e3:.2f} ms = {100 * own / tot:.1f} %.\n")
    print("| kernel / family | launches | total ms | share | avg us |")
    print("|---|---:|---:|---:|---:|")
    for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{k}` | {n} | {us / 1e3:.3f} | {100 * us / tot:.1f} % | {us / n:.1f} |")


if __name__ == "__main__":
    main()
