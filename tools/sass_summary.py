"""Per-kernel counts of the SASS mnemonics that prove tcgen05 / TMEM / TMA / warp-level MMA in the shipped library.
    python tools/sass_summary.py > profiles/r2_sass_summary.md      (needs cuobjdump; no GPU)"""
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "roma_b200" / "libroma_b200.so"
OPS = ["UTCHMMA", "UTCBAR", "LDTM", "UTMALDG", "UTMASTG", "SYNCS", "HMMA", "FFMA2", "LDGSTS", "MUFU"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True, check=True).stdout
    rows, cur, counts = [], None, None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            if cur:
                rows.append((cur, counts))
            cur, counts = m.group(1), dict.fromkeys(OPS, 0)
            continue
        if cur:
            m = re.search(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
            if m:
                op = m.group(1)
                for o in OPS:
                    if op == o or op.startswith(o + "."):
                        counts[o] += 1
    if cur:
        rows.append((cur, counts))
    demangle = subprocess.run(["c++filt"], input="\n".join(r[0] for r in rows), capture_output=True, text=True).stdout.splitlines()
    print("# SASS evidence — `cuobjdump -sass roma_b200/libroma_b200.so` (sm_100a), per kernel\n")
    print("UTCHMMA = tcgen05.mma, UTCBAR = tcgen05.commit, LDTM = tcgen05.ld (TMEM -> registers), UTMALDG = cp.async.bulk.tensor (TMA load),")
    print("SYNCS = mbarrier ops, HMMA = mma.sync (warp-level tensor core), FFMA2 = packed fp32 FMA.  Kernels with none of these are omitted.\n")
    print("| kernel | " + " | ".join(OPS) + " |")
    print("|---|" + "---:|" * len(OPS))
    tot = dict.fromkeys(OPS, 0)
    for (name, c), dn in zip(rows, demangle):
        for o in OPS:
            tot[o] += c[o]
        if not any(c[o] for o in OPS if o not in ("MUFU",)):
            continue
        short = re.sub(r"\(.*", "", dn)
        short = short.replace("void ", "")
        print(f"| `{short[:110]}` | " + " | ".join(str(c[o]) if c[o] else "" for o in OPS) + " |")
    print("| **total** | " + " | ".join(str(tot[o]) for o in OPS) + " |")
    print(f"\n{len(rows)} kernels in the library.")


if __name__ == "__main__":
    sys.exit(main())
