"""Time ops.topk_indices / ops.multinomial_without_replacement against torch.topk / torch.multinomial (one B200)."""
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from roma_b200 import ops  # noqa: E402


def timed(fn, iters=50):
    for _ in range(5):
        fn()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tot = 0.0
    for _ in range(iters):
        flush.zero_()
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        tot += s.elapsed_time(e)
    return tot / iters * 1e3


def main():
    g = torch.Generator(device="cuda").manual_seed(0)
    for n, k, dt in ((1492992, 40000, torch.float32), (40000, 10000, torch.float16), (3612672, 40000, torch.float32)):
        p = torch.where(torch.rand(n, generator=g, device="cuda") > 0.4, torch.ones((), device="cuda"),
                        torch.rand(n, generator=g, device="cuda") * 0.05).to(dt)
        q = (p.float() / torch.empty(n, device="cuda").exponential_(1, generator=g)).to(dt)
        row = dict(n=n, k=k, dtype=str(dt),
                   topk_own_us=timed(lambda: ops.topk_indices(q, k)), topk_aten_us=timed(lambda: torch.topk(q, k)),
                   multinomial_own_us=timed(lambda: ops.multinomial_without_replacement(p, k, generator=g)),
                   multinomial_aten_us=timed(lambda: torch.multinomial(p, k, replacement=False, generator=g)))
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
