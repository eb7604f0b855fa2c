"""cuBLAS time of the refiner 1x1 convolutions (NHWC GEMMs with bias) at the step's shapes, against their HBM floors."""
import torch, json
import torch.nn.functional as F
dev = torch.device("cuda:0")
flush = torch.zeros(256 * 1024 * 1024 // 4, device=dev)
def t(fn, it=10):
    for _ in range(3): fn()
    torch.cuda.synchronize(); tot = 0.0
    for _ in range(it):
        flush.add_(1.0)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize(); tot += s.elapsed_time(e)
    return tot / it
for M, C in [(16 * 432 * 432, 144), (16 * 280 * 280, 144), (16 * 216 * 216, 576), (16 * 140 * 140, 576), (16 * 108 * 108, 1144), (16 * 70 * 70, 1144),
             (16 * 40 * 40, 1384)]:
    x = torch.randn(M, C, device=dev, dtype=torch.bfloat16)
    w = torch.randn(C, C, device=dev, dtype=torch.bfloat16) * 0.05
    b = torch.randn(C, device=dev, dtype=torch.bfloat16)
    ms = t(lambda: F.linear(x, w, b))
    byt = 2 * M * C * 2
    fl = 2.0 * M * C * C
    print(json.dumps({"M": M, "C": C, "ms": round(ms, 4), "GBps": round(byt / ms / 1e6, 1), "TFLOPs": round(fl / ms / 1e9, 1),
                      "hbm_floor_ms": round(byt / 6.55e9, 4)}))
