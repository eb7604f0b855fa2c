"""Accuracy + speed of the tcgen05 NT GEMM building block vs fp64: python tools/tc_gemm_check.py"""
import sys, json
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from roma_b200 import ops
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
for (batch, M, N, K) in ((2, 1600, 1600, 512), (16, 1600, 1600, 512), (16, 1536, 128, 1408), (16, 3712, 128, 128)):
    a = torch.randn(batch, M, K, generator=g); b = torch.randn(batch, N, K, generator=g)
    want = (a[:1].double().to(dev) @ b[:1].double().to(dev).mT)
    scale = want.abs().max().item()
    for dt, split in ((torch.float32, 3), (torch.float32, 2), (torch.bfloat16, 1), (torch.float16, 1)):
        ad, bd = a.to(dt).to(dev), b.to(dt).to(dev)
        w = want if dt == torch.float32 else (ad[:1].double() @ bd[:1].double().mT)
        c = ops.tc_gemm_nt(ad, bd, split)
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(5):
            c = ops.tc_gemm_nt(ad, bd, split)
        e.record(); torch.cuda.synchronize()
        ms = s.elapsed_time(e) / 5
        ref = (ad[:1].float() @ bd[:1].float().mT)
        print(json.dumps({"shape": [batch, M, N, K], "dtype": str(dt), "split": split, "ms": round(ms, 4),
                          "TFLOPs_alg": round(2.0 * batch * M * N * K / ms / 1e9, 1),
                          "max_err_rel_to_max": (c[:1].double() - w).abs().max().item() / scale,
                          "torch_fp32_matmul_err": (ref.double() - w).abs().max().item() / scale}))
