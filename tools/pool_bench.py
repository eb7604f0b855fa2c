import torch, sys
sys.path.insert(0, '.')
from roma_b200 import ops
for C, s in ((64, 864), (128, 432), (256, 216), (64, 560)):
    x = torch.randn(16, C, s, s, device='cuda', dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    for _ in range(3): ops.maxpool2x2(x)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for _ in range(10): ops.maxpool2x2(x)
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 10
    print(C, s, round(ms * 1e3, 1), 'us', round(x.numel() * 2 * 1.25 / ms / 1e6), 'GB/s')
