"""torch.profiler view of one bench step (which library kernels the PyTorch bodies spend their time in)."""
import sys
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import roma_b200
from bench import synthetic_pairs
from torch.profiler import profile, ProfilerActivity

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
dt = getattr(torch, sys.argv[2]) if len(sys.argv) > 2 else torch.bfloat16
dev = torch.device("cuda:0")
w, d = roma_b200.random_init_weights(0)
m = roma_b200.roma_outdoor(dev, w, d, amp_dtype=dt)
A8, B8 = synthetic_pairs(B, (864, 864), 1234)
A, Bm = A8.to(dev), B8.to(dev)
for _ in range(2):
    warp, cert = m.match(A, Bm, batched=True, device=dev)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], record_shapes=False) as prof:
    warp, cert = m.match(A, Bm, batched=True, device=dev)
    for i in range(B):
        m.sample(warp[i], cert[i])
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=160, max_name_column_width=90))

if len(sys.argv) > 3 and sys.argv[3] == "shapes":
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], record_shapes=True) as prof2:
        warp, cert = m.match(A, Bm, batched=True, device=dev)
        for i in range(B):
            m.sample(warp[i], cert[i])
        torch.cuda.synchronize()
    rows = [e for e in prof2.key_averages(group_by_input_shape=True)
            if e.key in ("aten::copy_", "aten::cat", "aten::add_", "aten::mul", "aten::add", "aten::sub", "aten::div", "aten::clone",
                         "aten::contiguous", "aten::_to_copy", "aten::max_pool2d_with_indices", "aten::upsample_bicubic2d",
                         "aten::_upsample_bicubic2d_aa", "aten::index", "aten::where", "aten::sigmoid", "aten::topk", "aten::sort",
                         "aten::cumsum", "aten::exponential_", "aten::log", "aten::native_layer_norm", "aten::gelu", "aten::addmm",
                         "aten::mm", "aten::bmm", "aten::cudnn_convolution", "aten::cudnn_convolution_relu")]
    rows.sort(key=lambda e: -e.self_device_time_total)
    print("\n== by input shape (self CUDA us, calls, shapes) ==")
    for e in rows[:60]:
        print(f"{e.key:34s} {e.self_device_time_total:10.0f} {e.count:4d} {str(e.input_shapes)[:150]}")
