import sys, torch, time
sys.path.insert(0, '/root/repo')
import roma_b200
from roma_b200.host.encoder import VGG19
dev = torch.device('cuda:0')
vgg = VGG19(torch.bfloat16).to(dev).eval()
x = torch.randn(16, 3, 864, 864, device=dev)
def run(n=5):
    for _ in range(2): vgg(x)
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): vgg(x)
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / n
with torch.inference_mode():
    print("benchmark=False ms", run())
    torch.backends.cudnn.benchmark = True
    print("benchmark=True  ms", run())
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        vgg(x); torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=14, max_name_column_width=100))
