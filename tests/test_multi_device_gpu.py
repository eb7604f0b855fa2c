"""GPU (needs two devices; skipped otherwise): the C-ABI calls follow the TENSORS' device, not the process's current one —
ops._run makes it current around every call, and the library's shared-memory opt-ins are tracked per device."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_kernels_on_cuda1_while_current_device_is_0():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    from roma_b200 import ops
    torch.cuda.set_device(0)
    outs = []
    for dev in (torch.device("cuda:0"), torch.device("cuda:1")):
        g = torch.Generator().manual_seed(0)
        x = torch.randn(2, 256, 24, 24, generator=g).to(torch.bfloat16).to(dev).contiguous(memory_format=torch.channels_last)
        y = torch.randn(2, 256, 24, 24, generator=g).to(torch.bfloat16).to(dev).contiguous(memory_format=torch.channels_last)
        flow = (torch.rand(2, 2, 24, 24, generator=g) * 2.2 - 1.1).to(dev)
        pw, pb = torch.randn(64, 2, generator=g).to(dev), torch.randn(64, generator=g).to(dev)
        assert torch.cuda.current_device() == 0
        d = ops.refiner_inputs(x, y, flow, None, None, 1.0, 2, pad_to=8)          # tile kernel: > 48 KB shared-memory opt-in
        mu = ops.gp_posterior(x, y, pw, pb)                                       # tcgen05 GEMMs, two-lane schedule, plane GEMMs
        pts = (torch.rand(9000, 4, generator=g) * 2 - 1).to(dev)
        dens = ops.kde(pts, half=True)                                            # exact kde: table in > 48 KB of shared memory
        torch.cuda.synchronize(dev)
        assert d.device == dev and mu.device == dev and dens.device == dev
        outs.append((d.cpu(), mu.cpu(), dens.cpu()))
    assert torch.cuda.current_device() == 0
    for a, b in zip(outs[0], outs[1]):
        assert torch.equal(a, b)                                                  # same bits on both devices
