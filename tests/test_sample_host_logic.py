"""Host-side logic of the sampling path that runs without a GPU (matcher.py:468-495 restated in roma_b200/ops.py)."""
import pytest
import torch


def test_small_multinomial_defers_to_torch_on_any_device():
    """Draws outside the regime the own top-k claims (n < 20000 or k <= 4096) are torch.multinomial verbatim: same
    generator in, same indices out, same generator state afterwards."""
    from roma_b200 import ops
    p = torch.rand(5000, generator=torch.Generator().manual_seed(0))
    g1, g2 = torch.Generator().manual_seed(7), torch.Generator().manual_seed(7)
    want = torch.multinomial(p, 1000, replacement=False, generator=g1)
    got = ops.multinomial_without_replacement(p, 1000, generator=g2)
    assert torch.equal(want, got) and torch.equal(g1.get_state(), g2.get_state())
    # k <= 4096 with a long vector: still torch
    p = torch.rand(30000, generator=torch.Generator().manual_seed(1))
    g1, g2 = torch.Generator().manual_seed(9), torch.Generator().manual_seed(9)
    assert torch.equal(torch.multinomial(p, 4096, generator=g1), ops.multinomial_without_replacement(p, 4096, generator=g2))


def test_large_multinomial_has_no_cpu_fallback():
    """In the regime served by roma_topk_desc a CPU tensor is refused loudly (the product path never computes on the host)."""
    from roma_b200 import ops
    p = torch.rand(40000, generator=torch.Generator().manual_seed(2))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.multinomial_without_replacement(p, 10000)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.topk_indices(p, 10000)


def test_sample_refuses_cpu_tensors():
    from roma_b200 import ops
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.sample(torch.rand(64, 4), torch.rand(64), num=8)
