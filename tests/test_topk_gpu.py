"""roma_topk_desc / ops.multinomial_without_replacement against ATen on the same GPU: indices identical, ties included
(reference: the two torch.multinomial draws of RegressionMatcher.sample, romatch/models/matcher.py:482-494)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _scores(kind, n, dtype, g):
    if kind == "exp":              # what multinomial feeds topk: p / Exp(1), p mostly 1 (thresholded certainty)
        p = torch.where(torch.rand(n, generator=g, device="cuda") > 0.4, torch.ones((), device="cuda"),
                        torch.rand(n, generator=g, device="cuda") * 0.05)
        q = torch.empty(n, device="cuda").exponential_(1, generator=g)
        return (p / q).to(dtype)
    if kind == "ties":             # few distinct values: every boundary is a tie
        return (torch.randint(0, 37, (n,), generator=g, device="cuda").float() * 0.25).to(dtype)
    if kind == "zeros":            # fewer non-zero scores than k: the tail is taken from the zeros in index order
        x = torch.zeros(n, device="cuda")
        idx = torch.randperm(n, generator=g, device="cuda")[: n // 50]
        x[idx] = torch.rand(idx.numel(), generator=g, device="cuda") + 0.1
        return x.to(dtype)
    if kind == "signed":
        return (torch.randn(n, generator=g, device="cuda") * 3).to(dtype)
    if kind == "inf":
        x = torch.rand(n, generator=g, device="cuda")
        x[:: 977] = float("inf")
        return x.to(dtype)
    raise ValueError(kind)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16, torch.bfloat16])
@pytest.mark.parametrize("kind", ["exp", "ties", "zeros", "signed", "inf"])
@pytest.mark.parametrize("n,k", [(1492992, 40000), (40000, 10000), (40000, 40000), (20000, 4097), (3612672, 40000),
                                 (123457, 50001)])
def test_topk_indices_match_aten(kind, n, k, dtype):
    from roma_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(n + k)
    q = _scores(kind, n, dtype, g)
    ref = torch.topk(q, k)
    got = ops.topk_indices(q, k)
    assert got.dtype == torch.int64 and got.shape == (k,)
    same = got == ref.indices
    if not bool(same.all()):
        bad = (~same).nonzero().flatten()
        i = int(bad[0])
        raise AssertionError(f"{int(bad.numel())} of {k} indices differ, first at {i}: own {int(got[i])} (value {float(q[got[i]])}) "
                             f"vs ATen {int(ref.indices[i])} (value {float(ref.values[i])})")


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16])
def test_multinomial_without_replacement_is_torch_multinomial(dtype):
    """Same generator state in, same indices out, same generator state afterwards."""
    from roma_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(5)
    for n, k in ((1492992, 40000), (40000, 10000)):
        p = torch.rand(n, generator=g, device="cuda")
        p = torch.where(p > 0.5, torch.ones_like(p), p * 0.05).to(dtype)
        g1 = torch.Generator(device="cuda").manual_seed(11)
        g2 = torch.Generator(device="cuda").manual_seed(11)
        ref = torch.multinomial(p, k, replacement=False, generator=g1)
        got = ops.multinomial_without_replacement(p, k, generator=g2)
        assert torch.equal(ref, got)
        assert torch.equal(g1.get_state(), g2.get_state())
        assert torch.equal(torch.rand(4, device="cuda", generator=g1), torch.rand(4, device="cuda", generator=g2))


def test_small_problems_stay_on_aten():
    from roma_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(3)
    p = torch.rand(5000, generator=g, device="cuda")
    g1 = torch.Generator(device="cuda").manual_seed(1)
    g2 = torch.Generator(device="cuda").manual_seed(1)
    assert torch.equal(torch.multinomial(p, 1000, generator=g1), ops.multinomial_without_replacement(p, 1000, generator=g2))


def test_topk_rejects_bad_arguments():
    from roma_b200 import ops
    with pytest.raises(RuntimeError):
        ops.topk_indices(torch.zeros(8), 2)                       # CPU tensor: no fallback
    with pytest.raises(ValueError):
        ops.topk_indices(torch.zeros(4, 4, device="cuda"), 2)
    with pytest.raises(Exception):
        ops.topk_indices(torch.zeros(8, device="cuda"), 9)       # k > n
