"""CPU: host-side logic of the autograd glue that needs no kernel (the kernels themselves are GPU tests)."""
import torch

from s2p_b200 import autograd as A


def test_split_channels_matches_torch_split_without_the_arena():
    # consumers that are not ConvFn never write into the gradient arena: the backward falls back to a concatenation
    g = torch.Generator().manual_seed(0)
    x = torch.randn(2, 3, 4, 12, generator=g, requires_grad=True)
    w = [torch.randn(4, generator=g) for _ in range(3)]
    outs = A.split_channels(x, 4)
    assert len(outs) == 3 and all(o.shape == (2, 3, 4, 4) for o in outs)
    assert all(getattr(o, "_s2p_grad_slot", None) is not None for o in outs)
    (sum((o * wi).sum() for o, wi in zip(outs, w)) + (outs[1] ** 2).sum()).backward()
    x2 = x.detach().clone().requires_grad_(True)
    ref = torch.split(x2, 4, dim=3)
    (sum((o * wi).sum() for o, wi in zip(ref, w)) + (ref[1] ** 2).sum()).backward()
    assert torch.equal(x.grad, x2.grad)


def test_split_channels_unused_slice_gets_zeros_and_arena_slices_line_up():
    x = torch.randn(1, 2, 2, 8, requires_grad=True)
    outs = A.split_channels(x, 4)
    outs[0].sum().backward()                      # the second slice has no consumer
    assert torch.equal(x.grad[..., :4], torch.ones(1, 2, 2, 4)) and torch.equal(x.grad[..., 4:], torch.zeros(1, 2, 2, 4))
    arena = A._GradArena(8)
    like = torch.empty(1, 2, 2, 4)
    a, b = arena.slice_for(0, like), arena.slice_for(1, like)
    assert a.data_ptr() == arena.buf.data_ptr() and b.data_ptr() == arena.buf.data_ptr() + 4 * 4
    assert a.stride() == arena.buf[..., :4].stride()


def test_split_channels_without_grad_is_a_plain_split():
    x = torch.randn(1, 2, 2, 8)
    outs = A.split_channels(x, 4)
    assert torch.equal(outs[1], x[..., 4:]) and getattr(outs[0], "_s2p_grad_slot", None) is None
