"""The GradSlot hand-over protocol (ops/gemm.py) without any kernel: a slot filled during backward must be taken by a
later node of the same backward pass, otherwise the pass raises instead of silently dropping a gradient."""
import pytest
import torch

from databricks_distributed_deep_learning_b200.ops import _lib as L
from databricks_distributed_deep_learning_b200.ops.gemm import GradSlot


class _Skip(torch.autograd.Function):
    """y = x (the skip branch): its gradient goes into the slot, autograd gets none."""

    @staticmethod
    def forward(ctx, x, slot):
        ctx.slot = slot
        return x.clone()

    @staticmethod
    def backward(ctx, g):
        ctx.slot.put(g)
        return None, None


class _Main(torch.autograd.Function):
    """y = 2 x (the main branch): adds whatever the slot holds to its own input gradient."""

    @staticmethod
    def forward(ctx, x, slot):
        ctx.slot = slot
        return 2 * x

    @staticmethod
    def backward(ctx, g):
        add = ctx.slot.take() if ctx.slot is not None else None
        return (2 * g if add is None else 2 * g + add), None


def test_slot_sums_both_branches():
    x = torch.randn(5, requires_grad=True)
    slot = GradSlot()
    h = _Main.apply(x, slot)
    # the skip consumer sits downstream of the main branch, so its backward runs first
    y = h * 3 + _Skip.apply(x, slot) * 5
    y.sum().backward()
    torch.testing.assert_close(x.grad, torch.full((5,), 2.0 * 3 + 5))
    assert slot.grad is None and not GradSlot._pending


def test_unconsumed_slot_raises():
    x = torch.randn(5, requires_grad=True)
    slot = GradSlot()
    y = _Skip.apply(x, slot) * 5 + x  # nobody takes the slot
    with pytest.raises(L.DdlError, match="GradSlot"):
        y.sum().backward()
    assert slot.grad is None and not GradSlot._pending
    # the next backward pass starts clean
    x2 = torch.randn(3, requires_grad=True)
    s2 = GradSlot()
    (_Main.apply(x2, s2) + _Skip.apply(x2, s2)).sum().backward()
    torch.testing.assert_close(x2.grad, torch.full((3,), 3.0))


# ------------------------------------------------------------------ residual_fork (ops/functional.py), same idea for NHWC tensors
class _Producer(torch.autograd.Function):
    """Stands in for an op whose backward kernel accepts a second gradient (BN+residual+ReLU, max-pool)."""

    @staticmethod
    def forward(ctx, x):
        return x * 1.0

    @staticmethod
    def backward(ctx, g):
        from databricks_distributed_deep_learning_b200.ops import functional as OF

        g2 = OF._take_addend(g)
        _Producer.saw_second = g2 is not None
        return g if g2 is None else g + g2


def _cl(t):
    return t.contiguous(memory_format=torch.channels_last)


def test_residual_fork_hands_both_gradients_to_the_producer():
    from databricks_distributed_deep_learning_b200.ops import functional as OF

    x = _cl(torch.randn(2, 4, 3, 3)).requires_grad_()
    p = OF._accepts_second_grad(_Producer.apply(x))
    a, b = OF.residual_fork(p)
    assert a is not p  # the fork is active for tagged tensors
    w2, w3 = _cl(torch.full((2, 4, 3, 3), 2.0)), _cl(torch.full((2, 4, 3, 3), 3.0))  # channels-last gradients, as from NHWC kernels
    ((a * w2).sum() + (b * w3).sum()).backward()
    assert _Producer.saw_second
    torch.testing.assert_close(x.grad, torch.full_like(x, 5.0))
    assert not OF._pending_addend
    # untagged tensors are passed through (autograd sums as usual)
    y = _cl(torch.randn(2, 4, 3, 3)).requires_grad_()
    c, d = OF.residual_fork(y)
    assert c is y and d is y


def test_residual_fork_raises_when_the_skip_gradient_would_be_lost():
    from databricks_distributed_deep_learning_b200.ops import functional as OF

    x = _cl(torch.randn(2, 4, 3, 3)).requires_grad_()
    p = OF._accepts_second_grad(x * 1.0)  # tagged, but produced by an op that knows nothing about pending gradients
    a, b = OF.residual_fork(p)
    w2, w3 = _cl(torch.full((2, 4, 3, 3), 2.0)), _cl(torch.full((2, 4, 3, 3), 3.0))
    with pytest.raises(L.DdlError, match="residual_fork"):
        ((a * w2).sum() + (b * w3).sum()).backward()
    assert not OF._pending_addend
