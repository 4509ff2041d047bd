"""tcgen05 / TMEM / TMA GEMM front end (K8) and the linear-layer autograd
functions built on it (``csrc/gemm_tcgen05.cu``).

``gemm_nt`` is the raw call: ``C[M,N] = epilogue(A . B^T)`` with either operand K-major or
MN-major, so that forward, dgrad and wgrad of a linear layer all run on the same kernel
without transposes:

    forward  y  = x W^T + b          A = x  [M,K]           B = W [N,K]
    dgrad    dx = dy W               A = dy [M,N]           B = W read as [K=N, N=Kin] (MN-major)
    wgrad    dW = dy^T x             A = dy read as [K=tok, M=N], B = x read as [K=tok, N=Kin]
                                     (both MN-major; split-K with fp32 red.add for skinny outputs)

:func:`linear` fuses bias / GELU / residual into the epilogue; :func:`ffn` is the whole
transformer MLP (GEMM+bias+GELU -> GEMM+bias) with dGELU fused into the dgrad epilogue.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch.autograd.function import once_differentiable

from . import _lib as L

EPI_NONE, EPI_GELU, EPI_DGELU, EPI_RELU = 0, 1, 2, 3
_NUM_SMS = 148


def set_cluster_mode(mode: int) -> None:
    """0 = auto, 1 = single CTA, 2 = 2-CTA cluster sharing B by TMA multicast, 3 = CTA pairs with
    ``tcgen05.mma.cta_group::2`` (256 x 256 tiles)."""
    L.check(L.lib().ddl_gemm_set_cluster(int(mode)), "ddl_gemm_set_cluster")


def _apply_env_mode() -> None:
    import os

    m = os.environ.get("DDL_GEMM_CLUSTER")
    if m is not None and available():
        set_cluster_mode(int(m))


def available() -> bool:
    try:
        return hasattr(L.lib(), "ddl_gemm_bf16")
    except L.DdlError:
        return False


_env_applied = False
_acc_cache = {}  # (M, N, device index) -> zeroed fp32 split-K accumulator


def usable(x: torch.Tensor, weight: torch.Tensor) -> bool:
    """Shapes/dtypes the tcgen05 kernel takes: bf16, CUDA, inner dims multiples of 8."""
    return (x.is_cuda and x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16 and x.dim() == 2
            and x.shape[1] % 8 == 0 and weight.shape[0] % 8 == 0 and x.shape[0] >= 1 and available())


def gemm_nt(a: torch.Tensor, b: torch.Tensor, M: int, N: int, K: int, *, a_mn: bool = False, b_mn: bool = False,
            bias: Optional[torch.Tensor] = None, residual: Optional[torch.Tensor] = None, aux: Optional[torch.Tensor] = None,
            epilogue: int = EPI_NONE, out: Optional[torch.Tensor] = None, split_k: int = 1, block_n: int = 0,
            col_stats: Optional[torch.Tensor] = None) -> torch.Tensor:
    """C[M,N] = epilogue(A . B^T).  ``a``/``b`` are 2-D row-major bf16 tensors holding the operand
    as [M,K]/[N,K] (default) or as [K,M]/[K,N] (``a_mn``/``b_mn``)."""
    global _env_applied
    if not _env_applied:
        _env_applied = True
        _apply_env_mode()
    dev = a.device
    if out is None:
        out = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
    acc = None
    rezero = 0
    if split_k > 1:
        # fp32 accumulator of the split-K partial products.  Outside graph capture it is cached per shape and handed back
        # zeroed by the conversion kernel (flag 2), which saves one memset kernel per split-K GEMM (100 per BERT-large step);
        # buffers first needed during a capture are not cached (they would belong to that graph's memory pool).
        key = (M, N, dev.index)
        acc = _acc_cache.get(key)
        if acc is None:
            acc = torch.zeros(M, N, dtype=torch.float32, device=dev)
            if not torch.cuda.is_current_stream_capturing():
                _acc_cache[key] = acc
        if key in _acc_cache:
            rezero = 2
    try:
        L.call("ddl_gemm_bf16_ex", a.data_ptr(), int(a_mn), a.stride(0), b.data_ptr(), int(b_mn), b.stride(0), out.data_ptr(), out.stride(0),
               M, N, K, L.ptr(bias), L.ptr(residual), L.ptr(aux), epilogue, split_k, L.ptr(acc), L.ptr(col_stats), block_n, L.stream_ptr(dev))
        if split_k > 1:
            L.call("ddl_f32_to_bf16", acc.data_ptr(), out.data_ptr(), M * N, rezero, L.stream_ptr(dev))
    except Exception:
        _acc_cache.pop((M, N, dev.index), None)  # never reuse an accumulator that may not be zero
        raise
    return out


def pick_split_k(M: int, N: int, K: int, block_n: int = 256) -> int:
    """Split the reduction when the output has too few tiles to fill 148 SMs."""
    tiles = ((M + 127) // 128) * ((N + block_n - 1) // block_n)
    kblocks = (K + 63) // 64
    if tiles >= _NUM_SMS or kblocks < 8:
        return 1
    best, best_eff = 1, tiles / _NUM_SMS
    for s in range(2, 17):
        if s * 4 > kblocks:
            break
        t = tiles * s
        eff = t / (((t + _NUM_SMS - 1) // _NUM_SMS) * _NUM_SMS)
        if eff > best_eff + 0.05:
            best, best_eff = s, eff
    return best


def colsum(x: torch.Tensor) -> torch.Tensor:
    """Column sums of a [rows, cols] bf16 matrix (bias gradient), deterministic."""
    rows, cols = x.shape
    out = torch.empty(cols, dtype=torch.bfloat16, device=x.device)
    nbytes = L.lib().ddl_colsum_scratch_bytes(rows, cols)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
    L.call("ddl_colsum_bf16", x.data_ptr(), rows, cols, x.stride(0), out.data_ptr(), scratch.data_ptr(), nbytes, L.stream_ptr(x.device))
    return out


# Column sums that an upstream backward kernel produced together with a gradient tensor (LayerNorm backward: the sums of
# the gradient of its first input).  Keyed by the gradient's address; the entry keeps the tensor alive until it is taken or
# the backward pass ends, so the address cannot be handed to another tensor in between.
_colsum_offers = {}


def _clear_colsum_offers() -> None:
    _colsum_offers.clear()


def offer_colsum(grad: torch.Tensor, sums: torch.Tensor) -> None:
    if not _colsum_offers:
        torch.autograd.Variable._execution_engine.queue_callback(_clear_colsum_offers)
    _colsum_offers[grad.data_ptr()] = (grad, sums)


def _bias_grad(dy: torch.Tensor) -> torch.Tensor:
    """Column sums of ``dy``: taken from the kernel that produced ``dy`` when it offered them, else a column-sum pass."""
    ent = _colsum_offers.pop(dy.data_ptr(), None)
    if ent is not None and ent[0].shape == dy.shape and ent[0].stride() == dy.stride() and ent[1].shape[0] == dy.shape[1]:
        return ent[1]
    return colsum(dy)


# Where a weight's gradient should be written: id(parameter) -> (weak reference to it, 2-D bf16 view of its slot in the
# flat gradient bucket), registered by the DDP / FSDP wrappers.  The weight-gradient GEMM then writes there directly,
# autograd adopts that view as .grad, and the wrapper's gradient collection has nothing left to copy for this parameter.
_grad_dest = {}
# destinations already handed out during the current backward pass: a weight used by two linear nodes (tied weights) must
# not have both weight-gradient GEMMs write the same view (autograd would then sum two aliases of one buffer)
_handed_out = set()


def _clear_handed_out() -> None:
    _handed_out.clear()


def register_grad_destination(weight: torch.Tensor, view: torch.Tensor) -> None:
    import weakref

    if weight.dim() == 2 and view.dtype == torch.bfloat16 and view.shape == weight.shape and view.is_contiguous():
        _grad_dest[id(weight)] = (weakref.ref(weight), view)


def _grad_destination(weight: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    if weight is None or not _grad_dest:
        return None
    ent = _grad_dest.get(id(weight))
    if ent is None or ent[0]() is not weight:
        return None
    # only when nothing has been accumulated yet (first backward after zero_grad(set_to_none=True)) and only ONCE per
    # backward pass: a second use of the same weight gets a fresh tensor and autograd sums the two
    if weight.grad is not None or id(weight) in _handed_out:
        return None
    if not _handed_out:
        torch.autograd.Variable._execution_engine.queue_callback(_clear_handed_out)
    _handed_out.add(id(weight))
    return ent[1]


def _wgrad(dy: torch.Tensor, x: torch.Tensor, weight: Optional[torch.Tensor] = None) -> torch.Tensor:
    """dW[N, Kin] = dy^T x  with dy [tok, N], x [tok, Kin]."""
    tok, N = dy.shape
    Kin = x.shape[1]
    out = _grad_destination(weight)
    if out is not None and (out.shape[0] != N or out.shape[1] != Kin or out.device != dy.device):
        out = None
    dw = gemm_nt(dy, x, N, Kin, tok, a_mn=True, b_mn=True, split_k=pick_split_k(N, Kin, tok), out=out)
    # a FRESH view object: autograd adopts a gradient without copying only if nobody else holds a reference to it
    return dw if out is None else dw.view(N, Kin)


class GradSlot:
    """Hand-over point for the gradient of a residual connection.

    In ``z = LN(f(x) + x)`` the tensor ``x`` receives two gradients: one through ``f`` (a linear layer's data gradient,
    a GEMM) and one through the skip.  Instead of letting autograd sum them with an elementwise kernel, the op that
    produces the skip gradient (LayerNorm backward, or a linear layer whose epilogue added the residual) ``put``s it
    here and reports "no gradient" for that input; the linear layer at the head of ``f`` - whose backward always runs
    later - ``take``s it and adds it in the epilogue of its data-gradient GEMM.  A slot still full when the backward
    pass ends means a gradient was dropped: the pass raises."""

    __slots__ = ("grad",)
    _pending = set()

    def __init__(self):
        self.grad = None

    def put(self, g: torch.Tensor) -> None:
        if not GradSlot._pending:
            torch.autograd.Variable._execution_engine.queue_callback(GradSlot._check_all_taken)
        self.grad = g
        GradSlot._pending.add(self)

    def take(self) -> Optional[torch.Tensor]:
        g, self.grad = self.grad, None
        GradSlot._pending.discard(self)
        return g

    @staticmethod
    def _check_all_taken() -> None:
        left = [s for s in GradSlot._pending if s.grad is not None]
        for s in left:
            s.grad = None
        GradSlot._pending.clear()
        if left:
            raise L.DdlError(f"GradSlot: {len(left)} residual gradient(s) were never added to a data gradient "
                             "(the slot's consumer did not run in this backward pass)")


class _Linear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, residual, act, in_slot=None, res_slot=None):
        x = x.contiguous()
        M, K = x.shape
        N = weight.shape[0]
        epi = {None: EPI_NONE, "gelu": EPI_GELU, "relu": EPI_RELU}[act]
        pre = torch.empty(M, N, dtype=torch.bfloat16, device=x.device) if act == "gelu" else None
        y = gemm_nt(x, weight, M, N, K, bias=bias, residual=residual, aux=pre, epilogue=epi)
        ctx.save_for_backward(x, weight, pre if act == "gelu" else (y if act == "relu" else None))
        ctx.act, ctx.has_bias, ctx.has_res = act, bias is not None, residual is not None
        ctx.in_slot, ctx.res_slot = in_slot, res_slot
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        x, weight, saved = ctx.saved_tensors
        dy = dy.contiguous()
        M, K = x.shape
        N = weight.shape[0]
        dres = dy if (ctx.has_res and ctx.needs_input_grad[3]) else None
        if dres is not None and ctx.res_slot is not None:
            ctx.res_slot.put(dres)  # summed into the data gradient of the layer that consumes the slot
            dres = None
        if ctx.act == "gelu":
            # elementwise dGELU on the incoming gradient (the fused form lives in `ffn`)
            dy = _dgelu_mul(dy, saved)
        elif ctx.act == "relu":
            dy = dy * (saved > 0)
        addend = ctx.in_slot.take() if ctx.in_slot is not None else None
        dx = None
        if ctx.needs_input_grad[0]:
            dx = gemm_nt(dy, weight, M, K, N, b_mn=True, residual=addend)  # + skip gradient in the epilogue
        dw = _wgrad(dy, x, weight) if ctx.needs_input_grad[1] else None
        db = _bias_grad(dy) if ctx.has_bias and ctx.needs_input_grad[2] else None
        return dx, dw, db, dres, None, None, None


def _dgelu_mul(dy: torch.Tensor, pre: torch.Tensor) -> torch.Tensor:
    """``dy * gelu'(pre)`` in one native pass (the fused-into-the-GEMM form lives in :func:`ffn`)."""
    dy, pre = dy.contiguous(), pre.contiguous()
    out = torch.empty_like(dy)
    L.call("ddl_dgelu_mul", dy.data_ptr(), pre.data_ptr(), out.data_ptr(), dy.numel(), L.stream_ptr(dy.device))
    return out


def linear(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None, activation: Optional[str] = None,
           residual: Optional[torch.Tensor] = None, input_grad_slot: Optional[GradSlot] = None,
           residual_grad_slot: Optional[GradSlot] = None) -> torch.Tensor:
    """``act(x @ weight.T + bias) + residual`` on [M, K] bf16 rows, all in one kernel.  ``input_grad_slot``: a skip
    gradient put there during backward is added to d(x) in the GEMM epilogue; ``residual_grad_slot``: the gradient of
    ``residual`` is put there instead of being returned to autograd (see :class:`GradSlot`)."""
    return _Linear.apply(x, weight, bias, residual, activation, input_grad_slot, residual_grad_slot)


class _FFN(torch.autograd.Function):
    """out = gelu(x W1^T + b1) W2^T + b2 (+ residual).  Five GEMM launches in backward, with
    dGELU fused into the epilogue of the dgrad of the second linear."""

    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2, residual, in_slot=None, res_slot=None):
        ctx.res_is_input = residual is x  # out = f(x) + x: both gradients of x are summed in the last dgrad epilogue
        x = x.contiguous()
        M, H = x.shape
        I = w1.shape[0]
        pre = torch.empty(M, I, dtype=torch.bfloat16, device=x.device)
        h = gemm_nt(x, w1, M, I, H, bias=b1, aux=pre, epilogue=EPI_GELU)
        y = gemm_nt(h, w2, M, w2.shape[0], I, bias=b2, residual=residual)
        ctx.save_for_backward(x, w1, w2, pre, h)
        ctx.has_res = residual is not None
        ctx.in_slot, ctx.res_slot = in_slot, res_slot
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        x, w1, w2, pre, h = ctx.saved_tensors
        dy = dy.contiguous()
        M, H = x.shape
        I = w1.shape[0]
        O = w2.shape[0]
        dw2 = _wgrad(dy, h, w2)
        db2 = _bias_grad(dy)
        # (dy W2) * gelu'(pre); the epilogue also accumulates the column sums of what it stores = the bias gradient of W1
        stats = torch.zeros(2, I, dtype=torch.float32, device=x.device)
        dpre = gemm_nt(dy, w2, M, I, O, b_mn=True, aux=pre, epilogue=EPI_DGELU, col_stats=stats)
        dw1 = _wgrad(dpre, x, w1)
        db1 = stats[0].to(torch.bfloat16)
        dres = dy if (ctx.has_res and ctx.needs_input_grad[5]) else None
        addend = ctx.in_slot.take() if ctx.in_slot is not None else None
        if dres is not None and ctx.res_is_input and addend is None and ctx.needs_input_grad[0]:
            addend, dres = dres, None  # the skip gradient of x rides in the epilogue of d(x)
        elif dres is not None and ctx.res_slot is not None:
            ctx.res_slot.put(dres)
            dres = None
        dx = gemm_nt(dpre, w1, M, H, I, b_mn=True, residual=addend) if ctx.needs_input_grad[0] else None
        return dx, dw1, db1, dw2, db2, dres, None, None


def ffn(x, w1, b1, w2, b2, residual=None, input_grad_slot: Optional[GradSlot] = None,
        residual_grad_slot: Optional[GradSlot] = None) -> torch.Tensor:
    """``gelu(x W1^T + b1) W2^T + b2 (+ residual)``; the slots work as in :func:`linear`."""
    return _FFN.apply(x, w1, b1, w2, b2, residual, input_grad_slot, residual_grad_slot)


# ----------------------------------------------------------------------------- K6: all-gather fused into the GEMM
class GatheredLinear:
    """``y = act(x @ W^T + b)`` where W [N, K] is row-sharded over the ranks in symmetric memory.

    One kernel per rank performs the all-gather AND the GEMM (``ddl_allgather_gemm_bf16``): copier warps pull
    the peers' shards into ``full_w`` while the tensor cores already work on the shards that have arrived
    (local shard first).  ``full_w`` afterwards holds the whole weight, so the layer's backward can use
    the plain kernels.  Forward-only helper (the FSDP wrapper drives backward itself).
    """

    def __init__(self, comm, w_shard: torch.Tensor, n_total: int):
        import ctypes as C

        self.comm = comm
        self.w_shard = w_shard                       # [N / W, K] bf16 in the symmetric heap
        self.N, self.K = n_total, w_shard.shape[1]
        if self.N % (comm.world_size * 128) != 0:
            raise L.DdlError("GatheredLinear: out_features must be a multiple of world_size * 128")
        self.full_w = torch.empty(self.N, self.K, dtype=torch.bfloat16, device=w_shard.device)
        self.flags = torch.zeros(16, dtype=torch.int32, device=w_shard.device)
        sb, off = comm.lookup(w_shard)
        self._ptrs = sb.ptr_array(off)
        self._C = C

    def __call__(self, x: torch.Tensor, bias: Optional[torch.Tensor] = None, activation: Optional[str] = None,
                 residual: Optional[torch.Tensor] = None) -> torch.Tensor:
        M = x.shape[0]
        out = torch.empty(M, self.N, dtype=torch.bfloat16, device=x.device)
        epi = {None: EPI_NONE, "gelu": EPI_GELU, "relu": EPI_RELU}[activation]
        L.call("ddl_allgather_gemm_bf16", self._C.byref(self.comm._ct), self._ptrs, self.full_w.data_ptr(), self.flags.data_ptr(),
               x.data_ptr(), x.stride(0), out.data_ptr(), out.stride(0), M, self.N, self.K, L.ptr(bias), L.ptr(residual), None, epi,
               L.stream_ptr(x.device))
        return out
