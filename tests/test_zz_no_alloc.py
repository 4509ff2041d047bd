"""Allocation-free BatchNorm entry point, called straight through the C ABI.

``ddl_bn_act_fwd_train_noalloc`` is for host programs that cannot allocate device memory
themselves: a server whose own arena already owns the whole GPU, or a latency-critical
loop that must stay clear of ``cudaMalloc``'s device-wide synchronisation.  It takes the
same inputs and outputs as ``ddl_bn_act_fwd_train`` but keeps its cross-block partial sums
in a 2 MiB array that is part of the library image.

What is checked here: the output agrees with a plain PyTorch fp32 ``F.batch_norm`` (+ residual, ReLU) of the same
inputs, including the statistics and the running averages; and on two calls in a row (different data, different shapes, so the
second call finds the first call's leftovers in the shared array) the outputs agree with
the ordinary entry point given a caller-allocated scratch buffer, within the tolerance
the BatchNorm tests in ``test_gpu_kernels.py`` use; and a shape whose scratch need exceeds
the built-in array is refused with a readable message.

The file name sorts last on purpose: a ``pytest -x`` run has already executed every other
test before it gets here.  It also passes when run alone in a fresh process (the first
call on the device then performs the one-time address look-up).
"""
import ctypes as C
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "databricks_distributed_deep_learning_b200", "ops", "libb200ddl.so")
MAX_BYTES = 2 << 20  # DDL_BN_NOALLOC_MAX_BYTES in include/b200ddl.h

vp, ll, i32, f32, sz = C.c_void_p, C.c_longlong, C.c_int, C.c_float, C.c_size_t


def _load():
    lib = C.CDLL(LIB)
    lib.ddl_last_error.restype = C.c_char_p
    lib.ddl_bn_scratch_bytes.restype = sz
    lib.ddl_bn_scratch_bytes.argtypes = [ll, i32]
    lib.ddl_bn_act_fwd_train.restype = i32
    lib.ddl_bn_act_fwd_train.argtypes = [vp] * 12 + [sz, ll, i32, f32, f32, i32, vp]
    lib.ddl_bn_act_fwd_train_noalloc.restype = i32
    lib.ddl_bn_act_fwd_train_noalloc.argtypes = [vp] * 11 + [ll, i32, f32, f32, i32, vp]
    return lib


def _case(N, Cc, HW, with_res, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    bf = torch.bfloat16
    M = N * HW * HW
    mk = lambda *s: torch.randn(*s, device="cuda", generator=g)
    return {
        "M": M, "C": Cc,
        "x": (mk(M, Cc) * 1.5 + 0.3).to(bf), "res": mk(M, Cc).to(bf) if with_res else None,
        "gamma": (1 + 0.2 * mk(Cc)).to(bf), "beta": (0.2 * mk(Cc)).to(bf),
    }


def _outputs(case):
    M, Cc = case["M"], case["C"]
    f = lambda: torch.empty(Cc, dtype=torch.float32, device="cuda")
    return {"y": torch.empty(M, Cc, dtype=torch.bfloat16, device="cuda"), "mean": f(), "invstd": f(), "scale": f(), "shift": f(),
            "rm": torch.zeros(Cc, device="cuda"), "rv": torch.ones(Cc, device="cuda")}


def _p(t):
    return None if t is None else t.data_ptr()


def _run_pair(lib, case, relu):
    M, Cc = case["M"], case["C"]
    stream = torch.cuda.current_stream().cuda_stream
    a, b = _outputs(case), _outputs(case)
    nbytes = lib.ddl_bn_scratch_bytes(M, Cc)
    assert 0 < nbytes <= MAX_BYTES
    scratch = torch.empty(nbytes, dtype=torch.uint8, device="cuda")  # the ordinary entry point's caller-owned scratch
    rc = lib.ddl_bn_act_fwd_train(_p(case["x"]), _p(case["res"]), _p(case["gamma"]), _p(case["beta"]), _p(a["rm"]), _p(a["rv"]),
                                  _p(a["y"]), _p(a["mean"]), _p(a["invstd"]), _p(a["scale"]), _p(a["shift"]), _p(scratch), nbytes,
                                  M, Cc, 1e-5, 0.1, int(relu), stream)
    assert rc == 0, lib.ddl_last_error().decode()
    rc = lib.ddl_bn_act_fwd_train_noalloc(_p(case["x"]), _p(case["res"]), _p(case["gamma"]), _p(case["beta"]), _p(b["rm"]),
                                          _p(b["rv"]), _p(b["y"]), _p(b["mean"]), _p(b["invstd"]), _p(b["scale"]), _p(b["shift"]),
                                          M, Cc, 1e-5, 0.1, int(relu), stream)
    assert rc == 0, lib.ddl_last_error().decode()
    torch.cuda.synchronize()
    for k in a:
        # same tolerance as test_bn_act_train_fwd_bwd in test_gpu_kernels.py
        assert torch.allclose(a[k].float(), b[k].float(), rtol=2e-2, atol=3e-2), (k, (a[k].float() - b[k].float()).abs().max().item())
    assert torch.isfinite(b["y"].float()).all()
    _check_against_torch(case, b, relu)
    return b


def _check_against_torch(case, out, relu):
    """fp32 oracle: the same batch norm written with torch ops on the [M, C] matrix (channels = columns)."""
    import torch.nn.functional as F

    x = case["x"].float()
    rm, rv = torch.zeros(case["C"], device="cuda"), torch.ones(case["C"], device="cuda")
    ref = F.batch_norm(x, rm, rv, case["gamma"].float(), case["beta"].float(), training=True, momentum=0.1, eps=1e-5)
    if case["res"] is not None:
        ref = ref + case["res"].float()
    if relu:
        ref = F.relu(ref)
    assert torch.allclose(out["y"].float(), ref, rtol=2e-2, atol=3e-2), (out["y"].float() - ref).abs().max().item()
    assert torch.allclose(out["mean"], x.mean(0), rtol=1e-3, atol=1e-3)
    assert torch.allclose(out["invstd"], (x.var(0, unbiased=False) + 1e-5).rsqrt(), rtol=1e-3, atol=1e-3)
    # running statistics: momentum 0.1, unbiased variance, as torch updates them
    assert torch.allclose(out["rm"], rm, rtol=1e-3, atol=1e-3)
    assert torch.allclose(out["rv"], rv, rtol=1e-3, atol=1e-3)


def test_noalloc_matches_ordinary_entry_point_on_two_calls_in_a_row():
    lib = _load()
    # call 1: wide channels, many blocks -> fills a large part of the shared array
    first = _run_pair(lib, _case(8, 256, 28, with_res=True, seed=1), relu=True)
    # call 2: different shape and data, no residual: must not be disturbed by call 1's leftovers
    second = _run_pair(lib, _case(4, 64, 56, with_res=False, seed=2), relu=False)
    assert first["y"].shape != second["y"].shape


def test_noalloc_rejects_sizes_beyond_the_builtin_scratch():
    lib = _load()
    M, Cc = 256 * 7 * 7 * 4, 2048  # ResNet-50's last stage at a large batch: ~4.9 MB of partial sums
    need = lib.ddl_bn_scratch_bytes(M, Cc)
    assert need > MAX_BYTES
    x = torch.zeros(8, dtype=torch.bfloat16, device="cuda")
    o = torch.full((Cc,), 7.0, dtype=torch.float32, device="cuda")
    rc = lib.ddl_bn_act_fwd_train_noalloc(_p(x), None, None, None, None, None, _p(x), _p(o), _p(o), _p(o), _p(o), M, Cc, 1e-5, 0.1, 0,
                                          torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.ddl_last_error().decode()
    assert "noalloc" in msg and "scratch" in msg and str(need) in msg, msg
    torch.cuda.synchronize()
    assert bool((o == 7.0).all()), "a refused call must not launch anything"
