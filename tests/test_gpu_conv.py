"""Implicit-GEMM convolutions (csrc/conv_igemm.cu) against a plain PyTorch fp32 reference of the same op: forward with
the BatchNorm-statistics epilogue, data gradient and weight gradient, for every convolution shape of ResNet-50
(model at reference notebooks/cv/onnx_experiments.py:19) at small, ragged batch sizes."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

# (Cin, HW_in, Cout, k, stride): the 22 non-stem configurations of SURVEY.md section 2.E.2
RESNET50_CONVS = [
    (64, 56, 64, 1, 1), (64, 56, 64, 3, 1), (64, 56, 256, 1, 1), (256, 56, 64, 1, 1), (256, 56, 128, 1, 1), (128, 56, 128, 3, 2),
    (128, 28, 512, 1, 1), (256, 56, 512, 1, 2), (512, 28, 128, 1, 1), (128, 28, 128, 3, 1), (512, 28, 256, 1, 1), (256, 28, 256, 3, 2),
    (256, 14, 1024, 1, 1), (512, 28, 1024, 1, 2), (1024, 14, 256, 1, 1), (256, 14, 256, 3, 1), (1024, 14, 512, 1, 1), (512, 14, 512, 3, 2),
    (512, 7, 2048, 1, 1), (1024, 14, 2048, 1, 2), (2048, 7, 512, 1, 1), (512, 7, 512, 3, 1),
]


def conv_case(N, cin, hw, cout, k, stride, seed=0):
    """Runs the three native kernels on one shape; returns a dict of relative errors against fp32 autograd."""
    from databricks_distributed_deep_learning_b200.ops import _lib as L

    lib = L.lib()
    bf, cl = torch.bfloat16, torch.channels_last
    pad = k // 2
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn(N, cin, hw, hw, device="cuda", generator=g).to(bf).contiguous(memory_format=cl)
    w = (torch.randn(cout, cin, k, k, device="cuda", generator=g) * (cin * k * k) ** -0.5).to(bf).contiguous(memory_format=cl)
    geom = (N, hw, hw, cin, cout, k, k, stride, pad)
    assert lib.ddl_conv_supported(*geom)
    xr, wr = x.float().requires_grad_(), w.float().requires_grad_()
    ref = F.conv2d(xr, wr, None, stride, pad)
    oh = ref.shape[2]
    dy = torch.randn(N, cout, oh, oh, device="cuda", generator=g).to(bf).contiguous(memory_format=cl)
    ref.backward(dy.float())
    st = L.stream_ptr(x.device)
    # forward + statistics
    y = torch.full((N, cout, oh, oh), float("nan"), dtype=bf, device="cuda").contiguous(memory_format=cl)
    part = torch.full((lib.ddl_conv_fwd_groups(*geom) * 2 * cout,), float("nan"), dtype=torch.float32, device="cuda")
    L.call("ddl_conv_fwd", x.data_ptr(), w.data_ptr(), y.data_ptr(), part.data_ptr(), *geom, st)
    # data gradient (output pre-filled with NaN: every element must be written)
    dx = torch.full((N, cin, hw, hw), float("nan"), dtype=bf, device="cuda").contiguous(memory_format=cl)
    L.call("ddl_conv_dgrad", dy.data_ptr(), w.data_ptr(), dx.data_ptr(), *geom, st)
    # weight gradient
    dw = torch.full((cout, cin, k, k), float("nan"), dtype=bf, device="cuda").contiguous(memory_format=cl)
    scratch = torch.empty(max(lib.ddl_conv_wgrad_scratch_bytes(*geom), 16), dtype=torch.uint8, device="cuda")
    L.call("ddl_conv_wgrad", x.data_ptr(), dy.data_ptr(), dw.data_ptr(), *dw.stride(), scratch.data_ptr(), scratch.numel(), *geom, st)
    torch.cuda.synchronize()

    def rel(a, b):
        return ((a.float() - b).abs().max() / b.abs().max().clamp_min(1e-6)).item()

    extra = {}
    M = N * hw * hw
    if k == 1 and stride == 1 and lib.ddl_conv1x1_supported(M, cin, cout):
        # weight-stationary kernel (csrc/conv1x1.cu) in its data-gradient mode
        dx2 = torch.full((N, cin, hw, hw), float("nan"), dtype=bf, device="cuda").contiguous(memory_format=cl)
        L.call("ddl_conv1x1_dgrad", dy.data_ptr(), w.data_ptr(), dx2.data_ptr(), M, cin, cout, st)
        torch.cuda.synchronize()
        extra["ws_dgrad"] = rel(dx2, xr.grad)

    sums = part.view(-1, 2, cout).sum(0)
    yf = y.float()
    return {
        "fwd": rel(y, ref.detach()),
        "dgrad": rel(dx, xr.grad),
        "wgrad": rel(dw, wr.grad),
        "sum": ((sums[0] - yf.sum((0, 2, 3))).abs().max() / yf.sum((0, 2, 3)).abs().max().clamp_min(1.0)).item(),
        "sumsq": ((sums[1] - yf.square().sum((0, 2, 3))).abs().max() / yf.square().sum((0, 2, 3)).abs().max()).item(),
        **extra,
    }


def _check(errs):
    # bf16 output rounding is 2^-9 relative to the element; the bound is relative to the tensor's maximum
    assert errs["fwd"] < 8e-3, errs
    assert errs["dgrad"] < 8e-3, errs
    assert errs["wgrad"] < 8e-3, errs
    assert errs["sum"] < 2e-3 and errs["sumsq"] < 2e-3, errs
    assert errs.get("ws_dgrad", 0.0) < 8e-3, errs


@pytest.mark.parametrize("cin,hw,cout,k,stride", RESNET50_CONVS)
def test_conv_resnet50_shapes(cin, hw, cout, k, stride):
    # 3 images: ragged against every box (2 images per box at 7x7, 128-pixel tiles of the flattened 1x1 layers)
    _check(conv_case(3, cin, hw, cout, k, stride))


@pytest.mark.parametrize("N,cin,hw,cout,k,stride", [(1, 64, 8, 64, 3, 1), (5, 64, 10, 128, 3, 2), (2, 128, 20, 64, 3, 1), (7, 192, 6, 192, 1, 1),
                                                    (4, 64, 12, 64, 1, 2), (9, 256, 7, 256, 3, 1), (2, 64, 32, 64, 3, 1)])
def test_conv_odd_shapes(N, cin, hw, cout, k, stride):
    _check(conv_case(N, cin, hw, cout, k, stride, seed=1))


def test_conv_autograd_matches_library_path():
    """conv_bn_act through the native kernels against the same module on ATen ops (fp32)."""
    import torch.nn as nn
    from databricks_distributed_deep_learning_b200.ops import functional as OF

    torch.manual_seed(0)
    conv = nn.Conv2d(64, 128, 3, 2, 1, bias=False).cuda()
    bn = nn.BatchNorm2d(128).cuda()
    x = torch.randn(4, 64, 16, 16, device="cuda")
    ref = F.relu(bn(conv(x)))
    ref.square().sum().backward()
    gw_ref, conv.weight.grad = conv.weight.grad.clone(), None
    conv2 = nn.Conv2d(64, 128, 3, 2, 1, bias=False).cuda()
    conv2.weight.data.copy_(conv.weight.data)
    conv2 = conv2.to(torch.bfloat16).to(memory_format=torch.channels_last)
    bn2 = nn.BatchNorm2d(128).cuda()
    bn2.weight.data = bn2.weight.data.to(torch.bfloat16)
    bn2.bias.data = bn2.bias.data.to(torch.bfloat16)
    xb = x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_()
    assert OF.conv_native_usable(xb, conv2)
    out = OF.conv_bn_act(xb, conv2, bn2, relu=True)
    out.float().square().sum().backward()
    assert ((out.float() - ref).abs().max() / ref.abs().max()).item() < 3e-2
    assert ((conv2.weight.grad.float() - gw_ref).abs().max() / gw_ref.abs().max()).item() < 6e-2
    torch.testing.assert_close(bn2.running_mean, bn.running_mean, rtol=2e-2, atol=2e-3)
