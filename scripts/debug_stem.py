import torch, torch.nn.functional as F, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from databricks_distributed_deep_learning_b200.models import prepare_for_fused, resnet18
from databricks_distributed_deep_learning_b200.ops import functional as OF
torch.manual_seed(0)
bf = torch.bfloat16
m = prepare_for_fused(resnet18(num_classes=10, impl="fused"), "cuda")
ref = resnet18(num_classes=10, impl="torch").cuda().float()
ref.load_state_dict({k: v.float() for k, v in m.state_dict().items()})
x = torch.randn(16, 3, 64, 64, device="cuda")
def stages(model, x, fused, pad):
    out = []
    if fused:
        model.stem_pad = pad
        x = OF.to_nhwc_bf16(x, pad_channels=pad)
        if x.shape[1] != 3:
            c = model.conv1
            y = F.conv2d(x, OF.pad_in_channels(c.weight, x.shape[1]), None, c.stride, c.padding)
            out.append(("conv1", y))
            x = OF.bn_act(y, model.bn1, relu=True)
        else:
            y = F.conv2d(x, model.conv1.weight, None, model.conv1.stride, model.conv1.padding)
            out.append(("conv1", y))
            x = OF.bn_act(y, model.bn1, relu=True)
        out.append(("bn1", x))
        x = OF.maxpool3x3s2(x)
    else:
        y = model.conv1(x); out.append(("conv1", y))
        x = F.relu(model.bn1(y)); out.append(("bn1", x))
        x = F.max_pool2d(x, 3, 2, 1)
    out.append(("pool", x))
    for i, l in enumerate([model.layer1, model.layer2, model.layer3, model.layer4]):
        x = l(x); out.append((f"layer{i+1}", x))
    if fused:
        f = OF.global_avgpool(x); out.append(("avg", f))
        o = OF.linear(f, model.fc.weight, model.fc.bias)
    else:
        f = torch.flatten(F.adaptive_avg_pool2d(x, 1), 1); out.append(("avg", f))
        o = model.fc(f)
    out.append(("out", o))
    return out
with torch.no_grad():
    r = stages(ref, x, False, 0)
    for pad in (0, 8):
        o = stages(m, x, True, pad)
        for (n, a), (_, b) in zip(o, r):
            a = a.float()
            print(pad, n, tuple(a.shape), "absmax", float(a.abs().max()), "ref", float(b.abs().max()), "rel", float((a - b).abs().max() / b.abs().max()))
