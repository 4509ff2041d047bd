"""Model zoo: the two workload families BASELINE.json:5 names (ResNet, BERT)."""
from .bert import BertConfig, BertForSequenceClassification, bert_base, bert_large, bert_tiny  # noqa: F401
from .resnet import ResNet, resnet18, resnet50  # noqa: F401


def prepare_for_fused(model, device=None, dtype=None, channels_last: bool = True):
    """Put a model into the form the native kernels take: parameters in ``dtype`` (bf16), convolution
    models in channels-last memory, BatchNorm running statistics kept in fp32 (the BN kernels update
    them in fp32 and refuse any other dtype).  Returns the model."""
    import torch
    import torch.nn as nn

    dtype = dtype or torch.bfloat16
    model = model.to(device, dtype) if device is not None else model.to(dtype)
    if channels_last and any(isinstance(m, nn.Conv2d) for m in model.modules()):
        model = model.to(memory_format=torch.channels_last)
    for m in model.modules():
        if isinstance(m, nn.modules.batchnorm._BatchNorm) and m.running_mean is not None:
            m.running_mean.data = m.running_mean.data.float()
            m.running_var.data = m.running_var.data.float()
    return model


def build_model(name: str, **kw):
    table = {"resnet18": resnet18, "resnet50": resnet50, "bert-base": bert_base,
             "bert-large": bert_large, "bert-tiny": bert_tiny}
    if name not in table:
        raise KeyError(f"unknown model {name!r}; choose from {sorted(table)}")
    return table[name](**kw)
