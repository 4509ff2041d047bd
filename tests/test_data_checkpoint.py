import os

import pytest
import torch

from databricks_distributed_deep_learning_b200 import DDP
from databricks_distributed_deep_learning_b200.models import bert_tiny, resnet18
from databricks_distributed_deep_learning_b200.utils import checkpoint as ck
from databricks_distributed_deep_learning_b200.utils.data import (
    DevicePrefetcher, SyntheticTokenLoader, cifar_shape_loader, imagenet_shape_loader)


def test_image_loader_shapes_and_sharding():
    a = cifar_shape_loader(8, steps=3, rank=0, world_size=2, pin_memory=False)
    b = cifar_shape_loader(8, steps=3, rank=1, world_size=2, pin_memory=False)
    xa, ya = next(iter(a))
    xb, _ = next(iter(b))
    assert xa.shape == (8, 3, 32, 32) and xa.dtype == torch.float32 and ya.dtype == torch.int64
    assert int(ya.max()) < 10 and not torch.equal(xa, xb)
    assert len(list(a)) == 3
    a2 = cifar_shape_loader(8, steps=3, rank=0, world_size=2, pin_memory=False)
    assert torch.equal(next(iter(a2))[0], xa), "loader must be deterministic per (seed, rank)"
    im = imagenet_shape_loader(2, steps=1, pool=1, pin_memory=False)
    x, y = next(iter(im))
    assert x.shape == (2, 3, 224, 224) and int(y.max()) < 1000
    assert im.bytes_per_batch == 2 * 3 * 224 * 224 * 4 + 2 * 8


def test_token_loader_padding_layout():
    l = SyntheticTokenLoader(4, seq_len=32, min_len=8, steps=2, pin_memory=False, vocab_size=512)
    b = next(iter(l))
    assert b["input_ids"].shape == (4, 32) and b["labels"].shape == (4,)
    m = b["attention_mask"]
    assert ((m[:, :-1] - m[:, 1:]) >= 0).all(), "right padded"
    assert (b["input_ids"][m == 0] == 0).all() and (b["input_ids"][:, 0] == 101).all()


def test_prefetcher_cpu_passthrough():
    l = cifar_shape_loader(2, steps=4, pin_memory=False)
    assert len(list(DevicePrefetcher(l, torch.device("cpu")))) == 4


def test_state_dict_checkpoint_roundtrip_through_ddp(tmp_path):
    m = resnet18(num_classes=10, small_input=True)
    ddp = DDP(m)
    opt = torch.optim.SGD(ddp.parameters(), lr=0.1, momentum=0.9)
    ddp(torch.randn(2, 3, 32, 32)).sum().backward()
    opt.step()
    p = str(tmp_path / "ckpt.pt")
    assert ck.save_checkpoint(p, ddp, opt, step=7, extra={"note": "x"}) == p
    blob = torch.load(p, weights_only=False)
    assert not any(k.startswith("module.") for k in blob["model"])
    fresh = resnet18(num_classes=10, small_input=True)
    meta = ck.load_checkpoint(p, fresh)
    assert meta["step"] == 7 and meta["extra"] == {"note": "x"}
    for (k, a), (_, b) in zip(m.state_dict().items(), fresh.state_dict().items()):
        assert torch.equal(a, b), k
    # resume into a wrapped model + optimizer, and accept stock-DDP style prefixed files
    ddp2 = DDP(resnet18(num_classes=10, small_input=True))
    opt2 = torch.optim.SGD(ddp2.parameters(), lr=0.1, momentum=0.9)
    ck.load_checkpoint(p, ddp2, opt2)
    assert opt2.state_dict()["state"].keys() == opt.state_dict()["state"].keys()
    torch.save(ddp.state_dict(), str(tmp_path / "stock.pt"))
    ck.load_checkpoint(str(tmp_path / "stock.pt"), resnet18(num_classes=10, small_input=True))


def test_safetensors_roundtrip_and_hf_interop(tmp_path):
    m = bert_tiny()
    p = str(tmp_path / "model.safetensors")
    ck.save_safetensors(p, DDP(m))
    m2 = bert_tiny()
    ck.load_safetensors(p, m2)
    for (k, a), (_, b) in zip(m.state_dict().items(), m2.state_dict().items()):
        assert torch.equal(a, b), k
    tr = pytest.importorskip("transformers")
    ck.save_pretrained(str(tmp_path / "hf"), m, config={
        "architectures": ["BertForSequenceClassification"], "model_type": "bert", "vocab_size": 512,
        "hidden_size": 128, "num_hidden_layers": 2, "num_attention_heads": 2, "intermediate_size": 512,
        "max_position_embeddings": 128})
    hf = tr.BertForSequenceClassification.from_pretrained(str(tmp_path / "hf")).eval()
    ids = torch.randint(0, 512, (2, 12))
    with torch.no_grad():
        assert torch.allclose(m.eval()(ids), hf(input_ids=ids).logits, rtol=1e-4, atol=1e-5)
