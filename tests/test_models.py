import pytest
import torch

from databricks_distributed_deep_learning_b200.models import build_model, bert_tiny, resnet18, resnet50
from databricks_distributed_deep_learning_b200.models.bert import BertConfig, BertForSequenceClassification


def test_resnet50_matches_reference_inventory():
    # 25,557,032 params in 161 tensors; 53 convs, 53 BNs (§2.E.2)
    m = resnet50()
    assert sum(p.numel() for p in m.parameters()) == 25_557_032
    assert len(list(m.parameters())) == 161
    assert sum(isinstance(x, torch.nn.Conv2d) for x in m.modules()) == 53
    assert sum(isinstance(x, torch.nn.BatchNorm2d) for x in m.modules()) == 53
    assert sum(b.numel() for b in m.buffers()) == 53_173


def test_resnet_keys_and_numerics_match_torchvision():
    tv = pytest.importorskip("torchvision")
    for ours, theirs in ((resnet18(), tv.models.resnet18()), (resnet50(), tv.models.resnet50())):
        assert list(ours.state_dict().keys()) == list(theirs.state_dict().keys())
        theirs.load_state_dict(ours.state_dict())
        ours.eval(), theirs.eval()
        x = torch.randn(1, 3, 64, 64)
        with torch.no_grad():
            assert torch.allclose(ours(x), theirs(x), rtol=1e-5, atol=1e-4)


def test_resnet18_cifar_shape():
    m = resnet18(num_classes=10, small_input=True)
    assert m(torch.randn(2, 3, 32, 32)).shape == (2, 10)


def test_bert_sizes():
    assert sum(p.numel() for p in build_model("bert-base").parameters()) == 109_483_778  # 109.48 M
    cfg = BertConfig.large()
    assert (cfg.hidden_size, cfg.num_hidden_layers, cfg.num_attention_heads, cfg.intermediate_size) == (1024, 24, 16, 4096)


def test_bert_state_dict_is_hf_layout_and_roundtrips():
    m = bert_tiny()
    sd = m.state_dict()
    assert "bert.encoder.layer.0.attention.self.query.weight" in sd
    assert "bert.encoder.layer.0.attention.self.qkv_weight" not in sd
    m2 = bert_tiny()
    m2.load_state_dict(sd)
    for (k1, v1), (k2, v2) in zip(sd.items(), m2.state_dict().items()):
        assert k1 == k2 and torch.equal(v1, v2)


def test_bert_matches_huggingface():
    tr = pytest.importorskip("transformers")
    hcfg = tr.BertConfig(vocab_size=512, hidden_size=128, num_hidden_layers=2, num_attention_heads=2,
                         intermediate_size=512, max_position_embeddings=128)
    hf = tr.BertForSequenceClassification(hcfg).eval()
    ours = bert_tiny().eval()
    assert set(hf.state_dict()) == set(ours.state_dict())
    ours.load_state_dict(hf.state_dict())
    ids = torch.randint(0, 512, (3, 24))
    mask = torch.ones(3, 24, dtype=torch.long)
    mask[1, 15:] = 0
    with torch.no_grad():
        a = ours(ids, mask)
        b = hf(input_ids=ids, attention_mask=mask).logits
    assert torch.allclose(a, b, rtol=1e-4, atol=1e-5)


def test_bert_loss_and_backward():
    m = bert_tiny().train()
    ids = torch.randint(0, 512, (2, 16))
    loss, logits = m(ids, torch.ones_like(ids), labels=torch.tensor([0, 1]))
    loss.backward()
    assert logits.shape == (2, 2) and all(p.grad is not None for p in m.parameters())


def test_prepare_for_fused_keeps_bn_statistics_fp32():
    import torch

    from databricks_distributed_deep_learning_b200.models import prepare_for_fused, resnet18

    m = prepare_for_fused(resnet18(num_classes=10, small_input=True))
    assert m.conv1.weight.dtype == torch.bfloat16
    assert m.conv1.weight.is_contiguous(memory_format=torch.channels_last)
    assert m.bn1.running_mean.dtype == torch.float32 and m.bn1.running_var.dtype == torch.float32
    assert m.bn1.weight.dtype == torch.bfloat16
