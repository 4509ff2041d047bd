"""Part A parity: the reference's own checks (SURVEY.md §4.2 "Export parity")."""
import os

import numpy as np
import pytest
import torch

from databricks_distributed_deep_learning_b200.models import resnet18, resnet50
from databricks_distributed_deep_learning_b200.utils import export as ex


def test_numpy_softmax_is_global_like_reference():
    x = np.array([[1.0, 2.0], [3.0, 4.0]], dtype=np.float32)
    s = ex.numpy_softmax(x)
    assert abs(s.sum() - 1.0) < 1e-6  # one distribution over the whole array (ref:85-88)
    t = torch.softmax(torch.tensor(x).flatten(), 0).numpy()
    assert np.allclose(s.flatten(), t, atol=1e-6)


def test_preprocess_shape_and_normalisation():
    img = ex.sample_image(seed=1)
    x = ex.preprocess(img)
    assert x.shape == (1, 3, 224, 224) and x.dtype == torch.float32
    assert -3 < float(x.min()) and float(x.max()) < 3


def test_topk_and_categories():
    cats = ex.load_categories(path="/nonexistent", n=1000)
    assert len(cats) == 1000
    logits = np.zeros((1, 1000), np.float32)
    logits[0, 7] = 5.0
    top = ex.topk_labels(logits, cats)
    assert top[0][0] == "class_0007" and len(top) == 5 and top[0][1] > top[1][1]


def test_export_parity_pickle_torchscript_eager(tmp_path):
    torch.manual_seed(0)
    cmp = ex.InferenceComparison(model=resnet50(), workdir=str(tmp_path))
    rep = cmp.run(warmup=1, samples=2)
    assert rep["exported"]["pickle"] and rep["exported"]["torchscript"]
    for rt in ("torchscript", "pickle-reload"):
        assert rep["parity"][rt]["allclose_1e-5_1e-4"], rt  # the reference's tolerance (ref:144)
    sizes = rep["sizes"]
    assert sizes[cmp.paths["pickle"]] > 100_000_000 and sizes[cmp.paths["torchscript"]] > 100_000_000
    assert rep["latency_ms"]["pytorch-eager"]["samples"] == 2
    assert rep["top5"]["pytorch-eager"][0][0] == rep["top5"]["torchscript"][0][0]
    if not ex.onnx_available():
        assert rep["exported"]["onnx"] is None and "onnxruntime" in rep["skipped"]


def test_onnx_export_guard():
    if ex.onnx_available():
        pytest.skip("onnx stack present")
    with pytest.raises(RuntimeError, match="onnx"):
        ex.export_onnx(resnet18(), "/tmp/should_not_exist.onnx")


def test_bn_folding_is_exact_in_fp32():
    torch.manual_seed(0)
    m = resnet18(num_classes=10, small_input=True)
    m.train()
    with torch.no_grad():
        m(torch.randn(8, 3, 32, 32))  # move running stats off their init (the reference's Q1 quirk)
    m.eval()
    f = ex.fold_batchnorm(m)
    x = torch.randn(2, 3, 32, 32)
    with torch.no_grad():
        assert torch.allclose(m(x), f(x), rtol=1e-4, atol=1e-4)
    assert not any(isinstance(t, torch.nn.BatchNorm2d) for t in f.modules())
