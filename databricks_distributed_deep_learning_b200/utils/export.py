"""Export + single-image inference comparison (target layer T9).

This is the part of the framework that corresponds line by line to what the mounted
reference actually does (notebooks/cv/onnx_experiments.py, cited below as ``ref:N``):

==========================  =====================================================
reference cell              here
==========================  =====================================================
ref:19   pretrained load    :func:`load_resnet50` (random init offline; ``weights=`` path)
ref:22-26 label table       :func:`load_categories` (file if present, else synthetic)
ref:29-42 ONNX export       :func:`export_onnx` (import-guarded: no onnx stack offline)
ref:47-50 sample image      :func:`sample_image` (synthetic PIL image, or a path)
ref:58-66 preprocessing     :func:`preprocess` (Resize256/CenterCrop224/Normalize)
ref:69-72 device placement  :class:`InferenceComparison` keeps model+input together
                            (fixes the device mismatches SURVEY.md calls Q3/Q4)
ref:81-104 ORT run + top5   :class:`OnnxRuntimeRunner`, :func:`run_sample`
ref:85-88 numpy softmax     :func:`numpy_softmax` (global max, as the reference)
ref:120-140 OpenVINO        :class:`OpenVinoRunner` (import-guarded)
ref:142-144 parity check    :func:`outputs_close` (rtol 1e-5, atol 1e-4)
ref:159-184 eager top-5     :class:`EagerRunner`, :func:`topk_labels`
ref:198  pickle             :func:`export_pickle`
ref:209-215 TorchScript     :func:`export_torchscript`, :class:`TorchScriptRunner`
ref:194,202,219 ls -all     :func:`size_report`
==========================  =====================================================

Added: a proper latency protocol (warm-up + N samples, median/min) next to the
reference's one-sample/no-warm-up one, and a B200 runner
(:class:`B200Runner`) - NHWC bf16, batch-norm folded into the convolution weights (what
ONNX constant folding at ref:39 does inside the libraries), timed with CUDA events.
"""
from __future__ import annotations

import os
import time
from dataclasses import dataclass, field
from typing import Any, Callable, Dict, List, Optional, Sequence

import numpy as np
import torch
import torch.nn as nn

IMAGENET_MEAN = (0.485, 0.456, 0.406)  # ref:63
IMAGENET_STD = (0.229, 0.224, 0.225)
PARITY_RTOL, PARITY_ATOL = 1e-05, 1e-04  # ref:144


# --------------------------------------------------------------------------- assets
def load_resnet50(weights: Optional[str] = None, eval_mode: bool = False) -> nn.Module:
    """ResNet-50 as at ref:19.  No network here, so weights are random unless a
    ``state_dict`` / safetensors path is given.  Like the reference the module is returned
    in TRAIN mode unless ``eval_mode``."""
    from ..models.resnet import resnet50

    m = resnet50()
    if weights:
        if weights.endswith(".safetensors"):
            from .checkpoint import load_safetensors

            load_safetensors(weights, m)
        else:
            from .checkpoint import load_checkpoint

            load_checkpoint(weights, m)
    return m.eval() if eval_mode else m


def load_categories(path: str = "/tmp/imagenet_classes.txt", n: int = 1000) -> List[str]:
    """ref:22-26: one label per line; synthetic names when the file is not there."""
    if os.path.exists(path):
        with open(path) as f:
            cats = [s.strip() for s in f.readlines()]
        if cats:
            return cats
    return [f"class_{i:04d}" for i in range(n)]


def sample_image(path: Optional[str] = None, size=(640, 480), seed: int = 0):
    """ref:47-50 downloads dog.jpg; offline we synthesise a deterministic RGB image."""
    from PIL import Image

    if path and os.path.exists(path):
        return Image.open(path).convert("RGB")
    rng = np.random.default_rng(seed)
    w, h = size
    yy, xx = np.mgrid[0:h, 0:w]
    base = np.stack([(xx * 255 // w), (yy * 255 // h), ((xx + yy) * 255 // (w + h))], -1).astype(np.float32)
    noise = rng.normal(0, 12, size=(h, w, 3))
    return Image.fromarray(np.clip(base + noise, 0, 255).astype(np.uint8), "RGB")


def preprocess(image, resize: int = 256, crop: int = 224) -> torch.Tensor:
    """ref:58-66 / ref:161-170: Resize -> CenterCrop -> ToTensor -> Normalize -> unsqueeze."""
    from torchvision import transforms as T

    pipeline = T.Compose([
        T.Resize(resize),
        T.CenterCrop(crop),
        T.ToTensor(),
        T.Normalize(mean=list(IMAGENET_MEAN), std=list(IMAGENET_STD)),
    ])
    return pipeline(image).unsqueeze(0)


# --------------------------------------------------------------------------- post-processing
def numpy_softmax(x: np.ndarray) -> np.ndarray:
    """ref:85-88 - softmax over the whole array with one global max."""
    e_x = np.exp(x - np.max(x))
    return e_x / e_x.sum()


def topk_labels(logits, categories: Sequence[str], k: int = 5):
    """ref:96-100: flatten -> softmax -> argsort(-p)[:k]."""
    out = numpy_softmax(np.asarray(logits, dtype=np.float32).flatten())
    idx = np.argsort(-out)[:k]
    return [(categories[i] if i < len(categories) else str(i), float(out[i])) for i in idx]


def outputs_close(a, b, rtol: float = PARITY_RTOL, atol: float = PARITY_ATOL) -> bool:
    """ref:142-144 (there printed twice and never asserted)."""
    return bool(np.allclose(np.asarray(a), np.asarray(b), rtol=rtol, atol=atol))


# --------------------------------------------------------------------------- exporters
def export_pickle(model: nn.Module, path: str = "/tmp/model.bin") -> str:
    """ref:198 - pickles the module object itself."""
    torch.save(model, path)
    return path


def export_state_dict(model: nn.Module, path: str) -> str:
    torch.save(model.state_dict(), path)
    return path


def export_torchscript(model: nn.Module, example: torch.Tensor, path: str = "/tmp/traced_resnet_model.pt") -> str:
    """ref:209-215 - eval(), trace, save."""
    model.eval()
    traced = torch.jit.trace(model, example)
    traced.save(path)
    return path


def onnx_available() -> bool:
    try:
        import onnx  # noqa: F401
        import onnxscript  # noqa: F401
    except Exception:
        return False
    return True


def export_onnx(model: nn.Module, path: str = "/tmp/resnet50.onnx", image_size: int = 224,
                opset_version: int = 12, reproduce_train_mode_forward: bool = False,
                seed: Optional[int] = 0) -> str:
    """ref:29-42 with the same arguments (static shapes, opset 12, constant folding,
    names 'input'/'output').  ``reproduce_train_mode_forward`` runs the reference's
    pre-export forward in train mode, which perturbs BN statistics; the
    reference is unseeded (Q6), ``seed`` makes it repeatable."""
    if not onnx_available():
        raise RuntimeError("ONNX export needs the 'onnx' and 'onnxscript' packages, which are not "
                           "installed in this offline image (SURVEY.md §6)")
    if seed is not None:
        torch.manual_seed(seed)
    x = torch.randn(1, 3, image_size, image_size, requires_grad=True)
    if reproduce_train_mode_forward:
        model(x)
    torch.onnx.export(model, x, path, export_params=True, opset_version=opset_version,
                      do_constant_folding=True, input_names=["input"], output_names=["output"])
    return path


def size_report(paths: Sequence[str]) -> Dict[str, Optional[int]]:
    """ref:194,202,219 (`ls -all`): bytes per artifact, None if missing."""
    return {p: (os.path.getsize(p) if os.path.exists(p) else None) for p in paths}


# --------------------------------------------------------------------------- runners
class Runner:
    name = "abstract"

    def available(self) -> bool:
        return True

    def prepare(self) -> None:
        pass

    def __call__(self, x: torch.Tensor) -> np.ndarray:
        raise NotImplementedError


class EagerRunner(Runner):
    name = "pytorch-eager"

    def __init__(self, model: nn.Module, device: str = "cpu"):
        self.model, self.device = model, torch.device(device)

    def prepare(self):
        self.model.eval().to(self.device)

    def __call__(self, x):
        with torch.no_grad():  # ref:173
            return self.model(x.to(self.device)).float().cpu().numpy()


class TorchScriptRunner(Runner):
    name = "torchscript"

    def __init__(self, path: str):
        self.path, self.mod = path, None

    def available(self):
        return os.path.exists(self.path)

    def prepare(self):
        self.mod = torch.jit.load(self.path).eval()

    def __call__(self, x):
        with torch.no_grad():
            return self.mod(x.cpu()).numpy()


class PickleRunner(Runner):
    name = "pickle-reload"

    def __init__(self, path: str):
        self.path, self.mod = path, None

    def available(self):
        return os.path.exists(self.path)

    def prepare(self):
        self.mod = torch.load(self.path, weights_only=False).eval()

    def __call__(self, x):
        with torch.no_grad():
            return self.mod(x.cpu()).numpy()


class OnnxRuntimeRunner(Runner):
    """ref:77-94 (CPU execution provider)."""

    name = "onnxruntime"

    def __init__(self, path: str, providers: Sequence[str] = ("CPUExecutionProvider",)):
        self.path, self.providers, self.session = path, list(providers), None

    def available(self):
        try:
            import onnxruntime  # noqa: F401
        except Exception:
            return False
        return os.path.exists(self.path)

    def prepare(self):
        import onnxruntime

        self.session = onnxruntime.InferenceSession(self.path, providers=self.providers)

    def __call__(self, x):
        return self.session.run([], {"input": x.cpu().detach().numpy()})[0]


class OpenVinoRunner(Runner):
    """ref:120-137 (CPU plugin); ``.cpu()`` added so it also works on a GPU node (Q3)."""

    name = "openvino"

    def __init__(self, path: str):
        self.path, self.compiled, self.out = path, None, None

    def available(self):
        try:
            from openvino.runtime import Core  # noqa: F401
        except Exception:
            return False
        return os.path.exists(self.path)

    def prepare(self):
        from openvino.runtime import Core

        ie = Core()
        self.compiled = ie.compile_model(model=ie.read_model(model=self.path), device_name="CPU")
        self.out = next(iter(self.compiled.outputs))

    def __call__(self, x):
        req = self.compiled.create_infer_request()
        req.infer(inputs={"input": x.cpu().detach().numpy()})
        return np.array(req.get_output_tensor(self.out.index).data)


def fold_batchnorm(model: nn.Module) -> nn.Module:
    """Fold every eval-mode BatchNorm2d of a ResNet into the preceding convolution
    (the constant folding ref:39 asks the exporter for).  Returns a deep copy."""
    import copy

    from ..models.resnet import BasicBlock, Bottleneck, ResNet

    m = copy.deepcopy(model).eval()

    def fold(conv: nn.Conv2d, bn: nn.BatchNorm2d) -> nn.Conv2d:
        scale = bn.weight.detach() / torch.sqrt(bn.running_var + bn.eps)
        new = nn.Conv2d(conv.in_channels, conv.out_channels, conv.kernel_size, conv.stride,
                        conv.padding, bias=True)
        new.weight.data = conv.weight.detach() * scale.view(-1, 1, 1, 1)
        new.bias.data = bn.bias.detach() - bn.running_mean * scale
        return new

    class _Id(nn.Module):
        def __init__(self, c):
            super().__init__()
            self.weight = nn.Parameter(torch.ones(c))
            self.bias = nn.Parameter(torch.zeros(c))
            self.register_buffer("running_mean", torch.zeros(c))
            self.register_buffer("running_var", torch.ones(c))
            self.eps, self.momentum, self.training = 0.0, 0.0, False

        def forward(self, x):
            return x

    assert isinstance(m, ResNet)
    m.conv1, m.bn1 = fold(m.conv1, m.bn1), _Id(64)
    for blk in m.modules():
        if isinstance(blk, (BasicBlock, Bottleneck)):
            names = ["1", "2"] + (["3"] if isinstance(blk, Bottleneck) else [])
            for n in names:
                conv, bn = getattr(blk, "conv" + n), getattr(blk, "bn" + n)
                setattr(blk, "conv" + n, fold(conv, bn))
                setattr(blk, "bn" + n, _Id(conv.out_channels))
            if blk.downsample is not None:
                c, b = blk.downsample[0], blk.downsample[1]
                blk.downsample = nn.Sequential(fold(c, b), _Id(c.out_channels))
    return m.set_impl("torch")


class B200Runner(Runner):
    """Batch-1 inference on the B200: NHWC bf16, BatchNorm folded into the convolution weights (the constant folding
    ref:39 asks the exporter for), every layer on the native kernels (``ops/inference.py``: tcgen05 implicit-GEMM
    convolutions + one bias / residual / ReLU pass per layer), the whole forward replayed as one CUDA graph."""

    name = "b200-bf16-folded"

    def __init__(self, model: nn.Module, dtype: torch.dtype = torch.bfloat16, use_graph: bool = True):
        self.src, self.dtype, self.use_graph = model, dtype, use_graph
        self.graph = None

    def available(self):
        return torch.cuda.is_available()

    def prepare(self):
        from ..ops.inference import FoldedResNet

        self.net = FoldedResNet(fold_batchnorm(self.src), "cuda", self.dtype)
        self.static_in = torch.zeros(1, 3, 224, 224, device="cuda", dtype=torch.float32)
        with torch.no_grad():
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                for _ in range(3):
                    self.static_out = self.net(self.static_in)
            torch.cuda.current_stream().wait_stream(s)
            torch.cuda.synchronize()
            if self.use_graph:
                self.graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(self.graph):
                    self.static_out = self.net(self.static_in)

    def __call__(self, x):
        with torch.no_grad():
            self.static_in.copy_(x.to("cuda", torch.float32, non_blocking=True))
            if self.graph is not None:
                self.graph.replay()
            else:
                self.static_out = self.net(self.static_in)
            return self.static_out.float().cpu().numpy()


# --------------------------------------------------------------------------- timing
@dataclass
class LatencyResult:
    runtime: str
    reference_protocol_ms: float          # ref:92-95 / ref:133-138: 1 sample, no warm-up
    median_ms: Optional[float] = None     # warm-up + N samples
    min_ms: Optional[float] = None
    samples: int = 0
    top5: List[Any] = field(default_factory=list)
    logits: Optional[np.ndarray] = None


def run_sample(runner: Callable[[torch.Tensor], np.ndarray], inputs: torch.Tensor, latency: List[float]) -> np.ndarray:
    """ref:91-95: time one call with ``time.time()`` and append the seconds to ``latency``."""
    start = time.time()
    out = runner(inputs)
    latency.append(time.time() - start)
    return out


def measure(runner: Runner, x: torch.Tensor, categories: Sequence[str], warmup: int = 3, samples: int = 10) -> LatencyResult:
    runner.prepare()
    lat: List[float] = []
    logits = run_sample(runner, x, lat)  # the reference's protocol: first call, cold
    ref_ms = sum(lat) * 1000 / len(lat)  # ref:104
    for _ in range(warmup):
        runner(x)
    ts: List[float] = []
    for _ in range(samples):
        run_sample(runner, x, ts)
    ts_ms = sorted(t * 1000 for t in ts)
    return LatencyResult(
        runtime=runner.name, reference_protocol_ms=ref_ms,
        median_ms=ts_ms[len(ts_ms) // 2] if ts_ms else None, min_ms=ts_ms[0] if ts_ms else None,
        samples=len(ts_ms), top5=topk_labels(logits, categories), logits=logits,
    )


class InferenceComparison:
    """End-to-end restatement of the reference notebook as one object.

    >>> cmp = InferenceComparison(workdir="/tmp/ddl_export")
    >>> report = cmp.run()          # export -> run every available runtime -> parity -> sizes
    """

    def __init__(self, model: Optional[nn.Module] = None, workdir: str = "/tmp/ddl_export",
                 image_path: Optional[str] = None, seed: int = 0):
        torch.manual_seed(seed)
        self.model = model if model is not None else load_resnet50()
        self.workdir = workdir
        os.makedirs(workdir, exist_ok=True)
        self.categories = load_categories()
        self.input_batch = preprocess(sample_image(image_path, seed=seed))
        self.paths = {
            "onnx": os.path.join(workdir, "resnet50.onnx"),
            "pickle": os.path.join(workdir, "model.bin"),
            "torchscript": os.path.join(workdir, "traced_resnet_model.pt"),
        }

    def export_all(self) -> Dict[str, Optional[str]]:
        done: Dict[str, Optional[str]] = {}
        try:
            done["onnx"] = export_onnx(self.model, self.paths["onnx"])
        except RuntimeError as e:
            done["onnx"] = None
            self.onnx_skip_reason = str(e)
        self.model.eval()  # ref:55
        done["pickle"] = export_pickle(self.model, self.paths["pickle"])
        done["torchscript"] = export_torchscript(self.model, self.input_batch, self.paths["torchscript"])
        return done

    def runners(self) -> List[Runner]:
        return [
            OnnxRuntimeRunner(self.paths["onnx"]),
            OpenVinoRunner(self.paths["onnx"]),
            EagerRunner(self.model, "cpu"),
            TorchScriptRunner(self.paths["torchscript"]),
            PickleRunner(self.paths["pickle"]),
            B200Runner(self.model),
        ]

    def run(self, warmup: int = 3, samples: int = 10) -> Dict[str, Any]:
        exported = self.export_all()
        results: Dict[str, LatencyResult] = {}
        skipped: Dict[str, str] = {}
        for r in self.runners():
            if not r.available():
                skipped[r.name] = "runtime or artifact not available in this image"
                continue
            results[r.name] = measure(r, self.input_batch, self.categories, warmup, samples)
        oracle = results["pytorch-eager"].logits
        parity = {}
        for name, res in results.items():
            if name == "pytorch-eager":
                continue
            if name.startswith("b200"):  # bf16: compare decisions + a bf16-sized tolerance
                parity[name] = {
                    "top1_equal": int(np.argmax(res.logits)) == int(np.argmax(oracle)),
                    "allclose_bf16": outputs_close(res.logits, oracle, rtol=5e-2, atol=5e-2),
                }
            else:
                parity[name] = {"allclose_1e-5_1e-4": outputs_close(res.logits, oracle)}
        return {
            "exported": exported,
            "latency_ms": {k: {"reference_protocol": v.reference_protocol_ms, "median": v.median_ms,
                               "min": v.min_ms, "samples": v.samples} for k, v in results.items()},
            "top5": {k: v.top5 for k, v in results.items()},
            "parity": parity,
            "sizes": size_report(list(self.paths.values())),
            "skipped": skipped,
        }
