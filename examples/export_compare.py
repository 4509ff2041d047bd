#!/usr/bin/env python
"""The reference notebook (notebooks/cv/onnx_experiments.py) as a script: export ResNet-50 to pickle / TorchScript
(/ ONNX when the onnx stack is installed), run every available runtime on one image, compare logits at the
reference's tolerance (rtol 1e-5, atol 1e-4), report latencies (reference protocol + warm N-sample protocol) and
file sizes.  On a B200 it adds the bf16 / folded-BN / CUDA-graph runner."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from databricks_distributed_deep_learning_b200.utils.export import InferenceComparison  # noqa: E402

if __name__ == "__main__":
    report = InferenceComparison(workdir=sys.argv[1] if len(sys.argv) > 1 else "/tmp/ddl_export").run()
    print(json.dumps(report, indent=1, default=str))
