"""Peer-memory collectives and the GPU DDP path (needs >= 2 GPUs: `gpurun --gpus 2`)."""
import pytest
import torch

import _gpu_workers as GW
from databricks_distributed_deep_learning_b200 import Distributor

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _world():
    return min(torch.cuda.device_count(), 8)


def test_allreduce_variants_match_sum():
    out = Distributor(num_processes=_world(), use_gpu=True, timeout_s=600).run(GW.w_allreduce_variants)
    assert out["ok"], out


def test_allgather_reducescatter_broadcast():
    out = Distributor(num_processes=_world(), use_gpu=True, timeout_s=600).run(GW.w_other_collectives)
    assert out["ok"], out


def test_ddp_p2p_matches_stock_ddp_nccl():
    out = Distributor(num_processes=_world(), use_gpu=True, timeout_s=900).run(GW.w_ddp_equivalence, 6)
    assert out["ok"], out


def test_trainer_graph_multi_gpu_ranks_stay_in_sync():
    out = Distributor(num_processes=_world(), use_gpu=True, timeout_s=900).run(GW.w_trainer_graph_sync, 5)
    assert out["ok"], out


@pytest.mark.parametrize("workload", ["resnet50", "bert-base"])
def test_trainer_20_steps_tracks_stock_ddp_nccl(workload):
    out = Distributor(num_processes=_world(), use_gpu=True, timeout_s=1200).run(GW.w_trainer_vs_stock, workload, 20)
    print({k: out[k] for k in ("mean_rel", "max_rel", "ranks_identical", "comm")})
    assert out["ok"], out


def test_fsdp_gpu_matches_ddp():
    out = Distributor(num_processes=_world(), use_gpu=True, timeout_s=900).run(GW.w_fsdp_gpu, 4)
    assert out["ok"], out


def test_allgather_fused_into_gemm():
    out = Distributor(num_processes=_world(), use_gpu=True, timeout_s=600).run(GW.w_allgather_gemm)
    print(out["report"])
    assert out["ok"], out


def test_flag_protocol_stress_with_random_skew():
    out = Distributor(num_processes=_world(), use_gpu=True, timeout_s=600).run(GW.w_allreduce_stress, 200)
    assert out["ok"], out
