import time

import pytest

import _workers as W
from databricks_distributed_deep_learning_b200 import DistributedError, Distributor, HorovodRunner


def test_run_returns_rank0_result_and_env():
    out = Distributor(num_processes=2, use_gpu=False, timeout_s=120).run(W.w_rank_info, "hello")
    assert out["rank"] == 0 and out["world"] == 2 and out["env_rank"] == "0"
    assert out["gathered"] == [(0, "hello"), (1, "hello")]


def test_without_auto_init():
    assert Distributor(num_processes=2, use_gpu=False, auto_init=False, timeout_s=60).run(W.w_no_pg) == 20


def test_closure_payload():
    k = 5
    assert Distributor(num_processes=1, use_gpu=False, auto_init=False).run(lambda a: a * k, 3) == 15


def test_failure_names_rank_and_kills_siblings():
    t0 = time.time()
    with pytest.raises(DistributedError) as ei:
        Distributor(num_processes=2, use_gpu=False, timeout_s=120).run(W.w_fail_on_rank, 1)
    assert ei.value.rank == 1
    assert "boom from rank 1" in ei.value.remote_traceback
    assert time.time() - t0 < 25, "siblings were not killed promptly"


def test_crash_without_report_is_detected():
    t0 = time.time()
    with pytest.raises(DistributedError) as ei:
        Distributor(num_processes=2, use_gpu=False, timeout_s=120).run(W.w_die_on_rank, 0)
    assert ei.value.rank == 0 and ei.value.exitcode == 7
    assert time.time() - t0 < 25


def test_timeout():
    with pytest.raises(DistributedError) as ei:
        Distributor(num_processes=1, use_gpu=False, auto_init=False, timeout_s=1.0).run(time.sleep, 20)
    assert "timeout" in ei.value.remote_traceback


def test_horovod_runner_surface():
    assert HorovodRunner(np=-2, use_gpu=False, auto_init=False, timeout_s=60).run(W.w_no_pg) == 20


def test_bad_args():
    with pytest.raises(ValueError):
        Distributor(num_processes=0)
