"""Process launcher: one worker process per GPU (or per CPU rank in local mode).

Public surface mirrors the two Databricks launchers named in BASELINE.json:5
(``TorchDistributor(num_processes, local_mode, use_gpu).run(train_fn, *args)`` and
``HorovodRunner(np=N).run(main, **kwargs)``).  The mounted reference itself has no
launcher at all (every cell of notebooks/cv/onnx_experiments.py runs in
one driver process), so the behaviour here is specified by BASELINE.json, not by a
reference file.

Design (B200-first, single node, NVSwitch domain):

* one OS process per GPU, started with the ``spawn`` context so that no CUDA
  context is ever forked;
* env-style rendezvous on ``127.0.0.1`` (RANK / LOCAL_RANK / WORLD_SIZE /
  MASTER_ADDR / MASTER_PORT), the same contract ``torchrun`` uses, so a ``train_fn``
  runs unchanged under either launcher;
* the payload (function + arguments) travels as a cloudpickle blob, which lets
  closures and notebook-defined functions through;
* rank 0's return value comes back over a pipe; every rank reports failures with its
  traceback; the parent raises on the first failure, names the rank and kills the
  surviving siblings (SURVEY.md §4.2 "Fault" row, §5 "Failure detection").
"""
from __future__ import annotations

import multiprocessing as mp
import os
import pickle
import socket
import sys
import time
import traceback
from dataclasses import dataclass
from typing import Any, Callable, List, Optional

try:  # closures / lambdas / notebook cells
    import cloudpickle as _pickler
except Exception:  # pragma: no cover - cloudpickle is in the image
    _pickler = pickle


class DistributedError(RuntimeError):
    """Raised in the parent when a worker fails.  Carries rank and remote traceback."""

    def __init__(self, rank: int, exitcode: Optional[int], remote_traceback: str):
        self.rank = rank
        self.exitcode = exitcode
        self.remote_traceback = remote_traceback
        super().__init__(
            f"worker rank {rank} failed (exitcode={exitcode})\n"
            f"----- remote traceback (rank {rank}) -----\n{remote_traceback}"
        )


def find_free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@dataclass
class _WorkerSpec:
    rank: int
    world_size: int
    master_port: int
    use_gpu: bool
    backend: Optional[str]
    auto_init: bool
    payload: bytes
    timeout_s: float


def _worker_main(spec: _WorkerSpec, conn) -> None:
    """Entry point of every spawned rank (module-level so ``spawn`` can import it)."""
    os.environ["RANK"] = str(spec.rank)
    os.environ["LOCAL_RANK"] = str(spec.rank)
    os.environ["WORLD_SIZE"] = str(spec.world_size)
    os.environ["LOCAL_WORLD_SIZE"] = str(spec.world_size)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(spec.master_port)
    os.environ.setdefault("OMP_NUM_THREADS", str(max(1, (os.cpu_count() or 1) // spec.world_size)))
    status, value = "ok", None
    try:
        fn, args, kwargs = _pickler.loads(spec.payload)
        if spec.auto_init:
            from .parallel import process_group as pg

            pg.init_process_group(
                backend=spec.backend, use_gpu=spec.use_gpu, timeout_s=spec.timeout_s
            )
        result = fn(*args, **kwargs)
        if spec.rank == 0:
            value = _pickler.dumps(result)
        if spec.auto_init:
            from .parallel import process_group as pg

            pg.destroy_process_group()
    except BaseException:  # noqa: BLE001 - we forward everything to the parent
        status, value = "error", traceback.format_exc()
    try:
        conn.send((status, value))
        conn.close()
    finally:
        # hard exit on error: a rank stuck in a collective must not keep the job alive
        if status != "ok":
            sys.stdout.flush()
            sys.stderr.flush()
            os._exit(13)


class Distributor:
    """TorchDistributor-style launcher (BASELINE.json:5).

    Parameters
    ----------
    num_processes: number of ranks == number of GPUs used (or CPU ranks).
    local_mode:    kept for API compatibility; this framework is single-node
                   (multi-node is out of scope), so both values run
                   the ranks on this host.
    use_gpu:       ``None`` = use GPUs iff ``torch.cuda.is_available()``.  ``False`` is
                   the BASELINE.json:7 plumbing config (gloo on CPU).
    backend:       override the process-group backend ("gloo" / "nccl").  The gradient
                   path on GPU does not use it (hand-written P2P kernels do that work);
                   it is the bootstrap/barrier/object-exchange channel.
    auto_init:     initialise the process group + device before calling ``train_fn``.
    timeout_s:     wall-clock budget for the whole ``run``.
    """

    def __init__(
        self,
        num_processes: int = 1,
        local_mode: bool = True,
        use_gpu: Optional[bool] = None,
        backend: Optional[str] = None,
        auto_init: bool = True,
        timeout_s: float = 1800.0,
    ):
        if num_processes < 1:
            raise ValueError("num_processes must be >= 1")
        self.num_processes = int(num_processes)
        self.local_mode = bool(local_mode)
        self.use_gpu = use_gpu
        self.backend = backend
        self.auto_init = auto_init
        self.timeout_s = float(timeout_s)

    # -- public API ---------------------------------------------------------------
    def run(self, train_fn: Callable[..., Any], *args: Any, **kwargs: Any) -> Any:
        """Run ``train_fn(*args, **kwargs)`` on every rank; return rank 0's result."""
        use_gpu = self.use_gpu
        if use_gpu is None:
            import torch

            use_gpu = torch.cuda.is_available()
        if use_gpu:
            import torch

            if torch.cuda.device_count() < self.num_processes:
                raise RuntimeError(
                    f"num_processes={self.num_processes} but only "
                    f"{torch.cuda.device_count()} GPU(s) are visible"
                )
        payload = _pickler.dumps((train_fn, args, kwargs))
        port = find_free_port()
        ctx = mp.get_context("spawn")
        procs: List[mp.Process] = []
        conns = []
        for rank in range(self.num_processes):
            parent_conn, child_conn = ctx.Pipe(duplex=False)
            spec = _WorkerSpec(
                rank=rank,
                world_size=self.num_processes,
                master_port=port,
                use_gpu=bool(use_gpu),
                backend=self.backend,
                auto_init=self.auto_init,
                payload=payload,
                timeout_s=self.timeout_s,
            )
            p = ctx.Process(target=_worker_main, args=(spec, child_conn), daemon=False)
            p.start()
            child_conn.close()
            procs.append(p)
            conns.append(parent_conn)
        return self._supervise(procs, conns)

    # -- supervision --------------------------------------------------------------
    def _supervise(self, procs, conns) -> Any:
        deadline = time.monotonic() + self.timeout_s
        n = len(procs)
        done = [False] * n
        result_blob = None
        failure: Optional[DistributedError] = None
        try:
            while not all(done) and failure is None:
                progressed = False
                for r in range(n):
                    if done[r]:
                        continue
                    msg = None
                    was_alive = procs[r].is_alive()  # sample BEFORE polling: no lost message
                    if conns[r].poll(0):
                        try:
                            msg = conns[r].recv()
                        except (EOFError, OSError):
                            msg = None  # closed pipe polls readable; handled as a crash below
                    if msg is not None:
                        status, value = msg
                        done[r] = True
                        progressed = True
                        if status == "ok":
                            if r == 0:
                                result_blob = value
                        else:
                            procs[r].join(5)
                            failure = DistributedError(r, procs[r].exitcode, value)
                            break
                    elif not was_alive:
                        # died without reporting: killed, segfault, os._exit, OOM ...
                        # (it was already dead before the poll above, so nothing can still
                        # be in flight on the pipe)
                        procs[r].join(1)
                        done[r] = True
                        progressed = True
                        failure = DistributedError(
                            r,
                            procs[r].exitcode,
                            f"rank {r} exited with code {procs[r].exitcode} without "
                            "reporting (killed or crashed)",
                        )
                        break
                if failure is None and not progressed:
                    if time.monotonic() > deadline:
                        alive = [r for r in range(n) if not done[r]]
                        failure = DistributedError(
                            alive[0] if alive else -1,
                            None,
                            f"timeout after {self.timeout_s:.0f}s; ranks still running: {alive}",
                        )
                        break
                    time.sleep(0.02)
        finally:
            if failure is not None:
                for p in procs:
                    if p.is_alive():
                        p.terminate()
                for p in procs:
                    p.join(5)
                    if p.is_alive():
                        p.kill()
                        p.join(5)
            else:
                for p in procs:
                    p.join(30)
            for c in conns:
                c.close()
        if failure is not None:
            raise failure
        return None if result_blob is None else _pickler.loads(result_blob)


class TorchDistributor(Distributor):
    """Alias with the Databricks class name (``pyspark.ml.torch.distributor``)."""


class HorovodRunner:
    """HorovodRunner-style surface: ``HorovodRunner(np=N).run(main, **kwargs)``.

    ``np < 0`` means "run ``-np`` ranks on the driver" in Databricks; here every run is
    on this host, so ``abs(np)`` ranks are spawned.  Keyword arguments only, as in
    HorovodRunner.  Inside ``main`` use :mod:`..parallel.process_group` (``rank()``,
    ``world_size()``) where Horovod code would call ``hvd.rank()`` / ``hvd.size()``.
    """

    def __init__(self, np: int = 1, use_gpu: Optional[bool] = None, **distributor_kwargs: Any):
        self.np = int(np)
        self._d = Distributor(num_processes=max(1, abs(self.np)), use_gpu=use_gpu, **distributor_kwargs)

    def run(self, main: Callable[..., Any], **kwargs: Any) -> Any:
        return self._d.run(main, **kwargs)
