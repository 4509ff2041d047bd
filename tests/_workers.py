"""Module-level worker functions for the multi-process tests (importable by spawn)."""
import os
import time

import torch


def w_rank_info(tag):
    from databricks_distributed_deep_learning_b200.parallel import process_group as pg

    got = pg.all_gather_object((pg.rank(), tag))
    return {"rank": pg.rank(), "world": pg.world_size(), "gathered": got, "env_rank": os.environ["RANK"]}


def w_fail_on_rank(bad_rank):
    from databricks_distributed_deep_learning_b200.parallel import process_group as pg

    if pg.rank() == bad_rank:
        raise ValueError(f"boom from rank {pg.rank()}")
    time.sleep(30)  # siblings would hang here; the parent must kill them
    return "survived"


def w_die_on_rank(bad_rank):
    from databricks_distributed_deep_learning_b200.parallel import process_group as pg

    if pg.rank() == bad_rank:
        os._exit(7)  # simulated crash mid-step: no traceback, no report
    time.sleep(30)
    return "survived"


def w_no_pg():
    return int(os.environ["WORLD_SIZE"]) * 10 + int(os.environ["RANK"])


def _make_resnet18():
    from databricks_distributed_deep_learning_b200.models import resnet18

    torch.manual_seed(1234)
    return resnet18(num_classes=10, small_input=True)


def w_ddp_resnet18(steps, per_rank_batch, bucket_cap_mb, use_no_sync=False):
    """BASELINE.json:7 - ResNet-18, 32x32 synthetic, DDP over gloo, world_size=2."""
    import torch.nn.functional as F

    from databricks_distributed_deep_learning_b200 import DDP
    from databricks_distributed_deep_learning_b200.parallel import process_group as pg
    from databricks_distributed_deep_learning_b200.utils.data import cifar_shape_loader

    torch.set_num_threads(2)
    rank, world = pg.rank(), pg.world_size()
    model = _make_resnet18()
    if rank != 0:  # perturb: the init broadcast must repair this
        with torch.no_grad():
            for p in model.parameters():
                p.add_(1.0)
    # BN in eval mode so that the 2-rank run is exactly comparable to one process with the
    # concatenated batch (batch statistics would differ by construction)
    model.eval()
    ddp = DDP(model, bucket_cap_mb=bucket_cap_mb)
    opt = torch.optim.SGD(ddp.parameters(), lr=0.05)
    loader = cifar_shape_loader(per_rank_batch, rank=rank, world_size=world, seed=3, steps=steps, pin_memory=False)
    losses = []
    for x, y in loader:
        ddp.zero_grad()
        if use_no_sync:
            h = per_rank_batch // 2
            with ddp.no_sync():
                (F.cross_entropy(ddp(x[:h]), y[:h]) * 0.5).backward()
            (F.cross_entropy(ddp(x[h:]), y[h:]) * 0.5).backward()
            losses.append(0.0)
        else:
            loss = F.cross_entropy(ddp(x), y)
            loss.backward()
            losses.append(float(loss.detach()))
        opt.step()
    flat = torch.cat([p.detach().reshape(-1) for p in ddp.module.parameters()])
    gathered = pg.all_gather_object(flat)
    keys = list(ddp.state_dict().keys())
    return {"params": gathered, "losses": losses, "keys": keys, "nbuckets": len(ddp.buckets)}


def single_process_resnet18(steps, per_rank_batch, world):
    import torch.nn.functional as F

    from databricks_distributed_deep_learning_b200.utils.data import cifar_shape_loader

    torch.set_num_threads(2)
    model = _make_resnet18().eval()
    opt = torch.optim.SGD(model.parameters(), lr=0.05)
    loaders = [iter(cifar_shape_loader(per_rank_batch, rank=r, world_size=world, seed=3, steps=steps, pin_memory=False))
               for r in range(world)]
    for _ in range(steps):
        batches = [next(l) for l in loaders]
        x = torch.cat([b[0] for b in batches])
        y = torch.cat([b[1] for b in batches])
        opt.zero_grad()
        F.cross_entropy(model(x), y).backward()
        opt.step()
    return torch.cat([p.detach().reshape(-1) for p in model.parameters()])


def w_fsdp_bert_tiny(steps, reshard):
    """FSDP (ZeRO-3) over gloo, world_size=2: ranks stay consistent and match a 1-process AdamW run."""
    from databricks_distributed_deep_learning_b200.models import bert_tiny
    from databricks_distributed_deep_learning_b200.parallel import process_group as pg
    from databricks_distributed_deep_learning_b200.parallel.fsdp import FullyShardedDataParallel
    from databricks_distributed_deep_learning_b200.utils.data import SyntheticTokenLoader

    torch.set_num_threads(2)
    rank, world = pg.rank(), pg.world_size()
    torch.manual_seed(7)
    model = bert_tiny(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
    if rank != 0:
        with torch.no_grad():
            for p in model.parameters():
                p.add_(0.5)  # must be repaired by the rank-0 broadcast at wrap time
    fsdp = FullyShardedDataParallel(model, lr=1e-3, weight_decay=0.01, reshard_after_forward=reshard)
    loader = SyntheticTokenLoader(4, seq_len=16, vocab_size=512, rank=rank, world_size=world, seed=2, steps=steps, pin_memory=False, pool=steps)
    losses = []
    for batch in loader:
        fsdp.zero_grad()
        loss, _ = fsdp(**batch)
        loss.backward()
        fsdp.step()
        losses.append(float(loss.detach()))
    sd = fsdp.full_state_dict()
    keys = sorted(sd)
    flat = torch.cat([sd[k].reshape(-1).float() for k in keys])
    return {"flat": pg.all_gather_object(flat), "keys": keys, "units": len(fsdp.units), "losses": losses}


def single_process_bert_tiny(steps, world):
    from databricks_distributed_deep_learning_b200.models import bert_tiny
    from databricks_distributed_deep_learning_b200.utils.data import SyntheticTokenLoader

    torch.set_num_threads(2)
    torch.manual_seed(7)
    model = bert_tiny(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
    opt = torch.optim.AdamW(model.parameters(), lr=1e-3, weight_decay=0.01)
    loaders = [iter(SyntheticTokenLoader(4, seq_len=16, vocab_size=512, rank=r, world_size=world, seed=2, steps=steps, pin_memory=False, pool=steps))
               for r in range(world)]
    for _ in range(steps):
        bs = [next(l) for l in loaders]
        batch = {k: torch.cat([b[k] for b in bs]) for k in bs[0]}
        opt.zero_grad()
        loss, _ = model(**batch)
        loss.backward()
        opt.step()
    sd = model.state_dict()
    return torch.cat([sd[k].reshape(-1).float() for k in sorted(sd)])


def w_fsdp_resume(reshard):
    """FSDP over gloo, world 2: (a) save after 2 steps, load into a fresh wrapper, 2 more steps == 4 uninterrupted steps;
    (b) load_full_state_dict reproduces the loaded weights on every unit, also with shared full-parameter buffers."""
    import os
    import tempfile

    from databricks_distributed_deep_learning_b200.models import bert_tiny
    from databricks_distributed_deep_learning_b200.parallel import process_group as pg
    from databricks_distributed_deep_learning_b200.parallel.fsdp import FullyShardedDataParallel
    from databricks_distributed_deep_learning_b200.utils import checkpoint as ck
    from databricks_distributed_deep_learning_b200.utils.data import SyntheticTokenLoader

    torch.set_num_threads(2)
    rank, world = pg.rank(), pg.world_size()

    def make(seed):
        torch.manual_seed(seed)
        return FullyShardedDataParallel(bert_tiny(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0), lr=1e-3,
                                        weight_decay=0.01, reshard_after_forward=reshard)

    def run(f, batches):
        for batch in batches:
            f.zero_grad()
            f(**batch)[0].backward()
            f.step()

    batches = list(SyntheticTokenLoader(4, seq_len=16, vocab_size=512, rank=rank, world_size=world, seed=2, steps=4, pin_memory=False, pool=4))
    ref = make(7)
    run(ref, batches)
    a = make(7)
    run(a, batches[:2])
    path = pg.broadcast_object(os.path.join(tempfile.mkdtemp(), "fsdp.pt") if rank == 0 else None)
    ck.save_checkpoint(path, a, a, step=2)
    b = make(11)
    ck.load_checkpoint(path, b, b)
    run(b, batches[2:])
    sr, sb = ref.full_state_dict(), b.full_state_dict()
    resume_err = max(float((sr[k].float() - sb[k].float()).abs().max()) for k in sr)
    c = make(13)
    c.load_full_state_dict(sr)
    sc = c.full_state_dict()
    load_err = max(float((sr[k].float() - sc[k].float()).abs().max()) for k in sr)
    errs = pg.all_gather_object((resume_err, load_err))
    return {"resume_err": max(e[0] for e in errs), "load_err": max(e[1] for e in errs)}


def w_profile_comm(steps):
    """Trainer.profile_comm on the gloo path: per-bucket all-reduce records, and the training state is rolled back."""
    from databricks_distributed_deep_learning_b200.engine import Trainer
    from databricks_distributed_deep_learning_b200.parallel import process_group as pg

    torch.set_num_threads(2)
    torch.manual_seed(0)
    tr = Trainer(_make_resnet18(), lr=1e-3, device=torch.device("cpu"), bucket_cap_mb=1.0)
    x = torch.randn(4, 3, 32, 32)
    y = torch.randint(0, 10, (4,))
    before = torch.cat([p.detach().flatten() for p in tr.model.parameters()]).clone()
    out = tr.profile_comm((x, y), steps=steps)
    after = torch.cat([p.detach().flatten() for p in tr.model.parameters()])
    loss = float(tr.step((x, y)))
    return {"stats": out, "rolled_back": bool(torch.equal(before, after)), "nbuckets": len(tr.ddp.buckets), "loss": loss,
            "rank": pg.rank()}


def w_fit_with_logging(path):
    """Trainer.fit with a JSONL logger: one comm record (per-bucket all-reduce metrics) and step records with throughput."""
    from databricks_distributed_deep_learning_b200.engine import Trainer
    from databricks_distributed_deep_learning_b200.utils.data import cifar_shape_loader
    from databricks_distributed_deep_learning_b200.utils.timing import JsonlLogger
    from databricks_distributed_deep_learning_b200.parallel import process_group as pg

    torch.set_num_threads(2)
    torch.manual_seed(0)
    tr = Trainer(_make_resnet18(), lr=1e-3, device=torch.device("cpu"), bucket_cap_mb=1.0)
    loader = cifar_shape_loader(4, rank=pg.rank(), world_size=pg.world_size(), seed=1, steps=4, pin_memory=False)
    out = tr.fit(loader, log_every=2, logger=JsonlLogger(path), log_comm=True)
    return {"steps": out["steps"], "nlosses": len(out["losses"])}
