"""Worker functions for the multi-GPU tests (one process per GPU)."""
import torch


def _pg():
    from databricks_distributed_deep_learning_b200.parallel import process_group as pg

    return pg


def w_allreduce_variants():
    from databricks_distributed_deep_learning_b200.parallel.collectives import PeerMemoryComm

    pg = _pg()
    dev = pg.device()
    rank, world = pg.rank(), pg.world_size()
    comm = PeerMemoryComm(dev)
    report = {"ok": True, "cases": []}
    for dtype in (torch.bfloat16, torch.float32):
        for n in (1024, 8 * 1024 * 3, 1 << 20, (1 << 22) + 8192):
            n = n // (8 * world) * (8 * world)
            buf = comm.alloc_flat(n, dtype, dev)
            for algo in ("oneshot", "twoshot") + (("nvls",) if comm.heap.multicast_supported else ()):
                if algo == "oneshot" and n * buf.element_size() > 2 << 20:
                    continue
                for rep in range(3):  # back-to-back launches exercise epoch reuse
                    g = torch.Generator(device="cpu").manual_seed(100 + rep)
                    full = torch.randn(world, n, generator=g)
                    buf.copy_(full[rank].to(dtype))
                    pg.barrier()
                    comm.all_reduce_mean_(buf, algo=algo)
                    torch.cuda.synchronize()
                    ref = full.to(dtype).float().mean(0)
                    err = (buf.float().cpu() - ref).abs().max().item()
                    tol = 2e-2 if dtype == torch.bfloat16 else 1e-5
                    ok = err < tol
                    report["cases"].append((str(dtype), n, algo, rep, err))
                    report["ok"] &= ok
                    pg.barrier()
        # fused cast: fp32 in -> bf16 out (out of place)
    n = 8 * world * 4096
    src = comm.alloc_flat(n, torch.float32, dev)
    dst = comm.alloc_flat(n, torch.bfloat16, dev)
    report["nvls_tested"] = bool(comm.heap.multicast_supported)
    if comm.heap.multicast_supported and world in (2, 4, 8):
        # hybrid (switch reduction + peer loads / stores in one kernel) against the fp32 mean of the bf16 inputs
        for n_h, permille in (((1 << 20) + 8192, 600), (8 * world * 64, 500), (1 << 22, 0), (1 << 21, 1000)):
            n_h = n_h // (8 * world) * (8 * world)
            hb = comm.alloc_flat(n_h, torch.bfloat16, dev)
            for rep in range(2):
                g = torch.Generator(device="cpu").manual_seed(300 + rep)
                full = torch.randn(world, n_h, generator=g)
                hb.copy_(full[rank].to(torch.bfloat16))
                pg.barrier()
                comm.all_reduce_hybrid_(hb, permille, 4, 8)
                torch.cuda.synchronize()
                err = (hb.float().cpu() - full.to(torch.bfloat16).float().mean(0)).abs().max().item()
                report["cases"].append(("hybrid", n_h, permille, rep, err))
                report["ok"] &= err < 2e-2
                pg.barrier()
    for algo in ("oneshot", "twoshot"):
        g = torch.Generator(device="cpu").manual_seed(7)
        full = torch.randn(world, n, generator=g)
        src.copy_(full[rank])
        dst.zero_()
        pg.barrier()
        comm.all_reduce_mean_(src, out=dst, algo=algo)
        torch.cuda.synchronize()
        err = (dst.float().cpu() - full.mean(0)).abs().max().item()
        report["cases"].append(("f32->bf16", n, algo, 0, err))
        report["ok"] &= err < 2e-2
        pg.barrier()
    gathered = pg.all_gather_object(report["ok"])
    report["ok"] = all(gathered)
    return report


def w_other_collectives():
    from databricks_distributed_deep_learning_b200.parallel.collectives import PeerMemoryComm

    pg = _pg()
    dev = pg.device()
    rank, world = pg.rank(), pg.world_size()
    comm = PeerMemoryComm(dev)
    ok = True
    notes = []
    n = 8 * 5000
    shard = comm.alloc_flat(n, torch.bfloat16, dev)
    shard.copy_(torch.full((n,), float(rank + 1)).to(torch.bfloat16))
    full = torch.empty(world * n, dtype=torch.bfloat16, device=dev)
    pg.barrier()
    comm.all_gather_into(full, shard)
    torch.cuda.synchronize()
    for r in range(world):
        ok &= bool((full[r * n:(r + 1) * n].float() == r + 1).all())
    notes.append(("allgather", ok))
    fullg = comm.alloc_flat(world * n, torch.float32, dev)
    g = torch.Generator(device="cpu").manual_seed(5)
    data = torch.randn(world, world * n, generator=g)
    fullg.copy_(data[rank])
    out = torch.empty(n, dtype=torch.float32, device=dev)
    pg.barrier()
    comm.reduce_scatter_mean_into(out, fullg)
    torch.cuda.synchronize()
    ref = data.mean(0)[rank * n:(rank + 1) * n]
    ok &= bool((out.cpu() - ref).abs().max() < 1e-5)
    notes.append(("reducescatter", ok))
    # fused reduce-scatter + AdamW vs unfused
    master = torch.randn(n, device=dev)
    m = torch.zeros(n, device=dev)
    v = torch.zeros(n, device=dev)
    pout = torch.empty(n, dtype=torch.bfloat16, device=dev)
    hyper = torch.tensor([1e-2, 0.9, 0.999, 1e-8, 0.01, 1.0, 0, 0], device=dev)
    ref_p = torch.nn.Parameter(master.clone())
    opt = torch.optim.AdamW([ref_p], lr=1e-2, weight_decay=0.01)
    pg.barrier()
    comm.reduce_scatter_adamw_(fullg, master, m, v, pout, hyper)
    torch.cuda.synchronize()
    ref_p.grad = out.clone()
    opt.step()
    ok &= bool(torch.allclose(master, ref_p.data, rtol=1e-5, atol=1e-5))
    ok &= bool(torch.allclose(pout.float(), ref_p.data, rtol=1e-2, atol=1e-2))
    notes.append(("rs_adamw", ok))
    # broadcast (symmetric and staged)
    b = comm.alloc_flat(4096, torch.float32, dev)
    b.fill_(float(rank))
    pg.barrier()
    comm.broadcast_(b, src=0)
    torch.cuda.synchronize()
    ok &= bool((b == 0).all())
    t = torch.full((1000,), float(rank + 3), device=dev)
    comm.broadcast_(t, src=0)
    torch.cuda.synchronize()
    ok &= bool((t == 3).all())
    notes.append(("broadcast", ok))
    comm.barrier()
    torch.cuda.synchronize()
    return {"ok": all(pg.all_gather_object(ok)), "notes": notes}


def _make_model():
    from databricks_distributed_deep_learning_b200.models import resnet18

    torch.manual_seed(11)
    return resnet18(num_classes=10, small_input=True)


def w_ddp_equivalence(steps):
    """Our wrapper (peer-memory all-reduce) vs stock DDP + NCCL: same parameters after K steps."""
    import torch.nn.functional as F

    from databricks_distributed_deep_learning_b200 import DDP
    from databricks_distributed_deep_learning_b200.utils.data import cifar_shape_loader

    pg = _pg()
    dev = pg.device()
    rank, world = pg.rank(), pg.world_size()
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cudnn.deterministic = True
    # BN in eval mode: batch statistics make a tiny-batch fp32 run chaotic (two stock runs already
    # differ by ~1e-2 after a few steps), which would hide what this test is about
    ours = DDP(_make_model().to(dev).eval(), bucket_cap_mb=1.0, comm="p2p")
    stock = torch.nn.parallel.DistributedDataParallel(_make_model().to(dev).eval(), device_ids=[dev.index], bucket_cap_mb=1.0)
    o1 = torch.optim.SGD(ours.parameters(), lr=0.05)
    o2 = torch.optim.SGD(stock.parameters(), lr=0.05)
    loader = cifar_shape_loader(16, rank=rank, world_size=world, seed=5, steps=steps, pin_memory=False)
    maxdiff = 0.0
    first_grad_diff = None
    for x, y in loader:
        x, y = x.to(dev), y.to(dev)
        ours.zero_grad()
        F.cross_entropy(ours(x), y).backward()
        o2.zero_grad()
        F.cross_entropy(stock(x), y).backward()
        if first_grad_diff is None:
            first_grad_diff = max((a.grad - b.grad).abs().max().item() for a, b in zip(ours.module.parameters(), stock.module.parameters()))
        o1.step()
        o2.step()
    for a, b in zip(ours.module.parameters(), stock.module.parameters()):
        maxdiff = max(maxdiff, (a - b).abs().max().item())
    flat = torch.cat([p.detach().reshape(-1) for p in ours.module.parameters()]).cpu()
    same = all(torch.equal(flat, o) for o in pg.all_gather_object(flat))
    ok = maxdiff < 2e-3 and same and first_grad_diff < 1e-4
    return {"ok": all(pg.all_gather_object(ok)), "maxdiff": maxdiff, "first_step_grad_diff": first_grad_diff,
            "ranks_identical": same, "comm": ours.comm.name}


def w_trainer_graph_sync(steps):
    from databricks_distributed_deep_learning_b200.engine import Trainer
    from databricks_distributed_deep_learning_b200.utils.data import cifar_shape_loader

    pg = _pg()
    dev = pg.device()
    rank, world = pg.rank(), pg.world_size()
    tr = Trainer(_make_model(), lr=1e-3, device=dev, use_graph=True, bucket_cap_mb=1.0)
    loader = cifar_shape_loader(16, rank=rank, world_size=world, seed=9, steps=steps)
    out = tr.fit(loader)
    flat = torch.cat([p.detach().float().reshape(-1) for p in tr.model.parameters()]).cpu()
    same = all(torch.equal(flat, o) for o in pg.all_gather_object(flat))
    finite = all(l == l for l in out["losses"])
    return {"ok": same and finite, "losses": out["losses"], "comm": tr.ddp.comm.name}


def w_fsdp_gpu(steps):
    """FSDP Trainer (peer-memory all-gather + fused reduce-scatter/AdamW) vs the DDP Trainer."""
    from databricks_distributed_deep_learning_b200.engine import Trainer
    from databricks_distributed_deep_learning_b200.models import bert_tiny
    from databricks_distributed_deep_learning_b200.utils.data import SyntheticTokenLoader

    pg = _pg()
    dev = pg.device()
    rank, world = pg.rank(), pg.world_size()
    results = {}
    for mode, kw in (("ddp", {}), ("fsdp", {}), ("fsdp_reshard", {"reshard_after_forward": True})):
        torch.manual_seed(3)
        model = bert_tiny(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
        tr = Trainer(model, lr=1e-3, device=dev, use_graph=True, parallelism=mode.split("_")[0], bucket_cap_mb=1.0, **kw)
        loader = SyntheticTokenLoader(8, seq_len=128, vocab_size=512, rank=rank, world_size=world, seed=4, steps=steps, pool=steps)
        out = tr.fit(loader)
        sd = tr.ddp.full_state_dict() if mode.startswith("fsdp") else {k: v.detach().cpu() for k, v in tr.model.state_dict().items()}
        flat = torch.cat([sd[k].reshape(-1).float() for k in sorted(sd)])
        same = all(torch.equal(flat, o) for o in pg.all_gather_object(flat))
        results[mode] = {"losses": out["losses"], "flat": flat, "same": same}
    ok = all(r["same"] for r in results.values())
    ref = results["ddp"]["flat"]
    diffs = {m: float((results[m]["flat"] - ref).abs().max()) for m in ("fsdp", "fsdp_reshard")}
    # bf16 training: identical math, different reduction order/rounding points -> small drift
    ok = ok and all(d < 2e-2 for d in diffs.values())
    ok = ok and all(l == l for r in results.values() for l in r["losses"])
    return {"ok": all(pg.all_gather_object(ok)), "diffs": diffs, "losses": {m: r["losses"] for m, r in results.items()}}


def w_allgather_gemm():
    """K6: fused all-gather + GEMM vs (all-gather kernel, then GEMM) vs fp32 reference; also timed."""
    from databricks_distributed_deep_learning_b200.ops import gemm as G
    from databricks_distributed_deep_learning_b200.parallel.collectives import PeerMemoryComm

    pg = _pg()
    dev = pg.device()
    rank, world = pg.rank(), pg.world_size()
    comm = PeerMemoryComm(dev)
    ok, report = True, {}
    for (M, N, K) in ((512, 1024, 1024), (8192, 4096, 1024), (300, 2048, 768)):
        N = N // (world * 128) * (world * 128)
        g = torch.Generator(device="cpu").manual_seed(11)
        w_full = (torch.randn(N, K, generator=g) * 0.05).to(torch.bfloat16)
        x = torch.randn(M, K, generator=torch.Generator().manual_seed(100 + rank)).to(torch.bfloat16).to(dev)
        bias = torch.randn(N, generator=g).to(torch.bfloat16).to(dev)
        rows = N // world
        shard = comm.alloc_flat(rows * K, torch.bfloat16, dev).view(rows, K)
        shard.copy_(w_full[rank * rows:(rank + 1) * rows])
        layer = G.GatheredLinear(comm, shard, N)
        torch.cuda.synchronize()
        pg.barrier()
        for rep in range(3):  # back-to-back launches: flag reset + epoch reuse
            y = layer(x, bias=bias)
        torch.cuda.synchronize()
        ref = x.float() @ w_full.float().to(dev).t() + bias.float()
        err = ((y.float() - ref).abs().max() / ref.abs().max()).item()
        same_w = bool(torch.equal(layer.full_w.cpu(), w_full))
        ok &= err < 2e-2 and same_w
        # timing: fused vs all-gather kernel followed by the GEMM
        full = torch.empty(N, K, dtype=torch.bfloat16, device=dev)

        def unfused():
            comm.all_gather_into(full.view(-1), shard.view(-1), num_blocks=64)
            return G.gemm_nt(x, full, M, N, K, bias=bias)

        def timeit(fn):
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            pg.barrier()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            for _ in range(10):
                fn()
            e.record()
            torch.cuda.synchronize()
            return pg.max_over_ranks(s.elapsed_time(e) / 10)

        t_f, t_u = timeit(lambda: layer(x, bias=bias)), timeit(unfused)
        report[f"{M}x{N}x{K}"] = {"rel_err": err, "gathered_w_exact": same_w, "fused_us": t_f * 1e3, "unfused_us": t_u * 1e3}
    return {"ok": all(pg.all_gather_object(ok)), "report": report}


def w_allreduce_stress(iters):
    """Flag-protocol stress (SURVEY.md 5 "Race detection"; compute-sanitizer is closed on this pool): random
    per-rank delays in front of back-to-back collectives of random size / algorithm on ONE buffer, every
    result checked.  A lost or early flag shows up as a wrong sum."""
    import random

    from databricks_distributed_deep_learning_b200.parallel.collectives import PeerMemoryComm

    pg = _pg()
    dev = pg.device()
    rank, world = pg.rank(), pg.world_size()
    comm = PeerMemoryComm(dev)
    nmax = 1 << 21
    buf = comm.alloc_flat(nmax, torch.float32, dev)
    rng = random.Random(1234)          # same sequence on every rank
    delay_rng = random.Random(99 + rank)  # different delays per rank
    algos = ["oneshot", "twoshot"] + (["nvls"] if comm.heap.multicast_supported else [])
    bad = 0
    # per-element random data: every rank can rebuild every other rank's contribution (same generator), so a wrong lane,
    # a mis-sliced granule or a stale read shows up against the fp32 sum computed from the W inputs on this GPU
    gen = torch.Generator(device=dev).manual_seed(4242)
    pool = torch.randn(world, nmax, device=dev, generator=gen)  # identical on every rank
    for it in range(iters):
        n = rng.choice([8 * world, 1024, 8192, 1 << 16, 1 << 18, nmax]) // (8 * world) * (8 * world)
        algo = rng.choice(algos)
        if algo == "oneshot" and n * 4 > (1 << 20):
            algo = "twoshot"
        blocks = rng.choice([1, 4, 16, 64])
        view = buf[:n]
        shift = (it * 977) % (nmax - n + 1) // 8 * 8
        view.copy_(pool[rank, shift:shift + n])
        if delay_rng.random() < 0.5:
            torch.cuda._sleep(int(delay_rng.random() * 2_000_000))  # up to ~1 ms of skew
        comm.all_reduce_mean_(view, algo=algo, num_blocks=blocks, scale=1.0)
        if it % 8 == 7 or it == iters - 1:
            expect = pool[:, shift:shift + n].sum(0)
            # fp32 sums in a different order: a few ulps; anything structural is O(1)
            bad += int(((view - expect).abs() > 1e-4 * (1 + expect.abs())).any().item())
        if it % 8 == 7:
            comm.barrier()
    torch.cuda.synchronize()
    return {"ok": all(pg.all_gather_object(bad == 0)), "bad": bad, "algos": algos}


def w_trainer_vs_stock(workload, steps):
    """SURVEY.md section 4.2 "DDP equivalence (GPU)": the Trainer (fused kernels, whole-step CUDA graph, bf16, peer-memory
    gradient plane) against stock DistributedDataParallel + NCCL + autocast on the same batches from the same initial
    weights: loss curves within a bf16 tolerance, replicas bit-identical."""
    import torch.nn.functional as F

    from databricks_distributed_deep_learning_b200.engine import Trainer
    from databricks_distributed_deep_learning_b200.models import build_model
    from databricks_distributed_deep_learning_b200.utils.data import SyntheticTokenLoader, imagenet_shape_loader

    pg = _pg()
    dev = pg.device()
    rank, world = pg.rank(), pg.world_size()
    torch.manual_seed(21)
    if workload == "resnet50":
        import torchvision

        ours_model = build_model("resnet50")
        stock_model = torchvision.models.resnet50()
        stock_model.load_state_dict(ours_model.state_dict())
        stock_model = stock_model.to(dev).to(memory_format=torch.channels_last)
        batches = list(imagenet_shape_loader(16, rank=rank, world_size=world, seed=6, steps=steps, pool=steps))
        lr = 1e-4
    else:
        import transformers

        ours_model = build_model("bert-base", hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
        cfg = transformers.BertConfig(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, num_labels=2)
        stock_model = transformers.BertForSequenceClassification(cfg)
        stock_model.load_state_dict(ours_model.state_dict())
        stock_model = stock_model.to(dev)
        batches = list(SyntheticTokenLoader(8, seq_len=128, rank=rank, world_size=world, seed=6, steps=steps, pool=steps))
        lr = 2e-5
    stock_model.train()
    stock = torch.nn.parallel.DistributedDataParallel(stock_model, device_ids=[dev.index]) if world > 1 else stock_model
    sopt = torch.optim.AdamW(stock.parameters(), lr=lr, weight_decay=0.01)
    tr = Trainer(ours_model, lr=lr, weight_decay=0.01, device=dev, use_graph=True)
    ours_losses, stock_losses = [], []
    for b in batches:
        tr.step(b)
        sopt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            if workload == "resnet50":
                x, y = b
                loss = F.cross_entropy(stock(x.to(dev).contiguous(memory_format=torch.channels_last)), y.to(dev))
            else:
                loss = stock(**{k: v.to(dev) for k, v in b.items()}).loss
        loss.backward()
        sopt.step()
        stock_losses.append(float(loss))
    ours_losses = tr.loss_history(steps)
    flat = torch.cat([p.detach().float().reshape(-1) for p in tr.model.parameters()]).cpu()
    same = all(torch.equal(flat, o) for o in pg.all_gather_object(flat))
    rel = [abs(a - b) / max(abs(b), 1e-3) for a, b in zip(ours_losses, stock_losses)]
    finite = all(l == l and abs(l) < 1e4 for l in ours_losses)
    # Both arms compute in bf16, but ours also STORES the compute weights in bf16 (fp32 masters behind them) while stock
    # autocasts fp32 weights.  The first step sees identical weights, so its loss must agree tightly.  Later steps see
    # weights that moved in bf16-sized increments on our side: with ResNet-50 at lr 1e-4 the curves stay within a fraction of
    # a per cent (measured 0.2 % mean); with BERT at the fine-tuning rate 2e-5 an AdamW step is below the bf16 resolution of
    # most weights, the compute copy follows the masters in coarser steps and the per-batch losses of a random-label task
    # differ by several per cent without either run being wrong.  A broken gradient plane or kernel shows up as a step-0
    # mismatch, non-identical replicas, or O(1) / diverging losses.
    first_ok = rel[0] < 0.01
    if workload == "resnet50":
        ok = same and finite and first_ok and sum(rel) / len(rel) < 0.05 and max(rel) < 0.20
    else:
        ok = same and finite and first_ok and sum(rel) / len(rel) < 0.20 and max(ours_losses) < 2.0 * max(stock_losses)
    return {"ok": all(pg.all_gather_object(ok)), "ranks_identical": same, "mean_rel": sum(rel) / len(rel), "max_rel": max(rel),
            "ours": ours_losses, "stock": stock_losses, "comm": tr.ddp.comm.name}
