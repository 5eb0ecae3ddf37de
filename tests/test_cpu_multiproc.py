"""Multi-process host-side logic on CPU (gloo, world_size 2): vocabulary consensus and client weights."""
import os
import socket

import torch.multiprocessing as mp


def _worker(rank, world, port, out):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from gfedntm_b200.parallel.consensus import consensus_vocabulary

    tokens = [["wd1", "wd5", "apple"], ["wd5", "zebra", "wd2", "wd2"]][rank]
    vocab, weights = consensus_vocabulary(tokens, [30, 90][rank])
    out.put((rank, vocab, weights))
    dist.barrier()
    dist.destroy_process_group()


def test_vocabulary_consensus_gloo():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in range(2)])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (_, v0, w0), (_, v1, w1) = res
    assert v0 == v1 == {"apple": 0, "wd1": 1, "wd2": 2, "wd5": 3, "zebra": 4}
    assert w0 == w1 == [0.25, 0.75]


def _agg_worker(rank, world, port, out):
    import torch
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from gfedntm_b200.parallel.host_fabric import HostAggregator, owner_slice

    n = 1003  # not a multiple of the 4-element granule nor of the world size
    weights = [0.25, 0.75]
    g0 = torch.Generator().manual_seed(0)
    p_init = torch.randn(n, generator=g0)
    # --- grad mode: 3 steps, every rank has its own gradients
    params = p_init.clone()
    stats = torch.full((8,), float(rank + 1))
    agg = HostAggregator(n, weights, lr=2e-3, betas=(0.99, 0.99))
    grads = []
    for step in range(3):
        gl = torch.randn(n, generator=torch.Generator().manual_seed(100 * step + rank))
        grads.append(gl)
        agg.aggregate_grad(params, gl * weights[rank], stats if step == 0 else None)
    lo, hi = owner_slice(n, world, rank)
    assert agg.m.numel() == hi - lo  # the optimiser state really is sharded
    # --- param mode
    pm = torch.full((n,), float(rank + 1))
    agg.aggregate_param(pm)
    out.put((rank, params, stats, pm, (lo, hi)))
    dist.barrier()
    dist.destroy_process_group()


def test_host_fabric_matches_single_process_adam():
    """Sharded reduce-scatter / Adam / all-gather over gloo == torch.optim.Adam on the weighted gradient sum;
    parameter mode == the reference's weighted mean (server.py:477-487)."""
    import torch

    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_agg_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in range(2)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (_, p0, s0, pm0, sl0), (_, p1, s1, pm1, sl1) = res
    assert sl0 == (0, 504) and sl1 == (504, 1003)
    assert torch.equal(p0, p1) and torch.equal(s0, s1) and torch.equal(pm0, pm1)
    # oracle
    n, weights = 1003, [0.25, 0.75]
    ref = torch.randn(n, generator=torch.Generator().manual_seed(0)).requires_grad_(True)
    opt = torch.optim.Adam([ref], lr=2e-3, betas=(0.99, 0.99))
    for step in range(3):
        g = sum(weights[r] * torch.randn(n, generator=torch.Generator().manual_seed(100 * step + r)) for r in range(2))
        ref.grad = g
        opt.step()
    torch.testing.assert_close(p0, ref.detach(), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(s0, torch.full((8,), 0.25 * 1 + 0.75 * 2))
    torch.testing.assert_close(pm0, torch.full((n,), 0.25 * 1 + 0.75 * 2))


def _subset_worker(rank, world, port, out):
    import torch
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from gfedntm_b200.parallel.host_fabric import HostAggregator

    n, weights = 64, [0.25, 0.75]
    shared = torch.zeros(n, dtype=torch.bool)
    shared[16:40] = True
    params = torch.randn(n, generator=torch.Generator().manual_seed(0))
    agg = HostAggregator(n, weights, lr=1e-2, betas=(0.9, 0.99), shared=shared)
    for step in range(2):
        gl = torch.randn(n, generator=torch.Generator().manual_seed(10 * step + rank))
        agg.aggregate_grad(params, gl * weights[rank])
    pm = torch.full((n,), float(rank + 1))
    agg.aggregate_param(pm)
    out.put((rank, params, pm))
    dist.barrier()
    dist.destroy_process_group()


def test_host_fabric_subset_sharing_keeps_unshared_tensors_local():
    """grads_to_share subset (federated_model.py:98-131): shared elements follow the global Adam on the weighted
    gradient sum, unshared ones the client's own Adam on its own gradient; parameter mode averages the shared
    elements only."""
    import torch

    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_subset_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in range(2)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (_, p0, pm0), (_, p1, pm1) = res
    n, weights = 64, [0.25, 0.75]
    shared = torch.zeros(n, dtype=torch.bool)
    shared[16:40] = True
    assert torch.equal(p0[shared], p1[shared]) and not torch.equal(p0[~shared], p1[~shared])
    init = torch.randn(n, generator=torch.Generator().manual_seed(0))
    refs = []
    for who in ("global", 0, 1):
        ref = init.clone().requires_grad_(True)
        opt = torch.optim.Adam([ref], lr=1e-2, betas=(0.9, 0.99))
        for step in range(2):
            gs = [torch.randn(n, generator=torch.Generator().manual_seed(10 * step + r)) for r in range(2)]
            ref.grad = sum(weights[r] * gs[r] for r in range(2)) if who == "global" else gs[who]
            opt.step()
        refs.append(ref.detach())
    torch.testing.assert_close(p0[shared], refs[0][shared], rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(p0[~shared], refs[1][~shared], rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(p1[~shared], refs[2][~shared], rtol=1e-5, atol=1e-6)
    assert torch.equal(pm0[shared], torch.full((24,), 1.75)) and torch.equal(pm0[~shared], torch.full((40,), 1.0))
    assert torch.equal(pm1[~shared], torch.full((40,), 2.0))
