"""Fused aggregation kernel (reduce-scatter + weighting + sharded Adam + all-gather of parameters and
bf16 shadows) — all logical ranks emulated on ONE GPU as co-resident blocks of a single launch."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup(world, model_type="prodLDA", V=700, K=12, B=48, share=None):
    from gfedntm_b200.engine.fused import FusedEngine
    from gfedntm_b200.models.networks import DecoderNetwork
    from gfedntm_b200.ops.decoder import CsrBatch
    from gfedntm_b200.parallel.fabric import EmulatedArena, FusedAggregator

    torch.manual_seed(0)
    base = DecoderNetwork(V, K, model_type, (24, 24), "softplus", 0.2, True)
    arena = EmulatedArena(world)
    engines, batches = [], []
    g = torch.Generator(device="cuda").manual_seed(11)
    for r in range(world):
        model = copy.deepcopy(base)
        eng = FusedEngine(model, max_batch=B, max_nnz=B * 30, lr=2e-3, betas=(0.99, 0.99), seed=5 + r,
                          arena=arena if r == 0 else arena.view_as(r))
        if r == 0:
            arena = eng.flat.arena  # committed
        engines.append(eng)
        x = torch.zeros(B, V, device="cuda")
        x.scatter_add_(1, torch.randint(0, V, (B, 30), device="cuda", generator=g), torch.ones(B, 30, device="cuda"))
        batches.append(CsrBatch.from_dense(x))
    weights = [float(i + 1) for i in range(world)]
    tot = sum(weights)
    weights = [w / tot for w in weights]
    fl = engines[0].flat
    ranges, s_mask, beta_shared = None, 63, True
    if share is not None:
        from gfedntm_b200.federated.trainer import FederatedTrainer

        import types

        planner = FederatedTrainer.__new__(FederatedTrainer)
        planner.grads_to_share = list(share)
        planner.model, planner.agg, planner.world, planner.arena = types.SimpleNamespace(engine=engines[0]), "grad", 1, arena
        ranges, s_mask, beta_shared = planner._share_plan(fl)
    agg = FusedAggregator(world=world, weights=weights, flat_n=fl.beta_off, beta_off=fl.beta_off, s_n=fl.s_n,
                          K=K, V=V, Kp=engines[0].Kp,
                          ldb=engines[0].beta_bf.stride(0) if engines[0].is_prod else 0, lr=2e-3,
                          betas=(0.99, 0.99), ranges=ranges, s_mask=s_mask, beta_shared=beta_shared)
    for r, eng in enumerate(engines):
        agg.add_rank(arena.view_as(r) if r else arena, r, M=eng.flat.M, V2=eng.flat.V2, has_shadows=eng.is_prod)
    return engines, batches, weights, agg


@pytest.mark.parametrize("world", [2, 4, 8])
def test_fused_grad_aggregation_matches_global_adam(world):
    engines, batches, weights, agg = _setup(world)
    names = [n for n, _ in engines[0].model.named_parameters()]
    ref = {n: engines[0].flat.view(engines[0].flat.P, n).detach().clone() for n in names}
    for n in ref:
        ref[n].requires_grad_(True)
    opt = torch.optim.Adam([ref[n] for n in names], lr=2e-3, betas=(0.99, 0.99))
    for step in range(1, 4):
        for r, eng in enumerate(engines):
            eng.forward_backward(batches[r], training=True, gscale=weights[r])
        gsum = {n: sum(e.flat.view(e.flat.G, n).detach().clone() for e in engines) for n in names}
        s_mean = sum(w * e.flat.S.detach().clone() for w, e in zip(weights, engines))
        agg.launch(mode="grad", step=step)
        torch.cuda.synchronize()
        for n in names:
            ref[n].grad = gsum[n]
        opt.step()
        for r, eng in enumerate(engines):
            for n in names:
                torch.testing.assert_close(eng.flat.view(eng.flat.P, n), ref[n].detach(), rtol=2e-4, atol=2e-6,
                                           msg=lambda m: f"rank {r} {n} step {step}: {m}")
            torch.testing.assert_close(eng.flat.S, s_mean, rtol=1e-5, atol=1e-6)
            K, V = eng.K, eng.V
            # the bf16 operand shadows are the rounding of the rank's own fp32 beta (compared with the oracle above):
            # exact, where a comparison with the rounded ORACLE flips an ulp whenever a value sits on a bf16 boundary
            own = eng.flat.view(eng.flat.P, "beta").detach()
            torch.testing.assert_close(eng.beta_bf[:K, :V].float(), own.to(torch.bfloat16).float(), rtol=0, atol=0)
            torch.testing.assert_close(eng.betaT_bf[:, :K].float(), own.t().to(torch.bfloat16).float(), rtol=0, atol=0)
            eng.step_count = step


def test_fused_param_aggregation_matches_weighted_mean():
    world = 4
    engines, batches, weights, agg = _setup(world, model_type="LDA")
    names = [n for n, _ in engines[0].model.named_parameters()]
    for step in range(1, 3):
        for r, eng in enumerate(engines):
            eng.forward_backward(batches[r], training=True)
            eng.optimizer_step()
        want = {n: sum(w * e.flat.view(e.flat.P, n).detach().clone() for w, e in zip(weights, engines)) for n in names}
        agg.launch(mode="param", step=step)
        torch.cuda.synchronize()
        for eng in engines:
            for n in names:
                torch.testing.assert_close(eng.flat.view(eng.flat.P, n), want[n], rtol=1e-5, atol=1e-7)


@pytest.mark.parametrize("mode", ["grad", "param"])
def test_subset_sharing_keeps_unshared_tensors_local(mode):
    """grads_to_share subset (reference: federated_model.py:98-131, server default server.py:71): shared tensors
    follow the global update, every other tensor the client's own optimiser; only shared BatchNorm statistics are
    averaged."""
    world = 2
    share = ["prior_mean", "prior_variance", "beta", "inf_net.f_mu.weight", "beta_batchnorm.running_var"]
    engines, batches, weights, agg = _setup(world, share=share)
    names = [n for n, _ in engines[0].model.named_parameters()]
    init = {n: engines[0].flat.view(engines[0].flat.P, n).detach().clone() for n in names}
    # oracles: one global Adam (shared tensors) + one private Adam per client (unshared tensors)
    glob = {n: init[n].clone().requires_grad_(True) for n in names}
    gopt = torch.optim.Adam([glob[n] for n in names], lr=2e-3, betas=(0.99, 0.99))
    loc = [{n: init[n].clone().requires_grad_(True) for n in names} for _ in range(world)]
    lopt = [torch.optim.Adam([l[n] for n in names], lr=2e-3, betas=(0.99, 0.99)) for l in loc]
    for step in range(1, 3):
        if mode == "grad":
            for r, eng in enumerate(engines):
                agg.args.local_gscale = 1.0  # emulation: one argument block for every rank -> unit weights below
                eng.forward_backward(batches[r], training=True, gscale=1.0)
            g = [{n: e.flat.view(e.flat.G, n).detach().clone() for n in names} for e in engines]
            s_before = [e.flat.S.detach().clone() for e in engines]
            agg.launch(mode="grad", step=step)
            torch.cuda.synchronize()
            for n in names:
                glob[n].grad = sum(gr[n] for gr in g)
            gopt.step()
            for r in range(world):
                for n in names:
                    loc[r][n].grad = g[r][n]
                lopt[r].step()
            for r, eng in enumerate(engines):
                for n in names:
                    want = glob[n] if n in share else loc[r][n]
                    torch.testing.assert_close(eng.flat.view(eng.flat.P, n), want.detach(), rtol=3e-4, atol=3e-6,
                                               msg=lambda m: f"rank {r} {n} step {step}: {m}")
                eng.step_count = step
        else:
            for r, eng in enumerate(engines):
                eng.forward_backward(batches[r], training=True)
                eng.optimizer_step()
            mine = [{n: e.flat.view(e.flat.P, n).detach().clone() for n in names} for e in engines]
            s_before = [e.flat.S.detach().clone() for e in engines]
            agg.launch(mode="param", step=step)
            torch.cuda.synchronize()
            for r, eng in enumerate(engines):
                for n in names:
                    want = sum(w * m[n] for w, m in zip(weights, mine)) if n in share else mine[r][n]
                    torch.testing.assert_close(eng.flat.view(eng.flat.P, n), want, rtol=1e-5, atol=1e-7)
        K, V = engines[0].K, engines[0].V
        s_mean = sum(w * s for w, s in zip(weights, s_before))
        for r, eng in enumerate(engines):
            # only beta_batchnorm.running_var (last block) is shared
            torch.testing.assert_close(eng.flat.S[4 * K + V:4 * K + 2 * V], s_mean[4 * K + V:4 * K + 2 * V],
                                       rtol=1e-5, atol=1e-6)
            torch.testing.assert_close(eng.flat.S[:4 * K + V], s_before[r][:4 * K + V], rtol=0, atol=0)


@pytest.mark.parametrize("agg", ["grad", "param"])
def test_emulated_federation_trains_identical_clients(agg):
    """EmulatedFederation: N logical clients on one GPU through the same aggregation kernel; after every step the
    clients hold bit-identical shared state and the loss goes down."""
    import numpy as np

    from gfedntm_b200.federated.emulated import EmulatedFederation
    from gfedntm_b200.models import AVITM, BOWDataset

    rng = np.random.default_rng(0)
    V, K = 600, 8
    datasets, trainers = [], []
    for r in range(3):
        X = np.zeros((64 * (r + 1), V), dtype=np.float32)
        for b in range(X.shape[0]):
            blk = b % K
            idx = rng.integers(blk * (V // K), (blk + 1) * (V // K), 60)
            np.add.at(X[b], idx, 1.0)
        datasets.append(BOWDataset(X, None))
        trainers.append(AVITM(input_size=V, n_components=K, hidden_sizes=(32, 32), batch_size=64, num_epochs=1,
                              verbose=False, device="cuda:0", seed=11))
    fed = EmulatedFederation(trainers, datasets, agg=agg)
    assert abs(sum(fed.weights) - 1.0) < 1e-9 and fed.weights[2] > fed.weights[0]
    first = sum(float(l.item()) for l in fed.step())
    for _ in range(60):
        last = sum(float(l.item()) for l in fed.step())
    torch.cuda.synchronize()
    assert last < 0.97 * first, (first, last)
    for e in fed.engines[1:]:
        assert torch.equal(e.flat.P, fed.engines[0].flat.P) and torch.equal(e.flat.S, fed.engines[0].flat.S)
        assert torch.equal(e.beta_bf, fed.engines[0].beta_bf)
