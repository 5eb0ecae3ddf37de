"""Fused engine (all kernels of one training step) against eager PyTorch autograd on the same
parameters, noise and dropout masks."""
import copy

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _oracle(model, x, x_ctx, eps, mask_enc, mask_theta, kl_weight=1.0):
    """Eager forward with injected noise/masks; returns loss (sum over batch)."""
    from gfedntm_b200.models.networks import elbo_terms

    inf = model.inf_net
    h = inf.activation(inf.input_layer(inf.first_layer_input(x, x_ctx, None)))
    h = inf.hiddens(h)
    h = h * mask_enc
    mu = inf.f_mu_batchnorm(inf.f_mu(h))
    logvar = inf.f_sigma_batchnorm(inf.f_sigma(h))
    theta = F.softmax(mu + eps * torch.exp(0.5 * logvar), dim=1) * mask_theta
    wd = model.decode(theta)
    kl, rl = elbo_terms(x, wd, model.prior_mean, model.prior_variance, mu, torch.exp(logvar), logvar)
    return (kl_weight * kl + rl).sum(), theta


def _masks(B, H, K, p_enc, p_theta, g):
    me = (torch.rand(B, H, device="cuda", generator=g) > p_enc).float() / (1 - p_enc)
    mt = (torch.rand(B, K, device="cuda", generator=g) > p_theta).float() / (1 - p_theta)
    return me, mt


@pytest.mark.parametrize("kind,model_type,act,hidden", [
    ("bow", "prodLDA", "softplus", (50, 50)),
    ("bow", "prodLDA", "relu", (64,)),
    ("bow", "LDA", "softplus", (50, 50)),
    ("combined", "prodLDA", "softplus", (50, 50)),
    ("zeroshot", "prodLDA", "tanh", (40, 30, 20)),
])
def test_engine_step_matches_autograd(kind, model_type, act, hidden):
    _autograd_case(kind, model_type, act, hidden, 96)


@pytest.mark.parametrize("kind,model_type,act,hidden", [
    ("bow", "prodLDA", "softplus", (50, 50)),
    ("zeroshot", "prodLDA", "tanh", (40, 30, 20)),
    ("bow", "LDA", "relu", (64,)),
])
def test_engine_step_matches_autograd_large_batch(kind, model_type, act, hidden):
    """The same comparison at a batch that spans many blocks of every per-document kernel (partial last block)."""
    _autograd_case(kind, model_type, act, hidden, 650)


def _autograd_case(kind, model_type, act, hidden, B):
    from gfedntm_b200.engine.fused import FusedEngine
    from gfedntm_b200.models.networks import DecoderNetwork
    from gfedntm_b200.ops.decoder import CsrBatch

    torch.manual_seed(0)
    V, K, C = 1500, 20, 48
    g = torch.Generator(device="cuda").manual_seed(5)
    model = DecoderNetwork(V, K, model_type, hidden, act, 0.2, True, infnet=kind, contextual_size=C if kind != "bow" else 0)
    # bigger beta than xavier so the decoder is not dominated by the BN epsilon
    with torch.no_grad():
        model.beta.mul_(20.0)
    ref = copy.deepcopy(model).cuda().train()
    x = torch.zeros(B, V, device="cuda")
    idx = torch.randint(0, V, (B, 60), device="cuda", generator=g)
    x.scatter_add_(1, idx, torch.ones(B, 60, device="cuda"))
    x_ctx = torch.randn(B, C, device="cuda", generator=g) if kind != "bow" else None
    eps = torch.randn(B, K, device="cuda", generator=g)
    me, mt = _masks(B, hidden[-1], K, 0.2, 0.2, g)

    eng = FusedEngine(model, max_batch=B, max_nnz=B * 60, lr=2e-3, seed=1)
    batch = CsrBatch.from_dense(x)
    loss = eng.forward_backward(batch, x_ctx, training=True, eps_in=eps, mask_enc_in=me, mask_theta_in=mt)
    torch.cuda.synchronize()

    ref_loss, ref_theta = _oracle(ref, x, x_ctx, eps, me, mt)
    ref_loss.backward()
    assert abs(loss.item() - ref_loss.item()) <= 3e-3 * abs(ref_loss.item()), (loss.item(), ref_loss.item())
    torch.testing.assert_close(eng.theta[:B], ref_theta, rtol=1e-3, atol=1e-5)
    bad = []
    for name, p in ref.named_parameters():
        got = eng.flat.view(eng.flat.G, name)
        scale = p.grad.abs().max().item()
        err = (got - p.grad).abs().max().item()
        # absolute slack: d prior_mean and the biases feeding a BatchNorm have exactly-zero gradients
        # (BatchNorm removes them), so both sides hold rounding noise of ~1e-3 there
        if err > 3e-2 * scale + 2e-3 * max(1.0, B / 96):
            bad.append(f"{name}: err {err:.3e} scale {scale:.3e}")
    assert not bad, "\n".join(bad)
    # running statistics of the two encoder BatchNorms follow PyTorch's update rule
    torch.testing.assert_close(model.inf_net.f_mu_batchnorm.running_mean, ref.inf_net.f_mu_batchnorm.running_mean,
                               rtol=1e-3, atol=1e-5)
    torch.testing.assert_close(model.inf_net.f_sigma_batchnorm.running_var, ref.inf_net.f_sigma_batchnorm.running_var,
                               rtol=1e-3, atol=1e-5)
    sd = model.state_dict()
    assert int(sd["beta_batchnorm.num_batches_tracked"]) == 1


def test_engine_adam_matches_torch():
    from gfedntm_b200.engine.fused import FusedEngine
    from gfedntm_b200.models.networks import DecoderNetwork
    from gfedntm_b200.ops.decoder import CsrBatch

    torch.manual_seed(1)
    B, V, K = 64, 800, 10
    model = DecoderNetwork(V, K, "prodLDA", (32, 32), "softplus", 0.2, True)
    eng = FusedEngine(model, max_batch=B, max_nnz=B * 30, lr=2e-3, betas=(0.99, 0.99), seed=3)
    x = torch.zeros(B, V, device="cuda")
    x.scatter_add_(1, torch.randint(0, V, (B, 30), device="cuda"), torch.ones(B, 30, device="cuda"))
    batch = CsrBatch.from_dense(x)
    params = {n: p.detach().clone().requires_grad_(True) for n, p in model.named_parameters()}
    opt = torch.optim.Adam(list(params.values()), lr=2e-3, betas=(0.99, 0.99))
    for _ in range(3):
        eng.forward_backward(batch, training=True)
        for n, p in params.items():
            p.grad = eng.flat.view(eng.flat.G, n).detach().clone()
        eng.optimizer_step()
        opt.step()
    torch.cuda.synchronize()
    for n, p in params.items():
        torch.testing.assert_close(eng.flat.view(eng.flat.P, n), p.detach(), rtol=1e-4, atol=1e-6)
    # bf16 shadows track the fp32 master copy of beta
    torch.testing.assert_close(eng.beta_bf[:K, :V].float(), model.beta.detach().to(torch.bfloat16).float())
    torch.testing.assert_close(eng.betaT_bf[:, :K].float(), model.beta.detach().t().to(torch.bfloat16).float())


def test_engine_loss_decreases_and_theta_inference():
    from gfedntm_b200.engine.fused import FusedEngine
    from gfedntm_b200.models.networks import DecoderNetwork
    from gfedntm_b200.ops.decoder import CsrBatch

    torch.manual_seed(2)
    B, V, K = 128, 1000, 8
    # documents drawn from 8 disjoint word blocks -> learnable structure
    x = torch.zeros(B, V, device="cuda")
    for b in range(B):
        blk = b % K
        idx = torch.randint(blk * (V // K), (blk + 1) * (V // K), (80,), device="cuda")
        x[b].scatter_add_(0, idx, torch.ones(80, device="cuda"))
    model = DecoderNetwork(V, K, "prodLDA", (50, 50), "softplus", 0.2, True)
    eng = FusedEngine(model, max_batch=B, max_nnz=int((x > 0).sum().item()), lr=2e-3, seed=0)
    batch = CsrBatch.from_dense(x)
    losses = [eng.train_step(batch).item() for _ in range(150)]
    assert all(l == l for l in losses)
    assert losses[-1] < 0.9 * losses[0], (losses[0], losses[-1])
    # every step also leaves its loss in a pinned host slot (copied by the step's own CUDA graph on the side stream)
    for _ in range(3):
        dev_loss = eng.train_step(batch)
        torch.cuda.synchronize()
        assert eng.loss_slot.is_pinned() and float(eng.loss_slot[0]) == float(dev_loss.item())
    theta = eng.sample_theta(batch, n_samples=5)
    torch.testing.assert_close(theta.sum(1), torch.ones(B, device="cuda"), rtol=1e-4, atol=1e-4)
    val = eng.forward_backward(batch, training=False, sample=True, need_grad=False)
    assert torch.isfinite(val).all()


def test_theta_mc_matches_per_draw_softmax_statistics():
    """gf_theta_mc (all Monte-Carlo draws + threshold / renormalisation in one launch) against the same
    computation in PyTorch: the draws are random, so the mean over many draws must match the analytic
    expectation estimated with an independent large sample; the epilogue must match exactly."""
    from gfedntm_b200.ops import kernels as K_

    torch.manual_seed(0)
    B, K = 37, 50
    dd = dict(dtype=torch.float32, device="cuda")
    mu_pre, ls_pre = torch.randn(B, K, **dd), torch.randn(B, K, **dd) * 0.3 - 1.0
    mu_rm, ls_rm = torch.randn(K, **dd) * 0.1, torch.randn(K, **dd) * 0.1
    mu_rv, ls_rv = torch.rand(K, **dd) + 0.5, torch.rand(K, **dd) + 0.5
    mu = (mu_pre - mu_rm) * torch.rsqrt(mu_rv + 1e-5)
    sd = torch.exp(0.5 * (ls_pre - ls_rm) * torch.rsqrt(ls_rv + 1e-5))
    n = 4000
    out = torch.empty(B, K, **dd)
    K_.theta_mc(mu_pre, ls_pre, mu_rm, mu_rv, ls_rm, ls_rv, out, B=B, K=K, n_samples=n, seed=123, thr=0.0)
    ref = torch.zeros(B, K, **dd)
    for _ in range(40):
        ref += torch.softmax(mu + torch.randn(100, B, K, **dd) * sd, dim=-1).sum(0)
    ref /= 4000
    torch.testing.assert_close(out.sum(1), torch.ones(B, **dd), rtol=1e-4, atol=1e-4)
    assert float((out - ref).abs().max()) < 0.02  # two independent 4000-draw estimates of a value in [0,1]
    # different seeds give different draws, the same seed the same result
    out2, out3 = torch.empty_like(out), torch.empty_like(out)
    K_.theta_mc(mu_pre, ls_pre, mu_rm, mu_rv, ls_rm, ls_rv, out2, B=B, K=K, n_samples=20, seed=1, thr=0.0)
    K_.theta_mc(mu_pre, ls_pre, mu_rm, mu_rv, ls_rm, ls_rv, out3, B=B, K=K, n_samples=20, seed=1, thr=0.0)
    assert torch.equal(out2, out3)
    K_.theta_mc(mu_pre, ls_pre, mu_rm, mu_rv, ls_rm, ls_rv, out3, B=B, K=K, n_samples=20, seed=2, thr=0.0)
    assert not torch.equal(out2, out3)
    # threshold + L1 renormalisation epilogue (federated_model.py:170-173) on the same draws
    thr = 3e-3
    K_.theta_mc(mu_pre, ls_pre, mu_rm, mu_rv, ls_rm, ls_rv, out3, B=B, K=K, n_samples=20, seed=1, thr=thr)
    want = out2.clone()
    want[want < thr] = 0
    want = want / want.sum(1, keepdim=True)
    torch.testing.assert_close(out3, want, rtol=1e-5, atol=1e-7)


@pytest.mark.parametrize("solver", ["sgd", "adagrad", "adadelta", "rmsprop"])
def test_engine_other_solvers_match_torch(solver):
    """The non-Adam solvers of the reference's optimiser factory (avitm.py:141-153) on the fused path: three steps of
    the engine's fused update against torch.optim with torch's defaults on the same gradients."""
    from gfedntm_b200.engine.fused import FusedEngine
    from gfedntm_b200.models.networks import DecoderNetwork
    from gfedntm_b200.ops.decoder import CsrBatch

    torch.manual_seed(1)
    B, V, K = 64, 800, 10
    lr, mom = 2e-3, 0.9
    model = DecoderNetwork(V, K, "prodLDA", (32, 32), "softplus", 0.2, True)
    eng = FusedEngine(model, max_batch=B, max_nnz=B * 30, lr=lr, betas=(mom, 0.99), seed=3, solver=solver)
    x = torch.zeros(B, V, device="cuda")
    x.scatter_add_(1, torch.randint(0, V, (B, 30), device="cuda"), torch.ones(B, 30, device="cuda"))
    batch = CsrBatch.from_dense(x)
    params = {n: p.detach().clone().requires_grad_(True) for n, p in model.named_parameters()}
    ps = list(params.values())
    opt = {"sgd": lambda: torch.optim.SGD(ps, lr=lr, momentum=mom), "adagrad": lambda: torch.optim.Adagrad(ps, lr=lr),
           "adadelta": lambda: torch.optim.Adadelta(ps, lr=lr),
           "rmsprop": lambda: torch.optim.RMSprop(ps, lr=lr, momentum=mom)}[solver]()
    for _ in range(3):
        eng.forward_backward(batch, training=True)
        for n, p in params.items():
            p.grad = eng.flat.view(eng.flat.G, n).detach().clone()
        eng.optimizer_step()
        opt.step()
    torch.cuda.synchronize()
    for n, p in params.items():
        torch.testing.assert_close(eng.flat.view(eng.flat.P, n), p.detach(), rtol=2e-4, atol=2e-6, msg=lambda m: f"{n}: {m}")
    torch.testing.assert_close(eng.beta_bf[:K, :V].float(), model.beta.detach().to(torch.bfloat16).float())


def test_non_adam_solver_and_plateau_run_on_the_fused_backend():
    """AVITM(solver='rmsprop', reduce_on_plateau=True) no longer falls back to eager PyTorch; the plateau rule lowers
    the engine's learning rate and drops the captured graphs."""
    import numpy as np

    from gfedntm_b200.models import AVITM, BOWDataset
    from gfedntm_b200.models.base import PlateauLR

    rng = np.random.default_rng(0)
    X = rng.poisson(0.05, (256, 900)).astype(np.float32)
    X[X.sum(1) == 0, 0] = 1
    m = AVITM(input_size=900, n_components=8, hidden_sizes=(16, 16), batch_size=64, num_epochs=2, solver="rmsprop",
              reduce_on_plateau=True, verbose=False, device="cuda:0", seed=0)
    assert m.backend == "fused" and isinstance(m.scheduler, PlateauLR)
    m.fit(BOWDataset(X, None))
    assert m.engine.solver_name == "rmsprop" and m.engine.solver == 4
    lr0 = m.lr
    for _ in range(12):
        m.scheduler.step(1.0)   # no improvement for more than `patience` calls
    assert abs(m.lr - 0.1 * lr0) < 1e-12 and abs(m.engine.lr - 0.1 * lr0) < 1e-12 and not m.engine._graphs


def test_single_gpu_overlapped_beta_update_matches_separate_kernel(monkeypatch):
    """GFEDNTM_SELF_OVERLAP=1 runs the aggregation role of the d-theta sweep with a world of one (TMA-fed update
    loop, bf16 gradient): three training steps must track the default path (separate fused Adam on fp32 gradients)."""
    import numpy as np

    from gfedntm_b200.engine.fused import FusedEngine
    from gfedntm_b200.models.networks import DecoderNetwork
    from gfedntm_b200.ops.decoder import CsrBatch

    monkeypatch.setenv("GFEDNTM_DECODER_V1", "0")   # folded sweeps at a small shape
    B, V, K = 192, 2048, 24
    g = torch.Generator(device="cuda").manual_seed(2)
    x = torch.zeros(B, V, device="cuda")
    x.scatter_add_(1, torch.randint(0, V, (B, 50), device="cuda", generator=g), torch.ones(B, 50, device="cuda"))
    batch = CsrBatch.from_dense(x)
    eps = torch.randn(B, K, device="cuda", generator=g)
    me, mt = torch.ones(B, 32, device="cuda"), torch.ones(B, K, device="cuda")
    res = {}
    for flag in ("0", "1"):
        monkeypatch.setenv("GFEDNTM_SELF_OVERLAP", flag)
        torch.manual_seed(0)
        model = DecoderNetwork(V, K, "prodLDA", (32, 32), "softplus", 0.2, True)
        with torch.no_grad():
            model.beta.mul_(10.0)
        eng = FusedEngine(model, max_batch=B, max_nnz=batch.nnz, lr=2e-3, betas=(0.99, 0.99), seed=1)
        assert (eng._self_overlap is not None) == (flag == "1")
        losses = [float(eng.train_step(batch, eps_in=eps, mask_enc_in=me, mask_theta_in=mt).item()) for _ in range(3)]
        torch.cuda.synchronize()
        res[flag] = (losses, model.beta.detach().clone(), eng.beta_bf.detach().clone(), eng.betaT_bf.detach().clone())
    np.testing.assert_allclose(res["1"][0], res["0"][0], rtol=2e-5)
    # Adam's first steps move every weight by ~lr whatever the gradient magnitude: a bf16-rounded gradient may flip
    # the direction of weights whose gradient is ~0, so compare in units of the step size
    diff = (res["1"][1] - res["0"][1]).abs()
    assert float(diff.mean()) < 0.05 * 2e-3 * 3 and float((diff > 2e-3).float().mean()) < 0.02
    assert torch.equal(res["1"][2][:K, :V].float(), res["1"][1].to(torch.bfloat16).float())
    assert torch.equal(res["1"][3][:, :K].float(), res["1"][1].t().to(torch.bfloat16).float())


@pytest.mark.parametrize("kind", ["combined", "zeroshot"])
def test_engine_supervised_head_matches_autograd(kind):
    """label_size > 0 on the fused path: labels as extra first-layer inputs, Linear(K -> L) head on theta with the mean
    cross-entropy added to the loss (decoding_network.py:84-85,156-159, ctm.py:292-296) -- against eager autograd."""
    from gfedntm_b200.engine.fused import FusedEngine
    from gfedntm_b200.models.networks import DecoderNetwork, elbo_terms
    from gfedntm_b200.ops.decoder import CsrBatch

    torch.manual_seed(0)
    B, V, K, C, L = 64, 1200, 16, 40, 7
    g = torch.Generator(device="cuda").manual_seed(5)
    model = DecoderNetwork(V, K, "prodLDA", (32, 32), "softplus", 0.2, True, infnet=kind, contextual_size=C, label_size=L)
    with torch.no_grad():
        model.beta.mul_(20.0)
    ref = copy.deepcopy(model).cuda().train()
    x = torch.zeros(B, V, device="cuda")
    x.scatter_add_(1, torch.randint(0, V, (B, 50), device="cuda", generator=g), torch.ones(B, 50, device="cuda"))
    x_ctx = torch.randn(B, C, device="cuda", generator=g)
    labels = F.one_hot(torch.randint(0, L, (B,), device="cuda", generator=g), L).float()
    eps = torch.randn(B, K, device="cuda", generator=g)
    me, mt = _masks(B, 32, K, 0.2, 0.2, g)
    eng = FusedEngine(model, max_batch=B, max_nnz=B * 50, lr=2e-3, seed=1)
    loss = eng.forward_backward(CsrBatch.from_dense(x), x_ctx, training=True, eps_in=eps, mask_enc_in=me,
                                mask_theta_in=mt, labels=labels)
    torch.cuda.synchronize()
    inf = ref.inf_net
    h = inf.hiddens(inf.activation(inf.input_layer(inf.first_layer_input(x, x_ctx, labels)))) * me
    mu = inf.f_mu_batchnorm(inf.f_mu(h))
    logvar = inf.f_sigma_batchnorm(inf.f_sigma(h))
    theta = F.softmax(mu + eps * torch.exp(0.5 * logvar), dim=1) * mt
    kl, rl = elbo_terms(x, ref.decode(theta), ref.prior_mean, ref.prior_variance, mu, torch.exp(logvar), logvar)
    ref_loss = (kl + rl).sum() + torch.nn.CrossEntropyLoss()(ref.label_classification(theta), torch.argmax(labels, 1))
    ref_loss.backward()
    assert abs(loss.item() - ref_loss.item()) <= 3e-3 * abs(ref_loss.item()), (loss.item(), ref_loss.item())
    bad = []
    for name, p in ref.named_parameters():
        got = eng.flat.view(eng.flat.G, name)
        scale = p.grad.abs().max().item()
        err = (got - p.grad).abs().max().item()
        if err > 3e-2 * scale + 2e-3:
            bad.append(f"{name}: err {err:.3e} scale {scale:.3e}")
    assert not bad, "\n".join(bad)
    # the head's own gradients are exact fp32 work: tighter
    for name in ("label_classification.weight", "label_classification.bias"):
        torch.testing.assert_close(eng.flat.view(eng.flat.G, name), dict(ref.named_parameters())[name].grad, rtol=2e-3, atol=1e-6)
    # one optimiser step moves every tensor, label rows of the first layer included
    before = eng.flat.P.clone()
    eng.optimizer_step()
    torch.cuda.synchronize()
    w = eng.flat.segments["inf_net.input_layer.weight"]
    lab = slice(w.offset + w.numel - L * 32, w.offset + w.numel)          # storage [fan_in, H0]: the last L rows
    assert float((eng.flat.P[lab] - before[lab]).abs().min()) > 0
    ctx_rows = slice(w.offset + w.numel - (L + C) * 32, w.offset + w.numel - L * 32) if kind == "zeroshot" else \
        slice(w.offset + V * 32, w.offset + 2 * V * 32)
    assert float((eng.flat.P[ctx_rows] - before[ctx_rows]).abs().max()) > 0


@pytest.mark.parametrize("rows,V,k", [(7, 1000, 10), (50, 20011, 15), (3, 33, 32)])
def test_row_topk_matches_torch(rows, V, k):
    """Top words per topic (get_topics; ref avitm.py:574 torch.topk per row)."""
    from gfedntm_b200.ops import kernels as K_

    g = torch.Generator(device="cuda").manual_seed(rows * 131 + V)
    mat = torch.randn(rows, V, device="cuda", generator=g)
    val, idx = K_.row_topk(mat, k)
    ref_val, ref_idx = torch.topk(mat, k, dim=1)
    torch.testing.assert_close(val, ref_val, rtol=0, atol=0)
    assert torch.equal(idx, ref_idx)
