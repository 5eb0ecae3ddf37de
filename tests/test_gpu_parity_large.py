"""Numerical parity at the shape that is benchmarked (prodlda_200k: B=4096, V=200 000, K=200).

(a) the folded tcgen05 decoder (bf16 operands, fp32 accumulate) against a float64 oracle built from the UNROUNDED
    theta and beta, chunked over the vocabulary so that nothing of size [B,V] has to exist in float64 at once;
(b) 50 training steps of the whole fused engine against eager fp32 PyTorch (autograd + torch.optim.Adam) with the
    same initial weights, reparameterisation noise and dropout masks: the two loss curves must stay together.

The bounds asserted here are the bf16 budget quoted in DESIGN.md.  Ref ops: decoder_network.py:121-126,
avitm.py:203-229 and their autograd; federated_avitm.py:51-83 for the step.
"""
import copy
import json
import os

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

B, V, K = 4096, 200_000, 200


def _corpus(n_docs, seed):
    from gfedntm_b200.utils.synthetic import generate_corpus_gpu

    counts, _ = generate_corpus_gpu(n_docs, V, K, eta=0.01, seed=seed, device="cuda")
    return counts.tocsr()


def _csr_batch(counts, rows):
    from gfedntm_b200.ops.decoder import CsrBatch

    sub = counts[rows]
    return CsrBatch(torch.from_numpy(sub.indptr.astype("int32")).cuda(), torch.from_numpy(sub.indices.astype("int32")).cuda(),
                    torch.from_numpy(sub.data.astype("float32")).cuda(), sub.shape[0], V), sub


def _dense_cols(sub, c0, c1, dtype):
    blk = sub[:, c0:c1].tocoo()
    x = torch.zeros(sub.shape[0], c1 - c0, dtype=dtype, device="cuda")
    x.index_put_((torch.from_numpy(blk.row).long().cuda(), torch.from_numpy(blk.col).long().cuda()),
                 torch.from_numpy(blk.data).to(dtype).cuda(), accumulate=True)
    return x


def _oracle_decoder(theta, beta, sub, eps=1e-5, chunk=8192):
    """float64 forward + backward of softmax_V(BN_V(theta.beta)) with the BoW NLL, two passes over V chunks."""
    th = theta.double()
    Bn = th.shape[0]
    m = torch.full((Bn,), -float("inf"), dtype=torch.float64, device="cuda")
    s = torch.zeros(Bn, dtype=torch.float64, device="cuda")
    xy = torch.zeros(Bn, dtype=torch.float64, device="cuda")
    nb = torch.zeros(Bn, dtype=torch.float64, device="cuda")
    stats = []
    for c0 in range(0, V, chunk):
        c1 = min(V, c0 + chunk)
        L = th @ beta[:, c0:c1].double()
        mean, var = L.mean(0), L.var(0, unbiased=False)
        rstd = torch.rsqrt(var + eps)
        y = (L - mean) * rstd
        stats.append((mean, rstd))
        x = _dense_cols(sub, c0, c1, torch.float64)
        xy += (x * y).sum(1)
        nb += x.sum(1)
        cm = y.max(1).values
        nm = torch.maximum(m, cm)
        s = s * torch.exp(m - nm) + torch.exp(y - nm[:, None]).sum(1)
        m = nm
    lse = m + torch.log(s)
    rl = nb * lse - xy
    dtheta = torch.zeros(Bn, K, dtype=torch.float64, device="cuda")
    dbeta = torch.empty(K, V, dtype=torch.float32, device="cuda")
    for i, c0 in enumerate(range(0, V, chunk)):
        c1 = min(V, c0 + chunk)
        bc = beta[:, c0:c1].double()
        mean, rstd = stats[i]
        y = (th @ bc - mean) * rstd
        x = _dense_cols(sub, c0, c1, torch.float64)
        dy = nb[:, None] * torch.exp(y - lse[:, None]) - x
        dl = rstd * (dy - dy.mean(0) - y * (dy * y).mean(0))
        dtheta += dl @ bc.t()
        dbeta[:, c0:c1] = (th.t() @ dl).float()
    return rl, dtheta, dbeta


def test_folded_decoder_at_benchmark_shape_against_float64_oracle():
    from gfedntm_b200.ops.decoder import CscWorkspace, ProdLDADecoder, make_beta_shadows, refresh_beta_shadows

    torch.manual_seed(0)
    counts = _corpus(B, seed=11)
    batch, sub = _csr_batch(counts, slice(0, B))
    g = torch.Generator(device="cuda").manual_seed(3)
    theta = torch.softmax(torch.randn(B, K, device="cuda", generator=g) * 1.5, dim=1)
    # xavier-uniform-sized topic-word matrix (the reference's initialisation) times a few to leave the BN epsilon
    beta = (torch.rand(K, V, device="cuda", generator=g) - 0.5) * 0.2
    beta_bf, betaT_bf = make_beta_shadows(K, V, "cuda")
    refresh_beta_shadows(beta, beta_bf, betaT_bf)
    dec = ProdLDADecoder(B, V, K, batch.nnz, "cuda")
    assert dec.folded, "the benchmark shape must take the folded tcgen05 formulation"
    csc = CscWorkspace(V, batch.nnz, "cuda")
    csc.build(batch)
    dbeta = torch.zeros(K, V, device="cuda")
    loss = torch.zeros(1, dtype=torch.float64, device="cuda")
    rm, rv = torch.zeros(V, device="cuda"), torch.ones(V, device="cuda")
    rl, dtheta = dec.forward_backward(theta, theta.sum(0), beta_bf, betaT_bf, batch, csc, training=True,
                                      running_mean=rm, running_var=rv, dbeta=dbeta, loss=loss)
    torch.cuda.synchronize()
    rl_ref, dth_ref, db_ref = _oracle_decoder(theta, beta, sub)
    rel_loss = abs(loss.item() - rl_ref.sum().item()) / abs(rl_ref.sum().item())
    rel_rl = float(((rl.double() - rl_ref).abs() / rl_ref.abs()).max())
    dth = dtheta[:, :K].double()
    e_t = float((dth - dth_ref).abs().max() / dth_ref.abs().max())
    e_t_rms = float(torch.sqrt(((dth - dth_ref) ** 2).mean()) / torch.sqrt((dth_ref ** 2).mean()))
    e_b = float((dbeta - db_ref).abs().max() / db_ref.abs().max())
    e_b_rms = float(torch.sqrt(((dbeta - db_ref).double() ** 2).mean()) / torch.sqrt((db_ref.double() ** 2).mean()))
    report = {"shape": [B, V, K], "loss_rel_err": rel_loss, "rl_per_doc_max_rel_err": rel_rl, "dtheta_max_err_over_max": e_t,
              "dtheta_rms_rel_err": e_t_rms, "dbeta_max_err_over_max": e_b, "dbeta_rms_rel_err": e_b_rms}
    print("PARITY_DECODER " + json.dumps(report))
    _dump("parity_decoder_200k.json", report)
    # bf16 operands (8 mantissa bits): the budget quoted in DESIGN.md.  Measured: loss 7e-6, per-document RL
    # 1.6e-3, d theta 0.27 % rms / 5.6 % of max at the single worst entry, d beta 0.51 % rms / 12 % of max.  The
    # worst entries belong to the document with the largest topic weight (theta = 0.66): the BatchNorm-backward
    # projection subtracts two terms ~15-50x larger than their difference there, so the 2^-9 operand rounding is
    # amplified at those few entries while the rms stays at the operand precision.
    assert rel_loss < 1e-4 and rel_rl < 4e-3, report
    assert e_t < 0.10 and e_t_rms < 6e-3, report
    assert e_b < 0.20 and e_b_rms < 1e-2, report


def _dump(name, obj):
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    try:
        os.makedirs(out, exist_ok=True)
        with open(os.path.join(out, name), "w") as f:
            json.dump(obj, f, indent=1)
    except OSError:
        pass


def _eager_loss(model, x, eps, mask_enc, mask_theta):
    from gfedntm_b200.models.networks import elbo_terms

    inf = model.inf_net
    h = inf.hiddens(inf.activation(inf.input_layer(x))) * mask_enc
    mu = inf.f_mu_batchnorm(inf.f_mu(h))
    logvar = inf.f_sigma_batchnorm(inf.f_sigma(h))
    theta = F.softmax(mu + eps * torch.exp(0.5 * logvar), dim=1) * mask_theta
    kl, rl = elbo_terms(x, model.decode(theta), model.prior_mean, model.prior_variance, mu, torch.exp(logvar), logvar)
    return (kl + rl).sum()


def test_fused_training_tracks_eager_fp32_for_50_steps_at_benchmark_shape():
    from gfedntm_b200.engine.fused import FusedEngine
    from gfedntm_b200.models.networks import DecoderNetwork

    torch.manual_seed(0)
    n_steps, n_mb = 50, 2
    counts = _corpus(B * n_mb, seed=5)
    batches = [_csr_batch(counts, slice(i * B, (i + 1) * B)) for i in range(n_mb)]
    dense = [_dense_cols(sub, 0, V, torch.float32) for _, sub in batches]
    model = DecoderNetwork(V, K, "prodLDA", (50, 50), "softplus", 0.2, True)
    ref = copy.deepcopy(model).cuda().train()
    opt = torch.optim.Adam(ref.parameters(), lr=2e-3, betas=(0.99, 0.99))
    eng = FusedEngine(model, max_batch=B, max_nnz=max(b.nnz for b, _ in batches), lr=2e-3, betas=(0.99, 0.99), seed=1)
    g = torch.Generator(device="cuda").manual_seed(9)
    fused, eager = [], []
    for step in range(n_steps):
        eps = torch.randn(B, K, device="cuda", generator=g)
        me = (torch.rand(B, 50, device="cuda", generator=g) > 0.2).float() / 0.8
        mt = (torch.rand(B, K, device="cuda", generator=g) > 0.2).float() / 0.8
        batch, _ = batches[step % n_mb]
        # the path the benchmark runs: beta's gradient as bf16, its optimiser step overlapped with the d-theta sweep
        loss = eng.train_step(batch, eps_in=eps, mask_enc_in=me, mask_theta_in=mt)
        fused.append(float(loss.item()) / B)
        opt.zero_grad(set_to_none=True)
        ref_loss = _eager_loss(ref, dense[step % n_mb], eps, me, mt)
        ref_loss.backward()
        opt.step()
        eager.append(float(ref_loss.item()) / B)
    rel = [abs(a - b) / abs(b) for a, b in zip(fused, eager)]
    drift = float((model.beta.detach() - ref.beta.detach()).abs().max() / ref.beta.detach().abs().max())
    report = {"steps": n_steps, "loss_per_doc_fused_first_last": [fused[0], fused[-1]],
              "loss_per_doc_eager_first_last": [eager[0], eager[-1]], "max_rel_loss_diff": max(rel),
              "rel_loss_diff_last": rel[-1], "beta_max_drift_over_max": drift,
              "eager_loss_decrease": eager[0] - eager[-1]}
    print("PARITY_TRAINING " + json.dumps(report))
    _dump("parity_training_200k.json", report)
    assert all(l == l for l in fused)
    assert eager[-1] < eager[0], "the oracle itself must be learning"
    # the two curves stay within a small fraction of the distance the loss travels in 50 steps
    assert max(rel) < 2e-3, report
    assert abs(fused[-1] - eager[-1]) < 0.05 * abs(eager[0] - eager[-1]) + 1e-3 * abs(eager[-1]), report
