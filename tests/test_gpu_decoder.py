"""GPU numerics of the tcgen05 decoder sweeps against plain PyTorch fp32 references."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _mk(B, V, K, seed=0, nnz_per_doc=40):
    from gfedntm_b200.ops import kernels as Kx
    from gfedntm_b200.ops.decoder import CsrBatch, make_beta_shadows, refresh_beta_shadows

    g = torch.Generator(device="cuda").manual_seed(seed)
    dev = "cuda"
    theta = torch.softmax(torch.randn(B, K, device=dev, generator=g) * 1.5, dim=1)
    beta = (torch.rand(K, V, device=dev, generator=g) - 0.5) * 0.6
    x = torch.zeros(B, V, device=dev)
    idx = torch.randint(0, V, (B, nnz_per_doc), device=dev, generator=g)
    x.scatter_add_(1, idx, torch.ones(B, nnz_per_doc, device=dev))
    beta_bf, betaT_bf = make_beta_shadows(K, V, dev)
    refresh_beta_shadows(beta, beta_bf, betaT_bf)
    return theta, beta, x, beta_bf, betaT_bf, CsrBatch.from_dense(x), Kx


@pytest.mark.parametrize("B,V,K", [(64, 1000, 10), (200, 5000, 50), (300, 2000, 100), (130, 3000, 200)])
@pytest.mark.parametrize("mode", ["row", "col"])
def test_logits_gemm(B, V, K, mode):
    theta, beta, x, beta_bf, betaT_bf, batch, Kx = _mk(B, V, K)
    Kp = beta_bf.shape[0]
    th_bf = torch.zeros(B, Kp, dtype=torch.bfloat16, device="cuda")
    th_bf[:, :K] = theta.to(torch.bfloat16)
    thT_bf = torch.zeros(Kp, Kx.round_up(B, 8), dtype=torch.bfloat16, device="cuda")
    thT_bf[:, :B] = th_bf.t()
    out = torch.full((B, V), float("nan"), device="cuda")
    Vpad = Kx.round_up(V, 128)
    Bpad = Kx.round_up(B, 128)
    sb = torch.zeros(max(Vpad, Bpad), 4, device="cuda")
    Kx.dec_sweep(Kx.ROW_LOGITS if mode == "row" else Kx.COL_LOGITS, B=B, V=V, K=K, theta_bf=thT_bf,
                 beta_bf=beta_bf, sb_stream=sb, n_chunks=min(3, (V + 127) // 128), out=out)
    torch.cuda.synchronize()
    ref = th_bf[:, :K].float() @ beta_bf[:K, :V].float()
    assert torch.isfinite(out).all()
    torch.testing.assert_close(out, ref, rtol=1e-3, atol=1e-4)
    # the bf16 transposed shadow must agree with the row-major one
    torch.testing.assert_close(betaT_bf[:, :K].float(), beta_bf[:K, :V].float().t())


def _reference(theta, beta_bf, x, K, V, eps=1e-5):
    thc = (theta - theta.mean(0, keepdim=True)).to(torch.bfloat16).float().requires_grad_(True)
    bb = beta_bf[:K, :V].float().clone().requires_grad_(True)
    logits = thc @ bb
    y = F.batch_norm(logits, None, None, training=True, eps=eps)
    p = torch.softmax(y, dim=1)
    rl = -(x * torch.log(p + 1e-10)).sum(1)
    rl.sum().backward()
    var = logits.detach().var(0, unbiased=False)
    return rl.detach(), thc.grad, bb.grad, var


@pytest.mark.parametrize("folded", [True, False])
@pytest.mark.parametrize("B,V,K", [(64, 1000, 10), (200, 5000, 50), (257, 2100, 100), (130, 3000, 200), (300, 1500, 13)])
def test_prodlda_decoder_fwd_bwd(B, V, K, folded):
    from gfedntm_b200.ops.decoder import CscWorkspace, ProdLDADecoder

    theta, beta, x, beta_bf, betaT_bf, batch, Kx = _mk(B, V, K, seed=1)
    dec = ProdLDADecoder(B, V, K, batch.nnz, "cuda", folded=folded)
    assert dec.folded == folded
    csc = CscWorkspace(V, batch.nnz, "cuda")
    csc.build(batch)
    dbeta = torch.zeros(K, V, device="cuda")
    loss = torch.zeros(1, dtype=torch.float64, device="cuda")
    rm = torch.zeros(V, device="cuda")
    rv = torch.ones(V, device="cuda")
    rl, dtheta = dec.forward_backward(theta, theta.sum(0), beta_bf, betaT_bf, batch, csc, training=True,
                                      running_mean=rm, running_var=rv, dbeta=dbeta, loss=loss)
    torch.cuda.synchronize()
    rl_ref, dth_ref, db_ref, var_ref = _reference(theta, beta_bf, x, K, V)
    # BatchNorm statistics
    rstd = dec.colc[:V, 0] * 0.6931471805599453
    torch.testing.assert_close(rstd, torch.rsqrt(var_ref + 1e-5), rtol=2e-3, atol=1e-6)
    torch.testing.assert_close(rl, rl_ref, rtol=2e-3, atol=1e-2)
    assert abs(loss.item() - rl_ref.sum().item()) <= 2e-3 * abs(rl_ref.sum().item())
    scale_t = dth_ref.abs().max().item()
    scale_b = db_ref.abs().max().item()
    err_t = (dtheta[:, :K] - dth_ref).abs().max().item()
    err_b = (dbeta - db_ref).abs().max().item()
    assert err_t <= 2e-2 * scale_t, (err_t, scale_t)
    assert err_b <= 2e-2 * scale_b, (err_b, scale_b)
    # running statistics (momentum 0.1, unbiased variance)
    mean_ref = (theta.mean(0, keepdim=True) @ beta_bf[:K, :V].float()).squeeze(0)
    torch.testing.assert_close(rm, 0.1 * mean_ref, rtol=1e-2, atol=1e-4)
    torch.testing.assert_close(rv, 0.9 + 0.1 * var_ref * B / (B - 1), rtol=1e-2, atol=1e-6)


def test_csr_to_csc_roundtrip():
    theta, beta, x, beta_bf, betaT_bf, batch, Kx = _mk(77, 999, 10, seed=3)
    from gfedntm_b200.ops.decoder import CscWorkspace

    csc = CscWorkspace(999, batch.nnz, "cuda")
    csc.build(batch)
    torch.cuda.synchronize()
    colptr = csc.colptr.cpu()
    assert colptr[0] == 0 and colptr[-1] == batch.nnz
    counts_per_col = (x > 0).sum(0).cpu()
    assert torch.equal(colptr[1:] - colptr[:-1], counts_per_col.to(torch.int32))
    # every CSC entry points back at a CSR entry of the right column / row
    src = csc.src[: batch.nnz].long()
    assert torch.equal(torch.sort(src).values.cpu(), torch.arange(batch.nnz))
    cols = torch.repeat_interleave(torch.arange(999), (colptr[1:] - colptr[:-1]).long()).cuda()
    assert torch.equal(batch.indices.long()[src], cols)
    rows_of_src = torch.repeat_interleave(torch.arange(77, device="cuda"),
                                          (batch.indptr[1:] - batch.indptr[:-1]).long())
    assert torch.equal(rows_of_src[src], csc.row[: batch.nnz].long())


@pytest.mark.parametrize("folded", [True, False])
@pytest.mark.parametrize("B,V,K", [(96, 1300, 20), (257, 2100, 100)])
def test_prodlda_decoder_eval_mode(B, V, K, folded):
    """Inference path: BatchNorm with running statistics (non-zero bias rows in the folded operands)."""
    from gfedntm_b200.ops.decoder import ProdLDADecoder

    theta, beta, x, beta_bf, betaT_bf, batch, Kx = _mk(B, V, K, seed=5)
    g = torch.Generator(device="cuda").manual_seed(11)
    rm = torch.randn(V, device="cuda", generator=g) * 0.05
    rv = torch.rand(V, device="cuda", generator=g) * 0.02 + 0.005
    dec = ProdLDADecoder(B, V, K, batch.nnz, "cuda", folded=folded)
    loss = torch.zeros(1, dtype=torch.float64, device="cuda")
    rm0, rv0 = rm.clone(), rv.clone()
    rl, dth = dec.forward_backward(theta, theta.sum(0), beta_bf, betaT_bf, batch, None, training=False,
                                   running_mean=rm, running_var=rv, dbeta=None, loss=loss)
    torch.cuda.synchronize()
    assert dth is None
    torch.testing.assert_close(rm, rm0)  # eval mode must not touch the running statistics
    torch.testing.assert_close(rv, rv0)
    logits = theta.to(torch.bfloat16).float() @ beta_bf[:K, :V].float()
    y = F.batch_norm(logits, rm0, rv0, training=False, eps=1e-5)
    rl_ref = -(x * torch.log(torch.softmax(y, 1) + 1e-10)).sum(1)
    torch.testing.assert_close(rl, rl_ref, rtol=3e-3, atol=2e-2)
    assert abs(loss.item() - rl_ref.sum().item()) <= 3e-3 * abs(rl_ref.sum().item())


def test_csr_to_csc_large_vocabulary():
    """Many scan segments (V = 300k -> 74 blocks of 4096) and the dynamic-nnz variant."""
    from gfedntm_b200.ops.decoder import CscWorkspace, CsrBatch

    B, V, nnz_doc = 512, 300_001, 64
    g = torch.Generator(device="cuda").manual_seed(9)
    idx = torch.randint(0, V, (B, nnz_doc), device="cuda", generator=g).sort(1).values.to(torch.int32)
    indptr = (torch.arange(B + 1, device="cuda") * nnz_doc).to(torch.int32)
    batch = CsrBatch(indptr, idx.reshape(-1).contiguous(), torch.ones(B * nnz_doc, device="cuda"), B, V)
    for dyn in (False, True):
        csc = CscWorkspace(V, batch.nnz + (1000 if dyn else 0), "cuda")
        csc.build(batch, dynamic_nnz=dyn)
        torch.cuda.synchronize()
        ref = torch.zeros(V + 1, dtype=torch.int64, device="cuda")
        ref[1:] = torch.cumsum(torch.bincount(batch.indices.long(), minlength=V), 0)
        assert torch.equal(csc.colptr.long(), ref)
        src = csc.src[: batch.nnz].long()
        assert torch.equal(torch.sort(src).values, torch.arange(batch.nnz, device="cuda"))
        cols = torch.repeat_interleave(torch.arange(V, device="cuda"), (ref[1:] - ref[:-1]))
        assert torch.equal(batch.indices.long()[src], cols)


@pytest.mark.parametrize("V,K", [(1000, 10), (70_000, 200), (5000, 124)])
def test_gram_gemm_matches_torch(V, K):
    """Split-V tcgen05 Gram GEMM (+ finish): M = X Y^T summed over the CTAs' partials, stored as bf16 hi/lo."""
    from gfedntm_b200.ops import kernels as Kx

    Kp = Kx.padded_topics(K)
    g = torch.Generator(device="cuda").manual_seed(4)
    ld = Kx.round_up(V, 8)
    x = torch.zeros(Kp, ld, dtype=torch.bfloat16, device="cuda")
    y = torch.zeros(Kp, ld, dtype=torch.bfloat16, device="cuda")
    x[: K + 1, :V] = (torch.randn(K + 1, V, device="cuda", generator=g) * 0.3).to(torch.bfloat16)
    y[:K, :V] = (torch.randn(K, V, device="cuda", generator=g) * 0.3).to(torch.bfloat16)
    parts = Kx.gram_parts()
    partial = torch.full((parts * Kp * Kp,), float("nan"), device="cuda")
    n_ext = (Kp + 127) // 128
    out = torch.zeros(Kp, 2 * n_ext * 128, dtype=torch.bfloat16, device="cuda")
    Kx.gram(x, y, partial, V=V)
    Kx.gram_finish(partial, parts, out, K=K, col_hi=0, col_lo=n_ext * 128)
    torch.cuda.synchronize()
    ref = x[: K + 1, :V].float() @ y[:K, :V].float().t()
    got = out[: K + 1, :K].float() + out[: K + 1, n_ext * 128: n_ext * 128 + K].float()
    torch.testing.assert_close(got, ref, rtol=1e-4, atol=1e-4 * ref.abs().max().item())
    assert torch.count_nonzero(out[K + 1:]) == 0


def test_csc_corr_transposed_matches_row_major():
    from gfedntm_b200.ops import kernels as Kx
    from gfedntm_b200.ops.decoder import CscWorkspace

    theta, beta, x, beta_bf, betaT_bf, batch, _ = _mk(150, 1234, 50, seed=8)
    Kp = beta_bf.shape[0]
    csc = CscWorkspace(1234, batch.nnz, "cuda")
    csc.build(batch)
    th_bf = torch.zeros(150, Kp, dtype=torch.bfloat16, device="cuda")
    th_bf[:, :50] = theta.to(torch.bfloat16)
    tn = torch.rand(batch.nnz, device="cuda")
    corr = torch.zeros(1234, Kp, device="cuda")
    corrT = torch.full((Kp, Kx.round_up(1234, 8)), 7.0, device="cuda")
    Kx.csc_corr(csc.colptr, csc.row, csc.src, tn, th_bf, corr, V=1234)
    Kx.csc_corr_t(csc.colptr, csc.row, csc.src, tn, th_bf, corrT, V=1234)
    torch.cuda.synchronize()
    nonempty = (csc.colptr[1:] > csc.colptr[:-1])
    torch.testing.assert_close(corrT[:, :1234].t()[nonempty], corr[nonempty])
    assert (corrT[:, :1234].t()[~nonempty] == 7.0).all()  # words without entries are left untouched
