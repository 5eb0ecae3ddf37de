"""Torch-tensor front end of the C-ABI kernels.

Every function takes CUDA tensors, checks shapes/dtypes, fills the argument struct and launches
on the current CUDA stream.  There is no CPU or PyTorch fallback in this module.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import torch

from . import _lib
from ._lib import (GF_MAX_HIDDEN, GfDecSparseArgs, GfEncArgs, GfEncBwdArgs, GfLatentArgs,
                   GfLatentBwdArgs, GfSweepArgs, check)

(ROW_LOGITS, ROW_LSE, ROW_GRAD, COL_STATS, COL_GRAD, COL_LOGITS, ROW_LSE2, COL_GRAD2, ROW_GRAD2,
 COL_STATS2) = range(10)
FOLD_ROWS = 4  # spare operand rows above K: bias hi/lo and the -e_b hi/lo pair of the folded sweeps
TILE = 128
LOG2E = 1.4426950408889634
LN2 = 0.6931471805599453

# solver -> (GfSolver id, b1 = momentum?, b2, eps): torch defaults of the reference's optimiser factory (avitm.py:141-153)
SOLVERS = {"adam": 0, "sgd": 1, "adagrad": 2, "adadelta": 3, "rmsprop": 4}


def solver_constants(solver: str, momentum: float):
    """(id, b1, b2, eps) of a solver: Adam(betas=(momentum, 0.99)), SGD(momentum), Adagrad(eps 1e-10), Adadelta(rho 0.9,
    eps 1e-6), RMSprop(alpha 0.99, eps 1e-8, momentum)."""
    return {"adam": (0, momentum, 0.99, 1e-8), "sgd": (1, momentum, 0.0, 0.0), "adagrad": (2, 0.0, 0.0, 1e-10),
            "adadelta": (3, 0.0, 0.9, 1e-6), "rmsprop": (4, momentum, 0.99, 1e-8)}[solver]


ACTIVATIONS = {"softplus": 0, "relu": 1, "sigmoid": 2, "tanh": 3, "leakyrelu": 4, "rrelu": 5,
               "elu": 6, "selu": 7}


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    if t is None:
        return None
    assert t.is_cuda, "kernel arguments must be CUDA tensors"
    return t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


def padded_topics(k: int) -> int:
    """Topics plus the four constant rows of the folded sweeps, padded to the UMMA K granularity (16)."""
    return round_up(k + FOLD_ROWS, 16)


def num_sms() -> int:
    return int(_lib.load().gf_num_sms())


def default_row_chunks(B: int, V: int, second_mma: bool = False, Kp: int = 64, grid: Optional[int] = None) -> int:
    """Vocabulary split of the ROW sweeps: smallest chunk count that fills the persistent grid in
    whole waves (computed by the library, which knows the grid each mode launches).  ``grid``: number of
    sweep CTAs when some are set aside (overlapped aggregation)."""
    lib = _lib.load()
    if grid is not None:
        return int(lib.gf_dec_row_chunks_grid(B, V, grid))
    return int(lib.gf_dec_row_chunks(ROW_GRAD if second_mma else ROW_LSE, B, V, Kp))


# ----------------------------------------------------------------------------------------------
# decoder sweeps
# ----------------------------------------------------------------------------------------------
def dec_sweep(mode: int, *, B: int, V: int, K: int, theta_bf: torch.Tensor, beta_bf: torch.Tensor,
              sb_stream: Optional[torch.Tensor] = None, n_chunks: int = 1, out=None, lse_part=None,
              doc_const=None, dtheta=None, colc=None, theta_mean=None, running_mean=None, running_var=None,
              bn_eps: float = 1e-5, bn_momentum: float = 0.1, use_running: bool = False, asp=None,
              csp=None, dbeta=None, corr=None, colptr=None, theta_cols: int = 0, beta_cols: int = 0,
              ext_col0: int = 0, betaT_bf=None, beta3=None, fed: int = 0, fed_ctas: int = 0, fed_tmaps: int = 0,
              dbeta_bf=None) -> None:
    lib = _lib.load()
    Kp = beta_bf.shape[0]
    assert theta_bf.dtype == torch.bfloat16 and beta_bf.dtype == torch.bfloat16
    # theta_bf is the TRANSPOSED centred theta: [Kp, ld >= B], documents contiguous
    assert theta_bf.shape[0] == Kp and theta_bf.shape[1] >= B and theta_bf.stride(1) == 1
    assert beta_bf.stride(1) == 1
    a = GfSweepArgs()
    a.mode, a.B, a.V, a.K, a.Kp, a.n_chunks = mode, B, V, K, Kp, n_chunks
    a.theta_bf, a.ld_theta = _ptr(theta_bf), theta_bf.stride(0)
    a.beta_bf, a.ld_beta = _ptr(beta_bf), beta_bf.stride(0)
    a.sb_stream = _ptr(sb_stream)
    a.out, a.lse_part, a.doc_const, a.dtheta = _ptr(out), _ptr(lse_part), _ptr(doc_const), _ptr(dtheta)
    a.colc, a.theta_mean = _ptr(colc), _ptr(theta_mean)
    a.running_mean, a.running_var = _ptr(running_mean), _ptr(running_var)
    a.bn_eps, a.bn_momentum, a.use_running = bn_eps, bn_momentum, int(use_running)
    a.asp, a.csp = _ptr(asp), _ptr(csp)
    a.dbeta, a.ld_dbeta = _ptr(dbeta), (dbeta.stride(0) if dbeta is not None else 0)
    a.corr, a.colptr = _ptr(corr), _ptr(colptr)
    a.theta_cols, a.beta_cols, a.ext_col0 = theta_cols, beta_cols, ext_col0
    a.betaT_bf, a.beta3 = _ptr(betaT_bf), _ptr(beta3)
    a.ld_beta3 = beta3.stride(0) if beta3 is not None else 0
    a.ld_corr = corr.stride(0) if corr is not None else 0
    a.fed, a.fed_ctas, a.fed_tmaps = (fed or None), fed_ctas, (fed_tmaps or None)
    a.dbeta_bf = _ptr(dbeta_bf)
    a.ld_dbeta_bf = dbeta_bf.stride(0) if dbeta_bf is not None else 0
    check(lib.gf_dec_sweep(C.byref(a), _stream()), "gf_dec_sweep")


# ----------------------------------------------------------------------------------------------
# encoder / latent
# ----------------------------------------------------------------------------------------------
def _fill_ptr_array(arr, tensors: Sequence[Optional[torch.Tensor]]) -> None:
    for i, t in enumerate(tensors):
        arr[i] = _ptr(t)


def enc_fwd(*, B: int, V: int, K: int, hidden: Sequence[int], act: int, training: bool,
            indptr, indices, counts, W1T, b1, xdense, WdT, Wh, bh, Wmu, bmu, Wsig, bsig, pre,
            enc_drop_p: float, mask_enc_in, mask_enc, mu_pre, ls_pre, bn_sums, seed: int, step_ctr=None,
            pre0_ext=None) -> None:
    lib = _lib.load()
    a = GfEncArgs()
    a.B, a.V, a.K, a.L = B, V, K, len(hidden)
    for i, h in enumerate(hidden):
        a.H[i] = h
    a.act, a.training = act, int(training)
    a.indptr, a.indices, a.counts = _ptr(indptr), _ptr(indices), _ptr(counts)
    a.W1T, a.b1 = _ptr(W1T), _ptr(b1)
    a.xdense, a.C, a.WdT = _ptr(xdense), (xdense.shape[1] if xdense is not None else 0), _ptr(WdT)
    _fill_ptr_array(a.Wh, Wh)
    _fill_ptr_array(a.bh, bh)
    a.Wmu, a.bmu, a.Wsig, a.bsig = _ptr(Wmu), _ptr(bmu), _ptr(Wsig), _ptr(bsig)
    _fill_ptr_array(a.pre, pre)
    a.enc_drop_p = enc_drop_p
    a.mask_enc_in, a.mask_enc = _ptr(mask_enc_in), _ptr(mask_enc)
    a.mu_pre, a.ls_pre, a.bn_sums = _ptr(mu_pre), _ptr(ls_pre), _ptr(bn_sums)
    a.seed = seed & 0xFFFFFFFFFFFFFFFF
    a.step_ctr = _ptr(step_ctr)
    a.pre0_ext = _ptr(pre0_ext)
    check(lib.gf_enc_fwd(C.byref(a), _stream()), "gf_enc_fwd")


def latent_fwd(*, B: int, K: int, training: bool, sample: bool, mu_pre, ls_pre, bn_sums, mu_rm,
               mu_rv, ls_rm, ls_rv, bn_eps: float, bn_momentum: float, prior_mean, prior_var, eps_in,
               mask_theta_in, theta_drop_p: float, mu, logvar, eps, theta_s, mask_theta, theta,
               theta_colsum, kl, loss, kl_weight: float, seed: int, step_ctr=None) -> None:
    lib = _lib.load()
    a = GfLatentArgs()
    a.B, a.K, a.training, a.sample = B, K, int(training), int(sample)
    a.mu_pre, a.ls_pre, a.bn_sums = _ptr(mu_pre), _ptr(ls_pre), _ptr(bn_sums)
    a.mu_rm, a.mu_rv, a.ls_rm, a.ls_rv = _ptr(mu_rm), _ptr(mu_rv), _ptr(ls_rm), _ptr(ls_rv)
    a.bn_eps, a.bn_momentum = bn_eps, bn_momentum
    a.prior_mean, a.prior_var = _ptr(prior_mean), _ptr(prior_var)
    a.eps_in, a.mask_theta_in, a.theta_drop_p = _ptr(eps_in), _ptr(mask_theta_in), theta_drop_p
    a.mu, a.logvar, a.eps, a.theta_s = _ptr(mu), _ptr(logvar), _ptr(eps), _ptr(theta_s)
    a.mask_theta, a.theta = _ptr(mask_theta), _ptr(theta)
    a.theta_colsum, a.kl, a.loss, a.kl_weight = _ptr(theta_colsum), _ptr(kl), _ptr(loss), kl_weight
    a.seed = seed & 0xFFFFFFFFFFFFFFFF
    a.step_ctr = _ptr(step_ctr)
    check(lib.gf_latent_fwd(C.byref(a), _stream()), "gf_latent_fwd")


def theta_center(theta, colsum, out_bf16, outT_bf16, mean_out, *, B: int, K: int, center: bool) -> None:
    """theta_c as bf16 in both layouts: out [B, Kp] and outT [Kp, ldT]."""
    lib = _lib.load()
    check(lib.gf_theta_center(_ptr(theta), _ptr(colsum), _ptr(out_bf16), _ptr(outT_bf16), _ptr(mean_out), B, K,
                              out_bf16.stride(0), outT_bf16.stride(0), int(center), _stream()), "gf_theta_center")


def theta_center_fold(theta, colsum, out_bf16, outT_bf16, mean_out, *, B: int, K: int, Bpad: int,
                      center: bool) -> None:
    """theta_center plus the constant rows of the folded sweeps (ones, zeroed e rows, padded-doc mask)."""
    lib = _lib.load()
    check(lib.gf_theta_center_fold(_ptr(theta), _ptr(colsum), _ptr(out_bf16), _ptr(outT_bf16), _ptr(mean_out),
                                   B, K, out_bf16.stride(0), outT_bf16.stride(0), Bpad, int(center), _stream()),
          "gf_theta_center_fold")


def scale_beta(beta_bf, colc, out, *, K: int, V: int) -> None:
    """out = bf16(beta * rstd*log2e) + bias / ones rows (operand of the folded sweeps)."""
    lib = _lib.load()
    check(lib.gf_scale_beta(_ptr(beta_bf), beta_bf.stride(0), _ptr(colc), _ptr(out), out.stride(0), K,
                            out.shape[0], V, out.shape[1], _stream()), "gf_scale_beta")


def theta_gram(theta_bf, thetaT_bf, c_scratch, *, B: int, K: int, ext_col0: int, lo_col0: int) -> None:
    """C = theta_c^T theta_c as bf16 hi/lo pseudo-document blocks of theta_c^T (c_scratch stays zeroed)."""
    lib = _lib.load()
    check(lib.gf_theta_gram(_ptr(theta_bf), theta_bf.stride(0), B, K, _ptr(thetaT_bf), thetaT_bf.stride(0),
                            ext_col0, lo_col0, _ptr(c_scratch), _stream()), "gf_theta_gram")


def gram_parts() -> int:
    return int(_lib.load().gf_gram_parts())


def gram(x_bf, y_bf, partial, *, V: int) -> None:
    lib = _lib.load()
    check(lib.gf_gram(_ptr(x_bf), x_bf.stride(0), _ptr(y_bf), y_bf.stride(0), V, x_bf.shape[0], _ptr(partial),
                      _stream()), "gf_gram")


def gram_finish(partial, n_parts: int, out_bf, *, K: int, col_hi: int, col_lo: int) -> None:
    """Sums the Gram partials into M (row K = a) and stores it as bf16 hi/lo pseudo-word blocks of ``out_bf``."""
    lib = _lib.load()
    check(lib.gf_gram_finish(_ptr(partial), n_parts, K, out_bf.shape[0], _ptr(out_bf), out_bf.stride(0), col_hi,
                             col_lo, _stream()), "gf_gram_finish")


def latent_bwd(*, B: int, K: int, dtheta, mask_theta, theta_s, mu, logvar, eps, prior_mean,
               prior_var, kl_weight: float, gscale: float, dmu, dlv, bn_bsums, d_prior_mean,
               d_prior_var) -> None:
    lib = _lib.load()
    a = GfLatentBwdArgs()
    a.B, a.K = B, K
    a.dtheta, a.ld_dtheta = _ptr(dtheta), dtheta.stride(0)
    a.mask_theta, a.theta_s, a.mu, a.logvar, a.eps = (_ptr(mask_theta), _ptr(theta_s), _ptr(mu),
                                                       _ptr(logvar), _ptr(eps))
    a.prior_mean, a.prior_var = _ptr(prior_mean), _ptr(prior_var)
    a.kl_weight, a.gscale = kl_weight, gscale
    a.dmu, a.dlv, a.bn_bsums = _ptr(dmu), _ptr(dlv), _ptr(bn_bsums)
    a.d_prior_mean, a.d_prior_var = _ptr(d_prior_mean), _ptr(d_prior_var)
    check(lib.gf_latent_bwd(C.byref(a), _stream()), "gf_latent_bwd")


def enc_bwd(*, B: int, V: int, K: int, hidden: Sequence[int], act: int, dmu, dlv, mu, logvar,
            bn_sums, bn_bsums, bn_eps: float, pre, mask_enc, Wh, Wmu, Wsig, dWh, dbh, dWmu, dbmu,
            dWsig, dbsig, db1, dpre0, xdense, dWdT, seed: int, step_ctr=None) -> None:
    lib = _lib.load()
    a = GfEncBwdArgs()
    a.B, a.V, a.K, a.L = B, V, K, len(hidden)
    for i, h in enumerate(hidden):
        a.H[i] = h
    a.act = act
    a.dmu, a.dlv, a.mu, a.logvar = _ptr(dmu), _ptr(dlv), _ptr(mu), _ptr(logvar)
    a.bn_sums, a.bn_bsums, a.bn_eps = _ptr(bn_sums), _ptr(bn_bsums), bn_eps
    _fill_ptr_array(a.pre, pre)
    a.mask_enc = _ptr(mask_enc)
    _fill_ptr_array(a.Wh, Wh)
    a.Wmu, a.Wsig = _ptr(Wmu), _ptr(Wsig)
    _fill_ptr_array(a.dWh, dWh)
    _fill_ptr_array(a.dbh, dbh)
    a.dWmu, a.dbmu, a.dWsig, a.dbsig, a.db1 = _ptr(dWmu), _ptr(dbmu), _ptr(dWsig), _ptr(dbsig), _ptr(db1)
    a.dpre0 = _ptr(dpre0)
    a.xdense, a.C, a.dWdT = _ptr(xdense), (xdense.shape[1] if xdense is not None else 0), _ptr(dWdT)
    a.seed = seed & 0xFFFFFFFFFFFFFFFF
    a.step_ctr = _ptr(step_ctr)
    check(lib.gf_enc_bwd(C.byref(a), _stream()), "gf_enc_bwd")


# ----------------------------------------------------------------------------------------------
# sparse side
# ----------------------------------------------------------------------------------------------
def csr_to_csc(indptr, indices, *, B: int, V: int, nnz: int, colptr, cursor, csc_row, csc_src,
               dynamic_nnz: bool = False) -> None:
    """``dynamic_nnz``: ``nnz`` is only a capacity; the real count is read from ``indptr[B]`` on the device
    (needed when the launch is replayed from a CUDA graph on other minibatches)."""
    lib = _lib.load()
    fn = lib.gf_csr_to_csc_dyn if dynamic_nnz else lib.gf_csr_to_csc
    check(fn(_ptr(indptr), _ptr(indices), B, V, nnz, _ptr(colptr), _ptr(cursor), _ptr(csc_row), _ptr(csc_src),
             _stream()), "gf_csr_to_csc")


def step_begin_zero(ctr, inc_rng: int, inc_opt: int, buffers) -> None:
    """``step_begin`` and the zero fill of ``buffers`` (contiguous tensors, at most 6) in one launch."""
    import ctypes as C

    bufs = [b for b in buffers if b is not None and b.numel() > 0]
    assert len(bufs) <= 6 and all(b.is_contiguous() for b in bufs)
    ptrs = (C.c_void_p * max(len(bufs), 1))(*[b.data_ptr() for b in bufs])
    nbytes = (C.c_longlong * max(len(bufs), 1))(*[b.numel() * b.element_size() for b in bufs])
    check(_lib.load().gf_step_begin_zero(_ptr(ctr), inc_rng, inc_opt, ptrs, nbytes, len(bufs), _stream()),
          "gf_step_begin_zero")


def step_begin(ctr: torch.Tensor, inc_rng: int = 1, inc_opt: int = 0) -> None:
    lib = _lib.load()
    check(lib.gf_step_begin(_ptr(ctr), inc_rng, inc_opt, _stream()), "gf_step_begin")


def dec_sparse_doc(*, B: int, n_chunks: int, indptr, indices, counts, theta_bf, betaT_bf, colc,
                   lse_part, gscale: float, rl, loss, asp, csp, tnnz, dtheta, doc_const, sb_doc,
                   thetaT_bf=None, K: int = 0) -> None:
    lib = _lib.load()
    a = GfDecSparseArgs()
    a.B, a.Kp, a.n_chunks = B, betaT_bf.shape[1], n_chunks
    a.indptr, a.indices, a.counts = _ptr(indptr), _ptr(indices), _ptr(counts)
    a.theta_bf, a.ld_theta = _ptr(theta_bf), theta_bf.stride(0)
    a.betaT_bf, a.colc, a.lse_part = _ptr(betaT_bf), _ptr(colc), _ptr(lse_part)
    a.gscale = gscale
    a.rl, a.loss, a.asp, a.csp, a.tnnz = _ptr(rl), _ptr(loss), _ptr(asp), _ptr(csp), _ptr(tnnz)
    a.dtheta, a.doc_const, a.sb_doc = _ptr(dtheta), _ptr(doc_const), _ptr(sb_doc)
    a.thetaT_bf = _ptr(thetaT_bf)
    a.ldT, a.K = (thetaT_bf.stride(0) if thetaT_bf is not None else 0), K
    check(lib.gf_dec_sparse_doc(C.byref(a), _stream()), "gf_dec_sparse_doc")


def csc_corr(colptr, csc_row, csc_src, tnnz, theta_bf, corr, *, V: int) -> None:
    lib = _lib.load()
    check(lib.gf_csc_corr(_ptr(colptr), _ptr(csc_row), _ptr(csc_src), _ptr(tnnz), _ptr(theta_bf),
                          theta_bf.stride(0), _ptr(corr), V, corr.shape[1], _stream()), "gf_csc_corr")


def csc_corr_t(colptr, csc_row, csc_src, tnnz, theta_bf, corrT, *, V: int) -> None:
    """csc_corr with the result transposed: corrT [Kp, ldc >= V]."""
    lib = _lib.load()
    check(lib.gf_csc_corr_t(_ptr(colptr), _ptr(csc_row), _ptr(csc_src), _ptr(tnnz), _ptr(theta_bf),
                            theta_bf.stride(0), _ptr(corrT), corrT.stride(0), V, corrT.shape[0], _stream()),
          "gf_csc_corr_t")


def csc_spmm(colptr, csc_row, csc_src, val, dense, out, *, V: int, N: int, accumulate: bool = False) -> None:
    lib = _lib.load()
    check(lib.gf_csc_spmm(_ptr(colptr), _ptr(csc_row), _ptr(csc_src), _ptr(val), _ptr(dense),
                          dense.stride(0), _ptr(out), out.stride(0), V, N, int(accumulate), _stream()),
          "gf_csc_spmm")


def lda_beta_fwd(beta, *, K: int, V: int, eps: float, momentum: float, training: bool, running_mean,
                 running_var, xhat, col_rstd, bp) -> None:
    lib = _lib.load()
    check(lib.gf_lda_beta_fwd(_ptr(beta), beta.stride(0), K, V, eps, momentum, int(training),
                              _ptr(running_mean), _ptr(running_var), _ptr(xhat), _ptr(col_rstd), _ptr(bp),
                              _stream()), "gf_lda_beta_fwd")


def lda_doc(indptr, indices, counts, theta, bp, *, B: int, K: int, V: int, gscale: float, rl, loss,
            dtheta, gnnz) -> None:
    lib = _lib.load()
    check(lib.gf_lda_doc(_ptr(indptr), _ptr(indices), _ptr(counts), _ptr(theta), _ptr(bp), B, K, V, gscale,
                         _ptr(rl), _ptr(loss), _ptr(dtheta), (dtheta.stride(0) if dtheta is not None else 0),
                         _ptr(gnnz), _stream()), "gf_lda_doc")


def lda_beta_bwd(bp, dbpT, xhat, col_rstd, dxhat, dbeta, *, K: int, V: int, training: bool) -> None:
    lib = _lib.load()
    check(lib.gf_lda_beta_bwd(_ptr(bp), _ptr(dbpT), dbpT.stride(0), _ptr(xhat), _ptr(col_rstd), _ptr(dxhat),
                              _ptr(dbeta), dbeta.stride(0), K, V, int(training), _stream()), "gf_lda_beta_bwd")


def row_topk(mat: torch.Tensor, k: int):
    """Indices and values of the k largest entries of every row (k <= 32), sorted descending."""
    lib = _lib.load()
    rows, V = mat.shape
    idx = torch.empty(rows, k, dtype=torch.int32, device=mat.device)
    val = torch.empty(rows, k, dtype=torch.float32, device=mat.device)
    mat = mat.float()
    if mat.stride(1) != 1:
        mat = mat.contiguous()
    check(lib.gf_row_topk(_ptr(mat), mat.stride(0), rows, V, k, _ptr(idx), _ptr(val), _stream()), "gf_row_topk")
    return val, idx.long()


# ----------------------------------------------------------------------------------------------
# optimiser
# ----------------------------------------------------------------------------------------------
def adam_flat(p, g, m, v, *, lr: float, b1: float, b2: float, eps: float, wd: float, step: int,
              gscale: float = 1.0, step_ctr=None, solver: int = 0) -> None:
    lib = _lib.load()
    n = p.numel()
    assert g.numel() == n and m.numel() == n and v.numel() == n
    check(lib.gf_adam_flat(_ptr(p), _ptr(g), _ptr(m), _ptr(v), n, lr, b1, b2, eps, wd, step, gscale,
                           _ptr(step_ctr), solver, _stream()), "gf_adam_flat")


def adam_beta(p, g, m, v, beta_bf, betaT_bf, *, K: int, V: int, lr: float = 0.0, b1: float = 0.9,
              b2: float = 0.999, eps: float = 1e-8, wd: float = 0.0, step: int = 1, gscale: float = 1.0,
              do_update: bool = True, step_ctr=None, solver: int = 0) -> None:
    lib = _lib.load()
    check(lib.gf_adam_beta(_ptr(p), _ptr(g), _ptr(m), _ptr(v), K, V, p.stride(0),
                           (g.stride(0) if g is not None else 0), _ptr(beta_bf), beta_bf.stride(0),
                           _ptr(betaT_bf), betaT_bf.shape[1], lr, b1, b2, eps, wd, step, gscale,
                           int(do_update), _ptr(step_ctr), solver, _stream()), "gf_adam_beta")


def adam_beta_bg(p, g, m, v, beta_bf, betaT_bf, *, K: int, V: int, lr: float, b1: float, b2: float, eps: float,
                 wd: float = 0.0, step: int = 1, gscale: float = 1.0, step_ctr=None, solver: int = 0,
                 blocks_per_sm: int = 2) -> None:
    """``adam_beta`` for the stream that runs next to the d-theta sweep (no shared memory, small register footprint)."""
    check(_lib.load().gf_adam_beta_bg(_ptr(p), _ptr(g), _ptr(m), _ptr(v), K, V, p.stride(0), g.stride(0), _ptr(beta_bf),
                                      beta_bf.stride(0), _ptr(betaT_bf), betaT_bf.shape[1], lr, b1, b2, eps, wd, step,
                                      gscale, _ptr(step_ctr), solver, blocks_per_sm, _stream()), "gf_adam_beta_bg")


# ----------------------------------------------------------------------------------------------
# CombinedTM: collapsed adapt_bert layer (csrc/ctm_gemm.cu)
# ----------------------------------------------------------------------------------------------
def ctm_split(src: torch.Tensor, out_hi: torch.Tensor, out_lo: torch.Tensor, *, extra_col=None, tail_row=None) -> None:
    """bf16 hi / lo shadows of ``src`` [rows, cols] (+ one extra column / row), zero padded to ``out`` width."""
    assert src.dim() == 2 and src.stride(1) == 1 and out_hi.stride(1) == 1 and out_hi.shape == out_lo.shape
    rows, cols = src.shape
    assert out_hi.shape[0] >= rows + (1 if tail_row is not None else 0)
    check(_lib.load().gf_ctm_split(_ptr(src), src.stride(0), _ptr(extra_col), _ptr(tail_row), rows, cols,
                                   _ptr(out_hi), _ptr(out_lo), out_hi.stride(0), _stream()), "gf_ctm_split")


def ctm_init_out(out: torch.Tensor, *, bias_row: int, bias) -> None:
    check(_lib.load().gf_ctm_init_out(_ptr(out), out.shape[0], out.shape[1], out.stride(0), bias_row, _ptr(bias),
                                      _stream()), "gf_ctm_init_out")


def ctm_vc(x_hi, x_lo, y_hi, y_lo, out, *, m_total: int, n_valid: int) -> None:
    """out[m, n] += sum_r x[r, m] y[r, n]  (contraction over the rows of both operands)."""
    R, ldx = x_hi.shape[0], x_hi.stride(0)
    nb = y_hi.stride(0) // 64
    assert y_hi.shape[0] == R and y_hi.stride(0) % 64 == 0
    check(_lib.load().gf_ctm_vc(_ptr(x_hi), _ptr(x_lo), ldx, m_total, _ptr(y_hi), _ptr(y_lo), nb, R, _ptr(out),
                                out.stride(0), n_valid, _stream()), "gf_ctm_vc")


def ctm_rc(x_hi, x_lo, z_hi, z_lo, out, *, k_total: int, n_valid: int) -> None:
    """out[r, n] = sum_k x[r, k] z[k, n]."""
    R, ldx = x_hi.shape[0], x_hi.stride(0)
    nb = z_hi.stride(0) // 64
    check(_lib.load().gf_ctm_rc(_ptr(x_hi), _ptr(x_lo), ldx, k_total, _ptr(z_hi), _ptr(z_lo), nb, z_hi.shape[0], R,
                                _ptr(out), out.stride(0), n_valid, _stream()), "gf_ctm_rc")


def ctm_rr(x_hi, x_lo, z_hi, z_lo, out, *, n_valid: int, extra=None) -> None:
    """out[r, c] = sum_k x[r, k] z[c, k] for c < n_valid; ``extra[r]`` receives column ``n_valid``."""
    R = x_hi.shape[0]
    nb = x_hi.stride(0) // 64
    assert z_hi.stride(0) == x_hi.stride(0)
    check(_lib.load().gf_ctm_rr(_ptr(x_hi), _ptr(x_lo), _ptr(z_hi), _ptr(z_lo), nb, z_hi.shape[0], R, _ptr(out),
                                out.stride(0), n_valid, _ptr(extra), _stream()), "gf_ctm_rr")


def adam_split(p, g, m, v, out_hi, out_lo, *, p_col=None, g_col=None, m_col=None, v_col=None, lr: float = 0.0,
               b1: float = 0.9, b2: float = 0.999, eps: float = 1e-8, wd: float = 0.0, step: int = 1, gscale: float = 1.0,
               do_update: bool = True, step_ctr=None, solver: int = 0) -> None:
    """Adam on a row-major [rows, cols] block (+ per-row scalar ``p_col`` as one more column) fused with the refresh
    of its bf16 hi / lo shadows; ``do_update=False`` only refreshes the shadows."""
    assert p.dim() == 2 and p.stride(1) == 1
    rows, cols = p.shape
    check(_lib.load().gf_adam_split(_ptr(p), _ptr(g), _ptr(m), _ptr(v), p.stride(0), rows, cols, _ptr(p_col), _ptr(g_col),
                                    _ptr(m_col), _ptr(v_col), _ptr(out_hi), _ptr(out_lo), out_hi.stride(0), lr, b1, b2, eps,
                                    wd, step, gscale, int(do_update), _ptr(step_ctr), solver, _stream()), "gf_adam_split")


def theta_mc(mu_pre, ls_pre, mu_rm, mu_rv, ls_rm, ls_rv, out, *, B: int, K: int, n_samples: int, seed: int,
             bn_eps: float = 1e-5, thr: float = 0.0) -> None:
    """Mean of ``n_samples`` reparameterised soft-max draws (eval-mode BatchNorm) in one launch, with the
    ``thr`` threshold + L1 renormalisation epilogue of the reference's result extraction (``thr <= 0``: off)."""
    check(_lib.load().gf_theta_mc(_ptr(mu_pre), _ptr(ls_pre), _ptr(mu_rm), _ptr(mu_rv), _ptr(ls_rm), _ptr(ls_rv), bn_eps,
                                  B, K, n_samples, seed & 0xFFFFFFFFFFFFFFFF, float(thr), _ptr(out), _stream()),
          "gf_theta_mc")


def bow_gather(indptr, indices, counts, W1T, out, *, B: int, accumulate: bool = False) -> None:
    """out[b,:] (+)= sum_e counts[e] * W1T[indices[e],:] over the non-zeros of document b (first encoder layer)."""
    assert W1T.stride(1) == 1 and W1T.stride(0) == W1T.shape[1] and out.stride(0) == W1T.shape[1]
    check(_lib.load().gf_bow_gather(_ptr(indptr), _ptr(indices), _ptr(counts), _ptr(W1T), W1T.shape[1], B, _ptr(out),
                                    int(accumulate), _stream()), "gf_bow_gather")


def small_linear_fwd(x, W, out, *, accumulate: bool) -> None:
    """out[b,:] (+)= x[b,:] @ W  with x [B,L], W [L,H] (label rows of the first encoder layer)."""
    B, L = x.shape
    assert W.shape[0] == L and W.stride(1) == 1 and W.stride(0) == W.shape[1] and x.is_contiguous()
    check(_lib.load().gf_small_linear_fwd(_ptr(x), _ptr(W), _ptr(out), B, L, W.shape[1], int(accumulate), _stream()),
          "gf_small_linear_fwd")


def small_linear_wgrad(x, g, dW) -> None:
    """dW [L,H] = x^T g  with x [B,L], g [B,H]."""
    B, L = x.shape
    assert x.is_contiguous() and g.is_contiguous() and dW.stride(1) == 1 and dW.stride(0) == g.shape[1]
    check(_lib.load().gf_small_linear_wgrad(_ptr(x), _ptr(g), _ptr(dW), B, L, g.shape[1], _stream()),
          "gf_small_linear_wgrad")


def label_head(theta, Wc, bc, labels, *, B: int, gscale: float, loss, dtheta=None, dWc=None, dbc=None) -> None:
    """Classification head of the supervised CTM: mean cross-entropy into ``loss`` and (optionally) its gradients."""
    L, K = Wc.shape
    assert theta.stride(0) == K and labels.is_contiguous() and labels.shape[1] == L and Wc.is_contiguous()
    check(_lib.load().gf_label_head(_ptr(theta), _ptr(Wc), _ptr(bc), _ptr(labels), B, K, L, gscale, _ptr(loss),
                                    _ptr(dtheta), dtheta.stride(0) if dtheta is not None else 0, _ptr(dWc), _ptr(dbc),
                                    _stream()), "gf_label_head")
