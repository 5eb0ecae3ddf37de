"""ProdLDA decoder (theta.beta -> BatchNorm(V) -> softmax(V) -> BoW NLL) forward + backward,
composed from the sm_100a kernels.  Nothing of size [B,V] is ever materialised.

Ref: avitm_network/decoder_network.py:121-126, avitm.py:225 and their autograd
(`loss.backward()` in federated_avitm.py:75).

Kernel sequence of one training step (see csrc/dec_sweep.cu for the operand families and
csrc/fold_ops.cu for the algebra that folds every epilogue constant into the GEMM operands):

  theta_center_fold  theta_c = bf16(theta - mean_b theta), plus the constant rows of theta_c^T
  theta_gram         C = theta_c^T theta_c appended to theta_c^T as pseudo documents (bf16 hi + lo)
  COL_STATS2         per-word batch variance beta_v^T C beta_v / B -> rstd, running stats
  scale_beta         beta' = bf16(beta * rstd*log2e) with bias / ones rows
  ROW_LSE2           per-document log-sum-exp partials over vocabulary chunks
  dec_sparse_doc     non-zeros: RL, S_b, sparse d theta, sparse BN-backward sums, -e_b rows
  csc_corr           sparse correction of d beta (through the CSC of the minibatch)
  COL_GRAD2          one pass: BN-backward constants and d beta; beta3 = gamma-scaled beta
  gram, gram_finish  M = beta diag(gamma) beta^T, a = beta.alpha (split-V tcgen05 GEMM) -> pseudo-word tiles
  ROW_GRAD2          d theta += ln2 * P . beta'^T - (a + theta_c M)   (dense part)

Small problems (B*V < 2^24) use the earlier side-band sweeps (ROW_LSE / COL_GRAD / ROW_GRAD), which need
fewer launches; ``GFEDNTM_DECODER_V1=1`` / ``=0`` forces one formulation.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Optional

import torch

from . import kernels as K_
from .kernels import TILE, padded_topics, round_up


@dataclass
class CsrBatch:
    """A minibatch of bag-of-words rows in CSR form on the device."""
    indptr: torch.Tensor   # int32 [B+1]
    indices: torch.Tensor  # int32 [nnz]
    counts: torch.Tensor   # float32 [nnz]
    B: int
    V: int

    @property
    def nnz(self) -> int:
        return self.indices.numel()

    @staticmethod
    def from_dense(x: torch.Tensor) -> "CsrBatch":
        """Helper for tests / dense-input API calls."""
        B, V = x.shape
        nz = x != 0
        lens = nz.sum(1)
        indptr = torch.zeros(B + 1, dtype=torch.int32, device=x.device)
        indptr[1:] = torch.cumsum(lens, 0).to(torch.int32)
        rows, cols = nz.nonzero(as_tuple=True)  # row-major order == CSR order with sorted columns
        return CsrBatch(indptr, cols.to(torch.int32), x[rows, cols].to(torch.float32), B, V)

    def to_dense(self) -> torch.Tensor:
        lens = (self.indptr[1:] - self.indptr[:-1]).long()
        rows = torch.repeat_interleave(torch.arange(self.B, device=self.indptr.device), lens)
        out = torch.zeros(self.B, self.V, dtype=torch.float32, device=self.indptr.device)
        out.index_put_((rows, self.indices.long()), self.counts, accumulate=True)
        return out


class CscWorkspace:
    """Per-step transpose of the minibatch (colptr / row ids / source positions)."""

    def __init__(self, V: int, max_nnz: int, device) -> None:
        self.V = V
        self.colptr = torch.zeros(V + 1, dtype=torch.int32, device=device)
        self.cursor = torch.zeros(V, dtype=torch.int32, device=device)
        self.row = torch.zeros(max(max_nnz, 1), dtype=torch.int32, device=device)
        self.src = torch.zeros(max(max_nnz, 1), dtype=torch.int32, device=device)

    def build(self, batch: CsrBatch, dynamic_nnz: bool = False) -> None:
        assert batch.nnz <= self.row.numel(), "CSC workspace too small for this minibatch"
        nnz = self.row.numel() if dynamic_nnz else batch.nnz
        K_.csr_to_csc(batch.indptr, batch.indices, B=batch.B, V=self.V, nnz=nnz, colptr=self.colptr,
                      cursor=self.cursor, csc_row=self.row, csc_src=self.src, dynamic_nnz=dynamic_nnz)


class ProdLDADecoder:
    """Workspace + launcher for the fused ProdLDA decoder at fixed (max B, V, K)."""

    # below this many logits per step the extra small kernels of the folded formulation cost more than
    # its leaner sweeps save (measured: B=64, V=10k: 0.248 vs 0.21 ms per training step)
    FOLD_MIN_LOGITS = 1 << 24

    def __init__(self, max_B: int, V: int, K: int, max_nnz: int, device, bn_eps: float = 1e-5,
                 bn_momentum: float = 0.1, folded: Optional[bool] = None) -> None:
        self.max_B, self.V, self.K = max_B, V, K
        self.Kp = padded_topics(K)
        if self.Kp > 240:
            raise ValueError("the fused decoder supports n_components <= 236 (four operand rows are reserved)")
        self.bn_eps, self.bn_momentum = bn_eps, bn_momentum
        dev = device
        f32 = dict(dtype=torch.float32, device=dev)
        self.Bpad = round_up(max_B, TILE)
        self.Vpad = round_up(V, TILE)
        env = os.environ.get("GFEDNTM_DECODER_V1")
        if folded is None:
            folded = (env == "0") if env in ("0", "1") else (max_B * V >= self.FOLD_MIN_LOGITS)
        self.folded = bool(folded)
        self.n_ext = (self.Kp + TILE - 1) // TILE          # pseudo-document tiles holding C = theta_c^T theta_c
        self.ext_col0 = self.Bpad
        bf = dict(dtype=torch.bfloat16, device=dev)
        self.theta_bf = torch.zeros(max_B, self.Kp, **bf)                                # sparse kernels
        self.thetaT_bf = torch.zeros(self.Kp, self.Bpad + 2 * self.n_ext * TILE, **bf)  # sweeps (+ C hi, C lo)
        if self.folded:
            # scaled operand shadow (per step) + pseudo-word blocks holding M hi / lo for ROW_GRAD2
            self.beta_s = torch.zeros(self.Kp, self.Vpad + 2 * self.n_ext * TILE, **bf)
            self.corrT = torch.zeros(self.Kp, round_up(V, 8), **f32)
            self.beta3 = torch.zeros(self.Kp, round_up(V, 8), **bf)       # gamma-scaled beta (per step)
            self.gram_parts = K_.gram_parts()
            self.gram_partial = torch.zeros(self.gram_parts * self.Kp * self.Kp, **f32)
            self.c_scratch = torch.zeros(K * K, **f32)
        self.theta_mean = torch.zeros(K, **f32)
        self.colc = torch.zeros(self.Vpad, 4, **f32)
        self.sb_doc = torch.zeros(self.Bpad, 4, **f32)
        self.doc_const = torch.zeros(max_B, 2, **f32)
        self.max_chunks = min((V + TILE - 1) // TILE, 2048)
        self.lse_part = torch.zeros(max_B * self.max_chunks, 2, **f32)
        self.rl = torch.zeros(max_B, **f32)
        self.asp_csp = torch.zeros(2, V, **f32)
        self.tnnz = torch.zeros(max(max_nnz, 1), **f32)
        self.corr = torch.zeros(V, self.Kp, **f32)
        self.dtheta = torch.zeros(max_B, self.Kp, **f32)

    # -- forward statistics + per-document reconstruction loss (+ everything backward needs)
    def forward_backward(self, theta: torch.Tensor, theta_colsum: torch.Tensor, beta_bf: torch.Tensor,
                         betaT_bf: torch.Tensor, batch: CsrBatch, csc: Optional[CscWorkspace], *,
                         training: bool, running_mean: Optional[torch.Tensor],
                         running_var: Optional[torch.Tensor], dbeta: Optional[torch.Tensor],
                         loss: Optional[torch.Tensor], gscale: float = 1.0, need_grad: bool = True, fed=None,
                         on_dbeta_final=None, sums_zeroed: bool = False):
        """Runs the sweeps.  Returns (rl[B], dtheta[B,Kp] or None).  ``dbeta`` [K,ld] is overwritten.

        ``fed`` (folded formulation only): an object with ``signal_beta_ready()``, ``dev_args`` (device
        pointer of the aggregation arguments), ``tmaps_ptr``, ``gb`` (bf16 buffer the beta gradient is
        written to INSTEAD of ``dbeta``) and ``ctas``; the beta part of the federated aggregation then runs
        on ``ctas`` CTAs of the d-theta sweep plus every sweep CTA that has finished (parallel/fabric.py)."""
        B, V, Kt = batch.B, self.V, self.K
        assert B <= self.max_B
        if not self.folded:
            return self._forward_backward_v1(theta, theta_colsum, beta_bf, betaT_bf, batch, csc, training=training,
                                             running_mean=running_mean, running_var=running_var, dbeta=dbeta,
                                             loss=loss, gscale=gscale, need_grad=need_grad, sums_zeroed=sums_zeroed)
        tcols = self.thetaT_bf.shape[1]
        K_.theta_center_fold(theta, theta_colsum, self.theta_bf, self.thetaT_bf, self.theta_mean, B=B, K=Kt,
                             Bpad=round_up(B, TILE), center=training)
        K_.theta_gram(self.theta_bf, self.thetaT_bf, self.c_scratch, B=B, K=Kt, ext_col0=self.ext_col0,
                      lo_col0=self.ext_col0 + self.n_ext * TILE)
        K_.dec_sweep(K_.COL_STATS2, B=B, V=V, K=Kt, theta_bf=self.thetaT_bf, beta_bf=beta_bf, colc=self.colc,
                     theta_mean=self.theta_mean if training else None,
                     running_mean=running_mean, running_var=running_var, bn_eps=self.bn_eps,
                     bn_momentum=self.bn_momentum, use_running=not training, theta_cols=tcols,
                     ext_col0=self.ext_col0, betaT_bf=betaT_bf)
        K_.scale_beta(beta_bf, self.colc, self.beta_s[:, :self.Vpad], K=Kt, V=V)
        n_chunks = K_.default_row_chunks(B, V, Kp=self.Kp)
        K_.dec_sweep(K_.ROW_LSE2, B=B, V=V, K=Kt, theta_bf=self.thetaT_bf, beta_bf=self.beta_s,
                     n_chunks=n_chunks, lse_part=self.lse_part, theta_cols=tcols, beta_cols=self.Vpad)
        bcols = self.beta_s.shape[1]
        grad = need_grad and training
        if grad and not sums_zeroed:  # (the engine zeroes the column sums with its other fills)
            self.asp_csp.zero_()
        K_.dec_sparse_doc(B=B, n_chunks=n_chunks, indptr=batch.indptr, indices=batch.indices,
                          counts=batch.counts, theta_bf=self.theta_bf, betaT_bf=betaT_bf, colc=self.colc,
                          lse_part=self.lse_part, gscale=gscale, rl=self.rl, loss=loss,
                          asp=self.asp_csp[0] if grad else None, csp=self.asp_csp[1] if grad else None,
                          tnnz=self.tnnz if grad else None, dtheta=self.dtheta if grad else None,
                          doc_const=None, sb_doc=None, thetaT_bf=self.thetaT_bf if grad else None, K=Kt)
        if not grad:
            return self.rl[:B], None
        assert csc is not None and dbeta is not None
        K_.csc_corr_t(csc.colptr, csc.row, csc.src, self.tnnz, self.theta_bf, self.corrT, V=V)
        K_.dec_sweep(K_.COL_GRAD2, B=B, V=V, K=Kt, theta_bf=self.thetaT_bf, beta_bf=self.beta_s,
                     colc=self.colc, asp=self.asp_csp[0], csp=self.asp_csp[1], dbeta=dbeta, corr=self.corrT,
                     colptr=csc.colptr, theta_cols=tcols, beta_cols=self.Vpad, ext_col0=self.ext_col0,
                     betaT_bf=betaT_bf, beta3=self.beta3,
                     dbeta_bf=fed.gb if fed is not None else None)  # federated exchange format: bf16
        K_.gram(self.beta3, beta_bf, self.gram_partial, V=V)
        K_.gram_finish(self.gram_partial, self.gram_parts, self.beta_s, K=Kt, col_hi=self.Vpad,
                       col_lo=self.Vpad + self.n_ext * TILE)
        if on_dbeta_final is not None:
            # d beta is final and the unscaled beta shadows are no longer read by this step (ROW_GRAD2 reads the scaled
            # copy): the caller may start the optimiser step of beta now (engine: side stream)
            on_dbeta_final()
        if fed is not None:
            # d beta is final and this rank no longer reads its unscaled beta shadows: peers may push
            fed.signal_beta_ready()
            n_chunks_g = K_.default_row_chunks(B, V, second_mma=True, Kp=self.Kp, grid=K_.num_sms() - fed.ctas)
            K_.dec_sweep(K_.ROW_GRAD2, B=B, V=V, K=Kt, theta_bf=self.thetaT_bf, beta_bf=self.beta_s,
                         n_chunks=n_chunks_g, dtheta=self.dtheta, theta_cols=tcols, beta_cols=bcols,
                         ext_col0=self.Vpad, fed=fed.dev_args, fed_ctas=fed.ctas, fed_tmaps=fed.tmaps_ptr)
            return self.rl[:B], self.dtheta[:B]
        n_chunks_g = K_.default_row_chunks(B, V, second_mma=True, Kp=self.Kp)
        K_.dec_sweep(K_.ROW_GRAD2, B=B, V=V, K=Kt, theta_bf=self.thetaT_bf, beta_bf=self.beta_s,
                     n_chunks=n_chunks_g, dtheta=self.dtheta, theta_cols=tcols, beta_cols=bcols,
                     ext_col0=self.Vpad)
        return self.rl[:B], self.dtheta[:B]

    def _forward_backward_v1(self, theta, theta_colsum, beta_bf, betaT_bf, batch, csc, *, training,
                             running_mean, running_var, dbeta, loss, gscale, need_grad, sums_zeroed=False):
        """Earlier formulation: epilogue constants through the float4 side bands."""
        B, V, Kt = batch.B, self.V, self.K
        K_.theta_center(theta, theta_colsum, self.theta_bf, self.thetaT_bf, self.theta_mean, B=B, K=Kt,
                        center=training)
        K_.dec_sweep(K_.COL_STATS, B=B, V=V, K=Kt, theta_bf=self.thetaT_bf, beta_bf=beta_bf,
                     sb_stream=self.sb_doc, colc=self.colc,
                     theta_mean=self.theta_mean if training else None,
                     running_mean=running_mean, running_var=running_var, bn_eps=self.bn_eps,
                     bn_momentum=self.bn_momentum, use_running=not training)
        n_chunks = K_.default_row_chunks(B, V, Kp=self.Kp)
        K_.dec_sweep(K_.ROW_LSE, B=B, V=V, K=Kt, theta_bf=self.thetaT_bf, beta_bf=beta_bf,
                     sb_stream=self.colc, n_chunks=n_chunks, lse_part=self.lse_part)
        grad = need_grad and training
        if grad and not sums_zeroed:  # (the engine zeroes the column sums with its other fills)
            self.asp_csp.zero_()
        K_.dec_sparse_doc(B=B, n_chunks=n_chunks, indptr=batch.indptr, indices=batch.indices,
                          counts=batch.counts, theta_bf=self.theta_bf, betaT_bf=betaT_bf, colc=self.colc,
                          lse_part=self.lse_part, gscale=gscale, rl=self.rl, loss=loss,
                          asp=self.asp_csp[0] if grad else None, csp=self.asp_csp[1] if grad else None,
                          tnnz=self.tnnz if grad else None, dtheta=self.dtheta if grad else None,
                          doc_const=self.doc_const if grad else None, sb_doc=self.sb_doc if grad else None)
        if not grad:
            return self.rl[:B], None
        assert csc is not None and dbeta is not None
        K_.csc_corr(csc.colptr, csc.row, csc.src, self.tnnz, self.theta_bf, self.corr, V=V)
        K_.dec_sweep(K_.COL_GRAD, B=B, V=V, K=Kt, theta_bf=self.thetaT_bf, beta_bf=beta_bf,
                     sb_stream=self.sb_doc, colc=self.colc, asp=self.asp_csp[0], csp=self.asp_csp[1],
                     dbeta=dbeta, corr=self.corr, colptr=csc.colptr)
        n_chunks_g = K_.default_row_chunks(B, V, second_mma=True, Kp=self.Kp)
        K_.dec_sweep(K_.ROW_GRAD, B=B, V=V, K=Kt, theta_bf=self.thetaT_bf, beta_bf=beta_bf,
                     sb_stream=self.colc, n_chunks=n_chunks_g, doc_const=self.doc_const,
                     dtheta=self.dtheta)
        return self.rl[:B], self.dtheta[:B]


def make_beta_shadows(K: int, V: int, device):
    """bf16 operand shadows of beta: [Kp, ldb] (rows >= K zero) and beta^T [V, Kp]."""
    Kp = padded_topics(K)
    beta_bf = torch.zeros(Kp, round_up(V, 8), dtype=torch.bfloat16, device=device)
    betaT_bf = torch.zeros(V, Kp, dtype=torch.bfloat16, device=device)
    return beta_bf, betaT_bf


def refresh_beta_shadows(beta: torch.Tensor, beta_bf: torch.Tensor, betaT_bf: torch.Tensor) -> None:
    K, V = beta.shape
    K_.adam_beta(beta, None, None, None, beta_bf, betaT_bf, K=K, V=V, do_update=False)
