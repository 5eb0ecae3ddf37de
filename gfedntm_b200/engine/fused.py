"""Fused sm_100a training / inference engine for the VAE topic models.

The engine owns flat fp32 buffers for parameters, gradients and Adam moments, re-homes the
``nn.Module`` parameters into them (so ``state_dict()`` / ``load_state_dict()`` keep working on
the live storage) and runs one minibatch step as a fixed sequence of hand-written kernels:

  csr_to_csc -> enc_fwd -> latent_fwd -> [ProdLDA decoder sweeps | NeuralLDA sparse decoder]
  -> latent_bwd -> enc_bwd -> csc_spmm (d W1^T) -> fused Adam (+ bf16 operand shadows)

Ref: the per-minibatch body of ``FederatedAVITM.train_mb_delta`` (federated_avitm.py:51-83) /
``AVITM._train_epoch`` (avitm.py:231-277) and their CTM counterparts.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Dict, Optional, Tuple

import torch

from ..models.networks import DecoderNetwork
from ..ops import kernels as K_
from ..ops.decoder import (CscWorkspace, CsrBatch, ProdLDADecoder, make_beta_shadows,
                           refresh_beta_shadows)

BN_EPS = 1e-5
BN_MOMENTUM = 0.1


@dataclass
class Segment:
    name: str
    offset: int
    numel: int
    shape: Tuple[int, ...]
    transposed: bool  # stored as [in, out] (transposed w.r.t. the nn.Linear weight)


class FlatParams:
    """Flat fp32 parameter / gradient / Adam-moment buffers with named segments.

    Order: every small tensor first (one contiguous ``small`` range), then the first-layer weight
    stored transposed ``[fan_in, H0]`` (row gather for the sparse BoW input), then
    ``adapt_bert.weight`` (CombinedTM), and ``beta`` last.
    """

    BIG = ("inf_net.input_layer.weight", "inf_net.adapt_bert.weight", "beta")

    def __init__(self, model: DecoderNetwork, device, arena) -> None:
        named = [(n, p) for n, p in model.named_parameters()]
        small = [(n, p) for n, p in named if n not in self.BIG]
        big = [(n, p) for n in self.BIG for (m, p) in named if m == n]
        self.segments: Dict[str, Segment] = {}
        off = 0
        for n, p in small + big:
            off = (off + 3) // 4 * 4  # 16-byte alignment of every segment
            self.segments[n] = Segment(n, off, p.numel(), tuple(p.shape), n == "inf_net.input_layer.weight")
            off += p.numel()
        self.total = (off + 3) // 4 * 4
        first_big = big[0][0]
        self.small_end = self.segments[first_big].offset
        self.beta_off = self.segments["beta"].offset
        f32 = dict(dtype=torch.float32, device=device)
        # BatchNorm running statistics: one flat buffer as well (federated clients average them)
        K, V = model.n_components, model.input_size
        self.s_n = (4 * K + 2 * V + 3) // 4 * 4
        Kp = K_.padded_topics(K)
        ldb = (V + 7) // 8 * 8
        self.has_shadows = model.model_type.lower() == "prodlda"
        if "P" not in arena.offsets:
            arena.reserve("P", 4 * self.total)
            arena.reserve("G", 4 * self.total)
            arena.reserve("S", 4 * self.s_n)
            arena.reserve("flags", 256)  # GF_FED_FLAG_BYTES
            if self.has_shadows:
                arena.reserve("beta_bf", 2 * Kp * ldb)
                arena.reserve("betaT_bf", 2 * V * Kp)
                # beta gradient in the exchange format of the overlapped update (bf16 [K, ldb]): pulled by the owners
                # in a federation, read by this GPU's own update CTAs otherwise
                arena.reserve("Gb", 2 * K * ldb)
            if getattr(arena, "world", 1) > 1 and getattr(model, "infnet", "") == "combined":
                # staging area of the small all-reduce of CombinedTM's collapsed-weight gradient
                self.ar_cap = ((model.contextual_size + 1) * model.hidden_sizes[0] + 3) // 4 * 4
                arena.reserve("AR", 4 * self.ar_cap * arena.world)
            arena.commit(device)
        self.arena = arena
        self.P = arena.tensor("P", torch.float32)
        self.G = arena.tensor("G", torch.float32)
        self.S = arena.tensor("S", torch.float32)
        self.M = torch.zeros(self.total, **f32)
        self.V2 = torch.zeros(self.total, **f32)
        self._params = dict(named)
        with torch.no_grad():
            for n, p in named:
                view = self.view(self.P, n)
                view.copy_(p.detach().to(device))
                p.data = view
                p.grad = self.view(self.G, n)
            off = 0
            for bn, n in ((model.inf_net.f_mu_batchnorm, K), (model.inf_net.f_sigma_batchnorm, K),
                          (model.beta_batchnorm, V)):
                for attr in ("running_mean", "running_var"):
                    view = self.S[off:off + n]
                    view.copy_(getattr(bn, attr).detach().to(device))
                    setattr(bn, attr, view)
                    off += n
        if self.has_shadows:
            self.beta_bf = arena.tensor("beta_bf", torch.bfloat16).view(Kp, ldb)
            self.betaT_bf = arena.tensor("betaT_bf", torch.bfloat16).view(V, Kp)

    def view(self, flat: torch.Tensor, name: str) -> torch.Tensor:
        s = self.segments[name]
        raw = flat[s.offset:s.offset + s.numel]
        if s.transposed:
            out_f, in_f = s.shape
            return raw.view(in_f, out_f).t()
        return raw.view(s.shape)

    def storage_view(self, flat: torch.Tensor, name: str) -> torch.Tensor:
        """The segment in its *storage* layout (first layer: [fan_in, H0])."""
        s = self.segments[name]
        raw = flat[s.offset:s.offset + s.numel]
        if s.transposed:
            out_f, in_f = s.shape
            return raw.view(in_f, out_f)
        return raw.view(s.shape)


class FusedEngine:
    """One model replica on one GPU, trained step-at-a-time through the kernel library."""

    def __init__(self, model: DecoderNetwork, *, max_batch: int, max_nnz: int, lr: float = 2e-3,
                 betas: Tuple[float, float] = (0.99, 0.99), adam_eps: float = 1e-8,
                 weight_decay: float = 0.0, kl_weight: float = 1.0, seed: int = 0, device=None,
                 arena=None, solver: str = "adam") -> None:
        if not torch.cuda.is_available():
            raise RuntimeError("FusedEngine needs a CUDA device (sm_100a); use the eager backend on CPU")
        self.L = int(model.label_size or 0)
        if self.L and model.infnet == "bow":
            raise NotImplementedError("labels belong to the contextualized models")
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.model = model.to(self.device)
        self.kind = model.infnet
        self.is_prod = model.model_type.lower() == "prodlda"
        self.V, self.K = model.input_size, model.n_components
        self.hidden = tuple(model.hidden_sizes)
        self.C = model.contextual_size
        self.act = K_.ACTIVATIONS[model.activation]
        self.theta_drop = float(model.dropout)
        self.enc_drop = float(model.inf_net.dropout)
        self.kl_weight = float(kl_weight)
        self.lr, self.betas, self.adam_eps, self.weight_decay = lr, betas, adam_eps, weight_decay
        self.solver_name = solver
        self.solver = 0
        if solver != "adam":  # (momentum travels as betas[0], like Adam's first beta in the reference's factory)
            self.solver, b1_, b2_, eps_ = K_.solver_constants(solver, betas[0])
            self.betas, self.adam_eps = (b1_, b2_), eps_
        self.seed = int(seed)
        self.step_count = 0       # optimiser steps taken
        self.rng_calls = 0        # forward passes (advances the Philox seed)
        self.max_batch, self.max_nnz = max_batch, max_nnz

        if arena is None:
            from ..parallel.fabric import LocalArena

            arena = LocalArena()
        self.flat = FlatParams(self.model, self.device, arena)
        dev = self.device
        f32 = dict(dtype=torch.float32, device=dev)
        B, K, V = max_batch, self.K, self.V
        HL = self.hidden[-1]
        self.Kp = K_.padded_topics(K)

        # forward/backward workspace
        self.pre = [torch.zeros(B, h, **f32) for h in self.hidden]
        self.mask_enc = torch.ones(B, HL, **f32)
        self.mu_pre = torch.zeros(B, K, **f32)
        self.ls_pre = torch.zeros(B, K, **f32)
        self.mu = torch.zeros(B, K, **f32)
        self.logvar = torch.zeros(B, K, **f32)
        self.eps = torch.zeros(B, K, **f32)
        self.theta_s = torch.zeros(B, K, **f32)
        self.mask_theta = torch.ones(B, K, **f32)
        self.theta = torch.zeros(B, K, **f32)
        self.kl = torch.zeros(B, **f32)
        self.dmu = torch.zeros(B, K, **f32)
        self.dlv = torch.zeros(B, K, **f32)
        self.dpre0 = torch.zeros(B, self.hidden[0], **f32)
        # buffers that must start at zero every step, zeroed with one memset each
        n_dm = K_.round_up((self.C + 1) * self.hidden[0], 4) if self.kind == "combined" else 0
        off_dm = K_.round_up(5 * K, 4)  # 16-byte aligned: the federated small all-reduce moves float4
        self.zpool = torch.zeros(off_dm + n_dm, **f32)
        self.bn_bsums = self.zpool[:4 * K]
        self.theta_colsum = self.zpool[4 * K:5 * K]
        if self.kind == "bow":
            self.bow_pre0 = torch.zeros(B, self.hidden[0], **f32)   # first-layer BoW contribution (gf_bow_gather)
        if self.kind in ("combined", "zeroshot"):
            # Dense contextual input on the tensor cores (csrc/ctm_gemm.cu): x_ctx [B,C] . W^T [C,H0] forward
            # (gf_ctm_rc) and d W^T = x_ctx^T . d pre0 backward (gf_ctm_vc), on bf16 hi/lo shadows.
            C_, H0 = self.C, self.hidden[0]
            bf = dict(dtype=torch.bfloat16, device=dev)
            self.ctx_Cp, self.ctx_Np = K_.round_up(C_, 64), K_.round_up(H0, 64)
            self.ctx_x_hi = torch.zeros(B, self.ctx_Cp, **bf)
            self.ctx_x_lo = torch.zeros(B, self.ctx_Cp, **bf)
            self.ctx_z_hi = torch.zeros(C_, self.ctx_Np, **bf)     # shadows of the dense first-layer weight [C, H0]
            self.ctx_z_lo = torch.zeros(C_, self.ctx_Np, **bf)
            self.ctx_g_hi = torch.zeros(B, self.ctx_Np, **bf)      # shadows of d pre0 [B, H0]
            self.ctx_g_lo = torch.zeros(B, self.ctx_Np, **bf)
            self.ctx_pre0 = torch.zeros(B, H0, **f32)
        if self.kind == "combined":
            # Collapsed adapt_bert layer (csrc/ctm_gemm.cu): bf16 hi/lo shadows of A' = [adapt_bert.weight |
            # adapt_bert.bias] and of W = input_layer.weight[:, V:2V]^T, the collapsed weight M'^T (+ effective
            # bias in its last row) and the gradient wrt it (+ d b1 in its last row).
            C_, H0 = self.C, self.hidden[0]
            self.ctm_Cp = K_.round_up(C_ + 1, 64)
            self.ctm_Np = K_.round_up(H0, 64)
            bf = dict(dtype=torch.bfloat16, device=dev)
            self.ctm_a_hi = torch.zeros(V, self.ctm_Cp, **bf)
            self.ctm_a_lo = torch.zeros(V, self.ctm_Cp, **bf)
            self.ctm_w_hi = torch.zeros(V, self.ctm_Np, **bf)
            self.ctm_w_lo = torch.zeros(V, self.ctm_Np, **bf)
            self.ctm_d_hi = torch.zeros(C_ + 1, self.ctm_Np, **bf)
            self.ctm_d_lo = torch.zeros(C_ + 1, self.ctm_Np, **bf)
            self.ctm_MTx = torch.zeros(C_ + 1, H0, **f32)
            self.ctm_dM_flat = self.zpool[off_dm:off_dm + n_dm]           # (padded to a multiple of 4 floats)
            self.ctm_dMx = self.ctm_dM_flat[:(C_ + 1) * H0].view(C_ + 1, H0)
        self.dpool = torch.zeros(4 * K + 1, dtype=torch.float64, device=dev)
        self.bn_sums = self.dpool[:4 * K]
        self.loss = self.dpool[4 * K:]

        self.csc = CscWorkspace(V, max_nnz, dev)
        if self.is_prod:
            self.decoder = ProdLDADecoder(B, V, K, max_nnz, dev, BN_EPS, BN_MOMENTUM)
            self.beta_bf, self.betaT_bf = self.flat.beta_bf, self.flat.betaT_bf
        else:
            self.xhat = torch.zeros(K, V, **f32)
            self.bp = torch.zeros(K, V, **f32)
            self.dxhat = torch.zeros(K, V, **f32)
            self.col_rstd = torch.zeros(V, **f32)
            self.dbpT = torch.zeros(V, K, **f32)
            self.gnnz = torch.zeros(max(max_nnz, 1), **f32)
            self.rl = torch.zeros(B, **f32)
            self.dtheta_lda = torch.zeros(B, self.Kp, **f32)
        # device-side counters (forward passes, optimiser steps): seeds / Adam bias correction / the
        # federated epoch flag are derived from them on the device, so a captured step can be replayed
        self.ctr = torch.zeros(2, dtype=torch.int64, device=dev)
        self._graphs = {}
        self._graph_launches = {}
        self.replayed_launches = 0   # kernels launched through graph replays (not seen by gf_launch_count)
        self._warm = set()
        self.use_graphs = os.environ.get("GFEDNTM_NO_GRAPHS", "0") != "1"
        self._nbt_pending = 0
        self.model.register_state_dict_pre_hook(lambda *a, **k: self.flush_counters())
        self.sync_shadows()
        self._self_overlap = None
        self._init_self_overlap(arena)
        # Side stream of the training step: the per-step CSC transpose (needed only by the backward pass) runs under
        # the encoder forward, and the optimiser step of beta (pure HBM traffic) starts as soon as d beta is final and
        # runs under the tail of the d-theta sweep and the small latency-bound kernels of the encoder backward.
        self._side = torch.cuda.Stream(self.device) if os.environ.get("GFEDNTM_SIDE_STREAM", "1") != "0" else None
        self._side_beta = False
        self._side_join = False
        # every training step also copies its loss into a pinned host slot (one slot per CUDA-graph key, so
        # consecutive steps on different input buffers never share one); the copy is part of the step's graph and
        # runs on the side stream behind the forward pass: reading the loss on the host needs no D2H copy between
        # two graph launches on the compute stream (measured: 43 us per step at the large configuration)
        self._loss_slots = {}
        self.loss_slot = None

    def _init_self_overlap(self, arena) -> None:
        """Single-GPU training, OPT-IN (``GFEDNTM_SELF_OVERLAP=1``): the beta part of the optimiser step runs on CTAs of
        the d-theta sweep -- the overlapped federated aggregation with a world of one; the gradient travels as bf16
        [K, ldb].  Measured at V=200k, K=200, B=4096: 3.52 ms/step against 3.32 with the separate ``adam_beta``
        kernel (235 us at 5.3 TB/s): a sweep CTA owns its SM (registers and shared memory), so the update cannot
        share an SM with the sweep, and as the tail after the sweep it moves ~3 TB/s.  Kept because it exercises the
        aggregation role of the sweep kernel on one GPU (profiling with ncu, tests)."""
        from ..parallel.fabric import FusedAggregator, LocalArena

        if not (self.is_prod and self.decoder.folded and isinstance(arena, LocalArena) and self.V % 8 == 0
                and "Gb" in arena.offsets and os.environ.get("GFEDNTM_SELF_OVERLAP", "0") == "1"):
            return
        fl = self.flat
        agg = FusedAggregator(world=1, weights=[1.0], flat_n=fl.beta_off, beta_off=fl.beta_off, s_n=fl.s_n, K=self.K,
                              V=self.V, Kp=self.Kp, ldb=self.beta_bf.stride(0), lr=self.lr, betas=self.betas,
                              eps=self.adam_eps, wd=self.weight_decay, solver=self.solver)
        agg.add_rank(arena, 0, M=fl.M, V2=fl.V2, has_shadows=True, ld_gb=self.beta_bf.stride(0))
        ctas = int(os.environ.get("GFEDNTM_SELF_AGG_CTAS", "16"))
        self._self_agg = agg
        self._self_overlap = agg.overlap_handle(self.ctr, ctas)

    def flush_counters(self) -> None:
        """Apply the pending ``num_batches_tracked`` increments of the three BatchNorm layers."""
        if self._nbt_pending:
            for bn in (self.model.inf_net.f_mu_batchnorm, self.model.inf_net.f_sigma_batchnorm,
                       self.model.beta_batchnorm):
                bn.num_batches_tracked += self._nbt_pending
            self._nbt_pending = 0

    def set_lr(self, lr: float) -> None:
        """New learning rate (plateau scheduler): the captured step graphs carry the old value, so they are dropped."""
        if float(lr) != self.lr:
            self.lr = float(lr)
            if self._self_overlap is not None:  # the update CTAs read lr from a device copy of the arguments
                self._self_agg.set_lr(self.lr, self.betas)
                self._self_overlap = self._self_agg.overlap_handle(self.ctr, self._self_overlap.ctas)
            self._graphs.clear()
            self._graph_launches.clear()
            self._warm.clear()

    def set_step_count(self, n: int) -> None:
        """Optimiser step counter (host mirror + device counter used by the Adam kernels)."""
        self.step_count = int(n)
        self.ctr[1] = int(n)

    # ------------------------------------------------------------------------------------------
    # parameter access helpers
    # ------------------------------------------------------------------------------------------
    def _p(self, name: str) -> torch.Tensor:
        return self.flat.view(self.flat.P, name)

    def _g(self, name: str) -> torch.Tensor:
        return self.flat.view(self.flat.G, name)

    def sync_shadows(self) -> None:
        """Rebuild the bf16 operand shadows after parameters were written from outside
        (load_state_dict, federated parameter averaging)."""
        if self.is_prod:
            refresh_beta_shadows(self._p("beta"), self.beta_bf, self.betaT_bf)
        self._refresh_ctm_shadows()
        self._refresh_ctx_shadows()

    def _refresh_ctx_shadows(self) -> None:
        """ZeroShotTM: bf16 hi/lo shadows of the dense first-layer weight [C, H0] (CombinedTM rebuilds them from the
        collapsed weight every step)."""
        if self.kind == "zeroshot":
            st = self.flat.storage_view(self.flat.P, "inf_net.input_layer.weight")
            K_.ctm_split(st[:self.C], self.ctx_z_hi, self.ctx_z_lo)

    def _refresh_ctm_shadows(self) -> None:
        """bf16 hi/lo operand shadows of the CombinedTM adapt_bert layer and of the contextual half of the first
        layer (rebuilt after every optimiser step / parameter load)."""
        if self.kind != "combined":
            return
        self.ctm_adam_step(do_update=False)

    def ctm_adam_step(self, gscale: float = 1.0, do_update: bool = True) -> None:
        """Adam on adapt_bert.{weight,bias} and on the contextual half of the first layer, fused with the refresh of
        their bf16 hi/lo operand shadows (``do_update=False``: shadows only).  The caller advanced the step counter."""
        fl, V = self.flat, self.V
        b1, b2 = self.betas
        kw = dict(lr=self.lr, b1=b1, b2=b2, eps=self.adam_eps, wd=self.weight_decay, step=max(self.step_count, 1),
                  gscale=gscale, do_update=do_update, step_ctr=self.ctr if do_update else None, solver=self.solver)
        A, ab = "inf_net.adapt_bert.weight", "inf_net.adapt_bert.bias"
        K_.adam_split(fl.view(fl.P, A), fl.view(fl.G, A), fl.view(fl.M, A), fl.view(fl.V2, A), self.ctm_a_hi,
                      self.ctm_a_lo, p_col=fl.view(fl.P, ab), g_col=fl.view(fl.G, ab), m_col=fl.view(fl.M, ab),
                      v_col=fl.view(fl.V2, ab), **kw)
        W = "inf_net.input_layer.weight"
        K_.adam_split(fl.storage_view(fl.P, W)[V:2 * V], fl.storage_view(fl.G, W)[V:2 * V],
                      fl.storage_view(fl.M, W)[V:2 * V], fl.storage_view(fl.V2, W)[V:2 * V], self.ctm_w_hi,
                      self.ctm_w_lo, **kw)

    def _flat_adam_pieces(self):
        """Slices of the flat region [0, beta_off) that the plain flat Adam updates (everything except the tensors
        ``ctm_adam_step`` owns)."""
        fl = self.flat
        if self.kind != "combined":
            return [(0, fl.beta_off)]
        ab = fl.segments["inf_net.adapt_bert.bias"]
        w = fl.segments["inf_net.input_layer.weight"]
        ab_end = (ab.offset + ab.numel + 3) // 4 * 4
        half = self.V * self.hidden[0]
        pieces = [(0, ab.offset), (ab_end, w.offset + half)]
        if w.numel > 2 * half:  # label rows of the first layer (supervised CTM) follow the contextual half
            pieces.append((w.offset + 2 * half, (w.offset + w.numel + 3) // 4 * 4))
        return pieces

    def _hidden_params(self, flat: torch.Tensor):
        Wh, bh = [], []
        for i in range(len(self.hidden) - 1):
            Wh.append(self.flat.view(flat, f"inf_net.hiddens.l_{i}.0.weight"))
            bh.append(self.flat.view(flat, f"inf_net.hiddens.l_{i}.0.bias"))
        pad = K_.GF_MAX_HIDDEN - len(Wh)
        return Wh + [None] * pad, bh + [None] * pad

    def _bn(self, which: str):
        bn = getattr(self.model.inf_net, which)
        return bn.running_mean, bn.running_var, bn.num_batches_tracked

    # ------------------------------------------------------------------------------------------
    # forward (+ backward) of one minibatch
    # ------------------------------------------------------------------------------------------
    def _first_layer(self, x_ctx: Optional[torch.Tensor], B: int):
        """Returns (W1T or None, b1, pre0_ext or None) for enc_fwd according to the encoder kind.  The dense
        contextual input is contracted on the tensor cores here; enc_fwd only adds the result."""
        st = self.flat.storage_view(self.flat.P, "inf_net.input_layer.weight")  # [fan_in, H0]
        b1 = self._p("inf_net.input_layer.bias")
        if self.kind == "bow":
            return st, b1, None
        H0, C_ = self.hidden[0], self.C
        K_.ctm_split(x_ctx[:B], self.ctx_x_hi, self.ctx_x_lo)
        if self.kind == "zeroshot":
            W1T = None   # (weight shadows ctx_z_* follow the optimiser, see _refresh_ctx_shadows)
        else:
            # combined: collapse adapt_bert into the first layer: [M^T ; b_eff] = [A | ab]^T W (+ b1 in the last
            # row), a split-V tcgen05 GEMM over the bf16 hi/lo shadows (csrc/ctm_gemm.cu)
            K_.ctm_init_out(self.ctm_MTx, bias_row=C_, bias=b1)
            K_.ctm_vc(self.ctm_a_hi, self.ctm_a_lo, self.ctm_w_hi, self.ctm_w_lo, self.ctm_MTx, m_total=C_ + 1,
                      n_valid=H0)
            K_.ctm_split(self.ctm_MTx[:C_], self.ctx_z_hi, self.ctx_z_lo)
            W1T, b1 = st[:self.V], self.ctm_MTx[C_]
        K_.ctm_rc(self.ctx_x_hi[:B], self.ctx_x_lo[:B], self.ctx_z_hi, self.ctx_z_lo, self.ctx_pre0[:B], k_total=C_,
                  n_valid=H0)
        return W1T, b1, self.ctx_pre0

    def forward_backward(self, batch: CsrBatch, x_ctx: Optional[torch.Tensor] = None, *,
                         training: bool = True, sample: bool = True, gscale: float = 1.0,
                         need_grad: bool = True, eps_in=None, mask_enc_in=None, mask_theta_in=None, fed=None,
                         ctm_reduce_hook=None, labels: Optional[torch.Tensor] = None, side_stream: bool = False,
                         inc_opt: int = 0, beta_on_side: bool = False, loss_host: Optional[torch.Tensor] = None):
        """Runs forward (and backward when ``need_grad``); returns the device scalar
        ``loss = sum_b (kl_weight*KL_b + RL_b)`` (float64, shape [1])."""
        B, K, V = batch.B, self.K, self.V
        if B > self.max_batch or batch.nnz > self.max_nnz:
            raise ValueError("minibatch exceeds the engine workspace (max_batch / max_nnz)")
        grad = need_grad and training
        seed = (self.seed * 0x9E3779B97F4A7C15 + 1) & ((1 << 64) - 1)
        # step counter + every zero fill of the step in one launch (the column sums of the decoder included)
        zero = [self.zpool, self.dpool]
        if grad:
            zero.append(self.flat.G[:self.flat.small_end])
            if self.is_prod:
                zero.append(self.decoder.asp_csp)
        # (inc_opt = 1: the caller wants the optimiser-step counter advanced here too, see train_step)
        K_.step_begin_zero(self.ctr, 1, inc_opt, zero)
        # side stream: the CSC transpose of the batch (needed by the backward pass only) runs next to the encoder
        # forward; with beta_on_side also beta's optimiser step, beside the d-theta sweep (folded decoder)
        side = self._side if (grad and side_stream) else None
        if grad:
            if side is not None:
                side.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(side):
                    self.csc.build(batch, dynamic_nnz=True)
            else:
                self.csc.build(batch, dynamic_nnz=True)

        W1T, b1, pre0_ext = self._first_layer(x_ctx, B)
        Wh, bh = self._hidden_params(self.flat.P)
        use_bow = self.kind in ("bow", "combined")
        if use_bow:
            # sparse BoW part of the first layer: several warps per document, added to the contextual part
            if pre0_ext is None:
                pre0_ext = self.bow_pre0
            K_.bow_gather(batch.indptr, batch.indices, batch.counts, W1T, pre0_ext, B=B,
                          accumulate=self.kind == "combined")
            W1T = None   # (enc_fwd gets no CSR input any more: everything of layer 0 arrives through pre0_ext)
        if self.L and labels is not None:
            # supervised CTM: the one-hot labels are the last label_size inputs of the first layer
            st_all = self.flat.storage_view(self.flat.P, "inf_net.input_layer.weight")
            K_.small_linear_fwd(labels[:B].contiguous(), st_all[st_all.shape[0] - self.L:], pre0_ext, accumulate=True)
        pad = [None] * (K_.GF_MAX_HIDDEN - len(self.hidden))
        K_.enc_fwd(B=B, V=V, K=K, hidden=self.hidden, act=self.act, training=training,
                   indptr=None, indices=None, counts=None, W1T=W1T, b1=b1, xdense=None, WdT=None,
                   pre0_ext=pre0_ext,
                   Wh=Wh, bh=bh, Wmu=self._p("inf_net.f_mu.weight"), bmu=self._p("inf_net.f_mu.bias"),
                   Wsig=self._p("inf_net.f_sigma.weight"), bsig=self._p("inf_net.f_sigma.bias"),
                   pre=self.pre + pad, enc_drop_p=self.enc_drop, mask_enc_in=mask_enc_in,
                   mask_enc=self.mask_enc, mu_pre=self.mu_pre, ls_pre=self.ls_pre,
                   bn_sums=self.bn_sums if training else None, seed=seed, step_ctr=self.ctr)
        mu_rm, mu_rv, mu_nbt = self._bn("f_mu_batchnorm")
        ls_rm, ls_rv, ls_nbt = self._bn("f_sigma_batchnorm")
        K_.latent_fwd(B=B, K=K, training=training, sample=sample, mu_pre=self.mu_pre, ls_pre=self.ls_pre,
                      bn_sums=self.bn_sums, mu_rm=mu_rm, mu_rv=mu_rv, ls_rm=ls_rm, ls_rv=ls_rv,
                      bn_eps=BN_EPS, bn_momentum=BN_MOMENTUM, prior_mean=self.model.prior_mean,
                      prior_var=self.model.prior_variance, eps_in=eps_in, mask_theta_in=mask_theta_in,
                      theta_drop_p=self.theta_drop, mu=self.mu, logvar=self.logvar, eps=self.eps,
                      theta_s=self.theta_s, mask_theta=self.mask_theta, theta=self.theta,
                      theta_colsum=self.theta_colsum, kl=self.kl, loss=self.loss,
                      kl_weight=self.kl_weight, seed=seed, step_ctr=self.ctr)
        bbn = self.model.beta_batchnorm
        if training and not torch.cuda.is_current_stream_capturing():
            self._nbt_pending += 1  # BatchNorm step counters are flushed lazily (state_dict hook)

        if side is not None:
            torch.cuda.current_stream().wait_stream(side)   # the CSC transpose is needed from here on
        if self.is_prod:
            hook = None
            if side is not None:
                if fed is None and beta_on_side and self.decoder.folded:
                    def hook():  # optimiser step of beta on the side stream (joined in optimizer_step)
                        side.wait_stream(torch.cuda.current_stream())
                        with torch.cuda.stream(side):
                            self._adam_beta_now(gscale=1.0, background=os.environ.get("GFEDNTM_BG_ADAM", "1") != "0")
                        self._side_beta = True
            _, dtheta = self.decoder.forward_backward(
                self.theta, self.theta_colsum, self.beta_bf, self.betaT_bf, batch, self.csc,
                training=training, running_mean=bbn.running_mean, running_var=bbn.running_var,
                dbeta=self._g("beta") if grad else None, loss=self.loss, gscale=gscale, need_grad=grad,
                fed=fed if (grad and self.decoder.folded) else None, on_dbeta_final=hook, sums_zeroed=grad)
        else:
            beta = self._p("beta")
            K_.lda_beta_fwd(beta, K=K, V=V, eps=BN_EPS, momentum=BN_MOMENTUM, training=training,
                            running_mean=bbn.running_mean, running_var=bbn.running_var, xhat=self.xhat,
                            col_rstd=self.col_rstd, bp=self.bp)
            K_.lda_doc(batch.indptr, batch.indices, batch.counts, self.theta, self.bp, B=B, K=K, V=V,
                       gscale=gscale, rl=self.rl, loss=self.loss,
                       dtheta=self.dtheta_lda if grad else None, gnnz=self.gnnz if grad else None)
            dtheta = self.dtheta_lda[:B] if grad else None
            if grad:
                K_.csc_spmm(self.csc.colptr, self.csc.row, self.csc.src, self.gnnz, self.theta, self.dbpT,
                            V=V, N=K)
                K_.lda_beta_bwd(self.bp, self.dbpT, self.xhat, self.col_rstd, self.dxhat, self._g("beta"),
                                K=K, V=V, training=training)
        if self.L and labels is not None:
            # classification head on theta: mean cross-entropy joins the loss, its gradient joins d theta
            wc, bcl = "label_classification.weight", "label_classification.bias"
            K_.label_head(self.theta, self._p(wc), self._p(bcl), labels[:B].contiguous(), B=B, gscale=gscale,
                          loss=self.loss, dtheta=dtheta if grad else None, dWc=self._g(wc) if grad else None,
                          dbc=self._g(bcl) if grad else None)
        if loss_host is not None:
            # the loss is final here: its copy to the pinned slot overlaps the rest of the backward pass
            if side is not None:
                side.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(side):
                    loss_host.copy_(self.loss, non_blocking=True)
                self._side_join = True
            else:
                loss_host.copy_(self.loss, non_blocking=True)
        if not grad:
            self.join_side()
            return self.loss

        learn = self.model.learn_priors
        K_.latent_bwd(B=B, K=K, dtheta=dtheta, mask_theta=self.mask_theta, theta_s=self.theta_s, mu=self.mu,
                      logvar=self.logvar, eps=self.eps, prior_mean=self.model.prior_mean,
                      prior_var=self.model.prior_variance, kl_weight=self.kl_weight, gscale=gscale,
                      dmu=self.dmu, dlv=self.dlv, bn_bsums=self.bn_bsums,
                      d_prior_mean=self._g("prior_mean") if learn else None,
                      d_prior_var=self._g("prior_variance") if learn else None)
        dWh, dbh = self._hidden_params(self.flat.G)
        gst = self.flat.storage_view(self.flat.G, "inf_net.input_layer.weight")
        K_.enc_bwd(B=B, V=V, K=K, hidden=self.hidden, act=self.act, dmu=self.dmu, dlv=self.dlv, mu=self.mu,
                   logvar=self.logvar, bn_sums=self.bn_sums, bn_bsums=self.bn_bsums, bn_eps=BN_EPS,
                   pre=self.pre + pad, mask_enc=self.mask_enc, Wh=Wh, Wmu=self._p("inf_net.f_mu.weight"),
                   Wsig=self._p("inf_net.f_sigma.weight"), dWh=dWh, dbh=dbh,
                   dWmu=self._g("inf_net.f_mu.weight"), dbmu=self._g("inf_net.f_mu.bias"),
                   dWsig=self._g("inf_net.f_sigma.weight"), dbsig=self._g("inf_net.f_sigma.bias"),
                   db1=self._g("inf_net.input_layer.bias"), dpre0=self.dpre0, xdense=None, dWdT=None,
                   seed=seed, step_ctr=self.ctr)
        if self.L and labels is not None:
            K_.small_linear_wgrad(labels[:B].contiguous(), self.dpre0[:B], gst[gst.shape[0] - self.L:])
        if self.kind in ("zeroshot", "combined"):
            # d W^T [C,H0] = x_ctx^T . d pre0 on the tensor cores (accumulates into a zeroed target)
            if self.kind == "zeroshot":
                dWdT = gst[:self.C]
                K_.ctm_init_out(dWdT, bias_row=-1, bias=None)
            else:
                dWdT = self.ctm_dMx[:self.C]  # (zeroed with the step's other accumulators)
            K_.ctm_split(self.dpre0[:B], self.ctx_g_hi, self.ctx_g_lo)
            K_.ctm_vc(self.ctx_x_hi[:B], self.ctx_x_lo[:B], self.ctx_g_hi[:B], self.ctx_g_lo[:B], dWdT, m_total=self.C,
                      n_valid=self.hidden[0])
        if use_bow:
            # d W1a^T[v,:] = sum_b x[b,v] * dpre0[b,:]  (all V rows written, zeros for unused words)
            K_.csc_spmm(self.csc.colptr, self.csc.row, self.csc.src, batch.counts, self.dpre0, gst[:V],
                        V=V, N=self.hidden[0])
        if self.kind == "combined":
            # gradient wrt the collapsed weight (+ d b1 as its last row) -> exact gradients of adapt_bert and of
            # the contextual half of the first layer.  Both products are linear in [dM^T ; dsum], so a federated
            # run sums THAT over the clients (ctm_reduce_hook: 150 KB) instead of the 38 M-element gradients.
            self.ctm_dMx[self.C].copy_(self._g("inf_net.input_layer.bias"))        # dsum = sum_b dpre0
            if ctm_reduce_hook is not None:
                ctm_reduce_hook(self.ctm_dM_flat)
            K_.ctm_split(self.ctm_dMx[:self.C], self.ctm_d_hi, self.ctm_d_lo, tail_row=self.ctm_dMx[self.C])
            K_.ctm_rc(self.ctm_a_hi, self.ctm_a_lo, self.ctm_d_hi, self.ctm_d_lo, gst[V:2 * V], k_total=self.C + 1,
                      n_valid=self.hidden[0])
            K_.ctm_rr(self.ctm_w_hi, self.ctm_w_lo, self.ctm_d_hi, self.ctm_d_lo,
                      self._g("inf_net.adapt_bert.weight"), n_valid=self.C, extra=self._g("inf_net.adapt_bert.bias"))
        return self.loss

    # ------------------------------------------------------------------------------------------
    # optimiser
    # ------------------------------------------------------------------------------------------
    def optimizer_step(self, gscale: float = 1.0, beta_done: bool = False, ctr_done: bool = False) -> None:
        """Fused Adam over the flat buffers; refreshes the bf16 shadows of beta.  ``beta_done``: the step counter was
        advanced and beta updated already (overlapped with the d-theta sweep, see ``train_step``)."""
        if not torch.cuda.is_current_stream_capturing():
            self.step_count += 1
        if not (beta_done or ctr_done):
            K_.step_begin(self.ctr, 0, 1)
        fl = self.flat
        b1, b2 = self.betas
        for lo, hi in self._flat_adam_pieces():
            if hi > lo:
                K_.adam_flat(fl.P[lo:hi], fl.G[lo:hi], fl.M[lo:hi], fl.V2[lo:hi], lr=self.lr, b1=b1, b2=b2,
                             eps=self.adam_eps, wd=self.weight_decay, step=self.step_count, gscale=gscale,
                             step_ctr=self.ctr, solver=self.solver)
        if self.kind == "combined":
            self.ctm_adam_step(gscale=gscale)
        self._refresh_ctx_shadows()
        K, V = self.K, self.V
        if self._side_beta:   # beta was stepped on the side stream during the backward pass: join it
            self._side_join = True
            self._side_beta = False
            self.join_side()
        elif beta_done:
            pass
        elif self.is_prod:
            K_.adam_beta(fl.view(fl.P, "beta"), fl.view(fl.G, "beta"), fl.view(fl.M, "beta"),
                         fl.view(fl.V2, "beta"), self.beta_bf, self.betaT_bf, K=K, V=V, lr=self.lr, b1=b1,
                         b2=b2, eps=self.adam_eps, wd=self.weight_decay, step=self.step_count, gscale=gscale,
                         step_ctr=self.ctr, solver=self.solver)
        else:
            s = fl.segments["beta"]
            sl = slice(s.offset, s.offset + s.numel)
            K_.adam_flat(fl.P[sl], fl.G[sl], fl.M[sl], fl.V2[sl], lr=self.lr, b1=b1, b2=b2,
                         eps=self.adam_eps, wd=self.weight_decay, step=self.step_count, gscale=gscale,
                         step_ctr=self.ctr, solver=self.solver)

    def join_side(self) -> None:
        """The compute stream waits for whatever this step put on the side stream (called at the end of a step)."""
        if self._side_join and self._side is not None:
            torch.cuda.current_stream().wait_stream(self._side)
        self._side_join = False

    def loss_slot_for(self, key) -> torch.Tensor:
        """The pinned host slot (float64[1]) that the training step with CUDA-graph key ``key`` writes its loss to."""
        slot = self._loss_slots.get(key)
        if slot is None:
            slot = torch.zeros(1, dtype=torch.float64)
            if self.device.type == "cuda":
                slot = slot.pin_memory()
            self._loss_slots[key] = slot
        return slot

    def _adam_beta_now(self, gscale: float = 1.0, background: bool = False) -> None:
        """The fused optimiser step of beta + shadows; the device step counter must already hold step t.
        ``background``: the variant that fits on the SMs beside the d-theta sweep (side stream)."""
        fl = self.flat
        b1, b2 = self.betas
        if background and self.betaT_bf.shape[1] % 16 == 0:
            K_.adam_beta_bg(fl.view(fl.P, "beta"), fl.view(fl.G, "beta"), fl.view(fl.M, "beta"), fl.view(fl.V2, "beta"),
                            self.beta_bf, self.betaT_bf, K=self.K, V=self.V, lr=self.lr, b1=b1, b2=b2, eps=self.adam_eps,
                            wd=self.weight_decay, step=self.step_count + 1, gscale=gscale, step_ctr=self.ctr,
                            solver=self.solver, blocks_per_sm=int(os.environ.get("GFEDNTM_BG_BLOCKS", "2")))
            return
        K_.adam_beta(fl.view(fl.P, "beta"), fl.view(fl.G, "beta"), fl.view(fl.M, "beta"), fl.view(fl.V2, "beta"),
                     self.beta_bf, self.betaT_bf, K=self.K, V=self.V, lr=self.lr, b1=b1, b2=b2, eps=self.adam_eps,
                     wd=self.weight_decay, step=self.step_count + 1, gscale=gscale, step_ctr=self.ctr, solver=self.solver)

    def optimizer_step_beta(self, gscale: float = 1.0) -> None:
        """Adam on beta alone (federated runs that do not share beta); the caller advanced the step counter."""
        fl = self.flat
        b1, b2 = self.betas
        if self.is_prod:
            K_.adam_beta(fl.view(fl.P, "beta"), fl.view(fl.G, "beta"), fl.view(fl.M, "beta"),
                         fl.view(fl.V2, "beta"), self.beta_bf, self.betaT_bf, K=self.K, V=self.V, lr=self.lr, b1=b1,
                         b2=b2, eps=self.adam_eps, wd=self.weight_decay, step=self.step_count + 1, gscale=gscale,
                         step_ctr=self.ctr, solver=self.solver)
        else:
            sg = fl.segments["beta"]
            sl = slice(sg.offset, sg.offset + sg.numel)
            K_.adam_flat(fl.P[sl], fl.G[sl], fl.M[sl], fl.V2[sl], lr=self.lr, b1=b1, b2=b2, eps=self.adam_eps,
                         wd=self.weight_decay, step=self.step_count + 1, gscale=gscale, step_ctr=self.ctr, solver=self.solver)

    def train_step(self, batch: CsrBatch, x_ctx: Optional[torch.Tensor] = None, labels: Optional[torch.Tensor] = None,
                   **kw) -> torch.Tensor:
        """Forward + backward + Adam on one minibatch.  The kernel sequence is captured into a CUDA
        graph per (input buffers, batch size) and replayed: the first call on a new key runs eagerly
        (warm-up), the second captures, later ones replay.  Injected noise/masks force the eager path."""
        key = ("train", batch.indptr.data_ptr(), batch.indices.data_ptr(), batch.B,
               x_ctx.data_ptr() if x_ctx is not None else 0, labels.data_ptr() if labels is not None else 0)
        self.loss_slot = slot = self.loss_slot_for(key if not kw else "eager")
        kw = dict(kw, loss_host=slot)

        def body():
            # the optimiser-step counter advances with the first launch of the step (inc_opt=1): the update of beta
            # inside / beside the d-theta sweep reads step t, and the step saves a launch
            if self._self_overlap is not None:
                self.forward_backward(batch, x_ctx, training=True, fed=self._self_overlap, labels=labels, inc_opt=1, **kw)
                self.optimizer_step(beta_done=True, ctr_done=True)
            elif self._side is not None and self.is_prod and self.decoder.folded:
                self.forward_backward(batch, x_ctx, training=True, labels=labels, side_stream=True, inc_opt=1,
                                      beta_on_side=True, **kw)
                self.optimizer_step(beta_done=True, ctr_done=True)
            else:
                self.forward_backward(batch, x_ctx, training=True, labels=labels, inc_opt=1,
                                      side_stream=self._side is not None, **kw)
                self.optimizer_step(ctr_done=True)
            self.join_side()

        if len(kw) > 1 or not self.use_graphs:
            body()
            return self.loss
        return self.run_graphed(key, body)

    def run_graphed(self, key, body, n_opt_steps: int = 1) -> torch.Tensor:
        """Runs ``body`` (a pure sequence of launches on static buffers) through a cached CUDA graph."""
        graph = self._graphs.get(key)
        if graph is None:
            if key not in self._warm:
                self._warm.add(key)
                body()          # eager warm-up (this is a real step; host counters advance inside)
                return self.loss
            from ..ops import _lib

            graph = torch.cuda.CUDAGraph()
            torch.cuda.synchronize()
            n0 = _lib.launch_count()
            with torch.cuda.graph(graph):
                body()          # capture only: nothing executes, host counters are left alone
            self._graph_launches[key] = _lib.launch_count() - n0
            self._graphs[key] = graph
        graph.replay()
        self.replayed_launches += self._graph_launches[key]
        self.step_count += n_opt_steps
        self._nbt_pending += 1
        return self.loss

    # ------------------------------------------------------------------------------------------
    # inference
    # ------------------------------------------------------------------------------------------
    def sample_theta(self, batch: CsrBatch, x_ctx: Optional[torch.Tensor] = None, n_samples: int = 1,
                     thr: float = 0.0, labels: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Mean of ``n_samples`` stochastic draws of theta in eval mode (running BN statistics, no dropout): ONE
        encoder pass and ONE Monte-Carlo kernel for all draws; ``thr > 0`` also applies the result extraction's
        threshold + L1 renormalisation on the device.  Ref: avitm.py:470-523, federated_model.py:170-173."""
        B, K, V = batch.B, self.K, self.V
        acc = torch.empty(B, K, dtype=torch.float32, device=self.device)
        W1T, b1, pre0_ext = self._first_layer(x_ctx, B)
        Wh, bh = self._hidden_params(self.flat.P)
        use_bow = self.kind in ("bow", "combined")
        if use_bow:
            # sparse BoW part of the first layer: several warps per document, added to the contextual part
            if pre0_ext is None:
                pre0_ext = self.bow_pre0
            K_.bow_gather(batch.indptr, batch.indices, batch.counts, W1T, pre0_ext, B=B,
                          accumulate=self.kind == "combined")
            W1T = None   # (enc_fwd gets no CSR input any more: everything of layer 0 arrives through pre0_ext)
        if self.L and labels is not None:
            # supervised CTM: the one-hot labels are the last label_size inputs of the first layer
            st_all = self.flat.storage_view(self.flat.P, "inf_net.input_layer.weight")
            K_.small_linear_fwd(labels[:B].contiguous(), st_all[st_all.shape[0] - self.L:], pre0_ext, accumulate=True)
        pad = [None] * (K_.GF_MAX_HIDDEN - len(self.hidden))
        K_.enc_fwd(B=B, V=V, K=K, hidden=self.hidden, act=self.act, training=False,
                   indptr=None, indices=None, counts=None, W1T=W1T, b1=b1, xdense=None, WdT=None,
                   pre0_ext=pre0_ext,
                   Wh=Wh, bh=bh, Wmu=self._p("inf_net.f_mu.weight"), bmu=self._p("inf_net.f_mu.bias"),
                   Wsig=self._p("inf_net.f_sigma.weight"), bsig=self._p("inf_net.f_sigma.bias"),
                   pre=self.pre + pad, enc_drop_p=0.0, mask_enc_in=None, mask_enc=None,
                   mu_pre=self.mu_pre, ls_pre=self.ls_pre, bn_sums=None, seed=0)
        mu_rm, mu_rv, _ = self._bn("f_mu_batchnorm")
        ls_rm, ls_rv, _ = self._bn("f_sigma_batchnorm")
        seed = (self.seed * 0x9E3779B97F4A7C15 + self.rng_calls * 0xD1B54A32D192ED03 + 7) & ((1 << 64) - 1)
        self.rng_calls += n_samples
        K_.theta_mc(self.mu_pre, self.ls_pre, mu_rm, mu_rv, ls_rm, ls_rv, acc, B=B, K=K, n_samples=n_samples, seed=seed,
                    bn_eps=BN_EPS, thr=thr)
        return acc
