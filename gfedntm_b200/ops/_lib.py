"""ctypes binding of ``libgfedntm_kernels.so`` (C ABI declared in ``include/gfedntm_kernels.h``).

The library is loaded lazily.  On a machine with a CUDA device the kernels are the only code path:
if the shared object is missing or its struct layout does not match, loading raises instead of
silently falling back to PyTorch.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path
from typing import Optional

GF_MAX_HIDDEN = 8

_p = C.c_void_p
_i = C.c_int
_f = C.c_float
_u64 = C.c_ulonglong


class GfSweepArgs(C.Structure):
    _fields_ = [
        ("mode", _i), ("B", _i), ("V", _i), ("K", _i), ("Kp", _i), ("n_chunks", _i),
        ("theta_bf", _p), ("ld_theta", _i),
        ("beta_bf", _p), ("ld_beta", _i),
        ("sb_stream", _p),
        ("out", _p), ("lse_part", _p), ("doc_const", _p), ("dtheta", _p), ("colc", _p),
        ("theta_mean", _p), ("running_mean", _p), ("running_var", _p),
        ("bn_eps", _f), ("bn_momentum", _f), ("use_running", _i),
        ("asp", _p), ("csp", _p), ("dbeta", _p), ("ld_dbeta", _i),
        ("corr", _p), ("colptr", _p),
        ("theta_cols", _i), ("beta_cols", _i), ("ext_col0", _i),
        ("betaT_bf", _p), ("beta3", _p), ("ld_beta3", _i), ("ld_corr", _i),
        ("fed", _p), ("fed_ctas", _i),
        ("fed_tmaps", _p), ("dbeta_bf", _p), ("ld_dbeta_bf", _i),
    ]


class GfEncArgs(C.Structure):
    _fields_ = [
        ("B", _i), ("V", _i), ("K", _i), ("L", _i),
        ("H", _i * GF_MAX_HIDDEN),
        ("act", _i), ("training", _i),
        ("indptr", _p), ("indices", _p), ("counts", _p),
        ("W1T", _p), ("b1", _p),
        ("xdense", _p), ("C", _i), ("WdT", _p),
        ("Wh", _p * GF_MAX_HIDDEN), ("bh", _p * GF_MAX_HIDDEN),
        ("Wmu", _p), ("bmu", _p), ("Wsig", _p), ("bsig", _p),
        ("pre", _p * GF_MAX_HIDDEN),
        ("enc_drop_p", _f),
        ("mask_enc_in", _p), ("mask_enc", _p),
        ("mu_pre", _p), ("ls_pre", _p),
        ("bn_sums", _p),
        ("seed", _u64),
        ("step_ctr", _p),
        ("w_smem", _i),
        ("pre0_ext", _p),
    ]


class GfLatentArgs(C.Structure):
    _fields_ = [
        ("B", _i), ("K", _i), ("training", _i), ("sample", _i),
        ("mu_pre", _p), ("ls_pre", _p), ("bn_sums", _p),
        ("mu_rm", _p), ("mu_rv", _p), ("ls_rm", _p), ("ls_rv", _p),
        ("bn_eps", _f), ("bn_momentum", _f),
        ("prior_mean", _p), ("prior_var", _p),
        ("eps_in", _p), ("mask_theta_in", _p),
        ("theta_drop_p", _f),
        ("mu", _p), ("logvar", _p), ("eps", _p), ("theta_s", _p), ("mask_theta", _p), ("theta", _p),
        ("theta_colsum", _p), ("kl", _p), ("loss", _p),
        ("kl_weight", _f),
        ("seed", _u64),
        ("step_ctr", _p),
    ]


class GfLatentBwdArgs(C.Structure):
    _fields_ = [
        ("B", _i), ("K", _i),
        ("dtheta", _p), ("ld_dtheta", _i),
        ("mask_theta", _p), ("theta_s", _p), ("mu", _p), ("logvar", _p), ("eps", _p),
        ("prior_mean", _p), ("prior_var", _p),
        ("kl_weight", _f), ("gscale", _f),
        ("dmu", _p), ("dlv", _p), ("bn_bsums", _p),
        ("d_prior_mean", _p), ("d_prior_var", _p),
    ]


class GfEncBwdArgs(C.Structure):
    _fields_ = [
        ("B", _i), ("V", _i), ("K", _i), ("L", _i),
        ("H", _i * GF_MAX_HIDDEN),
        ("act", _i),
        ("dmu", _p), ("dlv", _p), ("mu", _p), ("logvar", _p),
        ("bn_sums", _p), ("bn_bsums", _p), ("bn_eps", _f),
        ("pre", _p * GF_MAX_HIDDEN), ("mask_enc", _p),
        ("Wh", _p * GF_MAX_HIDDEN), ("Wmu", _p), ("Wsig", _p),
        ("dWh", _p * GF_MAX_HIDDEN), ("dbh", _p * GF_MAX_HIDDEN),
        ("dWmu", _p), ("dbmu", _p), ("dWsig", _p), ("dbsig", _p), ("db1", _p),
        ("dpre0", _p),
        ("xdense", _p), ("C", _i), ("dWdT", _p),
        ("seed", _u64),
        ("step_ctr", _p),
    ]


class GfDecSparseArgs(C.Structure):
    _fields_ = [
        ("B", _i), ("Kp", _i), ("n_chunks", _i),
        ("indptr", _p), ("indices", _p), ("counts", _p),
        ("theta_bf", _p), ("ld_theta", _i),
        ("betaT_bf", _p), ("colc", _p), ("lse_part", _p),
        ("gscale", _f),
        ("rl", _p), ("loss", _p), ("asp", _p), ("csp", _p), ("tnnz", _p), ("dtheta", _p),
        ("doc_const", _p), ("sb_doc", _p),
        ("thetaT_bf", _p), ("ldT", _i), ("K", _i),
    ]


GF_MAX_RANKS = 8


class GfFabricRank(C.Structure):
    _fields_ = [
        ("rank", _i),
        ("G", _p * GF_MAX_RANKS), ("P", _p * GF_MAX_RANKS), ("S", _p * GF_MAX_RANKS),
        ("beta_bf", _p * GF_MAX_RANKS), ("betaT_bf", _p * GF_MAX_RANKS), ("flags", _p * GF_MAX_RANKS),
        ("M", _p), ("V2", _p), ("done_counter", _p),
        ("P_mc", _p), ("S_mc", _p), ("beta_bf_mc", _p), ("betaT_bf_mc", _p),
        ("Gb", _p * GF_MAX_RANKS), ("ld_gb", _i), ("agg_counter", _p),
        ("AR", _p * GF_MAX_RANKS), ("ar_cap", C.c_longlong),
    ]


GF_FED_MAX_RANGES = 48
GF_FED_FLAG_SLOTS = 8
GF_FED_FLAG_BYTES = GF_FED_FLAG_SLOTS * GF_MAX_RANKS * 4


class GfFedRange(C.Structure):
    _fields_ = [("off", C.c_longlong), ("len", C.c_longlong), ("shared", _i), ("pad_", _i)]


class GfFedArgs(C.Structure):
    _fields_ = [
        ("world", _i), ("mode", _i), ("epoch", _i), ("n_rank_ctx", _i),
        ("flat_n", C.c_longlong), ("beta_off", C.c_longlong), ("s_n", C.c_longlong),
        ("K", _i), ("V", _i), ("Kp", _i), ("ldb", _i),
        ("w", _f * GF_MAX_RANKS),
        ("lr", _f), ("b1", _f), ("b2", _f), ("eps", _f), ("wd", _f),
        ("step", _i),
        ("step_ctr", _p),
        ("parts", _i),
        ("n_ranges", _i),
        ("ranges", GfFedRange * GF_FED_MAX_RANGES),
        ("s_mask", _i), ("beta_shared", _i), ("local_gscale", _f), ("solver", _i),
        ("ranks", GfFabricRank * GF_MAX_RANKS),
    ]


_STRUCTS = [GfSweepArgs, GfEncArgs, GfLatentArgs, GfLatentBwdArgs, GfEncBwdArgs, GfDecSparseArgs,
            GfFabricRank, GfFedArgs, GfFedRange]

LIB_PATH = Path(__file__).resolve().parent / "libgfedntm_kernels.so"
_lib: Optional[C.CDLL] = None


class KernelLibraryError(RuntimeError):
    pass


def _declare(lib: C.CDLL) -> None:
    ll = C.c_longlong
    lib.gf_abi_version.restype = _i
    lib.gf_launch_count.restype = ll
    lib.gf_sizeof.argtypes = [_i]
    lib.gf_sizeof.restype = _i
    lib.gf_dec_sweep.argtypes = [C.POINTER(GfSweepArgs), _p]
    lib.gf_enc_fwd.argtypes = [C.POINTER(GfEncArgs), _p]
    lib.gf_latent_fwd.argtypes = [C.POINTER(GfLatentArgs), _p]
    lib.gf_theta_center.argtypes = [_p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _p]
    lib.gf_theta_center_fold.argtypes = [_p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _i, _p]
    lib.gf_scale_beta.argtypes = [_p, _i, _p, _p, _i, _i, _i, _i, _i, _p]
    lib.gf_theta_gram.argtypes = [_p, _i, _i, _i, _p, _i, _i, _i, _p, _p]
    lib.gf_gram_parts.argtypes = []
    lib.gf_gram_parts.restype = _i
    lib.gf_gram.argtypes = [_p, _i, _p, _i, _i, _i, _p, _p]
    lib.gf_gram_finish.argtypes = [_p, _i, _i, _i, _p, _i, _i, _i, _p]
    lib.gf_csc_corr_t.argtypes = [_p, _p, _p, _p, _p, _i, _p, _i, _i, _i, _p]
    for fn in ("gf_theta_center_fold", "gf_scale_beta", "gf_theta_gram", "gf_gram", "gf_gram_finish", "gf_csc_corr_t"):
        getattr(lib, fn).restype = _i
    lib.gf_ctm_split.argtypes = [_p, ll, _p, _p, _i, _i, _p, _p, _i, _p]
    lib.gf_ctm_init_out.argtypes = [_p, _i, _i, _i, _i, _p, _p]
    lib.gf_ctm_vc.argtypes = [_p, _p, _i, _i, _p, _p, _i, _i, _p, _i, _i, _p]
    lib.gf_ctm_rc.argtypes = [_p, _p, _i, _i, _p, _p, _i, _i, _i, _p, _i, _i, _p]
    lib.gf_ctm_rr.argtypes = [_p, _p, _p, _p, _i, _i, _i, _p, _i, _i, _p, _p]
    lib.gf_adam_split.argtypes = [_p, _p, _p, _p, ll, _i, _i, _p, _p, _p, _p, _p, _p, _i, _f, _f, _f, _f, _f, _i, _f, _i,
                                  _p, _i, _p]
    for fn in ("gf_ctm_split", "gf_ctm_init_out", "gf_ctm_vc", "gf_ctm_rc", "gf_ctm_rr", "gf_adam_split"):
        getattr(lib, fn).restype = _i
    lib.gf_small_linear_fwd.argtypes = [_p, _p, _p, _i, _i, _i, _i, _p]
    lib.gf_small_linear_wgrad.argtypes = [_p, _p, _p, _i, _i, _i, _p]
    lib.gf_label_head.argtypes = [_p, _p, _p, _p, _i, _i, _i, _f, _p, _p, _i, _p, _p, _p]
    for fn in ("gf_small_linear_fwd", "gf_small_linear_wgrad", "gf_label_head"):
        getattr(lib, fn).restype = _i
    lib.gf_bow_gather.argtypes = [_p, _p, _p, _p, _i, _i, _p, _i, _p]
    lib.gf_bow_gather.restype = _i
    lib.gf_theta_mc.argtypes = [_p, _p, _p, _p, _p, _p, _f, _i, _i, _i, _u64, _f, _p, _p]
    lib.gf_theta_mc.restype = _i
    lib.gf_fed_signal.argtypes = [C.POINTER(GfFedArgs), _i, _p]
    lib.gf_fed_signal.restype = _i
    lib.gf_fed_make_tmaps.argtypes = [C.POINTER(GfFedArgs), _p]
    lib.gf_fed_make_tmaps.restype = _i
    lib.gf_fed_gather_beta.argtypes = [C.POINTER(GfFedArgs), _p]
    lib.gf_fed_gather_beta.restype = _i
    lib.gf_fed_allreduce_small.argtypes = [C.POINTER(GfFedArgs), _p, C.c_longlong, _p]
    lib.gf_fed_allreduce_small.restype = _i
    lib.gf_dec_row_chunks_grid.argtypes = [_i, _i, _i]
    lib.gf_dec_row_chunks_grid.restype = _i
    lib.gf_num_sms.argtypes = []
    lib.gf_num_sms.restype = _i
    lib.gf_dec_row_chunks.argtypes = [_i, _i, _i, _i]
    lib.gf_dec_row_chunks.restype = _i
    lib.gf_latent_bwd.argtypes = [C.POINTER(GfLatentBwdArgs), _p]
    lib.gf_enc_bwd.argtypes = [C.POINTER(GfEncBwdArgs), _p]
    lib.gf_csr_to_csc.argtypes = [_p, _p, _i, _i, _i, _p, _p, _p, _p, _p]
    lib.gf_step_begin.argtypes = [_p, _i, _i, _p]
    lib.gf_step_begin.restype = _i
    lib.gf_step_begin_zero.argtypes = [_p, _i, _i, C.POINTER(C.c_void_p), C.POINTER(C.c_longlong), _i, _p]
    lib.gf_step_begin_zero.restype = _i
    lib.gf_csr_to_csc_dyn.argtypes = [_p, _p, _i, _i, _i, _p, _p, _p, _p, _p]
    lib.gf_csr_to_csc_dyn.restype = _i
    lib.gf_csr_to_csc_static.argtypes = [_p, _p, _i, _i, _i, _p, _p, _p, _p]
    lib.gf_csr_to_csc_static.restype = _i
    lib.gf_dec_sparse_doc.argtypes = [C.POINTER(GfDecSparseArgs), _p]
    lib.gf_csc_corr.argtypes = [_p, _p, _p, _p, _p, _i, _p, _i, _i, _p]
    lib.gf_csc_spmm.argtypes = [_p, _p, _p, _p, _p, _i, _p, _i, _i, _i, _i, _p]
    lib.gf_lda_beta_fwd.argtypes = [_p, _i, _i, _i, _f, _f, _i, _p, _p, _p, _p, _p, _p]
    lib.gf_lda_doc.argtypes = [_p, _p, _p, _p, _p, _i, _i, _i, _f, _p, _p, _p, _i, _p, _p]
    lib.gf_lda_beta_bwd.argtypes = [_p, _p, _i, _p, _p, _p, _p, _i, _i, _i, _i, _p]
    lib.gf_row_topk.argtypes = [_p, _i, _i, _i, _i, _p, _p, _p]
    lib.gf_adam_flat.argtypes = [_p, _p, _p, _p, ll, _f, _f, _f, _f, _f, _i, _f, _p, _i, _p]
    lib.gf_adam_beta.argtypes = [_p, _p, _p, _p, _i, _i, _i, _i, _p, _i, _p, _i,
                                 _f, _f, _f, _f, _f, _i, _f, _i, _p, _i, _p]
    lib.gf_adam_beta_bg.argtypes = [_p, _p, _p, _p, _i, _i, _i, _i, _p, _i, _p, _i,
                                    _f, _f, _f, _f, _f, _i, _f, _p, _i, _i, _p]
    lib.gf_adam_beta_bg.restype = _i
    lib.gf_fed_aggregate.argtypes = [C.POINTER(GfFedArgs), _p]
    lib.gf_fed_aggregate.restype = _i
    lib.gf_host_gather_csr.argtypes = [_p, _p, _p, _p, _i, _p, _p, _p, ll, _i]
    lib.gf_host_gather_rows.argtypes = [_p, ll, _p, _i, _p]
    for fn in ("gf_host_gather_csr", "gf_host_gather_rows", "gf_dec_sweep", "gf_enc_fwd", "gf_latent_fwd", "gf_theta_center", "gf_latent_bwd",
               "gf_enc_bwd", "gf_csr_to_csc", "gf_dec_sparse_doc", "gf_csc_corr", "gf_csc_spmm",
               "gf_lda_beta_fwd", "gf_lda_doc", "gf_lda_beta_bwd", "gf_row_topk", "gf_adam_flat",
               "gf_adam_beta"):
        getattr(lib, fn).restype = _i


def load(build_if_missing: bool = True) -> C.CDLL:
    """Load (building first if needed and possible) the kernel library; raises on failure."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists() and build_if_missing:
        from .build import build

        build()
    if not LIB_PATH.exists():
        raise KernelLibraryError(
            f"{LIB_PATH} is missing: run `python -m gfedntm_b200.ops.build` (needs nvcc)")
    # GFEDNTM_KERNEL_LIB: load an alternative build of the same ABI (kernel A/B experiments)
    lib = C.CDLL(os.environ.get("GFEDNTM_KERNEL_LIB") or str(LIB_PATH))
    _declare(lib)
    for idx, st in enumerate(_STRUCTS):
        want = lib.gf_sizeof(idx)
        if want != C.sizeof(st):
            raise KernelLibraryError(
                f"struct layout mismatch for {st.__name__}: library {want} bytes, ctypes {C.sizeof(st)}")
    # optional entry points added after the first ABI revision are declared by their modules
    _lib = lib
    return lib


def is_loaded() -> bool:
    return _lib is not None


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise KernelLibraryError(f"{what} failed with code {rc} (see stderr for the CUDA/argument error)")


def launch_count() -> int:
    return int(load().gf_launch_count())
