"""tcgen05 skinny GEMMs of the collapsed CombinedTM encoder (csrc/ctm_gemm.cu) against fp32 PyTorch.

Ref op: ctm_network/inference_network.py:160-193 (adapt_bert -> concat -> input_layer) and its autograd."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _split(x, ld, extra_col=None, tail_row=None):
    from gfedntm_b200.ops import kernels as K_

    rows = x.shape[0] + (1 if tail_row is not None else 0)
    hi = torch.full((rows, ld), 7.0, dtype=torch.bfloat16, device="cuda")  # garbage: the kernel must overwrite it
    lo = torch.full((rows, ld), 7.0, dtype=torch.bfloat16, device="cuda")
    K_.ctm_split(x, hi, lo, extra_col=extra_col, tail_row=tail_row)
    return hi, lo


@pytest.mark.parametrize("V,C,H", [(1000, 192, 50), (5000, 768, 50), (700, 100, 100), (50000, 768, 50)])
def test_ctm_gemms_match_fp32(V, C, H):
    from gfedntm_b200.ops import kernels as K_

    torch.manual_seed(V + C)
    dd = dict(dtype=torch.float32, device="cuda")
    A = torch.randn(V, C, **dd) * 0.05
    ab = torch.randn(V, **dd) * 0.1
    W = torch.randn(V, H, **dd) * 0.1
    b1 = torch.randn(H, **dd)
    dMT = torch.randn(C, H, **dd)
    dsum = torch.randn(H, **dd)
    Cp = (C + 1 + 63) // 64 * 64
    Np = (H + 63) // 64 * 64
    a_hi, a_lo = _split(A, Cp, extra_col=ab)
    w_hi, w_lo = _split(W, Np)
    # the shadows reproduce the fp32 values to ~16 mantissa bits
    torch.testing.assert_close((a_hi.float() + a_lo.float())[:, :C], A, rtol=2e-5, atol=1e-7)
    torch.testing.assert_close((a_hi.float() + a_lo.float())[:, C], ab, rtol=2e-5, atol=1e-7)
    assert float(a_hi[:, C + 1:].abs().max()) == 0.0 if Cp > C + 1 else True

    # ---- forward: [M^T ; b_eff] = [A | ab]^T W (+ b1 in the last row)
    MTx = torch.empty(C + 1, H, **dd)
    K_.ctm_init_out(MTx, bias_row=C, bias=b1)
    K_.ctm_vc(a_hi, a_lo, w_hi, w_lo, MTx, m_total=C + 1, n_valid=H)
    want_MT = (A.double().t() @ W.double()).float()
    want_b = (b1.double() + ab.double() @ W.double()).float()
    scale = float(want_MT.abs().max())
    assert float((MTx[:C] - want_MT).abs().max()) < 2e-5 * scale + 1e-6
    assert float((MTx[C] - want_b).abs().max()) < 2e-5 * float(want_b.abs().max()) + 1e-6

    # ---- backward
    d_hi, d_lo = _split(dMT, Np, tail_row=dsum)   # [C+1, Np]: rows = contextual dims (+ dsum), hidden contiguous
    dW = torch.full((V, H), 3.0, **dd)
    K_.ctm_rc(a_hi, a_lo, d_hi, d_lo, dW, k_total=C + 1, n_valid=H)
    want_dW = (A.double() @ dMT.double() + torch.outer(ab.double(), dsum.double())).float()
    assert float((dW - want_dW).abs().max()) < 2e-5 * float(want_dW.abs().max()) + 1e-6
    dA = torch.full((V, C), 3.0, **dd)
    dab = torch.full((V,), 3.0, **dd)
    K_.ctm_rr(w_hi, w_lo, d_hi, d_lo, dA, n_valid=C, extra=dab)
    want_dA = (W.double() @ dMT.double().t()).float()
    want_dab = (W.double() @ dsum.double()).float()
    assert float((dA - want_dA).abs().max()) < 2e-5 * float(want_dA.abs().max()) + 1e-6
    assert float((dab - want_dab).abs().max()) < 2e-5 * float(want_dab.abs().max()) + 1e-6
