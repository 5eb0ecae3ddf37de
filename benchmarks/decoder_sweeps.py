#!/usr/bin/env python
"""Micro-benchmark / profiling driver for the ProdLDA decoder (the four tcgen05 sweeps + sparse parts).

    python benchmarks/decoder_sweeps.py --B 4096 --V 200000 --K 200 --iters 5

Prints CUDA-event times per decoder forward+backward; keep it short when running under ncu."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--B", type=int, default=4096)
    ap.add_argument("--V", type=int, default=200000)
    ap.add_argument("--K", type=int, default=200)
    ap.add_argument("--nnz", type=int, default=180)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--breakdown", action="store_true", help="CUDA-event time of every kernel call of one step")
    a = ap.parse_args()
    from gfedntm_b200.ops.decoder import (CscWorkspace, CsrBatch, ProdLDADecoder, make_beta_shadows,
                                          refresh_beta_shadows)

    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(0)
    B, V, K = a.B, a.V, a.K
    theta = torch.softmax(torch.randn(B, K, device=dev, generator=g), 1)
    beta = torch.randn(K, V, device=dev, generator=g) * 0.3
    idx = torch.randint(0, V, (B, a.nnz), device=dev, generator=g).sort(1).values
    indptr = (torch.arange(B + 1, device=dev) * a.nnz).to(torch.int32)
    batch = CsrBatch(indptr, idx.reshape(-1).to(torch.int32), torch.ones(B * a.nnz, device=dev), B, V)
    beta_bf, betaT_bf = make_beta_shadows(K, V, dev)
    refresh_beta_shadows(beta, beta_bf, betaT_bf)
    dec = ProdLDADecoder(B, V, K, batch.nnz, dev)
    csc = CscWorkspace(V, batch.nnz, dev)
    csc.build(batch)
    dbeta = torch.zeros(K, V, device=dev)
    rm, rv = torch.zeros(V, device=dev), torch.ones(V, device=dev)
    colsum = theta.sum(0)

    def run():
        dec.forward_backward(theta, colsum, beta_bf, betaT_bf, batch, csc, training=True, running_mean=rm,
                             running_var=rv, dbeta=dbeta, loss=None)

    run()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(a.iters):
        run()
    e.record()
    torch.cuda.synchronize()
    ms = s.elapsed_time(e) / a.iters
    n_gemm = 6 if dec.folded else 7  # COL_STATS 1, ROW_LSE 1, COL_GRAD 2 (3 in the two-pass form), ROW_GRAD 2
    flops = 2.0 * B * V * dec.Kp * n_gemm
    print(f"decoder fwd+bwd B={B} V={V} K={K}: {ms:.3f} ms  ({flops / ms / 1e9:.1f} TFLOP/s tensor-core work)")
    if a.breakdown:
        import gfedntm_b200.ops.decoder as D
        from gfedntm_b200.ops import kernels as KK

        names = ["theta_center", "theta_center_fold", "dec_sweep", "scale_beta", "dec_sparse_doc", "csc_corr",
                 "csc_corr_t", "theta_gram", "gram", "gram_finish"]
        modes = {v: k for k, v in vars(KK).items() if k.startswith(("ROW_", "COL_")) and isinstance(v, int)}
        events = []

        class Timed:
            pass

        proxy = Timed()
        for n in dir(KK):
            setattr(proxy, n, getattr(KK, n))

        def wrap(name, fn):
            def inner(*args, **kw):
                s0, e0 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s0.record()
                r = fn(*args, **kw)
                e0.record()
                label = name if name != "dec_sweep" else modes.get(args[0], str(args[0]))
                events.append((label, s0, e0))
                return r
            return inner

        for n in names:
            setattr(proxy, n, wrap(n, getattr(KK, n)))
        D.K_ = proxy
        tot = {}
        for _ in range(a.iters):
            run()
        torch.cuda.synchronize()
        for label, s0, e0 in events:
            tot[label] = tot.get(label, 0.0) + s0.elapsed_time(e0) * 1e3 / a.iters
        D.K_ = KK
        print("  ".join(f"{k} {v:.0f}us" for k, v in tot.items()))


if __name__ == "__main__":
    main()
