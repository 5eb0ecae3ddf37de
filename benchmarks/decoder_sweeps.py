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
    flops = 2.0 * B * V * dec.Kp * 7  # COL_STATS 1, ROW_LSE 1, COL_GRAD 3, ROW_GRAD 2 tile GEMMs
    print(f"decoder fwd+bwd B={B} V={V} K={K}: {ms:.3f} ms  ({flops / ms / 1e9:.1f} TFLOP/s tensor-core work)")


if __name__ == "__main__":
    main()
