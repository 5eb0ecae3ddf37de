"""Multi-GPU check of the fused NVLink aggregation, launched by torch.distributed.run (one rank per GPU).

Every rank trains on its own minibatches; after each global step the fused kernel's result is compared
with a plain NCCL all-reduce of the gradients followed by torch.optim.Adam on a private copy.
Exits non-zero on mismatch."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from gfedntm_b200.federated.trainer import FederatedTrainer
    from gfedntm_b200.models import AVITM, BOWDataset
    from gfedntm_b200.ops import kernels as Kx

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=device)
    agg = sys.argv[1] if len(sys.argv) > 1 else "grad"
    V, K, B = 3000, 20, 64
    rng = np.random.default_rng(100 + rank)
    n_docs = B * (3 + rank)  # different corpus sizes -> non-uniform client weights
    X = rng.poisson(0.02, (n_docs, V)).astype(np.float32)
    X[X.sum(1) == 0, 0] = 1
    ds = BOWDataset(X, None)
    model = AVITM(input_size=V, n_components=K, hidden_sizes=(32, 32), batch_size=B, num_epochs=1, verbose=False,
                  device=device, seed=3)
    tr = FederatedTrainer(model, n_samples=n_docs, world_size=world, rank=rank, agg=agg)
    loader = model.make_loader(ds, shuffle=True)
    tr.prepare(loader)
    eng = model.engine
    names = [n for n, _ in model.model.named_parameters()]
    ref = {n: eng.flat.view(eng.flat.P, n).detach().clone().contiguous().requires_grad_(True) for n in names}
    opt = torch.optim.Adam([ref[n] for n in names], lr=model.lr, betas=(model.momentum, 0.99))
    wsum = sum(tr.weights)
    assert abs(wsum - 1.0) < 1e-6 and len(set(round(w, 6) for w in tr.weights)) == world
    ok = True
    it = iter(loader)
    losses = []
    for step in range(6):
        try:
            mb = next(it)
        except StopIteration:
            it = iter(loader)
            mb = next(it)
        if agg == "grad":
            # replicate what trainer.step does, but keep the gradients for the oracle
            if tr._overlap is not None:
                # overlapped path: the beta part of the aggregation runs inside the d-theta sweep (the
                # gradient buffers are only read by it, so they can still be cloned for the oracle)
                Kx.step_begin(eng.ctr, 0, 1)
                loss = eng.forward_backward(mb.bow, None, training=True, gscale=tr.weights[rank], fed=tr._overlap)
                g = {n: eng.flat.view(eng.flat.G, n).detach().clone().contiguous() for n in names}
                s_local = eng.flat.S.detach().clone() * tr.weights[rank]
                eng.step_count += 1
                tr._launch("grad", eng.step_count, parts=1)
            else:
                loss = eng.forward_backward(mb.bow, None, training=True, gscale=tr.weights[rank])
                g = {n: eng.flat.view(eng.flat.G, n).detach().clone().contiguous() for n in names}
                s_local = eng.flat.S.detach().clone() * tr.weights[rank]
                eng.step_count += 1
                Kx.step_begin(eng.ctr, 0, 1)
                tr._launch("grad", eng.step_count)
            for n in names:
                dist.all_reduce(g[n])
                ref[n].grad = g[n]
            dist.all_reduce(s_local)
            opt.step()
            torch.cuda.synchronize()
            for n in names:
                got = eng.flat.view(eng.flat.P, n)
                if not torch.allclose(got, ref[n].detach(), rtol=3e-4, atol=3e-6):
                    print(f"[rank {rank}] step {step} {n}: max err {(got - ref[n].detach()).abs().max().item():.3e}")
                    ok = False
            if not torch.allclose(eng.flat.S, s_local, rtol=1e-5, atol=1e-6):
                print(f"[rank {rank}] step {step}: BatchNorm statistics differ")
                ok = False
        else:
            loss = tr.step(mb.bow, None)
            torch.cuda.synchronize()
        losses.append(float(loss.item()) / mb.size)
        # every rank must hold bit-identical parameters and shadows after the step
        mine = eng.flat.P.detach().clone()
        root = mine.clone()
        dist.broadcast(root, src=0)
        if not torch.equal(mine, root):
            print(f"[rank {rank}] step {step}: parameters differ from rank 0")
            ok = False
        sh = eng.beta_bf.detach().clone().contiguous().view(torch.int32)
        sh0 = sh.clone()
        dist.broadcast(sh0, src=0)
        if not torch.equal(sh, sh0):
            print(f"[rank {rank}] step {step}: bf16 shadows differ from rank 0")
            ok = False
        if not torch.equal(eng.beta_bf[:K, :V].float(), model.model.beta.detach().to(torch.bfloat16).float()):
            print(f"[rank {rank}] step {step}: shadow does not match beta")
            ok = False
    flag = torch.tensor([1.0 if ok else 0.0], device=device)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(f"overlap={'on' if tr._overlap is not None else 'off'} ", end="")
        print(f"agg={agg} world={world} losses/doc={['%.1f' % l for l in losses]} agg_us={tr.last_aggregation_us():.1f} "
              f"{'OK' if flag.item() == 1.0 else 'MISMATCH'}")
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 1.0 else 1)


if __name__ == "__main__":
    main()
