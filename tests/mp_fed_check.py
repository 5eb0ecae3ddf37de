"""Multi-GPU check of the fused NVLink aggregation, launched by torch.distributed.run (one rank per GPU).

Every rank trains on its own minibatches (different corpus sizes, so non-uniform client weights); the fused
result is compared with the NCCL + torch oracle of ``gfedntm_b200.federated.verify`` and every rank must end
up with bit-identical replicated state.  ``argv[1]``: grad | param; ``argv[2]`` (optional): comma-separated
``grads_to_share`` subset.  Exits non-zero on mismatch."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from gfedntm_b200.federated.trainer import FederatedTrainer
    from gfedntm_b200.federated.verify import ranks_identical, verify_against_oracle
    from gfedntm_b200.models import AVITM, BOWDataset

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=device)
    # stage 1 on the box: vocabulary consensus through the symmetric heap (peer-mapped reads, host merge)
    from gfedntm_b200.parallel import consensus

    toks = [f"tok{j}" for j in range(rank, 40 + 3 * rank)] + ["ünïcode", f"only{rank}"]
    vocab, cw = consensus.consensus_vocabulary(toks, 10 * (rank + 1))
    want = sorted(set(t for r in range(world) for t in [f"tok{j}" for j in range(r, 40 + 3 * r)] + ["ünïcode", f"only{r}"]))
    assert list(vocab) == want and [vocab[t] for t in want] == list(range(len(want))), "vocabulary consensus mismatch"
    assert abs(sum(cw) - 1.0) < 1e-9 and abs(cw[rank] - 10 * (rank + 1) / sum(10 * (r + 1) for r in range(world))) < 1e-9
    assert consensus.LAST_TRANSPORT == "symmetric_heap", consensus.LAST_TRANSPORT
    agg = sys.argv[1] if len(sys.argv) > 1 else "grad"
    share = sys.argv[2].split(",") if len(sys.argv) > 2 and sys.argv[2] else None
    kind = sys.argv[3] if len(sys.argv) > 3 else "avitm"   # avitm | combined | zeroshot
    V, K, B = 3000, 20, 64
    rng = np.random.default_rng(100 + rank)
    n_docs = B * (3 + rank)  # different corpus sizes -> non-uniform client weights
    X = rng.poisson(0.02, (n_docs, V)).astype(np.float32)
    X[X.sum(1) == 0, 0] = 1
    if kind == "avitm":
        ds = BOWDataset(X, None)
        model = AVITM(input_size=V, n_components=K, hidden_sizes=(32, 32), batch_size=B, num_epochs=1, verbose=False,
                      device=device, seed=3)
    else:
        from gfedntm_b200.models import CTM, CTMDataset

        ctx = rng.standard_normal((n_docs, 96)).astype(np.float32)
        ds = CTMDataset(ctx, X, None)
        model = CTM(input_size=V, contextual_size=96, inference_type=kind, n_components=K, hidden_sizes=(32, 32),
                    batch_size=B, num_epochs=1, verbose=False, device=device, seed=3)
    tr = FederatedTrainer(model, n_samples=n_docs, world_size=world, rank=rank, agg=agg, grads_to_share=share)
    loader = model.make_loader(ds, shuffle=True)
    tr.prepare(loader)
    eng = model.engine
    assert abs(sum(tr.weights) - 1.0) < 1e-6 and len(set(round(w, 6) for w in tr.weights)) == world
    ok = True
    losses = []
    it = iter(loader)
    for step in range(5):  # eager warm-up, graph capture, replays
        try:
            mb = next(it)
        except StopIteration:
            it = iter(loader)
            mb = next(it)
        loss = tr.step(mb.bow, mb.x_ctx)
        losses.append(float(loss.item()) / mb.size)
    torch.cuda.synchronize()
    it.close()
    if share is None:
        res = verify_against_oracle(tr, loader, steps=3)
        if not res["ranks_identical"]:
            print(f"[rank {rank}] replicated state differs between ranks")
            ok = False
        if not res["oracle_max_rel_err"] < 5e-4:
            print(f"[rank {rank}] oracle error {res['oracle_max_rel_err']:.3e}")
            ok = False
        tr.materialize()
        if eng.is_prod and not torch.equal(eng.beta_bf[:K, :V].float(),
                                           model.model.beta.detach().to(torch.bfloat16).float()):
            print(f"[rank {rank}] shadow does not match beta")
            ok = False
    else:
        # subset sharing: shared tensors identical on every rank, unshared ones must have diverged
        tr.materialize()
        sd = {k: v.detach().clone().contiguous() for k, v in model.model.state_dict().items()
              if v.dtype.is_floating_point}
        res = {"oracle_max_rel_err": float("nan")}
        for k, v in sd.items():
            ref = v.clone()
            dist.broadcast(ref, src=0)
            same = bool(torch.equal(ref, v))
            flag = torch.tensor([1.0 if same else 0.0], device=device)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            same_everywhere = flag.item() == 1.0
            if k in share and not same_everywhere:
                print(f"[rank {rank}] shared tensor {k} differs between ranks")
                ok = False
            if k not in share and same_everywhere and v.numel() > 1 and "prior" not in k:
                print(f"[rank {rank}] unshared tensor {k} is identical on every rank (was it aggregated?)")
                ok = False
    if not all(np.isfinite(losses)):
        ok = False
    flag = torch.tensor([1.0 if ok else 0.0], device=device)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(f"overlap={'on' if tr._overlap is not None else 'off'} agg={agg} share={'all' if share is None else share} "
              f"world={world} losses/doc={['%.1f' % l for l in losses]} "
              f"oracle_err={res['oracle_max_rel_err']:.2e} {'OK' if flag.item() == 1.0 else 'MISMATCH'}")
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 1.0 else 1)


if __name__ == "__main__":
    main()
