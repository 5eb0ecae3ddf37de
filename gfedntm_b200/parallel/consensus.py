"""Vocabulary consensus inside one ``torch.distributed`` job (stage 1 of the federation when the clients
are ranks of one box).  Ref: server.py:270-279 — the global vocabulary is the sorted union of the client
vocabularies; client-side indices are ignored."""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple


LAST_TRANSPORT = None  # "symmetric_heap" or "all_gather_object": what the last consensus_vocabulary call used


def merge_vocabularies(vocabs: Sequence[Sequence[str]]) -> Dict[str, int]:
    """``sorted(set union)`` -> ``{token: global index}``."""
    merged = sorted(set().union(*[set(v) for v in vocabs]))
    return {t: i for i, t in enumerate(merged)}


def consensus_vocabulary(local_tokens: Sequence[str], n_local_docs: int, group=None
                         ) -> Tuple[Dict[str, int], List[float]]:
    """All-gathers every rank's token list and corpus size; returns the global vocabulary and the client
    weights ``n_c / sum n`` (identical on every rank)."""
    import torch.distributed as dist

    world = dist.get_world_size(group)
    global LAST_TRANSPORT
    mine = sorted(set(local_tokens))
    gathered = None
    if dist.get_backend(group) == "nccl":
        try:
            gathered = _allgather_through_symmetric_heap(mine, int(n_local_docs), group)
        except Exception:  # no peer access / symmetric memory unavailable: plain object all-gather below
            gathered = None
    LAST_TRANSPORT = "symmetric_heap" if gathered is not None else "all_gather_object"
    if gathered is None:
        gathered = [None] * world
        dist.all_gather_object(gathered, (mine, int(n_local_docs)), group=group)
    vocab = merge_vocabularies([g[0] for g in gathered])
    tot = float(sum(g[1] for g in gathered))
    return vocab, [g[1] / tot for g in gathered]


def _allgather_through_symmetric_heap(tokens: List[str], n_docs: int, group=None):
    """On-box all-gather of variable-length UTF-8 token blobs through NVLink peer memory: every rank
    writes ``[n_docs, n_bytes, blob]`` into its slot of a symmetric allocation, and after a barrier reads its peers'
    slots directly (peer-mapped loads, no NCCL data movement); the merge itself stays on the host."""
    import numpy as np
    import torch
    import torch.distributed as dist
    import torch.distributed._symmetric_memory as symm_mem

    group = group if group is not None else dist.group.WORLD
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dev = torch.device("cuda", torch.cuda.current_device())
    blob = "\n".join(tokens).encode("utf-8")
    n = torch.tensor([len(blob)], dtype=torch.int64, device=dev)
    dist.all_reduce(n, op=dist.ReduceOp.MAX, group=group)          # slot size (plumbing: one scalar)
    cap = (int(n.item()) + 16 + 15) // 16 * 16
    buf = symm_mem.empty(cap, dtype=torch.uint8, device=dev)
    hdl = symm_mem.rendezvous(buf, group)
    head = np.array([n_docs, len(blob)], dtype=np.int64).tobytes()
    payload = torch.frombuffer(bytearray(head + blob), dtype=torch.uint8)
    buf[: payload.numel()].copy_(payload.to(dev))
    torch.cuda.synchronize()
    hdl.barrier()
    out = []
    for r in range(world):
        peer = hdl.get_buffer(r, (cap,), torch.uint8)              # peer-mapped view of rank r's slot
        raw = peer.cpu().numpy().tobytes()
        nd, nb = np.frombuffer(raw[:16], dtype=np.int64)
        text = raw[16:16 + int(nb)].decode("utf-8")
        out.append((text.split("\n") if text else [], int(nd)))
    torch.cuda.synchronize()
    hdl.barrier()                                                 # nobody frees its slot while a peer still reads it
    return out
