"""Vocabulary consensus inside one ``torch.distributed`` job (stage 1 of the federation when the clients
are ranks of one box).  Ref: server.py:270-279 — the global vocabulary is the sorted union of the client
vocabularies; client-side indices are ignored."""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple


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
    gathered = [None] * world
    dist.all_gather_object(gathered, (sorted(set(local_tokens)), int(n_local_docs)), group=group)
    vocab = merge_vocabularies([g[0] for g in gathered])
    tot = float(sum(g[1] for g in gathered))
    return vocab, [g[1] / tot for g in gathered]
