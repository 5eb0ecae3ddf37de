"""Multi-process host-side logic on CPU (gloo, world_size 2): vocabulary consensus and client weights."""
import os
import socket

import torch.multiprocessing as mp


def _worker(rank, world, port, out):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from gfedntm_b200.parallel.consensus import consensus_vocabulary

    tokens = [["wd1", "wd5", "apple"], ["wd5", "zebra", "wd2", "wd2"]][rank]
    vocab, weights = consensus_vocabulary(tokens, [30, 90][rank])
    out.put((rank, vocab, weights))
    dist.barrier()
    dist.destroy_process_group()


def test_vocabulary_consensus_gloo():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in range(2)])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (_, v0, w0), (_, v1, w1) = res
    assert v0 == v1 == {"apple": 0, "wd1": 1, "wd2": 2, "wd5": 3, "zebra": 4}
    assert w0 == w1 == [0.25, 0.75]
