#!/usr/bin/env python
"""Headline benchmark: federated ProdLDA/CTM training throughput (documents / second).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config NAME] [--impl ours|reference]

One federated client per GPU (one process per GPU; for N > 1 launch through
``python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N``).  Every client runs
one minibatch step per global step on its own synthetic LDA corpus; with N > 1 the clients
aggregate every step (weighted by corpus size) through the fused NVLink collective.

Prints ONE JSON line on rank 0.  ``value`` is the device-timed (CUDA events, max over ranks)
aggregate docs/s with the minibatches already resident on the device; ``e2e`` is the same metric
through the public API (``CsrMinibatchLoader.cycle()`` + ``FederatedTrainer.step``), wall clock, max over
ranks: every step's minibatch is gathered from the host corpus into pinned memory and copied to the
device (``h2d_bytes_per_step``), and every step's loss reaches the host inside the timed region -- the
step's own CUDA graph copies it into a pinned slot (``engine.loss_slot``, 8 bytes) which the host reads
one step behind the GPU.  ``e2e.gpu_ms_per_step`` is the same loop timed with CUDA events.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONFIGS = {
    # BASELINE.json configs (hidden sizes / optimiser from config/dft_params.cf of the reference)
    "prodlda_10k": dict(family="avitm", model_type="prodLDA", V=10000, K=50, hidden=(50, 50), B=64,
                        desc="ProdLDA K=50 V=10k, batch 64/client"),
    "neurallda_20k": dict(family="avitm", model_type="LDA", V=20000, K=50, hidden=(50, 50), B=64,
                          desc="NeuralLDA K=50 V=20k, batch 64/client"),
    "ctm_50k": dict(family="ctm", model_type="prodLDA", V=50000, K=100, hidden=(50, 50), B=64, C=768,
                    inference_type="combined", desc="CombinedTM K=100 V=50k C=768, batch 64/client"),
    "zeroshot_50k": dict(family="ctm", model_type="prodLDA", V=50000, K=100, hidden=(50, 50), B=64, C=768,
                         inference_type="zeroshot", desc="ZeroShotTM K=100 V=50k C=768, batch 64/client"),
    "prodlda_200k": dict(family="avitm", model_type="prodLDA", V=200000, K=200, hidden=(50, 50), B=4096,
                         desc="large-vocab ProdLDA V=200k K=200, batch 4096/client"),
}
DEFAULT_CONFIG = "prodlda_200k"
L2_BYTES = 126 * 1024 * 1024


# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons every 50 ms while the timed regions run."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int) -> None:
        self.gpu_index = gpu_index
        self.proc = None
        self.lines = []

    def start(self) -> None:
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "50", "-i", str(self.gpu_index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self) -> None:
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                smax.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def make_corpus(cfg: dict, n_docs: int, seed: int, device: str):
    from gfedntm_b200.utils.synthetic import generate_corpus_gpu

    counts, _ = generate_corpus_gpu(n_docs, cfg["V"], cfg["K"], eta=0.01, seed=seed, device=device)
    ctx = None
    if cfg["family"] == "ctm":
        ctx = np.random.default_rng(seed).standard_normal((n_docs, cfg["C"])).astype(np.float32)
    return counts, ctx


def allreduce_max(x: float, world: int, device) -> float:
    if world == 1:
        return x
    import torch
    import torch.distributed as dist

    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


# ------------------------------------------------------------------------------------------------
class _Ctx:
    """Per-process state shared by the measurements of one bench run."""

    def __init__(self) -> None:
        import torch
        import torch.distributed as dist

        self.rank, self.world, self.local = dist_env()
        torch.cuda.set_device(self.local)
        self.device = torch.device("cuda", self.local)
        if self.world > 1 and not dist.is_initialized():
            dist.init_process_group("nccl", device_id=self.device)

    def barrier(self) -> None:
        import torch
        import torch.distributed as dist

        torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize()


def _trace(ctx, msg: str) -> None:
    """Progress lines on stderr (GFEDNTM_BENCH_TRACE=1): where a multi-rank run is when it stalls."""
    if os.environ.get("GFEDNTM_BENCH_TRACE", "0") == "1":
        print(f"[bench r{ctx.rank} {time.strftime('%H:%M:%S')}] {msg}", file=sys.stderr, flush=True)


def build_trainer(ctx: _Ctx, cfg: dict, agg: str):
    from gfedntm_b200.federated.trainer import FederatedTrainer
    from gfedntm_b200.models import AVITM, CTM, BOWDataset, CTMDataset

    B, V, K = cfg["B"], cfg["V"], cfg["K"]
    n_batches = 4
    counts, xctx = make_corpus(cfg, B * n_batches, seed=1234 + ctx.rank, device=f"cuda:{ctx.local}")
    if cfg["family"] == "ctm":
        ds = CTMDataset(xctx, counts, None)
        model = CTM(input_size=V, contextual_size=cfg["C"], inference_type=cfg["inference_type"],
                    n_components=K, model_type=cfg["model_type"], hidden_sizes=cfg["hidden"], batch_size=B,
                    num_epochs=1, verbose=False, device=ctx.device, seed=7)
    else:
        ds = BOWDataset(counts, None)
        model = AVITM(input_size=V, n_components=K, model_type=cfg["model_type"], hidden_sizes=cfg["hidden"],
                      batch_size=B, num_epochs=1, verbose=False, device=ctx.device, seed=7)
    assert model.backend == "fused", "bench must run the sm_100a kernel path"
    trainer = FederatedTrainer(model, n_samples=len(ds), world_size=ctx.world, rank=ctx.rank, agg=agg)
    loader = model.make_loader(ds, shuffle=True)
    trainer.prepare(loader)
    return model, trainer, loader


def measure_ours(ctx: _Ctx, cfg: dict, steps: int, warmup: int, agg: str, sampler=None, verify: bool = False) -> dict:
    """Device-timed and end-to-end throughput of one configuration on the ranks of ``ctx``."""
    import torch

    from gfedntm_b200.ops import _lib

    world, device = ctx.world, ctx.device
    B = cfg["B"]
    model, trainer, loader = build_trainer(ctx, cfg, agg)
    lib = _lib.load()
    eng = model.engine

    # ---- device-resident minibatches for the kernel-only number
    staged = []
    for mb in loader:
        staged.append((mb.bow.__class__(mb.bow.indptr.clone(), mb.bow.indices.clone(), mb.bow.counts.clone(),
                                        mb.bow.B, mb.bow.V),
                       mb.x_ctx.clone() if mb.x_ctx is not None else None))
    torch.cuda.synchronize()
    small_ws = trainer.working_set_bytes() < 2 * L2_BYTES
    flush = torch.empty(2 * L2_BYTES // 4, dtype=torch.float32, device=device) if small_ws else None

    # untimed: the engine captures one CUDA graph per input buffer set (eager warm-up + capture call),
    # so every staged minibatch is visited twice before the W warm-up steps proper
    _trace(ctx, "trainer built, warm-up")
    for i in range(2 * len(staged) + warmup):
        bow, xc = staged[i % len(staged)]
        trainer.step(bow, xc)
        if i < 3:
            torch.cuda.synchronize()
            _trace(ctx, f"warm-up step {i} done")
    ctx.barrier()
    _trace(ctx, "timed region")
    if sampler is not None and ctx.rank == 0:
        sampler.start()
    launches0 = lib.gf_launch_count() + eng.replayed_launches
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    if flush is None:
        starts[0].record()
    for i in range(steps):
        bow, xc = staged[i % len(staged)]
        if flush is not None:
            flush.add_(1.0)  # evict L2 between timed steps (write > L2 bytes)
            starts[i].record()
        trainer.step(bow, xc)
        if flush is not None:
            ends[i].record()
    if flush is None:
        ends[-1].record()
    ctx.barrier()
    launches = lib.gf_launch_count() + eng.replayed_launches - launches0
    if flush is None:
        ms = starts[0].elapsed_time(ends[-1])
    else:
        ms = sum(s.elapsed_time(e) for s, e in zip(starts, ends))
    ms = allreduce_max(ms, world, device)
    _trace(ctx, f"device-timed {ms / steps:.3f} ms/step; end-to-end loop")

    # ---- end to end through the public API: host assembly + pinned H2D + step + loss D2H.  The loss of
    # step i is copied to pinned host memory right behind the step and read by the host while step i+1
    # runs (two slots): every step's loss reaches the host inside the timed region, without draining the
    # GPU once per step.
    loss_host = [None, None]
    loss_ev = [torch.cuda.Event() for _ in range(2)]
    it = loader.cycle()   # endless: the staging pipeline keeps running across the epochs of the small bench corpus
    h2d = 0

    def next_mb():
        return next(it)

    for _ in range(max(warmup, 9)):  # three staging slots x (eager warm-up + graph capture) + slack
        mb = next_mb()
        trainer.step(mb.bow, mb.x_ctx)
    ctx.barrier()
    loss_sum = 0.0
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    t0 = time.perf_counter()
    for i in range(steps):
        mb = next_mb()
        trainer.step(mb.bow, mb.x_ctx)
        # the step's own graph copies its loss into a pinned host slot (one slot per staging-buffer set, so the next
        # step writes another one): the host reads step i-1's loss once that step's event has completed
        loss_host[i & 1] = eng.loss_slot
        loss_ev[i & 1].record()
        if i > 0:
            loss_ev[(i - 1) & 1].synchronize()
            loss_sum += float(loss_host[(i - 1) & 1][0])
        h2d = mb.h2d_bytes
    ev1.record()
    loss_ev[(steps - 1) & 1].synchronize()
    loss_sum += float(loss_host[(steps - 1) & 1][0])
    ev1.synchronize()
    e2e_dev_ms = ev0.elapsed_time(ev1) / steps   # the same loop seen from the GPU (first launch to last kernel)
    ctx.barrier()
    e2e_s = allreduce_max(time.perf_counter() - t0, world, device)
    assert loss_sum == loss_sum, "NaN loss in the end-to-end loop"
    it.close()  # stops the loader's staging thread and releases its slots (verification iterates again)

    docs = world * B * steps
    out = {"value": docs / (ms / 1e3), "ms_per_step": ms / steps, "steps": steps,
           "e2e": {"value": docs / e2e_s, "unit": "docs/s", "h2d_bytes_per_step": int(h2d),
                   "d2h_bytes_per_step": 8, "ms_per_step": 1e3 * e2e_s / steps,
                   "loss_per_doc": loss_sum / (steps * B), "gpu_ms_per_step": e2e_dev_ms},
           "gpu_launches": int(launches),
           "l2": "flush 252 MB between timed steps" if flush is not None else
                 "working set (params+shadows+grads) larger than L2",
           "global_batch": world * B}
    if world > 1:
        out["aggregation"] = trainer.describe_aggregation()
        _trace(ctx, "verification")
        if verify:
            from gfedntm_b200.federated.verify import verify_against_oracle

            out["verify"] = verify_against_oracle(trainer, loader, steps=2)
        # exposed cost of the aggregation = federated step - the same step trained locally (N=1 path:
        # local fused Adam instead of the NVLink aggregation).  Runs last: it lets the ranks diverge.
        ctx.barrier()
        _trace(ctx, "local steps")
        for i in range(2 * len(staged) + 3):
            eng.train_step(*staged[i % len(staged)])
        ctx.barrier()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n_loc = max(3, min(steps, 10))
        loc_ms = 0.0
        for i in range(n_loc):
            if flush is not None:
                flush.add_(1.0)
            s.record()
            eng.train_step(*staged[i % len(staged)])
            e.record()
            e.synchronize()
            loc_ms += s.elapsed_time(e)
        loc_ms = allreduce_max(loc_ms / n_loc, world, device)
        out["local_ms_per_step"] = loc_ms
        out["aggregation_us_per_step"] = 1e3 * (ms / steps - loc_ms)
    del trainer, model, loader, staged, eng
    import gc

    gc.collect()
    torch.cuda.empty_cache()
    return out


def run_ours(args, cfg: dict) -> dict:
    import torch.distributed as dist

    ctx = _Ctx()
    world = ctx.world
    sampler = ClockSampler(ctx.local)
    main = measure_ours(ctx, cfg, args.steps, args.warmup, args.agg, sampler=sampler, verify=True)
    clocks = sampler.stop() if ctx.rank == 0 else None
    B, V, K = cfg["B"], cfg["V"], cfg["K"]
    out = {
        "metric": "ProdLDA/CTM docs/sec (federated, one client per GPU)", "value": main["value"], "unit": "docs/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": main["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16 operands, fp32 accumulate/master",
        "data": "synthetic LDA corpus (torch-sampled), random-init weights",
        "config": {"model": cfg["desc"], "name": args.config, "global_batch": world * B, "seq_len": None,
                   "vocab": V, "topics": K, "hidden": list(cfg["hidden"]),
                   "parallelism": f"fed{world}" + (f"-{args.agg}" if world > 1 else ""),
                   "aggregation": main.get("aggregation"), "l2": main["l2"]},
        "e2e": main["e2e"], "gpu_launches": main["gpu_launches"],
        "aggregation_us_per_step": main.get("aggregation_us_per_step"),
        "local_ms_per_step": main.get("local_ms_per_step"), "clocks": clocks, "impl": "ours",
    }
    if "verify" in main:
        out.update(main["verify"])
    # the other BASELINE.json configurations, a few hundred short steps each (same protocol, same ranks)
    if args.all_configs:
        extra = {}
        for name, c in CONFIGS.items():
            if name == args.config:
                continue
            try:
                r = measure_ours(ctx, c, steps=max(args.steps, 50), warmup=args.warmup, agg=args.agg)
                extra[name] = {k: r[k] for k in ("value", "ms_per_step", "steps", "e2e", "gpu_launches", "l2",
                                                 "global_batch", "aggregation_us_per_step", "local_ms_per_step")
                               if k in r}
                extra[name]["model"] = c["desc"]
            except Exception as ex:  # noqa: BLE001 -- the headline number must survive a side config
                extra[name] = {"error": f"{type(ex).__name__}: {ex}".splitlines()[0][:200]}
        out["configs"] = extra
    if world > 1:
        dist.destroy_process_group()
    return out


# ------------------------------------------------------------------------------------------------
def run_reference(args, cfg: dict) -> dict:
    from baseline.ref_arm import run_reference_arm

    return run_reference_arm(args, cfg, ClockSampler, dist_env, make_corpus, allreduce_max)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--config", type=str, default=DEFAULT_CONFIG, choices=sorted(CONFIGS))
    ap.add_argument("--impl", type=str, default="ours", choices=["ours", "reference"])
    ap.add_argument("--agg", type=str, default="grad", choices=["grad", "param"],
                    help="federated aggregation: weighted gradients + global Adam, or weighted parameters")
    ap.add_argument("--all-configs", dest="all_configs", type=int, default=None,
                    help="1/0: also measure the other BASELINE.json configurations into `configs` "
                         "(default: on for the default configuration)")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    cfg = CONFIGS[args.config]
    if args.all_configs is None:
        args.all_configs = int(args.config == DEFAULT_CONFIG and args.impl == "ours")
    rank, world, _ = dist_env()
    if world != args.gpus and args.gpus > 1 and world == 1:
        # convenience: spawn the ranks ourselves when not started through torch.distributed.run
        port = 29500 + (os.getpid() % 2000)
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    try:
        out = run_reference(args, cfg) if args.impl == "reference" else run_ours(args, cfg)
    except Exception as ex:
        if args.impl != "reference":
            raise
        # the driver's contract for this arm: one JSON line and exit 0 even when it cannot run
        out = {"impl": "reference", "unavailable": f"{type(ex).__name__}: {ex}".splitlines()[0][:300]}
    if rank == 0:
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
