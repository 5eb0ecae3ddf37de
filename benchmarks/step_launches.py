"""One configuration, a few eager (non-graph) training steps: the command profiled for the per-kernel launch
lists under profiles/ (ncu --metrics gpu__time_duration.sum).  Usage: step_launches.py CONFIG [STEPS]"""
import os
import sys

os.environ["GFEDNTM_NO_GRAPHS"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import bench  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else bench.DEFAULT_CONFIG
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    ctx = bench._Ctx()
    model, trainer, loader = bench.build_trainer(ctx, bench.CONFIGS[name], "grad")
    batches = [(mb.bow.__class__(mb.bow.indptr.clone(), mb.bow.indices.clone(), mb.bow.counts.clone(), mb.bow.B,
                                 mb.bow.V), mb.x_ctx.clone() if mb.x_ctx is not None else None) for mb in loader]
    torch.cuda.synchronize()
    for i in range(steps):
        trainer.step(*batches[i % len(batches)])
    torch.cuda.synchronize()
    print(f"{name}: {steps} eager steps done")


if __name__ == "__main__":
    main()
