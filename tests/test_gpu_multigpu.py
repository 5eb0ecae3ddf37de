"""Real multi-GPU runs of the fused aggregation (skipped with fewer than two GPUs)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


SUBSET = "prior_mean,prior_variance,beta"  # the reference server's default grads_to_share (server.py:71)


@pytest.mark.parametrize("agg,folded,share", [("grad", False, ""), ("param", False, ""), ("grad", True, ""),
                                              ("grad", False, SUBSET), ("param", False, SUBSET),
                                              ("grad", True, SUBSET)])
def test_fused_aggregation_across_gpus(agg, folded, share):
    """``folded=True`` forces the folded decoder, which turns on the overlapped beta aggregation (spare CTAs of
    the d-theta sweep) in grad mode; the check against the NCCL + torch.Adam oracle is the same."""
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    world = 2 if n < 4 else (4 if n < 8 else 8)
    port = 29600 + (os.getpid() % 300) + (0 if agg == "grad" else 1) + (2 if folded else 0) + (4 if share else 0)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "mp_fed_check.py"), agg,
           share]
    env = dict(os.environ, GFEDNTM_DECODER_V1="0" if folded else "1")
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "OK" in res.stdout


@pytest.mark.parametrize("agg,kind", [("grad", "combined"), ("param", "combined"), ("grad", "zeroshot")])
def test_fused_aggregation_ctm_across_gpus(agg, kind):
    """CombinedTM in gradient mode sums the gradient of the collapsed weight (150 KB) instead of the adapt_bert
    gradient and applies the identical Adam step on every rank; the result must equal the oracle that all-reduces
    every gradient."""
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    world = 2 if n < 4 else (4 if n < 8 else 8)
    port = 29700 + (os.getpid() % 200) + (0 if agg == "grad" else 1) + (2 if kind == "zeroshot" else 0)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "mp_fed_check.py"), agg,
           "", kind]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "OK" in res.stdout
