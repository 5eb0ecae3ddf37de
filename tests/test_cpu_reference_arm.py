"""The reference arm of bench.py must be importable after ``__graft_entry__.build()``."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _have_reference():
    return os.path.isdir("/root/reference/src") or os.path.isdir(os.path.join(ROOT, "baseline", "_ref", "src"))


@pytest.mark.skipif(not _have_reference(), reason="no reference checkout and no installed copy")
def test_reference_arm_imports_after_ensure_installed():
    from baseline import ref_arm

    path = ref_arm.ensure_installed()
    assert os.path.isfile(os.path.join(path, "src", "models", "federated", "federated_avitm.py"))
    classes = ref_arm._import_reference()
    assert len(classes) == 6 and classes[0].__name__ == "AVITM" and classes[4].__name__ == "FederatedAVITM"
    # the classes come from the installed copy, not from this repository's package
    assert classes[0].__module__.startswith("src.models.base.pytorchavitm")


@pytest.mark.skipif(not _have_reference(), reason="no reference checkout and no installed copy")
def test_reference_copy_is_unmodified():
    """Every file under baseline/_ref/src is byte-identical to the reference checkout (when it is mounted)."""
    if not os.path.isdir("/root/reference/src"):
        pytest.skip("reference tree not mounted")
    from baseline import ref_arm

    dst = os.path.join(ref_arm.ensure_installed(), "src")
    n = 0
    for dirpath, _, files in os.walk(dst):
        for f in files:
            if f.endswith(".pyc"):
                continue
            p = os.path.join(dirpath, f)
            q = os.path.join("/root/reference/src", os.path.relpath(p, dst))
            with open(p, "rb") as a, open(q, "rb") as b:
                assert a.read() == b.read(), p
            n += 1
    assert n > 30


def test_bench_reference_cli_reports_unavailable_or_runs_without_gpu():
    """Without a GPU the arm must still exit 0 with one JSON line (the driver's contract)."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("CPU-only contract test")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "3", "--config", "prodlda_10k"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    import json

    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference"
