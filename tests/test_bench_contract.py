"""The driver-facing contract of bench.py that can be checked without a GPU: the reference arm's JSON line (keys,
types, one line from rank 0 under torchrun) and that the product arm fails loudly -- no CPU fallback -- when there is
no CUDA device."""
import json
import os
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
SMALL = ["--impl", "reference", "--images-per-step", "2", "--steps", "1", "--warmup", "0"]


def _json_lines(out: str):
    lines = []
    for ln in out.splitlines():
        ln = ln.strip()
        if ln.startswith("{") and ln.endswith("}"):
            try:
                lines.append(json.loads(ln))
            except ValueError:
                pass
    return lines


def _check_reference_line(d, n_gpus):
    assert d["impl"] == "reference"
    assert d["unit"] == "Msymbols/s" and "Msymbols/s" in d["metric"]
    assert isinstance(d["value"], float) and d["value"] > 0
    assert d["n_gpus"] == n_gpus and d["steps"] == 1 and d["warmup"] == 0
    assert d["ms_per_step"] > 0 and d["higher_is_better"] is True and d["scaling"] in ("weak", "strong")
    assert d["vs_baseline"] is None and d["data"] == "synthetic" and isinstance(d["dtype"], str)
    assert "workload" in d["config"] and "model" not in d["config"]
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["value"] == d["value"]
    assert cb["unit"] == d["unit"] and isinstance(cb["sample"], str) and cb["sample"]
    e = d["e2e"]
    assert e["value"] == d["value"] and e["unit"] == d["unit"]
    assert e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0


def test_reference_arm_prints_one_contract_line():
    r = subprocess.run([sys.executable, str(ROOT / "bench.py"), *SMALL], cwd=str(ROOT), capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = _json_lines(r.stdout)
    assert len(lines) == 1, r.stdout[-2000:]
    _check_reference_line(lines[0], 1)


def test_reference_arm_under_torchrun_only_rank0_prints():
    env = dict(os.environ, OMP_NUM_THREADS="1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29631", str(ROOT / "bench.py"), *SMALL,
                        "--gpus", "2"], cwd=str(ROOT), env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = _json_lines(r.stdout)
    assert len(lines) == 1, r.stdout[-2000:]
    _check_reference_line(lines[0], 2)


def test_product_arm_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    r = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--steps", "1", "--warmup", "3"], cwd=str(ROOT),
                       capture_output=True, text=True, timeout=600)
    assert r.returncode != 0
    assert not _json_lines(r.stdout), "no number may be printed without a GPU"
