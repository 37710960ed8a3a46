"""CPU: the measurement contract of bench.py that can be checked without a GPU - the reference arm
(`--impl reference`: the reference's own CPU code from oracle/_ref on the host cores) prints ONE JSON line with the
agreed keys, and the product arm refuses to run without a B200 instead of falling back to anything."""
import json
import os
import subprocess
import sys

import pytest

from oracle import ref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built")
def test_reference_arm_prints_the_contract_line():
    # --ref-n: a small sample of the C3 workload so the test takes seconds (the default picks ~10-30 s of CPU work)
    p = subprocess.run([sys.executable, "bench.py", "--impl", "reference", "--steps", "1", "--warmup", "1", "--ref-n", "1536"],
                       cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "gp_learn_fp64_tflops" and d["unit"] == "TFLOP/s"
    assert d["higher_is_better"] is True and d["dtype"] == "f64" and d["vs_baseline"] is None
    assert d["cpu_baseline"]["kind"] == "reference" and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert d["e2e"]["value"] == d["value"] > 0
    assert "workload" in d["config"]


def test_product_arm_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    p = subprocess.run([sys.executable, "bench.py", "--steps", "1", "--warmup", "1"], cwd=ROOT, capture_output=True,
                       text=True, timeout=600)
    assert p.returncode != 0
    assert not any(l.startswith("{") and '"value"' in l for l in p.stdout.splitlines())
