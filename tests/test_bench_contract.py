"""CPU: the parts of bench.py's contract that need no GPU -- the reference arm (`--impl reference`: the unmodified reference,
or for the P2 extension the oracle port, timed on the host cores) prints ONE JSON line with the agreed keys, alone on rank 0
under torchrun, and the product arm refuses to run without a CUDA device instead of falling back to anything."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HAVE_REF = os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libyfd_ref.so"))


def run(args, launcher=()):
    env = dict(os.environ, PYTHONPATH=ROOT)
    p = subprocess.run([sys.executable, *launcher, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True, cwd=ROOT,
                       env=env, timeout=600)
    return p


def check_reference_line(line, workload, n_gpus):
    d = json.loads(line)
    assert d["impl"] == "reference" and d["config"]["workload"] == workload and d["n_gpus"] == n_gpus
    for k in ("metric", "value", "unit", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype", "data"):
        assert k in d, k
    assert d["value"] > 0 and d["higher_is_better"] is True and d["vs_baseline"] is None and d["dtype"] == "f64"
    cb, e2e = d["cpu_baseline"], d["e2e"]
    assert cb["value"] == d["value"] and cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["sample"]
    assert e2e["value"] == d["value"] and e2e["unit"] == d["unit"] and e2e["h2d_bytes_per_step"] == 0 and e2e["d2h_bytes_per_step"] == 0
    return d


@pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref not built")
def test_reference_arm_single_process():
    p = run(["--impl", "reference", "--workload", "dg_sip_tri_k1_64x64", "--steps", "1", "--warmup", "0"])
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = check_reference_line(lines[0], "dg_sip_tri_k1_64x64", 1)
    assert d["cpu_baseline"]["kind"] == "reference" and d["metric"] == "assembled+solved DOFs/sec"


def test_reference_arm_of_the_p2_extension_is_the_oracle_port():
    p = run(["--impl", "reference", "--workload", "cfem_p2_4m", "--steps", "1", "--warmup", "0"])
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = check_reference_line(lines[0], "cfem_p2_4m", 1)
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] == 1


@pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref not built")
def test_reference_arm_under_torchrun_prints_once():
    p = run(["--impl", "reference", "--workload", "dg_sip_tri_k1_64x64", "--gpus", "2", "--steps", "1", "--warmup", "0"],
            launcher=("-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1", "--master-port", "29541"))
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    check_reference_line(lines[0], "dg_sip_tri_k1_64x64", 2)


def test_product_arm_needs_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    p = run(["--workload", "dg_sip_tri_k1_64x64", "--steps", "1", "--warmup", "0"])
    assert p.returncode != 0
    assert not [l for l in p.stdout.splitlines() if l.startswith("{")]       # no number without a device
