import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def yfd():
    import yaourt_fem_dg_b200 as y
    return y


@pytest.fixture(scope="session")
def ref():
    """The UNMODIFIED reference compiled as oracle/_ref/libyfd_ref.so (travels to the GPU box)."""
    import refbind
    if not refbind.available():
        pytest.skip("oracle/_ref/libyfd_ref.so not built (needs /root/reference at build time)")
    return refbind.Ref()


@pytest.fixture(scope="session")
def ctx(yfd):
    c = yfd.Context(0)
    yield c
    c.close()


def host_mesh_from_ref(yfd, m):
    return yfd.HostMesh(m["type"], m["points"], m["cells"], m["faces"], m["face_boundary"].astype(np.uint8))


def block_max_rel_err(a, b, rowptr, bs):
    """max over row blocks of |a-b|_inf / |b|_inf(block row): value parity relative to the block
    max-norm (explicit zeros are stored, so per-entry relative error is undefined)."""
    n = len(rowptr) - 1
    worst = 0.0
    for c in range(0, n, bs):
        s, e = rowptr[c], rowptr[min(c + bs, n)]
        scale = np.abs(b[s:e]).max()
        if scale == 0:
            scale = 1.0
        worst = max(worst, np.abs(a[s:e] - b[s:e]).max() / scale)
    return worst


GOLDEN = os.path.join(ROOT, "tests", "golden")


def golden(name):
    """A fixture file written by tools/make_golden.py from the UNMODIFIED reference."""
    return np.load(os.path.join(GOLDEN, name))


def golden_digests():
    out = {}
    with open(os.path.join(GOLDEN, "mesh_digests.txt")) as f:
        for line in f:
            k, v = line.rstrip("\n").split(" ", 1)
            out[k] = v
    return out


@pytest.fixture(scope="session")
def oracle():
    """oracle/libyfd_oracle.so, the plain-C restatement (built by __graft_entry__.build())."""
    import oraclebind
    if not oraclebind.available():
        import subprocess
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"])
    return oraclebind.Oracle()
