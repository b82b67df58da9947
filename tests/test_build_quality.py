"""CPU (no GPU needed): what the sm_100a build of the hot kernels looks like -- read from the shared library itself with
cuobjdump, so a change that starts spilling registers or loses the TMA path in a benchmarked instantiation fails here,
before any GPU time is spent (B200_PROFILING.md: check -res-usage and the SASS first)."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "yaourt-fem-dg_b200", "lib", "libyfd_b200.so")
CUOBJDUMP = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"

pytestmark = pytest.mark.skipif(not (os.path.exists(LIB) and os.path.exists(CUOBJDUMP)), reason="needs the built library and cuobjdump")


@pytest.fixture(scope="module")
def usage():
    out = subprocess.run([CUOBJDUMP, "-res-usage", LIB], capture_output=True, text=True, check=True).stdout
    res = {}
    for m in re.finditer(r"Function (\S+):\s*\n\s*REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", out):
        res[m.group(1)] = dict(reg=int(m.group(2)), stack=int(m.group(3)), shared=int(m.group(4)), local=int(m.group(5)))
    assert res, "cuobjdump -res-usage printed nothing"
    return res


def find(usage, *parts):
    hits = {k: v for k, v in usage.items() if all(p in k for p in parts)}
    assert hits, parts
    return hits


def test_arch_is_sm_100a():
    out = subprocess.run([CUOBJDUMP, "-lelf", LIB], capture_output=True, text=True, check=True).stdout
    assert "sm_100a" in out and "sm_90" not in out and "sm_80" not in out


def test_hot_kernels_do_not_spill(usage):
    # the instantiations bench.py measures: DG block SpMV bs = 10 (quad k=3) and 6 / 15 (tri k=2 / k=4), the CSR SpMV of the
    # continuous-FEM systems, the Maxwell stage kernel tri k=4 / quad k=3, and the CG vector kernels
    for parts in (("dg_spmv_tma", "ILi10E"), ("dg_spmv_tma", "ILi6E"), ("dg_spmv_tma", "ILi15E"), ("csr_spmv_tma",),
                  ("13maxwell_stageI", "ILi3ELi4E"), ("13maxwell_stageI", "ILi4ELi3E"), ("cg_update_xr",), ("cg_update_d",),
                  ("13dg_spmv_classI", "ILi10E", "ELb0EEE"), ("13dg_spmv_classI", "ILi6E", "ELb0EEE"), ("pmg_sweep", "ILi10E", "ELb0EEE"), ("pmg_sweep", "ILi6E", "ELb0EEE"), ("pmg_sweep", "ILi3E", "ELb0EEE")):
        for name, u in find(usage, *parts).items():
            assert u["stack"] == 0 and u["local"] == 0, (name, u)
    # two CTAs per SM for the Maxwell stage kernel at the benchmarked degrees need <= 128 registers (256 threads)
    for name, u in find(usage, "13maxwell_stageI", "ILi3ELi4E").items():
        assert u["reg"] <= 128, (name, u)
    # the class kernels run 2 CTAs of 256 threads per SM: <= 128 registers, at most a few spilled words at bs = 15
    for parts in (("13dg_spmv_classI",), ("pmg_sweep",)):
        for name, u in find(usage, *parts).items():
            assert u["reg"] <= 128 and u["stack"] <= 64, (name, u)
    # the tensor-pipe class kernel: 2 CTAs of 256 threads (16 warps per SM hide the operand latency), a few spilled words
    for name, u in find(usage, "17dg_spmv_class_mmaI").items():
        assert u["reg"] <= 128 and u["stack"] <= 96, (name, u)
    # the Maxwell class kernel holds two cells per lane (one CTA of 256 threads per SM: up to 255 registers)
    for name, u in find(usage, "maxwell_stage_class").items():
        assert u["stack"] <= 64, (name, u)


def test_spmv_kernels_use_the_tma_engine():
    """UBLKCP = cp.async.bulk global->shared; SYNCS = mbarrier (the mnemonics B200_PROFILING.md lists)"""
    out = subprocess.run([CUOBJDUMP, "-sass", LIB], capture_output=True, text=True, check=True).stdout
    for fun in ("dg_spmv_tma", "csr_spmv_tma"):
        blocks = [b for b in out.split("Function : ")[1:] if fun in b.split("\n", 1)[0]]
        assert blocks, fun
        for b in blocks:
            assert "UBLKCP" in b and "SYNCS" in b, b.split("\n", 1)[0]
    # the Maxwell stage kernel prefetches its operator slab into L2 with one bulk request per chunk
    blocks = [b for b in out.split("Function : ")[1:] if "13maxwell_stageI" in b.split("\n", 1)[0]]
    assert blocks and all("UBLKPF.L2" in b for b in blocks)          # cp.async.bulk.prefetch.L2
    # the class kernels read their block entries as 128-bit shared-memory broadcasts: one LDS.128 per two DFMA
    blocks = [b for b in out.split("Function : ")[1:] if "13dg_spmv_classILi10ELi5ELb1ELb0" in b.split("\n", 1)[0]]
    assert blocks
    for b in blocks:
        assert b.count("LDS.128") >= 250 and b.count("DFMA") >= 1000 and "DMMA" not in b        # two cells per lane: 4 DFMA per LDS.128
    # the tensor-pipe class kernel: mma.sync m8n8k4 f64 = DMMA.8x8x4; bs = 15 is two n-tiles x four k-steps x five slots x four m-tiles
    blocks = [b for b in out.split("Function : ")[1:] if "17dg_spmv_class_mmaILi15ELi5ELb1E" in b.split("\n", 1)[0]]
    assert blocks
    for b in blocks:
        assert b.count("DMMA") >= 160, b.count("DMMA")
    # the Maxwell stage on the tensor pipe: 16 block products per triangle (flux operands factored) x 2 n-tiles x 4 k-steps at k = 4
    blocks = [b for b in out.split("Function : ")[1:] if "23maxwell_stage_class_mmaILi3ELi4E" in b.split("\n", 1)[0]]
    assert len(blocks) == 2                                  # centred / upwind fluxes
    for b in blocks:
        assert 120 <= b.count("DMMA") <= 140, b.count("DMMA")
    assert any("maxwell_rk_vector" in b.split("\n", 1)[0] for b in out.split("Function : ")[1:])
    # bs = 10: one n-tile on the tensor pipe, rows 8 and 9 on DFMA
    blocks = [b for b in out.split("Function : ")[1:] if "17dg_spmv_class_mmaILi10ELi5ELb1E" in b.split("\n", 1)[0]]
    assert blocks
    for b in blocks:
        assert b.count("DMMA") >= 60 and b.count("DFMA") >= 120, (b.count("DMMA"), b.count("DFMA"))
