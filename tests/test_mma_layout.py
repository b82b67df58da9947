"""CPU: the fragment geometry of the FP64 tensor-pipe kernels (yfd_mma_core.cuh), checked on the host.

`mma_col<BS>(q, t)` numbers the contraction index of `mma.sync.m8n8k4.f64` so that a lane's operand entries are one 16-byte
load (yfd_class_mma.cuh).  Whatever the numbering, it must be a bijection: every column of the block appears in exactly one
(k-step, lane-in-quad) position, the rest are padding (>= BS, multiplied by stored zeros), and the run-time form used to build
the fragment tables (`mma_col_rt`) must agree with the compile-time form the kernels use.  The header is compiled as host code
by nvcc (no GPU needed)."""
import os
import subprocess

import pytest

from conftest import ROOT

NVCC = "/usr/local/cuda/bin/nvcc"

SRC = r"""
#include <cstdio>
#include "yfd_mma_core.cuh"
using namespace yfd;
template <int BS> int check()
{
    using sh = mma_shape<BS>;
    int seen[64] = {0}, bad = 0;
    const mma_shape_rt rt = mma_shape_of<BS>(4);
    for (int q = 0; q < sh::KS; q++)
        for (int t = 0; t < 4; t++) {
            const int j = mma_col<BS>(q, t);
            if (j != mma_col_rt(rt, q, t)) bad++;
            if (j >= 0 && j < BS) seen[j]++; else if (j < BS) bad++;
        }
    for (int j = 0; j < BS; j++) if (seen[j] != 1) bad++;
    /* rows: NTM tiles of 8 on the tensor pipe + LEFT rows on DFMA cover the block */
    if (8 * sh::NTM + sh::LEFT < BS) bad++;
    if (sh::QS != sh::NTM * 32 + (sh::LEFT ? 8 : 0)) bad++;
    printf("BS %d KS %d NTM %d LEFT %d QS %d bad %d\n", BS, sh::KS, sh::NTM, sh::LEFT, sh::QS, bad);
    return bad;
}
int main() { return check<6>() + check<10>() + check<15>() + check<3>() + check<21>() + check<28>(); }
"""


@pytest.mark.skipif(not os.path.exists(NVCC), reason="nvcc not found")
def test_mma_column_numbering_is_a_bijection(tmp_path):
    src = tmp_path / "t.cu"
    src.write_text(SRC)
    exe = tmp_path / "t"
    inc = os.path.join(ROOT, "yaourt-fem-dg_b200", "csrc")
    subprocess.run([NVCC, "-std=c++17", "-I", inc, "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(src)], check=True,
                   capture_output=True, text=True)
    out = subprocess.run([str(exe)], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout
    # the shapes the kernels are instantiated for
    assert "BS 6 KS 2 NTM 1 LEFT 0 QS 32 bad 0" in out.stdout
    assert "BS 10 KS 3 NTM 1 LEFT 2 QS 40 bad 0" in out.stdout
    assert "BS 15 KS 4 NTM 2 LEFT 0 QS 64 bad 0" in out.stdout
