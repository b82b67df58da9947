"""CPU: the multi-GPU host logic (yfd_partition.cpp) -- plan invariants for 1..8 ranks, and a world-size-2 gloo run
that drives a halo exchange + block SpMV + CG scalar all-reduce with the plan, checked against the global product."""
import os
import socket
import sys

import numpy as np
import pytest

from conftest import ROOT


def neighbours_from_oracle(oracle, t, r):
    m = oracle.mesh(t, r)
    cf, ow = oracle.connectivity(m["cells"], m["faces"])
    nc, nv = m["cells"].shape
    nbr = np.full((nc, nv), -1, np.int64)
    for e in range(nv):
        o = ow[cf[:, e]]
        other = np.where(o[:, 0] == np.arange(nc), o[:, 1], o[:, 0])
        nbr[:, e] = other
    return m, nbr


@pytest.mark.parametrize("t,r", [(0, 3), (1, 3), (0, 0)])
@pytest.mark.parametrize("nranks", [1, 2, 3, 4, 8])
def test_plan_invariants(yfd, oracle, t, r, nranks):
    m, nbr = neighbours_from_oracle(oracle, t, r)
    nc = len(nbr)
    if nranks > nc:
        pytest.skip("fewer cells than ranks")
    plans = [yfd.partition_plan(nbr, nranks, q) for q in range(nranks)]
    # contiguous cover of the cell order, balanced to one cell
    assert plans[0]["cell_begin"] == 0 and plans[-1]["cell_end"] == nc
    sizes = []
    for a, b in zip(plans[:-1], plans[1:]):
        assert a["cell_end"] == b["cell_begin"]
    for p in plans:
        sizes.append(p["cell_end"] - p["cell_begin"])
    assert max(sizes) - min(sizes) <= 1
    owner = np.concatenate([np.full(s, q) for q, s in enumerate(sizes)])
    for q, p in enumerate(plans):
        own = np.arange(p["cell_begin"], p["cell_end"])
        # halo = exactly the non-owned neighbours of owned cells, grouped by owner, ascending inside a group
        nb = nbr[own].reshape(-1)
        nb = nb[(nb >= 0) & ((nb < p["cell_begin"]) | (nb >= p["cell_end"]))]
        expect = np.array(sorted(set(nb.tolist()), key=lambda c: (owner[c], c)), np.int64)
        assert np.array_equal(p["halo_cells"], expect)
        assert p["recv_counts"].sum() == len(p["halo_cells"]) and p["send_counts"].sum() == len(p["send_cells"])
        assert np.all(np.diff(p["peers"]) > 0) and q not in p["peers"]
        # symmetry: my halo group owned by peer k == peer k's send list for me, entry for entry
        ro = np.concatenate([[0], np.cumsum(p["recv_counts"])])
        for k, peer in enumerate(p["peers"]):
            pp = plans[peer]
            kk = list(pp["peers"]).index(q)
            so = np.concatenate([[0], np.cumsum(pp["send_counts"])])
            assert np.array_equal(p["halo_cells"][ro[k]:ro[k + 1]], pp["send_cells"][so[kk]:so[kk + 1]])
    if nranks == 1:
        assert len(plans[0]["halo_cells"]) == 0 and len(plans[0]["peers"]) == 0


@pytest.mark.parametrize("t,r", [(0, 3), (1, 3)])
@pytest.mark.parametrize("nranks", [2, 3, 8])
def test_plan_from_cut_faces_is_the_same_plan(yfd, oracle, t, r, nranks):
    """yfd_partition_create_from_cuts (what the library feeds from the device's cut-face list) == yfd_partition_create,
    from ONE list of all cuts of the mesh, in any order, duplicates allowed."""
    m, nbr = neighbours_from_oracle(oracle, t, r)
    nc, nf = nbr.shape
    begins = [(nc * q) // nranks for q in range(nranks + 1)]
    owner = np.searchsorted(begins, np.arange(nc), side="right") - 1
    c = np.repeat(np.arange(nc), nf); n = nbr.reshape(-1)
    cut = (n >= 0) & (owner[c] != owner[np.maximum(n, 0)])
    cc, cn = c[cut], n[cut]
    rng = np.random.default_rng(5)
    order = rng.permutation(len(cc))
    cc, cn = np.concatenate([cc[order], cc[:3]]), np.concatenate([cn[order], cn[:3]])
    for q in range(nranks):
        a = yfd.partition_plan(nbr, nranks, q)
        b = yfd.partition_plan_from_cuts(nc, nranks, q, cc, cn)
        assert a["cell_begin"] == b["cell_begin"] and a["cell_end"] == b["cell_end"]
        for key in ("halo_cells", "send_cells", "peers", "recv_counts", "send_counts"):
            assert np.array_equal(a[key], b[key]), key
    with pytest.raises(yfd.YfdError):
        yfd.partition_plan_from_cuts(nc, nranks, 0, np.array([0]), np.array([nc]))


def test_plan_rejects_bad_arguments(yfd):
    nbr = np.full((4, 3), -1, np.int64)
    with pytest.raises(yfd.YfdError):
        yfd.partition_plan(nbr, 2, 2)
    with pytest.raises(yfd.YfdError):
        yfd.partition_plan(nbr, 0, 0)
    bad = nbr.copy(); bad[0, 0] = 99; bad[3, 0] = 0
    with pytest.raises(yfd.YfdError):
        yfd.partition_plan(bad, 2, 0)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, t, r, k, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    import scipy.sparse as sp
    import oraclebind
    import yaourt_fem_dg_b200 as yfd
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    oracle = oraclebind.Oracle()
    m, nbr = neighbours_from_oracle(oracle, t, r)
    bs = (k + 2) * (k + 1) // 2
    d = oracle.dg_sip(t, m["points"], m["cells"], m["faces"], k, 1.0)
    A = sp.csr_matrix((d["values"], d["colidx"], d["rowptr"]), shape=(d["n"], d["n"]))
    pl = yfd.partition_plan(nbr, world, rank)
    c0, c1 = pl["cell_begin"], pl["cell_end"]
    nown, nhalo = c1 - c0, len(pl["halo_cells"])
    # local numbering [owned | halo], as the device tables have it
    g2l = {c: c - c0 for c in range(c0, c1)}
    g2l.update({int(c): nown + i for i, c in enumerate(pl["halo_cells"])})
    rows = np.arange(c0 * bs, c1 * bs)
    Aloc = A[rows].tocoo()
    lcol = np.array([g2l[int(c) // bs] * bs + int(c) % bs for c in Aloc.col], np.int64)     # KeyError = a needed cell is missing from the halo
    Aloc = sp.csr_matrix((Aloc.data, (Aloc.row, lcol)), shape=(nown * bs, (nown + nhalo) * bs))

    rng = np.random.default_rng(5)
    xg = rng.standard_normal(d["n"])
    x = np.zeros((nown + nhalo) * bs)
    x[:nown * bs] = xg[rows]
    # halo exchange: one packed message per peer, received straight into the halo slots
    so = np.concatenate([[0], np.cumsum(pl["send_counts"])]); ro = np.concatenate([[0], np.cumsum(pl["recv_counts"])])
    reqs, recvs = [], []
    for kk, peer in enumerate(pl["peers"]):
        cells = pl["send_cells"][so[kk]:so[kk + 1]]
        buf = torch.from_numpy(np.concatenate([x[(c - c0) * bs:(c - c0 + 1) * bs] for c in cells]))
        reqs.append(dist.isend(buf, int(peer)))
        rb = torch.empty(int(pl["recv_counts"][kk]) * bs, dtype=torch.float64)
        recvs.append((rb, ro[kk]))
        reqs.append(dist.irecv(rb, int(peer)))
    for q in reqs:
        q.wait()
    for rb, off in recvs:
        x[(nown + off) * bs:(nown + off) * bs + len(rb)] = rb.numpy()
    y = Aloc @ x
    ok_spmv = np.allclose(y, (A @ xg)[rows], rtol=1e-13, atol=1e-13)
    # CG scalar: d.y summed over ranks == the global dot
    part = torch.tensor([float(x[:nown * bs] @ y)], dtype=torch.float64)
    dist.all_reduce(part)
    ok_dot = abs(part.item() - xg @ (A @ xg)) <= 1e-10 * abs(xg @ (A @ xg))
    with open(os.path.join(out_dir, "rank%d.txt" % rank), "w") as f:
        f.write("%d %d %d %d\n" % (ok_spmv, ok_dot, nhalo, len(pl["send_cells"])))
    dist.destroy_process_group()


@pytest.mark.parametrize("t,r,k", [(0, 3, 1), (1, 2, 2)])
def test_world_size_2_gloo_halo_exchange_and_spmv(tmp_path, t, r, k):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(2, port, t, r, k, str(tmp_path)), nprocs=2, join=True)
    for rank in range(2):
        ok_spmv, ok_dot, nhalo, nsend = map(int, open(tmp_path / ("rank%d.txt" % rank)).read().split())
        assert ok_spmv and ok_dot and nhalo > 0 and nsend > 0


def _exchange(dist, torch, pl, bs, vec, c0, nown):
    """fill the halo part of `vec` ([owned | halo], local numbering) from the owners: the explicit exchange"""
    so = np.concatenate([[0], np.cumsum(pl["send_counts"])]); ro = np.concatenate([[0], np.cumsum(pl["recv_counts"])])
    reqs, recvs = [], []
    for kk, peer in enumerate(pl["peers"]):
        cells = pl["send_cells"][so[kk]:so[kk + 1]]
        buf = torch.from_numpy(np.concatenate([vec[(c - c0) * bs:(c - c0 + 1) * bs] for c in cells]))
        reqs.append(dist.isend(buf, int(peer)))
        rb = torch.empty(int(pl["recv_counts"][kk]) * bs, dtype=torch.float64)
        recvs.append((rb, ro[kk]))
        reqs.append(dist.irecv(rb, int(peer)))
    for q in reqs:
        q.wait()
    for rb, off in recvs:
        vec[(nown + off) * bs:(nown + off) * bs + len(rb)] = rb.numpy()


def _worker_cg(rank, world, port, t, r, k, use_pc, out_dir):
    """The multi-GPU CG of yfd_cg.cu on the peer transport, restated in numpy over gloo: per iteration the owners hand out
    y = A d on the cells their neighbours hold as halo, and every rank advances its OWN halo copies of r, z, d with the
    owner's recurrences -- no exchange of d.  Checked: the halo copies equal the owners' values bit for bit at every
    iteration, and the iterates are those of the same CG run with an explicit exchange of d."""
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    import scipy.sparse as sp
    import oraclebind
    import yaourt_fem_dg_b200 as yfd
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    oracle = oraclebind.Oracle()
    m, nbr = neighbours_from_oracle(oracle, t, r)
    bs = (k + 2) * (k + 1) // 2
    sysd = oracle.dg_sip(t, m["points"], m["cells"], m["faces"], k, 1.0)
    A = sp.csr_matrix((sysd["values"], sysd["colidx"], sysd["rowptr"]), shape=(sysd["n"], sysd["n"]))
    pl = yfd.partition_plan(nbr, world, rank)
    c0, c1 = pl["cell_begin"], pl["cell_end"]
    nown, nhalo = c1 - c0, len(pl["halo_cells"])
    g2l = {c: c - c0 for c in range(c0, c1)}
    g2l.update({int(c): nown + i for i, c in enumerate(pl["halo_cells"])})
    rows = np.arange(c0 * bs, c1 * bs)
    Aloc = A[rows].tocoo()
    lcol = np.array([g2l[int(c) // bs] * bs + int(c) % bs for c in Aloc.col], np.int64)
    Aloc = sp.csr_matrix((Aloc.data, (Aloc.row, lcol)), shape=(nown * bs, (nown + nhalo) * bs))
    n, nl = nown * bs, (nown + nhalo) * bs
    pc = np.ones(nl)
    if use_pc:
        pc[:n] = 1.0 / A.diagonal()[rows]
        _exchange(dist, torch, pl, bs, pc, c0, nown)            # the Jacobi diagonal of the halo cells, once

    def allsum(v):
        tt = torch.tensor([float(v)], dtype=torch.float64)
        dist.all_reduce(tt)
        return tt.item()

    def run(exchange_free):
        x = np.zeros(n); rv = np.zeros(nl); dv = np.zeros(nl); yh = np.zeros(nl)
        rv[:n] = sysd["rhs"][rows]
        _exchange(dist, torch, pl, bs, rv, c0, nown)             # halo copies of r0, once
        z = pc * rv
        dv[:] = z
        rho = allsum(rv[:n] @ z[:n])
        same = True
        hist = []
        for it in range(25):
            if not exchange_free:
                _exchange(dist, torch, pl, bs, dv, c0, nown)
            y = Aloc @ dv
            alpha = rho / allsum(dv[:n] @ y)
            x += alpha * dv[:n]
            rv[:n] -= alpha * y
            z[:n] = pc[:n] * rv[:n]
            rr, rz = allsum(rv[:n] @ rv[:n]), allsum(rv[:n] @ z[:n])
            beta = rz / rho
            dv[:n] = z[:n] + beta * dv[:n]
            if exchange_free:
                yh[:n] = y
                _exchange(dist, torch, pl, bs, yh, c0, nown)     # the owners' y lands in the neighbours' zones
                rv[n:] -= alpha * yh[n:]
                z[n:] = pc[n:] * rv[n:]
                dv[n:] = z[n:] + beta * dv[n:]
                chk = dv.copy()
                _exchange(dist, torch, pl, bs, chk, c0, nown)    # what an explicit exchange of d would deliver
                same = same and np.array_equal(chk[n:], dv[n:])
            rho = rz
            hist.append(np.sqrt(rr))
        return x, np.array(hist), same

    xa, ha, same = run(True)
    xb, hb, _ = run(False)
    ok = same and np.array_equal(ha, hb) and np.array_equal(xa, xb) and nhalo > 0
    with open(os.path.join(out_dir, "cg_rank%d.txt" % rank), "w") as f:
        f.write("%d %d %.3e\n" % (ok, same, ha[-1] / ha[0]))
    dist.destroy_process_group()


@pytest.mark.parametrize("world,t,r,k,use_pc", [(2, 0, 3, 1, False), (2, 1, 2, 2, True), (3, 0, 3, 2, True)])
def test_gloo_exchange_free_cg_keeps_halo_copies_bit_identical(tmp_path, world, t, r, k, use_pc):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker_cg, args=(world, port, t, r, k, use_pc, str(tmp_path)), nprocs=world, join=True)
    for rank in range(world):
        ok, same, rel = open(tmp_path / ("cg_rank%d.txt" % rank)).read().split()
        assert int(same) == 1 and int(ok) == 1 and float(rel) < 1.0
