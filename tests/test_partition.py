"""Space-filling-curve partition and the multi-rank DSS host logic (CPU only).

The 2-rank test runs the product's own connectivity code (finalize_connectivity in
libthetal_b200.so, host side, no device) in two gloo processes, moves the halo with
torch.distributed send/recv in place of ncclSend/ncclRecv, and checks that the ghost-slot gather
reproduces the single-process DSS bit for bit.
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from e3sm_b200.partition import (Partition, face_mesh, factor, gen_space_curve, genspacepart,  # noqa: E402
                                 space_curve)


def test_factor_order():
    # spacecurve_mod.F90:874-950: all 2s, then 3s, then 5s
    assert factor(30) == [2, 3, 5]
    assert factor(120) == [2, 2, 2, 3, 5]
    assert factor(256) == [2] * 8
    assert factor(7) is None and factor(14) is None


def test_hilbert_level1_and_2_known_order():
    # hilbert(), spacecurve_mod.F90:717-803 with (ma,md,ja,jd) = (0,1,0,1): cells (0,0),(0,1),(1,1),(1,0)
    m = gen_space_curve(2)
    assert m.tolist() == [[0, 1], [3, 2]]
    m4 = gen_space_curve(4)
    assert m4[0, 0] == 0 and m4[3, 0] == 15          # enters bottom-left, leaves bottom-right


@pytest.mark.parametrize("n", [2, 3, 4, 5, 6, 9, 10, 15, 30, 60])
def test_face_curve_is_a_continuous_permutation(n):
    m = gen_space_curve(n)
    assert sorted(m.ravel().tolist()) == list(range(n * n))
    pos = np.zeros((n * n, 2), dtype=int)
    for i in range(n):
        for j in range(n):
            pos[m[i, j]] = (i, j)
    assert np.abs(np.diff(pos, axis=0)).sum(axis=1).max() == 1     # every move is to an edge neighbour
    assert tuple(pos[0]) == (0, 0) and tuple(pos[-1]) == (n - 1, 0)


def test_non_factorable_ne_uses_power_of_two_mesh():
    m = face_mesh(7)                                   # cube_mod.F90:1036-1102
    assert sorted(m.ravel().tolist()) == list(range(49))


@pytest.mark.parametrize("ne", [2, 4, 6, 30])
def test_cube_curve_is_continuous_across_faces(ne):
    from e3sm_b200.geometry import CubedSphere
    sc = space_curve(ne)
    assert sorted(sc.tolist()) == list(range(6 * ne * ne))
    order = np.argsort(sc)
    grid = CubedSphere(ne)                             # cube numbering
    nb = grid.edge_nbrs                                # [nelem,4] edge neighbours
    jumps = sum(1 for a, b in zip(order[:-1], order[1:]) if b not in nb[a])
    assert jumps == 0                                  # faces chained 1,2,6,4,5,3 end to start


@pytest.mark.parametrize("nelem,npart", [(96, 8), (96, 7), (5400, 8), (5400, 7), (150, 4)])
def test_genspacepart_remainder_rule(nelem, npart):
    proc = genspacepart(np.arange(nelem), npart)
    cnt = np.bincount(proc, minlength=npart)
    assert cnt.sum() == nelem and (np.diff(proc) >= 0).all()
    assert cnt.max() - cnt.min() <= (2 if nelem % npart else 0)
    if nelem % npart == 0:
        assert (cnt == nelem // npart).all()


def test_partition_slices():
    p = Partition(30, 8)
    assert p.start.tolist() == [675 * i for i in range(9)]
    assert sorted(p.order.tolist()) == list(range(5400))


# ------------------------------------------------------------------------------------------------
def _rank_main(rank, world, port, ne, nlev, q):
    import torch.distributed as td
    import torch
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    td.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from e3sm_b200.dycore import ThetaLDycore
        from e3sm_b200.geometry import CubedSphere, dss_numpy
        from e3sm_b200.hvcoord import HVCoord
        from e3sm_b200.params import config
        part = Partition(ne, world)
        grid = CubedSphere(ne, order=part.order)
        hv = HVCoord.from_table(nlev)
        p = config("C1").with_(ne=ne)
        e0, e1 = part.local_range(rank)
        d = ThetaLDycore(grid, hv, p, None, None, None, rank=rank, part=part, host_only=True)
        tb = d.halo_tables()
        nloc, ng, L = e1 - e0, tb["nghost"], 5
        rng = np.random.default_rng(7)
        field = rng.standard_normal((grid.nelem, L, 16))          # same on every rank
        tens = np.zeros((nloc + ng, L, 16))
        tens[:nloc] = field[e0:e1]
        # pack -> send/recv -> unpack (what halo_pack_kernel / ncclSend / ncclRecv / halo_unpack_kernel do)
        sidx, ridx = tb["send_idx"], tb["recv_idx"]
        sendbuf = tens[sidx // 16, :, sidx % 16]                   # [nsend, L]
        recvbuf = np.zeros((ridx.size, L))
        reqs, keep = [], []
        for (pr, so, sc_, ro, rc) in tb["peers"]:
            ts = torch.from_numpy(np.ascontiguousarray(sendbuf[so:so + sc_]))
            tr = torch.zeros((rc, L), dtype=torch.float64)
            keep.append((tr, ro, rc))
            reqs.append(td.isend(ts, dst=pr))
            reqs.append(td.irecv(tr, src=pr))
        for r in reqs:
            r.wait()
        for tr, ro, rc in keep:
            recvbuf[ro:ro + rc] = tr.numpy()
        tens[ridx // 16, :, ridx % 16] = recvbuf
        # pull-style DSS with the library's gather table (own value first, then the Fortran unpack order)
        out = tens[:nloc].copy()
        src = tb["dss_src"]
        for j in range(3):
            s = src[:, :, j]
            ok = s >= 0
            se, sp = np.where(ok, s // 16, 0), np.where(ok, s % 16, 0)
            contrib = np.where(ok[:, :, None], tens[se, :, sp], 0.0)      # [nloc,16,L]
            out = out + np.transpose(contrib, (0, 2, 1))
        ref = dss_numpy(field, grid.dss_src)[e0:e1]
        peers = {x[0]: (x[2], x[4]) for x in tb["peers"]}          # peer -> (send_cnt, recv_cnt)
        q.put((rank, bool(np.array_equal(out, ref)), ng, peers))
    finally:
        td.destroy_process_group()


@pytest.mark.parametrize("world,ne", [(2, 4), (3, 4)])
def test_multirank_dss_host_logic_gloo(world, ne):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000) + world
    procs = [ctx.Process(target=_rank_main, args=(r, world, port, ne, 26, q)) for r in range(world)]
    for pr in procs:
        pr.start()
    res = [q.get(timeout=300) for _ in range(world)]
    for pr in procs:
        pr.join(timeout=60)
        assert pr.exitcode == 0
    res.sort()
    table = {r[0]: r[3] for r in res}
    for rank, same, ng, peers in res:
        assert same, f"rank {rank}: partitioned DSS differs from the single-process DSS"
        assert ng > 0 and rank not in peers and len(peers) >= 1
        for peer, (ns, nr) in peers.items():          # what A sends to B is what B expects from A
            assert table[peer][rank] == (nr, ns)
