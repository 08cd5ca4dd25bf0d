"""Host-side logic of the multi-GPU path (partition, ownership, ghost layers, exchange plans) on the CPU, with
real two-process communication over torch.distributed / gloo."""
import os
import socket

import numpy as np
import pytest
import scipy.sparse as sp

from conftest import seeded_vector


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


@pytest.mark.parametrize("mesh,dim,refines,n_ranks", [("nsbench2.s35m", 2, 2, 2), ("nsbench2.s35m", 2, 1, 3),
                                                       ("cylinder3d.s35m", 3, 0, 4),
                                                       ("cylinderextruded.s35m", 3, 0, 8)])  # the bench mesh, 8 GPUs
def test_ownership_and_halo_plans_are_consistent(pkg, make_disc, mesh, dim, refines, n_ranks):
    disc = make_disc(mesh, dim, n_global_refines=refines)
    parts = [pkg.Part(disc, n_ranks, r) for r in range(n_ranks)]
    for which, n_global in (("vel", disc.nv), ("pres", disc.np_)):
        owned = [p.vel_global[:p.nv_owned] if which == "vel" else p.pres_global[:p.np_owned] for p in parts]
        allowned = np.concatenate(owned)
        assert len(allowned) == n_global and len(np.unique(allowned)) == n_global  # every DoF has exactly one owner
        for p in parts:
            glob = p.vel_global if which == "vel" else p.pres_global
            n_own = p.nv_owned if which == "vel" else p.np_owned
            assert np.all(np.diff(glob[:n_own]) > 0)  # owned DoFs keep the global (Cuthill-McKee) order
            nb, send_ptr, send_idx, recv_ptr = p.halo(which)
            assert recv_ptr[-1] == len(glob) - n_own
            for q, peer in enumerate(nb):
                # what this rank receives from `peer` is, entry by entry, what `peer` sends to it
                ghosts = glob[n_own + recv_ptr[q]:n_own + recv_ptr[q + 1]]
                pp = parts[peer]
                pglob = pp.vel_global if which == "vel" else pp.pres_global
                pnb, psend_ptr, psend_idx, _ = pp.halo(which)
                me = list(pnb).index(p.rank)
                assert np.array_equal(pglob[psend_idx[psend_ptr[me]:psend_ptr[me + 1]]], ghosts)
    # balanced cells, every cell assigned
    cell_rank = np.zeros(disc.n_cells, np.int32)
    import ctypes as C
    pkg.check_host(pkg.host_lib().s35_disc_partition_cells(disc.handle, n_ranks, cell_rank.ctypes.data_as(C.POINTER(C.c_int32))))
    counts = np.bincount(cell_rank, minlength=n_ranks)
    assert counts.sum() == disc.n_cells and counts.max() - counts.min() <= 1


def test_local_rows_are_the_global_rows(pkg, make_disc):
    disc = make_disc("cylinder3d.s35m", 3)
    rp, cl, val = disc.csr("velm")
    G = sp.csr_matrix((val, cl, rp), shape=(disc.nv, disc.nv))
    x = seeded_vector(disc.nv)
    y = G @ x
    for rank in range(3):
        p = pkg.Part(disc, 3, rank)
        lrp, lcl, lval = p.csr("velm")
        L = sp.csr_matrix((lval, lcl, lrp), shape=(p.nv_owned, p.nv))
        assert np.all(np.diff(lrp) == np.diff(rp)[p.vel_global[:p.nv_owned]])
        assert np.abs(L @ x[p.vel_global] - y[p.vel_global[:p.nv_owned]]).max() < 1e-12 * np.abs(y).max()
        # constraint lists are local indices of owned and ghost DoFs
        wall_g, _, _ = disc.constraints()
        prob = p.problem.contents
        wall_l = np.ctypeslib.as_array(prob.wall_dofs, shape=(prob.n_wall,))
        assert set(p.vel_global[wall_l]) == set(wall_g) & set(p.vel_global)


def test_single_rank_partition_is_the_whole_problem(pkg, make_disc):
    disc = make_disc("nsbench2.s35m", 2, n_global_refines=1)
    p = pkg.Part(disc, 1, 0)
    assert p.nv == p.nv_owned == disc.nv and p.np_ == p.np_owned == disc.np_
    assert np.array_equal(p.vel_global, np.arange(disc.nv))
    assert len(p.halo("vel")[0]) == 0


@pytest.mark.parametrize("mesh,dim,refines", [("nsbench2.s35m", 2, 2), ("cylinder3d.s35m", 3, 0)])
def test_two_ranks_over_gloo(pkg, ob, make_disc, tmp_path, mesh, dim, refines):
    """world_size 2 over gloo: distributed mat-vecs, dot product and block-ILU(0) CG against the global ones"""
    torch = pytest.importorskip("torch")
    import torch.multiprocessing as mp

    import dist_worker

    port = free_port()
    mp.spawn(dist_worker.run, args=(2, port, mesh, dim, refines, str(tmp_path)), nprocs=2, join=True)
    disc = make_disc(mesh, dim, n_global_refines=refines)
    rng = np.random.Generator(np.random.PCG64(5))
    xg_v, xg_p = rng.standard_normal(disc.nv), rng.standard_normal(disc.np_)

    def glob(which, k=None):
        rp, cl, val = disc.csr(which)
        if isinstance(val, list):
            val = val[k]
        shape = {"velm": (disc.nv, disc.nv), "grad": (disc.nv, disc.np_), "plap": (disc.np_, disc.np_)}[which]
        return sp.csr_matrix((val, cl, rp), shape=shape)

    ref = {"velm": glob("velm") @ xg_v, "plap": glob("plap") @ xg_p, "grad0": glob("grad", 0) @ xg_p,
           "gradT": sum(glob("grad", d).T @ xg_v for d in range(dim))}
    out = [np.load(tmp_path / ("rank%d.npz" % r)) for r in range(2)]
    for key, rows in (("velm", "owned_v"), ("grad0", "owned_v"), ("plap", "owned_p"), ("gradT", "owned_p")):
        got = np.empty_like(ref[key])
        for o in out:
            got[o[rows]] = o[key]
        assert np.abs(got - ref[key]).max() <= 1e-12 * np.abs(ref[key]).max()
    assert all(abs(float(o["dot"]) - xg_v @ xg_v) < 1e-9 * (xg_v @ xg_v) for o in out)
    # CG: both ranks stop at the same iteration; the solution solves the global system
    assert out[0]["cg_its"] == out[1]["cg_its"] < 1000
    xs = np.empty(disc.np_)
    for o in out:
        xs[o["owned_p"]] = o["cg_x"]
    b = out[0]["b_glob"]
    assert np.linalg.norm(glob("plap") @ xs - b) <= 1.001e-10 * np.linalg.norm(b)
    # ... and agrees with the single-process oracle solve (different preconditioner, same answer)
    rp, cl, val = disc.csr("plap")
    lu = ob.ilu0_factor(disc.np_, rp, cl, val)
    x1, st, ok = ob.cg(disc.np_, rp, cl, val, lu, np.zeros(disc.np_), b, 1e-10 * np.linalg.norm(b))
    assert ok and np.abs(xs - x1).max() <= 1e-6 * np.abs(x1).max()
    assert out[0]["cg_its"] >= st.iterations  # dropping the couplings across the cut cannot help
