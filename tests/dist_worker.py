"""Worker of the world_size-2 gloo tests (tests/test_partition.py): one rank of the mesh-partitioned path on
the CPU.  The ownership and halo plans come from the product library (host scaffolding); the ghost exchange and
the reductions run over torch.distributed with the gloo backend; the local arithmetic uses numpy / the oracle.
Nothing here runs on the GPU: it checks the host-side logic of the multi-GPU path."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def exchange(dist, torch, part, which, x):
    """ghost entries of x <- owners' values (the same plan the device path feeds to NCCL)"""
    nb, send_ptr, send_idx, recv_ptr = part.halo(which)
    n_owned = part.nv_owned if which == "vel" else part.np_owned
    reqs, bufs = [], []
    for q, peer in enumerate(nb):
        out = torch.from_numpy(np.ascontiguousarray(x[send_idx[send_ptr[q]:send_ptr[q + 1]]]))
        reqs.append(dist.isend(out, int(peer)))
        inp = torch.empty(int(recv_ptr[q + 1] - recv_ptr[q]), dtype=torch.float64)
        reqs.append(dist.irecv(inp, int(peer)))
        bufs.append((q, inp))
    for r in reqs:
        r.wait()
    for q, inp in bufs:
        x[n_owned + recv_ptr[q]:n_owned + recv_ptr[q + 1]] = inp.numpy()
    return x


def all_sum(dist, torch, value):
    t = torch.tensor([float(value)], dtype=torch.float64)
    dist.all_reduce(t)
    return float(t[0])


def run(rank, world, port, mesh, dim, refines, out_dir):
    import scipy.sparse as sp
    import torch
    import torch.distributed as dist

    import __graft_entry__ as ge
    import oracle_binding as ob
    from conftest import run_params

    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    pkg = ge.load_package()
    disc = pkg.Disc(os.path.join(ROOT, "meshes", mesh), dim, run_params(pkg, n_global_refines=refines))
    part = pkg.Part(disc, world, rank)
    rng = np.random.Generator(np.random.PCG64(5))
    xg_v, xg_p = rng.standard_normal(disc.nv), rng.standard_normal(disc.np_)  # same on every rank
    results = {}

    def local_matrix(which, k=None):
        rp, cl, val = part.csr(which)
        if isinstance(val, list):
            val = val[k]
        n_rows = len(rp) - 1
        n_cols = part.nv if which in ("velm", "velc", "gradT") else part.np_
        return sp.csr_matrix((val, cl, rp), shape=(n_rows, n_cols))

    # --- distributed mat-vecs: owned input values, ghosts by exchange, rows of the owned DoFs
    x = np.zeros(part.nv)
    x[:part.nv_owned] = xg_v[part.vel_global[:part.nv_owned]]
    exchange(dist, torch, part, "vel", x)
    assert np.array_equal(x, xg_v[part.vel_global])  # the exchange delivered exactly the owners' values
    results["velm"] = local_matrix("velm") @ x
    results["gradT"] = sum(local_matrix("gradT", d) @ x for d in range(dim))
    xp = np.zeros(part.np_)
    xp[:part.np_owned] = xg_p[part.pres_global[:part.np_owned]]
    exchange(dist, torch, part, "pres", xp)
    assert np.array_equal(xp, xg_p[part.pres_global])
    results["plap"] = local_matrix("plap") @ xp
    results["grad0"] = local_matrix("grad", 0) @ xp

    # --- distributed dot product
    results["dot"] = all_sum(dist, torch, x[:part.nv_owned] @ x[:part.nv_owned])

    # --- distributed CG on the pressure Laplacian with the block-local ILU(0) (couplings to ghosts dropped)
    A = local_matrix("plap")
    n_own = part.np_owned
    blk = sp.csr_matrix(A[:, :n_own])
    blk.sort_indices()
    lu = ob.ilu0_factor(n_own, blk.indptr, blk.indices, blk.data)
    b_glob = rng.standard_normal(disc.np_)
    _, _, proj = disc.constraints()
    b_glob[proj] = 0.0
    b = b_glob[part.pres_global[:n_own]]
    tol = 1e-10 * np.sqrt(b_glob @ b_glob)
    xs = np.zeros(part.np_)
    g = -b.copy()
    h = ob.ilu0_apply(n_own, blk.indptr, blk.indices, lu, g)
    dvec = np.zeros(part.np_)
    dvec[:n_own] = -h
    gh = all_sum(dist, torch, g @ h)
    its = 0
    while True:
        its += 1
        exchange(dist, torch, part, "pres", dvec)
        Ad = A @ dvec
        alpha = gh / all_sum(dist, torch, dvec[:n_own] @ Ad)
        xs[:n_own] += alpha * dvec[:n_own]
        g += alpha * Ad
        res = np.sqrt(all_sum(dist, torch, g @ g))
        if res <= tol or its >= 1000:
            break
        h = ob.ilu0_apply(n_own, blk.indptr, blk.indices, lu, g)
        gh_new = all_sum(dist, torch, g @ h)
        dvec[:n_own] = (gh_new / gh) * dvec[:n_own] - h
        gh = gh_new
    results["cg_x"] = xs[:n_own]
    results["cg_its"] = its
    results["b_glob"] = b_glob
    results["owned_v"] = part.vel_global[:part.nv_owned]
    results["owned_p"] = part.pres_global[:n_own]
    np.savez(os.path.join(out_dir, "rank%d.npz" % rank), **results)
    dist.barrier()
    dist.destroy_process_group()
