"""Generates tests/golden/*.npz: fields after a few time steps computed by the independent numpy/scipy
restatement (oracle/oracle_np.py), stored by support point so that they do not depend on a DoF numbering.

The reference ships no golden data and cannot be run (it needs deal.II); these vectors pin the C++ oracle and
the device path against a second implementation and against accidental change.  Run from the repo root:
    python tools/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import __graft_entry__ as ge  # noqa: E402
import oracle_np  # noqa: E402

# (mesh, dim, refines, steps, Reynolds, options other than parameter-file.prm's)
CASES = [("nsbench2", 2, 0, 8, 100.0, {}), ("nsbench2", 2, 1, 5, 100.0, {}), ("cylinder3d", 3, 0, 2, 20.0, {}),
         # the other branches of the time loop: standard (non-rotational) pressure update, strengthened ILU diagonal,
         # preconditioner refreshed every 3 steps
         ("nsbench2", 2, 1, 7, 100.0, {"rotational": 0, "diag_strength": 0.01, "update_prec": 3}),
         # the bench mesh itself (unrefined: 21 190 cells, 0.57 M DoF), first two steps
         ("cylinderextruded", 3, 0, 2, 100.0, {})]


def main():
    pkg = ge.load_package()
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    force = "--force" in sys.argv
    for mesh, dim, refines, steps, reynolds, opt in CASES:
        name = "%s_r%d_steps%d%s.npz" % (mesh, refines, steps, "_variant" if opt else "")
        if os.path.exists(os.path.join(out_dir, name)) and not force:
            print(name, "exists (pass --force to regenerate)")
            continue
        rotational, diag_strength, update_prec = opt.get("rotational", 1), opt.get("diag_strength", 0.0), opt.get("update_prec", 100)
        params = pkg.default_params(reynolds=reynolds, final_time=0.01, dt=5e-4, vel_eps=1e-8, vel_off_diagonals=0,
                                    vel_diag_strength=diag_strength, vel_update_prec=update_prec, rotational=rotational,
                                    n_global_refines=refines)
        disc = pkg.Disc(os.path.join(ROOT, "meshes", mesh + ".s35m"), dim, params)
        v, c, b, ids = disc.mesh()
        ref = oracle_np.NumpyStep35(dim, v, c, b, ids, dt=params.dt, reynolds=reynolds, eps=params.vel_eps,
                                    update_prec=update_prec, diag_strength=diag_strength, rotational=bool(rotational),
                                    vel_numbering=disc.vel_support, pres_numbering=disc.pres_support)
        ref.advance(2, steps)
        # a coarse subset keeps the 3D fixture small: every k-th support point
        kv = max(1, disc.nv // 4000)
        kp = max(1, disc.np_ // 4000)
        np.savez_compressed(
            os.path.join(out_dir, name), mesh=mesh, dim=dim, refines=refines, steps=steps, reynolds=reynolds, dt=params.dt,
            eps=params.vel_eps, rotational=rotational, diag_strength=diag_strength, update_prec=update_prec,
            vel_points=disc.vel_support[::kv], pres_points=disc.pres_support[::kp],
            u=np.stack([ref.u_n[d][::kv] for d in range(dim)]), p=ref.pres_n[::kp], phi=ref.phi_n[::kp],
            u_norm=np.array([np.linalg.norm(np.concatenate(ref.u_n)), np.abs(np.concatenate(ref.u_n)).max()]),
            p_norm=np.array([np.linalg.norm(ref.pres_n), np.abs(ref.pres_n).max()]),
            its_vel=np.array(ref.its_vel), its_proj=ref.its_proj)
        print(name, os.path.getsize(os.path.join(out_dir, name)), "bytes; GMRES its", ref.its_vel, "CG its", ref.its_proj)


if __name__ == "__main__":
    main()
