"""Rows-in-flight variants of the warp-per-row SpMV (S35_SPMV_ROWS) against the plain kernel: bit-identity of the
results through the C ABI, then time and bandwidth on the velocity matrix of the bench workload.

    python tools/tune_spmv.py [mesh=cylinderextruded] [dim=3] [refines=1]
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

VARIANTS = (0, 2, 4, 8)


def main():
    mesh = sys.argv[1] if len(sys.argv) > 1 else "cylinderextruded"
    dim = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    refines = int(sys.argv[3]) if len(sys.argv) > 3 else 1
    pkg = ge.load_package()
    out = {}
    small = pkg.Disc(os.path.join(ROOT, "meshes", "cylinder3d.s35m"), 3, pkg.default_params(vel_off_diagonals=0, n_global_refines=0))
    P = small.problem.contents
    n = P.n_vel
    rp = np.ctypeslib.as_array(P.velm_rowptr, (n + 1,))
    col = np.ctypeslib.as_array(P.velm_col, (int(rp[-1]),))
    val = np.ctypeslib.as_array(P.velm_val, (int(rp[-1]),))
    A = pkg.Matrix(n, n, rp, col, val)
    x = np.random.default_rng(35).standard_normal(n)
    ys = {}
    for r in VARIANTS:
        os.environ["S35_SPMV_ROWS"] = str(r)
        ys[r] = A.spmv(x)
    out["bit_identical_to_plain"] = {str(r): bool(np.array_equal(ys[r], ys[0])) for r in VARIANTS}
    out["nonzero_result"] = bool(np.abs(ys[0]).max() > 0)
    print(json.dumps(out), flush=True)
    disc = pkg.Disc(os.path.join(ROOT, "meshes", mesh + ".s35m"), dim, pkg.default_params(vel_off_diagonals=0, n_global_refines=refines))
    ctx = pkg.Ctx(disc.problem)
    out["spmv_vel"] = {}
    for r in VARIANTS:
        os.environ["S35_SPMV_ROWS"] = str(r)
        ms, nbytes, _ = ctx.bench_kernel(pkg.BENCH_SPMV_VEL, warmup=2, reps=10, flush_l2=True)
        out["spmv_vel"][str(r)] = {"ms": round(ms, 4), "GB/s": round(nbytes / ms / 1e6, 1)}
    print(json.dumps(out), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "tune_spmv.json"), "w") as f:
        json.dump(out, f)


if __name__ == "__main__":
    main()
