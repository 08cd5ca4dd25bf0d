"""Times the ILU(0) substitution sweeps under different launch shapes (run on the GPU box).

    python tools/tune_ilu.py <mesh> <dim> <refines> [variant-set]
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import sys, os, json
sys.path.insert(0, %r)
import __graft_entry__ as ge
pkg = ge.load_package()
mesh, dim, refines = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
params = pkg.default_params(reynolds=100.0, vel_eps=1e-8, vel_off_diagonals=0, vel_diag_strength=0.0, vel_update_prec=100,
                            n_global_refines=refines)
disc = pkg.Disc(os.path.join(%r, "meshes", mesh + ".s35m"), dim, params)
ctx = pkg.Ctx(disc.problem)
ctx.set_field(pkg.PRES_N, disc.initial_pressure)
ctx.set_field(pkg.PRES_N_MINUS_1, disc.initial_pressure)
ms_step = ctx.advance(2, 2) / 2
out = {"step_ms": round(ms_step, 2), "its": [ctx.solver_stats(d).iterations for d in range(dim)] + [ctx.solver_stats(-1).iterations]}
for name, k in (("vel", pkg.BENCH_ILU_APPLY_VEL), ("pres", pkg.BENCH_ILU_APPLY_PRES)):
    ms, nbytes, _ = ctx.bench_kernel(k, warmup=2, reps=5, flush_l2=False)
    out[name] = round(ms, 3)
print(json.dumps(out))
''' % (ROOT, ROOT)

VARIANT_SETS = {
    "default": [{}],
    "multi": [{}] + [{"S35_ILU_PER_LANE": str(p)} for p in (1, 2, 4)] + [{"S35_ILU_BLOCKS_PER_SM": str(b)} for b in (1, 2, 8)]
             + [{"S35_ILU_BACKOFF_NS": str(ns)} for ns in (32, 256)] + [{"S35_ILU_ORDER": "natural"}],
    # wavefront sweeps (one block per sweep): warp groups x rows per group, window, ring; the wide sweep beside them
    "wavefront": [{"S35_ILU_MODE": "wide"}, {}]
                 + [{"S35_WAVE_GROUPS": str(g), "S35_WAVE_GROUP_ROWS": str(w)} for g, w in ((2, 256), (4, 192), (6, 128))]
                 + [{"S35_WAVE_RING_KB": "112"}, {"S35_WAVE_WINDOW": "4096", "S35_WAVE_RING_KB": "176"}],
    # small factors: levels of ~16 rows; few warps keep the load queues short
    "narrow": [{}] + [{"S35_WAVE_GROUPS": str(g), "S35_WAVE_GROUP_ROWS": str(w)}
                      for g, w in ((3, 32), (5, 32), (7, 32), (3, 64), (5, 64), (3, 128), (5, 128))],
}


def main():
    mesh, dim, refines = sys.argv[1], sys.argv[2], sys.argv[3]
    variants = VARIANT_SETS[sys.argv[4] if len(sys.argv) > 4 else "default"]
    for v in variants:
        env = dict(os.environ)
        env.update(v)
        try:
            r = subprocess.run([sys.executable, "-c", CHILD, mesh, dim, refines], env=env, capture_output=True, text=True,
                               timeout=int(os.environ.get("S35_TUNE_TIMEOUT", "90")))
        except subprocess.TimeoutExpired:
            print(json.dumps(v), "TIMED OUT (hung kernel?)", flush=True)
            continue
        print(json.dumps(v), r.stdout.strip() or r.stderr[-800:], flush=True)


if __name__ == "__main__":
    main()
