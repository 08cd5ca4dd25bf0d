"""Runs a few ILU applications of the pressure / velocity preconditioner (profiling target)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
mesh, dim, refines = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
which = pkg.BENCH_ILU_APPLY_PRES if (len(sys.argv) < 5 or sys.argv[4] == "pres") else pkg.BENCH_ILU_APPLY_VEL
params = pkg.default_params(reynolds=100.0, vel_eps=1e-8, vel_off_diagonals=0, vel_diag_strength=0.0, vel_update_prec=100,
                            n_global_refines=refines)
disc = pkg.Disc(os.path.join(ROOT, "meshes", mesh + ".s35m"), dim, params)
ctx = pkg.Ctx(disc.problem)
ctx.set_field(pkg.PRES_N, disc.initial_pressure)
ctx.set_field(pkg.PRES_N_MINUS_1, disc.initial_pressure)
try:
    ctx.advance(2, 1)  # factorises; under a timing experiment that breaks the sweeps the solvers give up, which is fine here
except pkg.S35Error as err:
    print("advance:", str(err)[:100])
ms, nbytes, launches = ctx.bench_kernel(which, warmup=2, reps=3, flush_l2=False)
print("ms per apply", ms)
