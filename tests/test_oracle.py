"""Pins the CPU oracle (oracle/oracle.cpp).  The reference has no tests or golden vectors (parity unpinned,
see oracle/oracle.h), so the oracle is checked against (1) an independent numpy/scipy restatement,
(2) golden vectors that restatement produced (tests/golden, tools/make_golden.py), (3) known answers of the
solver kernels, (4) invariants of the scheme."""
import os

import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spla

from conftest import write_channel_inp, poiseuille, GOLDEN_DIR, rel_diff, run_params, seeded_vector, start_state


def tridiagonal(n):
    rowptr, col, val = [0], [], []
    for r in range(n):
        for c, v in ((r - 1, -1.0), (r, 2.5), (r + 1, -1.2)):
            if 0 <= c < n:
                col.append(c)
                val.append(v + 0.01 * ((r * 7 + c * 3) % 5))
        rowptr.append(len(col))
    return np.array(rowptr, np.int32), np.array(col, np.int32), np.array(val)


def test_ilu0_is_lu_on_a_tridiagonal_matrix(ob):
    n = 300
    rp, cl, val = tridiagonal(n)
    A = sp.csr_matrix((val, cl, rp), shape=(n, n))
    lu = ob.ilu0_factor(n, rp, cl, val)
    # rebuild L (unit) and U from the factor layout: strictly lower = L, upper = U, diagonal = 1/u_ii
    rows = np.repeat(np.arange(n), np.diff(rp))
    L = sp.csr_matrix((np.where(cl < rows, lu, 0.0), cl, rp), shape=(n, n)) + sp.identity(n)
    U = sp.csr_matrix((np.where(cl > rows, lu, 0.0) + np.where(cl == rows, 1.0 / lu, 0.0), cl, rp), shape=(n, n))
    assert abs(L @ U - A).max() < 1e-13
    b = seeded_vector(n)
    assert rel_diff(ob.ilu0_apply(n, rp, cl, lu, b), spla.spsolve(A.tocsc(), b)) < 1e-12


def test_ilu0_matches_the_definition_on_a_mesh_matrix(ob, make_disc):
    """(LU)_ij = a_ij on the pattern (the defining property of ILU(0)), strengthening included"""
    disc = make_disc("nsbench2.s35m", 2, n_global_refines=1)
    rp, cl, val = disc.csr("plap")
    n = disc.np_
    for strengthen in (0.0, 0.05):
        lu = ob.ilu0_factor(n, rp, cl, val, strengthen)
        rows = np.repeat(np.arange(n), np.diff(rp))
        L = sp.csr_matrix((np.where(cl < rows, lu, 0.0), cl, rp), shape=(n, n)) + sp.identity(n)
        U = sp.csr_matrix((np.where(cl > rows, lu, 0.0) + np.where(cl == rows, 1.0 / lu, 0.0), cl, rp), shape=(n, n))
        A = sp.csr_matrix((val, cl, rp), shape=(n, n)).tolil()
        if strengthen:
            offsum = np.asarray(abs(sp.csr_matrix((val, cl, rp), shape=(n, n))).sum(axis=1)).ravel() - np.abs(A.diagonal())
            A.setdiag(A.diagonal() + strengthen * offsum)
        P = sp.csr_matrix((np.ones_like(val), cl, rp), shape=(n, n))
        assert abs((L @ U - A.tocsr()).multiply(P)).max() < 1e-12


def test_gmres_known_answers(ob):
    n = 200
    rp, cl, val = tridiagonal(n)
    A = sp.csr_matrix((val, cl, rp), shape=(n, n))
    lu = ob.ilu0_factor(n, rp, cl, val)  # exact LU: one iteration
    xs = seeded_vector(n)
    b = A @ xs
    x, st, ok = ob.gmres(n, rp, cl, val, lu, np.zeros(n), b, 1e-12 * np.linalg.norm(b))
    assert ok and st.iterations == 1 and rel_diff(x, xs) < 1e-12
    # Jacobi-like preconditioner (diagonal-only "ILU"): several restarts with a small Krylov space
    lu_diag = np.where(cl == np.repeat(np.arange(n), np.diff(rp)), 1.0 / val, 0.0)
    x, st, ok = ob.gmres(n, rp, cl, val, lu_diag, np.zeros(n), b, 1e-10 * np.linalg.norm(b), krylov=6)
    assert ok and st.iterations > 4 and rel_diff(x, xs) < 1e-8
    # the criterion is the PRECONDITIONED residual: final_residual = ||P^-1 (b - A x)||
    r = ob.ilu0_apply(n, rp, cl, lu_diag, b - A @ x)
    assert abs(np.linalg.norm(r) - st.final_residual) < 1e-3 * st.tolerance + 1e-14
    # exact initial guess: zero iterations, success
    x, st, ok = ob.gmres(n, rp, cl, val, lu, xs, b, 1e-9 * np.linalg.norm(b))
    assert ok and st.iterations == 0
    # iteration limit -> failure, last iterate returned
    x, st, ok = ob.gmres(n, rp, cl, val, lu_diag, np.zeros(n), b, 0.0, max_its=3)
    assert not ok and st.iterations == 3


def test_gmres_and_cg_agree_with_the_independent_restatement(ob, make_disc):
    import oracle_np
    disc = make_disc("nsbench2.s35m", 2, n_global_refines=2)
    rp, cl, val = disc.csr("plap")
    n = disc.np_
    A = sp.csr_matrix((val, cl, rp), shape=(n, n))
    lu = ob.ilu0_factor(n, rp, cl, val)
    M = oracle_np.Ilu0(A)
    assert rel_diff(M.lu, lu) < 1e-13
    b = seeded_vector(n, 4)
    tol = 1e-9 * np.linalg.norm(b)
    x1, st, ok = ob.cg(n, rp, cl, val, lu, np.zeros(n), b, tol)
    x2, its, res, ok2 = oracle_np.pcg(A, M, np.zeros(n), b, tol, 1000)
    assert ok and ok2 and st.iterations == its and rel_diff(x1, x2) < 1e-9
    # nonsymmetric variant for GMRES: Laplacian + mass with a skewed upper triangle
    rows = np.repeat(np.arange(n), np.diff(rp))
    rpm, clm, vm = disc.csr("pmass")
    full = sp.csr_matrix((val, cl, rp), shape=(n, n)) * 1e-3 + sp.csr_matrix((vm, clm, rpm), shape=(n, n))
    full.sort_indices()
    rp, cl = full.indptr.astype(np.int32), full.indices.astype(np.int32)
    rows = np.repeat(np.arange(n), np.diff(rp))
    v2 = full.data * (1.0 + 0.4 * (cl > rows))
    A2 = sp.csr_matrix((v2, cl, rp), shape=(n, n))
    lu2 = ob.ilu0_factor(n, rp, cl, v2)
    x0 = seeded_vector(n, 5)
    for krylov in (30, 5):
        x1, st, ok = ob.gmres(n, rp, cl, v2, lu2, x0, b, tol, krylov=krylov)
        x2, its, res, ok2 = oracle_np.gmres_left(A2, oracle_np.Ilu0(A2), x0, b, tol, 1000, krylov)
        assert ok and ok2 and st.iterations == its
        assert rel_diff(x1, x2) < 1e-8


def test_cg_known_answers(ob):
    n = 150
    rp, cl, _ = tridiagonal(n)
    rows = np.repeat(np.arange(n), np.diff(rp))
    val = np.where(cl == rows, 2.0 + 0.01 * rows, -1.0)  # SPD
    A = sp.csr_matrix((val, cl, rp), shape=(n, n))
    xs = seeded_vector(n)
    b = A @ xs
    lu = ob.ilu0_factor(n, rp, cl, val)
    x, st, ok = ob.cg(n, rp, cl, val, lu, np.zeros(n), b, 1e-12 * np.linalg.norm(b))
    assert ok and st.iterations == 1 and rel_diff(x, xs) < 1e-11  # exact factor: one step
    x, st, ok = ob.cg(n, rp, cl, val, None, np.zeros(n), b, 1e-10 * np.linalg.norm(b), max_its=n + 5)
    assert ok and st.iterations <= n and rel_diff(x, xs) < 1e-7
    # the criterion is the recurrence residual of the UNPRECONDITIONED system
    assert abs(np.linalg.norm(A @ x - b) - st.final_residual) < 1e-6 * np.linalg.norm(b)
    x, st, ok = ob.cg(n, rp, cl, val, None, np.zeros(n), b, 0.0, max_its=4)
    assert not ok and st.iterations == 4


def test_condense_semantics(ob):
    """SURVEY A.4 on a small matrix: lifted rhs, cleared rows and columns, mean |diagonal| on constrained rows"""
    n = 6
    dense = np.arange(1.0, 37.0).reshape(6, 6) + 10 * np.eye(6)
    A = sp.csr_matrix(dense)
    A.sort_indices()
    rhs = np.arange(1.0, 7.0)
    dofs, g = np.array([1, 4], np.int32), np.array([2.0, -3.0])
    val, out = ob.condense(n, A.indptr, A.indices, A.data, rhs, dofs, g)
    C = sp.csr_matrix((val, A.indices, A.indptr), shape=(n, n)).toarray()
    avg = np.abs(np.diag(dense)).mean()
    free = [0, 2, 3, 5]
    assert np.array_equal(C[np.ix_(free, free)], dense[np.ix_(free, free)])
    assert np.all(C[:, [1, 4]][free] == 0) and np.all(C[[1, 4]][:, free] == 0)
    assert C[1, 1] == avg and C[4, 4] == avg and C[1, 4] == 0 and C[4, 1] == 0
    assert np.allclose(out[free], rhs[free] - dense[np.ix_(free, [1, 4])] @ g) and out[1] == 0 and out[4] == 0


CASES = [("nsbench2.s35m", 2, 0, 8, 100.0, 1e-8), ("nsbench2.s35m", 2, 1, 5, 100.0, 1e-8), ("cylinder3d.s35m", 3, 0, 2, 20.0, 1e-8)]


@pytest.mark.parametrize("mesh,dim,refines,steps,reynolds,tol", CASES)
def test_time_steps_against_the_independent_restatement(pkg, ob, make_disc, mesh, dim, refines, steps, reynolds, tol):
    import oracle_np
    disc = make_disc(mesh, dim, n_global_refines=refines, reynolds=reynolds)
    orc = ob.Oracle(disc.problem)
    start_state(pkg, disc, orc)
    v, c, b, ids = disc.mesh()
    ref = oracle_np.NumpyStep35(dim, v, c, b, ids, dt=5e-4, reynolds=reynolds, eps=1e-8, vel_numbering=disc.vel_support,
                                pres_numbering=disc.pres_support)
    nnz = int(disc.csr("velc")[0][-1])
    for n in range(2, 2 + steps):
        orc.advance(n, 1)
        ref.advance(n, 1)
        assert [orc.solver_stats(d).iterations for d in range(dim)] == ref.its_vel
        assert orc.solver_stats(-1).iterations == ref.its_proj
    # operators of the last step
    assert rel_diff(orc.get_operator_values(pkg.VEL_ADVECTION, nnz), ref.vel_advection.data) < 1e-12
    assert rel_diff(orc.get_operator_values(pkg.VEL_IT_MATRIX, nnz, dim - 1), ref.last_it_matrix.data) < 1e-12
    assert rel_diff(orc.get_operator_values(pkg.VEL_ILU, nnz, 0), ref.prec_vel[0].lu) < 1e-10
    u_o = np.concatenate([orc.get_field(pkg.U_N, d) for d in range(dim)])
    u_r = np.concatenate(ref.u_n)
    assert rel_diff(u_o, u_r) < tol and np.linalg.norm(u_o - u_r) < tol * np.linalg.norm(u_r)
    for f, r in ((pkg.PRES_N, ref.pres_n), (pkg.PHI_N, ref.phi_n)):
        assert rel_diff(orc.get_field(f), r) < tol and np.linalg.norm(orc.get_field(f) - r) < tol * np.linalg.norm(r)


def test_conditioning_limits_field_agreement_on_finer_meshes(pkg, ob, make_disc):
    """Two correct implementations stop their Krylov solvers at the same iteration and still differ by about
    eps * condition number: on nsbench2 r3 that is 1e-7 in the pressure -- the reason why the device tests
    carry a stated tolerance of 2e-6 for this mesh instead of the north star's 1e-8."""
    import oracle_np
    disc = make_disc("nsbench2.s35m", 2, n_global_refines=3)
    orc = ob.Oracle(disc.problem)
    start_state(pkg, disc, orc)
    v, c, b, ids = disc.mesh()
    ref = oracle_np.NumpyStep35(2, v, c, b, ids, dt=5e-4, reynolds=100.0, eps=1e-8, vel_numbering=disc.vel_support,
                                pres_numbering=disc.pres_support)
    orc.advance(2, 6)
    ref.advance(2, 6)
    dp = rel_diff(orc.get_field(pkg.PRES_N), ref.pres_n)
    du = rel_diff(np.concatenate([orc.get_field(pkg.U_N, d) for d in range(2)]), np.concatenate(ref.u_n))
    assert 1e-9 < dp < 2e-6 and du < 2e-6


@pytest.mark.parametrize("name", sorted(f for f in os.listdir(GOLDEN_DIR) if f.endswith(".npz")) if os.path.isdir(GOLDEN_DIR) else [])
def test_oracle_reproduces_the_golden_vectors(pkg, ob, make_disc, name):
    g = np.load(os.path.join(GOLDEN_DIR, name))
    dim, refines, steps = int(g["dim"]), int(g["refines"]), int(g["steps"])
    disc = make_disc(str(g["mesh"]) + ".s35m", dim, n_global_refines=refines, reynolds=float(g["reynolds"]), **golden_options(g))
    orc = ob.Oracle(disc.problem)
    start_state(pkg, disc, orc)
    orc.advance(2, steps)
    assert [orc.solver_stats(d).iterations for d in range(dim)] == list(g["its_vel"])
    check_against_golden(pkg, disc, orc, g, 1e-8)


def golden_options(g):
    """solver options a fixture was made with, where they differ from parameter-file.prm's"""
    opt = {}
    if "rotational" in g.files:
        opt = dict(rotational=int(g["rotational"]), vel_diag_strength=float(g["diag_strength"]), vel_update_prec=int(g["update_prec"]))
    return opt


def check_against_golden(pkg, disc, c, g, tol):
    """fields by support point (the fixtures hold every k-th point) relative to the stored field norms"""
    dim = int(g["dim"])
    kv, kp = max(1, disc.nv // 4000), max(1, disc.np_ // 4000)
    assert np.array_equal(disc.vel_support[::kv], g["vel_points"]) and np.array_equal(disc.pres_support[::kp], g["pres_points"])
    for d in range(dim):
        assert np.abs(c.get_field(pkg.U_N, d)[::kv] - g["u"][d]).max() <= tol * g["u_norm"][1]
    assert np.abs(c.get_field(pkg.PRES_N)[::kp] - g["p"]).max() <= tol * g["p_norm"][1]
    assert np.abs(c.get_field(pkg.PHI_N)[::kp] - g["phi"]).max() <= tol * max(np.abs(g["phi"]).max(), 1e-300)


def test_stepping_is_restartable_and_deterministic(pkg, ob, make_disc):
    disc = make_disc("nsbench2.s35m", 2, n_global_refines=1)
    a, b = ob.Oracle(disc.problem), ob.Oracle(disc.problem)
    start_state(pkg, disc, a, b)
    a.advance(2, 6)
    b.advance(2, 3)
    b.advance(5, 3)
    for f in (pkg.PRES_N, pkg.PHI_N):
        assert np.array_equal(a.get_field(f), b.get_field(f))
    for d in range(2):
        assert np.array_equal(a.get_field(pkg.U_N, d), b.get_field(pkg.U_N, d))


def test_threaded_oracle_is_bit_identical(pkg, ob, make_disc):
    disc = make_disc("nsbench2.s35m", 2, n_global_refines=2)
    a = ob.Oracle(disc.problem, threads=1)
    start_state(pkg, disc, a)
    a.advance(2, 3)
    ua, pa = [a.get_field(pkg.U_N, d) for d in range(2)], a.get_field(pkg.PRES_N)
    b = ob.Oracle(disc.problem, threads=4)
    start_state(pkg, disc, b)
    b.advance(2, 3)
    ob.lib().oracle_set_threads(1)
    assert np.array_equal(pa, b.get_field(pkg.PRES_N))
    for d in range(2):
        assert np.array_equal(ua[d], b.get_field(pkg.U_N, d))


def test_scheme_invariants(pkg, ob, make_disc):
    """boundary values are imposed exactly; the discrete divergence decays after the impulsive start;
    u_star is the second-order extrapolation"""
    disc = make_disc("nsbench2.s35m", 2, n_global_refines=2)
    orc = ob.Oracle(disc.problem)
    start_state(pkg, disc, orc)
    wall, flow, proj = disc.constraints()
    rp, cl, vals = disc.csr("grad")
    div_norms = []
    for n in range(2, 14):
        orc.advance(n, 1)
        u = [orc.get_field(pkg.U_N, d) for d in range(2)]
        um = [orc.get_field(pkg.U_N_MINUS_1, d) for d in range(2)]
        for d in range(2):
            assert np.all(u[d][wall] == 0.0)
            assert np.array_equal(u[d][flow[d][0]], flow[d][1])
            assert np.array_equal(orc.get_field(pkg.U_STAR, d), 2.0 * u[d] + -1.0 * um[d])
        assert np.all(orc.get_field(pkg.PHI_N)[proj] == 0.0)
        div = sum(ob.spmv(disc.nv, disc.np_, rp, cl, vals[d], u[d], transpose=True) for d in range(2))
        div_norms.append(np.linalg.norm(div))
    assert div_norms[-1] < 0.2 * div_norms[0]
    assert np.isfinite(orc.get_field(pkg.PRES_N)).all()


def test_oracle_reports_no_convergence(pkg, ob, make_disc):
    disc = make_disc("nsbench2.s35m", 2, n_global_refines=2, vel_max_iterations=2)
    orc = ob.Oracle(disc.problem)
    start_state(pkg, disc, orc)
    with pytest.raises(ob.OracleError):
        orc.advance(2, 1)


def test_config_1_hundred_steps_drag_and_lift(pkg, ob, make_disc):
    """BASELINE.json config 1 (nsbench2.inp, Re = 100, 100 time steps): both CPU restatements agree on the
    fields to 1e-8 and on the drag / lift coefficients to 1e-6 over the whole run."""
    import oracle_np
    disc = make_disc("nsbench2.s35m", 2, final_time=0.0505)
    orc = ob.Oracle(disc.problem)
    start_state(pkg, disc, orc)
    v, c, b, ids = disc.mesh()
    ref = oracle_np.NumpyStep35(2, v, c, b, ids, dt=5e-4, reynolds=100.0, eps=1e-8, vel_numbering=disc.vel_support,
                                pres_numbering=disc.pres_support)
    n_steps = int((0.0505 - 0.0) / 5e-4) - 1  # run() does steps n = 2 .. (T - t0)/dt (step-35a.cc:1174,1190-1191)
    assert n_steps == 100
    for first in range(2, 2 + n_steps, 25):
        orc.advance(first, 25)
        ref.advance(first, 25)
        u_o = np.concatenate([orc.get_field(pkg.U_N, d) for d in range(2)])
        u_r = np.concatenate(ref.u_n)
        assert rel_diff(u_o, u_r) < 1e-8 and rel_diff(orc.get_field(pkg.PRES_N), ref.pres_n) < 1e-8
        cd_o = disc.drag_lift(100.0, u_o, orc.get_field(pkg.PRES_N))
        cd_r = disc.drag_lift(100.0, u_r, ref.pres_n)
        assert np.abs(cd_o - cd_r).max() <= 1e-6 * max(1.0, np.abs(cd_r).max())
    # after the impulsive start has decayed the drag on the square obstacle is positive and O(1)
    assert 0.5 < cd_o[0] < 50 and abs(cd_o[1]) < 5


def test_block_ilu_alternative_in_the_oracle(pkg, ob, make_disc):
    """s35_problem.ilu_blocks (gated alternative, no counterpart in the reference): another preconditioner, the same
    converged fields to the solver tolerance, no fewer CG iterations."""
    disc = make_disc("nsbench2.s35m", 2, n_global_refines=1, vel_eps=1e-12)
    runs = []
    for k in (1, 4):
        orc = ob.Oracle(pkg.with_ilu_blocks(disc.problem, k))
        start_state(pkg, disc, orc)
        orc.advance(2, 3)
        runs.append(([orc.get_field(pkg.U_N, d) for d in range(2)] + [orc.get_field(pkg.PRES_N)],
                     orc.solver_stats(-1).iterations))
        orc.close()
    for a, b in zip(runs[0][0], runs[1][0]):
        assert not np.array_equal(a, b)
        assert np.max(np.abs(a - b)) <= 5e-6 * max(1.0, np.max(np.abs(a)))
    assert runs[1][1] >= runs[0][1]


def test_off_diagonals_widen_the_factor_pattern_in_the_oracle(pkg, ob, make_disc):
    """off_diagonals = k (SparseILU::AdditionalData::extra_off_diagonals, step-35a.cc:1070-1074, 1312-1319): the factor is
    the ILU(0) of the matrix on its pattern plus the entries (r, r - k .. r + k).  The widened pattern is rebuilt here with
    scipy, independently of the oracle's own construction; fewer iterations than with the plain pattern."""
    k = 5
    disc = make_disc("nsbench2.s35m", 2, n_global_refines=1, vel_off_diagonals=k)
    plain_disc = make_disc("nsbench2.s35m", 2, n_global_refines=1)
    orc, plain = ob.Oracle(disc.problem), ob.Oracle(plain_disc.problem)
    start_state(pkg, disc, orc, plain)
    orc.advance(2, 1)       # the first step factorises its own matrix: VEL_IT_MATRIX and VEL_ILU belong together
    plain.advance(2, 1)
    rp, cl, _ = disc.csr("velc")
    n, nnz = disc.nv, int(rp[-1])
    rows = np.repeat(np.arange(n), np.diff(rp))
    band = sp.diags([np.ones(n - abs(o)) for o in range(-k, k + 1)], list(range(-k, k + 1)), format="csr")
    mark = (sp.csr_matrix((np.ones(nnz), cl, rp), shape=(n, n)) + band).tocsr()   # structural union, all entries > 0
    mark.sort_indices()
    assert mark.nnz > nnz
    for d in range(2):
        a = orc.get_operator_values(pkg.VEL_IT_MATRIX, nnz, d)
        A = sp.csr_matrix((a, cl, rp), shape=(n, n))
        values = np.asarray(A[np.repeat(np.arange(n), np.diff(mark.indptr)), mark.indices]).ravel()
        lu = ob.ilu0_factor(n, mark.indptr.astype(np.int32), mark.indices.astype(np.int32), values, 0.0)
        at_matrix_positions = np.asarray(sp.csr_matrix((lu, mark.indices, mark.indptr), shape=(n, n))[rows, cl]).ravel()
        assert rel_diff(orc.get_operator_values(pkg.VEL_ILU, nnz, d), at_matrix_positions) < 1e-13
    orc.advance(3, 2)
    plain.advance(3, 2)
    for d in range(2):
        assert orc.solver_stats(d).iterations <= plain.solver_stats(d).iterations
    assert orc.solver_stats(-1).iterations < plain.solver_stats(-1).iterations


@pytest.mark.parametrize("reynolds,rotational", [(100.0, 1), (10.0, 1), (100.0, 0)])
def test_known_answer_poiseuille_flow_is_a_fixed_point(pkg, ob, tmp_path, reynolds, rotational):
    """Analytic known answer (the reference ships none): plane Poiseuille flow solves the Navier-Stokes equations with
    the reference's inflow profile, no-slip walls and its outflow condition, and Q2/Q1 represent it exactly.  Started
    from it, the time loop has to stay on it -- which needs the mass, stiffness, gradient and advection operators,
    their signs, the BDF2 weights, the constraints and the pressure update all to be right."""
    mesh = str(tmp_path / "channel.inp")
    write_channel_inp(mesh, 8, 4)
    params = run_params(pkg, reynolds=reynolds, vel_eps=1e-12, n_global_refines=1, rotational=rotational)
    disc = pkg.Disc(mesh, 2, params)
    u, p = poiseuille(disc, reynolds)
    orc = ob.Oracle(disc.problem)
    for field in (pkg.U_N, pkg.U_N_MINUS_1):
        for d in range(2):
            orc.set_field(field, u[d], d)
    orc.set_field(pkg.PRES_N, p)
    orc.set_field(pkg.PRES_N_MINUS_1, p)
    orc.advance(2, 10)
    assert np.abs(orc.get_field(pkg.U_N, 0) - u[0]).max() < 1e-9
    assert np.abs(orc.get_field(pkg.U_N, 1)).max() < 1e-9
    assert np.abs(orc.get_field(pkg.PRES_N) - p).max() < 1e-8 * max(1.0, np.abs(p).max())
    assert np.abs(orc.get_field(pkg.PHI_N)).max() < 1e-8
    # and it is not a fixed point of a wrong pressure: the same start with p = 0 moves away at once
    orc2 = ob.Oracle(disc.problem)
    for field in (pkg.U_N, pkg.U_N_MINUS_1):
        for d in range(2):
            orc2.set_field(field, u[d], d)
    orc2.advance(2, 1)
    assert np.abs(orc2.get_field(pkg.PHI_N)).max() > 1e-6
