"""Device kernels against the CPU oracle, through the C ABI (host buffers in and out)."""
import numpy as np
import pytest

from conftest import rel_diff, seeded_vector

pytestmark = pytest.mark.gpu

CASES = [("nsbench2.s35m", 2, 0), ("nsbench2.s35m", 2, 3), ("cylinder3d.s35m", 3, 0)]


def tridiagonal(n, lo=-1.0, di=2.5, up=-1.2):
    rowptr, col, val = [0], [], []
    for r in range(n):
        for c, v in ((r - 1, lo), (r, di), (r + 1, up)):
            if 0 <= c < n:
                col.append(c)
                val.append(v + 0.01 * ((r * 7 + c * 3) % 5))
        rowptr.append(len(col))
    return np.array(rowptr, np.int32), np.array(col, np.int32), np.array(val)


@pytest.mark.parametrize("mesh,dim,refines", CASES)
def test_spmv_matches_oracle(pkg, ob, make_disc, mesh, dim, refines):
    disc = make_disc(mesh, dim, n_global_refines=refines)
    for which, nr, nc in (("velm", disc.nv, disc.nv), ("plap", disc.np_, disc.np_), ("pmass", disc.np_, disc.np_),
                          ("grad", disc.nv, disc.np_)):
        rp, cl, val = disc.csr(which)
        vals = val if which == "grad" else [val]
        for v in vals:
            A = pkg.Matrix(nr, nc, rp, cl, v)
            x = seeded_vector(nc, 1)
            y0 = seeded_vector(nr, 2)
            ref = ob.spmv(nr, nc, rp, cl, v, x)
            scale = np.abs(ref).max()
            assert rel_diff(A.spmv(x), ref, scale) < 1e-13
            assert rel_diff(A.spmv(x, y=y0), ob.spmv(nr, nc, rp, cl, v, x, y=y0), scale + 1) < 1e-13
            xt = seeded_vector(nr, 3)
            reft = ob.spmv(nr, nc, rp, cl, v, xt, transpose=True)
            assert rel_diff(A.spmv(xt, transpose=True), reft) < 1e-13
            yt = seeded_vector(nc, 4)
            assert rel_diff(A.spmv(xt, y=yt, transpose=True), ob.spmv(nr, nc, rp, cl, v, xt, y=yt, transpose=True),
                            np.abs(reft).max() + 1) < 1e-13
            A.close()


def test_spmv_empty_rows_and_single_entry(pkg, ob):
    # ragged: empty rows, a one-entry row, a long row
    rowptr = np.array([0, 0, 1, 1, 9, 9], np.int32)
    col = np.array([3, 0, 1, 2, 3, 4, 5, 6, 7], np.int32)
    val = np.arange(1.0, 10.0)
    x = seeded_vector(8)
    A = pkg.Matrix(5, 8, rowptr, col, val)
    assert rel_diff(A.spmv(x), ob.spmv(5, 8, rowptr, col, val, x)) < 1e-15
    assert rel_diff(A.spmv(seeded_vector(5), transpose=True), ob.spmv(5, 8, rowptr, col, val, seeded_vector(5), transpose=True)) < 1e-15


def test_matrix_rejects_bad_input(pkg):
    with pytest.raises(pkg.S35Error) as e:
        pkg.Matrix(2, 2, np.array([0, 2, 3], np.int32), np.array([1, 0, 1], np.int32), np.ones(3))
    assert e.value.status == 1  # unsorted row
    with pytest.raises(pkg.S35Error):
        pkg.Matrix(2, 2, np.array([0, 1, 2], np.int32), np.array([0, 5], np.int32), np.ones(2))


@pytest.mark.parametrize("n", [0, 1, 31, 1000, 123457])
def test_vector_kernels(pkg, n):
    x, y, w = seeded_vector(n, 1), seeded_vector(n, 2), seeded_vector(n, 3)
    if n == 0:
        assert pkg.vec_dot(x, y) == 0.0
        return
    assert np.allclose(pkg.vec_lincomb(2.0, x, -0.5, y), 2.0 * x - 0.5 * y, rtol=0, atol=1e-15 * 4)
    assert np.array_equal(pkg.vec_lincomb(3.0, x), 3.0 * x)
    ref = float(np.dot(x, y))
    assert abs(pkg.vec_dot(x, y) - ref) <= 1e-13 * np.sqrt(n) * max(1.0, abs(ref))
    y2, r = pkg.vec_add_and_dot(-0.25, x, y, w)
    assert np.allclose(y2, y - 0.25 * x, rtol=0, atol=1e-15)
    assert abs(r - float(np.dot(y - 0.25 * x, w))) <= 1e-12 * np.sqrt(n)
    # w aliasing y: squared norm
    y3, r3 = pkg.vec_add_and_dot(0.5, x, y, y)
    assert abs(r3 - float(np.dot(y + 0.5 * x, y + 0.5 * x))) <= 1e-12 * n


def test_dot_is_deterministic(pkg):
    x, y = seeded_vector(300001, 5), seeded_vector(300001, 6)
    assert len({pkg.vec_dot(x, y) for _ in range(5)}) == 1


def test_ilu0_is_exact_lu_on_tridiagonal(pkg, ob):
    n = 2000
    rp, cl, val = tridiagonal(n)
    A = pkg.Matrix(n, n, rp, cl, val)
    P = pkg.Ilu(A)
    P.factor()
    assert P.levels() == (n, n)
    lu_ref = ob.ilu0_factor(n, rp, cl, val)
    assert rel_diff(P.values(), lu_ref) < 1e-14
    # ILU(0) = LU here: applying it solves the system
    xs = seeded_vector(n)
    b = ob.spmv(n, n, rp, cl, val, xs)
    assert rel_diff(P.apply(b), xs) < 1e-11


@pytest.mark.parametrize("mesh,dim,refines", CASES)
def test_ilu0_factor_and_apply_match_oracle(pkg, ob, make_disc, mesh, dim, refines):
    disc = make_disc(mesh, dim, n_global_refines=refines)
    for which, n in (("plap", disc.np_), ("velm", disc.nv)):
        rp, cl, val = disc.csr(which)
        A = pkg.Matrix(n, n, rp, cl, val)
        P = pkg.Ilu(A)
        for strengthen in (0.0, 0.01):
            P.factor(strengthen_diagonal=strengthen)
            lu_ref = ob.ilu0_factor(n, rp, cl, val, strengthen)
            assert rel_diff(P.values(), lu_ref) < 1e-12
            src = seeded_vector(n, 9)
            ref = ob.ilu0_apply(n, rp, cl, lu_ref, src)
            for _ in range(2):  # twice: the armed scratch vector must be restored by the first call
                assert rel_diff(P.apply(src), ref) < 1e-12
        P.close()
        A.close()


def test_ilu0_zero_pivot_is_reported(pkg):
    rp = np.array([0, 2, 4], np.int32)
    cl = np.array([0, 1, 0, 1], np.int32)
    A = pkg.Matrix(2, 2, rp, cl, np.array([1.0, 2.0, 2.0, 4.0]))  # u_22 = 4 - 2*2 = 0
    P = pkg.Ilu(A)
    with pytest.raises(pkg.S35Error) as e:
        P.factor()
    assert e.value.status == 4


def test_condense_matches_oracle(pkg, ob, make_disc):
    disc = make_disc("nsbench2.s35m", 2, n_global_refines=2)
    rp, cl, val = disc.csr("velm")
    _, flow, _ = disc.constraints()
    for dofs, g in flow:
        A = pkg.Matrix(disc.nv, disc.nv, rp, cl, val)
        rhs = seeded_vector(disc.nv, 11)
        rhs_dev = A.condense(rhs, dofs, g)
        val_ref, rhs_ref = ob.condense(disc.nv, rp, cl, val, rhs, dofs, g)
        assert rel_diff(A.get_values(), val_ref) < 1e-14
        assert rel_diff(rhs_dev, rhs_ref) < 1e-14
        A.close()


def test_condense_without_constraints_is_identity(pkg, make_disc):
    disc = make_disc("nsbench2.s35m", 2, n_global_refines=0)
    rp, cl, val = disc.csr("pmass")
    A = pkg.Matrix(disc.np_, disc.np_, rp, cl, val)
    rhs = seeded_vector(disc.np_)
    assert np.array_equal(A.condense(rhs, np.zeros(0, np.int32)), rhs)
    assert np.array_equal(A.get_values(), val)


@pytest.mark.parametrize("mesh,dim,refines", CASES)
def test_cg_matches_oracle_iteration_by_iteration(pkg, ob, make_disc, mesh, dim, refines):
    disc = make_disc(mesh, dim, n_global_refines=refines)
    rp, cl, val = disc.csr("plap")
    n = disc.np_
    A = pkg.Matrix(n, n, rp, cl, val)
    P = pkg.Ilu(A)
    P.factor()
    lu = ob.ilu0_factor(n, rp, cl, val)
    b = seeded_vector(n, 21)
    _, _, proj = disc.constraints()
    b[proj] = 0.0
    tol = 1e-8 * np.linalg.norm(b)
    x_ref, st_ref, ok_ref = ob.cg(n, rp, cl, val, lu, np.zeros(n), b, tol)
    x, st, ok = pkg.cg(A, P, np.zeros(n), b, tol)
    assert ok and ok_ref
    assert st.iterations == st_ref.iterations
    # ~100 iterations of rounding differences in the recurrences: the residuals agree to a few digits
    assert abs(st.final_residual - st_ref.final_residual) <= 1e-2 * st_ref.final_residual + 1e-300
    assert rel_diff(x, x_ref) < 1e-8
    # iteration cap: same iterate after a fixed number of steps, and the failure is reported
    # (a handful of steps: later CG iterates are sensitive to rounding long before the solution is)
    k = min(5, max(1, st_ref.iterations // 2))
    xk_ref, stk_ref, okk_ref = ob.cg(n, rp, cl, val, lu, np.zeros(n), b, 0.0, max_its=k)
    xk, stk, okk = pkg.cg(A, P, np.zeros(n), b, 0.0, max_iterations=k)
    assert not okk and not okk_ref and stk.iterations == k == stk_ref.iterations
    assert rel_diff(xk, xk_ref) < 1e-9
    # identity preconditioner
    xi_ref, sti_ref, _ = ob.cg(n, rp, cl, val, None, np.zeros(n), b, 0.0, max_its=10)
    xi, sti, _ = pkg.cg(A, None, np.zeros(n), b, 0.0, max_iterations=10)
    assert rel_diff(xi, xi_ref) < 1e-10


def nonsymmetric_from(disc, ob):
    """velocity-like nonsymmetric matrix: mass + scaled upper triangle."""
    rp, cl, val = disc.csr("velm")
    rows = np.repeat(np.arange(disc.nv), np.diff(rp))
    v = val * (1.0 + 0.3 * (cl > rows)) + 0.05 * val.max() * (cl == rows)
    return rp, cl, v


@pytest.mark.parametrize("mesh,dim,refines", CASES)
def test_gmres_matches_oracle(pkg, ob, make_disc, mesh, dim, refines):
    disc = make_disc(mesh, dim, n_global_refines=refines)
    rp, cl, val = nonsymmetric_from(disc, ob)
    n = disc.nv
    A = pkg.Matrix(n, n, rp, cl, val)
    P = pkg.Ilu(A)
    P.factor()
    lu = ob.ilu0_factor(n, rp, cl, val)
    b = seeded_vector(n, 31)
    x0 = seeded_vector(n, 32)
    tol = 1e-10 * np.linalg.norm(b)
    x_ref, st_ref, ok_ref = ob.gmres(n, rp, cl, val, lu, x0, b, tol)
    x, st, ok = pkg.gmres(A, P, x0, b, tol)
    assert ok and ok_ref
    assert st.iterations == st_ref.iterations
    assert rel_diff(x, x_ref) < 1e-9
    # restarts: a tiny Krylov space forces several cycles
    x_ref, st_ref, _ = ob.gmres(n, rp, cl, val, lu, x0, b, tol, krylov=4)
    x, st, _ = pkg.gmres(A, P, x0, b, tol, krylov_size=4)
    assert st.iterations == st_ref.iterations and st.iterations > 2
    assert rel_diff(x, x_ref) < 1e-8
    # iteration limit
    x_ref, st_ref, ok_ref = ob.gmres(n, rp, cl, val, lu, x0, b, 0.0, max_its=3)
    x, st, ok = pkg.gmres(A, P, x0, b, 0.0, max_iterations=3)
    assert not ok and not ok_ref and st.iterations == st_ref.iterations == 3
    assert rel_diff(x, x_ref) < 1e-10
    # converged initial guess: zero iterations
    x, st, ok = pkg.gmres(A, P, x_ref * 0, b * 0, 1e-30)
    assert ok and st.iterations == 0
