"""Second, independent CPU restatement of the step-35a time step (numpy / scipy / numba).

TEST INFRASTRUCTURE ONLY (see oracle/oracle.h).  Its job is to pin the C++ oracle: it is written from the
same sources (step-35a.cc:903-1545 and SURVEY.md Appendix A) but shares no code with it or with the host
library, and uses different formulations wherever the mathematics allows:

  * shape functions from numpy polynomial fits, assembly vectorised with einsum over all cells,
    global matrices by COO summation (scipy), not by per-cell scatter;
  * DoFs identified by their support-point coordinates (np.unique), not by mesh-entity hashing; they are
    then renumbered to the numbering of the library under test by matching coordinates, so that vectors
    and ILU factors can be compared entry by entry;
  * constraints applied with boolean masks on scipy matrices;
  * GMRES with an explicit small least-squares solve (numpy lstsq) instead of Givens rotations, CG in the
    textbook r = b - Ax form; both reproduce deal.II's iterates and stopping quantities in exact arithmetic;
  * the pressure mass solve with scipy's sparse LU (a direct solve, like the reference's UMFPACK).

Input: the refined mesh (vertices, cells, boundary faces + indicators) and the run parameters.
"""
import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla

try:
    import numba

    njit = numba.njit(cache=False)
except Exception:  # pragma: no cover - numba is present in the image; plain Python is just slower

    def njit(f):
        return f


def gauss01(n):
    x, w = np.polynomial.legendre.leggauss(n)
    return 0.5 * (x + 1.0), 0.5 * w


def lagrange_tables(degree, xq):
    """values[q, i], derivs[q, i] of the Lagrange basis on equidistant nodes of [0, 1]."""
    nodes = np.linspace(0.0, 1.0, degree + 1)
    vals = np.zeros((len(xq), degree + 1))
    ders = np.zeros((len(xq), degree + 1))
    for i in range(degree + 1):
        y = np.zeros(degree + 1)
        y[i] = 1.0
        c = np.polynomial.polynomial.polyfit(nodes, y, degree)
        vals[:, i] = np.polynomial.polynomial.polyval(xq, c)
        ders[:, i] = np.polynomial.polynomial.polyval(xq, np.polynomial.polynomial.polyder(c))
    return vals, ders


def tensor_tables(dim, degree, nq):
    """phi[q, i], gradref[q, i, d], weights[q]; lexicographic q and i (x fastest)."""
    xq, wq = gauss01(nq)
    v1, d1 = lagrange_tables(degree, xq)
    n1 = degree + 1
    if dim == 2:
        phi = np.einsum("bj,ai->baji", v1, v1).reshape(nq ** 2, n1 ** 2)
        gx = np.einsum("bj,ai->baji", v1, d1).reshape(nq ** 2, n1 ** 2)
        gy = np.einsum("bj,ai->baji", d1, v1).reshape(nq ** 2, n1 ** 2)
        grad = np.stack([gx, gy], axis=-1)
        w = np.einsum("b,a->ba", wq, wq).reshape(-1)
    else:
        phi = np.einsum("ck,bj,ai->cbakji", v1, v1, v1).reshape(nq ** 3, n1 ** 3)
        gx = np.einsum("ck,bj,ai->cbakji", v1, v1, d1).reshape(nq ** 3, n1 ** 3)
        gy = np.einsum("ck,bj,ai->cbakji", v1, d1, v1).reshape(nq ** 3, n1 ** 3)
        gz = np.einsum("ck,bj,ai->cbakji", d1, v1, v1).reshape(nq ** 3, n1 ** 3)
        grad = np.stack([gx, gy, gz], axis=-1)
        w = np.einsum("c,b,a->cba", wq, wq, wq).reshape(-1)
    return phi, grad, w


def node_points(dim, degree):
    """reference coordinates of the local nodes, lexicographic."""
    n1 = degree + 1
    t = np.linspace(0.0, 1.0, n1)
    if dim == 1:
        return t.reshape(-1, 1)
    if dim == 2:
        Y, X = np.meshgrid(t, t, indexing="ij")
        return np.stack([X.ravel(), Y.ravel()], axis=-1)
    Z, Y, X = np.meshgrid(t, t, t, indexing="ij")
    return np.stack([X.ravel(), Y.ravel(), Z.ravel()], axis=-1)


def q1_at(points):
    """Q1 shape values[p, v] and gradients[p, v, d] at reference points (lexicographic vertices)."""
    dim = points.shape[1]
    nv = 1 << dim
    val = np.ones((len(points), nv))
    grad = np.ones((len(points), nv, dim))
    for v in range(nv):
        for e in range(dim):
            bit = (v >> e) & 1
            f = points[:, e] if bit else 1.0 - points[:, e]
            df = 1.0 if bit else -1.0
            val[:, v] *= f
            for d in range(dim):
                grad[:, v, d] *= df if d == e else f
    return val, grad


def _keys(pts):
    return np.round(pts * 1e9).astype(np.int64)


class CoordinateIndex:
    def __init__(self, pts):
        self.table = {tuple(k): i for i, k in enumerate(_keys(pts))}
        if len(self.table) != len(pts):
            raise ValueError("support points are not unique")

    def lookup(self, pts):
        return np.array([self.table[tuple(k)] for k in _keys(pts)], dtype=np.int64)


@njit
def _ilu0(n, rowptr, col, diag, lu):
    iw = -np.ones(n, dtype=np.int64)
    for k in range(n):
        for j in range(rowptr[k], rowptr[k + 1]):
            iw[col[j]] = j
        for j in range(rowptr[k], rowptr[k + 1]):
            jrow = col[j]
            if jrow >= k:
                break
            t = lu[j] * lu[diag[jrow]]
            lu[j] = t
            for jj in range(diag[jrow] + 1, rowptr[jrow + 1]):
                jw = iw[col[jj]]
                if jw >= 0:
                    lu[jw] -= t * lu[jj]
        lu[diag[k]] = 1.0 / lu[diag[k]]
        for j in range(rowptr[k], rowptr[k + 1]):
            iw[col[j]] = -1


@njit
def _ilu_solve(n, rowptr, col, diag, lu, src):
    x = src.copy()
    for r in range(n):
        s = x[r]
        for k in range(rowptr[r], diag[r]):
            s -= lu[k] * x[col[k]]
        x[r] = s
    for r in range(n - 1, -1, -1):
        s = x[r]
        for k in range(diag[r] + 1, rowptr[r + 1]):
            s -= lu[k] * x[col[k]]
        x[r] = s * lu[diag[r]]
    return x


class Ilu0:
    """ILU(0) on the pattern of a scipy CSR matrix with sorted indices."""

    def __init__(self, A, strengthen=0.0):
        A = sp.csr_matrix(A)
        A.sort_indices()
        self.n = A.shape[0]
        self.rowptr = A.indptr.astype(np.int64)
        self.col = A.indices.astype(np.int64)
        rows = np.repeat(np.arange(self.n), np.diff(self.rowptr))
        self.diag = np.flatnonzero(rows == self.col).astype(np.int64)
        assert len(self.diag) == self.n
        self.lu = A.data.astype(np.float64).copy()
        if strengthen > 0.0:
            rowsum = np.bincount(rows, weights=np.abs(self.lu), minlength=self.n) - np.abs(self.lu[self.diag])
            self.lu[self.diag] += strengthen * rowsum
        _ilu0(self.n, self.rowptr, self.col, self.diag, self.lu)

    def solve(self, b):
        return _ilu_solve(self.n, self.rowptr, self.col, self.diag, self.lu, np.ascontiguousarray(b, dtype=np.float64))


def gmres_left(A, M, x, b, tol, max_its, n_tmp):
    """deal.II SolverGMRES (left preconditioning, preconditioned residual, restart length n_tmp - 2), SURVEY A.1,
    with the small least-squares problem solved explicitly.  Returns x, iterations, last residual."""
    x = x.copy()
    its = 0
    m = n_tmp - 2
    reorth = False
    while True:
        r = M.solve(b - A @ x)
        beta = np.linalg.norm(r)
        if beta <= tol:
            return x, its, beta, True
        if its >= max_its or beta != beta:
            return x, its, beta, False
        V = [r / beta]
        H = np.zeros((m + 1, m))
        res = beta
        k_used = 0
        done = False
        for j in range(m):
            its += 1
            w = M.solve(A @ V[j])
            consider = (not reorth) and (j % 5 == 4)
            norm_start = np.linalg.norm(w) if consider else 0.0
            for i in range(j + 1):
                H[i, j] = w @ V[i]
                w = w - H[i, j] * V[i]
            nw = np.linalg.norm(w)
            if consider and not (nw > 10.0 * norm_start * np.sqrt(np.finfo(float).eps)):
                reorth = True
            if reorth:
                for i in range(j + 1):
                    c = w @ V[i]
                    H[i, j] += c
                    w = w - c * V[i]
                nw = np.linalg.norm(w)
            H[j + 1, j] = nw
            V.append(w / nw if nw != 0.0 else w)
            k_used = j + 1
            e1 = np.zeros(k_used + 1)
            e1[0] = beta
            y, *_ = np.linalg.lstsq(H[: k_used + 1, :k_used], e1, rcond=None)
            res = np.linalg.norm(e1 - H[: k_used + 1, :k_used] @ y)
            if res <= tol:
                done = True
                break
            if its >= max_its:
                break
        for i in range(k_used):
            x = x + y[i] * V[i]
        if done:
            return x, its, res, True
        if its >= max_its:
            return x, its, res, False


def pcg(A, M, x, b, tol, max_its):
    """deal.II SolverCG in the textbook form; stops on the recurrence residual."""
    x = x.copy()
    r = b - A @ x
    res = np.linalg.norm(r)
    it = 0
    if res <= tol:
        return x, it, res, True
    z = M.solve(r)
    p = z.copy()
    rz = r @ z
    while True:
        it += 1
        Ap = A @ p
        alpha = rz / (p @ Ap)
        x = x + alpha * p
        r = r - alpha * Ap
        res = np.linalg.norm(r)
        if res <= tol:
            return x, it, res, True
        if it >= max_its:
            return x, it, res, False
        z = M.solve(r)
        rz_new = r @ z
        p = z + (rz_new / rz) * p
        rz = rz_new


class NumpyStep35:
    def __init__(self, dim, vertices, cells, bfaces, bface_ids, *, dt, reynolds, eps, max_its=1000, krylov=30,
                 update_prec=100, diag_strength=0.0, rotational=True, vel_numbering=None, pres_numbering=None):
        self.dim, self.dt, self.Re, self.eps = dim, dt, reynolds, eps
        self.max_its, self.krylov, self.update_prec = max_its, krylov, update_prec
        self.diag_strength, self.rotational = diag_strength, rotational
        X = vertices[cells]  # [c, v, dim]
        self.X = X
        nc = len(cells)
        # DoFs from support-point coordinates
        ref2, ref1 = node_points(dim, 2), node_points(dim, 1)
        N2, _ = q1_at(ref2)
        N1, _ = q1_at(ref1)
        pts2 = np.einsum("pv,cvk->cpk", N2, X).reshape(-1, dim)
        pts1 = np.einsum("pv,cvk->cpk", N1, X).reshape(-1, dim)
        self.vel_pts, self.vel_dofs = self._number(pts2, nc, vel_numbering)
        self.pres_pts, self.pres_dofs = self._number(pts1, nc, pres_numbering)
        self.nv, self.np_ = len(self.vel_pts), len(self.pres_pts)
        self.width = vertices[:, 1].max()
        self.height = vertices[:, 2].max() if dim == 3 else 0.0

        # boundary DoFs: nodes of boundary faces, located by coordinates
        vidx, pidx = CoordinateIndex(self.vel_pts), CoordinateIndex(self.pres_pts)
        fref2, fref1 = node_points(dim - 1, 2), node_points(dim - 1, 1)
        F2, _ = q1_at(fref2)
        F1, _ = q1_at(fref1)

        def face_dofs(ids, F, index):
            sel = np.isin(bface_ids, ids)
            if not sel.any():
                return np.zeros(0, dtype=np.int64)
            P = np.einsum("pv,fvk->fpk", F, vertices[bfaces[sel]]).reshape(-1, dim)
            return np.unique(index.lookup(P))

        inflow, outflow = face_dofs([2], F2, vidx), face_dofs([3], F2, vidx)
        self.wall = face_dofs([1, 4], F2, vidx)
        self.proj = face_dofs([3], F1, pidx)
        self.flow = []
        for d in range(dim):
            g = np.zeros(self.nv)
            mask = np.zeros(self.nv, dtype=bool)
            mask[inflow] = True
            if d == 0:
                y = self.vel_pts[inflow, 1]
                if dim == 2:
                    g[inflow] = 4.0 * 1.5 * y * (self.width - y) / self.width ** 2
                else:
                    z = self.vel_pts[inflow, 2]
                    g[inflow] = 4.0 * 1.5 * y * (self.width - y) * z * (self.height - z) / self.height ** 2
            else:
                mask[outflow] = True
            self.flow.append((mask, g))
        self.wall_mask = np.zeros(self.nv, dtype=bool)
        self.wall_mask[self.wall] = True
        self.proj_mask = np.zeros(self.np_, dtype=bool)
        self.proj_mask[self.proj] = True

        # tables
        self.phi_v, self.gref_v, self.w_v = tensor_tables(dim, 2, 3)
        phi_pv, _, _ = tensor_tables(dim, 1, 3)
        phi_p, gref_p, w_p = tensor_tables(dim, 1, 2)
        xq3 = self._qpoints(3)
        xq2 = self._qpoints(2)
        _, dN3 = q1_at(xq3)
        _, dN2 = q1_at(xq2)
        self.G_v, self.JxW_v = self._physical(dN3, self.gref_v, self.w_v)  # [c,q,i,k], [c,q]
        G_p, JxW_p = self._physical(dN2, gref_p, w_p)

        # one-off operators
        I = np.repeat(self.vel_dofs[:, :, None], self.vel_dofs.shape[1], axis=2)
        J = np.repeat(self.vel_dofs[:, None, :], self.vel_dofs.shape[1], axis=1)
        self._I, self._J = I.ravel(), J.ravel()
        Mloc = np.einsum("cq,qi,qj->cij", self.JxW_v, self.phi_v, self.phi_v)
        self.vel_mass = sp.csr_matrix((Mloc.ravel(), (self._I, self._J)), shape=(self.nv, self.nv))
        Ip = np.repeat(self.vel_dofs[:, :, None], self.pres_dofs.shape[1], axis=2).ravel()
        Jp = np.repeat(self.pres_dofs[:, None, :], self.vel_dofs.shape[1], axis=1).ravel()
        self.pres_diff = []
        for d in range(dim):
            B = -np.einsum("cq,cqi,qj->cij", self.JxW_v, self.G_v[..., d], phi_pv)
            self.pres_diff.append(sp.csr_matrix((B.ravel(), (Ip, Jp)), shape=(self.nv, self.np_)))
        Ipp = np.repeat(self.pres_dofs[:, :, None], self.pres_dofs.shape[1], axis=2).ravel()
        Jpp = np.repeat(self.pres_dofs[:, None, :], self.pres_dofs.shape[1], axis=1).ravel()
        Kloc = np.einsum("cq,cqik,cqjk->cij", JxW_p, G_p, G_p)
        Mp = np.einsum("cq,qi,qj->cij", JxW_p, phi_p, phi_p)
        self.pres_mass = sp.csr_matrix((Mp.ravel(), (Ipp, Jpp)), shape=(self.np_, self.np_))
        self.pres_laplace = self._constrained_sum(Kloc, self.pres_dofs, self.proj_mask, self.np_)
        self.prec_pres = Ilu0(self.pres_laplace, diag_strength)
        self.mass_lu = spla.splu(sp.csc_matrix(self.pres_mass))

        # state
        self.u_n = [np.zeros(self.nv) for _ in range(dim)]
        self.u_nm1 = [np.zeros(self.nv) for _ in range(dim)]
        self.u_star = [np.zeros(self.nv) for _ in range(dim)]
        self.pres_n = 25.0 - self.pres_pts[:, 0]
        self.pres_nm1 = self.pres_n.copy()
        self.phi_n = np.zeros(self.np_)
        self.phi_nm1 = np.zeros(self.np_)
        self.vel_advection = None
        self.prec_vel = [None] * dim
        self.its_vel = [0] * dim
        self.its_proj = 0

    def _number(self, pts, nc, numbering):
        keys = _keys(pts)
        _, first, inverse = np.unique(keys, axis=0, return_index=True, return_inverse=True)
        upts = pts[first]
        inverse = inverse.reshape(-1)
        if numbering is not None:
            target = CoordinateIndex(numbering)  # support points of the numbering to adopt
            perm = target.lookup(upts)           # own index -> adopted index
            new_pts = np.empty_like(upts)
            new_pts[perm] = upts
            return new_pts, perm[inverse].reshape(nc, -1)
        return upts, inverse.reshape(nc, -1)

    def _qpoints(self, nq):
        xq, _ = gauss01(nq)
        if self.dim == 2:
            Y, X = np.meshgrid(xq, xq, indexing="ij")
            return np.stack([X.ravel(), Y.ravel()], axis=-1)
        Z, Y, X = np.meshgrid(xq, xq, xq, indexing="ij")
        return np.stack([X.ravel(), Y.ravel(), Z.ravel()], axis=-1)

    def _physical(self, dN, gref, w):
        J = np.einsum("cvk,qvd->cqkd", self.X, dN)
        det = np.linalg.det(J)
        Jinv = np.linalg.inv(J)  # [c,q,d,k]
        G = np.einsum("cqdk,qid->cqik", Jinv, gref)
        return G, det * w[None, :]

    @staticmethod
    def _constrained_sum(loc, dofs, mask, n):
        """distribute_local_to_global with homogeneous Dirichlet constraints.  Built from
        triplets only, so that the result's stored pattern is the structural keep_constrained_dofs=false
        pattern (explicit zeros survive)."""
        nc, dpc, _ = loc.shape
        cm = mask[dofs]  # [c, i]
        keep = ((~cm)[:, :, None] & (~cm)[:, None, :]).ravel()
        I = np.repeat(dofs[:, :, None], dpc, axis=2).ravel()
        J = np.repeat(dofs[:, None, :], dpc, axis=1).ravel()
        diag_loc = np.abs(np.einsum("cii->ci", loc))
        avg = diag_loc.mean(axis=1, keepdims=True)
        diag_loc = np.where(diag_loc != 0.0, diag_loc, avg)
        cdofs = dofs[cm]
        rows = np.concatenate([I[keep], cdofs])
        cols = np.concatenate([J[keep], cdofs])
        vals = np.concatenate([loc.ravel()[keep], diag_loc[cm]])
        A = sp.coo_matrix((vals, (rows, cols)), shape=(n, n)).tocsr()
        A.sort_indices()
        return A

    # ---- the time step ----
    def interpolate_velocity(self):
        for d in range(self.dim):
            self.u_star[d] = 2.0 * self.u_n[d] - self.u_nm1[d]

    def assemble_advection_term(self):
        dim = self.dim
        ul = [self.u_star[d][self.vel_dofs] for d in range(dim)]  # [c, i]
        uq = np.stack([ul[d] @ self.phi_v.T for d in range(dim)], axis=-1)  # [c, q, k]
        div = sum(np.einsum("ci,cqi->cq", ul[d], self.G_v[..., d]) for d in range(dim))
        cq = 0.5 * div + 1.5 / self.dt
        adv = np.einsum("cqk,cqjk->cqj", uq, self.G_v) + cq[:, :, None] * self.phi_v[None, :, :]
        A = (1.0 / self.Re) * np.einsum("cq,cqik,cqjk->cij", self.JxW_v, self.G_v, self.G_v)
        A += np.einsum("cq,qi,cqj->cij", self.JxW_v, self.phi_v, adv)
        self.vel_advection = self._constrained_sum(A, self.vel_dofs, self.wall_mask, self.nv)

    def diffusion_step(self, reinit_prec):
        dim, dt = self.dim, self.dt
        pres_tmp = -self.pres_n - 4.0 / 3.0 * self.phi_n + 1.0 / 3.0 * self.phi_nm1
        for d in range(dim):
            force = self.vel_mass @ (2.0 / dt * self.u_n[d] - 0.5 / dt * self.u_nm1[d]) + self.pres_diff[d] @ pres_tmp
            self.u_nm1[d] = self.u_n[d].copy()
            A = sp.csr_matrix(self.vel_advection, copy=True)
            mask, g = self.flow[d]
            avg = np.abs(A.diagonal()).mean()
            # condense, in place on the stored entries so that the pattern is unchanged
            force = force - A @ np.where(mask, g, 0.0)
            rows = np.repeat(np.arange(self.nv), np.diff(A.indptr))
            touched = mask[rows] | mask[A.indices]
            A.data[touched] = 0.0
            A.data[touched & (rows == A.indices)] = avg
            force[mask] = 0.0
            if reinit_prec:
                self.prec_vel[d] = Ilu0(A, self.diag_strength)
            tol = self.eps * np.linalg.norm(force)
            x, its, _, ok = gmres_left(A, self.prec_vel[d], self.u_n[d], force, tol, self.max_its, self.krylov)
            if not ok:
                raise RuntimeError("GMRES: no convergence")
            self.its_vel[d] = its
            x[mask] = g[mask]
            x[self.wall_mask] = 0.0
            self.u_n[d] = x
            self.last_it_matrix = A

    def projection_step(self):
        rhs = sum(self.pres_diff[d].T @ self.u_n[d] for d in range(self.dim))
        self.phi_nm1 = self.phi_n.copy()
        tol = self.eps * np.linalg.norm(rhs)
        rhs = rhs.copy()
        rhs[self.proj_mask] = 0.0
        x, its, _, ok = pcg(self.pres_laplace, self.prec_pres, self.phi_n, rhs, tol, self.max_its)
        if not ok:
            raise RuntimeError("CG: no convergence")
        self.its_proj = its
        x[self.proj_mask] = 0.0
        self.phi_n = x * (1.5 / self.dt)
        self.pres_tmp = rhs

    def update_pressure(self):
        self.pres_nm1 = self.pres_n.copy()
        if self.rotational:
            self.pres_n = self.mass_lu.solve(self.pres_tmp) / self.Re + self.pres_nm1 + self.phi_n
        else:
            self.pres_n = self.pres_n + self.phi_n

    def advance(self, first_n, n_steps, start_n=2, start_offset_n=2):
        for n in range(first_n, first_n + n_steps):
            if n == start_offset_n:
                self.interpolate_velocity()
                self.assemble_advection_term()
            self.diffusion_step((n % self.update_prec == 0) or (n == start_n))
            self.interpolate_velocity()
            self.assemble_advection_term()
            self.projection_step()
            self.update_pressure()
