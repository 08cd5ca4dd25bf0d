"""Host scaffolding: parameter file, mesh readers, refinement, DoFs, constraints and the one-off operators
(everything the reference gets from deal.II at set-up, step-35a.cc:121-293, 738-1164)."""
import os

import numpy as np
import pytest
import scipy.sparse as sp

from conftest import MESH_DIR, REFERENCE_DIR, ROOT, run_params

SHIPPED_PRM = """
# First a global definition
set Method_Form = rotational
subsection Physical data
  set initial_time = 0.
  set final_time   = 0.01
  set Reynolds     = 100
  # set load_checkpoint_file_name = checkpoint_save.h5
  set save_checkpoint_file_name = checkpoint_save.h5
end
subsection Time step data
  set dt = 5e-4
end
subsection Space discretization
  set n_of_refines = 0
  set pressure_fe_degree = 2
end
subsection Data solve velocity
  set max_iterations = 1000  # maximal number of iterations that GMRES must make
  set eps            = 1e-8  # stopping criterion
  set Krylov_size    = 30
  set off_diagonals  = 0
  set diag_strength  = 0.00
  set update_prec    = 100
end
set output_interval = 999
set verbose = false
"""


def test_parameter_defaults_are_the_declared_ones(pkg):
    p = pkg.default_params()
    # step-35a.cc:157-234
    assert p.rotational == 1 and p.initial_time == 0.0 and p.final_time == 1.0 and p.reynolds == 1.0
    assert p.dt == 5e-4 and p.n_global_refines == 0 and p.pressure_degree == 1
    assert (p.vel_max_iterations, p.vel_krylov_size, p.vel_off_diagonals, p.vel_update_prec) == (1000, 30, 60, 15)
    assert p.vel_eps == 1e-12 and p.vel_diag_strength == 0.01
    assert p.verbose == 1 and p.output_interval == 1
    assert p.load_checkpoint_file_name == b"" and p.save_checkpoint_file_name == b""


def test_parameter_file_syntax(pkg):
    p = pkg.parse_params(SHIPPED_PRM)
    assert p.final_time == 0.01 and p.reynolds == 100 and p.pressure_degree == 2
    assert p.vel_eps == 1e-8 and p.vel_off_diagonals == 0 and p.vel_diag_strength == 0.0 and p.vel_update_prec == 100
    assert p.output_interval == 999 and p.verbose == 0
    assert p.save_checkpoint_file_name == b"checkpoint_save.h5" and p.load_checkpoint_file_name == b""
    assert pkg.parse_params("set Method_Form = standard").rotational == 0


@pytest.mark.skipif(not os.path.exists(os.path.join(REFERENCE_DIR, "parameter-file.prm")), reason="reference not mounted")
def test_reads_the_reference_parameter_file(pkg):
    p = pkg.read_params(os.path.join(REFERENCE_DIR, "parameter-file.prm"))
    q = pkg.parse_params(SHIPPED_PRM)
    for name, _ in pkg.Params._fields_:
        assert getattr(p, name) == getattr(q, name), name


@pytest.mark.parametrize("text", ["set max_iterations = 5", "set nonsense = 1",
                                  "subsection Data solve velocity\n set max_iterations = 1001\nend",
                                  "subsection Space discretization\n set n_of_refines = 16\nend",
                                  "subsection Space discretization\n set pressure_fe_degree = 0\nend",
                                  "set Method_Form = implicit", "subsection Time step data\n set dt = -1\nend",
                                  "subsection Time step data\n set dt = 1e-3", "set verbose = maybe"])
def test_parameter_patterns_are_enforced(pkg, text):
    with pytest.raises(pkg.S35Error):
        pkg.parse_params(text)


def test_missing_parameter_file(pkg):
    with pytest.raises(pkg.S35Error) as e:
        pkg.read_params("/nonexistent/parameter-file.prm")
    assert e.value.status == 6


# sizes from SURVEY.md Appendix B (computed from the shipped meshes)
SIZES = [("nsbench2.s35m", 2, 0, 60, 284, 82), ("nsbench2.s35m", 2, 3, 3840, 15712, 4016),
         ("cylinder3d.s35m", 3, 0, 6820, 60648, 8382), ("cylinderextruded.s35m", 3, 0, 21190, 183624, 24783)]


@pytest.mark.parametrize("mesh,dim,refines,cells,nv,np_", SIZES)
def test_dof_counts(pkg, make_disc, mesh, dim, refines, cells, nv, np_):
    disc = make_disc(mesh, dim, n_global_refines=refines)
    assert (disc.n_cells, disc.nv, disc.np_) == (cells, nv, np_)


def test_sparsity_sizes_match_the_survey(pkg, make_disc):
    d = make_disc("cylinder3d.s35m", 3)
    assert (d.info.nnz_vel_unconstrained, d.info.nnz_pres, d.info.nnz_grad) == (3587328, 197718, 889950)
    d = make_disc("cylinderextruded.s35m", 3)
    assert (d.info.nnz_vel_unconstrained, d.info.nnz_pres, d.info.nnz_grad) == (11071728, 603663, 2735895)
    d = make_disc("nsbench2.s35m", 2, n_global_refines=3)
    assert (d.info.nnz_vel_unconstrained, d.info.nnz_pres, d.info.nnz_grad) == (247168, 35088, 96880)


def to_sp(disc, which, k=None):
    rp, cl, val = disc.csr(which)
    if k is not None:
        val = val[k]
    shape = {"velm": (disc.nv, disc.nv), "grad": (disc.nv, disc.np_), "plap": (disc.np_, disc.np_),
             "pmass": (disc.np_, disc.np_)}[which]
    return sp.csr_matrix((val, cl, rp), shape=shape)


@pytest.mark.parametrize("mesh,dim,refines,volume", [("nsbench2.s35m", 2, 0, 25 * 4.1 - 0.5), ("nsbench2.s35m", 2, 2, 25 * 4.1 - 0.5),
                                                      ("cylinder3d.s35m", 3, 0, 25 * 4.5 * 4.5 - 1.0 * 4.5),
                                                      ("cylinderextruded.s35m", 3, 0, 279.299097)])
def test_mass_matrices_integrate_the_domain(pkg, make_disc, mesh, dim, refines, volume):
    disc = make_disc(mesh, dim, n_global_refines=refines)
    for which in ("velm", "pmass"):
        M = to_sp(disc, which)
        assert abs(M.sum() - volume) < 2e-6 * volume  # the 32-gon area is quoted to 6 digits in the survey
        assert abs(M - M.T).max() < 1e-14
        assert M.diagonal().min() > 0


def test_laplacian_annihilates_constants_and_gradient_operator_is_consistent(pkg, make_disc):
    disc = make_disc("cylinder3d.s35m", 3)
    K = to_sp(disc, "plap")
    _, _, proj = disc.constraints()
    mask = np.zeros(disc.np_, bool)
    mask[proj] = True
    touched = np.asarray(abs(to_sp(disc, "pmass"))[:, proj].sum(axis=1)).ravel() > 0
    r = K @ np.ones(disc.np_)
    assert np.abs(r[~touched]).max() < 1e-12
    assert abs(K - K.T).max() < 1e-13
    # constrained rows: only a positive diagonal
    for i in proj[:20]:
        row = K.getrow(i)
        assert row.nnz == 1 and row.indices[0] == i and row.data[0] > 0
    # sum_i (B_d)_{ij} over velocity DoFs = -int d_d(sum_i phi_i) psi_j = 0 (partition of unity)
    for d in range(3):
        B = to_sp(disc, "grad", d)
        assert np.abs(B.T @ np.ones(disc.nv)).max() < 1e-12
    # B_d applied to a linear pressure p = x_d against the mass matrix: sum_ij ... = -int d_d phi_i x_d
    # divergence theorem check: 1^T B_d^T x-coordinate vector of a *linear velocity* u = x e_d gives -|Omega|
    vol = to_sp(disc, "pmass").sum()
    for d in range(3):
        B = to_sp(disc, "grad", d)
        assert abs(np.ones(disc.np_) @ (B.T @ disc.vel_support[:, d]) + vol) < 1e-9 * vol


def test_constraints(pkg, make_disc):
    disc = make_disc("cylinderextruded.s35m", 3)
    wall, flow, proj = disc.constraints()
    X = disc.vel_support
    assert len(wall) == 26342 and len(flow[0][0]) == 777 and len(flow[1][0]) == 1554 and len(proj) == 209
    # inflow face x = 0 for every component, outflow x = 25 only for d >= 1
    assert np.all(X[flow[0][0], 0] == 0.0)
    assert set(np.unique(X[flow[1][0], 0])) == {0.0, 25.0}
    # the reference's 3D profile divides by height^2 only: peak 1.5 * w^2 / 4 (step-35a.cc:377-378)
    y, z = X[flow[0][0], 1], X[flow[0][0], 2]
    w, h = disc.info.width, disc.info.height
    assert (w, h) == (4.5, 2.5)
    assert np.allclose(flow[0][1], 4 * 1.5 * y * (w - y) * z * (h - z) / h ** 2, rtol=0, atol=1e-14)
    assert abs(flow[0][1].max() - 1.5 * w * w / 4) < 1e-12
    assert np.all(flow[1][1] == 0.0) and np.all(flow[2][1] == 0.0)
    assert np.all(disc.pres_support[proj, 0] == 25.0)
    assert np.allclose(disc.initial_pressure, 25.0 - disc.pres_support[:, 0])
    # 2D: parabola normalised with width^2
    d2 = make_disc("nsbench2.s35m", 2, n_global_refines=1)
    _, flow2, _ = d2.constraints()
    y = d2.vel_support[flow2[0][0], 1]
    assert np.allclose(flow2[0][1], 4 * 1.5 * y * (4.1 - y) / 4.1 ** 2, atol=1e-14) and abs(flow2[0][1].max() - 1.5) < 1e-12


def test_constrained_pattern_drops_wall_rows_and_columns(pkg, make_disc):
    disc = make_disc("nsbench2.s35m", 2, n_global_refines=1)
    rp, cl, _ = disc.csr("velc")
    wall, _, _ = disc.constraints()
    mask = np.zeros(disc.nv, bool)
    mask[wall] = True
    rows = np.repeat(np.arange(disc.nv), np.diff(rp))
    assert np.all(np.diff(rp)[wall] == 1) and np.all(cl[rp[wall]] == wall)
    off = rows != cl
    assert not mask[cl[off]].any() and not mask[rows[off]].any()
    # columns ascending, diagonal stored in every row
    for r in range(disc.nv):
        c = cl[rp[r]:rp[r + 1]]
        assert np.all(np.diff(c) > 0) and r in c


def test_colouring_is_conflict_free(pkg, make_disc):
    disc = make_disc("cylinder3d.s35m", 3)
    _, dofs, color, _ = disc.cell_tables()
    for c in range(color.max() + 1):
        d = dofs[color == c].ravel()
        assert len(np.unique(d)) == len(d)
    assert disc.info.n_colors == color.max() + 1 <= 20


def test_renumbering_reduces_the_bandwidth(pkg, make_disc):
    disc = make_disc("cylinderextruded.s35m", 3)
    rp, cl, _ = disc.csr("velm")
    rows = np.repeat(np.arange(disc.nv), np.diff(rp))
    assert np.abs(rows - cl).max() < 9000  # 8 126 with scipy's RCM; natural numbering is > 150 000


def test_refinement_keeps_boundary_indicators_and_geometry(pkg, make_disc):
    coarse, fine = make_disc("nsbench2.s35m", 2), make_disc("nsbench2.s35m", 2, n_global_refines=2)
    _, _, b0, id0 = coarse.mesh()
    v2, c2, b2, id2 = fine.mesh()
    assert len(c2) == 16 * 60 and len(b2) == 4 * len(b0)
    assert sorted(np.unique(id2)) == sorted(np.unique(id0)) == [1, 2, 3, 4]
    for i in (1, 2, 3, 4):
        assert (id2 == i).sum() == 4 * (id0 == i).sum()
    # boundary-id 2 faces stay on x = 0, id 3 on x = 25
    assert np.all(v2[b2[id2 == 2]][:, :, 0] == 0.0) and np.all(v2[b2[id2 == 3]][:, :, 0] == 25.0)
    assert abs(fine.info.min_cell_diameter * 4 - coarse.info.min_cell_diameter) < 0.3


@pytest.mark.skipif(not os.path.isdir(REFERENCE_DIR), reason="reference not mounted")
@pytest.mark.parametrize("name,dim", [("nsbench2.inp", 2), ("cylinder3d.msh", 3), ("cylinderextruded.msh", 3)])
def test_text_readers_against_the_reference_files(pkg, tmp_path, name, dim):
    """an independent Python parse of the reference's mesh file against the library's reader (via .s35m)"""
    path = os.path.join(REFERENCE_DIR, name)
    out = str(tmp_path / "m.s35m")
    pkg.mesh_convert(path, dim, out)
    packed = open(os.path.join(MESH_DIR, os.path.splitext(name)[0] + ".s35m"), "rb").read()
    assert open(out, "rb").read() == packed  # the shipped container is what the reader produces
    disc = pkg.Disc(out, dim, run_params(pkg))
    v, c, b, ids = disc.mesh()
    toks = open(path).read().split()
    if name.endswith(".inp"):
        nv, ne = int(toks[0]), int(toks[1])
        verts = np.array(toks[5:5 + 4 * nv], float).reshape(nv, 4)
        assert np.array_equal(verts[:, 1:1 + dim], v)
        rest = toks[5 + 4 * nv:]
        quads = [rest[i + 1:i + 5] for i in range(len(rest)) if rest[i] == "quad"]
        lines = [(rest[i - 1], rest[i + 1:i + 3]) for i in range(len(rest)) if rest[i] == "line"]
        assert len(quads) == len(c) and len(lines) == len(b)
        q = np.array(quads, int)
        assert np.array_equal(q[:, [0, 1, 3, 2]], c)  # counter-clockwise -> lexicographic
        assert np.array_equal(np.array([int(m) for m, _ in lines]), ids)
    else:
        i0 = toks.index("$Nodes")
        nv = int(toks[i0 + 1])
        verts = np.array(toks[i0 + 2:i0 + 2 + 4 * nv], float).reshape(nv, 4)
        assert np.array_equal(verts[:, 1:], v)
        i1 = toks.index("$Elements")
        ne = int(toks[i1 + 1])
        p = i1 + 2
        hexes, quads, qids = [], [], []
        for _ in range(ne):
            etype, ntags = int(toks[p + 1]), int(toks[p + 2])
            tags = toks[p + 3:p + 3 + ntags]
            nn = {3: 4, 5: 8}[etype]
            nodes = [int(t) - 1 for t in toks[p + 3 + ntags:p + 3 + ntags + nn]]
            if etype == 5:
                hexes.append(nodes)
            else:
                quads.append(nodes)
                qids.append(int(tags[0]))
            p += 3 + ntags + nn
        assert np.array_equal(np.array(hexes)[:, [0, 1, 3, 2, 4, 5, 7, 6]], c)
        assert np.array_equal(np.array(quads)[:, [0, 1, 3, 2]], b) and np.array_equal(np.array(qids), ids)


@pytest.mark.parametrize("name,dim,ext", [("nsbench2", 2, ".inp"), ("nsbench2", 2, ".msh"), ("cylinder3d", 3, ".msh"),
                                          ("cylinder3d", 3, ".inp")])
def test_mesh_writers_and_readers_round_trip(pkg, tmp_path, name, dim, ext):
    src = os.path.join(MESH_DIR, name + ".s35m")
    text = str(tmp_path / ("m" + ext))
    back = str(tmp_path / "back.s35m")
    pkg.mesh_convert(src, dim, text)
    pkg.mesh_convert(text, dim, back)
    a, b = pkg.Disc(src, dim, run_params(pkg)), pkg.Disc(back, dim, run_params(pkg))
    for x, y in zip(a.mesh(), b.mesh()):
        assert np.array_equal(x, y)


def test_bad_mesh_files_are_rejected(pkg, tmp_path):
    with pytest.raises(pkg.S35Error):
        pkg.Disc("/nonexistent.msh", 3, run_params(pkg))
    bad = tmp_path / "bad.inp"
    bad.write_text("4 1 0 0 0\n0 0 0 0\n1 1 0 0\n2 1 1 0\n3 0 1 0\n1 1 quad 0 3 2 1\n")  # clockwise quad
    with pytest.raises(pkg.S35Error):
        pkg.Disc(str(bad), 2, run_params(pkg))
    with pytest.raises(pkg.S35Error):
        pkg.Disc(os.path.join(MESH_DIR, "nsbench2.s35m"), 3, run_params(pkg))  # dimension mismatch
    with pytest.raises(pkg.S35Error):
        pkg.Disc(os.path.join(MESH_DIR, "nsbench2.s35m"), 2, run_params(pkg, pressure_degree=2))  # Q3/Q2 not built


def test_drag_lift_of_a_hydrostatic_and_a_shear_field(pkg, make_disc):
    """known answers of the force functional: constant pressure exerts no net force on a closed obstacle;
    p = x (pressure rising downstream) pushes the body upstream with F_x = -|obstacle| = -0.5 for the diamond of
    nsbench2, i.e. c_d = 2 F_x / (Ubar^2 D) = -1."""
    disc = make_disc("nsbench2.s35m", 2, n_global_refines=2)
    zero_u = np.zeros(2 * disc.nv)
    cd = disc.drag_lift(100.0, zero_u, np.ones(disc.np_))
    assert np.abs(cd).max() < 1e-12
    cd = disc.drag_lift(100.0, zero_u, disc.pres_support[:, 0].copy())
    # F = -int_S p n dS (n out of the body) = -int_body grad p dV = -(area, 0); Ubar = 1, frontal height 1
    assert abs(cd[0] + 2 * 0.5) < 1e-12 and abs(cd[1]) < 1e-12
    # rigid rotation u = (-y, x): symmetric gradient vanishes, no viscous force
    u = np.concatenate([-disc.vel_support[:, 1], disc.vel_support[:, 0]])
    assert np.abs(disc.drag_lift(1.0, u, np.zeros(disc.np_))).max() < 1e-11
