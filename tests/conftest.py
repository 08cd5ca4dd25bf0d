import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import __graft_entry__ as ge  # noqa: E402

MESH_DIR = os.path.join(ROOT, "meshes")
GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")
REFERENCE_DIR = "/root/reference"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    if not os.path.exists(os.path.join(ROOT, "dealii-step-35a_b200", "libstep35a_b200.so")):
        ge.build()
    return ge.load_package()


@pytest.fixture(scope="session")
def ob(pkg):
    import oracle_binding

    oracle_binding.lib()
    return oracle_binding


def run_params(pkg, **kw):
    """parameter-file.prm values of the reference with pressure_fe_degree = 1 (the Q2/Q1 north star)."""
    base = dict(reynolds=100.0, final_time=0.01, dt=5e-4, vel_eps=1e-8, vel_off_diagonals=0, vel_diag_strength=0.0,
                vel_update_prec=100, vel_krylov_size=30, vel_max_iterations=1000, n_global_refines=0, pressure_degree=1,
                rotational=1)
    base.update(kw)
    return pkg.default_params(**base)


_DISC_CACHE = {}


@pytest.fixture(scope="session")
def make_disc(pkg):
    def make(mesh, dim, **kw):
        key = (mesh, dim, tuple(sorted(kw.items())))
        if key not in _DISC_CACHE:
            _DISC_CACHE[key] = pkg.Disc(os.path.join(MESH_DIR, mesh), dim, run_params(pkg, **kw))
        return _DISC_CACHE[key]

    return make


def start_state(pkg, disc, *contexts):
    for c in contexts:
        c.set_field(pkg.PRES_N, disc.initial_pressure)
        c.set_field(pkg.PRES_N_MINUS_1, disc.initial_pressure)


def rel_diff(a, b, scale=None):
    """max |a - b| relative to max |b| (or to a given field scale)."""
    s = scale if scale is not None else max(np.abs(b).max(), 1e-300)
    return np.abs(np.asarray(a) - np.asarray(b)).max() / s


def seeded_vector(n, seed=35):
    return np.random.Generator(np.random.PCG64(seed)).standard_normal(n)


def write_channel_inp(path, nx, ny, length=8.0, height=4.1):
    """A plain channel [0, length] x [0, height] of nx x ny cells as a UCD file with the reference's boundary
    indicators (1 walls, 2 inflow at x = 0, 3 outflow at x = length; nsbench2.inp uses the same)."""
    vid = lambda i, j: j * (nx + 1) + i
    lines = ["%d %d 0 0 0" % ((nx + 1) * (ny + 1), nx * ny + 2 * nx + 2 * ny)]
    for j in range(ny + 1):
        for i in range(nx + 1):
            lines.append("%d %.17g %.17g 0" % (vid(i, j), length * i / nx, height * j / ny))
    k = 1
    for j in range(ny):
        for i in range(nx):
            lines.append("%d 1 quad %d %d %d %d" % (k, vid(i, j), vid(i + 1, j), vid(i + 1, j + 1), vid(i, j + 1)))
            k += 1
    for i in range(nx):
        for j in (0, ny):
            lines.append("%d 1 line %d %d" % (k, vid(i, j), vid(i + 1, j)))
            k += 1
    for j in range(ny):
        for i, ident in ((0, 2), (nx, 3)):
            lines.append("%d %d line %d %d" % (k, ident, vid(i, j), vid(i, j + 1)))
            k += 1
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")


def poiseuille(disc, reynolds, length=8.0, height=4.1, u_max=1.5):
    """The exact steady solution the reference's inflow profile (step-35a.cc:373) belongs to in a plain channel:
    u = (4 Um y (H - y) / H^2, 0), p = 8 Um (L - x) / (Re H^2).  Q2 / Q1 hold it exactly."""
    y = disc.vel_support[:, 1]
    u = [4.0 * u_max * y * (height - y) / height ** 2, np.zeros(disc.nv)]
    p = 8.0 * u_max * (length - disc.pres_support[:, 0]) / (reynolds * height ** 2)
    return u, p
