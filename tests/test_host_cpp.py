"""Host-side logic of the device library that needs no GPU, compiled from tests/cpp with g++:
the collective combiner's protocol (csrc/combiner.h; 2-4 ranks of three solver threads each against a mock transport
that matches the ranks' operations position by position) and the packing of the wavefront sweep (csrc/wavefront.h;
the packed streams walked the way the kernel walks them, against plain substitutions)."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_combiner_protocol_forms_the_same_rounds_on_every_rank(tmp_path):
    exe = str(tmp_path / "test_combiner")
    env = dict(os.environ)
    env.pop("CXX", None)
    subprocess.run(["g++", "-std=c++17", "-O1", "-pthread", "-I", os.path.join(ROOT, "dealii-step-35a_b200", "csrc"),
                    os.path.join(ROOT, "tests", "cpp", "test_combiner.cpp"), "-o", exe], check=True, env=env)
    r = subprocess.run([exe, "40"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "40 runs consistent" in r.stdout


def test_wavefront_schedule_reproduces_the_substitutions(tmp_path):
    exe = str(tmp_path / "test_wavefront")
    env = dict(os.environ)
    env.pop("CXX", None)
    subprocess.run(["g++", "-std=c++17", "-O1", "-I", os.path.join(ROOT, "dealii-step-35a_b200", "csrc"),
                    os.path.join(ROOT, "tests", "cpp", "test_wavefront.cpp"), "-o", exe], check=True, env=env)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "all schedules reproduce the substitutions" in r.stdout


def test_wavefront_schedule_on_the_matrices_of_the_meshes(tmp_path, pkg, make_disc):
    """The patterns the library really packs (pressure Laplace, velocity mass / iteration matrix of the 2-D and 3-D coarse
    meshes), in the two shapes it chooses.  Long Q2 rows make a warp's old and fresh runs come from different rows: the
    case in which a tile grows beyond its planned size."""
    import numpy as np
    exe = str(tmp_path / "test_wavefront")
    env = dict(os.environ)
    env.pop("CXX", None)
    subprocess.run(["g++", "-std=c++17", "-O2", "-I", os.path.join(ROOT, "dealii-step-35a_b200", "csrc"),
                    os.path.join(ROOT, "tests", "cpp", "test_wavefront.cpp"), "-o", exe], check=True, env=env)
    files = []
    for mesh, dim in (("nsbench2.s35m", 2), ("cylinder3d.s35m", 3)):
        disc = make_disc(mesh, dim)
        for which in ("plap", "velm", "velc"):
            rp, cl, _ = disc.csr(which)
            f = str(tmp_path / ("%s_%s.bin" % (mesh.split(".")[0], which)))
            np.concatenate([np.array([len(rp) - 1, len(cl)], np.int32), rp.astype(np.int32), cl.astype(np.int32)]).tofile(f)
            files.append(f)
    r = subprocess.run([exe] + files, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "all schedules reproduce the substitutions" in r.stdout
