"""The C-ABI shared library: it loads, exports every symbol the headers declare, and refuses to compute
without a CUDA device instead of falling back to the CPU."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import ROOT


def declared_symbols():
    names = set()
    for header in ("step35a_b200.h", "step35a_host.h"):
        text = open(os.path.join(ROOT, "include", header)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        names.update(re.findall(r"\b(s35_[a-z0-9_]+)\s*\(", text))
    return sorted(names)


def test_headers_declare_something():
    syms = declared_symbols()
    assert "s35_ctx_create" in syms and "s35_spmv" in syms and "s35_sim_run" in syms
    assert len(syms) > 50


def test_library_exports_every_declared_symbol(pkg):
    handle = ctypes.CDLL(pkg.LIB_PATH)
    missing = [s for s in declared_symbols() if not hasattr(handle, s)]
    assert missing == []


def test_host_only_library_exports_the_set_up_entry_points_and_links_no_cuda(pkg):
    """libstep35a_host.so: what the CPU reference arm of bench.py and the oracle tests load instead of the CUDA library"""
    import subprocess
    handle = ctypes.CDLL(pkg.HOST_LIB_PATH)
    assert [s for s in pkg.HOST_ONLY_SYMBOLS if not hasattr(handle, s)] == []
    assert not hasattr(handle, "s35_ctx_create") and not hasattr(handle, "s35_sim_run")
    out = subprocess.run(["ldd", pkg.HOST_LIB_PATH], capture_output=True, text=True).stdout
    assert "cuda" not in out.lower() and "nccl" not in out.lower()
    dynsyms = subprocess.run(["nm", "-D", "--undefined-only", pkg.HOST_LIB_PATH], capture_output=True, text=True).stdout
    assert "cuda" not in dynsyms.lower()


def test_setting_a_problem_up_does_not_load_the_cuda_library():
    """a fresh interpreter that builds a discretisation maps libstep35a_host.so only"""
    import subprocess
    import sys
    code = ("import sys, os; sys.path.insert(0, %r); import __graft_entry__ as ge; pkg = ge.load_package(); "
            "d = pkg.Disc(os.path.join(%r, 'meshes', 'nsbench2.s35m'), 2, pkg.default_params(vel_off_diagonals=0)); "
            "maps = open('/proc/self/maps').read(); print('libstep35a_b200' in maps, 'libstep35a_host' in maps)") % (ROOT, ROOT)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
    assert out.stdout.split() == ["False", "True"], out.stdout + out.stderr


def test_binding_covers_every_declared_symbol(pkg):
    assert sorted(set(pkg.DECLARED_SYMBOLS)) == declared_symbols()


def test_no_torch_types_in_the_abi():
    for header in ("step35a_b200.h", "step35a_host.h"):
        text = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", header)).read(), flags=re.S)
        assert "torch" not in text and "at::" not in text and "std::" not in text and "&" not in text


def test_product_does_not_touch_the_oracle():
    """only tests/, __graft_entry__.smoke() and bench.py's CPU legs may use oracle/"""
    pkg_dir = os.path.join(ROOT, "dealii-step-35a_b200")
    for base, _, files in os.walk(pkg_dir):
        if os.path.basename(base) == "build":
            continue
        for f in files:
            if f.endswith((".cu", ".cuh", ".cpp", ".h", ".py", "Makefile")):
                text = open(os.path.join(base, f), errors="ignore").read()
                for needle in ("oracle/", "liboracle", "oracle_", "import oracle", "oracle.h"):
                    assert needle not in text, (os.path.join(base, f), needle)
    text = open(os.path.join(ROOT, "include", "step35a_b200.h")).read() + open(os.path.join(ROOT, "include", "step35a_host.h")).read()
    assert "oracle_" not in text


def test_library_links_no_cpu_solver_fallback(pkg):
    """the product library depends on the CUDA runtime (static) and OpenMP only"""
    import subprocess
    out = subprocess.run(["ldd", pkg.LIB_PATH], capture_output=True, text=True).stdout
    assert "liboracle" not in out and "libtorch" not in out


@pytest.mark.skipif(os.path.exists("/dev/nvidia0"), reason="checks the behaviour on a machine without a GPU")
def test_compute_entry_points_fail_loudly_without_a_gpu(pkg):
    assert pkg.device_count() == 0
    rp = np.array([0, 1, 2], np.int32)
    cl = np.array([0, 1], np.int32)
    with pytest.raises(pkg.S35Error) as e:
        pkg.Matrix(2, 2, rp, cl, np.ones(2))
    assert e.value.status == 2 and "no CPU fallback" in str(e.value)
    with pytest.raises(pkg.S35Error) as e:
        pkg.vec_dot(np.ones(4), np.ones(4))
    assert e.value.status == 2
    params = pkg.default_params(vel_off_diagonals=0)
    disc = pkg.Disc(os.path.join(ROOT, "meshes", "nsbench2.s35m"), 2, params)
    with pytest.raises(pkg.S35Error) as e:
        pkg.Ctx(disc.problem)
    assert e.value.status == 2
