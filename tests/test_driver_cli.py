"""The step-35a program (the reference's main(), step-35a.cc:1712-1760) on a machine without a GPU: it reads the
parameter file and the mesh, prints the reference's progress lines, and -- there being no CPU fallback -- stops
with the reference's exception banner and exit code 1 when the device context cannot be created."""
import os
import subprocess

import pytest

from conftest import MESH_DIR, ROOT

BINARY = os.path.join(ROOT, "dealii-step-35a_b200", "step-35a")

PRM = """
set Method_Form = rotational
subsection Physical data
  set final_time = 0.005
  set Reynolds   = 100
end
subsection Space discretization
  set n_of_refines = 1
  set pressure_fe_degree = 1
end
subsection Data solve velocity
  set eps = 1e-8
  set off_diagonals = 0
  set diag_strength = 0.0
  set update_prec = 100
end
set output_interval = 999
set verbose = false
"""


@pytest.mark.skipif(os.path.exists("/dev/nvidia0"), reason="checks the behaviour on a machine without a GPU")
def test_program_fails_loudly_without_a_gpu(pkg, tmp_path):
    assert os.path.exists(BINARY)
    (tmp_path / "parameter-file.prm").write_text(PRM)
    out = subprocess.run([BINARY, "2", os.path.join(MESH_DIR, "nsbench2.s35m")], cwd=tmp_path, capture_output=True, text=True,
                         timeout=120)
    assert out.returncode == 1
    assert "Number of refines = 1" in out.stdout and "Number of active cells: 240" in out.stdout
    assert "width: 4.1" in out.stdout
    assert "Exception on processing" in out.stderr and "no CPU fallback" in out.stderr and "Aborting!" in out.stderr


def test_program_reports_a_missing_parameter_file(pkg, tmp_path):
    out = subprocess.run([BINARY, "2", os.path.join(MESH_DIR, "nsbench2.s35m")], cwd=tmp_path, capture_output=True, text=True,
                         timeout=60)
    assert out.returncode == 1 and "parameter-file.prm" in out.stderr


def test_program_rejects_a_time_step_out_of_range(pkg, tmp_path):
    (tmp_path / "parameter-file.prm").write_text(PRM.replace("set final_time = 0.005", "set final_time = 0.0005"))
    out = subprocess.run([BINARY, "2", os.path.join(MESH_DIR, "nsbench2.s35m")], cwd=tmp_path, capture_output=True, text=True,
                         timeout=60)
    # AssertThrow(!((dt <= 0.) || (dt > .5*T))) at step-35a.cc:731
    assert out.returncode == 1 and "is out of range" in out.stderr
