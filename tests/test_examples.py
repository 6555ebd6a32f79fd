"""The example scripts (drop-in versions of the reference's examples/performance/mlups_3d.py and
examples/cfd/flow_past_sphere_3d.py): they compile on CPU and run end to end on a GPU."""

import os
import py_compile
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXAMPLES = ["mlups_3d.py", "flow_past_sphere_3d.py"]


@pytest.mark.parametrize("name", EXAMPLES)
def test_examples_compile(name):
    py_compile.compile(os.path.join(ROOT, "examples", name), doraise=True)


@pytest.mark.gpu
def test_mlups_example_runs():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "examples", "mlups_3d.py"), "32", "20", "--omega", "1.7"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "MLUPS:" in r.stdout


@pytest.mark.gpu
def test_sphere_example_reports_positive_drag():
    r = subprocess.run(
        [sys.executable, os.path.join(ROOT, "examples", "flow_past_sphere_3d.py"), "--shape", "96", "48", "48", "--steps", "300", "--interval", "150"],
        capture_output=True, text=True, timeout=600,
    )
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    last = [ln for ln in r.stdout.splitlines() if ln.startswith("step ")][-1]
    cd = float(last.split("CD=")[1].split()[0])
    assert 0.0 < cd < 50.0, last  # the flow pushes the sphere downstream
