"""tools/compare_fields.py: the numpy restatement of the reference's scripts/compare_hdf.py and of the solver-test
tolerance rule (solverTests/testAll.py:23-36) over .npz dumps (h5py is absent on this image)."""
import io
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import compare_fields as cf


def _fields(seed=0):
    rng = np.random.default_rng(seed)
    return {"rho": rng.random((4, 5, 6)) + 1.0, "velocity": rng.standard_normal((4, 5, 6, 3)), "temperature": rng.random((4, 5, 6)) + 300.0}


def test_lines_and_locations_like_compare_hdf():
    a = _fields(); b = {k: v.copy() for k, v in a.items()}
    b["velocity"][2, 3, 1, 2] += 0.5
    out = io.StringIO()
    res = cf.compare(a, b, out)
    assert res["rho"] == (0.0, (0, 0, 0))
    assert res["velocity"][1] == (2, 3, 1, 2) and abs(res["velocity"][0] - 0.5) < 1e-15
    lines = out.getvalue().splitlines()
    assert lines[0] == "rho: max diff = 0.0 at (0, 0, 0)"      # the reference's format
    assert lines[1].startswith("velocity: max diff = 0.5 at (2, 3, 1, 2)")


def test_solver_tolerance_rule():
    ref = _fields(1)
    ok = {k: v*(1.0 + 5e-11) for k, v in ref.items()}           # inside 1e-10 relative
    assert cf.check_solution(ok, ref, out=io.StringIO())
    small = {k: v.copy() for k, v in ref.items()}
    ref2 = {k: v.copy() for k, v in ref.items()}
    ref2["velocity"][0, 0, 0, 0] = 0.0; small["velocity"][0, 0, 0, 0] = 9e-13      # the absolute floor 1e-12 near zero
    assert cf.check_solution(small, ref2, out=io.StringIO())
    bad = {k: v.copy() for k, v in ref.items()}
    bad["temperature"][1, 2, 3] *= 1.0 + 3e-10
    out = io.StringIO()
    assert not cf.check_solution(bad, ref, out=out)
    assert out.getvalue().startswith("Error on field: temperature, delta = ")


def test_command_line_on_npz(tmp_path):
    a = _fields(2); b = {k: v.copy() for k, v in a.items()}
    b["rho"][0, 0, 0] += 1e-3
    pa, pb = str(tmp_path/"a.npz"), str(tmp_path/"b.npz")
    np.savez(pa, **a); np.savez(pb, **b)
    tool = os.path.join(ROOT, "tools", "compare_fields.py")
    r = subprocess.run([sys.executable, tool, pa, pb], capture_output=True, text=True)
    assert r.returncode == 0 and "rho: max diff = " in r.stdout
    r = subprocess.run([sys.executable, tool, pa, pb, "--solver-tol"], capture_output=True, text=True)
    assert r.returncode == 1 and "Error on field: rho" in r.stdout
    r = subprocess.run([sys.executable, tool, pa, pa, "--solver-tol"], capture_output=True, text=True)
    assert r.returncode == 0
