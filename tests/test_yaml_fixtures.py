"""The hand-restated fixtures of tests/fixtures.py against the reference's own example files, read with
openbeam_b200.yaml_loader (pyyaml + an arithmetic evaluator over `parameters`).  Runs where /root/reference is present
(this container); the GPU box has no /root/reference and uses the restated fixtures, which this test proves equal."""
import glob
import os

import numpy as np
import pytest

import fixtures
from openbeam_b200 import yaml_loader
from oracle import pyoracle
from oracle.pyoracle import OracleProblem

EX = "/root/reference/examples-structures"
pytestmark = pytest.mark.skipif(not os.path.isdir(EX), reason="/root/reference is not on this machine")

PAIRS = {
    "demo1": "demo1.yaml", "demo1_tilded_plane": "demo1_tilded_plane.yaml", "demo2": "demo2.yaml",
    "demo2_bis": "demo2_bis.yaml", "demo3_frame": "demo3_frame.yaml", "demo3_frame_slider": "demo3_frame_slider.yaml",
    "demo3_frame_tilted_slider": "demo3_frame_tilted_slider.yaml", "ex2021_01": "ex2021_01.yaml",
    "old_demo3": "old_demo3.yaml",
}
FIELDS = ("node_xy", "elem_nodes", "elem_type", "elem_EAI", "con_node", "con_dof", "con_val", "nload_node", "nload_dof",
          "nload_val", "eload_elem", "eload_q", "eload_dir")


def test_evaluator():
    p = {"G": 9.81, "L": 4.0}
    assert yaml_loader.evaluate("1000*G", p) == 9810.0
    assert yaml_loader.evaluate("-(L/2)^2 + 3", p) == -1.0
    assert yaml_loader.evaluate(2.5e-3, p) == 2.5e-3
    with pytest.raises(ValueError):
        yaml_loader.evaluate("__import__('os')", p)
    with pytest.raises(ValueError):
        yaml_loader.evaluate("Q + 1", p)


@pytest.mark.parametrize("name", sorted(PAIRS))
def test_restated_fixture_equals_yaml(name):
    y = yaml_loader.load_structure(os.path.join(EX, PAIRS[name]))
    f = {**fixtures.ALL, **fixtures.ILL_CONDITIONED}[name]()
    for k in FIELDS:
        assert np.array_equal(np.asarray(getattr(y, k)), np.asarray(getattr(f, k))), (name, k)
    ry = np.zeros(y.n_nodes) if y.node_rotz is None else y.node_rotz
    rf = np.zeros(f.n_nodes) if f.node_rotz is None else f.node_rotz
    assert np.array_equal(ry, rf)


def test_every_example_file_loads_and_solves():
    """all YAML examples of the reference go through the loader and the oracle (the legacy *.txt files use the pre-YAML
    syntax the reference itself no longer parses, parse_file.cpp:36-48)"""
    files = sorted(glob.glob(os.path.join(EX, "*.yaml")))
    assert len(files) >= 9
    for path in files:
        st = yaml_loader.load_structure(path)
        o = OracleProblem(st).solveStatic().postProcCalcStress()
        assert np.all(np.isfinite(o.U_f)) and np.all(np.isfinite(o.F_b))
        if pyoracle.ref_available():
            r = OracleProblem(st, reference=True).solveStatic().postProcCalcStress()
            assert np.array_equal(o.free_dof_indices, r.free_dof_indices)
            den = np.abs(r.U_f).max()
            # old_demo3.yaml ties two coincident nodes with 1e20 penalty springs: cond(K_ff) ~ 1e16, two Cholesky orders
            # of the same matrix agree to ~1e-2 only (SURVEY.md 7 "hard parts"); its integer artefacts are still exact
            tol = 5e-2 if path.endswith("old_demo3.yaml") else 1e-9
            assert np.abs(o.U_f - r.U_f).max() <= tol * den, path
