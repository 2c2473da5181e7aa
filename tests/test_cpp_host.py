"""The C++ host mirror (include/obhost.hpp) over the C ABI: compiles examples/ob_solve_demo3.cpp with g++ and runs
it — symbolic-only on CPU, the full solve on the GPU box against SURVEY.md §4's closed-form-checked numbers."""
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def build(tmp_path):
    exe = str(tmp_path / "ob_solve_demo3")
    libdir = os.path.join(ROOT, "openbeam_b200")
    subprocess.check_call(["g++", "-std=c++17", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "examples", "ob_solve_demo3.cpp"), "-L" + libdir, "-lobgpu",
                           "-Wl,-rpath," + libdir, "-o", exe])
    return exe


def test_cpp_mirror_symbolic(tmp_path):
    out = subprocess.check_output([build(tmp_path), "-1"], text=True)
    assert "nDOFs 12" in out
    dofs = re.findall(r"dof (\d+) node (\d+) id (\d+)", out)
    assert [(int(n), int(d)) for _, n, d in dofs] == [(n, d) for n in range(4) for d in (0, 1, 5)]


@pytest.mark.gpu
def test_cpp_mirror_solve(tmp_path):
    out = subprocess.check_output([build(tmp_path), "0"], text=True)
    U = np.array([float(x) for x in re.findall(r"(?m)^U_f\[\d+\] (\S+)", out)])
    F = np.array([float(x) for x in re.findall(r"(?m)^F_b\[\d+\] (\S+)", out)])
    assert np.allclose(U, [2.6013078134695e-5, -6.5563909774436e-5, -1.29984937854503e-3,
                           -2.6013078134695e-5, -6.5563909774436e-5, 1.29984937854503e-3], rtol=1e-10)
    assert np.allclose(F, [7784.41363180755, 19620, -5136.62044771133, -7784.41363180755, 19620, 5136.62044771133],
                       rtol=1e-10)
    assert re.search(r"stress e1 f0 N -7784\.41", out)
    # mesh() through the C++ mirror: original nodes keep their ids and (Euler-Bernoulli, consistent loads) their
    # displacements
    assert "meshed nodes 17 elements 16 pieces_of_beam 8" in out
    Um = np.array([float(x) for x in re.findall(r"meshed U_f\[\d+\] (\S+)", out)])
    assert np.allclose(Um, U, rtol=1e-9)
    # nonLinearIterative through the C++ mirror against the same loop around the oracle
    import dataclasses
    import fixtures
    from oracle.pyoracle import OracleProblem
    st = fixtures.demo3_frame()
    xy, rz = st.node_xy.copy(), np.zeros(st.n_nodes)
    for _ in range(40):
        o = OracleProblem(dataclasses.replace(st, node_xy=xy.copy(), node_rotz=rz.copy())).solveStatic()
        for (node, dof), u in zip(o.problem_DoFs[np.asarray(o.free_dof_indices, np.int64)], np.asarray(o.U_f)):
            if dof < 3:
                xy[node, dof] += u
            elif dof == 5:
                rz[node] += u
    Ui = np.array([float(x) for x in re.findall(r"iterative U_f\[\d+\] (\S+)", out)])
    Fi = np.array([float(x) for x in re.findall(r"iterative F_b\[\d+\] (\S+)", out)])
    assert np.allclose(Ui, o.U_f, rtol=1e-9, atol=0) and np.allclose(Fi, o.F_b, rtol=1e-9, atol=1e-6)
    assert not np.allclose(Ui, U, rtol=1e-6)
    Ua = np.array([float(x) for x in re.findall(r"again U_f\[\d+\] (\S+)", out)])
    assert np.allclose(Ua, U, rtol=1e-12)


def test_cpp_batch_pipeline_builds(tmp_path):
    """include/obpipeline.hpp + its example compile warning-free and link against libobgpu.so (no GPU needed)"""
    exe = str(tmp_path / "ob_pipeline_frames")
    libdir = os.path.join(ROOT, "openbeam_b200")
    subprocess.check_call(["g++", "-std=c++17", "-Wall", "-Wextra", "-Werror", "-pthread", "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "examples", "ob_pipeline_frames.cpp"), "-L" + libdir, "-lobgpu",
                           "-Wl,-rpath," + libdir, "-o", exe])
    assert os.path.exists(exe)


@pytest.mark.gpu
def test_cpp_batch_pipeline(tmp_path):
    """include/obpipeline.hpp (the C++ twin of openbeam_b200/pipeline.py): sub-batches of a parametric sweep dealt to three
    contexts on host threads return the bits a single context returns, and the reactions balance the applied loads."""
    exe = str(tmp_path / "ob_pipeline_frames")
    libdir = os.path.join(ROOT, "openbeam_b200")
    subprocess.check_call(["g++", "-std=c++17", "-Wall", "-Werror", "-pthread", "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "examples", "ob_pipeline_frames.cpp"), "-L" + libdir, "-lobgpu",
                           "-Wl,-rpath," + libdir, "-o", exe])
    out = subprocess.check_output([exe, "0", "5", "2000", "3"], text=True)
    assert "structures 10000 sub_batches 5 workers 3" in out
    assert "pipeline_matches_single_context 1" in out
    assert float(re.search(r"vertical_equilibrium_relerr (\S+)", out).group(1)) < 1e-9
    assert 1e-5 < float(re.search(r"max_abs_U_f (\S+)", out).group(1)) < 1.0
