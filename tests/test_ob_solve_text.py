"""openbeam_b200/ob_solve.py: the text report of the reference's CLI (apps/ob-solve/ob-solve_main.cpp:590-638, 877-884,
996-1031).  Layout checks run on the CPU; the end-to-end run of the CLI on a YAML input is a GPU test against the oracle."""
import os
import re

import numpy as np
import pytest

from openbeam_b200 import ob_solve, yaml_loader
from oracle.pyoracle import OracleProblem

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
YAML = os.path.join(ROOT, "examples", "two_bay_frame.yaml")


def test_eigen_default_format():
    # Eigen IO.h, default IOFormat: every coefficient printed like operator<<(double) at precision 6, right-aligned to the
    # widest one of the WHOLE matrix, columns separated by one space, rows by a newline, vectors as columns
    assert ob_solve.eigen_str([1.5, -0.000123456789, 1e-10, 0.0, 123456789.0]) == (
        "         1.5\n-0.000123457\n       1e-10\n           0\n 1.23457e+08")
    assert ob_solve.eigen_str([[1, 2.5], [-3e5, 4]]) == "      1     2.5\n-300000       4"
    assert ob_solve.eigen_str([-0.0]) == "-0"
    assert ob_solve.eigen_str([]) == ""


def test_stress_table_layout():
    es = np.zeros((2, 2, 6))
    es[0, 0] = [1234.5, -2.0, 0, 0, 0, 3.25e-7]
    es[1, 1] = [-1e10, 0, 0, 0, 0, 9.999e5]
    t = ob_solve.stress_table(es).split("\n")
    assert t[2] == "ELEMENT STRESS:"
    assert t[4] == "Element|Face|     N    |     Vy   |     Vz   |     Mx   |     My   |     Mz   |"
    assert t[5] == "-------+----+----------+----------+----------+----------+----------+----------+"
    assert t[6] == "     0 |  0 | 1.234e+03|-2.000e+00| 0.000e+00| 0.000e+00| 0.000e+00| 3.250e-07|"
    assert t[7] == "       |  1 | 0.000e+00| 0.000e+00| 0.000e+00| 0.000e+00| 0.000e+00| 0.000e+00|"
    assert t[9] == "       |  1 |-1.000e+10| 0.000e+00| 0.000e+00| 0.000e+00| 0.000e+00| 9.999e+05|"
    assert len(t) == 11 and t[10] == ""


def _oracle(reference_if_any=True):
    from oracle import pyoracle
    st = yaml_loader.load_structure(YAML)
    return st, OracleProblem(st, reference=reference_if_any and pyoracle.ref_available())


def test_full_k_reads_like_the_reference():
    """ob-solve_main.cpp:600-628 reads K_ff(i1, i2) for i2 >= i1 although K_ff holds its LOWER triangle
    (assemble.cpp:264-270): the free-free off-diagonal of the printed K is zero.  Reproduced, on the oracle's matrices."""
    st, o = _oracle()
    o.assembleProblem()

    class Info:
        pass
    info = Info()
    info.free_dof_indices, info.bounded_dof_indices = o.free_dof_indices, o.bounded_dof_indices
    from oracle import pyoracle
    z = o.sizes()
    info.K_ff, info.K_bf = o.K_ff_dense_lower(), o.K_bf_dense()
    info.K_bb = o.getf(pyoracle.F_KBB).reshape(z["n_b"], z["n_b"]).T      # col-major, upper triangle filled
    K = ob_solve.full_k(info)
    n = len(info.free_dof_indices) + len(info.bounded_dof_indices)
    assert K.shape == (n, n) and np.array_equal(K, K.T)
    fi = np.asarray(info.free_dof_indices, np.int64)
    bi = np.asarray(info.bounded_dof_indices, np.int64)
    assert np.array_equal(K[np.ix_(fi, fi)], np.diag(np.diag(np.asarray(info.K_ff))))
    assert np.array_equal(K[np.ix_(bi, fi)], np.asarray(info.K_bf))
    kbb = np.asarray(info.K_bb)
    assert np.array_equal(np.triu(K[np.ix_(bi, bi)]), np.triu(kbb))


def _numbers(block):
    return np.array([float(x) for x in re.findall(r"[-+]?\d[\d.]*(?:e[-+]?\d+)?", block)])


@pytest.mark.gpu
def test_cli_report_against_oracle(tmp_path, capsys):
    out = str(tmp_path / "report.txt")
    assert ob_solve.main(["-i", YAML, "--show-U", "--stress", "--full-k", "-o", out]) == 0
    text = open(out).read()
    st, o = _oracle()
    o.solveStatic()
    o.postProcCalcStress()
    # section order and headers as in the reference's text mode
    i_k, i_r, i_s = text.index("Global stiffness matrix(K):\nK:\n"), text.index("\n\nRESULTS:\n------------------------\n"), \
        text.index("\n\nELEMENT STRESS:\n\n")
    assert i_k == 0 and i_k < i_r < i_s
    res = text[i_r:i_s]
    u_txt = res[res.index("Displacements (U_f):\n") + 21:res.index("Reactions (F_b):\n")]
    f_txt = res[res.index("Reactions (F_b):\n") + 17:]
    U, Fb = _numbers(u_txt), _numbers(f_txt)
    assert len(U) == len(o.U_f) and len(Fb) == len(o.F_b)
    # printed at 6 significant digits
    assert np.allclose(U, o.U_f, rtol=1e-5, atol=1e-5 * np.abs(o.U_f).max())
    assert np.allclose(Fb, o.F_b, rtol=1e-5, atol=1e-5 * np.abs(o.F_b).max())
    # the layout itself: the same formatter on the oracle's numbers gives the same lines wherever rounding agrees
    same = sum(a == b for a, b in zip(u_txt.strip("\n").split("\n"), ob_solve.eigen_str(o.U_f).split("\n")))
    assert same >= len(U) - 1
    rows = [l for l in text[i_s:].split("\n") if re.match(r"^( +\d+ |       )\| +[01] \|", l)]
    assert len(rows) == 2 * st.n_elems
    S = np.array([[float(x) for x in l.split("|")[2:8]] for l in rows]).reshape(st.n_elems, 2, 6)
    ref = np.asarray(o.element_stress).reshape(st.n_elems, 2, 6)
    assert np.allclose(S, ref, rtol=2e-3, atol=2e-3 * np.abs(ref).max())   # %10.3e
    # nothing on stdout when -o is given; an unknown solver name is refused with the reference's message
    assert capsys.readouterr().out == ""
    with pytest.raises(SystemExit) as e:
        ob_solve.main(["-i", YAML, "--solver", "Dense_QR"])
    assert "Invalid --solver algorithm name." in str(e.value)
