"""CStructureProblem::mesh(): the C-ABI generator `obgpu_mesh_create` against the oracle's
restatement (bit-exact: integer and coordinate work), the reference's piece-type and load-splitting rules including its
quirks (QM1, QM2), and physics invariants on the oracle solver; on the GPU, the meshed structures through the path."""
import numpy as np
import pytest

import fixtures
from openbeam_b200 import obgpu, workloads
from openbeam_b200.structure import build_structure
from oracle import pymesh
from oracle.pyoracle import OracleProblem

FIELDS = ("node_xy", "elem_nodes", "elem_type", "elem_EAI", "con_node", "con_dof", "con_val", "nload_node", "nload_dof",
          "nload_val", "eload_elem", "eload_q", "eload_dir")
E_STEEL = 2.1e11
SEC = {"S": (E_STEEL, 28.5e-4, 1940e-8)}


def assert_same_mesh(st, max_len):
    got, info = obgpu.mesh_structure(st, max_len)
    ref, e2e, e2n = pymesh.mesh(st, max_len)
    for k in FIELDS:
        a, b = getattr(got, k), ref[k]
        assert a.shape == b.shape and np.array_equal(a, b), k
    kinds = got.eload_type if got.eload_type is not None else np.zeros(len(got.eload_elem), np.uint8)
    q2 = got.eload_q2 if got.eload_q2 is not None else np.zeros(len(got.eload_elem))
    assert np.array_equal(kinds, ref["eload_type"])
    if ref["eload_type"].any():
        assert np.array_equal(q2, ref["eload_q2"])
    rz = got.node_rotz if got.node_rotz is not None else np.zeros(len(got.node_xy))
    assert np.array_equal(rz, ref["node_rotz"])
    assert info["element2elements"] == e2e and info["element2nodes"] == e2n
    return got, info


@pytest.mark.parametrize("name", list(fixtures.ALL))
@pytest.mark.parametrize("max_len", [10.0, 0.75, 0.1])
def test_mesh_matches_oracle_on_fixtures(name, max_len):
    assert_same_mesh(fixtures.ALL[name](), max_len)


def test_piece_types_and_reference_quirks():
    """CBaseElementBeam.cpp:374-420.  QM1: an un-subdivided RA / RD beam becomes RR and AA becomes AR."""
    nodes = [(0, 0), (3, 0), (6, 0), (9, 0), (12, 0), (12, 1)]
    els = [("BEAM2D_RR", 0, 1, "S"), ("BEAM2D_RA", 1, 2, "S"), ("BEAM2D_AR", 2, 3, "S"), ("BEAM2D_AA", 3, 4, "S"),
           ("BEAM2D_RD", 4, 5, "S")]
    st = build_structure(nodes, els, SEC, [(0, "DXDYRZ"), (5, "DXDYRZ")])
    m, info = assert_same_mesh(st, 1.0)
    t = m.elem_type
    RR, RA, AR, AA, RD = 0, 1, 2, 3, 4
    assert t[info["element2elements"][0]].tolist() == [RR, RR, RR]
    assert t[info["element2elements"][1]].tolist() == [RR, RR, RA]
    assert t[info["element2elements"][2]].tolist() == [AR, RR, RR]
    assert t[info["element2elements"][3]].tolist() == [AR, RR, RA]
    assert t[info["element2elements"][4]].tolist() == [RR]            # QM1: single piece, RD lost
    m1, _ = assert_same_mesh(st, 100.0)
    assert m1.elem_type.tolist() == [RR, RR, AR, AR, RR]              # QM1 on every beam
    # intermediate nodes: evenly spaced, appended after the original ones, element by element
    assert m.node_xy[:6].tolist() == st.node_xy.tolist()
    assert np.allclose(m.node_xy[6:8], [[1, 0], [2, 0]]) and info["element2nodes"][0] == [0, 6, 7, 1]


def test_load_splitting_rules():
    """meshLoad of the four load classes; QM2: the concentrated load lands on meshed element `piece index`."""
    nodes = [(0, 0), (4, 0), (8, 0)]
    els = [("BEAM2D_RR", 0, 1, "S"), ("BEAM2D_RR", 1, 2, "S")]
    loads = [(1, "TRIANGULAR", 1000.0, 5000.0, (0.0, -1.0, 0.0)), (1, "CONCENTRATED", 700.0, 2.5, (0.0, -1.0, 0.0)),
             (0, 2000.0, (0.0, -1.0, 0.0)), (0, "TEMPERATURE", 30.0)]
    st = build_structure(nodes, els, SEC, [(0, "DXDYRZ"), (2, "DXDYRZ")], [], loads)
    m, info = assert_same_mesh(st, 1.0)
    assert info["element2elements"] == [[0, 1, 2, 3], [4, 5, 6, 7]]
    rows = list(zip(m.eload_elem.tolist(), m.eload_type.tolist(), m.eload_q.tolist(), m.eload_q2.tolist()))
    # multimap order: the two loads of beam 0 first (uniform, temperature: copied to every piece) ...
    assert rows[:4] == [(k, 0, 2000.0, 0.0) for k in range(4)]
    assert [(r[0], r[1], r[2]) for r in rows[4:8]] == [(k, 3, 30.0) for k in range(4)]
    # ... then beam 1: trapezoid slices 1000..5000 in four steps of 1000
    assert rows[8:12] == [(4 + k, 1, 1000.0 + 1000.0 * k, 2000.0 + 1000.0 * k) for k in range(4)]
    # ... and its concentrated load at 2.5 m: piece 2 of beam 1, 0.5 m into it — attached to meshed element 2 (QM2)
    assert rows[12] == (2, 2, 700.0, 0.5)


def test_meshing_does_not_change_the_solution_at_original_nodes():
    """Euler-Bernoulli beams with consistent nodal loads are exact at the nodes: displacements of the original nodes and
    the reactions are unchanged by subdivision (RR frame, uniform + nodal loads)."""
    st = workloads.demo3_frame()
    st.nload_val[:] = [3000.0, -2000.0, 500.0, -1000.0, 4000.0, -800.0]
    o0 = OracleProblem(st).solveStatic()
    for max_len in (1.0, 0.25):
        m, _ = obgpu.mesh_structure(st, max_len)
        o1 = OracleProblem(m).solveStatic()
        dofs1 = np.asarray(o1.problem_DoFs)
        U1 = np.zeros(len(dofs1))
        U1[np.asarray(o1.free_dof_indices, int)] = o1.U_f
        U0 = np.zeros(len(o0.problem_DoFs))
        U0[np.asarray(o0.free_dof_indices, int)] = o0.U_f
        for k, (node, dof) in enumerate(np.asarray(o0.problem_DoFs)):
            j = np.nonzero((dofs1[:, 0] == node) & (dofs1[:, 1] == dof))[0][0]
            assert abs(U1[j] - U0[k]) <= 1e-9 * np.abs(U0).max()
        assert np.allclose(np.sort(o1.F_b), np.sort(o0.F_b), rtol=1e-9, atol=1e-9 * np.abs(o0.F_b).max())


def test_mesh_rejects_bad_input():
    st = workloads.demo3_frame()
    with pytest.raises(obgpu.ObgpuError):
        obgpu.mesh_structure(st, 0.0)


def tri_temp_frame():
    """Portal frame with a triangular load on the first column, temperature on the beam, UDL on the beam, a concentrated
    load on element 0 (the only place where the reference's meshLoad puts it on the right piece, QM2)."""
    nodes = [(0, 0), (0, 3), (5, 3), (5, 0)]
    els = [("BEAM2D_RR", 0, 1, "S"), ("BEAM2D_RA", 1, 2, "S"), ("BEAM2D_RR", 2, 3, "S")]
    loads = [(0, "TRIANGULAR", 500.0, 2500.0, (1.0, 0.0, 0.0)), (1, "TEMPERATURE", 25.0), (1, 4000.0, (0.0, -1.0, 0.0)),
             (0, "CONCENTRATED", 900.0, 1.73, (1.0, 0.0, 0.0))]
    return build_structure(nodes, els, SEC, [(0, "DXDYRZ"), (3, "DXDY")], [(1, "DX", 1500.0)], loads)


MESH_CASES = dict(fixtures.ALL, tri_temp_frame=tri_temp_frame)


@pytest.mark.gpu
@pytest.mark.parametrize("name,max_len", [("demo3_frame", 0.5), ("demo3_frame", 0.10), ("tri_temp_frame", 0.1),
                                           ("inclined_mixed", 0.2), ("demo3_frame_tilted_slider", 0.1),
                                           ("demo3_frame_slider", 0.25)])
def test_meshed_structures_through_the_gpu_path(name, max_len):
    """The generator's product through upload -> symbolic -> solve -> stress against the oracle on the same meshed
    structure (1e-9).  mesh() numbers the intermediate nodes after all original ones, so the band of K_ff in the
    reference's numbering grows with the subdivision; the batched-LDLt route factors in a reverse Cuthill-McKee order
    (capi_general.cu) and returns U_f / F_f in the reference's numbering."""
    from openbeam_b200.problem import CStructureProblem, MeshParams, StaticSolveProblemInfo, StressInfo
    p0 = CStructureProblem(MESH_CASES[name]())
    p, mo = p0.mesh(MeshParams(max_len))
    assert mo.num_original_elements == p0.getNumberOfElements()
    info, sinfo = StaticSolveProblemInfo(), StressInfo()
    p.solveStatic(info)
    p.postProcCalcStress(sinfo, info)
    o = OracleProblem(p.structure).solveStatic().postProcCalcStress()
    rel = lambda a, b: float(np.max(np.abs(np.asarray(a).ravel() - np.asarray(b).ravel())) / max(np.max(np.abs(b)), 1e-300))
    assert np.array_equal(info.build_info.free_dof_indices, o.free_dof_indices)
    assert rel(info.U_f, o.U_f) <= 1e-9 and rel(info.F_b, o.F_b) <= 1e-9
    assert rel(sinfo.element_stress, o.element_stress) <= 1e-9
    p.ctx.close()


@pytest.mark.gpu
def test_web_default_mesh_size_through_pcg():
    """demo3_frame meshed at the web app's 0.10 m (80 elements) through the sparse route as well; agreement with the
    oracle's direct solve is limited by conditioning there."""
    st, _ = obgpu.mesh_structure(workloads.demo3_frame(), 0.10)
    assert st.n_elems == 80 and st.n_nodes == 81
    ctx = obgpu.Context(0)
    try:
        ctx.upload(st)
        ctx.solve_static(method=obgpu.PCG_JACOBI, rtol=1e-13, max_iter=200000)
        U = ctx.download(obgpu.RES_UF)[0]
        o = OracleProblem(st).solveStatic()
        assert np.max(np.abs(U - o.U_f)) <= 1e-6 * np.max(np.abs(o.U_f))
    finally:
        ctx.close()
