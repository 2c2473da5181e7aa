"""CPU-side checks of libobgpu: the library loads, exports every symbol include/obgpu.h declares, and its
integer symbolic phase (K3: DOF numbering, free/bounded partition, K pattern, element->nnz map) equals the
oracle's bit for bit.  No compute entry point is called here (there is no GPU in this container); compute
calls on a symbolic-only ctx must fail with OBGPU_ECUDA — there is no CPU fallback."""
import json
import os
import re

import numpy as np
import pytest

import fixtures
from openbeam_b200 import obgpu, workloads
from openbeam_b200.problem import CStructureProblem, StaticSolveProblemInfo
from oracle.pyoracle import OracleProblem

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "obgpu.h")).read()
    declared = set(re.findall(r"\b(obgpu_[a-z_]+)\s*\(", hdr))
    assert declared == set(obgpu.ABI_SYMBOLS)
    L = obgpu.lib()
    for name in declared:
        assert hasattr(L, name), name


def test_status_strings():
    L = obgpu.lib()
    assert L.obgpu_status_string(obgpu.OK) == b"ok"
    # solver_static.cpp:102-104
    assert L.obgpu_status_string(obgpu.ESINGULAR) == \
        b"[CFiniteElementProblem::solveStatic] Ill-conditioned matrix. Too few restrictions?"


def test_no_cpu_fallback():
    ctx = obgpu.Context(-1)
    ctx.upload(fixtures.demo3_frame())
    ctx.symbolic()
    for call in (ctx.solve_static, ctx.assemble, ctx.synchronize):
        with pytest.raises(obgpu.ObgpuError) as e:
            call()
        assert e.value.status == obgpu.ECUDA
    with pytest.raises(RuntimeError):
        CStructureProblem(fixtures.demo1(), ctx=obgpu.Context(-1)).solveStatic(StaticSolveProblemInfo())


def _check_symbolic(st, ctx=None):
    ctx = ctx or obgpu.Context(-1)
    ctx.upload(st)
    ctx.symbolic()
    o = OracleProblem(st).assembleProblem(skip_subK=True)
    z, oz = ctx.sizes(), o.sizes()
    assert (z["Ntot"], z["Etot"], z["n_tot"], z["nf_tot"], z["nb_tot"], z["T_tot"], z["nnz_tot"]) == \
        (oz["N"], oz["E"], oz["n"], oz["n_f"], oz["n_b"], oz["T"], oz["nnz"])
    dofs = o.problem_DoFs
    assert np.array_equal(ctx.symbolic_array(obgpu.SYM_DOF_NODE), dofs[:, 0])
    assert np.array_equal(ctx.symbolic_array(obgpu.SYM_DOF_ID), dofs[:, 1])
    assert np.array_equal(ctx.symbolic_array(obgpu.SYM_FREE), o.free_dof_indices)
    assert np.array_equal(ctx.symbolic_array(obgpu.SYM_BOUNDED), o.bounded_dof_indices)
    from oracle import pyoracle
    assert np.array_equal(ctx.symbolic_array(obgpu.SYM_DOF_INDEX), o.geti(pyoracle.I_DOF_INDEX))
    outer, inner, _ = o.K_csc_upper()
    assert np.array_equal(ctx.symbolic_array(obgpu.SYM_K_ROWPTR), outer)
    assert np.array_equal(ctx.symbolic_array(obgpu.SYM_K_COL), inner)
    # element->nnz map: triplet t lands in the slot whose (row, col) are the triplet's
    tr, tc = o.geti(pyoracle.I_TRI_ROW), o.geti(pyoracle.I_TRI_COL)
    slot = ctx.symbolic_array(obgpu.SYM_TRI_NNZ)
    col_of_slot = np.repeat(np.arange(oz["n"]), np.diff(outer))
    assert np.array_equal(inner[slot], tr) and np.array_equal(col_of_slot[slot], tc)
    return ctx, o


@pytest.mark.parametrize("name", list(fixtures.ALL))
def test_symbolic_matches_oracle(name):
    ctx, _ = _check_symbolic(fixtures.ALL[name]())
    g = json.load(open(os.path.join(ROOT, "tests", "golden", name + ".json")))
    assert ctx.symbolic_array(obgpu.SYM_FREE).tolist() == g["free_dof_indices"]
    assert ctx.symbolic_array(obgpu.SYM_BOUNDED).tolist() == g["bounded_dof_indices"]
    assert ctx.symbolic_array(obgpu.SYM_K_ROWPTR).tolist() == g["K_outer"]
    assert ctx.symbolic_array(obgpu.SYM_K_COL).tolist() == g["K_inner"]


@pytest.mark.parametrize("index", [0, 1, 7, 123, 4567, 99999])
def test_symbolic_random_frames(index):
    _check_symbolic(workloads.c3_frame(index))


def test_symbolic_grid_and_batch():
    _check_symbolic(workloads.grid_frame(7, 5))
    _check_symbolic(workloads.grid_frame(6, 4, etype=3))
    # a batch of different structures: per-structure arrays are the concatenation of the single-structure ones
    sts = [workloads.c3_frame(i) for i in range(20)] + [fixtures.demo1(), fixtures.inclined_mixed()]
    ctx = obgpu.Context(-1)
    ctx.upload(sts)
    ctx.symbolic()
    fo, no = ctx.offsets(obgpu.OFF_FREE), ctx.offsets(obgpu.OFF_NNZ)
    free, col = ctx.symbolic_array(obgpu.SYM_FREE), ctx.symbolic_array(obgpu.SYM_K_COL)
    for s, st in enumerate(sts):
        o = OracleProblem(st).assembleProblem(skip_subK=True)
        assert np.array_equal(free[fo[s]:fo[s + 1]], o.free_dof_indices)
        assert np.array_equal(col[no[s]:no[s + 1]], o.K_csc_upper()[1])


def test_invalid_inputs_rejected():
    ctx = obgpu.Context(-1)
    st = fixtures.demo1()
    st.elem_nodes = st.elem_nodes.copy()
    st.elem_nodes[0, 1] = 99
    ctx.upload(st)
    with pytest.raises(obgpu.ObgpuError) as e:
        ctx.symbolic()
    assert e.value.status == obgpu.EINVAL
    st = fixtures.demo1()
    st.node_xy = st.node_xy.copy()
    st.node_xy[1] = st.node_xy[0]  # zero-length element: ASSERT_(L > 0)
    ctx.upload(st)
    with pytest.raises(obgpu.ObgpuError):
        ctx.symbolic()
