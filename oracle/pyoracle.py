"""ctypes wrapper over oracle/libob_oracle.so — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module (the product package openbeam_b200 never does).  PARITY UNPINNED:
see oracle/ob_oracle.h.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
_REF = None
REF_SOURCES = "/root/reference/libopenbeam"   # only needed to (re)build oracle/_ref; the built .so travels

I_DOF_NODE, I_DOF_ID, I_FREE, I_BOUNDED, I_DOF_INDEX, I_TRI_ROW, I_TRI_COL, I_K_OUTER, I_K_INNER = range(9)
(F_TRI_VAL, F_K_VAL, F_KFF, F_KBF, F_KBB, F_UB, F_FF, F_UF, F_FB, F_STRESS, F_ELEM_R, F_ELEM_KGLOBAL) = range(12)


class OracleError(RuntimeError):
    pass


def build(force=False):
    so = os.path.join(_HERE, "libob_oracle.so")
    srcs = [os.path.join(_HERE, f) for f in
            ("ob_oracle.cpp", "ob_assemble.ipp", "ob_solve.ipp", "ob_capi.ipp", "ob_batch.ipp", "ob_oracle.h")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s", "libob_oracle.so"])
    return so


def build_ref(force=False):
    """oracle/_ref/libob_ref.so: the reference's own unmodified hot-path sources compiled against oracle/ref_shim/ (see
    Makefile.ref).  Built only where /root/reference exists (this container); elsewhere the prebuilt file is used."""
    so = os.path.join(_HERE, "_ref", "libob_ref.so")
    if os.path.isdir(REF_SOURCES):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-j8", "-f", "Makefile.ref"] + (["-B"] if force else []))
    return so if os.path.exists(so) else None


def ref_available():
    return build_ref() is not None


def _bind(L):
    vp, i64, dp = C.c_void_p, C.c_int64, C.c_void_p
    L.obo_create.restype = vp
    L.obo_create.argtypes = [i64, dp, dp, i64, dp, dp, dp, i64, dp, dp, dp, i64, dp, dp, dp, i64, dp, dp, dp]
    L.obo_destroy.argtypes = [vp]
    L.obo_set_load_values.argtypes = [vp, dp, dp]
    L.obo_set_eload_kinds.argtypes = [vp, dp, dp]
    L.obo_set_elem_EAI.argtypes = [vp, dp]
    L.obo_assemble.argtypes = [vp, C.c_int]
    L.obo_solve_static.argtypes = [vp, C.c_int]
    L.obo_stress.argtypes = [vp]
    L.obo_last_error.argtypes = [vp]
    L.obo_last_error.restype = C.c_char_p
    L.obo_get_sizes.argtypes = [vp, dp]
    L.obo_get_i64.argtypes = [vp, C.c_int, dp, i64]
    L.obo_get_i64.restype = i64
    L.obo_get_f64.argtypes = [vp, C.c_int, dp, i64]
    L.obo_get_f64.restype = i64
    L.obo_run_cases.argtypes = [vp, i64, dp, dp, dp, C.c_int, dp, dp, dp]
    L.obo_run_cases.restype = C.c_double
    L.obo_run_batch.argtypes = [i64] + [dp] * 18 + [C.c_int, dp]
    L.obo_run_batch.restype = C.c_double
    return L


def run_batch(batch, n_threads=1, reference=False, want_summaries=False):
    """CPU baseline on a packed batch of different structures (obgpu.pack_batch / workloads.c3_batch dict): one problem
    set-up + solveStatic + postProcCalcStress per structure.  Returns (seconds, summaries [S,2] or None)."""
    L = ref_lib() if reference else lib()
    S = batch["S"]
    summ = np.zeros((S, 2)) if want_summaries else None
    keys = ("node_off", "elem_off", "con_off", "nload_off", "eload_off", "node_xy", "elem_nodes", "elem_type", "elem_EAI",
            "con_node", "con_dof", "con_val", "nload_node", "nload_dof", "nload_val", "eload_elem", "eload_q", "eload_dir")
    dt = L.obo_run_batch(S, *[_p(np.ascontiguousarray(batch[k])) for k in keys], n_threads, _p(summ))
    if dt < 0:
        raise OracleError("obo_run_batch failed")
    return dt, summ


def ref_lib():
    """the reference itself (oracle/_ref), same C ABI as the restated oracle"""
    global _REF
    if _REF is None:
        so = build_ref()
        if so is None:
            raise OracleError("oracle/_ref/libob_ref.so is not built and /root/reference is not here to build it")
        _REF = _bind(C.CDLL(so))
    return _REF


def lib():
    global _LIB
    if _LIB is None:
        _LIB = _bind(C.CDLL(build()))
    return _LIB


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class OracleProblem:
    """One structure inside the oracle; methods follow the reference's names."""

    def __init__(self, st, reference=False):
        """reference=True: run the reference's own sources (oracle/_ref) instead of the restatement"""
        self.st = st
        L = self._L = ref_lib() if reference else lib()
        self._keep = st
        self.h = L.obo_create(
            st.n_nodes, _p(st.node_xy), _p(st.node_rotz), st.n_elems, _p(st.elem_nodes), _p(st.elem_type),
            _p(st.elem_EAI), len(st.con_node), _p(st.con_node), _p(st.con_dof), _p(st.con_val),
            len(st.nload_node), _p(st.nload_node), _p(st.nload_dof), _p(st.nload_val),
            len(st.eload_elem), _p(st.eload_elem), _p(st.eload_q), _p(st.eload_dir))
        if not self.h:
            raise OracleError("obo_create failed (bad element type?)")
        if getattr(st, "eload_type", None) is not None:
            if L.obo_set_eload_kinds(self.h, _p(st.eload_type), _p(st.eload_q2)) != 0:
                raise OracleError("obo_set_eload_kinds failed (bad load type?)")

    def __del__(self):
        if getattr(self, "h", None):
            self._L.obo_destroy(self.h)
            self.h = None

    def _chk(self, rc):
        if rc != 0:
            raise OracleError(self._L.obo_last_error(self.h).decode())

    def assembleProblem(self, skip_subK=False):
        self._chk(self._L.obo_assemble(self.h, int(skip_subK)))
        return self

    def solveStatic(self, nonLinearIterative=False):
        self._chk(self._L.obo_solve_static(self.h, 2 if nonLinearIterative else 0))
        return self

    def postProcCalcStress(self):
        self._chk(self._L.obo_stress(self.h))
        return self

    def set_load_values(self, nl_val=None, el_q=None):
        nl = None if nl_val is None else np.ascontiguousarray(nl_val, np.float64)
        eq = None if el_q is None else np.ascontiguousarray(el_q, np.float64)
        self._L.obo_set_load_values(self.h, _p(nl), _p(eq))

    def set_elem_EAI(self, eai):
        eai = np.ascontiguousarray(eai, np.float64)
        self._L.obo_set_elem_EAI(self.h, _p(eai))

    def sizes(self):
        out = np.zeros(8, np.int64)
        self._L.obo_get_sizes(self.h, _p(out))
        return dict(zip(("N", "E", "n", "n_f", "n_b", "T", "nnz"), out[:7].tolist()))

    def geti(self, which):
        n = self._L.obo_get_i64(self.h, which, None, 0)
        out = np.zeros(max(n, 0), np.int64)
        self._L.obo_get_i64(self.h, which, _p(out), n)
        return out

    def getf(self, which):
        n = self._L.obo_get_f64(self.h, which, None, 0)
        out = np.zeros(max(n, 0), np.float64)
        self._L.obo_get_f64(self.h, which, _p(out), n)
        return out

    # named accessors (reference field names)
    @property
    def free_dof_indices(self): return self.geti(I_FREE)
    @property
    def bounded_dof_indices(self): return self.geti(I_BOUNDED)
    @property
    def problem_DoFs(self): return np.stack([self.geti(I_DOF_NODE), self.geti(I_DOF_ID)], 1)
    @property
    def U_f(self): return self.getf(F_UF)
    @property
    def F_b(self): return self.getf(F_FB)
    @property
    def F_f(self): return self.getf(F_FF)
    @property
    def U_b(self): return self.getf(F_UB)
    @property
    def element_stress(self): return self.getf(F_STRESS).reshape(-1, 2, 6)

    def K_csc_upper(self):
        return self.geti(I_K_OUTER), self.geti(I_K_INNER), self.getf(F_K_VAL)

    def K_ff_dense_lower(self):
        s = self.sizes()
        return self.getf(F_KFF).reshape(s["n_f"], s["n_f"]).T  # col-major -> [r,c]

    def K_bf_dense(self):
        s = self.sizes()
        return self.getf(F_KBF).reshape(s["n_f"], s["n_b"]).T

    def run_cases(self, n_cases, nl_val=None, el_q=None, elem_EAI=None, n_threads=1, want_outputs=True):
        """CPU-baseline driver: one solveStatic + postProcCalcStress per case.  Returns
        (seconds, U_f, F_b, stress)."""
        self.solveStatic()
        s = self.sizes()
        nl = None if nl_val is None else np.ascontiguousarray(nl_val, np.float64)
        eq = None if el_q is None else np.ascontiguousarray(el_q, np.float64)
        ea = None if elem_EAI is None else np.ascontiguousarray(elem_EAI, np.float64)
        U = np.zeros((n_cases, s["n_f"])) if want_outputs else None
        F = np.zeros((n_cases, s["n_b"])) if want_outputs else None
        S = np.zeros((n_cases, s["E"], 2, 6)) if want_outputs else None
        dt = self._L.obo_run_cases(self.h, n_cases, _p(nl), _p(eq), _p(ea), n_threads, _p(U), _p(F), _p(S))
        if dt < 0:
            raise OracleError("obo_run_cases failed")
        return dt, U, F, S
