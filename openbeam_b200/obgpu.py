"""ctypes binding of libobgpu.so (include/obgpu.h) plus the host-side mirror of the reference's
problem interface for the static-solve path:

    CFiniteElementProblem::assembleProblem()     libopenbeam/include/openbeam/CFiniteElementProblem.h:435
    CFiniteElementProblem::solveStatic()         ...:444-446
    CFiniteElementProblem::postProcCalcStress()  ...:453-454

No arithmetic of the path happens in this file: it packs SoA arrays, calls the C ABI and wraps
device results.  There is NO CPU fallback — if libobgpu.so is missing or no sm_100 device is
usable, every entry point raises.
"""
from __future__ import annotations

import ctypes as C
import os
import weakref

import numpy as np

from .structure import Structure

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("OBGPU_LIB") or os.path.join(_HERE, "libobgpu.so")  # OBGPU_LIB: kernel-variant experiments

OK, EINVAL, ECUDA, ESINGULAR, ENOTCONVERGED, EUNSUPPORTED, ESEQUENCE = range(7)
BATCHED_LDLT, PCG_JACOBI, PCG_MG = range(3)
PROF_SPMV, PROF_MG_SPMV = 0, 1
SOLVE_SKIP_FF, SOLVE_COMPACT_STRESS = 1, 2  # obgpu_solve_opts.flags
(RES_UF, RES_FB, RES_STRESS, RES_INFO, RES_FF, RES_UB, RES_KVAL, RES_ELEM_K, RES_STRESS_COMPACT, RES_KFF_DENSE,
 RES_KBF_DENSE, RES_KBB_DENSE, RES_KFF_CSR_VAL) = range(13)
(SYM_DOF_NODE, SYM_DOF_ID, SYM_FREE, SYM_BOUNDED, SYM_DOF_INDEX, SYM_K_ROWPTR, SYM_K_COL, SYM_TRI_NNZ, SYM_KFF_ROWPTR,
 SYM_KFF_COL) = range(10)
OFF_DOF, OFF_FREE, OFF_BOUNDED, OFF_NNZ, OFF_TRIPLET = range(5)
SIZE_NAMES = ("S", "L", "Ntot", "Etot", "n_tot", "nf_tot", "nb_tot", "T_tot", "nnz_tot", "Ctot", "NLtot", "ELtot")

# every symbol include/obgpu.h declares (tests check the library exports all of them)
ABI_SYMBOLS = (
    "obgpu_create", "obgpu_destroy", "obgpu_last_error", "obgpu_status_string", "obgpu_host_alloc",
    "obgpu_host_free", "obgpu_batch_upload", "obgpu_batch_upload_device", "obgpu_update_cases", "obgpu_symbolic", "obgpu_assemble",
    "obgpu_solve_static", "obgpu_stress", "obgpu_get_sizes", "obgpu_get_offsets", "obgpu_get_symbolic",
    "obgpu_download", "obgpu_result_count", "obgpu_symbolic_count", "obgpu_device_ptr", "obgpu_synchronize", "obgpu_solve_cases_host", "obgpu_solve_cases_host_flags",
    "obgpu_summaries_device", "obgpu_get_timings", "obgpu_pcg_info", "obgpu_profile_kernel", "obgpu_set_stress_length_poses", "obgpu_get_stream", "obgpu_kernel_info", "obgpu_bind_summaries", "obgpu_bind_summary_peers",
    "obgpu_mesh_create", "obgpu_mesh_desc", "obgpu_mesh_info", "obgpu_mesh_destroy", "obgpu_scan_offsets_static",
)


class BatchDesc(C.Structure):
    _fields_ = [
        ("n_structures", C.c_int64),
        ("node_off", C.c_void_p), ("elem_off", C.c_void_p), ("con_off", C.c_void_p),
        ("nload_off", C.c_void_p), ("eload_off", C.c_void_p),
        ("node_xy", C.c_void_p), ("node_rotz", C.c_void_p),
        ("elem_nodes", C.c_void_p), ("elem_type", C.c_void_p), ("elem_EAI", C.c_void_p),
        ("con_node", C.c_void_p), ("con_dof", C.c_void_p), ("con_val", C.c_void_p),
        ("nload_node", C.c_void_p), ("nload_dof", C.c_void_p), ("nload_val", C.c_void_p),
        ("eload_elem", C.c_void_p), ("eload_q", C.c_void_p), ("eload_dir", C.c_void_p),
        ("n_cases", C.c_int64),
        ("case_nload_val", C.c_void_p), ("case_eload_q", C.c_void_p), ("case_elem_EAI", C.c_void_p),
        ("eload_type", C.c_void_p), ("eload_q2", C.c_void_p),
    ]


class SolveOpts(C.Structure):
    _fields_ = [("method", C.c_int), ("with_stress", C.c_int), ("rtol", C.c_double), ("max_iter", C.c_int),
                ("flags", C.c_int)]


class ObgpuError(RuntimeError):
    def __init__(self, status, msg):
        super().__init__(msg)
        self.status = status


_LIB = None


def lib():
    """Load libobgpu.so (fails loudly when it has not been built: no fallback)."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise ObgpuError(ECUDA, f"{LIB_PATH} not built — run `python -c 'import __graft_entry__ as g; g.build()'`")
        L = C.CDLL(LIB_PATH)
        vp, i64 = C.c_void_p, C.c_int64
        L.obgpu_create.argtypes = [C.c_int, C.POINTER(vp)]
        L.obgpu_destroy.argtypes = [vp]
        L.obgpu_destroy.restype = None
        L.obgpu_kernel_info.argtypes = [vp]
        L.obgpu_kernel_info.restype = C.c_char_p
        L.obgpu_last_error.argtypes = [vp]
        L.obgpu_last_error.restype = C.c_char_p
        L.obgpu_status_string.argtypes = [C.c_int]
        L.obgpu_status_string.restype = C.c_char_p
        L.obgpu_host_alloc.argtypes = [C.POINTER(vp), C.c_size_t]
        L.obgpu_host_free.argtypes = [vp]
        L.obgpu_batch_upload.argtypes = [vp, C.POINTER(BatchDesc)]
        L.obgpu_batch_upload_device.argtypes = [vp, C.POINTER(BatchDesc)]
        L.obgpu_update_cases.argtypes = [vp, i64, vp, vp, vp]
        for f in ("obgpu_symbolic", "obgpu_assemble", "obgpu_stress", "obgpu_synchronize"):
            getattr(L, f).argtypes = [vp]
        L.obgpu_solve_static.argtypes = [vp, C.POINTER(SolveOpts)]
        L.obgpu_get_sizes.argtypes = [vp, vp]
        L.obgpu_get_offsets.argtypes = [vp, C.c_int, vp]
        L.obgpu_get_symbolic.argtypes = [vp, C.c_int, vp]
        L.obgpu_download.argtypes = [vp, C.c_int, vp]
        L.obgpu_device_ptr.argtypes = [vp, C.c_int, C.POINTER(vp)]
        L.obgpu_solve_cases_host.argtypes = [vp, i64, vp, vp, vp, vp, vp, vp, vp]
        L.obgpu_solve_cases_host_flags.argtypes = [vp, i64, vp, vp, vp, vp, vp, vp, vp, C.c_int]
        L.obgpu_summaries_device.argtypes = [vp, vp]
        L.obgpu_mesh_create.argtypes = [C.POINTER(BatchDesc), C.c_double, C.POINTER(vp)]
        L.obgpu_mesh_desc.argtypes = [vp, C.POINTER(BatchDesc)]
        L.obgpu_mesh_info.argtypes = [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]
        L.obgpu_mesh_destroy.argtypes = [vp]
        L.obgpu_mesh_destroy.restype = None
        L.obgpu_bind_summaries.argtypes = [vp, vp]
        L.obgpu_bind_summary_peers.argtypes = [vp, C.c_int, C.POINTER(C.c_void_p)]
        L.obgpu_get_timings.argtypes = [vp, vp]
        L.obgpu_pcg_info.argtypes = [vp, vp]
        L.obgpu_result_count.argtypes = [vp, C.c_int, vp]
        L.obgpu_symbolic_count.argtypes = [vp, C.c_int, vp]
        L.obgpu_set_stress_length_poses.argtypes = [vp, vp]
        L.obgpu_profile_kernel.argtypes = [vp, C.c_int, C.c_int, vp, vp]
        L.obgpu_get_stream.argtypes = [vp, C.POINTER(vp)]
        L.obgpu_scan_offsets_static.argtypes = [C.c_int, i64, vp, vp, vp, vp, vp, C.c_char_p, C.c_size_t]
        _LIB = L
    return _LIB


def _p(a):
    if a is None:
        return None
    if isinstance(a, int):
        return C.c_void_p(a)
    return a.ctypes.data_as(C.c_void_p)


def _free_pinned(addr):
    if _LIB is not None:
        _LIB.obgpu_host_free(C.c_void_p(addr))


def pinned_empty(shape, dtype=np.float64):
    """numpy array over cudaHostAlloc'ed memory (obgpu_host_alloc); freed when the last view dies."""
    count = int(np.prod(shape))
    n = max(count * np.dtype(dtype).itemsize, 8)
    ptr = C.c_void_p()
    if lib().obgpu_host_alloc(C.byref(ptr), n) != OK:
        raise ObgpuError(ECUDA, "obgpu_host_alloc failed")
    buf = (C.c_char * n).from_address(ptr.value)
    weakref.finalize(buf, _free_pinned, ptr.value)  # numpy keeps buf alive through the buffer protocol
    return np.frombuffer(buf, dtype=dtype, count=count).reshape(shape)


def scan_offsets_static(device, n_structures, d_counts, d_dof_off, d_free_off, d_bnd_off, stream=0):
    """obgpu_scan_offsets_static: the symbolic phase's offset scans with static device scratch (allocates nothing).  All
    array arguments are DEVICE addresses (ints): counts [S][4] int32 in, three [S+1] int64 offset arrays out."""
    err = C.create_string_buffer(256)
    rc = lib().obgpu_scan_offsets_static(int(device), int(n_structures), int(d_counts), int(d_dof_off), int(d_free_off),
                                         int(d_bnd_off), int(stream) or None, err, len(err))
    if rc != OK:
        raise ObgpuError(rc, err.value.decode() or lib().obgpu_status_string(rc).decode())


def pack_batch(structures):
    """CSR-of-structures SoA arrays of obgpu_batch_desc from a list of Structure."""
    S = len(structures)

    def off(counts):
        o = np.zeros(S + 1, np.int64)
        np.cumsum(counts, out=o[1:])
        return o

    def cat(arrs, dt, shape=None):
        a = np.concatenate([np.asarray(x, dt).reshape(-1) for x in arrs]) if arrs else np.zeros(0, dt)
        return np.ascontiguousarray(a, dt)

    b = dict(
        S=S,
        node_off=off([s.n_nodes for s in structures]),
        elem_off=off([s.n_elems for s in structures]),
        con_off=off([len(s.con_node) for s in structures]),
        nload_off=off([len(s.nload_node) for s in structures]),
        eload_off=off([len(s.eload_elem) for s in structures]),
        node_xy=cat([s.node_xy for s in structures], np.float64),
        elem_nodes=cat([s.elem_nodes for s in structures], np.uint32),
        elem_type=cat([s.elem_type for s in structures], np.uint8),
        elem_EAI=cat([s.elem_EAI for s in structures], np.float64),
        con_node=cat([s.con_node for s in structures], np.uint32),
        con_dof=cat([s.con_dof for s in structures], np.uint8),
        con_val=cat([s.con_val for s in structures], np.float64),
        nload_node=cat([s.nload_node for s in structures], np.uint32),
        nload_dof=cat([s.nload_dof for s in structures], np.uint8),
        nload_val=cat([s.nload_val for s in structures], np.float64),
        eload_elem=cat([s.eload_elem for s in structures], np.uint32),
        eload_q=cat([s.eload_q for s in structures], np.float64),
        eload_dir=cat([s.eload_dir for s in structures], np.float64),
        node_rotz=None, eload_type=None, eload_q2=None,
    )
    if any(s.eload_type is not None for s in structures):
        b["eload_type"] = cat([s.eload_type if s.eload_type is not None else np.zeros(len(s.eload_elem), np.uint8)
                               for s in structures], np.uint8)
        b["eload_q2"] = cat([s.eload_q2 if s.eload_q2 is not None else np.zeros(len(s.eload_elem))
                             for s in structures], np.float64)
    if any(s.node_rotz is not None for s in structures):
        b["node_rotz"] = cat([s.node_rotz if s.node_rotz is not None else np.zeros(s.n_nodes) for s in structures],
                             np.float64)
    return b


def _fill_desc(d, batch):
    d.n_structures = batch["S"]
    for k in ("node_off", "elem_off", "con_off", "nload_off", "eload_off", "node_xy", "node_rotz", "elem_nodes",
              "elem_type", "elem_EAI", "con_node", "con_dof", "con_val", "nload_node", "nload_dof", "nload_val",
              "eload_elem", "eload_q", "eload_dir", "eload_type", "eload_q2"):
        a = batch.get(k)
        setattr(d, k, None if a is None or a.size == 0 else a.ctypes.data)
    d.n_cases = 1


def mesh_structure(st, max_element_length):
    """``CStructureProblem::mesh()`` (CStructureProblem_mesh.cpp:37-132) through ``obgpu_mesh_create``: returns the meshed
    Structure and the MeshOutputInfo lists ``element2elements`` / ``element2nodes`` (one list per original element)."""
    batch = pack_batch([st])
    d = BatchDesc()
    _fill_desc(d, batch)
    h = C.c_void_p()
    rc = lib().obgpu_mesh_create(C.byref(d), float(max_element_length), C.byref(h))
    if rc != 0:
        raise ObgpuError(rc, "obgpu_mesh_create: " + lib().obgpu_status_string(rc).decode())
    try:
        o = BatchDesc()
        lib().obgpu_mesh_desc(h, C.byref(o))

        def arr(ptr, n, dt):
            if not ptr or n == 0:
                return np.zeros(0, dt)
            return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(np.ctypeslib.as_ctypes_type(dt))), shape=(n,)).copy()

        off = lambda ptr: int(arr(ptr, 2, np.int64)[1])
        N, E, Cn, NL, EL = off(o.node_off), off(o.elem_off), off(o.con_off), off(o.nload_off), off(o.eload_off)
        out = Structure(
            node_xy=arr(o.node_xy, 2 * N, np.float64), elem_nodes=arr(o.elem_nodes, 2 * E, np.uint32),
            elem_type=arr(o.elem_type, E, np.uint8), elem_EAI=arr(o.elem_EAI, 3 * E, np.float64),
            con_node=arr(o.con_node, Cn, np.uint32), con_dof=arr(o.con_dof, Cn, np.uint8),
            con_val=arr(o.con_val, Cn, np.float64), nload_node=arr(o.nload_node, NL, np.uint32),
            nload_dof=arr(o.nload_dof, NL, np.uint8), nload_val=arr(o.nload_val, NL, np.float64),
            eload_elem=arr(o.eload_elem, EL, np.uint32), eload_q=arr(o.eload_q, EL, np.float64),
            eload_dir=arr(o.eload_dir, 3 * EL, np.float64),
            node_rotz=arr(o.node_rotz, N, np.float64) if o.node_rotz else None,
            eload_type=arr(o.eload_type, EL, np.uint8) if o.eload_type else None,
            eload_q2=arr(o.eload_q2, EL, np.float64) if o.eload_q2 else None)
        p = [C.c_void_p() for _ in range(4)]
        lib().obgpu_mesh_info(h, *[C.byref(x) for x in p])
        E0 = st.n_elems if not callable(st.n_elems) else st.n_elems()
        e2e_off, e2n_off = arr(p[0].value, E0 + 1, np.int64), arr(p[2].value, E0 + 1, np.int64)
        e2e, e2n = arr(p[1].value, int(e2e_off[-1]), np.uint32), arr(p[3].value, int(e2n_off[-1]), np.uint32)
        info = {"element2elements": [e2e[e2e_off[i]:e2e_off[i + 1]].tolist() for i in range(E0)],
                "element2nodes": [e2n[e2n_off[i]:e2n_off[i + 1]].tolist() for i in range(E0)],
                "num_original_nodes": len(st.node_xy), "num_original_elements": E0}
        return out, info
    finally:
        lib().obgpu_mesh_destroy(h)


class Context:
    """One obgpu_ctx (one GPU).  Thin: every method is one C-ABI call."""

    def __init__(self, device=0):
        self.h = C.c_void_p()
        rc = lib().obgpu_create(int(device), C.byref(self.h))
        if rc != OK:
            self.h = None
            raise ObgpuError(rc, "obgpu_create: " + lib().obgpu_status_string(rc).decode())
        self.device = device
        self._keep = None

    def close(self):
        if getattr(self, "h", None):
            lib().obgpu_destroy(self.h)
            self.h = None

    __del__ = close

    def _chk(self, rc):
        if rc != OK:
            msg = lib().obgpu_last_error(self.h).decode() or lib().obgpu_status_string(rc).decode()
            raise ObgpuError(rc, msg)

    # ---- upload ----
    def upload(self, batch, n_cases=1, case_nload_val=None, case_eload_q=None, case_elem_EAI=None, device=False):
        """batch: a Structure, a list of them, or a packed dict (pack_batch / workloads.c3_batch).  device=True:
        obgpu_batch_upload_device — arrays straight to HBM, symbolic phase as device kernels (many small structures)."""
        if isinstance(batch, Structure):
            batch = pack_batch([batch])
        elif isinstance(batch, (list, tuple)):
            batch = pack_batch(list(batch))
        d = BatchDesc()
        d.n_structures = batch["S"]
        for k in ("node_off", "elem_off", "con_off", "nload_off", "eload_off", "node_xy", "node_rotz", "elem_nodes",
                  "elem_type", "elem_EAI", "con_node", "con_dof", "con_val", "nload_node", "nload_dof", "nload_val",
                  "eload_elem", "eload_q", "eload_dir", "eload_type", "eload_q2"):
            a = batch.get(k)
            setattr(d, k, None if a is None or a.size == 0 else a.ctypes.data)
        d.n_cases = n_cases
        cn, cq, ce = (None if a is None else np.ascontiguousarray(a, np.float64)
                      for a in (case_nload_val, case_eload_q, case_elem_EAI))
        d.case_nload_val = None if cn is None else cn.ctypes.data
        d.case_eload_q = None if cq is None else cq.ctypes.data
        d.case_elem_EAI = None if ce is None else ce.ctypes.data
        self._keep = (batch, cn, cq, ce)
        fn = lib().obgpu_batch_upload_device if device else lib().obgpu_batch_upload
        self._chk(fn(self.h, C.byref(d)))
        self._keep = None
        return self

    def update_cases(self, n_cases, case_nload_val=None, case_eload_q=None, case_elem_EAI=None):
        self._chk(lib().obgpu_update_cases(self.h, n_cases, _p(case_nload_val), _p(case_eload_q), _p(case_elem_EAI)))

    # ---- phases ----
    def symbolic(self):
        self._chk(lib().obgpu_symbolic(self.h))

    def assemble(self):
        self._chk(lib().obgpu_assemble(self.h))

    def bind_summaries(self, ptr):
        self._chk(lib().obgpu_bind_summaries(self.h, _p(int(ptr)) if ptr else None))

    def bind_summary_peers(self, ptrs):
        """ptrs: device pointers (ints) of this rank's slot in the other GPUs' gathered buffers; [] unbinds."""
        arr = (C.c_void_p * max(len(ptrs), 1))(*[int(p) for p in ptrs])
        self._chk(lib().obgpu_bind_summary_peers(self.h, len(ptrs), arr))

    def kernel_info(self):
        return lib().obgpu_kernel_info(self.h).decode()

    def solve_static(self, method=BATCHED_LDLT, with_stress=True, rtol=1e-12, max_iter=0, keep_Ff=True,
                     compact_stress=False):
        o = SolveOpts(method, int(with_stress), rtol, max_iter,
                      (0 if keep_Ff else SOLVE_SKIP_FF) | (SOLVE_COMPACT_STRESS if compact_stress else 0))
        self._chk(lib().obgpu_solve_static(self.h, C.byref(o)))

    def stress(self):
        self._chk(lib().obgpu_stress(self.h))

    def synchronize(self):
        self._chk(lib().obgpu_synchronize(self.h))

    def solve_cases_host(self, n_cases, case_nload_val, case_eload_q, case_elem_EAI, U_f, F_b, stress, info,
                         compact_stress=False):
        """stress: [n_cases, E, 2, 6], or [n_cases, E, 2, 3] = (N, Vy, Mz) with compact_stress"""
        self._chk(lib().obgpu_solve_cases_host_flags(self.h, n_cases, _p(case_nload_val), _p(case_eload_q),
                                                _p(case_elem_EAI), _p(U_f), _p(F_b), _p(stress), _p(info),
                                                SOLVE_COMPACT_STRESS if compact_stress else 0))

    def summaries_device(self, dptr):
        self._chk(lib().obgpu_summaries_device(self.h, _p(int(dptr))))

    # ---- read-back ----
    def sizes(self):
        out = np.zeros(12, np.int64)
        self._chk(lib().obgpu_get_sizes(self.h, _p(out)))
        return dict(zip(SIZE_NAMES, out.tolist()))

    def offsets(self, which):
        out = np.zeros(self.sizes()["S"] + 1, np.int64)
        self._chk(lib().obgpu_get_offsets(self.h, which, _p(out)))
        return out

    def symbolic_array(self, which):
        cnt = C.c_int64(0)
        self._chk(lib().obgpu_symbolic_count(self.h, which, C.byref(cnt)))
        n = cnt.value
        out = np.zeros(max(n, 1), np.int32)
        self._chk(lib().obgpu_get_symbolic(self.h, which, _p(out)))
        return out[:n]

    def result_shape(self, which):
        z = self.sizes()
        L = z["L"]
        return {RES_UF: ((L, z["nf_tot"]), np.float64), RES_FF: ((L, z["nf_tot"]), np.float64),
                RES_FB: ((L, z["nb_tot"]), np.float64), RES_UB: ((z["nb_tot"],), np.float64),
                RES_STRESS: ((L, z["Etot"], 2, 6), np.float64), RES_INFO: ((L, z["S"]), np.int32),
                RES_STRESS_COMPACT: ((L, z["Etot"], 2, 3), np.float64),
                RES_KVAL: ((-1, z["nnz_tot"]), np.float64), RES_ELEM_K: ((-1, z["Etot"], 21), np.float64)}[which]

    def result_count(self, which):
        cnt = C.c_int64(0)
        self._chk(lib().obgpu_result_count(self.h, which, C.byref(cnt)))
        return cnt.value

    def sub_matrices(self, k=0):
        """BuildProblemInfo::K_ff (lower-stored), K_bf, K_bb (upper-stored) per structure, dense like the reference builds
        them (assemble.cpp:249-297), for K copy k: three lists of 2-D arrays"""
        z = self.sizes()
        fo, bo = self.offsets(OFF_FREE), self.offsets(OFF_BOUNDED)
        LK = z["L"] if self.result_count(RES_KVAL) > z["nnz_tot"] else 1
        out = []
        for which, rows, cols in ((RES_KFF_DENSE, "f", "f"), (RES_KBF_DENSE, "b", "f"), (RES_KBB_DENSE, "b", "b")):
            flat = np.zeros(max(self.result_count(which), 1))
            self._chk(lib().obgpu_download(self.h, which, _p(flat)))
            per = self.result_count(which) // LK
            flat = flat[k * per:(k + 1) * per]
            mats, p = [], 0
            for s in range(z["S"]):
                nr = (fo if rows == "f" else bo)[s + 1] - (fo if rows == "f" else bo)[s]
                nc = (fo if cols == "f" else bo)[s + 1] - (fo if cols == "f" else bo)[s]
                mats.append(flat[p:p + nr * nc].reshape(nc, nr).T.copy())   # column-major blocks
                p += nr * nc
            out.append(mats)
        return out

    def download(self, which, out=None, lk=None):
        if which in (RES_KFF_DENSE, RES_KBF_DENSE, RES_KBB_DENSE, RES_KFF_CSR_VAL):
            if out is None:
                out = np.zeros(max(self.result_count(which), 1))
            self._chk(lib().obgpu_download(self.h, which, _p(out)))
            return out[:self.result_count(which)]
        shape, dt = self.result_shape(which)
        if shape[0] == -1:
            shape = ((lk or 1),) + shape[1:]
        if out is None:
            out = np.zeros(shape, dt)
        rc = lib().obgpu_download(self.h, which, _p(out))
        if rc != OK and not (which == RES_INFO and rc == ESINGULAR):
            self._chk(rc)
        return out

    def device_ptr(self, which):
        p = C.c_void_p()
        self._chk(lib().obgpu_device_ptr(self.h, which, C.byref(p)))
        return p.value

    def timings(self):
        out = np.zeros(8)
        self._chk(lib().obgpu_get_timings(self.h, _p(out)))
        return dict(zip(("symbolic", "assembleProblem", "solveStatic.2_solve", "postProcCalcStress", "H2D", "D2H",
                         "launches", "pcg_iters"), out.tolist()))

    def profile_kernel(self, which, reps=20):
        """(avg ms per launch, algorithmic bytes per launch) of one iteration kernel of the last PCG solve"""
        ms, by = C.c_double(0), C.c_double(0)
        self._chk(lib().obgpu_profile_kernel(self.h, which, reps, C.byref(ms), C.byref(by)))
        return ms.value, by.value

    def set_stress_length_poses(self, node_xy):
        """node_xy [Ntot,2] or None: poses the element lengths of the stress recovery are read from (nonLinearIterative)"""
        a = None if node_xy is None else np.ascontiguousarray(node_xy, np.float64)
        self._chk(lib().obgpu_set_stress_length_poses(self.h, _p(a)))

    def pcg_info(self):
        """iterations, worst true relative residual, restarts, multigrid levels of the last PCG solve"""
        out = np.zeros(4)
        self._chk(lib().obgpu_pcg_info(self.h, _p(out)))
        return dict(zip(("iters", "true_relres", "restarts", "levels"), out.tolist()))

    def stream(self):
        p = C.c_void_p()
        self._chk(lib().obgpu_get_stream(self.h, C.byref(p)))
        return p.value
