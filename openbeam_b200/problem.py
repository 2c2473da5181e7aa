"""Host-side mirror of the reference's problem interface for the static-solve path, over the C ABI.

Names, argument meaning and error behaviour follow libopenbeam (paths under
/root/reference/libopenbeam/include/openbeam/):
    BuildProblemInfo            CFiniteElementProblem.h:76-102
    StaticSolveProblemInfo      CFiniteElementProblem.h:105-123
    StaticSolverAlgorithm       CFiniteElementProblem.h:69-73   (extended with the GPU_* cases)
    StaticSolverOptions         CFiniteElementProblem.h:165-171
    StressInfo / FaceStress     CFiniteElementProblem.h:173-183, types.h:78-157
    assembleProblem/solveStatic/postProcCalcStress   CFiniteElementProblem.h:435-454
The C++ twin of this file is include/obhost.hpp.  Nothing here computes: all arithmetic is in libobgpu.
"""
from __future__ import annotations

import enum
from dataclasses import dataclass, field

import dataclasses
import numpy as np

from . import obgpu
from .structure import Structure

NPOS = np.iinfo(np.uint64).max  # std::string::npos, used by dof_types for "n/a" (assemble.cpp:83-97)


class StaticSolverAlgorithm(enum.IntEnum):
    # the reference's two CPU algorithms are NOT available here (no CPU fallback); selecting them raises
    Dense_LLT = -1
    Sparse_LLT = -2
    GPU_Batched_LDLT = obgpu.BATCHED_LDLT
    GPU_PCG_Jacobi = obgpu.PCG_JACOBI
    GPU_PCG_MG = obgpu.PCG_MG


@dataclass
class StaticSolverOptions:
    algorithm: StaticSolverAlgorithm = StaticSolverAlgorithm.GPU_Batched_LDLT
    nonLinearIterative: bool = False
    pcg_rtol: float = 1e-12
    pcg_max_iter: int = 0


@dataclass
class BuildProblemInfo:
    free_dof_indices: np.ndarray = None
    bounded_dof_indices: np.ndarray = None
    dof_types: np.ndarray = None          # [n,2] uint64: (bounded_index, free_index), NPOS = n/a
    K_outer: np.ndarray = None            # K: Eigen CSC(upper) == CSR(lower): outerIndexPtr
    K_inner: np.ndarray = None            #    innerIndexPtr
    K_values: np.ndarray = None           #    valuePtr
    K_ff: np.ndarray = None               # [n_f,n_f] dense-structured, LOWER triangle filled (assemble.cpp:264-270)
    K_bf: np.ndarray = None               # [n_b,n_f]                                       (assemble.cpp:273-286)
    K_bb: np.ndarray = None               # [n_b,n_b] UPPER triangle filled (quirk Q1)      (assemble.cpp:258-262)
    U_b: np.ndarray = None
    F_f: np.ndarray = None


@dataclass
class StaticSolveProblemInfo:
    build_info: BuildProblemInfo = field(default_factory=BuildProblemInfo)
    U_f: np.ndarray = None
    F_b: np.ndarray = None


@dataclass
class StressInfo:
    element_stress: np.ndarray = None     # [E][2] x {N,Vy,Vz,Mx,My,Mz}


@dataclass
class MeshParams:
    max_element_length: float = 50e-3     # CStructureProblem_mesh.cpp:34


@dataclass
class MeshOutputInfo:                     # CFiniteElementProblem.h MeshOutputInfo
    num_original_nodes: int = 0
    num_original_elements: int = 0
    element2nodes: list = None
    element2elements: list = None


class CStructureProblem:
    """One structure on one GPU context.  ``solveStatic`` raises RuntimeError with the reference's message
    when K_ff is not positive definite (solver_static.cpp:98-105)."""

    def __init__(self, structure: Structure, ctx: obgpu.Context | None = None, device: int = 0):
        self.structure = structure
        self.ctx = ctx or obgpu.Context(device)
        self.ctx.upload(structure)
        self._sym = False
        self._stale_upload = False

    # ---- getters named as in CFiniteElementProblem.h ----
    def getNumberOfNodes(self):
        return self.structure.n_nodes

    def getNumberOfElements(self):
        return self.structure.n_elems

    def updateAll(self):
        if self._stale_upload:            # the device still holds the last pass of a nonLinearIterative solve
            self.ctx.upload(self.structure)
            self._stale_upload = False
        self.ctx.symbolic()
        self._sym = True

    def mesh(self, params: "MeshParams | None" = None, ctx: obgpu.Context | None = None):
        """CStructureProblem::mesh (CStructureProblem_mesh.cpp:37-132): returns ``(out_fem, MeshOutputInfo)``; the meshed
        problem lives on ``ctx`` (default: this problem's context, which is then re-uploaded with the meshed structure)."""
        params = params or MeshParams()
        meshed, info = obgpu.mesh_structure(self.structure, params.max_element_length)
        mo = MeshOutputInfo(info["num_original_nodes"], info["num_original_elements"], info["element2nodes"],
                            info["element2elements"])
        return CStructureProblem(meshed, ctx=ctx or self.ctx), mo

    def getProblemDoFs(self):
        """m_problem_DoFs as [n,2] (nodeId, dof)."""
        if not self._sym:
            self.updateAll()
        return np.stack([self.ctx.symbolic_array(obgpu.SYM_DOF_NODE), self.ctx.symbolic_array(obgpu.SYM_DOF_ID)], 1)

    def getDOFIndex(self, node, dof):
        """CFiniteElementProblem.cpp:308-317 — index or NPOS."""
        if not self._sym:
            self.updateAll()
        k = int(self.ctx.symbolic_array(obgpu.SYM_DOF_INDEX)[6 * node + dof])
        return NPOS if k < 0 else k

    def _fill_lists(self, bi: BuildProblemInfo):
        c = self.ctx
        bi.free_dof_indices = c.symbolic_array(obgpu.SYM_FREE).astype(np.uint64)
        bi.bounded_dof_indices = c.symbolic_array(obgpu.SYM_BOUNDED).astype(np.uint64)
        n = c.sizes()["n_tot"]
        dt = np.full((n, 2), NPOS, np.uint64)
        dt[bi.bounded_dof_indices.astype(np.int64), 0] = np.arange(len(bi.bounded_dof_indices), dtype=np.uint64)
        dt[bi.free_dof_indices.astype(np.int64), 1] = np.arange(len(bi.free_dof_indices), dtype=np.uint64)
        bi.dof_types = dt
        bi.K_outer = c.symbolic_array(obgpu.SYM_K_ROWPTR)
        bi.K_inner = c.symbolic_array(obgpu.SYM_K_COL)

    def assembleProblem(self, out_info: BuildProblemInfo):
        self.updateAll()
        self.ctx.assemble()
        self._fill_lists(out_info)
        out_info.K_values = self.ctx.download(obgpu.RES_KVAL)[0]
        # BuildProblemInfo::K_ff / K_bf / K_bb (CFiniteElementProblem.h:86-91): dense-structured like the reference's
        # (K_ff lower-stored, K_bb upper-stored), kept for problems small enough that this is not absurd
        if self.ctx.sizes()["n_tot"] <= 4096:
            kff, kbf, kbb = self.ctx.sub_matrices()
            out_info.K_ff, out_info.K_bf, out_info.K_bb = kff[0], kbf[0], kbb[0]
        out_info.U_b = self.ctx.download(obgpu.RES_UB)
        out_info.F_f = self.ctx.download(obgpu.RES_FF)[0]
        return out_info

    def _solve_once(self, options):
        self.updateAll()
        # element end forces come back as compact (N, Vy, Mz) records: Vz, Mx, My are structurally zero in 2D
        self.ctx.solve_static(int(options.algorithm), True, options.pcg_rtol, options.pcg_max_iter, compact_stress=True)
        info = self.ctx.download(obgpu.RES_INFO)
        if np.any(info != 0):
            raise RuntimeError(obgpu.lib().obgpu_status_string(obgpu.ESINGULAR).decode())

    N_MAX_ITERS = 40                      # solver_static.cpp:63

    def _solveStaticIterative(self, out_info, options):
        """StaticSolverOptions::nonLinearIterative (solver_static.cpp:55-58,63-66,139-193): the reference's loop sits
        above assemble+solve, so it sits above the C-ABI here too.  Every pass re-assembles and re-solves on the GPU
        with the node poses moved by the previous pass's U_f — a translation DOF adds to the node's global
        coordinate as it is (also for a node with a rotated frame, like the reference), RZ composes into the node's
        frame angle.  As in the reference there is no convergence test: all 40 passes run, the poses are restored
        afterwards, and U_f / F_b / the stresses are those of the LAST pass (an increment on the last geometry)."""
        orig = self.structure
        xy = orig.node_xy.copy()
        rz = np.zeros(orig.n_nodes) if orig.node_rotz is None else orig.node_rotz.copy()
        try:
            for it in range(self.N_MAX_ITERS):
                if it:
                    self.structure = dataclasses.replace(orig, node_xy=xy.copy(), node_rotz=rz.copy())
                    self.ctx.upload(self.structure)
                if it == self.N_MAX_ITERS - 1:
                    # solver_static.cpp:190-193 restores the poses BEFORE the caller's postProcCalcStress: its element
                    # lengths come from the original poses, orientations / U / fixed-end terms from this last pass
                    self._stale_upload = False
                    self.ctx.symbolic()
                    self.ctx.set_stress_length_poses(orig.node_xy)
                self._solve_once(options)
                U_f = self.ctx.download(obgpu.RES_UF)[0]
                free = self.ctx.symbolic_array(obgpu.SYM_FREE).astype(np.int64)
                node = self.ctx.symbolic_array(obgpu.SYM_DOF_NODE)[free]
                dof = self.ctx.symbolic_array(obgpu.SYM_DOF_ID)[free]
                for d in (0, 1):            # DZ, RX, RY do not exist in a planar problem
                    m = dof == d
                    np.add.at(xy[:, d], node[m], U_f[m])
                m = dof == 5
                np.add.at(rz, node[m], U_f[m])
        finally:
            self.structure = orig         # solver_static.cpp:188-191 — the device keeps the last pass for stress()
            self._stale_upload = True
        self._fill_lists(out_info.build_info)
        out_info.build_info.U_b = self.ctx.download(obgpu.RES_UB)
        out_info.build_info.F_f = self.ctx.download(obgpu.RES_FF)[0]
        out_info.U_f = self.ctx.download(obgpu.RES_UF)[0]
        out_info.F_b = self.ctx.download(obgpu.RES_FB)[0]
        return out_info

    def solveStatic(self, out_info: StaticSolveProblemInfo, options: StaticSolverOptions | None = None):
        options = options or StaticSolverOptions()
        if options.algorithm < 0:
            raise RuntimeError("StaticSolverAlgorithm::%s is the reference's CPU path; libobgpu has no CPU fallback"
                               % options.algorithm.name)
        if options.nonLinearIterative:
            return self._solveStaticIterative(out_info, options)
        self._solve_once(options)
        self._fill_lists(out_info.build_info)
        out_info.build_info.U_b = self.ctx.download(obgpu.RES_UB)
        out_info.build_info.F_f = self.ctx.download(obgpu.RES_FF)[0]
        out_info.U_f = self.ctx.download(obgpu.RES_UF)[0]
        out_info.F_b = self.ctx.download(obgpu.RES_FB)[0]
        return out_info

    def postProcCalcStress(self, out_stress: StressInfo, solver_info: StaticSolveProblemInfo | None = None):
        self.ctx.stress()
        c = self.ctx.download(obgpu.RES_STRESS_COMPACT)[0]          # [E][2][3] = N, Vy, Mz
        full = np.zeros((c.shape[0], 2, 6))                           # FaceStress order N,Vy,Vz,Mx,My,Mz (types.h:78-157)
        full[:, :, 0], full[:, :, 1], full[:, :, 5] = c[:, :, 0], c[:, :, 1], c[:, :, 2]
        out_stress.element_stress = full
        return out_stress
