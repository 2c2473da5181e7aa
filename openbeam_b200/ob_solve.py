"""ob_solve — the plain-text report of the reference's command-line solver over the GPU path.

    python -m openbeam_b200.ob_solve -i examples-structures/demo3_frame.yaml --show-U --stress

prints what `ob-solve -i ... --show-U --stress` prints in text mode (apps/ob-solve/ob-solve_main.cpp):
  * warnings of the loader, one per line                                            (ob-solve_main.cpp:571-585)
  * --full-k      "Global stiffness matrix(K):" + the dense n x n matrix            (:590-638)
  * --show-U      "RESULTS:" + U_f and F_b as Eigen prints a column vector          (:877-884)
  * --stress      "ELEMENT STRESS:" table, one row per element face                 (:996-1031)
The numbers come from CStructureProblem.solveStatic / postProcCalcStress of the mirror (openbeam_b200/problem.py), i.e.
from libobgpu; there is no CPU path.  HTML output, SVG / 3D scenes, animations and stress plots are not produced
(out of scope); `--solver` takes the GPU algorithms {GPU_Batched_LDLT, GPU_PCG_MG, GPU_PCG_Jacobi}.

Text layout follows Eigen's default IOFormat (Eigen/src/Core/IO.h: stream precision 6, columns right-aligned to the widest
coefficient of the whole matrix, one space between columns) and the reference's printf formats.
"""
from __future__ import annotations

import argparse
import sys

import numpy as np


def eigen_str(m) -> str:
    """`std::cout << m` for an Eigen dense vector or matrix with the default IOFormat."""
    a = np.atleast_2d(np.asarray(m, dtype=np.float64))
    if np.asarray(m).ndim == 1:
        a = a.T                                   # Eigen vectors are columns
    if a.size == 0:
        return ""
    cells = [["%g" % v for v in row] for row in a]
    w = max(len(c) for row in cells for c in row)
    return "\n".join(" ".join(c.rjust(w) for c in row) for row in cells)


def full_k(info) -> np.ndarray:
    """The n x n matrix ob-solve builds from K_ff (upper reads of a lower-stored matrix are mirrored the way the reference
    does: K(ii1,ii2) = K(ii2,ii1) = Kxx(i1,i2) for i2 >= i1), K_bb and K_bf (ob-solve_main.cpp:592-628)."""
    fi, bi = info.free_dof_indices.astype(np.int64), info.bounded_dof_indices.astype(np.int64)
    n = len(fi) + len(bi)
    K = np.zeros((n, n))
    for i1 in range(len(fi)):
        for i2 in range(i1, len(fi)):
            K[fi[i1], fi[i2]] = K[fi[i2], fi[i1]] = info.K_ff[i1, i2]
    for i1 in range(len(bi)):
        for i2 in range(i1, len(bi)):
            K[bi[i1], bi[i2]] = K[bi[i2], bi[i1]] = info.K_bb[i1, i2]
    for i1 in range(len(bi)):
        for i2 in range(len(fi)):
            K[bi[i1], fi[i2]] = K[fi[i2], bi[i1]] = info.K_bf[i1, i2]
    return K


def stress_table(element_stress) -> str:
    """ob-solve_main.cpp:1000-1031 (text branch)."""
    out = ["\n\nELEMENT STRESS:\n\n",
           "Element|Face|     N    |     Vy   |     Vz   |     Mx   |     My   |     Mz   |\n",
           "-------+----+----------+----------+----------+----------+----------+----------+\n"]
    for i, faces in enumerate(np.asarray(element_stress)):
        for face, es in enumerate(faces):
            out.append("%s| %2u |%10.3e|%10.3e|%10.3e|%10.3e|%10.3e|%10.3e|\n"
                       % ((" %5u " % i) if face == 0 else "       ", face, es[0], es[1], es[2], es[3], es[4], es[5]))
    return "".join(out)


def report(problem, sInfo, stress, args) -> str:
    info = sInfo.build_info
    out = []
    if args.full_k:
        # reads the upper triangle of K_ff exactly like the reference: the mirror returns K_ff with only its LOWER triangle
        # filled (assemble.cpp:264-270), so the free-free block prints as the reference prints it
        out.append("Global stiffness matrix(K):\n")
        out.append("K:\n" + eigen_str(full_k(info)) + "\n\n")
    if args.show_U:
        out.append("\n\nRESULTS:\n------------------------\n")
        out.append("Displacements (U_f):\n" + eigen_str(sInfo.U_f) + "\n")
        out.append("Reactions (F_b):\n" + eigen_str(sInfo.F_b) + "\n\n")
    if args.stress:
        out.append(stress_table(stress.element_stress))
    return "".join(out)


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="ob_solve", description="ob-solve text report over libobgpu (B200)")
    ap.add_argument("-i", "--input", required=True, help="structure definition (.yaml, the reference's format)")
    ap.add_argument("-o", "--output", default="", help="write the report to a file instead of stdout")
    ap.add_argument("--mesh", action="store_true", help="Enable meshing (Default: no)")
    ap.add_argument("--mesh-resolution", type=float, default=0.05, help="Meshing resolution (Default: 0.05 m)")
    ap.add_argument("--stress", action="store_true", help="Show element stress as text")
    ap.add_argument("--show-U", dest="show_U", action="store_true", help="Text display of all displacements")
    ap.add_argument("--show-FR", dest="show_FR", action="store_true", help="(HTML-only in the reference: accepted, no text)")
    ap.add_argument("--show-FL", dest="show_FL", action="store_true", help="(HTML-only in the reference: accepted, no text)")
    ap.add_argument("--full-k", dest="full_k", action="store_true", help="display the full stiffness matrix K")
    ap.add_argument("--solver", default="GPU_Batched_LDLT", help="{GPU_Batched_LDLT|GPU_PCG_MG|GPU_PCG_Jacobi}")
    ap.add_argument("--nonlinear", action="store_true", help="StaticSolverOptions::nonLinearIterative")
    ap.add_argument("--device", type=int, default=0)
    args = ap.parse_args(argv)

    from . import yaml_loader
    from .problem import (BuildProblemInfo, CStructureProblem, MeshParams, StaticSolveProblemInfo, StaticSolverAlgorithm, StaticSolverOptions,
                          StressInfo)
    try:
        algorithm = StaticSolverAlgorithm[args.solver]
    except KeyError:
        raise SystemExit("Invalid --solver algorithm name.")          # ob-solve_main.cpp:345-346
    try:
        structure = yaml_loader.load_structure(args.input)
    except Exception as e:                                              # errMsg lines, return -1 (ob-solve_main.cpp:240-244)
        print(str(e), file=sys.stderr)
        return 255
    try:
        problem = CStructureProblem(structure, device=args.device)
        if args.mesh:
            problem, _ = problem.mesh(MeshParams(max_element_length=args.mesh_resolution))
        else:
            problem.updateAll()
        pre = None
        if args.full_k:   # BuildProblemInfo::K_ff / K_bf / K_bb, which the reference's solveStatic leaves in sInfo.build_info
            pre = problem.assembleProblem(BuildProblemInfo())
            if pre.K_ff is None:
                raise RuntimeError("--full-k: the structure is too large for a dense K")
        sInfo = StaticSolveProblemInfo()
        problem.solveStatic(sInfo, StaticSolverOptions(algorithm=algorithm, nonLinearIterative=args.nonlinear))
        if pre is not None:
            sInfo.build_info.K_ff, sInfo.build_info.K_bf, sInfo.build_info.K_bb = pre.K_ff, pre.K_bf, pre.K_bb
        stress = StressInfo()
        if args.stress:
            problem.postProcCalcStress(stress, sInfo)
        text = report(problem, sInfo, stress, args)
    except Exception as e:                                              # ob-solve_main.cpp:1540-1548
        print("Exception: %s" % e, file=sys.stderr)
        return 255
    if args.output:
        with open(args.output, "w") as f:
            f.write(text)
    else:
        sys.stdout.write(text)
    return 0


if __name__ == "__main__":
    sys.exit(main())
