// ob_solve_demo3.cpp — the programmatic use of the path, as examples-code/example1.cpp:34-172 does with the
// reference, here over include/obhost.hpp: builds examples-structures/demo3_frame.yaml:30-51 and prints what
// `ob-solve --show-U --show-FR --stress` prints (U_f / F_b dump, ob-solve_main.cpp:879-884).
//   usage: ob_solve_demo3 [device]      device = -1 -> symbolic-only ctx (no GPU): prints the DOF lists only
// build: g++ -std=c++17 -Iinclude examples/ob_solve_demo3.cpp -Lopenbeam_b200 -lobgpu -Wl,-rpath,$PWD/openbeam_b200
#include <cstdio>
#include <cstdlib>

#include "obhost.hpp"

int main(int argc, char** argv) {
  const int device = argc > 1 ? std::atoi(argv[1]) : 0;
  try {
    obhost::CStructureProblem p(device);
    const double L = 4.0, H = 2.0, E = 2.1e11, A = 28.5e-4, Iz = 1940e-8;
    p.insertNode(0, 0); p.insertNode(0, H); p.insertNode(L, H); p.insertNode(L, 0);
    p.insertElement(OBGPU_BEAM2D_RR, 0, 1, E, A, Iz);
    p.insertElement(OBGPU_BEAM2D_RR, 1, 2, E, A, Iz);
    p.insertElement(OBGPU_BEAM2D_RR, 2, 3, E, A, Iz);
    for (size_t n : {size_t(0), size_t(3)})
      for (uint8_t d : {uint8_t(OBGPU_DX), uint8_t(OBGPU_DY), uint8_t(OBGPU_RZ)}) p.insertConstraint(n, d);
    p.addLoadAtBeam(1, 1000 * 9.81, 0, -1, 0);
    const auto dofs = p.getProblemDoFs();
    std::printf("nDOFs %zu\n", dofs.size());
    for (size_t k = 0; k < dofs.size(); k++) std::printf("dof %zu node %zu id %d\n", k, dofs[k].nodeId, int(dofs[k].dof));
    if (device < 0) return 0;
    obhost::StaticSolveProblemInfo info;
    p.solveStatic(info);
    obhost::StressInfo st;
    p.postProcCalcStress(st, info);
    for (size_t i = 0; i < info.U_f.size(); i++) std::printf("U_f[%zu] %.17g\n", info.build_info.free_dof_indices[i], info.U_f[i]);
    for (size_t i = 0; i < info.F_b.size(); i++) std::printf("F_b[%zu] %.17g\n", info.build_info.bounded_dof_indices[i], info.F_b[i]);
    for (size_t e = 0; e < st.element_stress.size(); e++)
      for (int f = 0; f < 2; f++)
        std::printf("stress e%zu f%d N %.17g Vy %.17g Mz %.17g\n", e, f, st.element_stress[e][f].N,
                    st.element_stress[e][f].Vy, st.element_stress[e][f].Mz);
    // CStructureProblem::mesh at 0.5 m: 4 + 8 + 4 pieces; the displacements of the original nodes do not change
    obhost::CStructureProblem pm(device);
    obhost::MeshOutputInfo mo;
    obhost::MeshParams mp;
    mp.max_element_length = 0.5;
    p.mesh(pm, mo, mp);
    std::printf("meshed nodes %zu elements %zu pieces_of_beam %zu\n", pm.getNumberOfNodes(), pm.getNumberOfElements(),
                mo.element2elements[1].size());
    obhost::StaticSolveProblemInfo im;
    pm.solveStatic(im);
    for (size_t i = 0; i < 6; i++) std::printf("meshed U_f[%zu] %.17g\n", im.build_info.free_dof_indices[i], im.U_f[i]);
    // StaticSolverOptions::nonLinearIterative (solver_static.cpp:139-193): 40 passes, results of the last one; the
    // problem is back on its original poses afterwards
    obhost::StaticSolverOptions nl;
    nl.nonLinearIterative = true;
    obhost::StaticSolveProblemInfo in;
    p.solveStatic(in, nl);
    for (size_t i = 0; i < in.U_f.size(); i++) std::printf("iterative U_f[%zu] %.17g\n", in.build_info.free_dof_indices[i], in.U_f[i]);
    for (size_t i = 0; i < in.F_b.size(); i++) std::printf("iterative F_b[%zu] %.17g\n", in.build_info.bounded_dof_indices[i], in.F_b[i]);
    p.solveStatic(in);
    for (size_t i = 0; i < in.U_f.size(); i++) std::printf("again U_f[%zu] %.17g\n", in.build_info.free_dof_indices[i], in.U_f[i]);
  } catch (const std::exception& e) {
    std::printf("Exception: %s\n", e.what());  // ob-solve_main.cpp:1540-1548
    return -1;
  }
  return 0;
}
