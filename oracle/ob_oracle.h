/* ob_oracle.h — C ABI of the CPU oracle (TEST INFRASTRUCTURE, not product code).
 *
 * The oracle is a CPU restatement of libopenbeam's static-solve hot path
 *   assembleProblem()      libopenbeam/src/CFiniteElementProblem_assemble.cpp:37-437
 *   solveStatic()          libopenbeam/src/CFiniteElementProblem_solver_static.cpp:29-219
 *   postProcCalcStress()   libopenbeam/src/CFiniteElementProblem_postproc_stress.cpp:32-183
 * The reference's cmake build needs MRPT >= 2.4 and Eigen, both absent from this image, and it ships no tests or golden
 * vectors for this path.  PARITY PINNED TO THE REFERENCE'S OWN CODE: oracle/_ref/libob_ref.so is libopenbeam's unmodified
 * hot-path sources compiled where they lie against header stand-ins for the Eigen/MRPT subset they use
 * (oracle/Makefile.ref, oracle/ref_shim/, oracle/ref_harness.cpp) and exposes THIS SAME C ABI; tests/test_reference_pin.py
 * holds the restatement to it (integer artefacts and assembled matrices bit-exact, results to 1e-12) live and through
 * committed vectors (tests/golden/reference/).  Also pinned by closed-form statics and independent numerics
 * (tests/test_oracle.py).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library.  The product (openbeam_b200/) never does.
 */
#ifndef OB_ORACLE_H
#define OB_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct obo_problem obo_problem;

/* element types: CElementBeam_2D_{RR,RA,AR,AA} */
enum { OBO_RR = 0, OBO_RA = 1, OBO_AR = 2, OBO_AA = 3,
       /* SURVEY 8f N4; constants ride in elem_EAI: (E,A,Iz) for RD, (K,-,-) SPRING_1D / SPRING_TORSION,
        * (Kx,Ky,-) SPRING_XY, (Kx,Ky,KRz) SPRING_DXDYRZ */
       OBO_RD = 4, OBO_SPRING_1D = 5, OBO_SPRING_XY = 6, OBO_SPRING_DXDYRZ = 7, OBO_SPRING_TORSION = 8 };
/* solver algorithm: StaticSolverAlgorithm (CFiniteElementProblem.h:69-73) */
enum { OBO_DENSE_LLT = 0, OBO_SPARSE_SKIP_SUBK = 1,
       OBO_NONLINEAR_ITERATIVE = 2 /* OR-ed in: StaticSolverOptions::nonLinearIterative (solver_static.cpp:139-193) */ };

/* integer arrays readable with obo_get_i64 */
enum {
  OBO_I_DOF_NODE = 0,     /* m_problem_DoFs[k].nodeId            [n]   */
  OBO_I_DOF_ID = 1,       /* m_problem_DoFs[k].dof (0..5)        [n]   */
  OBO_I_FREE = 2,         /* free_dof_indices                    [n_f] */
  OBO_I_BOUNDED = 3,      /* bounded_dof_indices                 [n_b] */
  OBO_I_DOF_INDEX = 4,    /* m_problem_DoFs_inverse_list  [N*6], -1 = unused */
  OBO_I_TRI_ROW = 5,      /* triplet rows                        [T]   */
  OBO_I_TRI_COL = 6,      /* triplet cols                        [T]   */
  OBO_I_K_OUTER = 7,      /* Eigen CSC(upper) outerIndexPtr      [n+1] */
  OBO_I_K_INNER = 8       /* Eigen CSC(upper) innerIndexPtr      [nnz] */
};
/* double arrays readable with obo_get_f64 */
enum {
  OBO_F_TRI_VAL = 0,      /* triplet values                      [T]   */
  OBO_F_K_VAL = 1,        /* CSC(upper) values                   [nnz] */
  OBO_F_KFF_DENSE = 2,    /* dense K_ff lower (col-major n_f x n_f, upper part = 0) */
  OBO_F_KBF_DENSE = 3,    /* dense K_bf (col-major n_b x n_f)          */
  OBO_F_KBB_DENSE = 4,    /* dense K_bb upper (col-major n_b x n_b)    */
  OBO_F_UB = 5, OBO_F_FF = 6, OBO_F_UF = 7, OBO_F_FB = 8,
  OBO_F_STRESS = 9,       /* [E][2][6] = N,Vy,Vz,Mx,My,Mz per face     */
  OBO_F_ELEM_R = 10,      /* [E][9] element orientation, col-major      */
  OBO_F_ELEM_KGLOBAL = 11 /* [E][3][36] global K11,K22,K12, col-major 6x6 */
};
/* sizes: obo_get_sizes fills out[0..7] = N, E, n, n_f, n_b, T, nnz, reserved */

obo_problem* obo_create(int64_t n_nodes, const double* node_xy, const double* node_rotz /*nullable, radians*/,
                        int64_t n_elems, const uint32_t* elem_nodes, const uint8_t* elem_type,
                        const double* elem_EAI,
                        int64_t n_cons, const uint32_t* con_node, const uint8_t* con_dof, const double* con_val,
                        int64_t n_nloads, const uint32_t* nl_node, const uint8_t* nl_dof, const double* nl_val,
                        int64_t n_eloads, const uint32_t* el_elem, const double* el_q, const double* el_dir);
void obo_destroy(obo_problem*);
/* element-load kinds (call right after obo_create; default = every load DISTRIB_UNIFORM):
 * kind[i] in {0 uniform, 1 triangular (q=q_ini, q2=q_end), 2 concentrated (q=P, q2=dist), 3 temperature (q=dT)} */
int obo_set_eload_kinds(obo_problem*, const uint8_t* kind, const double* q2);
/* replace load values in place (same slots): node-load values and element-load q */
int obo_set_load_values(obo_problem*, const double* nl_val, const double* el_q);
/* replace every element's (E,A,Iz) */
int obo_set_elem_EAI(obo_problem*, const double* elem_EAI);

int obo_assemble(obo_problem*, int skip_subK);     /* assembleProblem();  0 = ok */
int obo_solve_static(obo_problem*, int algorithm); /* solveStatic() (calls assemble) */
int obo_stress(obo_problem*);                      /* postProcCalcStress() */
const char* obo_last_error(const obo_problem*);

void obo_get_sizes(const obo_problem*, int64_t out[8]);
int64_t obo_get_i64(const obo_problem*, int which, int64_t* out, int64_t cap);
int64_t obo_get_f64(const obo_problem*, int which, double* out, int64_t cap);

/* CPU-baseline driver: clone the problem per thread; for each case c set the load
 * values (case-major arrays: nl_val[c*n_nloads+..], el_q[c*n_eloads+..], optional
 * per-case elem_EAI[c*3E+..]) and run solveStatic + postProcCalcStress, writing
 * U_f [c][n_f], F_b [c][n_b], stress [c][E][2][6] (any may be NULL).
 * Returns wall seconds spent inside the threaded region, <0 on error. */
double obo_run_cases(const obo_problem* tmpl, int64_t n_cases, const double* nl_val, const double* el_q,
                     const double* elem_EAI, int n_threads, double* U_f, double* F_b, double* stress);

/* CPU-baseline driver for a packed batch of DIFFERENT structures (obgpu_batch_desc layout, config C3): per structure a
 * full problem set-up + solveStatic + postProcCalcStress, structures dealt to n_threads host threads; summ (nullable)
 * receives {max|U_f|, sum F_b} per structure.  Returns wall seconds, <0 on error.  (oracle/ob_batch.ipp) */
double obo_run_batch(int64_t S, const int64_t* node_off, const int64_t* elem_off, const int64_t* con_off,
                     const int64_t* nload_off, const int64_t* eload_off, const double* node_xy, const uint32_t* elem_nodes,
                     const uint8_t* elem_type, const double* elem_EAI, const uint32_t* con_node, const uint8_t* con_dof,
                     const double* con_val, const uint32_t* nl_node, const uint8_t* nl_dof, const double* nl_val,
                     const uint32_t* el_elem, const double* el_q, const double* el_dir, int n_threads, double* summ);

#ifdef __cplusplus
}
#endif
#endif
