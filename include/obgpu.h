/* obgpu.h — C ABI of libobgpu, the B200-native static-solve path behind libopenbeam's
 *   CFiniteElementProblem::assembleProblem(BuildProblemInfo&)          include/openbeam/CFiniteElementProblem.h:435
 *   CFiniteElementProblem::solveStatic(StaticSolveProblemInfo&, opts)  include/openbeam/CFiniteElementProblem.h:444-446
 *   CFiniteElementProblem::postProcCalcStress(StressInfo&, info)       include/openbeam/CFiniteElementProblem.h:453-454
 * (reference paths relative to /root/reference/libopenbeam/).  The reference has no FFI or plugin
 * interface for this path; these entry points are what a GPU_* case added to
 * StaticSolverOptions::algorithm (CFiniteElementProblem.h:69-73,165-171) would bind — see INTEGRATION.md
 * and include/obhost.hpp for the C++ mirror of the reference's method and field names.
 *
 * Plain pointers and sizes only; the caller owns every host buffer, the library owns device memory
 * inside the opaque ctx.  Calls on one ctx must be serialised by the caller; different ctxs may be
 * used from different threads/processes (one ctx per GPU).  There is no CPU fallback: every call
 * fails with OBGPU_ECUDA when no sm_100 device is usable.
 */
#ifndef OBGPU_H
#define OBGPU_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct obgpu_ctx obgpu_ctx;

/* status codes (the C++ shim rethrows the reference's exception strings, solver_static.cpp:98-105) */
enum {
  OBGPU_OK = 0,
  OBGPU_EINVAL = 1,        /* malformed description (MRPT ASSERT_ in the reference)             */
  OBGPU_ECUDA = 2,         /* CUDA error / no device                                           */
  OBGPU_ESINGULAR = 3,     /* >=1 system had a non-positive pivot (reference: runtime_error)    */
  OBGPU_ENOTCONVERGED = 4, /* PCG hit max_iter                                                  */
  OBGPU_EUNSUPPORTED = 5,  /* valid input outside what the GPU path implements (stated in text) */
  OBGPU_ESEQUENCE = 6      /* call order violated (e.g. solve before symbolic)                  */
};

/* element types: CElementBeam_2D_{RR,RA,AR,AA}.cpp */
enum { OBGPU_BEAM2D_RR = 0, OBGPU_BEAM2D_RA = 1, OBGPU_BEAM2D_AR = 2, OBGPU_BEAM2D_AA = 3,
       /* CElementBeam_2D_RD.cpp, CElementSpring.cpp, CElementSpringXY.cpp, CElementSpringDXDYRZ.cpp (with its K11(5,5)
        * assignment slip, reproduced), CElementTorsionSpring.cpp.  Their constants ride in elem_EAI:
        * RD (E, A, Iz) | SPRING_1D (K, -, -) | SPRING_XY (Kx, Ky, -) | SPRING_DXDYRZ (Kx, Ky, KRz) | SPRING_TORSION (K, -, -) */
       OBGPU_BEAM2D_RD = 4, OBGPU_SPRING_1D = 5, OBGPU_SPRING_XY = 6, OBGPU_SPRING_DXDYRZ = 7, OBGPU_SPRING_TORSION = 8 };
/* element loads (libopenbeam/include/openbeam/loads.h): CLoadDistributedUniform, CLoadDistributedTriangular (RR only),
 * CLoadConcentratedForce (RR only, perpendicular to the beam), CLoadConstTemperature */
enum { OBGPU_LOAD_UNIFORM = 0, OBGPU_LOAD_TRIANGULAR = 1, OBGPU_LOAD_CONCENTRATED = 2, OBGPU_LOAD_TEMPERATURE = 3 };
/* DoF_index, CFiniteElementProblem.h:53-67 */
enum { OBGPU_DX = 0, OBGPU_DY = 1, OBGPU_DZ = 2, OBGPU_RX = 3, OBGPU_RY = 4, OBGPU_RZ = 5 };

/* A batch of S independent structures (CSR-of-structures SoA) x L load cases.
 * System (s, c) = structure s under load case c; there are S*L systems.
 * Element/constraint/load node indices are LOCAL to their structure.
 * Constraints and node loads on DOFs that do not exist in the problem are ignored, as the
 * reference's parser does (parse_file.cpp:653-662,747-756). */
typedef struct obgpu_batch_desc {
  int64_t n_structures;       /* S >= 1 */
  const int64_t* node_off;    /* [S+1] */
  const int64_t* elem_off;    /* [S+1] */
  const int64_t* con_off;     /* [S+1] */
  const int64_t* nload_off;   /* [S+1] */
  const int64_t* eload_off;   /* [S+1] */
  const double* node_xy;      /* [2*Ntot] x,y  (m_node_poses[i].t, z = 0)                         */
  const double* node_rotz;    /* NULL or [Ntot] nodal frame yaw in radians (parse_file.cpp:467)   */
  const uint32_t* elem_nodes; /* [2*Etot] conected_nodes_ids[0..1]                                */
  const uint8_t* elem_type;   /* [Etot]   OBGPU_BEAM2D_* / OBGPU_SPRING_*                                */
  const double* elem_EAI;     /* [3*Etot] E, A, Iz  (G = J = 0 as CBaseElementBeam.h:45-46); spring constants */
  const uint32_t* con_node;   /* [Ctot] */
  const uint8_t* con_dof;     /* [Ctot] OBGPU_DX..RZ */
  const double* con_val;      /* [Ctot] prescribed displacement (m_DoF_constraints)               */
  const uint32_t* nload_node; /* [NLtot] */
  const uint8_t* nload_dof;   /* [NLtot] */
  const double* nload_val;    /* [NLtot] addLoadAtDOF values (m_loads_at_each_dof)                */
  const uint32_t* eload_elem; /* [ELtot] element index (local) carrying the load (m_loads_on_beams) */
  const double* eload_q;      /* [ELtot] first parameter: q (N/m) | q_ini | P (N) | temperature increment */
  const double* eload_dir;    /* [3*ELtot] unit direction (CLoadDistributedUniform::dir); unused for TEMPERATURE */
  /* load cases: case-major overrides of the value arrays above (NULL = use the base values) */
  int64_t n_cases;              /* L >= 1 */
  const double* case_nload_val; /* NULL or [L][NLtot] */
  const double* case_eload_q;   /* NULL or [L][ELtot] */
  const double* case_elem_EAI;  /* NULL or [L][3*Etot] per-case section sweep */
  /* element-load kinds other than DISTRIB_UNIFORM (both NULL = every load is uniform) */
  const uint8_t* eload_type;    /* NULL or [ELtot] OBGPU_LOAD_*                                          */
  const double* eload_q2;       /* NULL or [ELtot] second parameter: q_end (TRIANGULAR), dist (CONCENTRATED) */
} obgpu_batch_desc;

/* Solvers.  OBGPU_BATCHED_LDLT stands for StaticSolverAlgorithm::Dense_LLT (solver_static.cpp:84-109): one dense
 * factorisation per system.  The two CG variants stand for ::Sparse_LLT (solver_static.cpp:110-129) on structures far
 * beyond what a dense-structured K_ff can hold: OBGPU_PCG_JACOBI (diagonal preconditioner) and OBGPU_PCG_MG (aggregation
 * multigrid V-cycle over rigid-body-mode coarse spaces, Chebyshev smoother — 100x fewer iterations on large lattices;
 * the sparse route's recommended method).  Both stop on the TRUE residual ||F_f - K_ff U_f||, recomputed at exit. */
enum { OBGPU_BATCHED_LDLT = 0, OBGPU_PCG_JACOBI = 1, OBGPU_PCG_MG = 2 };
typedef struct obgpu_solve_opts {
  int method;        /* OBGPU_BATCHED_LDLT (default) or a PCG variant            */
  int with_stress;   /* fuse postProcCalcStress into the solve kernel (default 1) */
  double rtol;       /* PCG: ||F_f - K_ff U_f||_2 <= max(rtol*||F_f||_2, 16 eps || |K_ff||U_f| ||_2) (default 1e-12);
                        the second term is the FP64 rounding floor of the residual itself */
  int max_iter;      /* PCG: default 20*sqrt(n_f)+1000                            */
  int flags;         /* OBGPU_SOLVE_* bits (default 0)                            */
} obgpu_solve_opts;
/* OBGPU_SOLVE_SKIP_FF: do not keep BuildProblemInfo::F_f (the assembled free-DOF load vector is an intermediate of the
 * fused kernel; storing it costs 8*n_f bytes per case of HBM writes).  OBGPU_RES_FF is then unavailable after the solve. */
/* OBGPU_SOLVE_COMPACT_STRESS: element end forces are written as [Etot][2][3] = (N, Vy, Mz) per face instead of the
 * reference's six-component FaceStress (types.h:78-157): Vz, Mx and My are structurally zero for every 2D element
 * (postproc_stress.cpp:95-121 only feeds the planar DOFs), i.e. half of every stress record.  Read it back with
 * OBGPU_RES_STRESS_COMPACT; the host mirrors expand it into FaceStress.  OBGPU_RES_STRESS needs a solve without the flag. */
enum { OBGPU_SOLVE_SKIP_FF = 1, OBGPU_SOLVE_COMPACT_STRESS = 2 };

/* sizes: obgpu_get_sizes fills out[0..11] with
 * S, L, Ntot, Etot, n_tot, nf_tot, nb_tot, T_tot (triplets), nnz_tot (K upper), Ctot, NLtot, ELtot */
enum { OBGPU_NSIZES = 12 };
/* int64 offset arrays [S+1] readable with obgpu_get_offsets */
enum { OBGPU_OFF_DOF = 0, OBGPU_OFF_FREE = 1, OBGPU_OFF_BOUNDED = 2, OBGPU_OFF_NNZ = 3, OBGPU_OFF_TRIPLET = 4 };
/* int32 symbolic arrays readable with obgpu_get_symbolic (all indices local to their structure) */
enum {
  OBGPU_SYM_DOF_NODE = 0,   /* [n_tot]    m_problem_DoFs[k].nodeId                                  */
  OBGPU_SYM_DOF_ID = 1,     /* [n_tot]    m_problem_DoFs[k].dof                                     */
  OBGPU_SYM_FREE = 2,       /* [nf_tot]   BuildProblemInfo::free_dof_indices                        */
  OBGPU_SYM_BOUNDED = 3,    /* [nb_tot]   BuildProblemInfo::bounded_dof_indices                     */
  OBGPU_SYM_DOF_INDEX = 4,  /* [6*Ntot]   m_problem_DoFs_inverse_list[node].dof_index[d], -1 unused */
  OBGPU_SYM_K_ROWPTR = 5,   /* [n_tot+S]  CSR(lower) row pointers == Eigen CSC(upper) outerIndexPtr */
  OBGPU_SYM_K_COL = 6,      /* [nnz_tot]  CSR(lower) columns     == Eigen CSC(upper) innerIndexPtr  */
  OBGPU_SYM_TRI_NNZ = 7,    /* [T_tot]    element->nnz map: triplet t (reference emission order) -> nnz slot */
  OBGPU_SYM_KFF_ROWPTR = 8, /* [nf_tot+S] CSR(lower) row pointers of the true-sparse K_ff, local to each structure   */
  OBGPU_SYM_KFF_COL = 9     /* [nnz(K_ff lower)] columns (free indices local to the structure)                      */
};
/* device/host result arrays */
enum {
  OBGPU_RES_UF = 0,      /* [L][nf_tot]        StaticSolveProblemInfo::U_f                          */
  OBGPU_RES_FB = 1,      /* [L][nb_tot]        StaticSolveProblemInfo::F_b                          */
  OBGPU_RES_STRESS = 2,  /* [L][Etot][2][6]    StressInfo::element_stress[e][face].{N,Vy,Vz,Mx,My,Mz} */
  OBGPU_RES_INFO = 3,    /* int32 [L][S]       0 ok; k>0: first non-positive pivot at free row k-1  */
  OBGPU_RES_FF = 4,      /* [L][nf_tot]        BuildProblemInfo::F_f                                */
  OBGPU_RES_UB = 5,      /* [nb_tot]           BuildProblemInfo::U_b                                */
  OBGPU_RES_KVAL = 6,    /* [LK][nnz_tot]      values of K (upper, CSC order); LK = L if case_elem_EAI else 1 */
  OBGPU_RES_ELEM_K = 7,  /* [LK][Etot][21]     compact global element blocks        */
  OBGPU_RES_STRESS_COMPACT = 8, /* [L][Etot][2][3] element_stress[e][face].{N,Vy,Mz} (OBGPU_SOLVE_COMPACT_STRESS) */
  /* BuildProblemInfo::K_ff / K_bf / K_bb (CFiniteElementProblem.h:86-91) after obgpu_assemble, as the reference builds them
   * (assemble.cpp:249-297): DENSE-STRUCTURED per structure, column-major, structures concatenated —
   *   K_ff n_f x n_f with only the LOWER triangle filled (what Eigen::LLT reads, :264-270),
   *   K_bf n_b x n_f full (:273-286),  K_bb n_b x n_b with only the UPPER triangle filled (:258-262).
   * For small structures (the whole batch must fit 2^28 doubles per view); large ones read the true-sparse K_ff below. */
  OBGPU_RES_KFF_DENSE = 9,    /* [LK][sum_s n_f(s)^2]      */
  OBGPU_RES_KBF_DENSE = 10,   /* [LK][sum_s n_b(s) n_f(s)] */
  OBGPU_RES_KBB_DENSE = 11,   /* [LK][sum_s n_b(s)^2]      */
  /* true-sparse K_ff: the entries of K (pattern bit-exact with Eigen's, OBGPU_SYM_K_*) whose row and column are free, lower
   * triangle, CSR over free indices; pattern through OBGPU_SYM_KFF_ROWPTR / OBGPU_SYM_KFF_COL */
  OBGPU_RES_KFF_CSR_VAL = 12  /* [LK][nnz(K_ff lower)]     */
};

/* device >= 0: CUDA ordinal (must be sm_100).  OBGPU_DEVICE_NONE: symbolic-only ctx — batch_upload,
 * symbolic, get_sizes/offsets/symbolic work (integer host phase K3); every compute call returns OBGPU_ECUDA. */
enum { OBGPU_DEVICE_NONE = -1 };
int obgpu_create(int device, obgpu_ctx** out);
void obgpu_destroy(obgpu_ctx*);
const char* obgpu_last_error(const obgpu_ctx*);
/* human-readable text for a status code; for OBGPU_ESINGULAR it is the reference's message */
const char* obgpu_status_string(int status);

/* pinned host memory for callers that want full-speed PCIe transfers */
int obgpu_host_alloc(void** p, size_t bytes);
int obgpu_host_free(void* p);

/* upload a batch (copies everything it needs; host buffers may be freed afterwards) */
int obgpu_batch_upload(obgpu_ctx*, const obgpu_batch_desc*);
/* The same for batches of MANY small structures (parameter sweeps, config "4M random frames"): the arrays go straight from
 * the caller's buffers (pinned ones from obgpu_host_alloc copy at PCIe speed) to the device, the library keeps no host
 * copy beyond the five offset arrays, and obgpu_symbolic then runs the DOF numbering / partition / slot phase
 * (CFiniteElementProblem.cpp:179-245, assemble.cpp:74-99) as device kernels, thread per structure, including the
 * validation obgpu_batch_upload does on the host (a malformed structure refuses the batch with the same messages).
 * The batched-LDLt route, summaries and downloads work on such a batch as they are; a call that needs the host-side
 * integer artefacts (obgpu_get_symbolic, the PCG route) downloads the batch once and runs the host phase.  Band-reducing
 * renumbering (meshed structures) is not applied in this mode. */
int obgpu_batch_upload_device(obgpu_ctx*, const obgpu_batch_desc*);
/* replace the per-case value arrays of the uploaded batch (same S and topology; L may change).
 * Host pointers; NULL = keep base values.  This is the per-step H2D traffic of a load-case sweep. */
int obgpu_update_cases(obgpu_ctx*, int64_t n_cases, const double* case_nload_val, const double* case_eload_q,
                       const double* case_elem_EAI);
/* K3: DOF numbering, free/bounded partition, K pattern, element->nnz map (once per topology) */
int obgpu_symbolic(obgpu_ctx*);
/* K1+K2+K4: element stiffness + deterministic segmented-reduction assembly of K (CSC upper values) */
int obgpu_assemble(obgpu_ctx*);
/* K5+K6(+K10): BC gather, factor/solve, reactions, optionally element end forces */
int obgpu_solve_static(obgpu_ctx*, const obgpu_solve_opts* /*NULL = defaults*/);
/* K10 alone (when with_stress was 0) */
int obgpu_stress(obgpu_ctx*);

int obgpu_get_sizes(const obgpu_ctx*, int64_t out[OBGPU_NSIZES]);
int obgpu_get_offsets(const obgpu_ctx*, int which, int64_t* out);
int obgpu_get_symbolic(const obgpu_ctx*, int which, int32_t* out);
/* copy one result array to a host buffer (blocking); returns OBGPU_ESINGULAR if any info != 0 when which==INFO */
int obgpu_download(obgpu_ctx*, int which, void* host_out);
/* number of elements obgpu_download(which) writes / obgpu_get_symbolic(which) writes, for the ids whose extent is not one of
 * obgpu_get_sizes' entries (the K_ff / K_bf / K_bb views); also valid for every other id */
int obgpu_result_count(obgpu_ctx*, int which, int64_t* count);
int obgpu_symbolic_count(obgpu_ctx*, int which, int64_t* count);
/* device pointer of a result array (valid until the next upload/update), for zero-copy consumers */
int obgpu_device_ptr(obgpu_ctx*, int which, void** dptr);
/* wait for all work queued on the ctx */
int obgpu_synchronize(obgpu_ctx*);

/* End-to-end convenience for load-case sweeps over the uploaded structures, HOST buffers in and out:
 * streams chunks H2D -> fused solve -> D2H on the ctx's copy/compute streams.  Any output may be NULL. */
int obgpu_solve_cases_host(obgpu_ctx*, int64_t n_cases, const double* case_nload_val, const double* case_eload_q,
                           const double* case_elem_EAI, double* U_f, double* F_b, double* stress, int32_t* info);
/* the same with OBGPU_SOLVE_* flags: with OBGPU_SOLVE_COMPACT_STRESS `stress` receives [n_cases][Etot][2][3] records
 * (N, Vy, Mz per face) — 144 fewer bytes per demo3_frame case over PCIe and HBM than the six-component record */
int obgpu_solve_cases_host_flags(obgpu_ctx*, int64_t n_cases, const double* case_nload_val, const double* case_eload_q,
                            const double* case_elem_EAI, double* U_f, double* F_b, double* stress, int32_t* info, int flags);

/* per-structure summaries {max|U_f|, sum F_b over DX dofs, sum F_b over DY dofs, info} = 4 doubles per system,
 * written to a caller-provided DEVICE buffer [L*S*4] (e.g. a torch tensor that is then NCCL-gathered) */
int obgpu_summaries_device(obgpu_ctx*, double* d_out);
/* Binds a device buffer [L][S][4] (32-byte aligned, NULL to unbind): later obgpu_solve_static calls leave the summaries
 * there as part of the solve (the thread-per-case kernel writes them from registers instead of a second pass over
 * U_f / F_b), and obgpu_summaries_device(ctx, same pointer) then returns at once.  Routes without the fused form
 * still produce them with the separate kernel at obgpu_summaries_device time. */
int obgpu_bind_summaries(obgpu_ctx*, double* d_out);
/* The all-gather of the summaries as part of the solve (the job's only inter-GPU exchange): bufs[0..n)
 * (n <= 8, 32-byte aligned) are device pointers valid in THIS process to the destination blocks [L][S][4] in other
 * GPUs' memory (peer / symmetric-memory mappings, already offset to this rank's slot).  Later solves store their summary
 * block into each of them as well, from the solve kernel itself over NVLink while it computes (bulk copies; plain
 * device-to-device copies behind the kernels on routes without the fused form).  The caller orders the ranks
 * afterwards (stream sync + barrier).  n = 0 unbinds; needs obgpu_bind_summaries first. */
int obgpu_bind_summary_peers(obgpu_ctx*, int n, double* const* bufs);

/* timings of the last call of each phase, in ms, measured with CUDA events on the ctx stream; section
 * names follow the reference's CTimeLogger entries (assemble.cpp:39-346, solver_static.cpp:53-121):
 * out[0] symbolic (host), [1] assembleProblem, [2] solveStatic.2_solve, [3] postProcCalcStress, [4] H2D, [5] D2H,
 * [6] kernel launches in the last solve/assemble call (count), [7] PCG iterations of the last solve */
int obgpu_get_timings(const obgpu_ctx*, double out[8]);
/* StaticSolverOptions::nonLinearIterative support (solver_static.cpp:139-193).  The reference restores the node poses after
 * its 40 passes, so the postProcCalcStress() that follows reads element LENGTHS from the restored poses while orientations,
 * displacements and fixed-end terms are those of the last pass.  node_xy [2*Ntot] = the poses to take the stress lengths
 * from (NULL clears it; every upload clears it); the next obgpu_solve_static runs on the batched-LDLt route with it. */
int obgpu_set_stress_length_poses(obgpu_ctx*, const double* node_xy);
/* the last PCG solve (sparse route): out[0] iterations summed over the structures, [1] worst TRUE relative residual
 * ||F_f - K_ff U_f||_2 / ||F_f||_2 recomputed from U_f at exit (the reference's SimplicialLLT branch has no such figure:
 * it does not even check info(), solver_static.cpp:117-128), [2] restarts from the true residual, [3] multigrid levels */
int obgpu_pcg_info(const obgpu_ctx*, double out[4]);
/* measurement aid: times `reps` back-to-back launches of one iteration kernel of the last PCG solve (first structure's
 * rows) with CUDA events on the ctx stream; *alg_bytes = the algorithmic HBM bytes of one launch.  The
 * results of the solve are left untouched. */
enum { OBGPU_PROF_SPMV = 0 /* k_spmv_dot */, OBGPU_PROF_MG_SPMV = 1 /* k_mg_spmv, fine level */ };
int obgpu_profile_kernel(obgpu_ctx*, int which, int reps, double* ms_avg, double* alg_bytes);
/* ---- CStructureProblem::mesh(): the caller-side step right before the path ----
 * Subdivides every beam of every structure into ceil(L / max_element_length) pieces and splits the element loads, exactly
 * as CStructureProblem_mesh.cpp:37-132, CBaseElementBeam.cpp:317-422 and the loads' meshLoad methods do (including the
 * reference's treatment of un-subdivided RA/AA/RD beams and of concentrated loads).  Host
 * side (integer/coordinate generation, once per topology); needs no device and no ctx.  Load cases are not carried over
 * (n_cases = 1 in the result).  obgpu_mesh_desc fills a batch description whose arrays live in the mesh object until
 * obgpu_mesh_destroy; obgpu_mesh_info gives MeshOutputInfo::element2elements / element2nodes as CSR over the INPUT
 * elements (offsets [Etot_in + 1], values local to the meshed structure). */
typedef struct obgpu_mesh obgpu_mesh;
int obgpu_mesh_create(const obgpu_batch_desc* in, double max_element_length, obgpu_mesh** out);
int obgpu_mesh_desc(const obgpu_mesh*, obgpu_batch_desc* out);
int obgpu_mesh_info(const obgpu_mesh*, const int64_t** e2e_off, const uint32_t** e2e, const int64_t** e2n_off,
                    const uint32_t** e2n);
void obgpu_mesh_destroy(obgpu_mesh*);

/* diagnostic: which kernel carried the last solve (the run-time specialised one, or the precompiled fallback when
 * NVRTC / the driver library could not be loaded); benchmark reports cite it.  Never NULL. */
const char* obgpu_kernel_info(const obgpu_ctx*);
/* the CUDA stream the ctx launches on (a cudaStream_t), so callers can record their own events on it */
int obgpu_get_stream(obgpu_ctx*, void** stream);

/* The offset scans of the symbolic phase with NO allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  From d_counts [n_structures][4] int32 in DEVICE memory (per structure: n DOFs, n_f free, n_b bounded,
 * one unused word; 16-byte aligned) it writes the three exclusive-scan offset arrays obgpu_get_offsets returns after
 * obgpu_symbolic (OBGPU_OFF_DOF / OFF_FREE / OFF_BOUNDED), each [n_structures + 1] int64 in DEVICE memory, on `stream`
 * (a cudaStream_t, may be 0) of device `device`; it returns after the launches, without synchronising.  The per-chunk
 * partial sums live in a file-scope device array of 96 KB that every GPU loading the library holds until exit.
 * Limit: 1 <= n_structures <= 4 194 304; larger counts return OBGPU_EINVAL with the reason in err (nullable, err_cap bytes).
 * Calls no cudaMalloc / cudaFree and no allocating routine.  All calls share the one scratch array: calls that overlap on
 * the GPU (different streams) race on it, so issue them on one stream.  The first call on a device looks the array's
 * address up (cudaGetSymbolAddress): make that call outside graph capture. */
int obgpu_scan_offsets_static(int device, int64_t n_structures, const int32_t* d_counts, int64_t* d_dof_off,
                              int64_t* d_free_off, int64_t* d_bnd_off, void* stream, char* err, size_t err_cap);

#ifdef __cplusplus
}
#endif
#endif
