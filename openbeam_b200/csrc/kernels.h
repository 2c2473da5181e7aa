// kernels.h — launcher declarations shared between the .cu files and the C-ABI layer.
#pragma once
#include "obgpu_internal.h"

namespace obgpu {

struct TinyArgs {
  const int32_t* prog_i;   // HOST pointers: the topology program is copied into the kernel parameter (constant bank)
  const double* prog_d;
  int n_ints, n_dbls;
  int64_t n_cases;
  const double* case_nl;   // [L][NL] or null
  const double* case_q;    // [L][EL] or null
  const double* case_eai;  // [L][3E] or null
  double* U_f;             // [L][nf]
  double* F_b;             // [L][nb]
  double* stress;          // [L][E][12] or null
  double* F_f;             // [L][nf] or null
  int32_t* info;           // [L]
  int32_t* fail;           // [1] set to 1 when any system has info != 0
  double* summ;            // [L][4] or null: per-case summaries written by the same kernel (obgpu_bind_summaries)
  double* summ_peer[8];    // further copies of the same block (peer-mapped buffers of other GPUs: the all-gather of the
  int n_summ_peer;         // summaries done by the solve kernel's own stores over NVLink, obgpu_bind_summary_peers)
  const uint8_t* bnd_dofid;                 // HOST [nb]: DoF_index of each bounded DOF (used with summ)
  const struct SummaryArgs* summ_fallback;  // HOST: arguments of the separate k_summaries launch for kernels without it
  int stress_compact;      // stress records are [E][2][3] = (N, Vy, Mz) per face (OBGPU_SOLVE_COMPACT_STRESS); a launcher
                           // whose kernel only writes the full record answers cudaErrorNotYetImplemented
};
cudaError_t launch_tiny_fused(int nf, int n, const TinyArgs& a, int sm_count, cudaStream_t st, const char** which = nullptr);
cudaError_t launch_tiny_jit(int device, const TinyArgs& a, int sm_count, cudaStream_t st, const char** which = nullptr);  // tiny_jit.cu

// Per-structure record of the batched kernels (one 64-byte line per structure).
struct __align__(16) FrameMeta {
  int64_t n0, e0, nl0;   // first node / element / node-load of the structure in the batch arrays
  int64_t f0, b0;        // first free / bounded DOF in the result arrays (per case)
  int32_t N, E, NL, n, nf, nb, hb, pad;
};
static_assert(sizeof(FrameMeta) == 80, "FrameMeta layout");

// Batch arrays on the device (SoA, see obgpu_batch_desc) + symbolic slots.
struct FrameArgs {
  int64_t S, L;                 // structures x load cases; system = c*S + s
  int64_t Ntot, Etot, NLtot, ELtot, nf_tot, nb_tot;
  const FrameMeta* meta;        // [S]
  const double2* node_xy;       // [Ntot]
  const uint2* elem_nodes;      // [Etot] local node ids
  const uint8_t* elem_type;     // [Etot]
  const double* elem_EAI;       // [3*Etot]
  const int32_t* node_slot;     // [3*Ntot]
  const int32_t* nload_slot;    // [NLtot]
  const double* nload_val;      // [NLtot]
  const int32_t* el_ptr;        // [Etot+1]
  const int32_t* el_order;      // [ELtot]
  const double* eload_q;        // [ELtot]
  const double* eload_dir;      // [3*ELtot]
  const uint8_t* eload_type;    // [ELtot] OBGPU_LOAD_*, or null when every load is DISTRIB_UNIFORM
  const double* eload_q2;       // [ELtot] second load parameter (with eload_type)
  const double* U_b;            // [nb_tot] prescribed displacements (nodal frames), or null when all are zero
  const double* node_rot;       // [Ntot] nodal frame yaw (radians), or null when every node frame is the identity
  const double2* node_xy_len;   // [Ntot] or null: node positions the element LENGTHS of postProcCalcStress are taken from
                                // (obgpu_set_stress_length_poses: the reference's nonLinearIterative restores the node poses
                                // before the stresses are computed, while orientations stay those of the last pass)
  const double* case_nl;        // [L][NLtot] or null
  const double* case_q;         // [L][ELtot] or null
  const double* case_eai;       // [L][3*Etot] or null
  double* U_f;                  // [L][nf_tot]
  double* F_b;                  // [L][nb_tot]
  double* F_f;                  // [L][nf_tot] or null
  double* stress;               // [L][Etot][12] or null
  int32_t* info;                // [L][S]
  int32_t* fail;                // [1]
  double* erec;                 // [Etot][FRAME_EREC] element records of the three-kernel path, or null: fused kernel
  // size class served by this launch: structures order[o0 .. o1), whose banded K_ff, n_f, n_b and half bandwidth
  // are bounded by the class maxima below (shared memory per warp is sized from them)
  const int32_t* order;         // [S] structure ids sorted by size class
  const int32_t* free_inv;      // [nf_tot] or null: free index (reference numbering) of each internal (band-reduced) free slot
  int64_t o0, o1;
  int max_band;                 // max over the class of frame_ld(hb)*nf   (doubles)
  int max_nf, max_nb, max_hb;
};
constexpr int FRAME_EREC = 30;  // doubles per element record: 21 stiffness entries + 6 equivalent loads + 6 int32 slots
cudaError_t launch_frame_fused(const FrameArgs& a, int sm_count, cudaStream_t st);
cudaError_t launch_frame_pre(const FrameArgs& a, int sm_count, cudaStream_t st);                // FrameArgs::erec != null
cudaError_t launch_frame_post(const FrameArgs& a, int max_nb, int sm_count, cudaStream_t st);
size_t frame_smem_per_warp(int max_band, int max_nf, int max_nb);
constexpr int FRAME_WARPS = 4;  // warps per CTA of k_frame_fused

// Device view of the sparse route (global indices over the batch).
struct SparseDev {
  int64_t Etot, ELtot, nnz, nf, nb, nnz_ff, nnz_bf;
  const uint2* elem_nodes; const uint8_t* elem_type; const double2* node_xy;
  const int32_t* elem_node0;   // [Etot] first node of the element's structure
  const int32_t* elem_f0;      // [Etot] first free DOF of the element's structure
  const int32_t* elem_b0;      // [Etot] first bounded DOF of the element's structure
  const int32_t* node_f0;      // [Ntot] first free DOF of the node's structure
  const double* node_rot;      // [Ntot] nodal frame yaw, or null when every node frame is the identity
  const double* U_b;           // [nb] prescribed displacements (nodal frames), or null when all are zero
  int64_t Ntot;
  const int32_t *bb_rowptr, *bb_col, *bb_src;  // K_bb (upper, as the reference stores it): rows / cols = bounded indices
  double* bb_val;
  const int32_t* node_slot; const int32_t* el_ptr; const int32_t* el_order;
  const int32_t* eload_elem_g; // [ELtot] global element of each element load (caller slot order)
  const double* eload_dir;
  const uint8_t* eload_type;   // null when every load is DISTRIB_UNIFORM
  const double* eload_q2;
  const int32_t *blk_elem_ptr, *blk_elems;  // fused assembly plan (null: two-kernel K1 + K4)
  const uint16_t* seg_loc;
  int asm_max_loc;
  const int32_t *seg_ptr, *seg_src, *sl_ptr, *ff_col, *ff_src, *ff_diag, *bf_rowptr, *bf_col, *bf_src, *fseg_ptr, *fseg_src;
  double *elem_k, *kval, *ff_val, *bf_val, *eload_g, *Eq;
};
cudaError_t launch_elem_stiff(const SparseDev& d, const double* eai, cudaStream_t st);
cudaError_t launch_assemble(const SparseDev& d, cudaStream_t st);
cudaError_t launch_assemble_fused(const SparseDev& d, const double* eai, cudaStream_t st);
cudaError_t launch_bc_gather(const SparseDev& d, cudaStream_t st);
cudaError_t launch_gather(int64_t n, const int32_t* src, const double* in, double* out, cudaStream_t st);  // out[i] = in[src[i]] (-1: 0)
// FP32 copy (paired layout, 10 floats per block) of a 3x3 block-SELL matrix of n_blocks block slots whose FP64 gather map is src
cudaError_t launch_gather_blk_f32(int64_t n_blocks, const int32_t* src, const double* in, float* out, cudaStream_t st);
cudaError_t launch_load_vector(const SparseDev& d, const double* nload_val, const double* q, const double* eai, double* F_f,
                               cudaStream_t st);
cudaError_t launch_stress(const SparseDev& d, const double* eai, const double* q, const double* U_f, double* stress,
                          cudaStream_t st);

// Jacobi-PCG on rows [r0, r1) of the K_ff CSR (one structure).  Device scalars live in `scal` (16 doubles).
struct PcgArgs {
  int32_t r0, r1;
  const int32_t *slptr, *col;  // K_ff in SELL-32 (SparsePlan): slice offsets, column of every stored entry
  const double* val;
  const int32_t *bslptr, *bcol;  // the same matrix in 3x3 block-SELL (BlockSell, obgpu_internal.h), or null
  const double* bval;
  const double* minv;   // [nf] inverse diagonal
  const double* b;      // rhs (F_f)
  double *x, *r, *z, *p, *ap;
  double* scal;         // [0..1] rz ping-pong, [2] pAp, [3] rr, [4] bb
  double* partial;      // [2 * max blocks]
  unsigned int* counter;
  int ctas_per_sm;      // grid cap of the iteration kernels (0 = 8): small when several structures' PCGs share the GPU
  const struct MgJob* mg;  // HOST pointer: multigrid V-cycle as the preconditioner (OBGPU_PCG_MG); null = Jacobi
};
// ---- multigrid preconditioner (kernels_mg.cu; plan in symbolic_mg.cpp) ----
// One level of one structure: rows [r0, r1) of the level's SELL-32 matrix, its work vectors, and the transfer to the
// next coarser level (arrays indexed by THIS level's rows; member lists indexed by the coarser level's aggregates).
struct MgLevelDev {
  int32_t r0, r1;
  const int32_t *slptr, *col;
  const double *val, *minv;
  const int32_t *bslptr, *bcol;   // 3x3 block-SELL view used by the SpMV-class kernels when present (else null)
  const double* bval;
  const float* bval32;            // FP32 copy of bval: when set, the V-cycle's kernels stream it instead (the preconditioner
                                  // needs no more; the PCG's own SpMV and every vector stay FP64)
  double *x, *b, *r, *d0, *d1;
  const int32_t* f_agg; const uint8_t* f_kind; const double* f_w;
  const int32_t *mem_ptr, *mem;
  double c0, c1, c2;    // Chebyshev(2) smoother on [lambda_max/8, lambda_max] of D^-1 A: d0 = c0 D^-1 r, d1 = c1 d0 + c2 D^-1 r'
};
constexpr int MG_MAX_LEVELS = 16;
struct MgJob {
  int nlev = 1;
  bool dense = false;          // coarsest level applied as an explicit inverse
  const double* cinv = nullptr;  // [nc*nc] row-major inverse of the coarsest matrix
  double omega = 1.4;          // over-correction of the (unsmoothed-aggregation) coarse-grid step
  MgLevelDev lv[MG_MAX_LEVELS];
};
// z = M r for the job's structure (level-0 b = r, x = z); the last kernel also leaves r.z in *dot_out (partial/counter:
// the job's reduction scratch)
cudaError_t launch_mg_vcycle(const MgJob& J, double* dot_out, double* partial, unsigned int* counter, int sm_count,
                             int ctas_per_sm, cudaStream_t st);
// numeric setup of level `coarse` from level `fine` over the whole batch: A_c = P^T A P
struct MgGalerkinArgs {
  int64_t n_coarse;
  const int32_t *f_slptr, *f_col; const double* f_val;
  const int32_t* f_agg; const uint8_t* f_kind; const double* f_w;
  const int32_t *mem_ptr, *mem;
  const int32_t *c_slptr, *c_col; double* c_val;
};
cudaError_t launch_mg_galerkin(const MgGalerkinArgs& g, cudaStream_t st);
// minv = 1/diag (0 where the diagonal is not positive) and lam = max_i sum_j |a_ij| / a_ii (Gershgorin bound of D^-1 A)
cudaError_t launch_mg_diag(int64_t r0, int64_t r1, const int32_t* slptr, const int32_t* col, const double* val, double* minv,
                           double* lam, cudaStream_t st);
// dense explicit inverse of rows [r0, r0 + nc) of a SELL-32 matrix (cooperative launch; SPD, no pivoting, null rows
// skipped); prow = scratch of 2 * (nc + 2) doubles
cudaError_t launch_mg_dense_inverse(int32_t r0, int32_t nc, const int32_t* slptr, const int32_t* col, const double* val,
                                    double* inv, double* prow, cudaStream_t st);

cudaError_t launch_spmv_only(const PcgArgs& a, int sm_count, cudaStream_t st);              // k_spmv_dot alone (profiling)
cudaError_t launch_mg_cheb_only(const MgLevelDev& L, int sm_count, cudaStream_t st);         // k_mg_spmv<CHEB> alone (profiling)
cudaError_t launch_pcg_init(const PcgArgs& a, int sm_count, cudaStream_t st);
// r = b - A x on the job's rows (true residual, not the recurrence), scal[3] = r.r
cudaError_t launch_pcg_true_residual(const PcgArgs& a, int sm_count, cudaStream_t st);
// restart from the current r: z = M r, p = z
cudaError_t launch_pcg_restart(const PcgArgs& a, int it, int sm_count, cudaStream_t st);
cudaError_t launch_pcg_iter(const PcgArgs& a, int it, int sm_count, cudaStream_t st);
cudaError_t launch_inv_diag(int64_t nf, const int32_t* ff_diag, const double* ff_val, double* minv, cudaStream_t st);
cudaError_t launch_reactions(const SparseDev& d, const double* U_f, double* F_b, cudaStream_t st);
// rhs = F_f - K_bf^T U_b (solver_static.cpp:37-43), deterministic (bounded rows applied one after the other)
cudaError_t launch_rhs_prescribed(const SparseDev& d, const double* F_f, double* rhs, cudaStream_t st);

// ---- device-side symbolic phase (kernels_symbolic.cu) ----
enum { SYM_E_ELEM_NODE = 1, SYM_E_UNINIT, SYM_E_ELEM_TYPE, SYM_E_ZERO_LEN, SYM_E_NONPLANAR, SYM_E_CONSTRAINT, SYM_E_NLOAD,
       SYM_E_ELOAD, SYM_E_LOAD_DIR, SYM_E_LOAD_ELEM, SYM_E_NO_DOFS };
struct SymDev {
  int64_t S;
  const int64_t *node_off, *elem_off, *con_off, *nload_off, *eload_off;  // [S+1]
  const double2* node_xy; const double* node_rotz;                         // node_rotz null = identity frames
  const uint2* elem_nodes; const uint8_t* elem_type; const double* elem_EAI;
  const uint32_t* con_node; const uint8_t* con_dof; const double* con_val;
  const uint32_t* nload_node; const uint8_t* nload_dof;
  const uint32_t* eload_elem; const uint8_t* eload_type; const double* eload_dir; const double* eload_q2;  // type null = uniform
  uint8_t* node_mask;      // [Ntot] bits 0..2: DX, DY, RZ exist; bits 3..5: constrained
  double* node_cval;       // [3*Ntot] prescribed values of the constrained DOFs
  int32_t* counts;         // [S][4] n, n_f, n_b, half bandwidth
  int64_t *dof_off, *free_off, *bnd_off;  // [S+1]
  int32_t *node_slot, *nload_slot, *el_ptr, *el_order, *el_cursor;
  FrameMeta* meta;
  double* U_b; uint8_t* bnd_dofid;
  int32_t* error;          // [0] first error code (SYM_E_*), [1] its structure, [2] any U_b != 0
  int64_t* scan_partial;   // [3 * (chunks + 1)] scratch of the scans; the last triple = totals (n, n_f, n_b)
  uint8_t* cls;            // [S] size class of every structure
  int32_t* cls_stats;      // [FRAME_NCLS][5] count, max_band, max_nf, max_nb, max_hb (zeroed by the caller)
};
cudaError_t launch_symbolic_device(const SymDev& d, cudaStream_t st);
cudaError_t launch_symbolic_order(int64_t S, const uint8_t* cls, int32_t* cursor, int32_t* order, cudaStream_t st);

// Size classes of the warp-per-structure kernel: a structure's shared-memory need (band + vectors, in doubles) bucketed
constexpr int FRAME_NCLS = 18;
// Row stride (doubles) of the banded K_ff of k_frame_fused: hb + 1 entries per row.  (Rounding up to an even stride, so
// that a walk down a column hits 16 distinct banks, measured no faster and costs shared memory.)
__host__ __device__ inline int frame_ld(int hb) { return hb + 1; }
__host__ __device__ inline int frame_class_of(int hb, int nf, int nb) {
  const long long need = (long long)frame_ld(hb) * nf + 2LL * nf + 2LL * nb;
  const long long lim[FRAME_NCLS - 1] = {256, 384, 512, 640, 768, 1024, 1280, 1536, 2048, 2560, 3072, 4096, 6144, 8192, 12288,
                                          16384, 24576};
  int k = 0;
  while (k < FRAME_NCLS - 1 && need > lim[k]) k++;
  return k;
}

struct SummaryArgs {
  int64_t L, S, nf_tot, nb_tot;
  const int64_t* free_off;   // [S+1]
  const int64_t* bnd_off;    // [S+1]
  const uint8_t* bnd_dofid;  // [nb_tot] DoF_index of each bounded DOF
  const double* U_f;         // [L][nf_tot]
  const double* F_b;         // [L][nb_tot]
  const int32_t* info;       // [L][S]
  double* out;               // [L][S][4], 32-byte aligned
};
cudaError_t launch_summaries(const SummaryArgs& a, int sm_count, cudaStream_t st);
// full [n][2][6] element stress records -> compact [n][2][3] (N, Vy, Mz): for kernels that only write the full record
cudaError_t launch_compact_stress(const double* full, double* compact, int64_t n_elem_records, cudaStream_t st);

}  // namespace obgpu
