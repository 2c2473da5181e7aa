// ctx.h — the opaque obgpu_ctx.
#pragma once
#include <vector>

#include "kernels.h"

struct obgpu_ctx {
  int device = 0;
  int sm_count = 148;
  cudaStream_t stream = nullptr;
  static constexpr int NSLOT = 3;
  cudaStream_t slot_stream[NSLOT] = {nullptr, nullptr, nullptr};
  cudaEvent_t ev[8] = {};
  std::string err;

  obgpu::HostBatch hb;
  obgpu::Symbolic sym;
  obgpu::TinyProgram prog;
  bool uploaded = false;
  // device batch mode (obgpu_batch_upload_device, capi_devsym.cu): the arrays live in HBM only, K3 runs as kernels
  bool device_batch_mode = false, host_batch_valid = true;
  bool any_ub = false;                            // some prescribed displacement is non-zero
  obgpu::DevBuf<double> d_node_xy_len;            // obgpu_set_stress_length_poses
  bool have_xy_len = false;
  obgpu::DevBuf<int64_t> d_off_node, d_off_elem, d_off_con, d_off_nload, d_off_eload, d_dof_off;  // [S+1]
  obgpu::DevBuf<uint32_t> d_con_node, d_nload_node, d_eload_elem;
  obgpu::DevBuf<uint8_t> d_con_dof, d_nload_dof, d_sym_mask;
  obgpu::DevBuf<double> d_con_val, d_sym_cval;
  obgpu::DevBuf<int32_t> d_sym_counts, d_sym_cursor, d_sym_error, d_sym_cls_stats;
  obgpu::DevBuf<int64_t> d_sym_partial;
  obgpu::DevBuf<uint8_t> d_sym_cls;
  bool host_offsets_valid = true;                // device batch mode: dof/free/bounded offsets are fetched on demand
  int64_t L = 1;
  bool has_case_nl = false, has_case_q = false, has_case_eai = false;

  obgpu::DevBuf<double> d_case_nl, d_case_q, d_case_eai;
  obgpu::DevBuf<int32_t> d_prog_i;
  obgpu::DevBuf<double> d_prog_d;
  obgpu::DevBuf<double> d_erec;   // element records of the three-kernel frame path (kernels_frame.cu)
  obgpu::DevBuf<double> d_Uf, d_Fb, d_stress, d_stress_c, d_Ff, d_Ub;  // d_stress_c: compact (N,Vy,Mz) records
  obgpu::DevBuf<int32_t> d_info;
  obgpu::DevBuf<int64_t> d_free_off, d_bnd_off;  // [S+1]
  obgpu::DevBuf<uint8_t> d_bnd_dofid;            // [nb_tot]
  obgpu::DevBuf<int32_t> d_frame_slot, d_frame_nlslot;   // node / nodal-load slots in the band-reduced numbering
  obgpu::DevBuf<int32_t> d_frame_inv;             // [nf_tot] band-reducing renumbering of the batched-LDLt route (or unused)
  bool frame_renumbered = false;
  obgpu::DevBuf<int32_t> d_fail;  // [1] OR of per-system failures of the last solve
  bool solved = false, have_stress = false, have_stress_c = false, have_ff = true;
  int last_method = 0;   // method of the last solve (obgpu_stress re-runs the same route)
  double last_rtol = 1e-12; int last_max_iter = 0;
  std::vector<uint8_t> h_bnd_dofid;               // host copy of d_bnd_dofid
  double* d_summ = nullptr;                       // obgpu_bind_summaries target (caller's device buffer) or null
  double* summ_peer[8] = {};                      // obgpu_bind_summary_peers
  int n_summ_peer = 0;
  const double* summ_written = nullptr;           // where the last solve's own kernel left the summaries, if it did
  const char* kernel_info = "";  // dominant kernel of the last solve (obgpu_kernel_info)
  // general path: the batch arrays + symbolic slots on the device
  bool dev_batch = false;
  obgpu::DevBuf<double> d_node_rot, d_eload_q2;
  obgpu::DevBuf<uint8_t> d_eload_type;
  obgpu::DevBuf<double> d_node_xy, d_elem_EAI, d_nload_val, d_eload_q, d_eload_dir;
  obgpu::DevBuf<uint32_t> d_elem_nodes;
  obgpu::DevBuf<uint8_t> d_elem_type;
  obgpu::DevBuf<int32_t> d_node_slot, d_nload_slot, d_el_ptr, d_el_order;
  obgpu::DevBuf<obgpu::FrameMeta> d_meta;
  int max_band = 0, max_nf = 0, max_nb = 0;
  // size classes of the warp-per-structure kernel: structures sorted by shared-memory need
  struct FrameClass {
    int64_t o0, o1;
    int max_band, max_nf, max_nb, max_hb;
  };
  std::vector<FrameClass> frame_classes;
  obgpu::DevBuf<int32_t> d_frame_order;
  // sparse route (capi_sparse.cu)
  obgpu::SparsePlan plan;
  bool dev_plan = false, assembled = false;
  int64_t LK = 1;  // number of assembled K copies (L when sections vary per case, else 1)
  obgpu::DevBuf<int32_t> d_seg_ptr, d_seg_src, d_sl_ptr, d_ff_col, d_ff_src, d_ff_diag, d_bf_rowptr, d_bf_col, d_bf_src,
      d_fseg_ptr, d_fseg_src, d_elem_node0, d_elem_f0, d_elem_b0, d_node_f0, d_eload_elem_g, d_bb_rowptr, d_bb_col, d_bb_src;
  obgpu::DevBuf<double> d_elem_k, d_kval, d_ff_val, d_bf_val, d_bb_val, d_rhs, d_eload_g, d_Eq, d_scal, d_partial;
  obgpu::DevBuf<double> d_pcgvec;  // p | ap | r | z | minv in one slab: the L2 persistence window of the PCG loop
  obgpu::DevBuf<unsigned int> d_counter;
  obgpu::DevBuf<int32_t> d_blk_elem_ptr, d_blk_elems;
  obgpu::DevBuf<uint16_t> d_seg_loc;
  // multigrid preconditioner of the sparse route (OBGPU_PCG_MG): plan + per-level device arrays (index 0 unused:
  // level 0 is K_ff itself)
  obgpu::MgPlan mg;
  bool dev_mg = false;
  struct MgLevelBufs {
    obgpu::DevBuf<int32_t> sl_ptr, col, f_agg, mem_ptr, mem;
    obgpu::DevBuf<uint8_t> f_kind;
    obgpu::DevBuf<double> f_w, val, minv, vec;  // vec: x | b | r | d0 | d1
    obgpu::DevBuf<int32_t> bsl_ptr, bcol, bmap;  // 3x3 block-SELL view (empty when the level has none)
    obgpu::DevBuf<double> bval;
    obgpu::DevBuf<float> bval32;   // FP32 copy streamed by the V-cycle's kernels
    int64_t nblk = 0;
  };
  std::vector<MgLevelBufs> mg_lev;
  obgpu::DevBuf<double> d_mg_lam, d_mg_cinv, d_mg_prow;
  obgpu::DevBuf<int32_t> d_ff_bslptr, d_ff_bcol, d_ff_bmap;   // K_ff in 3x3 block-SELL (SparsePlan::ff_blk), when it has one
  obgpu::DevBuf<float> d_ff_bval32;  // FP32 copy of d_ff_bval for the V-cycle's fine-level kernels
  obgpu::DevBuf<double> d_ff_bval;   // Gershgorin bounds per level; explicit coarsest inverses, one per job slot
  obgpu::MgJob mg_job[8];
  obgpu::PcgArgs prof_args;          // obgpu_profile_kernel: iteration arguments of the last PCG solve's first structure
  obgpu::MgLevelDev prof_lv0;
  bool prof_valid = false, prof_mg = false;
  bool elem_k_valid = false;  // d_elem_k holds the element blocks of the current batch (only the two-kernel path fills it)
  double* h_scal = nullptr;  // pinned [8 * NPCG]
  static constexpr int NPCG = 8;                  // PCGs of different structures in flight (sparse route)
  cudaStream_t pcg_stream[NPCG] = {};
  cudaEvent_t pcg_ev = nullptr;
  double pcg_relres = 0, pcg_restarts = 0, pcg_levels = 1;

  // chunk slots of obgpu_solve_cases_host
  struct Slot {
    obgpu::DevBuf<double> nl, q, eai, Uf, Fb, stress, stress_full;
    obgpu::DevBuf<int32_t> info;
  } slot[NSLOT];

  double t_symbolic_ms = 0;
  bool t_have[6] = {false, false, false, false, false, false};  // event pairs recorded
  double launches = 0, pcg_iters = 0;
};

namespace obgpu {
int fail(obgpu_ctx* c, int code, const std::string& msg);
int cuda_fail(obgpu_ctx* c, cudaError_t e, const char* where);
SummaryArgs summary_args(obgpu_ctx* c, double* d_out);  // capi_solve.cu
}  // namespace obgpu

#define OB_CUDA(c, call)                                             \
  do {                                                               \
    cudaError_t _e = (call);                                         \
    if (_e != cudaSuccess) return obgpu::cuda_fail(c, _e, #call);    \
  } while (0)
// select the ctx's device; a symbolic-only ctx (OBGPU_DEVICE_NONE) skips it
#define OB_DEVICE(c)                                                      \
  do {                                                                    \
    if ((c)->device != OBGPU_DEVICE_NONE) OB_CUDA(c, cudaSetDevice((c)->device)); \
  } while (0)
// compute entry points: no device => fail loudly (there is no CPU fallback)
#define OB_NEED_DEVICE(c)                                                                                  \
  do {                                                                                                     \
    if ((c)->device == OBGPU_DEVICE_NONE)                                                                  \
      return obgpu::fail(c, OBGPU_ECUDA, "symbolic-only ctx: this call needs an sm_100 device (no CPU fallback)"); \
    OB_CUDA(c, cudaSetDevice((c)->device));                                                                \
  } while (0)
