// obgpu_internal.h — internal declarations of libobgpu (not part of the ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <limits.h>

#include <string>
#include <vector>

#include "../../include/obgpu.h"

namespace obgpu {

// Compact per-element global stiffness ("ELEM_K", 21 doubles): 2D beam elements only touch the
// node DOFs {DX,DY,RZ} (= 6x6 indices 0,1,5), so of the three 6x6 blocks the reference builds
// (CElement.h:37-41) only 3x3 sub-blocks are ever non-zero:
//   [0..5]   K11 upper: (0,0)(0,1)(0,5)(1,1)(1,5)(5,5)
//   [6..11]  K22 upper, same order
//   [12..20] K12 full row-major over {0,1,5}x{0,1,5}
constexpr int ELEM_K = 21;
constexpr int TINY_MAX_NF = 12;   // thread-per-system kernel: K_ff lives in registers up to this size
constexpr int TINY_MAX_N = 48;
constexpr int TINY_MAX_E = 32;
constexpr int TINY_MAX_NODES = 32, TINY_MAX_NL = 48, TINY_MAX_EL = 32;
constexpr int32_t SLOT_NONE = INT32_MIN;  // DOF absent from the problem

template <class T>
struct DevBuf {  // owning device buffer (grow-only)
  T* p = nullptr;
  size_t cap = 0;
  ~DevBuf() { release(); }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
  cudaError_t reserve(size_t n) {
    if (n <= cap) return cudaSuccess;
    release();
    if (n == 0) return cudaSuccess;
    cudaError_t e = cudaMalloc((void**)&p, n * sizeof(T));
    if (e == cudaSuccess) cap = n;
    return e;
  }
  cudaError_t upload(const T* h, size_t n, cudaStream_t s) {
    cudaError_t e = reserve(n);
    if (e != cudaSuccess || n == 0) return e;
    return cudaMemcpyAsync(p, h, n * sizeof(T), cudaMemcpyHostToDevice, s);
  }
  cudaError_t upload(const std::vector<T>& v, cudaStream_t s) { return upload(v.data(), v.size(), s); }
};

// Host copy of the uploaded batch (CSR-of-structures SoA; see obgpu_batch_desc)
struct HostBatch {
  int64_t S = 0, L = 1;
  std::vector<int64_t> node_off, elem_off, con_off, nload_off, eload_off;
  std::vector<double> node_xy, node_rotz, elem_EAI, con_val, nload_val, eload_q, eload_dir;
  std::vector<double> eload_q2;      // [ELtot] second load parameter (zeros when every load is uniform)
  std::vector<uint8_t> eload_type;   // [ELtot] OBGPU_LOAD_*
  bool has_load_kinds = false;       // any load other than DISTRIB_UNIFORM
  std::vector<uint32_t> elem_nodes, con_node, nload_node, eload_elem;
  std::vector<uint8_t> elem_type, con_dof, nload_dof;
  bool has_rotz = false;
  int64_t Ntot() const { return node_off.empty() ? 0 : node_off.back(); }
  int64_t Etot() const { return elem_off.empty() ? 0 : elem_off.back(); }
  int64_t Ctot() const { return con_off.empty() ? 0 : con_off.back(); }
  int64_t NLtot() const { return nload_off.empty() ? 0 : nload_off.back(); }
  int64_t ELtot() const { return eload_off.empty() ? 0 : eload_off.back(); }
};

// Result of the symbolic phase K3 (all integer; once per topology).
struct Symbolic {
  bool valid = false;
  std::vector<int64_t> dof_off, free_off, bnd_off, nnz_off, tri_off;  // [S+1]
  std::vector<int32_t> dof_node, dof_id;     // [n_tot]
  std::vector<int32_t> dof_type;             // [n_tot] >=0: free index, <0: -(bounded index)-1
  std::vector<int32_t> free_idx, bnd_idx;    // [nf_tot], [nb_tot]
  std::vector<int32_t> dof_index;            // [6*Ntot]
  std::vector<int32_t> k_rowptr;             // [n_tot+S]
  std::vector<int32_t> k_col;                // [nnz_tot]
  std::vector<int32_t> tri_nnz;              // [T_tot] triplet -> nnz slot (local to structure)
  std::vector<int32_t> tri_src;              // [T_tot] triplet -> e_local*21+cid, or -1 (structural zero)
  std::vector<double> U_b;                   // [nb_tot]
  // per-node compact slots for the planar DOFs (DX,DY,RZ): >= 0 free index, < 0 && != SLOT_NONE: -(bounded index)-1
  std::vector<int32_t> node_slot;            // [3*Ntot]
  std::vector<int32_t> nload_slot;           // [NLtot] same encoding for the loaded DOF
  std::vector<int32_t> el_ptr;               // [Etot+1] CSR over GLOBAL element index -> el_order
  std::vector<int32_t> el_order;             // [ELtot] global element-load slot, reference multimap order
  std::vector<int32_t> half_bw;              // [S] half bandwidth of K_ff in free numbering
  bool has_pattern = false;
  // K4 segmented-reduction program: for global nnz slot g, sources seg_src[seg_ptr[g]..seg_ptr[g+1])
  // (global element index*21+cid, or -1), in the reference's triplet (= element) order.
  std::vector<int64_t> seg_ptr;              // [nnz_tot+1]
  std::vector<int64_t> seg_src;              // [T_tot]
  int64_t n_tot() const { return dof_off.empty() ? 0 : dof_off.back(); }
  int64_t nf_tot() const { return free_off.empty() ? 0 : free_off.back(); }
  int64_t nb_tot() const { return bnd_off.empty() ? 0 : bnd_off.back(); }
  int64_t nnz_tot() const { return nnz_off.empty() ? 0 : nnz_off.back(); }
  int64_t T_tot() const { return tri_off.empty() ? 0 : tri_off.back(); }
};

// "Program" of one shared topology for the thread-per-system kernel (uploaded as one int32 blob +
// one double blob).  Layout documented in kernels_tiny.cu.
struct TinyProgram {
  bool valid = false;
  bool has_load_kinds = false;  // element loads other than uniform: only the run-time specialised kernel serves them
  int N = 0, E = 0, n = 0, nf = 0, nb = 0, NL = 0, EL = 0;
  std::vector<int32_t> ints;
  std::vector<double> dbls;
};

// 3x3 block-SELL view of a scalar SELL-32 matrix whose rows come in triples (block row I = rows 3I..3I+2): slices of 32
// block rows, block k of lane l at (sl_ptr[slice] + 32 k + l), its 9 values (row-major a*3+b) at
// 9*sl_ptr[slice] + (9 k + v)*32 + l — every load of a warp reads 32 consecutive words.  One column index per 9 values:
// 8.44 instead of 12 bytes per stored entry.  `map` gathers the values out of the scalar SELL value array (-1 = zero).
struct BlockSell {
  bool valid = false;
  std::vector<int32_t> sl_ptr;   // [n/96 + 1] in blocks
  std::vector<int32_t> col;      // [blocks] block column
  std::vector<int32_t> map;      // [9 * blocks] position in the scalar SELL arrays, -1 = structural zero / padding
};
// n rows (multiple of 96 unless `ragged_ok`), scalar pattern in SELL-32; false when the fill of the blocks is below 70 %
bool build_block_sell(int64_t n, const std::vector<int32_t>& sl_ptr, const std::vector<int32_t>& col,
                      const std::vector<int32_t>* src /* -1 = padding entry, may be null */, BlockSell& out);

// Plan of the sparse route (K4 segmented reduction, K5 BC gathers, PCG): all indices GLOBAL over the batch.
struct SparsePlan {
  bool valid = false;
  std::vector<int32_t> seg_ptr, seg_src;      // [nnz_tot+1], [T_tot]: K slot <- ELEM_K entries (elem*21+cid, -1 = zero)
  // fused K1+K4 (k_assemble_fused): K slots in blocks of ASM_SLOTS; per block the elements its slots draw from
  // (blk_elems[blk_elem_ptr[b] ..]) and per triplet the block-local source (local element * 21 + cid, 0xFFFF = zero)
  #ifndef OBGPU_ASM_SLOTS
#define OBGPU_ASM_SLOTS 512
#endif
  static constexpr int ASM_SLOTS = OBGPU_ASM_SLOTS;
  std::vector<int32_t> blk_elem_ptr, blk_elems;
  std::vector<uint16_t> seg_loc;
  int asm_max_loc = 0;                        // largest element list of a block; 0 = fused assembly not available
  // K_ff, full symmetric over global free indices, in SELL-32 order (see build_sparse_plan): sl_ptr[nf_tot/32 + 1]
  // slice offsets, ff_col / ff_src (K slot, -1 = padding) per stored entry; ff_rowptr is scratch of the builder
  std::vector<int32_t> ff_rowptr, ff_col, ff_src, sl_ptr;
  std::vector<int32_t> ff_diag;               // [nf_tot] position of the diagonal entry inside ff_col/ff_src
  int64_t ff_nnz = 0;                         // structural entries of K_ff (without padding)
  std::vector<int32_t> bf_rowptr, bf_col, bf_src;  // K_bf: rows = global bounded index, cols = global free index
  BlockSell ff_blk;                           // K_ff in 3x3 block-SELL when every structure's free DOFs come in triples
  std::vector<int32_t> bb_rowptr, bb_col, bb_src;  // K_bb as the reference stores it (upper triangle): row b(lo), col b(hi)
  // load vector gather: target t < nf_tot -> F_f[t]; t >= nf_tot -> Eq[t - nf_tot] (equivalent loads on bounded DOFs)
  // src >= 0: ELOAD_G entry (multimap position*6 + k); src < 0: nodal load -(l+1)
  std::vector<int32_t> fseg_ptr, fseg_src;
};

// Aggregation-multigrid hierarchy of the sparse route's preconditioner (symbolic_mg.cpp; OBGPU_PCG_MG).
// Level 0 is K_ff itself (SparsePlan's SELL-32 arrays).  Level l >= 1 has three rows per aggregate — the rigid-body
// modes (tx, ty, rot about the aggregate centre) of the level-(l-1) "nodes" it collects — and is stored in SELL-32
// too, every structure's rows padded to a multiple of 96 so that row/3 is a global aggregate index.
struct MgLevel {
  int64_t n = 0;                        // rows of this level over the batch (padded)
  std::vector<int32_t> row_off;         // [S+1] first row of each structure (multiples of 96)
  std::vector<int32_t> nrows;           // [S] rows in use (3 * aggregates); 0 = the structure stops above this level
  std::vector<int32_t> sl_ptr, col;     // SELL-32 pattern (padding: col = own row)
  std::vector<int32_t> diag;            // [n] position of the diagonal entry, -1 on padding rows
  // transfer between level l-1 (fine) and this level: per fine row its aggregate's first row here (-1: none),
  // its DOF kind (0 tx, 1 ty, 2 rot) and the coefficient of the aggregate's rotation mode (-dy, +dx, 1)
  std::vector<int32_t> f_agg;
  std::vector<uint8_t> f_kind;
  std::vector<double> f_w;
  std::vector<int32_t> mem_ptr, mem;    // [n/3 + 1] fine rows of each aggregate, ascending
  BlockSell blk;                        // the same matrix in 3x3 block-SELL (SpMV-class kernels)
};
struct MgPlan {
  bool valid = false;
  std::vector<MgLevel> lev;             // lev[0] unused (fine level), lev[1..]
  std::vector<int32_t> nlev;            // [S] number of levels of each structure, fine one included (>= 1)
  std::vector<uint8_t> dense;           // [S] coarsest level small enough for the explicit dense inverse
  static constexpr int DENSE_MAX = 768; // rows
};

int run_symbolic(const HostBatch& hb, Symbolic& sym, std::string& err, bool want_pattern);
int build_mg_plan(const HostBatch& hb, const Symbolic& sym, const SparsePlan& plan, MgPlan& mg, std::string& err);
struct SparsePlan;
int build_sparse_plan(const HostBatch& hb, const Symbolic& sym, SparsePlan& plan, std::string& err);
int build_tiny_program(const HostBatch& hb, const Symbolic& sym, TinyProgram& prog, std::string& err);

}  // namespace obgpu
