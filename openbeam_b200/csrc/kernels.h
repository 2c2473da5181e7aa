// kernels.h — launcher declarations shared between the .cu files and the C-ABI layer.
#pragma once
#include "obgpu_internal.h"

namespace obgpu {

struct TinyArgs {
  const int32_t* prog_i;
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
};
cudaError_t launch_tiny_fused(int nf, int n, const TinyArgs& a, int sm_count, cudaStream_t st);

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

}  // namespace obgpu
