// capi_general.cu — general (any-size, any-topology) path: K1/K2 element kernels, K4 segmented-reduction
// assembly, K5 BC gather, CTA-per-structure LDL^T, PCG.  (filled in incrementally)
#include "ctx.h"

namespace obgpu {
int solve_general(obgpu_ctx* c, const obgpu_solve_opts&) {
  return fail(c, OBGPU_EUNSUPPORTED, "general solve path not built yet");
}
int assemble_general(obgpu_ctx* c) { return fail(c, OBGPU_EUNSUPPORTED, "general assembly path not built yet"); }
int download_general(obgpu_ctx*, int, void*, bool& handled) {
  handled = false;
  return OBGPU_OK;
}
}  // namespace obgpu

extern "C" int obgpu_summaries_device(obgpu_ctx* c, double* d_out) {
  if (!c || !d_out) return OBGPU_EINVAL;
  OB_NEED_DEVICE(c);
  if (!c->solved) return obgpu::fail(c, OBGPU_ESEQUENCE, "obgpu_summaries_device before obgpu_solve_static");
  if (reinterpret_cast<uintptr_t>(d_out) % 32) return obgpu::fail(c, OBGPU_EINVAL, "summary buffer must be 32-byte aligned");
  obgpu::SummaryArgs a;
  a.L = c->L; a.S = c->hb.S; a.nf_tot = c->sym.nf_tot(); a.nb_tot = c->sym.nb_tot();
  a.free_off = c->d_free_off.p; a.bnd_off = c->d_bnd_off.p; a.bnd_dofid = c->d_bnd_dofid.p;
  a.U_f = c->d_Uf.p; a.F_b = c->d_Fb.p; a.info = c->d_info.p; a.out = d_out;
  OB_CUDA(c, obgpu::launch_summaries(a, c->sm_count, c->stream));
  c->launches += 1;
  return OBGPU_OK;
}
