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

extern "C" int obgpu_summaries_device(obgpu_ctx* c, double*) {
  return obgpu::fail(c, OBGPU_EUNSUPPORTED, "summaries not built yet");
}
