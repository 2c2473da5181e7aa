// capi_sparse.cu — global K assembly (K1+K4), BC gather (K5) and the PCG route (K7-K9) for large structures.
#include "ctx.h"

namespace obgpu {
int solve_sparse(obgpu_ctx* c, const obgpu_solve_opts&) { return fail(c, OBGPU_EUNSUPPORTED, "PCG route not built yet"); }
int assemble_general(obgpu_ctx* c) { return fail(c, OBGPU_EUNSUPPORTED, "general assembly path not built yet"); }
int download_general(obgpu_ctx*, int, void*, bool& handled) {
  handled = false;
  return OBGPU_OK;
}
}  // namespace obgpu
