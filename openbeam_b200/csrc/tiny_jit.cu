// tiny_jit.cpp — run-time specialisation of the thread-per-case kernel to one topology (NVRTC).
//
// The hot loop of k_tiny_fused interprets an integer "topology program"; half of its instructions are that
// interpretation (index loads, skip tests, block dispatch, shared-memory read-modify-write of K_ff / F columns).
// Here the program is printed as `__device__ constexpr` arrays in front of the hand-written kernel source
// tiny_jit_kernel.cuh and compiled for sm_100a with NVRTC, so the compiler sees every index as a constant and keeps
// K_ff, F and U in registers.  The floating-point work per case is unchanged (nothing case-invariant is baked:
// coordinates, sections and base loads stay run-time parameters).  One compile per distinct topology per process
// (≈ 1 s), cached.  libnvrtc and libcuda are dlopen'ed so that libobgpu.so still loads on a machine without a driver;
// if either is missing, or the compile fails, the caller falls back to the precompiled k_tiny_fused and says so once.
#include <dlfcn.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "kernels.h"
#include "tiny_program.h"

namespace obgpu {

extern const char* const kTinyJitSource;  // generated: tiny_jit_src.cpp (beam_math + ldlt + kernel, NVRTC-clean)

namespace {

// ---- minimal NVRTC / driver API surface, resolved at run time ----
typedef struct _nvrtcProgram* nvrtcProgram;
typedef int nvrtcResult;
typedef int CUresult;
typedef struct CUmod_st* CUmodule;
typedef struct CUfunc_st* CUfunction;
typedef struct CUstream_st* CUstream;

struct Api {
  bool ok = false;
  std::string why;
  nvrtcResult (*nvrtcCreateProgram)(nvrtcProgram*, const char*, const char*, int, const char* const*, const char* const*);
  nvrtcResult (*nvrtcCompileProgram)(nvrtcProgram, int, const char* const*);
  nvrtcResult (*nvrtcGetCUBINSize)(nvrtcProgram, size_t*);
  nvrtcResult (*nvrtcGetCUBIN)(nvrtcProgram, char*);
  nvrtcResult (*nvrtcGetProgramLogSize)(nvrtcProgram, size_t*);
  nvrtcResult (*nvrtcGetProgramLog)(nvrtcProgram, char*);
  nvrtcResult (*nvrtcDestroyProgram)(nvrtcProgram*);
  CUresult (*cuModuleLoadData)(CUmodule*, const void*);
  CUresult (*cuModuleGetFunction)(CUfunction*, CUmodule, const char*);
  CUresult (*cuFuncSetAttribute)(CUfunction, int, int);
  CUresult (*cuOccupancyMaxActiveBlocksPerMultiprocessor)(int*, CUfunction, int, size_t);
  CUresult (*cuLaunchKernel)(CUfunction, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, CUstream,
                             void**, void**);
};

template <class F>
bool sym(void* h, const char* name, F& f) {
  f = reinterpret_cast<F>(dlsym(h, name));
  return f != nullptr;
}

Api& api() {
  static Api a;
  static std::once_flag once;
  std::call_once(once, [] {
    void* rtc = nullptr;
    for (const char* n : {"libnvrtc.so.12", "libnvrtc.so", "/usr/local/cuda/lib64/libnvrtc.so.12"})
      if ((rtc = dlopen(n, RTLD_NOW | RTLD_LOCAL))) break;
    void* drv = dlopen("libcuda.so.1", RTLD_NOW | RTLD_LOCAL);
    if (!rtc || !drv) {
      a.why = !rtc ? "libnvrtc not found" : "libcuda.so.1 not found";
      return;
    }
    bool ok = sym(rtc, "nvrtcCreateProgram", a.nvrtcCreateProgram) && sym(rtc, "nvrtcCompileProgram", a.nvrtcCompileProgram) &&
              sym(rtc, "nvrtcGetCUBINSize", a.nvrtcGetCUBINSize) && sym(rtc, "nvrtcGetCUBIN", a.nvrtcGetCUBIN) &&
              sym(rtc, "nvrtcGetProgramLogSize", a.nvrtcGetProgramLogSize) && sym(rtc, "nvrtcGetProgramLog", a.nvrtcGetProgramLog) &&
              sym(rtc, "nvrtcDestroyProgram", a.nvrtcDestroyProgram) && sym(drv, "cuModuleLoadData", a.cuModuleLoadData) &&
              sym(drv, "cuModuleGetFunction", a.cuModuleGetFunction) && sym(drv, "cuFuncSetAttribute", a.cuFuncSetAttribute) &&
              sym(drv, "cuOccupancyMaxActiveBlocksPerMultiprocessor", a.cuOccupancyMaxActiveBlocksPerMultiprocessor) &&
              sym(drv, "cuLaunchKernel", a.cuLaunchKernel);
    a.ok = ok;
    if (!ok) a.why = "missing NVRTC / driver entry point";
  });
  return a;
}

struct Compiled {
  CUfunction fn = nullptr;
  int per_sm = 1;
  bool bulk = false;
  size_t smem_nl = 0, smem_q = 0, smem_base = 0;  // dynamic shared memory = base + (case_nl ? nl : 0) + (case_q ? q : 0)
  bool failed = false;
};

void arr(std::string& s, const char* name, const int32_t* v, int cnt) {
  s += "__device__ constexpr signed char ";
  s += name;
  s += "[] = {";
  for (int i = 0; i < (cnt > 0 ? cnt : 1); i++) {
    s += std::to_string(i < cnt ? v[i] : 0);
    s += ',';
  }
  s += "};\n";
}

constexpr int JIT_RS = 128 + 16;  // = TJ_RS of the kernel (record staging pitch of the non-bulk variant)
int jit_block() {            // threads per CTA (= TJ_BLOCK = TJ_CS of the kernel): 128, or OBGPU_TINY_JIT_BLOCK in {32, 64, 96, 128}
  static const int b = [] {
    const char* e = getenv("OBGPU_TINY_JIT_BLOCK");
    const int v = e ? atoi(e) : 128;
    return (v == 32 || v == 64 || v == 96 || v == 128) ? v : 128;
  }();
  return b;
}

}  // namespace

// Returns cudaSuccess after a launch; cudaErrorNotSupported when the specialised kernel is unavailable (caller falls
// back to the precompiled kernel).
cudaError_t launch_tiny_jit(int device, const TinyArgs& a, int sm_count, cudaStream_t st, const char** which) {
  static const bool enabled = !(getenv("OBGPU_TINY_JIT") && atoi(getenv("OBGPU_TINY_JIT")) == 0);
  if (!enabled) return cudaErrorNotSupported;
  Api& A = api();
  static bool told = false;
  if (!A.ok) {
    if (!told) fprintf(stderr, "[obgpu] topology-specialised kernel unavailable (%s): using the precompiled kernel\n", A.why.c_str());
    told = true;
    return cudaErrorNotSupported;
  }
  const int32_t* I = a.prog_i;
  const double* D = a.prog_d;
  const int N = I[0], E = I[1], n = I[2], nf = I[3], nb = I[4], NL = I[5], EL = I[6];
  const TinyLayout lay(N, E, n, nf, nb, NL, EL);
  static std::mutex mu;
  static std::map<std::string, Compiled> cache;
  std::string key(reinterpret_cast<const char*>(I), size_t(a.n_ints) * sizeof(int32_t));
  key += char('0' + device);
  key += a.F_f ? 'F' : '-';
  key += a.summ ? 'S' : '-';
  key += a.stress_compact ? 'C' : '-';
  if (a.summ) key.append(reinterpret_cast<const char*>(a.bnd_dofid), size_t(nb));  // the kernel is also specialised on whether F_f is kept (staging rows, occupancy)
  std::lock_guard<std::mutex> lock(mu);
  Compiled& c = cache[key];
  if (c.failed) return cudaErrorNotSupported;
  if (!c.fn) {
    const int nfp = nf * (nf + 1) / 2;
    int minb = nfp + n <= 40 ? 5 : nfp + n <= 64 ? 3 : 2;  // register budget: K_ff + F/U live in registers
    minb = minb * 128 / jit_block();  // the same register budget per thread whatever the CTA width
    if (getenv("OBGPU_TINY_JIT_MINB")) minb = atoi(getenv("OBGPU_TINY_JIT_MINB"));  // tuning experiments
    // result records through the bulk-copy engine when a CTA's staging rows (final layout, 4 warps x 32 cases) stay small
    const int sr = a.stress_compact ? 6 : 12;
    const size_t bulk_bytes = size_t(jit_block() / 32) * (size_t(32) * (sr * E + (a.F_f ? 2 : 1) * nf + nb + (a.summ ? 4 : 0)) + 16) * 8;
    bool& bulk = c.bulk;
    bulk = bulk_bytes <= 60 * 1024;
    if (const char* ev = getenv("OBGPU_TINY_JIT_BULK")) bulk = bulk && atoi(ev) != 0;
    if (a.stress_compact && !bulk) {  // only the bulk-store variant writes compact records: the caller compacts afterwards
      cache.erase(key);
      return cudaErrorNotYetImplemented;
    }
    std::string src;
    src += "#define TJ_BLOCK_SIZE " + std::to_string(jit_block()) + "\n";
    src += std::string("#define TJ_BULK ") + (bulk ? "1\n" : "0\n");
    src += std::string("#define TJ_FF ") + (a.F_f ? "1\n" : "0\n");
    src += std::string("#define TJ_SC ") + (a.stress_compact ? "1\n" : "0\n");
    const char* ef = getenv("OBGPU_TINY_JIT_EVICT_FIRST");
    src += std::string("#define TJ_EVICT_FIRST ") + (ef && atoi(ef) == 0 ? "0\n" : "1\n");
    src += "#define TJ_N " + std::to_string(N) + "\n#define TJ_E " + std::to_string(E) + "\n#define TJ_n " + std::to_string(n) +
           "\n#define TJ_NF " + std::to_string(nf) + "\n#define TJ_NB " + std::to_string(nb) + "\n#define TJ_NL " +
           std::to_string(NL) + "\n#define TJ_EL " + std::to_string(EL) + "\n#define TJ_MINB " + std::to_string(minb) + "\n";
    arr(src, "TJ_enode", I + lay.o_enode, 2 * E);
    arr(src, "TJ_etype", I + lay.o_etype, E);
    arr(src, "TJ_udof", I + lay.o_udof, 6 * E);
    arr(src, "TJ_edst", I + lay.o_edst, ELEM_K * E);
    arr(src, "TJ_elptr", I + lay.o_elptr, E + 1);
    arr(src, "TJ_elslot", I + lay.o_elslot, EL);
    arr(src, "TJ_nldst", I + lay.o_nldst, NL);
    arr(src, "TJ_freek", I + lay.o_freek, nf);
    arr(src, "TJ_bndk", I + lay.o_bndk, nb);
    arr(src, "TJ_elkind", I + lay.o_elkind, EL);
    {
      std::vector<int32_t> bd(size_t(std::max(nb, 1)), 0);
      for (int i = 0; i < nb && a.summ; i++) bd[size_t(i)] = a.bnd_dofid[i];
      arr(src, "TJ_bnddof", bd.data(), nb);
      src += std::string("#define TJ_SUMM ") + (a.summ ? "1\n" : "0\n");
    }
    if (const char* abl = getenv("OBGPU_TINY_JIT_ABL")) {  // ablation experiments (scripts/), never set in production
      if (atoi(abl) & 1) src += "#define TJ_ABL_NOSTORE 1\n";
      if (atoi(abl) & 2) src += "#define TJ_ABL_NOSOLVE 1\n";
    }
    src += kTinyJitSource;
    if (const char* dump = getenv("OBGPU_TINY_JIT_DUMP"))  // the generated translation unit, for offline nvcc/cuobjdump study
      if (FILE* f = fopen(dump, "w")) {
        fwrite(src.data(), 1, src.size(), f);
        fclose(f);
      }
    nvrtcProgram prog = nullptr;
    const char* opts[] = {"--gpu-architecture=sm_100a", "--std=c++17", "-lineinfo", "-default-device"};
    bool ok = A.nvrtcCreateProgram(&prog, src.c_str(), "k_tiny_jit.cu", 0, nullptr, nullptr) == 0;
    std::string log;
    if (ok) {
      ok = A.nvrtcCompileProgram(prog, 4, opts) == 0;
      size_t ls = 0;
      if (A.nvrtcGetProgramLogSize(prog, &ls) == 0 && ls > 1) {
        log.resize(ls);
        A.nvrtcGetProgramLog(prog, &log[0]);
      }
    }
    std::vector<char> cubin;
    if (ok) {
      size_t sz = 0;
      ok = A.nvrtcGetCUBINSize(prog, &sz) == 0 && sz > 0;
      if (ok) {
        cubin.resize(sz);
        ok = A.nvrtcGetCUBIN(prog, cubin.data()) == 0;
      }
    }
    if (prog) A.nvrtcDestroyProgram(&prog);
    CUmodule mod = nullptr;
    if (ok) ok = A.cuModuleLoadData(&mod, cubin.data()) == 0 && A.cuModuleGetFunction(&c.fn, mod, "k_tiny_jit") == 0;
    if (!ok) {
      fprintf(stderr, "[obgpu] topology-specialised kernel did not build: using the precompiled kernel\n%s\n", log.c_str());
      c.failed = true;
      c.fn = nullptr;
      return cudaErrorNotSupported;
    }
    const int xw = std::max(std::max(nf, nb), 1);
    c.smem_base = size_t(4 * E + 2 * std::max(EL, 1)) * 8 + (bulk ? bulk_bytes : size_t(12 + xw) * JIT_RS * 8);
    c.smem_nl = size_t(NL) * jit_block() * 8;
    c.smem_q = size_t(2 * EL) * jit_block() * 8;
    const size_t smem_max = c.smem_base + c.smem_nl + c.smem_q;
    if (A.cuFuncSetAttribute(c.fn, 8 /*CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES*/, int(smem_max)) != 0 ||
        A.cuOccupancyMaxActiveBlocksPerMultiprocessor(&c.per_sm, c.fn, jit_block(), smem_max) != 0 || c.per_sm < 1) {
      c.failed = true;
      return cudaErrorNotSupported;
    }
  }
  // run-time values, in the layout of TinyJitVals
  std::vector<double> V;
  V.insert(V.end(), D + lay.o_xy, D + lay.o_xy + 2 * N);
  V.insert(V.end(), D + lay.o_eai, D + lay.o_eai + 3 * E);
  for (int i = 0; i < std::max(NL, 1); i++) V.push_back(i < NL ? D[lay.o_nlval + i] : 0.0);
  for (int i = 0; i < std::max(EL, 1); i++) V.push_back(i < EL ? D[lay.o_elq + i] : 0.0);
  if (EL > 0)
    for (int i = 0; i < EL; i++) {
      V.push_back(D[lay.o_eldir + 3 * i]);
      V.push_back(D[lay.o_eldir + 3 * i + 1]);
    }
  else
    V.push_back(0.0);
  for (int i = 0; i < std::max(EL, 1); i++) V.push_back(i < EL ? D[lay.o_elq2 + i] : 0.0);
  const size_t smem = c.smem_base + (a.case_nl ? c.smem_nl : 0) + (a.case_q ? c.smem_q : 0);
  const int64_t need = (a.n_cases + jit_block() - 1) / jit_block();
  int per_sm = c.per_sm;
  if (getenv("OBGPU_TINY_JIT_CTAS")) per_sm = std::max(1, std::min(per_sm, atoi(getenv("OBGPU_TINY_JIT_CTAS"))));  // tuning experiments
  int64_t grid = int64_t(sm_count) * per_sm;  // persistent: a whole number of CTAs per SM, grid-stride over cases
  if (need < grid) grid = need;
  if (grid < 1) grid = 1;
  TinyArgs args = a;
  void* params[2] = {V.data(), &args};
  if (A.cuLaunchKernel(c.fn, unsigned(grid), 1, 1, jit_block(), 1, 1, unsigned(smem), reinterpret_cast<CUstream>(st), params,
                       nullptr) != 0)
    return cudaErrorLaunchFailure;
  if (which)
    *which = c.bulk ? "k_tiny_jit (NVRTC, specialised to the topology; thread per case, 128 threads/CTA, records stored by cp.async.bulk)"
                    : "k_tiny_jit (NVRTC, specialised to the topology; thread per case, 128 threads/CTA, staged coalesced stores)";
  return cudaSuccess;
}

}  // namespace obgpu
