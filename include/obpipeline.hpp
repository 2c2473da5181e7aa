// obpipeline.hpp — host-side pipeline above the C ABI (include/obgpu.h) for jobs of MANY small structures whose inputs and
// results live in host memory (BASELINE config C3, "parametric sweep").  Header-only C++17; the Python twin is
// openbeam_b200/pipeline.py.
//
// One obgpu_ctx runs upload -> symbolic -> solve -> download strictly in sequence, so a host-to-host job is bound by the
// sum of the two PCIe directions plus the kernels.  BatchPipeline deals the sub-batches of a job to a few contexts (one host
// thread and one CUDA stream each, all on the same GPU) and lets one sub-batch at a time use each of the three resources
// — the H2D copy engine, the SMs, the D2H copy engine — so the workers fall into a staggered rhythm in which all three
// are busy at once and the job approaches max(H2D, kernels, D2H).
//
// The reference solves one structure per process (apps/ob-solve/ob-solve_main.cpp); structures of different sub-batches
// are independent problems, so cutting the list anywhere changes no result.
#pragma once
#include <cstdint>
#include <cstring>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "obgpu.h"

namespace obhost {

// cudaHostAlloc'ed array (obgpu_host_alloc): copies to and from it run at PCIe speed and asynchronously
template <class T>
class PinnedArray {
 public:
  PinnedArray() = default;
  explicit PinnedArray(size_t n) { resize(n); }
  PinnedArray(const PinnedArray&) = delete;
  PinnedArray& operator=(const PinnedArray&) = delete;
  PinnedArray(PinnedArray&& o) noexcept : p_(o.p_), n_(o.n_), cap_(o.cap_) { o.p_ = nullptr; o.n_ = o.cap_ = 0; }
  PinnedArray& operator=(PinnedArray&& o) noexcept {
    if (this != &o) {
      release();
      p_ = o.p_; n_ = o.n_; cap_ = o.cap_;
      o.p_ = nullptr; o.n_ = o.cap_ = 0;
    }
    return *this;
  }
  ~PinnedArray() { release(); }
  void resize(size_t n) {  // contents are not kept when the array grows
    if (n > cap_) {
      release();
      void* q = nullptr;
      if (obgpu_host_alloc(&q, (n ? n : 1) * sizeof(T)) != OBGPU_OK) throw std::runtime_error("obgpu_host_alloc failed");
      p_ = static_cast<T*>(q);
      cap_ = n;
    }
    n_ = n;
  }
  T* data() { return p_; }
  const T* data() const { return p_; }
  size_t size() const { return n_; }
  T& operator[](size_t i) { return p_[i]; }
  const T& operator[](size_t i) const { return p_[i]; }

 private:
  void release() {
    if (p_) obgpu_host_free(p_);
    p_ = nullptr;
    n_ = cap_ = 0;
  }
  T* p_ = nullptr;
  size_t n_ = 0, cap_ = 0;
};

// results of one sub-batch (the arrays are reused when a later run has the same sizes)
struct SubBatchResult {
  PinnedArray<double> U_f, F_b, stress;  // [nf_tot], [nb_tot], [Etot][2][3] (compact) or [Etot][2][6]
  PinnedArray<int32_t> info;             // [S]
  std::vector<int64_t> free_off, bnd_off;  // [S+1], filled when want_offsets
  bool want_offsets = false;
};

class BatchPipeline {
 public:
  explicit BatchPipeline(int device = 0, int workers = 3) {
    for (int w = 0; w < (workers < 1 ? 1 : workers); w++) {
      obgpu_ctx* c = nullptr;
      const int rc = obgpu_create(device, &c);
      if (rc != OBGPU_OK) {
        close();
        throw std::runtime_error(std::string("obgpu_create: ") + obgpu_status_string(rc));
      }
      ctx_.push_back(c);
    }
  }
  BatchPipeline(const BatchPipeline&) = delete;
  BatchPipeline& operator=(const BatchPipeline&) = delete;
  ~BatchPipeline() { close(); }

  // Solve every sub-batch (obgpu_batch_desc over caller-owned arrays, pinned ones make the copies asynchronous) with the
  // batched-LDLt route; results[i] receives sub-batch i.  Throws std::runtime_error with the library's message (for a
  // non-positive pivot: the reference's, solver_static.cpp:98-105) after all workers have stopped.
  void run(const std::vector<obgpu_batch_desc>& batches, std::vector<SubBatchResult>& results, bool compact_stress = true) {
    if (results.size() != batches.size()) results.resize(batches.size());
    std::string error;
    std::mutex error_lock;
    auto work = [&](size_t w) {
      try {
        for (size_t i = w; i < batches.size(); i += ctx_.size()) one(ctx_[w], batches[i], results[i], compact_stress);
      } catch (const std::exception& e) {
        std::lock_guard<std::mutex> g(error_lock);
        if (error.empty()) error = e.what();
      }
    };
    std::vector<std::thread> threads;
    const size_t n = batches.size() < ctx_.size() ? batches.size() : ctx_.size();
    for (size_t w = 0; w < n; w++) threads.emplace_back(work, w);
    for (auto& t : threads) t.join();
    if (!error.empty()) throw std::runtime_error(error);
  }

 private:
  static void check(obgpu_ctx* c, int rc) {
    if (rc == OBGPU_OK) return;
    const char* m = obgpu_last_error(c);
    throw std::runtime_error(m && *m ? m : obgpu_status_string(rc));
  }
  void one(obgpu_ctx* c, const obgpu_batch_desc& d, SubBatchResult& r, bool compact) {
    {
      std::lock_guard<std::mutex> g(h2d_);
      check(c, obgpu_batch_upload_device(c, &d));
      check(c, obgpu_synchronize(c));
    }
    check(c, obgpu_symbolic(c));  // small kernels and host round trips: latency, runs beside the other stages
    {
      std::lock_guard<std::mutex> g(sm_);
      obgpu_solve_opts o;
      o.method = OBGPU_BATCHED_LDLT; o.with_stress = 1; o.rtol = 1e-12; o.max_iter = 0;
      o.flags = compact ? OBGPU_SOLVE_COMPACT_STRESS : 0;
      check(c, obgpu_solve_static(c, &o));
      check(c, obgpu_synchronize(c));
    }
    int64_t z[OBGPU_NSIZES];
    check(c, obgpu_get_sizes(c, z));
    const size_t S = size_t(z[0]), E = size_t(z[3]), nf = size_t(z[5]), nb = size_t(z[6]);
    r.U_f.resize(nf); r.F_b.resize(nb); r.stress.resize(E * (compact ? 6 : 12)); r.info.resize(S);
    {
      std::lock_guard<std::mutex> g(d2h_);
      check(c, obgpu_download(c, OBGPU_RES_UF, r.U_f.data()));
      check(c, obgpu_download(c, OBGPU_RES_FB, r.F_b.data()));
      check(c, obgpu_download(c, compact ? OBGPU_RES_STRESS_COMPACT : OBGPU_RES_STRESS, r.stress.data()));
      check(c, obgpu_download(c, OBGPU_RES_INFO, r.info.data()));  // OBGPU_ESINGULAR if any structure failed
    }
    if (r.want_offsets) {
      r.free_off.resize(S + 1); r.bnd_off.resize(S + 1);
      check(c, obgpu_get_offsets(c, OBGPU_OFF_FREE, r.free_off.data()));
      check(c, obgpu_get_offsets(c, OBGPU_OFF_BOUNDED, r.bnd_off.data()));
    }
  }
  void close() {
    for (obgpu_ctx* c : ctx_) obgpu_destroy(c);
    ctx_.clear();
  }
  std::vector<obgpu_ctx*> ctx_;
  std::mutex h2d_, sm_, d2h_;
};

}  // namespace obhost
