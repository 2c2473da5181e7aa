// ob_batch.ipp — CPU-baseline driver for batches of DIFFERENT structures (config C3), written against the oracle's own C
// ABI (obo_create / obo_solve_static / obo_stress), so the same code drives the restated oracle (ob_oracle.cpp) and the
// reference's own sources (ref_harness.cpp).  One full problem set-up + solveStatic + postProcCalcStress per structure,
// structures dealt to `n_threads` host threads.  TEST INFRASTRUCTURE (bench.py's cpu_baseline / --impl reference legs).
#include <atomic>
#include <chrono>
#include <cmath>
#include <thread>
#include <vector>

extern "C" double obo_run_batch(int64_t S, const int64_t* node_off, const int64_t* elem_off, const int64_t* con_off,
                                const int64_t* nload_off, const int64_t* eload_off, const double* node_xy,
                                const uint32_t* elem_nodes, const uint8_t* elem_type, const double* elem_EAI,
                                const uint32_t* con_node, const uint8_t* con_dof, const double* con_val,
                                const uint32_t* nl_node, const uint8_t* nl_dof, const double* nl_val, const uint32_t* el_elem,
                                const double* el_q, const double* el_dir, int n_threads, double* summ /* [S][2] or null */) {
  if (n_threads < 1) n_threads = 1;
  std::atomic<int64_t> next{0};
  std::atomic<int> failed{0};
  const auto t0 = std::chrono::steady_clock::now();
  auto work = [&] {
    std::vector<double> U, F;
    for (;;) {
      const int64_t s0 = next.fetch_add(64);
      if (s0 >= S) break;
      for (int64_t s = s0; s < S && s < s0 + 64; s++) {
        const int64_t n0 = node_off[s], e0 = elem_off[s], c0 = con_off[s], l0 = nload_off[s], q0 = eload_off[s];
        obo_problem* p = obo_create(node_off[s + 1] - n0, node_xy + 2 * n0, nullptr, elem_off[s + 1] - e0, elem_nodes + 2 * e0,
                                    elem_type + e0, elem_EAI + 3 * e0, con_off[s + 1] - c0, con_node + c0, con_dof + c0,
                                    con_val + c0, nload_off[s + 1] - l0, nl_node + l0, nl_dof + l0, nl_val + l0,
                                    eload_off[s + 1] - q0, el_elem + q0, el_q + q0, el_dir + 3 * q0);
        if (!p || obo_solve_static(p, 0) != 0 || obo_stress(p) != 0) {
          failed = 1;
          if (p) obo_destroy(p);
          continue;
        }
        if (summ) {
          const int64_t nf = obo_get_f64(p, OBO_F_UF, nullptr, 0), nb = obo_get_f64(p, OBO_F_FB, nullptr, 0);
          U.resize(size_t(nf)); F.resize(size_t(nb));
          obo_get_f64(p, OBO_F_UF, U.data(), nf);
          obo_get_f64(p, OBO_F_FB, F.data(), nb);
          double um = 0, fs = 0;
          for (double v : U) um = std::fmax(um, std::fabs(v));
          for (double v : F) fs += v;
          summ[2 * s] = um; summ[2 * s + 1] = fs;
        }
        obo_destroy(p);
      }
    }
  };
  std::vector<std::thread> th;
  for (int t = 1; t < n_threads; t++) th.emplace_back(work);
  work();
  for (auto& x : th) x.join();
  const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  return failed ? -1.0 : dt;
}
