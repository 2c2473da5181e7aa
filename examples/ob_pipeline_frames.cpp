// ob_pipeline_frames.cpp — a parametric sweep of small frames through obhost::BatchPipeline (include/obpipeline.hpp):
// host arrays in, host arrays out, sub-batches dealt to several contexts so that copies and kernels overlap.
// The same sub-batches are first solved one after the other on a single context; the pipeline must return the same bits.
//   usage: ob_pipeline_frames [device] [sub_batches] [frames_per_sub_batch] [workers]
// build: g++ -std=c++17 -pthread -Iinclude examples/ob_pipeline_frames.cpp -Lopenbeam_b200 -lobgpu -Wl,-rpath,$PWD/openbeam_b200
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <memory>

#include "obpipeline.hpp"

namespace {

struct Lcg {  // small deterministic generator for the sweep parameters
  uint64_t s;
  double next() {
    s = s * 6364136223846793005ull + 1442695040888963407ull;
    return double(s >> 11) * (1.0 / 9007199254740992.0);
  }
};

// one sub-batch: `count` grid frames of b bays x s storeys, bases fixed, UDL on every beam, wind on the left column nodes
struct SubBatch {
  obhost::PinnedArray<int64_t> node_off, elem_off, con_off, nload_off, eload_off;
  obhost::PinnedArray<double> node_xy, elem_EAI, con_val, nload_val, eload_q, eload_dir;
  obhost::PinnedArray<uint32_t> elem_nodes, con_node, nload_node, eload_elem;
  obhost::PinnedArray<uint8_t> elem_type, con_dof, nload_dof;
  obgpu_batch_desc desc{};

  SubBatch(int64_t first, int64_t count) {
    std::vector<int> B(count), S(count);
    int64_t N = 0, E = 0, C = 0, NL = 0, EL = 0;
    node_off.resize(count + 1); elem_off.resize(count + 1); con_off.resize(count + 1); nload_off.resize(count + 1);
    eload_off.resize(count + 1);
    for (int64_t k = 0; k < count; k++) {
      const int64_t g = first + k;
      B[k] = 1 + int(g % 3);
      S[k] = 1 + int((g / 3) % 3);
      node_off[k] = N; elem_off[k] = E; con_off[k] = C; nload_off[k] = NL; eload_off[k] = EL;
      N += (B[k] + 1) * (S[k] + 1);
      E += S[k] * (2 * B[k] + 1);
      C += 3 * (B[k] + 1);
      NL += S[k];
      EL += S[k] * B[k];
    }
    node_off[count] = N; elem_off[count] = E; con_off[count] = C; nload_off[count] = NL; eload_off[count] = EL;
    node_xy.resize(2 * N); elem_nodes.resize(2 * E); elem_type.resize(E); elem_EAI.resize(3 * E);
    con_node.resize(C); con_dof.resize(C); con_val.resize(C);
    nload_node.resize(NL); nload_dof.resize(NL); nload_val.resize(NL);
    eload_elem.resize(EL); eload_q.resize(EL); eload_dir.resize(3 * EL);
    for (int64_t k = 0; k < count; k++) {
      Lcg r{uint64_t(first + k) * 0x9E3779B97F4A7C15ull + 12345};
      const int b = B[k], s = S[k];
      const double bay = 3.0 + 3.0 * r.next(), storey = 2.5 + 1.5 * r.next();
      const double A = 28.5e-4 * (1.0 + r.next()), Iz = 1940e-8 * (1.0 + 2.0 * r.next());
      double* xy = node_xy.data() + 2 * node_off[k];
      for (int j = 0; j <= s; j++)
        for (int i = 0; i <= b; i++) {
          xy[2 * (j * (b + 1) + i)] = i * bay;
          xy[2 * (j * (b + 1) + i) + 1] = j * storey;
        }
      int64_t e = elem_off[k], el = eload_off[k];
      auto add = [&](int n0, int n1) {
        elem_nodes[2 * e] = uint32_t(n0); elem_nodes[2 * e + 1] = uint32_t(n1);
        elem_type[e] = OBGPU_BEAM2D_RR;
        elem_EAI[3 * e] = 2.1e11; elem_EAI[3 * e + 1] = A; elem_EAI[3 * e + 2] = Iz;
        return e++;
      };
      for (int j = 1; j <= s; j++) {
        for (int i = 0; i <= b; i++) add((j - 1) * (b + 1) + i, j * (b + 1) + i);  // columns of storey j
        for (int i = 0; i < b; i++) {                                               // beams of storey j, UDL downwards
          const int64_t eb = add(j * (b + 1) + i, j * (b + 1) + i + 1);
          eload_elem[el] = uint32_t(eb - elem_off[k]);
          eload_q[el] = 5e3 + 25e3 * r.next();
          eload_dir[3 * el] = 0.0; eload_dir[3 * el + 1] = -1.0; eload_dir[3 * el + 2] = 0.0;
          el++;
        }
      }
      for (int i = 0; i <= b; i++)
        for (int d = 0; d < 3; d++) {
          const int64_t c = con_off[k] + 3 * i + d;
          con_node[c] = uint32_t(i);
          con_dof[c] = uint8_t(d == 2 ? OBGPU_RZ : d);
          con_val[c] = 0.0;
        }
      for (int j = 1; j <= s; j++) {
        const int64_t l = nload_off[k] + j - 1;
        nload_node[l] = uint32_t(j * (b + 1));
        nload_dof[l] = OBGPU_DX;
        nload_val[l] = 20e3 * r.next();
      }
    }
    desc.n_structures = count;
    desc.node_off = node_off.data(); desc.elem_off = elem_off.data(); desc.con_off = con_off.data();
    desc.nload_off = nload_off.data(); desc.eload_off = eload_off.data();
    desc.node_xy = node_xy.data(); desc.node_rotz = nullptr;
    desc.elem_nodes = elem_nodes.data(); desc.elem_type = elem_type.data(); desc.elem_EAI = elem_EAI.data();
    desc.con_node = con_node.data(); desc.con_dof = con_dof.data(); desc.con_val = con_val.data();
    desc.nload_node = nload_node.data(); desc.nload_dof = nload_dof.data(); desc.nload_val = nload_val.data();
    desc.eload_elem = eload_elem.data(); desc.eload_q = eload_q.data(); desc.eload_dir = eload_dir.data();
    desc.n_cases = 1;
  }
};

template <class T>
bool same(const obhost::PinnedArray<T>& a, const obhost::PinnedArray<T>& b) {
  return a.size() == b.size() && std::memcmp(a.data(), b.data(), a.size() * sizeof(T)) == 0;
}

}  // namespace

int main(int argc, char** argv) {
  const int device = argc > 1 ? std::atoi(argv[1]) : 0;
  const int nsub = argc > 2 ? std::atoi(argv[2]) : 6;
  const int64_t per = argc > 3 ? std::atoll(argv[3]) : 3000;
  const int workers = argc > 4 ? std::atoi(argv[4]) : 3;
  try {
    std::vector<std::unique_ptr<SubBatch>> subs;
    std::vector<obgpu_batch_desc> descs;
    for (int i = 0; i < nsub; i++) {
      subs.emplace_back(new SubBatch(int64_t(i) * per, per));
      descs.push_back(subs.back()->desc);
    }
    std::vector<obhost::SubBatchResult> ref, got;
    {
      obhost::BatchPipeline one(device, 1);  // a single context: the sub-batches strictly one after the other
      one.run(descs, ref);
    }
    obhost::BatchPipeline pipe(device, workers);
    pipe.run(descs, got);  // first pass: contexts allocate their device buffers
    const auto t0 = std::chrono::steady_clock::now();
    const int reps = 5;
    for (int r = 0; r < reps; r++) pipe.run(descs, got);
    const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() / reps;
    bool ok = true;
    double umax = 0.0, sum_fy = 0.0, sum_q = 0.0;
    for (int i = 0; i < nsub; i++) {
      ok = ok && same(ref[i].U_f, got[i].U_f) && same(ref[i].F_b, got[i].F_b) && same(ref[i].stress, got[i].stress) &&
           same(ref[i].info, got[i].info);
      for (size_t k = 0; k < got[i].U_f.size(); k++) umax = std::max(umax, std::abs(got[i].U_f[k]));
      for (size_t k = 0; k < got[i].info.size(); k++) ok = ok && got[i].info[k] == 0;
      // global vertical equilibrium of the whole sub-batch: reactions F_b,y = total UDL (bases fixed: bounded DOFs come
      // in (x, y, rz) triples per base node)
      for (size_t k = 1; k < got[i].F_b.size(); k += 3) sum_fy += got[i].F_b[k];
      const SubBatch& sb = *subs[i];
      for (int64_t s = 0; s < per; s++)
        for (int64_t l = sb.eload_off[s]; l < sb.eload_off[s + 1]; l++) {
          const int64_t e = sb.elem_off[s] + sb.eload_elem[l];
          const double* xy = sb.node_xy.data() + 2 * sb.node_off[s];
          const double dx = xy[2 * sb.elem_nodes[2 * e + 1]] - xy[2 * sb.elem_nodes[2 * e]];
          sum_q += sb.eload_q[l] * dx;
        }
    }
    std::printf("structures %lld sub_batches %d workers %d\n", (long long)(per * nsub), nsub, workers);
    std::printf("pipeline_matches_single_context %d\n", ok ? 1 : 0);
    std::printf("max_abs_U_f %.6e\n", umax);
    std::printf("vertical_equilibrium_relerr %.3e\n", std::abs(sum_fy - sum_q) / sum_q);
    std::printf("structures_per_s %.4g\n", double(per * nsub) / dt);
    return ok ? 0 : 1;
  } catch (const std::exception& e) {
    std::printf("Exception: %s\n", e.what());
    return 255;
  }
}
