// cpu_engine.cpp — host-side oracle and mirror engine (no CUDA needed).
//
//  * b200knn_cpu_bruteforce: the test oracle. Independent of every other piece of this library
//    except the distance expression: all n squared distances per query, std::nth_element.
//    Role of verifyKNN in the reference's test driver
//    (submodules/cudaKDTree/testing/floatN-knn-and-fcp.cu:292-335), but exact instead of 1e-6.
//  * b200knn_cpu_knn: the GPU algorithm restated for the host — Morton sort, 32-point leaves,
//    32-ary box tree, group traversal with shared radius, strided 4-ary heaps. It validates the
//    culling logic on machines without a GPU and is the CPU backend of the executables.
//    Role of buildTree_host (cukd/builder_thrust.h:514-601), the reference's only CPU path.
//
// Compiled with -ffp-contract=off so that fmaf() calls are the only fused operations.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <numeric>
#include <string>
#include <thread>
#include <vector>
#include "../include/b200knn.h"
#include "knn_math.h"

namespace b200knn {
void set_last_error(const std::string& msg);

namespace {

template <typename F>
void parallel_for(int64_t n, int num_threads, F f) {
  if (num_threads <= 0) num_threads = (int)std::thread::hardware_concurrency();
  if (num_threads < 1) num_threads = 1;
  if ((int64_t)num_threads > n) num_threads = (int)(n > 0 ? n : 1);
  if (num_threads == 1) { f(0, n); return; }
  std::vector<std::thread> ts;
  const int64_t per = (n + num_threads - 1) / num_threads;
  for (int t = 0; t < num_threads; ++t) {
    const int64_t b = t * per, e = std::min<int64_t>(n, b + per);
    if (b >= e) break;
    ts.emplace_back([=] { f(b, e); });
  }
  for (auto& t : ts) t.join();
}

std::atomic<unsigned long long> g_cpu_stats[4];   // leaves, candidates, inserts, inner nodes

// Optional trace for offline modelling of the insertion phase (tools/sift_model.py). With
// B200KNN_CPU_TRACE=<file> and a single thread, every 64th group appends one record per visited leaf:
//   int64 group, int32 leaves so far, int32 lanes in the group, float top[32] (k-th distance per lane
//   before the leaf), float d2[32][32] (lane-major squared distances to the leaf's points, +inf padding).
FILE* trace_file() {
  static FILE* f = [] {
    const char* path = getenv("B200KNN_CPU_TRACE");
    return path && *path ? fopen(path, "wb") : (FILE*)nullptr;
  }();
  return f;
}

struct CpuTree {
  std::vector<Float4> pts;                 // sorted, padded to groups*32 with +inf points
  std::vector<Box> boxes[kMaxLevels];
  int levels = 0;
  int64_t n = 0;
  float lo[3], hi[3];
};

Box make_box(const float* lo, const float* hi, int cnt) {
  Box b;
  b.lox = lo[0]; b.loy = lo[1]; b.loz = lo[2];
  b.hix = hi[0]; b.hiy = hi[1]; b.hiz = hi[2];
  memcpy(&b.cnt_bits, &cnt, 4);
  b.pad = 0.f;
  return b;
}

void bounds_of(const float* xyz, int64_t n, float* lo, float* hi) {
  for (int a = 0; a < 3; ++a) { lo[a] = INFINITY; hi[a] = -INFINITY; }
  for (int64_t i = 0; i < n; ++i)
    for (int a = 0; a < 3; ++a) {
      lo[a] = fminf(lo[a], xyz[3 * i + a]);
      hi[a] = fmaxf(hi[a], xyz[3 * i + a]);
    }
}

std::vector<uint32_t> morton_order(const float* xyz, int64_t n, const float* lo, const float* hi) {
  float iv[3];
  for (int a = 0; a < 3; ++a) { const float e = hi[a] - lo[a]; iv[a] = e > 0.f ? 1.f / e : 0.f; }
  std::vector<uint64_t> keys(n);
  for (int64_t i = 0; i < n; ++i)
    keys[i] = morton63(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], lo[0], lo[1], lo[2], iv[0], iv[1], iv[2]);
  std::vector<uint32_t> order(n);
  std::iota(order.begin(), order.end(), 0u);
  std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return keys[a] < keys[b]; });
  return order;
}

void build_tree(const float* xyz, int64_t n, CpuTree& t) {
  t.n = n;
  bounds_of(xyz, n, t.lo, t.hi);
  if (n == 0) { t.levels = 0; return; }
  const std::vector<uint32_t> order = morton_order(xyz, n, t.lo, t.hi);
  const int64_t groups = div_up(n, kGroup);
  t.pts.assign(groups * kGroup, Float4{INFINITY, INFINITY, INFINITY, 0.f});
  for (int64_t i = 0; i < n; ++i) {
    const uint32_t s = order[i];
    Float4 p{xyz[3 * (size_t)s], xyz[3 * (size_t)s + 1], xyz[3 * (size_t)s + 2], 0.f};
    memcpy(&p.w, &s, 4);
    t.pts[i] = p;
  }
  // k-d refinement of every 1024-point Morton run (mirror of k_refine_runs in build.cu): exact
  // median splits along the widest axis, 5 levels, padding stays at the tail.
  if (n > kGroup) {
    for (int64_t base = 0; base < n; base += 1024) {
      const int nrun = (int)std::min<int64_t>(1024, n - base);
      for (int seg_size = 1024; seg_size > kGroup; seg_size >>= 1) {
        for (int s0 = 0; s0 < nrun; s0 += seg_size) {
          const int s1 = std::min(nrun, s0 + seg_size);
          float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
          for (int i = s0; i < s1; ++i) {
            const Float4& p = t.pts[base + i];
            lo[0] = fminf(lo[0], p.x); lo[1] = fminf(lo[1], p.y); lo[2] = fminf(lo[2], p.z);
            hi[0] = fmaxf(hi[0], p.x); hi[1] = fmaxf(hi[1], p.y); hi[2] = fmaxf(hi[2], p.z);
          }
          const float ex = hi[0] - lo[0], ey = hi[1] - lo[1], ez = hi[2] - lo[2];
          const int axis = (ex >= ey && ex >= ez) ? 0 : (ey >= ez ? 1 : 2);
          std::stable_sort(t.pts.begin() + base + s0, t.pts.begin() + base + s1,
                           [axis](const Float4& a, const Float4& b) {
                             const float ca = axis == 0 ? a.x : (axis == 1 ? a.y : a.z);
                             const float cb = axis == 0 ? b.x : (axis == 1 ? b.y : b.z);
                             return ca < cb;
                           });
        }
      }
    }
  }
  t.boxes[0].resize(groups);
  for (int64_t g = 0; g < groups; ++g) {
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    const int cnt = (int)std::min<int64_t>(kGroup, n - g * kGroup);
    for (int j = 0; j < cnt; ++j) {
      const Float4& p = t.pts[g * kGroup + j];
      lo[0] = fminf(lo[0], p.x); lo[1] = fminf(lo[1], p.y); lo[2] = fminf(lo[2], p.z);
      hi[0] = fmaxf(hi[0], p.x); hi[1] = fmaxf(hi[1], p.y); hi[2] = fmaxf(hi[2], p.z);
    }
    t.boxes[0][g] = make_box(lo, hi, cnt);
  }
  int levels = 1;
  while ((int64_t)t.boxes[levels - 1].size() > kFanout) {
    const auto& ch = t.boxes[levels - 1];
    const int64_t pc = div_up((int64_t)ch.size(), kFanout);
    t.boxes[levels].resize(pc);
    for (int64_t p = 0; p < pc; ++p) {
      float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
      const int cnt = (int)std::min<int64_t>(kFanout, (int64_t)ch.size() - p * kFanout);
      for (int j = 0; j < cnt; ++j) {
        const Box& c = ch[p * kFanout + j];
        lo[0] = fminf(lo[0], c.lox); lo[1] = fminf(lo[1], c.loy); lo[2] = fminf(lo[2], c.loz);
        hi[0] = fmaxf(hi[0], c.hix); hi[1] = fmaxf(hi[1], c.hiy); hi[2] = fmaxf(hi[2], c.hiz);
      }
      t.boxes[levels][p] = make_box(lo, hi, cnt);
    }
    ++levels;
  }
  t.levels = levels;
}

struct Group {
  float q[kGroup][3];
  int cnt = 0;
  float glo[3], ghi[3];
  float maxR2;
  std::vector<float> heap;   // [k][32]
  int k;
  unsigned long long leaves = 0, inserts = 0, inner = 0;
  unsigned active = 0xffffffffu;   // lanes taking part in the current pass (see traverse_group in traverse.cuh)
  int64_t trace_id = -1;           // >= 0: this group is written to the trace file
  bool on(int lane) const { return lane < cnt && ((active >> lane) & 1u); }
  float top(int lane) const { return heap[lane]; }
  void refresh_max() {
    maxR2 = 0.f;
    for (int l = 0; l < cnt; ++l) if (on(l)) maxR2 = fmaxf(maxR2, top(l));
  }
  void refresh_box() {
    for (int a = 0; a < 3; ++a) { glo[a] = INFINITY; ghi[a] = -INFINITY; }
    for (int l = 0; l < cnt; ++l) {
      if (!on(l)) continue;
      for (int a = 0; a < 3; ++a) { glo[a] = fminf(glo[a], q[l][a]); ghi[a] = fmaxf(ghi[a], q[l][a]); }
    }
  }
};

void leaf(const CpuTree& t, int64_t leaf_idx, Group& g) {
  const Box& b = t.boxes[0][leaf_idx];
  bool any = false;
  for (int l = 0; l < g.cnt && !any; ++l)
    any = g.on(l) && point_box_lb(g.q[l][0], g.q[l][1], g.q[l][2], b.lox, b.loy, b.loz, b.hix, b.hiy, b.hiz) < g.top(l);
  if (!any) return;
  if (g.trace_id >= 0) {
    FILE* f = trace_file();
    const int32_t head[2] = {(int32_t)g.leaves, (int32_t)g.cnt};
    float tops[kGroup], d2[kGroup][kGroup];
    for (int l = 0; l < kGroup; ++l) {
      tops[l] = g.on(l) ? g.top(l) : -INFINITY;       // lanes that sit this pass out accept nothing
      for (int j = 0; j < kGroup; ++j) {
        const Float4& p = t.pts[leaf_idx * kGroup + j];
        d2[l][j] = l < g.cnt ? dist2(g.q[l][0], g.q[l][1], g.q[l][2], p.x, p.y, p.z) : INFINITY;
      }
    }
    fwrite(&g.trace_id, sizeof(int64_t), 1, f);
    fwrite(head, sizeof(head), 1, f);
    fwrite(tops, sizeof(tops), 1, f);
    fwrite(d2, sizeof(d2), 1, f);
  }
  for (int l = 0; l < g.cnt; ++l) {
    if (!g.on(l)) continue;
    for (int j = 0; j < kGroup; ++j) {
      const Float4& p = t.pts[leaf_idx * kGroup + j];
      const float d2 = dist2(g.q[l][0], g.q[l][1], g.q[l][2], p.x, p.y, p.z);
      if (d2 < g.top(l)) { heap_replace_top(g.heap.data() + l, kGroup, g.k, d2); ++g.inserts; }
    }
  }
  ++g.leaves;
  g.refresh_max();
}

void visit(const CpuTree& t, int level, int64_t base, int count, Group& g) {
  float key[kFanout];
  bool todo[kFanout];
  for (int c = 0; c < count; ++c) {
    const Box& b = t.boxes[level][base + c];
    key[c] = box_box_lb(g.glo[0], g.glo[1], g.glo[2], g.ghi[0], g.ghi[1], g.ghi[2],
                        b.lox, b.loy, b.loz, b.hix, b.hiy, b.hiz);
    todo[c] = true;
  }
  for (;;) {
    int who = -1;
    for (int c = 0; c < count; ++c) {
      if (!todo[c]) continue;
      if (!(key[c] < g.maxR2)) { todo[c] = false; continue; }
      if (who < 0 || key[c] < key[who]) who = c;
    }
    if (who < 0) break;
    todo[who] = false;
    const int64_t child = base + who;
    if (level == 0) {
      leaf(t, child, g);
    } else {
      ++g.inner;
      const int64_t rem = (int64_t)t.boxes[level - 1].size() - child * kFanout;
      visit(t, level - 1, child * kFanout, (int)std::min<int64_t>(rem, kFanout), g);
    }
  }
}

// Best-first traversal: one frontier of (lower bound, node) over all levels, always expanding the
// globally nearest entry; stops when the nearest entry is no closer than the group radius.
// Mirror of the warp-level frontier queue in query.cu.
void best_first(const CpuTree& t, Group& g, int64_t seed_l1) {
  struct Entry { float key; int level; int64_t idx; };
  auto worse = [](const Entry& a, const Entry& b) { return a.key > b.key; };
  std::vector<Entry> q;
  if (t.levels >= 2) {
    if (seed_l1 < 0) {   // greedy descent to the nearest level-1 node (group visiting a foreign tree)
      int64_t base = 0;
      int64_t count = (int64_t)t.boxes[t.levels - 1].size();
      for (int L = t.levels - 1;; --L) {
        int64_t bestc = base; float bk = INFINITY;
        for (int64_t c = base; c < base + count; ++c) {
          const Box& b = t.boxes[L][c];
          const float key = box_box_lb(g.glo[0], g.glo[1], g.glo[2], g.ghi[0], g.ghi[1], g.ghi[2],
                                       b.lox, b.loy, b.loz, b.hix, b.hiy, b.hiz);
          if (key < bk) { bk = key; bestc = c; }
        }
        if (L == 1) { seed_l1 = bestc; break; }
        base = bestc * kFanout;
        count = std::min<int64_t>(kFanout, (int64_t)t.boxes[L - 1].size() - base);
      }
    }
    const int64_t rem = (int64_t)t.boxes[0].size() - seed_l1 * kFanout;
    visit(t, 0, seed_l1 * kFanout, (int)std::min<int64_t>(rem, kFanout), g);   // seed phase
  } else {
    seed_l1 = -1;
  }
  auto push_children = [&](int level, int64_t base, int count) {
    for (int c = 0; c < count; ++c) {
      if (level == 1 && base + c == seed_l1) continue;
      const Box& b = t.boxes[level][base + c];
      const float key = box_box_lb(g.glo[0], g.glo[1], g.glo[2], g.ghi[0], g.ghi[1], g.ghi[2],
                                   b.lox, b.loy, b.loz, b.hix, b.hiy, b.hiz);
      if (key < g.maxR2) { q.push_back({key, level, base + c}); std::push_heap(q.begin(), q.end(), worse); }
    }
  };
  const int top = t.levels - 1;
  push_children(top, 0, (int)t.boxes[top].size());
  while (!q.empty()) {
    std::pop_heap(q.begin(), q.end(), worse);
    const Entry e = q.back();
    q.pop_back();
    if (!(e.key < g.maxR2)) break;
    if (e.level == 0) {
      leaf(t, e.idx, g);
    } else {
      ++g.inner;
      const int64_t rem = (int64_t)t.boxes[e.level - 1].size() - e.idx * kFanout;
      push_children(e.level - 1, e.idx * kFanout, (int)std::min<int64_t>(rem, kFanout));
    }
  }
}

}  // namespace
}  // namespace b200knn

using namespace b200knn;

extern "C" int b200knn_cpu_bruteforce(const float* xyz, int64_t n, const float* queries, int64_t nq, int k,
                                      float max_radius, float* out, int num_threads) {
  if (k < 1 || n < 0 || nq < 0 || (!xyz && n) || (!queries && nq) || (!out && nq)) {
    set_last_error("cpu_bruteforce: invalid argument");
    return -2;
  }
  const float sent = sentinel_d2(max_radius);
  parallel_for(nq, num_threads, [&](int64_t b, int64_t e) {
    std::vector<float> d(n);
    for (int64_t qi = b; qi < e; ++qi) {
      const float qx = queries[3 * qi], qy = queries[3 * qi + 1], qz = queries[3 * qi + 2];
      for (int64_t i = 0; i < n; ++i) d[i] = dist2(qx, qy, qz, xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
      float kth = INFINITY;
      if (n >= k) {
        std::nth_element(d.begin(), d.begin() + (k - 1), d.end());
        kth = d[k - 1];
      }
      out[qi] = finalize_kth(fminf(kth, sent));
    }
  });
  return 0;
}

extern "C" int b200knn_cpu_neighbors(const float* xyz, int64_t n, const float* queries, int64_t nq, int k,
                                     float max_radius, int32_t* out_ids, float* out_dist, int num_threads) {
  if (k < 1 || n < 0 || nq < 0 || (!xyz && n) || (!queries && nq) || ((!out_ids || !out_dist) && nq)) {
    set_last_error("cpu_neighbors: invalid argument");
    return -2;
  }
  const float sent = sentinel_d2(max_radius);
  parallel_for(nq, num_threads, [&](int64_t b, int64_t e) {
    std::vector<std::pair<float, int32_t>> d(n);
    for (int64_t qi = b; qi < e; ++qi) {
      const float qx = queries[3 * qi], qy = queries[3 * qi + 1], qz = queries[3 * qi + 2];
      for (int64_t i = 0; i < n; ++i)
        d[i] = {dist2(qx, qy, qz, xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]), (int32_t)i};
      const int64_t m = std::min<int64_t>(k, n);
      std::partial_sort(d.begin(), d.begin() + m, d.end());
      for (int64_t s = 0; s < k; ++s) {
        const bool have = s < m && d[s].first < sent;      // a candidate must beat the cut-off, as in the heap
        out_ids[qi * k + s] = have ? d[s].second : -1;
        out_dist[qi * k + s] = have ? finalize_kth(d[s].first) : INFINITY;
      }
    }
  });
  return 0;
}

extern "C" int b200knn_cpu_stats(uint64_t out4[4], int reset) {
  for (int i = 0; i < 4; ++i) { out4[i] = g_cpu_stats[i].load(); if (reset) g_cpu_stats[i] = 0; }
  return 0;
}

extern "C" int b200knn_cpu_knn(const float* xyz, int64_t n, const float* queries, int64_t nq, int k,
                               float max_radius, float* out, int num_threads) {
  if (k < 1 || n < 0 || (!xyz && n)) {
    set_last_error("cpu_knn: invalid argument");
    return -2;
  }
  CpuTree tree;
  build_tree(xyz, n, tree);
  // queries: the indexed points themselves (sorted order = groups are leaves), or an external set
  // ordered along the tree's Morton grid.
  std::vector<Float4> qs;
  int64_t qn;
  if (queries == nullptr) {
    qs = tree.pts;
    qn = n;
  } else {
    qn = nq;
    const std::vector<uint32_t> order = morton_order(queries, nq, tree.lo, tree.hi);
    qs.resize(div_up(std::max<int64_t>(nq, 1), kGroup) * kGroup);
    for (int64_t i = 0; i < nq; ++i) {
      const uint32_t s = order[i];
      Float4 p{queries[3 * (size_t)s], queries[3 * (size_t)s + 1], queries[3 * (size_t)s + 2], 0.f};
      memcpy(&p.w, &s, 4);
      qs[i] = p;
    }
  }
  const float sent = sentinel_d2(max_radius);
  const int64_t groups = div_up(qn, kGroup);
  parallel_for(groups, num_threads, [&](int64_t gb, int64_t ge) {
    Group g;
    g.k = k;
    g.heap.resize((size_t)k * kGroup);
    for (int64_t gi = gb; gi < ge; ++gi) {
      g.cnt = (int)std::min<int64_t>(kGroup, qn - gi * kGroup);
      g.trace_id = (num_threads == 1 && gi % 64 == 0 && trace_file()) ? gi : -1;
      std::fill(g.heap.begin(), g.heap.end(), sent);
      for (int l = 0; l < g.cnt; ++l) {
        const Float4& p = qs[gi * kGroup + l];
        g.q[l][0] = p.x; g.q[l][1] = p.y; g.q[l][2] = p.z;
      }
      if (tree.levels > 0) {
        static const bool dfs = getenv("B200KNN_CPU_DFS") != nullptr;
        const int top = tree.levels - 1;
        // Same lane-set splitting as traverse_group (traverse.cuh): cut at the largest gap along the
        // widest axis while that gap exceeds two thirds of the extent; traverse the parts separately.
        const unsigned all = g.cnt >= 32 ? 0xffffffffu : ((1u << g.cnt) - 1u);
        std::vector<unsigned> work{all};
        while (!work.empty()) {
          const unsigned m = work.back();
          work.pop_back();
          g.active = m;
          g.refresh_box();
          g.refresh_max();
          const float ext[3] = {g.ghi[0] - g.glo[0], g.ghi[1] - g.glo[1], g.ghi[2] - g.glo[2]};
          const int axis = (ext[0] >= ext[1] && ext[0] >= ext[2]) ? 0 : (ext[1] >= ext[2] ? 1 : 2);
          float best_gap = 0.f, cut = 0.f;
          for (int l = 0; l < g.cnt; ++l) {
            if (!g.on(l)) continue;
            float succ = INFINITY;
            for (int o = 0; o < g.cnt; ++o) if (g.on(o) && g.q[o][axis] > g.q[l][axis]) succ = fminf(succ, g.q[o][axis]);
            if (succ < INFINITY && succ - g.q[l][axis] > best_gap) { best_gap = succ - g.q[l][axis]; cut = g.q[l][axis]; }
          }
          if (best_gap > 0.66f * ext[axis] && work.size() <= 6) {
            unsigned left = 0;
            for (int l = 0; l < g.cnt; ++l) if (g.on(l) && g.q[l][axis] <= cut) left |= 1u << l;
            const unsigned right = m & ~left;
            if (left && right) { work.push_back(right); work.push_back(left); continue; }
          }
          if (dfs) visit(tree, top, 0, (int)tree.boxes[top].size(), g);
          else best_first(tree, g, (queries == nullptr && m == all) ? gi / kFanout : -1);
        }
        g.active = 0xffffffffu;
      }
      for (int l = 0; l < g.cnt; ++l) {
        uint32_t id;
        memcpy(&id, &qs[gi * kGroup + l].w, 4);
        out[id] = finalize_kth(g.top(l));
      }
    }
    g_cpu_stats[0] += g.leaves; g_cpu_stats[1] += g.leaves * kGroup;
    g_cpu_stats[2] += g.inserts; g_cpu_stats[3] += g.inner;
  });
  if (num_threads == 1 && trace_file()) fflush(trace_file());
  return 0;
}
