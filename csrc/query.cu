// query.cu — the persistent kNN kernel: bounded-k-heap group traversal fused with the ring hand-off.
//
// What it replaces in the reference:
//   * runQuery (unorderedDataVariant.cu:75-87): one thread per query, stack-free k-d traversal
//     (cukd/traverse-stack-free.h:24-117), heap of k uint64 in *global* memory at an 8k-byte stride
//     (cukd/knn.h:328-389).
//   * the ring of MPI_Isend/Irecv/Waitall + cudaDeviceSynchronize around it
//     (unorderedDataVariant.cu:173-205), and extractFinalResult (:89-103).
//
// B200 design:
//   * a warp owns a GROUP of 32 spatially adjacent queries (one leaf of their home tree); lane = query.
//     The group walks the 32-ary box tree once. Inner nodes: lane = child box.
//   * traversal is BEST-FIRST over a per-warp frontier in shared memory (lower bound, node) spanning
//     all levels: always open the globally nearest entry (strided argmin + redux.sync), stop when the
//     nearest entry is no closer than the group's k-th radius. Measured (tools/cpu_stats.py, mirror
//     engine): 2.2x fewer leaves and 2.7x fewer heap insertions than per-node nearest-first DFS.
//     Frontier full -> purge stale entries -> if still full, open that node depth-first (always exact).
//   * leaf points arrive by one coalesced 128-bit load per lane and are broadcast from shared memory
//     (LDS.128). Leaf = two phases: (1) convergent: 32 distances -> bitmask of candidates under the
//     lane's current k-th distance; (2) divergent only over set bits: re-derive d2, sift into the heap.
//   * every query's k-heap lives in shared memory, one column per lane (bank == lane: conflict-free
//     at any heap index), 4-ary and padded with -inf to 4m+1 entries so a sift level is 4 unguarded
//     loads.
//   * ring: groups travel with their heaps. A group's state is read from the local inbox (which the
//     previous GPU filled with plain st.global over NVLink), traversed against the local tree, and
//     written straight into the NEXT GPU's inbox, followed by fence + st.release.sys of a sequence
//     number; the consumer polls with ld.acquire.sys and acknowledges the slot the same way. No
//     host round-trip, no collective call between hops; transfer of group g overlaps traversal of
//     every other group in flight.
#include "common.cuh"
#include "kernels.h"

namespace b200knn {

namespace {

constexpr int Q_MAX_WARPS = 16;        // per CTA
constexpr int kQueueCap = 256;         // frontier entries per warp (2 KB)
constexpr unsigned kNodeIdxBits = 28;  // node = level << 28 | index

__host__ __device__ constexpr int heap_padded(int k) { return 4 * ((k - 1 + 3) / 4) + 1; }
__host__ __device__ constexpr size_t warp_smem_floats(int k) {
  return (size_t)heap_padded(k) * 32 + 128 + 2 * kQueueCap + 32;
}

struct GroupCtx {
  float qx, qy, qz;     // this lane's query
  float top;            // this lane's current k-th squared distance (heap root)
  float glx, gly, glz, ghx, ghy, ghz;   // bounding box of the group's valid queries (warp-uniform)
  float maxR2;          // max over valid lanes of top (warp-uniform)
  float* heap;          // shared: heap[i*32] is entry i of this lane
  float4* stage;        // shared: 32 leaf points
  unsigned* qkey;       // shared: frontier lower bounds (float bits; non-negative => uint-ordered)
  unsigned* qnode;      // shared: frontier nodes
  unsigned* frames;     // shared: frame stack of the depth-first fallback
  int kp;               // padded heap size
  bool valid;
  unsigned n_leaves, n_inner, n_inserts, n_fallback, n_purge;
};

// Replace the root (precondition e < root) and sift down. Children of i: 4i+1..4i+4, all present.
__device__ __forceinline__ float heap_replace_padded(float* h, int kp, float e) {
  int i = 0;
  for (;;) {
    const int c0 = 4 * i + 1;
    if (c0 >= kp) break;
    const float a0 = h[c0 * 32], a1 = h[(c0 + 1) * 32], a2 = h[(c0 + 2) * 32], a3 = h[(c0 + 3) * 32];
    const float m01 = fmaxf(a0, a1), m23 = fmaxf(a2, a3);
    const int i01 = a1 > a0 ? c0 + 1 : c0, i23 = a3 > a2 ? c0 + 3 : c0 + 2;
    const float best = fmaxf(m01, m23);
    const int bi = m23 > m01 ? i23 : i01;
    if (!(best > e)) break;
    h[i * 32] = best;
    i = bi;
  }
  h[i * 32] = e;
  return h[0];
}

__device__ __forceinline__ void process_leaf(const TreeView& t, int leaf, int cnt, GroupCtx& g) {
  const int lane = lane_id();
  float4 p = make_float4(INFINITY, INFINITY, INFINITY, 0.f);   // padding never beats any top
  if (lane < cnt) p = __ldg(reinterpret_cast<const float4*>(t.pts) + (size_t)leaf * kGroup + lane);
  g.stage[lane] = p;
  __syncwarp();

  // phase 1: convergent distance sweep
  unsigned mask = 0;
  const float top0 = g.top;
#pragma unroll
  for (int j = 0; j < kGroup; ++j) {
    const float4 c = g.stage[j];
    const float d2 = dist2(g.qx, g.qy, g.qz, c.x, c.y, c.z);
    mask |= (d2 < top0) ? (1u << j) : 0u;
  }
  if (!g.valid) mask = 0;

  // phase 2: only lanes with survivors loop, only over their survivors
  while (mask) {
    const int j = __ffs(mask) - 1;
    mask &= mask - 1;
    const float4 c = g.stage[j];
    const float d2 = dist2(g.qx, g.qy, g.qz, c.x, c.y, c.z);
    if (d2 < g.top) {
      g.top = heap_replace_padded(g.heap, g.kp, d2);
      ++g.n_inserts;
    }
  }
  __syncwarp();
  g.maxR2 = __uint_as_float(__reduce_max_sync(kFullMask, g.valid ? __float_as_uint(g.top) : 0u));
  ++g.n_leaves;
}

// Exact per-query test against the leaf box; skip the leaf if no lane can improve.
__device__ __forceinline__ void maybe_leaf(const TreeView& t, int leaf, float lx, float ly, float lz,
                                           float hx, float hy, float hz, int cnt, GroupCtx& g) {
  const float lb = point_box_lb(g.qx, g.qy, g.qz, lx, ly, lz, hx, hy, hz);
  if (__any_sync(kFullMask, g.valid && (lb < g.top))) process_leaf(t, leaf, cnt, g);
}

// ---- traversal ------------------------------------------------------------------------------------
// One loop, two sources of "next node", one leaf call site:
//   * FRAMES (depth >= 0): depth-first, nearest child first, explicit frame stack in shared memory
//     (level, base, count, todo mask); boxes are re-loaded when control returns to a frame. Used (a) for
//     the SEED phase — the 32 leaves of the level-1 node the group lives in (or the nearest one, for a
//     group visiting a foreign tree) are swept first so that the radius is tight before anything is
//     put on the frontier — and (b) when the frontier overflows (any visiting order is exact).
//   * FRONTIER: best-first over (lower bound, node) entries of all levels, seed node excluded.
__device__ __forceinline__ unsigned full_mask_of(int count) {
  return count >= 32 ? 0xffffffffu : ((1u << count) - 1u);
}

__device__ __forceinline__ void push_frame(unsigned* fr, int depth, int level, int base, int count) {
  fr[depth * 4 + 0] = (unsigned)level;
  fr[depth * 4 + 1] = (unsigned)base;
  fr[depth * 4 + 2] = (unsigned)count;
  fr[depth * 4 + 3] = full_mask_of(count);
}

// Greedy descent to the level-1 node nearest to the group box (-1 if the tree has no level 1).
__device__ __forceinline__ int find_seed_l1(const TreeView& t, const GroupCtx& g) {
  if (t.levels < 2) return -1;
  const int lane = lane_id();
  int base = 0, count = t.count[t.levels - 1];
  for (int L = t.levels - 1;; --L) {
    float key = INFINITY;
    if (lane < count) {
      const float4* bp = reinterpret_cast<const float4*>(t.boxes[L]) + 2 * (size_t)(base + lane);
      const float4 a = __ldg(bp), b = __ldg(bp + 1);
      key = box_box_lb(g.glx, g.gly, g.glz, g.ghx, g.ghy, g.ghz, a.x, a.y, a.z, b.x, b.y, b.z);
    }
    const unsigned kb = __float_as_uint(key);
    const unsigned m = __reduce_min_sync(kFullMask, kb);
    const int child = base + __ffs(__ballot_sync(kFullMask, kb == m)) - 1;
    if (L == 1) return child;
    base = child * kFanout;
    const int rem = t.count[L - 1] - base;
    count = rem < kFanout ? rem : kFanout;
  }
}

// Push the children [base, base+count) of `level` that lie inside the group radius. On overflow
// (after purging stale entries) the children are handed to the frame stack instead.
__device__ __forceinline__ void frontier_expand(const TreeView& t, int level, int base, int count, int seed_l1,
                                                int& qcount, int& depth, GroupCtx& g) {
  const int lane = lane_id();
  const bool have = lane < count;
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
  if (have) {
    const float4* bp = reinterpret_cast<const float4*>(t.boxes[level]) + 2 * (size_t)(base + lane);
    a = __ldg(bp);
    b = __ldg(bp + 1);
  }
  const float key = box_box_lb(g.glx, g.gly, g.glz, g.ghx, g.ghy, g.ghz, a.x, a.y, a.z, b.x, b.y, b.z);
  const bool pass = have && (key < g.maxR2) && !(level == 1 && base + lane == seed_l1);
  const unsigned mask = __ballot_sync(kFullMask, pass);
  const int np = __popc(mask);
  if (np == 0) return;
  const unsigned lt = (1u << lane) - 1u;
  if (qcount + np > kQueueCap) {
    // purge entries that the shrinking radius has made irrelevant (in-place, order-preserving)
    const unsigned maxbits = __float_as_uint(g.maxR2);
    int kept = 0;
    for (int s0 = 0; s0 < qcount; s0 += 32) {
      const int s = s0 + lane;
      const bool in = s < qcount;
      const unsigned kx = in ? g.qkey[s] : 0xffffffffu;
      const unsigned nd = in ? g.qnode[s] : 0u;
      const bool keep = in && (kx < maxbits);
      const unsigned km = __ballot_sync(kFullMask, keep);
      __syncwarp();
      if (keep) {
        const int pos = kept + __popc(km & lt);
        g.qkey[pos] = kx;
        g.qnode[pos] = nd;
      }
      kept += __popc(km);
      __syncwarp();
    }
    qcount = kept;
    ++g.n_purge;
  }
  if (qcount + np > kQueueCap) {
    ++g.n_fallback;
    depth = 0;
    if (lane == 0) push_frame(g.frames, 0, level, base, count);
    __syncwarp();
    return;
  }
  if (pass) {
    const int pos = qcount + __popc(mask & lt);
    g.qkey[pos] = __float_as_uint(key);
    g.qnode[pos] = ((unsigned)level << kNodeIdxBits) | (unsigned)(base + lane);
  }
  qcount += np;
  __syncwarp();
}

__device__ __forceinline__ void traverse(const TreeView& t, GroupCtx& g, int seed_l1) {
  const int lane = lane_id();
  unsigned* fr = g.frames;
  const int top_level = t.levels - 1;
  int depth = -1, qcount = 0;
  bool started = false;
  if (seed_l1 >= 0) {
    const int rem = t.count[0] - seed_l1 * kFanout;
    if (lane == 0) push_frame(fr, 0, 0, seed_l1 * kFanout, rem < kFanout ? rem : kFanout);
    depth = 0;
    __syncwarp();
  }
  for (;;) {
    int leaf, cnt;
    float lx, ly, lz, hx, hy, hz;
    if (depth >= 0) {
      // ---- frame step
      const int L = (int)fr[depth * 4 + 0], B = (int)fr[depth * 4 + 1], Cn = (int)fr[depth * 4 + 2];
      const unsigned todo_in = fr[depth * 4 + 3];
      const bool have = lane < Cn;
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
      if (have) {
        const float4* bp = reinterpret_cast<const float4*>(t.boxes[L]) + 2 * (size_t)(B + lane);
        a = __ldg(bp);
        b = __ldg(bp + 1);
      }
      const float key = box_box_lb(g.glx, g.gly, g.glz, g.ghx, g.ghy, g.ghz, a.x, a.y, a.z, b.x, b.y, b.z);
      const bool pass = have && ((todo_in >> lane) & 1u) && (key < g.maxR2) &&
                        !(L == 1 && B + lane == seed_l1);
      const unsigned pm = __ballot_sync(kFullMask, pass);
      if (pm == 0u) { --depth; continue; }
      const unsigned kb = pass ? __float_as_uint(key) : 0xffffffffu;
      const unsigned m = __reduce_min_sync(kFullMask, kb);
      const int who = __ffs(__ballot_sync(kFullMask, kb == m)) - 1;
      const int child = B + who;
      __syncwarp();
      if (lane == 0) fr[depth * 4 + 3] = pm & ~(1u << who);
      if (L != 0) {
        ++g.n_inner;
        ++depth;
        const int rem = t.count[L - 1] - child * kFanout;
        if (lane == 0) push_frame(fr, depth, L - 1, child * kFanout, rem < kFanout ? rem : kFanout);
        __syncwarp();
        continue;
      }
      __syncwarp();
      leaf = child;
      lx = __shfl_sync(kFullMask, a.x, who); ly = __shfl_sync(kFullMask, a.y, who);
      lz = __shfl_sync(kFullMask, a.z, who); hx = __shfl_sync(kFullMask, b.x, who);
      hy = __shfl_sync(kFullMask, b.y, who); hz = __shfl_sync(kFullMask, b.z, who);
      cnt = __float_as_int(__shfl_sync(kFullMask, a.w, who));
    } else {
      // ---- frontier step
      if (!started) {
        started = true;
        frontier_expand(t, top_level, 0, t.count[top_level], seed_l1, qcount, depth, g);
        continue;
      }
      if (qcount == 0) break;
      unsigned best = 0xffffffffu;
      int bslot = 0;
      for (int s = lane; s < qcount; s += 32) {
        const unsigned kx = g.qkey[s];
        if (kx < best) { best = kx; bslot = s; }
      }
      const unsigned m = __reduce_min_sync(kFullMask, best);
      if (m >= __float_as_uint(g.maxR2)) break;          // nothing left inside the group radius
      const int wl = __ffs(__ballot_sync(kFullMask, best == m)) - 1;
      const int slot = __shfl_sync(kFullMask, bslot, wl);
      const unsigned node = g.qnode[slot];
      __syncwarp();
      if (lane == 0) {
        g.qkey[slot] = g.qkey[qcount - 1];
        g.qnode[slot] = g.qnode[qcount - 1];
      }
      --qcount;
      __syncwarp();
      const int level = (int)(node >> kNodeIdxBits);
      const int idx = (int)(node & ((1u << kNodeIdxBits) - 1u));
      if (level != 0) {
        ++g.n_inner;
        const int rem = t.count[level - 1] - idx * kFanout;
        frontier_expand(t, level - 1, idx * kFanout, rem < kFanout ? rem : kFanout, seed_l1, qcount, depth, g);
        continue;
      }
      const float4* bp = reinterpret_cast<const float4*>(t.boxes[0]) + 2 * (size_t)idx;
      const float4 a = __ldg(bp), b = __ldg(bp + 1);
      leaf = idx;
      lx = a.x; ly = a.y; lz = a.z; hx = b.x; hy = b.y; hz = b.z;
      cnt = __float_as_int(a.w);
    }
    maybe_leaf(t, leaf, lx, ly, lz, hx, hy, hz, cnt, g);
  }
}

__global__ void __launch_bounds__(Q_MAX_WARPS * 32, 1)
knn_ring_kernel(const RingParams p) {
  extern __shared__ __align__(16) float q_smem[];
  const int lane = lane_id();
  const int warp = threadIdx.x >> 5;
  const int k = p.k;
  const int kp = heap_padded(k);
  float* warp_base = q_smem + warp * warp_smem_floats(k);

  GroupCtx g;
  g.heap = warp_base + lane;
  g.stage = reinterpret_cast<float4*>(warp_base + (size_t)kp * 32);
  g.qkey = reinterpret_cast<unsigned*>(warp_base + (size_t)kp * 32 + 128);
  g.qnode = g.qkey + kQueueCap;
  g.frames = g.qnode + kQueueCap;
  g.kp = kp;
  g.n_leaves = g.n_inner = g.n_inserts = g.n_fallback = g.n_purge = 0;

  const int R = p.R;
  const int Gc = p.groups_per_chunk;
  const unsigned long long first_task = (unsigned long long)p.step_begin * Gc;
  const unsigned long long total = (unsigned long long)(p.step_end - p.step_begin) * Gc;

  for (;;) {
    unsigned long long t = 0;
    if (lane == 0) t = atomicAdd(p.task_counter, 1ull);
    t = __shfl_sync(kFullMask, t, 0);
    if (t >= total) break;
    t += first_task;
    const int s = (int)(t / Gc);
    const int grp = (int)(t % Gc);
    const int hop = s % R;
    const int chunk = s / R;
    const int par = s & 1;
    const int64_t home_group = (int64_t)chunk * Gc + grp;

    // ------------------------------------------------------------------ input
    int cnt;
    if (hop == 0) {
      const int64_t rem = p.home_n - home_group * kGroup;
      cnt = rem <= 0 ? 0 : (rem < kGroup ? (int)rem : kGroup);
      if (cnt > 0) {
        const int src = lane < cnt ? lane : cnt - 1;
        const float4 q = __ldg(reinterpret_cast<const float4*>(p.home_queries) + home_group * kGroup + src);
        g.qx = q.x; g.qy = q.y; g.qz = q.z;
        const float sent = p.sentinel;
        for (int i = 0; i < k; ++i) g.heap[i * 32] = sent;
        g.top = sent;
      }
    } else {
      if (p.use_flags) {
        if (lane == 0) wait_flag_ge(p.flag_in[par] + grp, p.seq_base + (unsigned)s);
        __syncwarp();
      }
      cnt = ldcg_i(p.inbox_cnt[par] + grp);
      if (cnt > 0) {
        const float4 q = ldcg_f4(reinterpret_cast<const float4*>(p.inbox_pos[par]) + (size_t)grp * kGroup + lane);
        g.qx = q.x; g.qy = q.y; g.qz = q.z;
        const float* src = p.inbox_heap[par] + (size_t)grp * k * 32 + lane;
        for (int i = 0; i < k; ++i) g.heap[i * 32] = ldcg_f(src + (size_t)i * 32);
        g.top = g.heap[0];
      }
      if (p.use_flags) {
        __syncwarp();   // every lane's loads have returned (their values were stored to smem)
        if (lane == 0) st_release_sys(p.prev_ack[par] + grp, p.seq_base + (unsigned)s);
      }
    }
    g.valid = lane < cnt;

    // ------------------------------------------------------------------ traverse the local tree
    if (cnt > 0 && p.tree.n > 0) {
      for (int i = k; i < kp; ++i) g.heap[i * 32] = -INFINITY;   // heap padding
      float lx = g.valid ? g.qx : INFINITY, ly = g.valid ? g.qy : INFINITY, lz = g.valid ? g.qz : INFINITY;
      float hx = g.valid ? g.qx : -INFINITY, hy = g.valid ? g.qy : -INFINITY, hz = g.valid ? g.qz : -INFINITY;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        lx = fminf(lx, __shfl_xor_sync(kFullMask, lx, o));
        ly = fminf(ly, __shfl_xor_sync(kFullMask, ly, o));
        lz = fminf(lz, __shfl_xor_sync(kFullMask, lz, o));
        hx = fmaxf(hx, __shfl_xor_sync(kFullMask, hx, o));
        hy = fmaxf(hy, __shfl_xor_sync(kFullMask, hy, o));
        hz = fmaxf(hz, __shfl_xor_sync(kFullMask, hz, o));
      }
      g.glx = lx; g.gly = ly; g.glz = lz; g.ghx = hx; g.ghy = hy; g.ghz = hz;
      g.maxR2 = __uint_as_float(__reduce_max_sync(kFullMask, g.valid ? __float_as_uint(g.top) : 0u));
      // seed = the level-1 node holding the group's own leaf (self queries at home), else the nearest
      const int seed = (hop == 0 && p.home_is_tree) ? (p.tree.levels >= 2 ? (int)(home_group / kFanout) : -1)
                                                    : find_seed_l1(p.tree, g);
      traverse(p.tree, g, seed);
    }

    // ------------------------------------------------------------------ output
    if (hop == R - 1) {
      if (g.valid) p.result_sorted[home_group * kGroup + lane] = finalize_kth(g.top);
    } else {
      const int op = par ^ 1;
      if (p.use_flags) {
        // The slot we are about to fill was last filled by our step s2 (same parity, a writing
        // hop) and is free once the consumer acknowledged step s2+1.
        int s2 = s - 2;
        while (s2 >= p.step_begin && (s2 % R) == R - 1) s2 -= 2;
        if (s2 >= p.step_begin) {
          if (lane == 0) wait_flag_ge(p.ack_in[op] + grp, p.seq_base + (unsigned)(s2 + 1));
          __syncwarp();
        }
      }
      if (lane == 0) p.next_cnt[op][grp] = cnt;
      if (cnt > 0) {
        reinterpret_cast<float4*>(p.next_pos[op])[(size_t)grp * kGroup + lane] =
            make_float4(g.qx, g.qy, g.qz, 0.f);
        float* dst = p.next_heap[op] + (size_t)grp * k * 32 + lane;
        for (int i = 0; i < k; ++i) dst[(size_t)i * 32] = g.heap[i * 32];
      }
      if (p.use_flags) {
        __syncwarp();
        if (lane == 0) {
          __threadfence_system();
          st_release_sys(p.next_flag[op] + grp, p.seq_base + (unsigned)(s + 1));
        }
      }
    }
    __syncwarp();
  }

  if (p.stats != nullptr) {
    unsigned ins = g.n_inserts;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ins += __shfl_xor_sync(kFullMask, ins, o);
    if (lane == 0) {
      atomicAdd(p.stats + 0, (unsigned long long)g.n_leaves);
      atomicAdd(p.stats + 1, (unsigned long long)g.n_leaves * kGroup);
      atomicAdd(p.stats + 2, (unsigned long long)ins);
      atomicAdd(p.stats + 3, (unsigned long long)g.n_inner);
      atomicAdd(p.stats + 4, (unsigned long long)g.n_fallback);
      atomicAdd(p.stats + 5, (unsigned long long)g.n_purge);
    }
  }
}

constexpr size_t kSmemPerSM = 233472;     // 228 KB
constexpr size_t kSmemPerCtaMax = 232448; // 227 KB
constexpr size_t kSmemCtaReserve = 1024;

}  // namespace

size_t ring_smem_bytes(int k, int warps_per_cta) {
  return (size_t)warps_per_cta * warp_smem_floats(k) * sizeof(float);
}

// Shared memory decides the shape: maximise resident warps per SM over (CTAs per SM, warps per
// CTA); prefer more, smaller CTAs on ties. k = 100 -> one 15-warp CTA per SM. Returns 0 if not even
// one warp fits (k > ~1790).
int ring_pick_warps_per_cta(int k) {
  static int reg_warps = 0;     // resident warps the register file allows for this kernel
  if (reg_warps == 0) {
    cudaFuncAttributes fa;
    reg_warps = 32;
    if (cudaFuncGetAttributes(&fa, knn_ring_kernel) == cudaSuccess && fa.numRegs > 0) {
      const int regs = (fa.numRegs + 7) / 8 * 8;
      reg_warps = 65536 / (32 * regs);
      if (reg_warps > 32) reg_warps = 32;
      if (reg_warps < 1) reg_warps = 1;
    }
  }
  const size_t per_warp = ring_smem_bytes(k, 1);
  int best_w = 0, best_total = 0;
  for (int ctas = 4; ctas >= 1; --ctas) {
    for (int w = Q_MAX_WARPS; w >= 1; --w) {
      const size_t cta_bytes = (size_t)w * per_warp;
      if (cta_bytes > kSmemPerCtaMax) continue;
      if ((size_t)ctas * (cta_bytes + kSmemCtaReserve) > kSmemPerSM) continue;
      if (ctas * w > reg_warps) continue;
      if (ctas * w > best_total) { best_total = ctas * w; best_w = w; }
      break;
    }
  }
  return best_w;
}

int ring_launch(const RingParams& p, int num_sms, cudaStream_t stream) {
  KNN_REQUIRE(p.k >= 1, "k must be >= 1");
  const int warps = ring_pick_warps_per_cta(p.k);
  KNN_REQUIRE(warps >= 1, "k too large for the shared-memory heap (k <= 1790)");
  const size_t smem = ring_smem_bytes(p.k, warps);
  KNN_CUDA_TRY(cudaFuncSetAttribute(knn_ring_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  KNN_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, knn_ring_kernel, warps * 32, smem));
  KNN_REQUIRE(occ >= 1, "query kernel does not fit on an SM");
  const unsigned long long total = (unsigned long long)(p.step_end - p.step_begin) * p.groups_per_chunk;
  if (total == 0) return 0;
  unsigned long long want = (total + warps - 1) / warps;
  unsigned long long cap = (unsigned long long)num_sms * occ;
  const int grid = (int)(want < cap ? want : cap);
  KNN_CUDA_TRY(cudaMemsetAsync(p.task_counter, 0, sizeof(unsigned long long), stream));
  knn_ring_kernel<<<grid, warps * 32, smem, stream>>>(p);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

}  // namespace b200knn
