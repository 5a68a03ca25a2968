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
//     The group walks the 32-ary box tree once (lane = child box, nearest-first via redux.sync), so
//     node tests cost one instruction stream per 32 queries and leaf points arrive by one coalesced
//     128-bit load per lane, then are broadcast from shared memory (LDS.128, no bank conflicts).
//   * every query's k-heap lives in shared memory, one column per lane (bank == lane: conflict-free
//     at any heap index), 4-ary to halve the dependent LDS chain.
//   * leaf = two phases: (1) convergent: 32 distances -> bitmask of candidates under the lane's
//     current k-th distance; (2) divergent only over set bits: re-derive d2, sift into the heap.
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

constexpr int Q_THREADS = 256;
constexpr int Q_WARPS = Q_THREADS / 32;

struct GroupCtx {
  float qx, qy, qz;     // this lane's query
  float top;            // this lane's current k-th squared distance (heap root)
  float glx, gly, glz, ghx, ghy, ghz;   // bounding box of the group's valid queries (warp-uniform)
  float maxR2;          // max over valid lanes of top (warp-uniform)
  float* heap;          // shared: heap[i*32] is entry i of this lane
  float4* stage;        // shared: 32 leaf points
  int k;
  bool valid;
  unsigned n_leaves, n_inner, n_inserts;
};

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
      g.top = heap_replace_top(g.heap, 32, g.k, d2);
      ++g.n_inserts;
    }
  }
  __syncwarp();
  g.maxR2 = __uint_as_float(__reduce_max_sync(kFullMask, g.valid ? __float_as_uint(g.top) : 0u));
  ++g.n_leaves;
}

template <int LEVEL>
__device__ void visit(const TreeView& t, int base, int count, GroupCtx& g) {
  const int lane = lane_id();
  const bool have = lane < count;
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
  if (have) {
    const float4* bp = reinterpret_cast<const float4*>(t.boxes[LEVEL]) + 2 * (size_t)(base + lane);
    a = __ldg(bp);
    b = __ldg(bp + 1);
  }
  const float key = box_box_lb(g.glx, g.gly, g.glz, g.ghx, g.ghy, g.ghz, a.x, a.y, a.z, b.x, b.y, b.z);
  unsigned todo = __ballot_sync(kFullMask, have);
  for (;;) {
    // the group radius only shrinks, so a child that fails once never needs re-testing
    const bool pass = ((todo >> lane) & 1u) && (key < g.maxR2);
    todo = __ballot_sync(kFullMask, pass);
    if (todo == 0u) break;
    const unsigned kb = pass ? __float_as_uint(key) : 0xffffffffu;
    const unsigned m = __reduce_min_sync(kFullMask, kb);
    const int who = __ffs(__ballot_sync(kFullMask, kb == m)) - 1;   // nearest child first
    todo &= ~(1u << who);
    const int child = base + who;
    if constexpr (LEVEL == 0) {
      const float lx = __shfl_sync(kFullMask, a.x, who), ly = __shfl_sync(kFullMask, a.y, who);
      const float lz = __shfl_sync(kFullMask, a.z, who), hx = __shfl_sync(kFullMask, b.x, who);
      const float hy = __shfl_sync(kFullMask, b.y, who), hz = __shfl_sync(kFullMask, b.z, who);
      const int cnt = __float_as_int(__shfl_sync(kFullMask, a.w, who));
      // exact per-query test against the leaf box; skip the leaf if no lane can improve
      const float lb = point_box_lb(g.qx, g.qy, g.qz, lx, ly, lz, hx, hy, hz);
      if (__any_sync(kFullMask, g.valid && (lb < g.top))) process_leaf(t, child, cnt, g);
    } else {
      ++g.n_inner;
      const int rem = t.count[LEVEL - 1] - child * kFanout;
      visit<LEVEL - 1>(t, child * kFanout, rem < kFanout ? rem : kFanout, g);
    }
  }
}

__device__ __forceinline__ void traverse(const TreeView& t, GroupCtx& g) {
  const int top_level = t.levels - 1;
  switch (top_level) {
    case 0: visit<0>(t, 0, t.count[0], g); break;
    case 1: visit<1>(t, 0, t.count[1], g); break;
    case 2: visit<2>(t, 0, t.count[2], g); break;
    case 3: visit<3>(t, 0, t.count[3], g); break;
    case 4: visit<4>(t, 0, t.count[4], g); break;
    case 5: visit<5>(t, 0, t.count[5], g); break;
    default: break;
  }
}

__global__ void __launch_bounds__(Q_THREADS, 2)
knn_ring_kernel(const RingParams p) {
  extern __shared__ __align__(16) float q_smem[];
  const int lane = lane_id();
  const int warp = threadIdx.x >> 5;
  const int k = p.k;
  const size_t per_warp_floats = (size_t)k * 32 + 128;
  float* heap_base = q_smem + warp * per_warp_floats;

  GroupCtx g;
  g.heap = heap_base + lane;
  g.stage = reinterpret_cast<float4*>(heap_base + (size_t)k * 32);
  g.k = k;
  g.n_leaves = g.n_inner = g.n_inserts = 0;

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
      traverse(p.tree, g);
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

  if (p.stats != nullptr && lane == 0) {
    atomicAdd(p.stats + 0, (unsigned long long)g.n_leaves);
    atomicAdd(p.stats + 1, (unsigned long long)g.n_leaves * kGroup);
    atomicAdd(p.stats + 3, (unsigned long long)g.n_inner);
  }
  if (p.stats != nullptr) {
    unsigned ins = g.n_inserts;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ins += __shfl_xor_sync(kFullMask, ins, o);
    if (lane == 0) atomicAdd(p.stats + 2, (unsigned long long)ins);
  }
}

}  // namespace

size_t ring_smem_bytes(int k, int warps_per_cta) {
  return (size_t)warps_per_cta * ((size_t)k * 32 + 128) * sizeof(float);
}

// Warps per CTA: as many as fit the 227 KB shared-memory limit, at most 8 (two 8-warp CTAs per SM at
// k = 100). Returns 0 if even one warp does not fit (k > 1811).
int ring_pick_warps_per_cta(int k) {
  const size_t per_warp = ring_smem_bytes(k, 1);
  size_t w = (size_t)(227 * 1024) / per_warp;
  if (w > (size_t)Q_WARPS) w = Q_WARPS;
  return (int)w;
}

int ring_launch(const RingParams& p, int num_sms, cudaStream_t stream) {
  KNN_REQUIRE(p.k >= 1, "k must be >= 1");
  const int warps = ring_pick_warps_per_cta(p.k);
  KNN_REQUIRE(warps >= 1, "k too large for the shared-memory heap (k <= 1811)");
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
