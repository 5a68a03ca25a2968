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
//   * leaf points arrive by one coalesced 128-bit load per lane and are broadcast from shared memory.
//     32 distances per lane on the packed fp32x2 pipe; each one that beats the lane's k-th distance is
//     appended to the lane's buffer by a predicated store.
//   * every query's k candidates live in shared memory, one column per lane (bank == lane:
//     conflict-free at any index), as a DESCENDING sorted list; buffered candidates are sorted in
//     registers and merged in by the whole warp in lock step (traverse.cuh, "sorted list + append
//     buffer") — no per-lane sift loops. k == 1 and k beyond shared memory keep the 4-ary heap.
//   * ring: groups travel with their heaps. A group's state is read from the local inbox (which the
//     previous GPU filled with plain st.global over NVLink), traversed against the local tree, and
//     written straight into the NEXT GPU's inbox, followed by fence + st.release.sys of a sequence
//     number; the consumer polls with ld.acquire.sys and acknowledges the slot the same way. No
//     host round-trip, no collective call between hops; transfer of group g overlaps traversal of
//     every other group in flight.
#include "traverse.cuh"

namespace b200knn {

namespace {

// SORTED = true      : the product path, sorted lists + append buffers in shared memory (2 <= k <= ~1700).
// SORTED = false     : 4-ary heaps; in shared memory for k == 1 (keeps the winner's id for
//                      find-closest-point), or with
// GLOBAL_HEAP = true : one [kp][32] heap block per resident warp in global memory (rows are coalesced
//                      128-byte lines) for k beyond the shared-memory limit.
template <bool GLOBAL_HEAP, bool SORTED>
__global__ void __launch_bounds__(Q_MAX_WARPS * 32, 1)
knn_ring_kernel_t(const RingParams p) {
  extern __shared__ __align__(16) float q_smem[];
  const int lane = lane_id();
  const int k = p.k;
  GroupCtx g;
  if constexpr (GLOBAL_HEAP) {
    const size_t warp_global = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    group_ctx_init(g, q_smem, threadIdx.x >> 5, lane, k, p.global_heaps + warp_global * (size_t)heap_padded(k) * 32);
  } else {
    group_ctx_init(g, q_smem, threadIdx.x >> 5, lane, k, nullptr, SORTED);
  }
  const int kp = g.kp;

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
        g.last_id = 0xffffffffu;
        g.live = 0;
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
        g.live = k;     // an arriving list may be full
      }
      if (p.use_flags) {
        __syncwarp();   // every lane's loads have returned (their values were stored to smem)
        if (lane == 0) st_release_sys(p.prev_ack[par] + grp, p.seq_base + (unsigned)s);
      }
    }
    g.valid = lane < cnt;

    // ------------------------------------------------------------------ traverse the local tree
    const long long clock_before = p.stats != nullptr ? clock64() : 0ll;
    if (cnt > 0 && p.tree.n > 0) {
      for (int i = k; i < kp; ++i) g.heap[i * 32] = -INFINITY;   // heap padding
      // seed = the level-1 node holding the group's own leaf (self queries at home), else the nearest (-2)
      const int own_seed = (hop == 0 && p.home_is_tree) ? (p.tree.levels >= 2 ? (int)(home_group / kFanout) : -1) : -2;
      traverse_group_knn<SORTED>(p.tree, g, cnt, own_seed, k, (hop == 0 && p.fill) ? k / kGroup : 0);
    }

    if (p.stats != nullptr) {
      // diagnostics (b200knn_index_enable_stats): counters are flushed per group, and counter 7 keeps the slowest
      // group as (kilo-cycles << 32) | group index, counters 8..13 what that group did
      const unsigned long long kc = (unsigned long long)((clock64() - clock_before) >> 10);
      const unsigned long long packed = (kc << 32) | (unsigned long long)(home_group & 0xffffffffll);
      if (lane == 0 && atomicMax(p.stats + 7, packed) < packed) {
        p.stats[8] = g.n_leaves; p.stats[9] = g.n_inner; p.stats[10] = g.n_flush;
        p.stats[11] = g.n_fallback; p.stats[12] = g.n_purge; p.stats[13] = g.n_inserts;
      }
      flush_stats(g, p.stats);
      g.n_leaves = g.n_inner = g.n_inserts = g.n_fallback = g.n_purge = g.n_flush = 0;
    }
    // ------------------------------------------------------------------ output
    if (hop == R - 1) {
      if (g.valid) {
        p.result_sorted[home_group * kGroup + lane] = p.raw_result ? g.top : finalize_kth(g.top);
        if (p.result_ids != nullptr) p.result_ids[home_group * kGroup + lane] = g.last_id;
      }
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
  flush_stats(g, p.stats);
}

}  // namespace

size_t group_smem_bytes(int k, int warps_per_cta, bool sorted) {
  return (size_t)warps_per_cta * warp_smem_floats(k, sorted) * sizeof(float);
}

// Shared memory decides the shape: maximise resident warps per SM over (CTAs per SM, warps per
// CTA); prefer more, smaller CTAs on ties. k = 100 -> two 5-warp CTAs per SM. Returns 0 if not even
// one warp fits.
int group_pick_warps(int k, const void* kernel, bool sorted) {
  int reg_warps = 32;     // resident warps the register file allows for this kernel
  cudaFuncAttributes fa;
  if (cudaFuncGetAttributes(&fa, kernel) == cudaSuccess && fa.numRegs > 0) {
    const int regs = (fa.numRegs + 7) / 8 * 8;
    reg_warps = 65536 / (32 * regs);
    if (reg_warps > 32) reg_warps = 32;
    if (reg_warps < 1) reg_warps = 1;
  }
  const size_t per_warp = group_smem_bytes(k, 1, sorted);
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

namespace {
template <bool SORTED>
int launch_shared(RingParams& p, unsigned long long total, int num_sms, cudaStream_t stream) {
  auto kernel = knn_ring_kernel_t<false, SORTED>;
  const int warps = group_pick_warps(p.k, (const void*)kernel, SORTED);
  KNN_REQUIRE(warps >= 1, "query kernel: k does not fit in shared memory");
  const size_t smem = group_smem_bytes(p.k, warps, SORTED);
  KNN_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  KNN_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, warps * 32, smem));
  KNN_REQUIRE(occ >= 1, "query kernel does not fit on an SM");
  const unsigned long long want = (total + warps - 1) / warps;
  const unsigned long long cap = (unsigned long long)num_sms * occ;
  const int grid = (int)(want < cap ? want : cap);
  KNN_CUDA_TRY(cudaMemsetAsync(p.task_counter, 0, sizeof(unsigned long long), stream));
  kernel<<<grid, warps * 32, smem, stream>>>(p);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}
}  // namespace

int ring_pick_warps_per_cta(int k) {
  if (use_sorted_list(k)) return group_pick_warps(k, (const void*)knn_ring_kernel_t<false, true>, true);
  if (k == 1) return group_pick_warps(k, (const void*)knn_ring_kernel_t<false, false>, false);
  return 0;
}

int ring_launch(const RingParams& p_in, int num_sms, cudaStream_t stream) {
  RingParams p = p_in;
  KNN_REQUIRE(p.k >= 1, "k must be >= 1");
  const unsigned long long total = (unsigned long long)(p.step_end - p.step_begin) * p.groups_per_chunk;
  if (total == 0) return 0;
  if (use_sorted_list(p.k)) return launch_shared<true>(p, total, num_sms, stream);
  if (p.k == 1) return launch_shared<false>(p, total, num_sms, stream);
  // k does not fit in shared memory: heaps go to a stream-ordered global scratch buffer
  KNN_REQUIRE(p.allow_global_heap, "k too large for shared memory on this path");
  const int warps = 8;
  const size_t smem = (size_t)warps * warp_smem_floats_noheap() * sizeof(float);
  int occ = 0;
  KNN_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, knn_ring_kernel_t<true, false>, warps * 32, smem));
  KNN_REQUIRE(occ >= 1, "query kernel does not fit on an SM");
  if (occ > 2) occ = 2;
  unsigned long long want = (total + warps - 1) / warps;
  unsigned long long cap = (unsigned long long)num_sms * occ;
  const int grid = (int)(want < cap ? want : cap);
  const size_t heap_bytes = (size_t)grid * warps * heap_padded(p.k) * 32 * sizeof(float);
  float* heaps = nullptr;
  KNN_CUDA_TRY(cudaMallocAsync(&heaps, heap_bytes, stream));
  p.global_heaps = heaps;
  KNN_CUDA_TRY(cudaMemsetAsync(p.task_counter, 0, sizeof(unsigned long long), stream));
  knn_ring_kernel_t<true, false><<<grid, warps * 32, smem, stream>>>(p);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  KNN_CUDA_TRY(cudaFreeAsync(heaps, stream));
  return 0;
}

}  // namespace b200knn
