// neighbors.cu — k nearest NEIGHBOURS (ids and distances), not only the k-th distance.
//
// Role of cukd::stackFree::knn with a candidate list the caller reads afterwards
// (cudaKDTree/cukd/knn.h:617-631 + FlexHeapCandidateList::get_pointID / get_dist2, :311-326): the
// reference's two programs only ever read entry 0 (the k-th distance), users of the library read the
// whole list.
//
// Design: two passes instead of a heap of (distance, id) pairs.
//   pass 1  the product's own kernel (query.cu) — heaps of bare floats in shared memory — yields each
//           query's exact k-th squared distance L (raw_result = 1);
//   pass 2  (this file) a fixed-radius traversal with no heap at all: every point with d2 < L is a
//           neighbour, points with d2 == L fill the row up to k. Lists come out unordered, like the
//           reference's heap-ordered lists; the Python wrapper sorts rows by distance.
// A (distance, id) heap would double the shared memory per query — half the resident warps on the
// kernel that is already limited by exactly that — and pay the id traffic on every sift; pass 2 costs
// one more traversal with a tight, known radius (about a third of pass 1's leaves) and no sifting.
#include "traverse.cuh"

namespace b200knn {
namespace {

// Per-lane output cursor: strictly closer points are appended from the front of the row, ties with the
// k-th distance from the back; a tie gives way when a strictly closer point needs its slot.
struct CollectLeafOp {
  static constexpr bool kBuffered = false;
  float limit;        // this lane's k-th squared distance (or the sentinel if fewer than k points are in range)
  bool inclusive;     // ties with `limit` are neighbours (false when limit is the sentinel)
  int k, front, ties;
  int32_t* ids;       // this lane's row
  float* dist;

  __device__ __forceinline__ void operator()(const float4& p, float lx, float ly, float lz, float hx,
                                             float hy, float hz, GroupCtx& g) {
    const float lb = point_box_lb(g.qx, g.qy, g.qz, lx, ly, lz, hx, hy, hz);
    if (!__any_sync(kFullMask, g.valid && (lb < g.top))) return;
    const int lane = lane_id();
    g.stage[lane] = p;
    g.soa[lane] = p.x;
    g.soa[32 + lane] = p.y;
    g.soa[64 + lane] = p.z;
    __syncwarp();
    unsigned mask = 0;
    const float top0 = g.top;      // limit, or the next float above it when ties count
    const unsigned long long qx2 = pack2(g.qx, g.qx), qy2 = pack2(g.qy, g.qy), qz2 = pack2(g.qz, g.qz);
    const unsigned long long* sx = reinterpret_cast<const unsigned long long*>(g.soa);
    const unsigned long long* sy = reinterpret_cast<const unsigned long long*>(g.soa + 32);
    const unsigned long long* sz = reinterpret_cast<const unsigned long long*>(g.soa + 64);
#pragma unroll
    for (int jj = 0; jj < kGroup / 2; ++jj) {
      const unsigned long long dx = sub2(qx2, sx[jj]), dy = sub2(qy2, sy[jj]), dz = sub2(qz2, sz[jj]);
      const unsigned long long d2 = fma2(dz, dz, fma2(dx, dx, mul2(dy, dy)));
      mask |= (__uint_as_float((unsigned)(d2 & 0xffffffffull)) < top0) ? (1u << (2 * jj)) : 0u;
      mask |= (__uint_as_float((unsigned)(d2 >> 32)) < top0) ? (2u << (2 * jj)) : 0u;
    }
    if (!g.valid) mask = 0;
    while (mask) {
      const int j = __ffs(mask) - 1;
      mask &= mask - 1;
      const float4 c = g.stage[j];
      const float d2 = dist2(g.qx, g.qy, g.qz, c.x, c.y, c.z);
      if (d2 < limit) {
        if (front + ties == k && ties > 0) --ties;        // a tie gives up its slot
        if (front + ties < k) {
          ids[front] = (int32_t)__float_as_uint(c.w);
          dist[front] = finalize_kth(d2);
          ++front;
        }
      } else if (inclusive && d2 == limit && front + ties < k) {
        ++ties;
        ids[k - ties] = (int32_t)__float_as_uint(c.w);
        dist[k - ties] = finalize_kth(d2);
      }
    }
    __syncwarp();
    ++g.n_leaves;
  }
};

__global__ void __launch_bounds__(Q_MAX_WARPS * 32, 1)
knn_collect_kernel(const CollectParams p) {
  extern __shared__ __align__(16) float q_smem[];
  const int lane = lane_id();
  GroupCtx g;
  // no heap: the shared-memory slice of the global-heap variant, with a heap pointer nobody follows
  group_ctx_init(g, q_smem, threadIdx.x >> 5, lane, 1, q_smem);
  g.heap = nullptr;
  const unsigned long long total = (unsigned long long)((p.nq + kGroup - 1) / kGroup);
  for (;;) {
    unsigned long long t = 0;
    if (lane == 0) t = atomicAdd(p.task_counter, 1ull);
    t = __shfl_sync(kFullMask, t, 0);
    if (t >= total) break;
    const int64_t group = (int64_t)t;
    const int64_t rem = p.nq - group * kGroup;
    const int cnt = rem < kGroup ? (int)rem : kGroup;
    const int src = lane < cnt ? lane : cnt - 1;
    const float4 q = __ldg(reinterpret_cast<const float4*>(p.queries) + group * kGroup + src);
    g.qx = q.x; g.qy = q.y; g.qz = q.z;
    g.valid = lane < cnt;
    CollectLeafOp op;
    op.limit = __ldg(p.limit_sorted + group * kGroup + src);
    op.inclusive = op.limit < p.sentinel;
    op.k = p.k;
    op.front = 0;
    op.ties = 0;
    const size_t row = (size_t)(__float_as_uint(q.w) - p.row_offset);
    op.ids = p.out_ids + row * (size_t)p.k;
    op.dist = p.out_dist + row * (size_t)p.k;
    // traversal radius: strict "<" everywhere, so ties need the next float above the limit
    g.top = op.inclusive ? __uint_as_float(__float_as_uint(op.limit) + 1u) : op.limit;
    if (p.tree.n > 0) {
      const int own_seed = p.home_is_tree ? (p.tree.levels >= 2 ? (int)(group / kFanout) : -1) : -2;
      traverse_group(p.tree, g, cnt, own_seed, 0, op);
    }
    if (g.valid) {
      for (int s = op.front; s < p.k - op.ties; ++s) {     // fewer than k points in range
        op.ids[s] = -1;
        op.dist[s] = INFINITY;
      }
    }
    __syncwarp();
  }
}

}  // namespace

int collect_launch(const CollectParams& p, int num_sms, cudaStream_t stream) {
  KNN_REQUIRE(p.k >= 1, "k must be >= 1");
  if (p.nq == 0) return 0;
  const int warps = 8;
  const size_t smem = (size_t)warps * warp_smem_floats_noheap() * sizeof(float);
  int occ = 0;
  KNN_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, knn_collect_kernel, warps * 32, smem));
  KNN_REQUIRE(occ >= 1, "collect kernel does not fit on an SM");
  const unsigned long long groups = (unsigned long long)div_up(p.nq, (int64_t)kGroup);
  const unsigned long long want = (groups + warps - 1) / warps, cap = (unsigned long long)num_sms * occ;
  KNN_CUDA_TRY(cudaMemsetAsync(p.task_counter, 0, sizeof(unsigned long long), stream));
  knn_collect_kernel<<<(int)(want < cap ? want : cap), warps * 32, smem, stream>>>(p);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

}  // namespace b200knn
