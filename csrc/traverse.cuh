// traverse.cuh — device-side group traversal shared by the query/ring kernel (query.cu) and the
// culled-exchange kernels (pv.cu): per-warp context, padded 4-ary heap, leaf sweep, seed phase,
// best-first frontier with depth-first overflow. See the header comment of query.cu for the design.
#pragma once
#include "common.cuh"
#include "kernels.h"

namespace b200knn {
namespace {

constexpr int Q_MAX_WARPS = 16;        // per CTA
constexpr int kQueueCap = 128;         // frontier entries per warp (1 KB)
constexpr unsigned kNodeIdxBits = 28;  // node = level << 28 | index

__host__ __device__ constexpr int heap_padded(int k) { return 4 * ((k - 1 + 3) / 4) + 1; }
// Sorted candidate list (the product path, see "sorted list + append buffer" below): k rows, one -inf
// terminator row, then the append buffer.
constexpr int kFlushRows = 32;         // buffered candidates merged per flush (one register sort)
constexpr int kBufRows = 2 * kFlushRows;   // a leaf (32 candidates) always fits behind a not-yet-flushed batch
__host__ __device__ constexpr int list_rows(int k) { return k + 1; }
// frontier | frames | soa | worklist, preceded by the float4 stage of the heap kernels (the sorted path has no use for it)
__host__ __device__ constexpr size_t warp_aux_floats(bool with_stage = true) {
  return (with_stage ? 128 : 0) + 2 * kQueueCap + 32 + 96 + 8;
}
__host__ __device__ constexpr size_t warp_smem_floats(int k, bool sorted) {
  return sorted ? (size_t)(list_rows(k) + kBufRows + 1) * 32 + warp_aux_floats(false)
                : (size_t)heap_padded(k) * 32 + warp_aux_floats(true);
}

struct GroupCtx {
  float qx, qy, qz;     // this lane's query
  float top;            // this lane's current k-th squared distance (heap root)
  float glx, gly, glz, ghx, ghy, ghz;   // bounding box of the group's valid queries (warp-uniform)
  float maxR2;          // max over valid lanes of top (warp-uniform)
  float* heap;          // shared: heap[i*32] is entry i of this lane
  float4* stage;        // shared: 32 leaf points (x,y,z,id)
  float* soa;           // shared: the same points as x[32] | y[32] | z[32]
  unsigned* qkey;       // shared: frontier lower bounds (float bits; non-negative => uint-ordered)
  unsigned* qnode;      // shared: frontier nodes
  unsigned* frames;     // shared: frame stack of the depth-first fallback
  unsigned* work;       // shared: lane-mask worklist of traverse_group (8 entries)
  float* buf;           // shared (sorted mode): buf[i*32] is buffered candidate i of this lane
  int cnt;              // sorted mode: candidates in this lane's buffer
  int live;             // sorted mode: upper bound of the list entries that are not the sentinel any more
  bool flush_pending;   // sorted mode, warp-uniform: some lane's buffer cannot take another leaf
  float cull_scale;     // 1 + eps of an approximate search (1 for the exact one), see TreeView
  int kp;               // rows of the candidate structure: padded heap size, or k + 1 in sorted mode
  bool valid;
  unsigned last_id;     // id of the most recently accepted candidate (find-closest-point uses it, k == 1)
  unsigned n_leaves, n_inner, n_inserts, n_fallback, n_purge, n_flush;
};

// Sift e down from position i. Children of i: 4i+1..4i+4, all present (padding is -inf).
__device__ __forceinline__ void heap_sift_down(float* h, int kp, int i, float e) {
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
}

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

// ---- packed fp32x2 arithmetic (Blackwell FADD2/FMUL2/FFMA2) ----
__device__ __forceinline__ unsigned long long pack2(float lo, float hi) {
  return (unsigned long long)__float_as_uint(lo) | ((unsigned long long)__float_as_uint(hi) << 32);
}
#if defined(B200KNN_SIMT_HOST)   // host emulation of the kernels (tests/simt): the same IEEE operations per half
__device__ __forceinline__ float half_lo(unsigned long long v) { return __uint_as_float((unsigned)(v & 0xffffffffull)); }
__device__ __forceinline__ float half_hi(unsigned long long v) { return __uint_as_float((unsigned)(v >> 32)); }
__device__ __forceinline__ unsigned long long sub2(unsigned long long a, unsigned long long b) {
  return pack2(half_lo(a) - half_lo(b), half_hi(a) - half_hi(b));
}
__device__ __forceinline__ unsigned long long mul2(unsigned long long a, unsigned long long b) {
  return pack2(half_lo(a) * half_lo(b), half_hi(a) * half_hi(b));
}
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
  return pack2(fmaf(half_lo(a), half_lo(b), half_lo(c)), fmaf(half_hi(a), half_hi(b), half_hi(c)));
}
#else
__device__ __forceinline__ unsigned long long sub2(unsigned long long a, unsigned long long b) {
  unsigned long long r;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ unsigned long long mul2(unsigned long long a, unsigned long long b) {
  unsigned long long r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
#endif

// This lane's point of a leaf (+inf padding beyond the leaf's count never beats any threshold). The
// traversal issues this load as soon as it knows the leaf, next to the load of the leaf's box, so that
// the two global-memory latencies overlap instead of following each other.
__device__ __forceinline__ float4 load_leaf_point(const TreeView& t, int leaf, int cnt) {
  float4 p = make_float4(INFINITY, INFINITY, INFINITY, 0.f);
  if (lane_id() < cnt) p = __ldg(reinterpret_cast<const float4*>(t.pts) + (size_t)leaf * kGroup + lane_id());
  return p;
}
__device__ __forceinline__ int leaf_point_count(const TreeView& t, int leaf) {
  const int rem = t.n - leaf * kGroup;
  return rem < kGroup ? rem : kGroup;
}

__device__ __forceinline__ void process_leaf(const float4& p, GroupCtx& g) {
  const int lane = lane_id();
  g.stage[lane] = p;
  g.soa[lane] = p.x;          // x[32] | y[32] | z[32]: pairs of candidates load as one 64-bit word
  g.soa[32 + lane] = p.y;
  g.soa[64 + lane] = p.z;
  __syncwarp();

  // phase 1: convergent distance sweep, two candidates per instruction with the packed fp32x2
  // pipe (FADD2 / FMUL2 / FFMA2 on sm_100a): each half is an independent IEEE operation, so the
  // result is bit-identical to dist2().
  unsigned mask = 0;
  const float top0 = g.top;
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

  // phase 2: only lanes with survivors loop, only over their survivors
  while (mask) {
    const int j = __ffs(mask) - 1;
    mask &= mask - 1;
    const float4 c = g.stage[j];
    const float d2 = dist2(g.qx, g.qy, g.qz, c.x, c.y, c.z);
    if (d2 < g.top) {
      g.top = heap_replace_padded(g.heap, g.kp, d2);
      g.last_id = __float_as_uint(c.w);     // for k == 1 this is the id of the closest point so far
      ++g.n_inserts;
    }
  }
  __syncwarp();
  g.maxR2 = __uint_as_float(__reduce_max_sync(kFullMask, g.valid ? __float_as_uint(g.top) : 0u));
  ++g.n_leaves;
}

// ---- convergent fill ------------------------------------------------------------------------------
// A fresh heap (all entries = sentinel) accepts its first k candidates whatever they are, and every
// one of them would pay a full-depth sift (the new value is smaller than the sentinels below it).
// Instead the first floor(k/32) leaves of a fresh group are appended as they come — 32 plain stores
// per lane, no test, no divergence — and one Floyd heapify turns the array into the heap. Values are
// clamped to the sentinel, so a max-radius cut-off and +inf padding points behave as "no candidate".
// The result is the same set of k smallest distances, hence bit-identical.
__device__ __forceinline__ void fill_leaf(const float4& p, GroupCtx& g, int pos0) {
  const int lane = lane_id();
  g.soa[lane] = p.x;
  g.soa[32 + lane] = p.y;
  g.soa[64 + lane] = p.z;
  __syncwarp();
  if (g.valid) {
    const float sent = g.top;          // nothing has been sifted yet: top is still the sentinel
    const unsigned long long qx2 = pack2(g.qx, g.qx), qy2 = pack2(g.qy, g.qy), qz2 = pack2(g.qz, g.qz);
    const unsigned long long* sx = reinterpret_cast<const unsigned long long*>(g.soa);
    const unsigned long long* sy = reinterpret_cast<const unsigned long long*>(g.soa + 32);
    const unsigned long long* sz = reinterpret_cast<const unsigned long long*>(g.soa + 64);
    float* h = g.heap + (size_t)pos0 * 32;
#pragma unroll
    for (int jj = 0; jj < kGroup / 2; ++jj) {
      const unsigned long long dx = sub2(qx2, sx[jj]), dy = sub2(qy2, sy[jj]), dz = sub2(qz2, sz[jj]);
      const unsigned long long d2 = fma2(dz, dz, fma2(dx, dx, mul2(dy, dy)));
      h[(2 * jj) * 32] = fminf(__uint_as_float((unsigned)(d2 & 0xffffffffull)), sent);
      h[(2 * jj + 1) * 32] = fminf(__uint_as_float((unsigned)(d2 >> 32)), sent);
    }
    g.n_inserts += kGroup;
  }
  __syncwarp();
  ++g.n_leaves;
}

__device__ __forceinline__ void fill_finish(GroupCtx& g) {
  if (g.valid) {
    for (int start = (g.kp - 2) / 4; start >= 0; --start) heap_sift_down(g.heap, g.kp, start, g.heap[start * 32]);
    g.top = g.heap[0];
  }
  __syncwarp();
  g.maxR2 = __uint_as_float(__reduce_max_sync(kFullMask, g.valid ? __float_as_uint(g.top) : 0u));
}

// ---- sorted list + append buffer (the product path, k >= 2 in shared memory) ---------------------------
// Role of the reference's candidate lists (cukd/knn.h:214-258 FixedCandidateList / HeapCandidateList,
// :328-389 FlexHeapCandidateList::push): keep the k smallest distances seen so far and hand back the
// current k-th as the cull radius. The reference — and round 1 of this kernel — insert one candidate at
// a time by sifting through a heap; on a SIMT machine that loop runs at 13 of 32 lanes (lanes differ in
// how many survivors they have and how deep each sift goes) and was 55 % of all issued instructions.
// Here nothing is data-dependent in control flow:
//   * a lane's candidates live as a DESCENDING sorted column list[0..k) (list[0] = current k-th distance,
//     list[k] = -inf terminator), one column per lane, bank == lane;
//   * a leaf is 32 distances per lane, each compared with the lane's k-th distance as of the last flush
//     and, if smaller, appended to the lane's buffer column by a predicated store (append_leaf): no
//     branch, no second pass over survivors. A stale (too large) threshold only lets extra candidates
//     in; they fall out again at the merge;
//   * when some lane holds more than 32 buffered values (a further leaf might not fit), the whole warp
//     flushes (sorted_flush): every lane sorts its first 32 buffered values in REGISTERS with a fixed
//     191-comparator network (missing entries are +inf), writes them back, and merges them into its
//     list in place — skip the 32 largest of (list, batch), then k output steps of "store the larger
//     head, advance that side". It is the same instruction stream for all lanes; only addresses and
//     predicates differ. A lane with more than 32 buffered values flushes its last 32.
// The set of the k smallest distances is the same as with any other insertion order, so the k-th
// distance is bit-identical to the heap's.
//
// With 10 resident warps per SM (the lists pin 23 KB of shared memory per warp) the kernel lives on
// instruction-level parallelism inside each warp, so the serial parts are written for latency:
//   * a leaf's 32 distances are all computed before the first append (a store to a computed address
//     would otherwise fence the next pair's loads);
//   * the 32 largest of (list, batch) are found by a 6-step merge-path bisection, not 32 merge steps;
//   * the merge keeps the next four entries of either side in registers, so the shared-memory load
//     issued at step s is not needed before step s + 3: the loop-carried chain is compare -> move.
// The two inner loops are predicated PTX: left to the compiler, "advance whichever side held the larger
// head" became 14 instructions per step with address re-derivation on both arms.
struct MergeHeads {
  float a0, a1, a2, a3, b0, b1;    // a0/b0 = current heads, the others are prefetched successors
};
#if defined(B200KNN_SIMT_HOST)   // host emulation of the kernels (tests/simt): same steps in plain C++
typedef float* smem_col_t;
__device__ __forceinline__ smem_col_t smem_col(float* p) { return p; }
__device__ __forceinline__ void append_if_less(smem_col_t& p, float v, float thr) {
  if (v < thr) { *p = v; p += 32; }
}
// pa / pb: rows of a3 / b1. Returns the larger head and advances the side it came from (ties: b side).
__device__ __forceinline__ float merge_step(smem_col_t& pa, smem_col_t& pb, MergeHeads& h) {
  const float o = fmaxf(h.a0, h.b0);
  if (h.a0 > h.b0) { h.a0 = h.a1; h.a1 = h.a2; h.a2 = h.a3; pa += 32; h.a3 = *pa; }
  else             { h.b0 = h.b1; pb += 32; h.b1 = *pb; }
  return o;
}
__device__ __forceinline__ int rows_between(smem_col_t from, smem_col_t to) { return (int)((to - from) >> 5); }
#else
typedef unsigned smem_col_t;     // 32-bit shared-window address of one lane's column; rows are 128 bytes apart
__device__ __forceinline__ smem_col_t smem_col(float* p) { return (unsigned)__cvta_generic_to_shared(p); }
// if (v < thr) { *p = v; next row }. Readers of the buffer sit behind a __syncwarp().
__device__ __forceinline__ void append_if_less(smem_col_t& p, float v, float thr) {
  asm volatile("{\n\t.reg .pred q;\n\tsetp.lt.f32 q, %1, %2;\n\t@q st.shared.f32 [%0], %1;\n\t@q add.u32 %0, %0, 128;\n\t}"
               : "+r"(p) : "f"(v), "f"(thr));
}
__device__ __forceinline__ float merge_step(smem_col_t& pa, smem_col_t& pb, MergeHeads& h) {
  float o;
  asm volatile("{\n\t.reg .pred q;\n\tsetp.gt.f32 q, %1, %5;\n\tmax.f32 %0, %1, %5;\n\t"
               "@q mov.f32 %1, %2;\n\t@q mov.f32 %2, %3;\n\t@q mov.f32 %3, %4;\n\t"
               "@q ld.shared.f32 %4, [%7+128];\n\t@q add.u32 %7, %7, 128;\n\t"
               "@!q mov.f32 %5, %6;\n\t"
               "@!q ld.shared.f32 %6, [%8+128];\n\t@!q add.u32 %8, %8, 128;\n\t}"
               : "=f"(o), "+f"(h.a0), "+f"(h.a1), "+f"(h.a2), "+f"(h.a3), "+f"(h.b0), "+f"(h.b1), "+r"(pa), "+r"(pb));
  return o;
}
__device__ __forceinline__ int rows_between(smem_col_t from, smem_col_t to) { return (int)((to - from) >> 7); }
#endif

__device__ __forceinline__ void sorted_flush(GroupCtx& g, int k) {
  // This flush's batch = the lane's LAST (up to) 32 buffered rows [base, n); rows [0, base) simply stay
  // where they are for the next flush. Row base + 32 (free: it is row n, or row 32 of a lane with
  // n <= 32) takes the batch's -inf terminator.
  const int n = g.cnt;
  const int base = n > kFlushRows ? n - kFlushRows : 0;
  const int m = n - base;
  float* bb = g.buf + base * 32;
  float r[kFlushRows];
#pragma unroll
  for (int i = 0; i < kFlushRows; ++i) r[i] = i < m ? bb[i * 32] : INFINITY;
#define KNN_CE(a, b) { const float hi_ = fmaxf(r[a], r[b]); r[b] = fminf(r[a], r[b]); r[a] = hi_; }
#include "sort32_network.inc"
#undef KNN_CE
#pragma unroll
  for (int i = 0; i < kFlushRows; ++i) bb[i * 32] = r[i];
  bb[kFlushRows * 32] = -INFINITY;

  // While the lists are still filling up, their first rows are sentinels and stay sentinels: with at
  // most `live` real entries in any lane's list, output rows [0, t0) need not be touched.
  const int live_max = (int)__reduce_max_sync(kFullMask, (unsigned)g.live);
  const int t0 = k - kFlushRows - live_max > 0 ? k - kFlushRows - live_max : 0;
  const int diag = kFlushRows + t0;
  // The union has k + 32 entries (+inf fillers included); the `diag` largest are left alone: jb of them
  // from the batch, diag - jb from the list, jb = the largest j with batch[j-1] >= list[diag-j] (merge
  // path; ties take the batch side first, so fillers go before sentinels). Rows past the list's
  // terminator read as the terminator.
  int jb = 0;
  {
    int hi = kFlushRows;
#pragma unroll
    for (int it = 0; it < 6; ++it) {
      const int mid = (jb + hi + 1) >> 1;          // jb < mid <= hi whenever jb < hi
      const int ai = diag - mid < k ? diag - mid : k;
      const bool ok = mid > jb && bb[(mid - 1) * 32] >= g.heap[ai * 32];
      if (mid > jb) { if (ok) jb = mid; else hi = mid - 1; }
    }
  }
  const int ja = diag - jb;
  MergeHeads h;
  {
    const float* la = g.heap + ja * 32;
    const float* lb = bb + jb * 32;
    h.a0 = la[0]; h.a1 = la[32]; h.a2 = la[64]; h.a3 = la[96];
    h.b0 = lb[0]; h.b1 = lb[32];
  }
  __syncwarp();                             // the C++ stores above are in place before merge_step's loads
  smem_col_t pa = smem_col(g.heap + (ja + 3) * 32), pb = smem_col(bb + (jb + 1) * 32);
  // Output position t never passes the list head (the heads start 32 entries ahead in total and only
  // batch entries let t catch up), reads run ahead of writes, and where they meet the value is the same:
  // the merge is in place.
#pragma unroll 4
  for (int t = t0; t < k; ++t) g.heap[t * 32] = merge_step(pa, pb, h);
  __syncwarp();

  g.live = g.live + m < k ? g.live + m : k;
  g.cnt = base;
  g.top = g.heap[0];
  g.maxR2 = __uint_as_float(__reduce_max_sync(kFullMask, g.valid ? __float_as_uint(g.top) : 0u));
  g.flush_pending = false;                  // base <= 32 on every lane
  ++g.n_flush;
}

// One leaf against the lanes' thresholds: survivors are appended to the lanes' buffers (see above).
// Precondition: every lane has at most 32 buffered values (the traversal flushes before the next leaf).
__device__ __forceinline__ void append_leaf(const float4& p, GroupCtx& g) {
  const int lane = lane_id();
  g.soa[lane] = p.x;
  g.soa[32 + lane] = p.y;
  g.soa[64 + lane] = p.z;
  __syncwarp();
  const float top0 = g.valid ? g.top : -INFINITY;
  const unsigned long long qx2 = pack2(g.qx, g.qx), qy2 = pack2(g.qy, g.qy), qz2 = pack2(g.qz, g.qz);
  const unsigned long long* sx = reinterpret_cast<const unsigned long long*>(g.soa);
  const unsigned long long* sy = reinterpret_cast<const unsigned long long*>(g.soa + 32);
  const unsigned long long* sz = reinterpret_cast<const unsigned long long*>(g.soa + 64);
  unsigned long long d2[kGroup / 2];
#pragma unroll
  for (int jj = 0; jj < kGroup / 2; ++jj) {
    const unsigned long long dx = sub2(qx2, sx[jj]), dy = sub2(qy2, sy[jj]), dz = sub2(qz2, sz[jj]);
    d2[jj] = fma2(dz, dz, fma2(dx, dx, mul2(dy, dy)));
  }
  const smem_col_t first = smem_col(g.buf + g.cnt * 32);
  smem_col_t pos = first;
#pragma unroll
  for (int jj = 0; jj < kGroup / 2; ++jj) {
    append_if_less(pos, __uint_as_float((unsigned)(d2[jj] & 0xffffffffull)), top0);
    append_if_less(pos, __uint_as_float((unsigned)(d2[jj] >> 32)), top0);
  }
  const int added = rows_between(first, pos);
  g.n_inserts += (unsigned)added;
  g.cnt += added;
  __syncwarp();
  ++g.n_leaves;
  g.flush_pending = __any_sync(kFullMask, g.cnt > kFlushRows);
}

// Exact per-query test against the leaf box; skip the leaf if no lane can improve.
__device__ __forceinline__ void maybe_leaf(const float4& p, float lx, float ly, float lz,
                                           float hx, float hy, float hz, GroupCtx& g) {
  const float lb = g.cull_scale * point_box_lb(g.qx, g.qy, g.qz, lx, ly, lz, hx, hy, hz);
  if (__any_sync(kFullMask, g.valid && (lb < g.top))) process_leaf(p, g);
}

// What traverse() does with a leaf that survives the group-level test. The default keeps the k
// smallest distances in the lanes' heaps; other kernels (neighbors.cu) plug in their own.
struct HeapLeafOp {
  static constexpr bool kBuffered = false;
  __device__ __forceinline__ void flush(GroupCtx&) const {}
  __device__ __forceinline__ void operator()(const float4& p, float lx, float ly, float lz, float hx,
                                             float hy, float hz, GroupCtx& g) const {
    maybe_leaf(p, lx, ly, lz, hx, hy, hz, g);
  }
};

// The sorted-list variant: same per-query test, survivors go to the lanes' buffers; the traversal calls
// flush() before the next leaf whenever g.flush_pending, and before it concludes that nothing is left.
struct SortedLeafOp {
  static constexpr bool kBuffered = true;
  int k;
  __device__ __forceinline__ void flush(GroupCtx& g) const { sorted_flush(g, k); }
  __device__ __forceinline__ void operator()(const float4& p, float lx, float ly, float lz, float hx,
                                             float hy, float hz, GroupCtx& g) const {
    const float lb = g.cull_scale * point_box_lb(g.qx, g.qy, g.qz, lx, ly, lz, hx, hy, hz);
    if (__any_sync(kFullMask, g.valid && (lb < g.top))) append_leaf(p, g);
  }
};

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
// The group-level test (group box against node box, largest radius of any lane) is what makes one warp's
// walk serve 32 queries; it is also loose exactly when the group is not one compact thing: a group of two
// clumps (a leaf at a joint of the Morton curve, or between two clusters) has a box that spans the space in
// between, and everything in that space passes. Splitting such groups up front (traverse_group) only
// catches the blatant cases — measured on clustered data: one group with a gap of 0.63 of its extent, just
// under the splitting rule, kept a warp busy for 460 ms while the rest of the GPU idled. So whenever the
// group's box is wide compared with its radius, children that pass the group-level test are confirmed by
// the exact one: is the node's box inside the ball of at least one lane? (lane = query, one child at a
// time). A compact group never pays for this; a loose one then costs what its lanes need, not what its
// bounding box covers.
__device__ __forceinline__ bool group_is_loose(const GroupCtx& g) {
  const float e = fmaxf(g.ghx - g.glx, fmaxf(g.ghy - g.gly, g.ghz - g.glz));
  return e * e > 4.f * g.maxR2;          // wider than two radii (a compact leaf-sized group is about one to two
                                         // radii wide for k >= 32); never while the radius is still +inf
}
// Not inlined: it is the rare path, and three inlined copies of it made the kernel's hot loop a quarter longer
// (instruction cache). All arguments travel in registers.
#if defined(B200KNN_SIMT_HOST)
#define KNN_NOINLINE __attribute__((noinline))
#else
#define KNN_NOINLINE __noinline__
#endif
__device__ KNN_NOINLINE unsigned confirm_children_call(float qx, float qy, float qz, float top, bool valid, unsigned pm,
                                                       float ax, float ay, float az, float bx, float by, float bz) {
  unsigned keep = 0u;
  while (pm != 0u) {
    const int bi = __ffs(pm) - 1;
    pm &= pm - 1;
    const float lb = point_box_lb(qx, qy, qz, __shfl_sync(kFullMask, ax, bi), __shfl_sync(kFullMask, ay, bi),
                                  __shfl_sync(kFullMask, az, bi), __shfl_sync(kFullMask, bx, bi),
                                  __shfl_sync(kFullMask, by, bi), __shfl_sync(kFullMask, bz, bi));
    if (__any_sync(kFullMask, valid && (lb < top))) keep |= 1u << bi;
  }
  return keep;
}
__device__ __forceinline__ unsigned confirm_children(const GroupCtx& g, unsigned pm, const float4& a, const float4& b) {
  return confirm_children_call(g.qx, g.qy, g.qz, g.top, g.valid, pm, a.x, a.y, a.z, b.x, b.y, b.z);
}

// This lane's child box of the node whose children are [base, base+count) on `level` (zeros beyond count).
__device__ __forceinline__ void load_child_box(const TreeView& t, int level, int base, int count, float4& a, float4& b) {
  a = make_float4(0.f, 0.f, 0.f, 0.f);
  b = a;
  if (lane_id() < count) {
    const float4* bp = reinterpret_cast<const float4*>(t.boxes[level]) + 2 * (size_t)(base + lane_id());
    a = __ldg(bp);
    b = __ldg(bp + 1);
  }
}

// (a, b) = this lane's child box, loaded by the caller (load_child_box) — early, so that the latency is
// behind other work by the time the keys are needed.
__device__ __forceinline__ void frontier_expand(int level, int base, int count, const float4& a, const float4& b,
                                                int seed_l1, int qcap, float cull_scale, int& qcount, int& depth,
                                                GroupCtx& g) {
  const int lane = lane_id();
  const bool have = lane < count;
  // keys are stored already multiplied by (1 + eps): every later comparison with the radius inherits it
  const float key = cull_scale * box_box_lb(g.glx, g.gly, g.glz, g.ghx, g.ghy, g.ghz, a.x, a.y, a.z, b.x, b.y, b.z);
  bool pass = have && (key < g.maxR2) && !(level == 1 && base + lane == seed_l1);
  unsigned mask = __ballot_sync(kFullMask, pass);
  if (mask != 0u && group_is_loose(g)) {
    mask = confirm_children(g, mask, a, b);
    pass = (mask >> lane) & 1u;
  }
  const int np = __popc(mask);
  if (np == 0) return;
  const unsigned lt = (1u << lane) - 1u;
  if (qcount + np > qcap) {
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
  if (qcount + np > qcap) {
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

// fill_leaves > 0: the valid lanes' heaps are fresh; the first fill_leaves leaves that come up are
// appended without a test (see fill_leaf). fill_leaves * 32 <= k.
//
// Structure of the walk:
//   * the FRONTIER (shared memory, best-first) holds inner nodes only;
//   * when a level-1 node is opened, its 32 leaf boxes stay in the lanes' registers as a LEAF RUN: the
//     leaves inside the radius are swept nearest first, one per loop iteration, "pop" = one redux.sync
//     over the lanes' keys; a run ends when its nearest remaining leaf is outside the current radius.
//     Leaves of different level-1 nodes are therefore not interleaved by distance; each leaf is still
//     checked against the radius of the moment, and any visiting order is exact;
//   * the walk starts with a run over the seed level-1 node (the group's own, or the nearest), and the
//     root is expanded only when that run is done (and flushed), so the radius is tight before any
//     inner node is queued.
// With ~10 resident warps per SM nothing else hides a dependent global-memory access, so the loop is
// software-pipelined by one item: the next leaf of a run is chosen, and this lane's point of it
// requested, before the current leaf is swept; the next frontier entry is popped, and this lane's child
// box of it requested, before the current inner node is expanded. Popping early is harmless: between
// two flushes the radius does not change anyway and an item is re-checked when its turn comes.
// Buffered leaf ops (LeafOp::kBuffered) are flushed at ONE place, the top of the loop: when a buffer is
// about to overflow, and when the walk would otherwise stop — the flush tightens the radius, after
// which the stop test is simply evaluated again. Frontier overflow falls back to the depth-first frame
// stack (rare; not pipelined).
constexpr unsigned kNoNode = 0xffffffffu;

template <class LeafOp>
__device__ __forceinline__ void traverse(const TreeView& t, GroupCtx& g, int seed_l1, int fill_leaves, LeafOp& leaf_op) {
  const int lane = lane_id();
  int fill_pos = 0;
  unsigned* fr = g.frames;
  const int top_level = t.levels - 1;
  int depth = -1, qcount = 0;
  bool started = false;
  // the inner node selected one iteration ago: node, its frontier key, this lane's child box
  unsigned cur_node = kNoNode, cur_key = 0u;
  float4 cur_a = make_float4(0.f, 0.f, 0.f, 0.f), cur_b = cur_a;
  // the leaf run: this lane's leaf box and key (0xffffffff = not / no longer a candidate), the leaf
  // whose turn is next and this lane's point of it
  float4 run_a = cur_a, run_b = cur_a, run_pt = cur_a;
  unsigned run_key = 0xffffffffu;
  int run_base = 0, run_who = -1;              // run_who < 0: no run

  g.cull_scale = t.cull_scale;
  int budget = t.node_budget;                  // only consulted by unbuffered leaf ops (find-closest-point)
  auto run_select = [&]() {                    // nearest remaining leaf of the run; request its points
    const unsigned m = __reduce_min_sync(kFullMask, run_key);
    run_who = m == 0xffffffffu ? -1 : __ffs(__ballot_sync(kFullMask, run_key == m)) - 1;
    if (run_who >= 0) run_pt = load_leaf_point(t, run_base + run_who, leaf_point_count(t, run_base + run_who));
  };
  auto run_start = [&](int base, int count, const float4& a, const float4& b) {
    run_a = a; run_b = b; run_base = base;
    const float key = t.cull_scale * box_box_lb(g.glx, g.gly, g.glz, g.ghx, g.ghy, g.ghz, a.x, a.y, a.z, b.x, b.y, b.z);
    bool pass = lane < count && key < g.maxR2;
    if (group_is_loose(g)) {
      const unsigned pm = __ballot_sync(kFullMask, pass);
      if (pm != 0u) pass = (confirm_children(g, pm, a, b) >> lane) & 1u;
    }
    run_key = pass ? __float_as_uint(key) : 0xffffffffu;
    run_select();
  };

  if (seed_l1 >= 0) {
    const int base = seed_l1 * kFanout, rem = t.count[0] - base;
    const int count = rem < kFanout ? rem : kFanout;
    float4 a, b;
    load_child_box(t, 0, base, count, a, b);
    run_start(base, count, a, b);
  }
  for (;;) {
    if constexpr (LeafOp::kBuffered) {
      if (g.flush_pending) leaf_op.flush(g);
    }
    if (depth >= 0) {
      // ---- frame step (overflow fallback): depth-first, nearest child first, boxes re-loaded per visit
      const int L = (int)fr[depth * 4 + 0], B = (int)fr[depth * 4 + 1], Cn = (int)fr[depth * 4 + 2];
      const unsigned todo_in = fr[depth * 4 + 3];
      float4 a, b;
      load_child_box(t, L, B, Cn, a, b);
      const float key = t.cull_scale * box_box_lb(g.glx, g.gly, g.glz, g.ghx, g.ghy, g.ghz, a.x, a.y, a.z, b.x, b.y, b.z);
      const bool pass = lane < Cn && ((todo_in >> lane) & 1u) && (key < g.maxR2) &&
                        !(L == 1 && B + lane == seed_l1);
      const unsigned pm = __ballot_sync(kFullMask, pass);
      if (pm == 0u) { --depth; continue; }
      const unsigned kb = pass ? __float_as_uint(key) : 0xffffffffu;
      const unsigned m = __reduce_min_sync(kFullMask, kb);
      const int who = __ffs(__ballot_sync(kFullMask, kb == m)) - 1;
      const int child = B + who;
      __syncwarp();
      if (lane == 0) fr[depth * 4 + 3] = pm & ~(1u << who);
      // a loose group confirms the chosen child with the exact test (one child per visit, not all of them)
      if (group_is_loose(g) && confirm_children(g, 1u << who, a, b) == 0u) { __syncwarp(); continue; }
      if (L != 0) {
        ++g.n_inner;
        ++depth;
        const int rem = t.count[L - 1] - child * kFanout;
        if (lane == 0) push_frame(fr, depth, L - 1, child * kFanout, rem < kFanout ? rem : kFanout);
        __syncwarp();
        continue;
      }
      const float4 pt = load_leaf_point(t, child, leaf_point_count(t, child));
      __syncwarp();
      const float lx = __shfl_sync(kFullMask, a.x, who), ly = __shfl_sync(kFullMask, a.y, who);
      const float lz = __shfl_sync(kFullMask, a.z, who), hx = __shfl_sync(kFullMask, b.x, who);
      const float hy = __shfl_sync(kFullMask, b.y, who), hz = __shfl_sync(kFullMask, b.z, who);
      if (fill_leaves > 0) {
        fill_leaf(pt, g, fill_pos);
        fill_pos += kGroup;
        if (--fill_leaves == 0) fill_finish(g);
      } else {
        leaf_op(pt, lx, ly, lz, hx, hy, hz, g);
      }
      continue;
    }

    if (run_who >= 0) {
      // ---- leaf run step: take the selected leaf, select + request the next one, then sweep
      const int who = run_who;
      const float4 pt = run_pt;
      const unsigned kw = __shfl_sync(kFullMask, run_key, who);
      const float lx = __shfl_sync(kFullMask, run_a.x, who), ly = __shfl_sync(kFullMask, run_a.y, who);
      const float lz = __shfl_sync(kFullMask, run_a.z, who), hx = __shfl_sync(kFullMask, run_b.x, who);
      const float hy = __shfl_sync(kFullMask, run_b.y, who), hz = __shfl_sync(kFullMask, run_b.z, who);
      if (!(kw < __float_as_uint(g.maxR2))) { run_who = -1; continue; }   // nearest first: the rest is farther still
      if constexpr (!LeafOp::kBuffered) {
        if (started && t.node_budget >= 0 && --budget < 0) break;          // approximate search: inspection budget spent
      }
      if (lane == who) run_key = 0xffffffffu;
      run_select();
      if (fill_leaves > 0) {
        fill_leaf(pt, g, fill_pos);
        fill_pos += kGroup;
        if (--fill_leaves == 0) fill_finish(g);
      } else {
        leaf_op(pt, lx, ly, lz, hx, hy, hz, g);
      }
      continue;
    }

    // ---- select: pop the nearest frontier entry inside the radius and request this lane's child box
    unsigned nx_node = kNoNode, nx_key = 0u;
    float4 nx_a = make_float4(0.f, 0.f, 0.f, 0.f), nx_b = nx_a;
    if (qcount > 0) {
      unsigned best = 0xffffffffu;
      int bslot = 0;
      for (int s = lane; s < qcount; s += 32) {
        const unsigned kx = g.qkey[s];
        if (kx < best) { best = kx; bslot = s; }
      }
      const unsigned m = __reduce_min_sync(kFullMask, best);
      if (m < __float_as_uint(g.maxR2)) {
        const int wl = __ffs(__ballot_sync(kFullMask, best == m)) - 1;
        const int slot = __shfl_sync(kFullMask, bslot, wl);
        nx_node = g.qnode[slot];
        nx_key = m;
        __syncwarp();
        if (lane == 0) {
          g.qkey[slot] = g.qkey[qcount - 1];
          g.qnode[slot] = g.qnode[qcount - 1];
        }
        --qcount;
        __syncwarp();
        const int level = (int)(nx_node >> kNodeIdxBits);
        const int idx = (int)(nx_node & ((1u << kNodeIdxBits) - 1u));
        const int rem = t.count[level - 1] - idx * kFanout;
        load_child_box(t, level - 1, idx * kFanout, rem < kFanout ? rem : kFanout, nx_a, nx_b);
      }
    }

    // ---- process the inner node selected one iteration ago
    if (cur_node != kNoNode) {
      if constexpr (!LeafOp::kBuffered) {
        if (t.node_budget >= 0 && --budget < 0) break;    // approximate search: inspection budget spent
      }
      if (cur_key < __float_as_uint(g.maxR2)) {          // the radius may have shrunk since it was popped
        const int level = (int)(cur_node >> kNodeIdxBits);
        const int idx = (int)(cur_node & ((1u << kNodeIdxBits) - 1u));
        ++g.n_inner;
        const int rem = t.count[level - 1] - idx * kFanout;
        const int count = rem < kFanout ? rem : kFanout;
        if (level == 1) run_start(idx * kFanout, count, cur_a, cur_b);
        else frontier_expand(level - 1, idx * kFanout, count, cur_a, cur_b, seed_l1, t.queue_cap, t.cull_scale, qcount, depth, g);
      }
    } else if (nx_node == kNoNode) {
      // nothing in flight and nothing left inside the radius
      if (fill_pos > 0 && fill_leaves > 0) { fill_finish(g); fill_leaves = 0; continue; }   // ran out of leaves during the fill
      if (!started) {
        // the seed run is over: open the root (with the radius as of the last flush; whatever that
        // lets in too generously is dropped again when its turn comes)
        started = true;
        float4 a, b;
        load_child_box(t, top_level, 0, t.count[top_level], a, b);
        if (top_level == 0) { if (seed_l1 < 0) run_start(0, t.count[0], a, b); }
        else frontier_expand(top_level, 0, t.count[top_level], a, b, seed_l1, t.queue_cap, t.cull_scale, qcount, depth, g);
        continue;
      }
      if constexpr (LeafOp::kBuffered) {
        if (__any_sync(kFullMask, g.cnt > 0)) { g.flush_pending = true; continue; }
      }
      break;
    }
    cur_node = nx_node; cur_key = nx_key; cur_a = nx_a; cur_b = nx_b;
  }
}

// Shared-memory floats per warp when the heap lives in GLOBAL memory (k too large for shared memory).
__host__ __device__ constexpr size_t warp_smem_floats_noheap() { return warp_aux_floats(); }

// Carve one warp's slice of dynamic shared memory: candidate structure first, then the staging and
// frontier areas.
//   sorted               list (k + 1 rows) | append buffer (kBufRows + 1 rows) | aux
//   heap in shared       padded heap | aux
//   heap in global       (global_heap != nullptr) this warp's [kp][32] block of `global_heap`; shared
//                        memory only holds the aux areas.
__device__ __forceinline__ void group_ctx_init(GroupCtx& g, float* smem, int warp, int lane, int k,
                                               float* global_heap = nullptr, bool sorted = false) {
  const int kp = sorted ? list_rows(k) : heap_padded(k);
  float* aux;
  g.buf = nullptr;
  if (global_heap != nullptr) {
    aux = smem + warp * warp_smem_floats_noheap();
    g.heap = global_heap + lane;
  } else {
    float* warp_base = smem + warp * warp_smem_floats(k, sorted);
    g.heap = warp_base + lane;
    aux = warp_base + (size_t)kp * 32;
    if (sorted) {
      g.buf = aux + lane;
      aux += (size_t)(kBufRows + 1) * 32;
    }
  }
  g.stage = sorted ? nullptr : reinterpret_cast<float4*>(aux);
  g.qkey = reinterpret_cast<unsigned*>(aux + (sorted ? 0 : 128));
  g.qnode = g.qkey + kQueueCap;
  g.frames = g.qnode + kQueueCap;
  g.soa = reinterpret_cast<float*>(g.frames + 32);
  g.work = reinterpret_cast<unsigned*>(g.soa + 96);
  g.kp = kp;
  g.cull_scale = 1.f;
  g.cnt = 0;
  g.live = k;
  g.flush_pending = false;
  g.n_leaves = g.n_inner = g.n_inserts = g.n_fallback = g.n_purge = g.n_flush = 0;
}

// Group bounding box + group radius from the lanes' queries and heap tops.
__device__ __forceinline__ void group_bounds(GroupCtx& g) {
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
}

// ---- robust entry point: split incoherent groups --------------------------------------------------
// A group is 32 consecutive points of a k-d refined Morton run. When the run straddles a jump of the
// Z-curve (or the gap between two clusters) a leaf can hold points from two far-apart places. As a
// QUERY group such a leaf is poison: its box covers everything in between, so every node passes the
// group-level test and the lanes on the far side keep a huge radius until the random walk happens to
// reach their real neighbourhood — measured: one such group runs for tens of milliseconds and, if it
// is claimed late, the whole GPU waits for it (local phase of one rank 188 ms instead of 131 ms).
// Fix: before traversing, cut the lane set at the largest gap along an axis whenever that gap is more
// than two thirds of the extent along that axis, recursively, and traverse each coherent sub-group on its own
// (the other lanes are simply not valid during that pass). Exactness is unaffected.
// fill_leaves: see traverse(); pass k / 32 for a group whose heaps are all fresh, else 0.
template <class LeafOp>
__device__ __forceinline__ void traverse_group(const TreeView& t, GroupCtx& g, int cnt, int own_seed_l1,
                                               int fill_leaves, LeafOp& leaf_op) {
  const int lane = lane_id();
  const unsigned all = cnt >= 32 ? 0xffffffffu : ((1u << cnt) - 1u);
  int nwork = 1;
  __syncwarp();
  if (lane == 0) g.work[0] = all;
  __syncwarp();
  while (nwork > 0) {
    const unsigned m = g.work[--nwork];
    __syncwarp();
    g.valid = (m >> lane) & 1u;
    group_bounds(g);
    // Largest gap between consecutive coordinates along each axis: cut where a gap takes up more than two
    // thirds of the set's extent along its axis (and is not negligible next to the widest extent) — the
    // widest such gap if several axes qualify. 32 evenly spread points never qualify, nor do the mildly
    // two-lobed leaves at the joints of a Morton run (cutting those costs a second walk over nearly the
    // same leaves: tried, +13 % on uniform data); what qualifies is a set in two places, whose second
    // place would otherwise start with a useless radius — the distance to the first.
    const float ex = g.ghx - g.glx, ey = g.ghy - g.gly, ez = g.ghz - g.glz;
    const float emax = fmaxf(ex, fmaxf(ey, ez));
    float best_gap = 0.f, best_c = 0.f;
    int best_wl = -1;
#pragma unroll 1
    for (int axis = 0; axis < 3; ++axis) {
      const float c = g.valid ? (axis == 0 ? g.qx : (axis == 1 ? g.qy : g.qz)) : INFINITY;
      const float extent = axis == 0 ? ex : (axis == 1 ? ey : ez);
      float succ = INFINITY;
#pragma unroll 4
      for (int j = 0; j < 32; ++j) {
        const float cj = __shfl_sync(kFullMask, c, j);
        if (cj > c) succ = fminf(succ, cj);
      }
      const float gap = (g.valid && succ < INFINITY) ? succ - c : 0.f;
      const unsigned gbits = __float_as_uint(gap);
      const unsigned gmax = __reduce_max_sync(kFullMask, gbits);
      const float gm = __uint_as_float(gmax);
      if (gm > best_gap && gm > 0.66f * extent && gm > 0.33f * emax) {     // warp-uniform
        best_gap = gm;
        best_wl = __ffs(__ballot_sync(kFullMask, gbits == gmax)) - 1;
        best_c = c;
      }
    }
    if (__popc(m) >= 2 && nwork <= 6 && best_wl >= 0) {
      const float cut = __shfl_sync(kFullMask, best_c, best_wl);
      const unsigned left = __ballot_sync(kFullMask, g.valid && best_c <= cut);
      const unsigned right = m & ~left;
      if (left != 0u && right != 0u) {
        if (lane == 0) { g.work[nwork] = right; g.work[nwork + 1] = left; }
        nwork += 2;
        __syncwarp();
        continue;
      }
    }
    // coherent enough: one pass. The group's own level-1 node is the natural seed only for the
    // unsplit group; a split part looks for the node nearest to it.
    const int seed = (m == all && own_seed_l1 != -2) ? own_seed_l1 : find_seed_l1(t, g);
    traverse(t, g, seed, fill_leaves, leaf_op);
  }
  g.valid = lane < cnt;
}

// SORTED = true: candidates in sorted lists behind append buffers (k >= 2, shared memory);
// SORTED = false: the heap (k == 1 with ids, or the global-memory fallback for very large k).
template <bool SORTED>
__device__ __forceinline__ void traverse_group_knn(const TreeView& t, GroupCtx& g, int cnt, int own_seed_l1, int k,
                                                   int fill_leaves) {
  if constexpr (SORTED) {
    SortedLeafOp op{k};
    traverse_group(t, g, cnt, own_seed_l1, 0, op);
  } else {
    HeapLeafOp op;
    traverse_group(t, g, cnt, own_seed_l1, fill_leaves, op);
  }
}

__device__ __forceinline__ void flush_stats(const GroupCtx& g, unsigned long long* stats) {
  if (stats == nullptr) return;
  unsigned ins = g.n_inserts;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) ins += __shfl_xor_sync(kFullMask, ins, o);
  if (lane_id() == 0) {
    atomicAdd(stats + 0, (unsigned long long)g.n_leaves);
    atomicAdd(stats + 1, (unsigned long long)g.n_leaves * kGroup);
    atomicAdd(stats + 2, (unsigned long long)ins);
    atomicAdd(stats + 3, (unsigned long long)g.n_inner);
    atomicAdd(stats + 4, (unsigned long long)g.n_fallback);
    atomicAdd(stats + 5, (unsigned long long)g.n_purge);
    atomicAdd(stats + 6, (unsigned long long)g.n_flush);
  }
}

constexpr size_t kSmemPerSM = 233472;     // 228 KB
constexpr size_t kSmemPerCtaMax = 232448; // 227 KB
constexpr size_t kSmemCtaReserve = 1024;

}  // namespace

// Launch shape of any kernel built on GroupCtx: warps per CTA and dynamic shared memory bytes.
size_t group_smem_bytes(int k, int warps_per_cta, bool sorted);
int group_pick_warps(int k, const void* kernel, bool sorted);
// Which candidate structure a k-th-distance kernel uses for this k: the sorted list wherever it fits in
// shared memory (k >= 2); the heap for k == 1 (it also tracks the winner's id) and, in global memory,
// for k beyond the shared-memory limit. Every kernel that touches a travelling group must agree on it.
static inline bool use_sorted_list(int k) {
  return k >= 2 && (size_t)warp_smem_floats(k, true) * sizeof(float) + 1024 <= 232448;
}

}  // namespace b200knn
