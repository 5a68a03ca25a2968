// redist.cu — device-driven redistribution of an unordered cloud to spatial owners, and the way back.
//
// No counterpart in the reference (it never redistributes); in this repository it replaces
// the first version of dist.cu::knn_redistribute, which sorted every rank's points by Morton key (4 radix
// passes + gather), took splitters from sorted samples on the HOST (five bootstrap all-gathers, three host
// barriers, >= 10 stream synchronisations) and sorted the (id, value) pairs again for the way back.
// An owner sorts what it receives anyway, so the order inside a run is irrelevant in both directions:
// the sender only has to PARTITION. Everything below is stream-ordered; the host synchronises once, to
// learn how large the receive buffers must be.
//
//   rd_hist        Morton key of every local point -> two histograms of its leading bits: A (12 bits,
//                  4096 bins, privatised in shared memory) and B (21 bits, 2 Mi bins, L2 atomics).
//   (device barrier: everybody's histograms are complete and visible)
//   rd_split       one CTA per rank, identical result on every rank: global A histogram = sum of all
//                  ranks' A (peer loads) -> the A bin in which the global count crosses each target, then
//                  the 512 B bins of that A bin summed over all ranks -> crossing to 1/2 Mi of key space.
//                  Each splitter is the coarsest Morton-cell boundary within +-0.5 % of N around its
//                  quantile (regions made of whole coarse cells keep the ranks' boxes from overlapping).
//                  Also this rank's own row of the count matrix (points it will send to each owner).
//   (small device all-gather of the rows -> count matrix M on every rank)
//   rd_plan        offsets of this rank's runs inside every owner's receive / result buffers, sizes.
//   rd_scatter     every point -> its owner: a block partitions 1024 points by destination in shared memory,
//                  reserves one range per destination with one atomic, and writes each destination's run
//                  with consecutive lanes (coalesced 16-byte peer stores).
//   rd_return      k-th distances go home the same way: (global id, value) pairs partitioned by owner.
#include "common.cuh"
#include "kernels.h"

namespace b200knn {
namespace {

constexpr int kABits = 12, kBBits = 21;
constexpr int kABins = 1 << kABits;                 // 4096
constexpr int kBPerA = 1 << (kBBits - kABits);      // 512
constexpr int kKeyBits = 63;

struct KeyFrame { float lox, loy, loz, ivx, ivy, ivz; };

// Global bounding box = union of the ranks' boxes (order-preserving uint encoding, words 0..5 of each rank's info).
__device__ __forceinline__ KeyFrame key_frame(const unsigned* __restrict__ info_all, int R) {
  unsigned lo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, hi[3] = {0u, 0u, 0u};
  for (int r = 0; r < R; ++r) {
    const unsigned* w = info_all + (size_t)r * kRdInfoWords;
    if ((w[6] | w[7]) == 0u) continue;               // an empty rank has no box
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      lo[a] = w[a] < lo[a] ? w[a] : lo[a];
      hi[a] = w[3 + a] > hi[a] ? w[3 + a] : hi[a];
    }
  }
  KeyFrame f;
  const float lx = ord2f(lo[0]), ly = ord2f(lo[1]), lz = ord2f(lo[2]);
  const float ex = ord2f(hi[0]) - lx, ey = ord2f(hi[1]) - ly, ez = ord2f(hi[2]) - lz;
  f.lox = lx; f.loy = ly; f.loz = lz;
  f.ivx = ex > 0.f ? 1.f / ex : 0.f;
  f.ivy = ey > 0.f ? 1.f / ey : 0.f;
  f.ivz = ez > 0.f ? 1.f / ez : 0.f;
  return f;
}

__global__ void __launch_bounds__(256)
rd_hist_kernel(const float* __restrict__ xyz, int64_t n, const unsigned* __restrict__ info_all, int R,
               unsigned* __restrict__ histA, unsigned* __restrict__ histB) {
  __shared__ unsigned sA[kABins];
  for (int i = threadIdx.x; i < kABins; i += blockDim.x) sA[i] = 0u;
  const KeyFrame f = key_frame(info_all, R);
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint64_t key = morton63(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], f.lox, f.loy, f.loz, f.ivx, f.ivy, f.ivz);
    atomicAdd(&sA[(unsigned)(key >> (kKeyBits - kABits))], 1u);
    atomicAdd(&histB[(unsigned)(key >> (kKeyBits - kBBits))], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < kABins; i += blockDim.x)
    if (sA[i]) atomicAdd(&histA[i], sA[i]);
}

// Block-wide inclusive scan of one long long per thread (1024 threads).
__device__ __forceinline__ long long block_scan_incl(long long v, long long* warp_sums) {
  const int lane = lane_id(), w = threadIdx.x >> 5;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const long long u = __shfl_up_sync(kFullMask, v, o);
    if (lane >= o) v += u;
  }
  __syncthreads();
  if (lane == 31) warp_sums[w] = v;
  __syncthreads();
  if (w == 0) {
    long long s = warp_sums[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const long long u = __shfl_up_sync(kFullMask, s, o);
      if (lane >= o) s += u;
    }
    warp_sums[lane] = s;
  }
  __syncthreads();
  if (w > 0) v += warp_sums[w - 1];
  return v;
}

// One CTA of 1024 threads. B-bin index (21 bits) of the first bin whose global inclusive prefix exceeds `want`
// is what "crossing" means below: all keys in bins before it number <= want.
__global__ void __launch_bounds__(1024)
rd_split_kernel(const RdSplitParams p) {
  __shared__ long long prefA[kABins];        // inclusive global prefix over A bins
  __shared__ long long warp_sums[32];
  __shared__ long long binB[kBPerA];
  __shared__ int crossA[3 * kMaxRanks];
  __shared__ unsigned crossB[3 * kMaxRanks];
  __shared__ unsigned split_bin[kMaxRanks + 1];
  const int t = threadIdx.x, R = p.R;

  // ---- global A histogram and its prefix (thread t owns bins 4t..4t+3)
  long long a4[4] = {0, 0, 0, 0};
  for (int r = 0; r < R; ++r) {
    const uint4 v = __ldcg(reinterpret_cast<const uint4*>(p.histA[r]) + t);
    a4[0] += v.x; a4[1] += v.y; a4[2] += v.z; a4[3] += v.w;
  }
  const long long mine = a4[0] + a4[1] + a4[2] + a4[3];
  const long long incl = block_scan_incl(mine, warp_sums);
  long long run = incl - mine;
#pragma unroll
  for (int j = 0; j < 4; ++j) { run += a4[j]; prefA[4 * t + j] = run; }
  __syncthreads();
  const long long N = prefA[kABins - 1];
  const long long window = N / 200;

  // ---- for each splitter j = 1..R-1 and each of (target - w, target, target + w): the A bin of the crossing
  const int ncross = 3 * (R - 1);
  if (t < ncross) {
    const int j = t / 3 + 1, which = t % 3;
    long long want = N * j / R + (which - 1) * window;          // N < 2^36, j <= 16
    want = want < 0 ? 0 : (want >= N ? N - 1 : want);
    int lo = 0, hi = kABins - 1;               // first bin with prefA > want
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (prefA[mid] > want) hi = mid; else lo = mid + 1;
    }
    crossA[t] = lo;
  }
  __syncthreads();
  // ---- refine every crossing inside its A bin with the global B histogram of that bin
  for (int c = 0; c < ncross; ++c) {
    const int a = crossA[c];
    if (t < kBPerA) {
      long long s = 0;
      for (int r = 0; r < R; ++r) s += __ldcg(p.histB[r] + (size_t)a * kBPerA + t);
      binB[t] = s;
    }
    __syncthreads();
    if (t < 32) {                               // warp 0: lane l owns B bins 16l..16l+15 of this A bin
      const int j = c / 3 + 1, which = c % 3;
      long long want = N * j / R + (which - 1) * window;          // N < 2^36, j <= 16
      want = want < 0 ? 0 : (want >= N ? N - 1 : want);
      const long long before_a = a > 0 ? prefA[a - 1] : 0;
      long long mine16 = 0;
      for (int i = 0; i < kBPerA / 32; ++i) mine16 += binB[t * (kBPerA / 32) + i];
      long long incl16 = mine16;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const long long u = __shfl_up_sync(kFullMask, incl16, o);
        if (t >= o) incl16 += u;
      }
      const unsigned over = __ballot_sync(kFullMask, before_a + incl16 > want);
      const int owner = over ? __ffs(over) - 1 : 31;
      if (t == owner) {
        long long acc = before_a + incl16 - mine16;
        int b = t * (kBPerA / 32);
        const int last = b + kBPerA / 32 - 1;
        for (; b < last; ++b) {
          if (acc + binB[b] > want) break;
          acc += binB[b];
        }
        crossB[c] = (unsigned)a * kBPerA + (unsigned)b;
      }
    }
    __syncthreads();
  }
  // ---- snap: the coarsest cell boundary in (lo, hi], else the bin of the exact quantile
  if (t < R - 1) {
    const unsigned lo = crossB[3 * t], mid = crossB[3 * t + 1], hi = crossB[3 * t + 2];
    unsigned best = mid;
    for (int b = kBBits - 1; b >= 0; --b) {
      const unsigned cand = (hi >> b) << b;
      if (cand >= lo && cand > 0u) { best = cand; break; }
    }
    split_bin[t + 1] = best;
  }
  if (t == 0) { split_bin[0] = 0u; }
  __syncthreads();
  // splitters must not decrease (degenerate inputs: everything in one bin)
  if (t == 0) {
    for (int j = 2; j < R; ++j) if (split_bin[j] < split_bin[j - 1]) split_bin[j] = split_bin[j - 1];
    for (int j = 1; j < R; ++j) p.split_keys[j - 1] = (uint64_t)split_bin[j] << (kKeyBits - kBBits);
  }
  __syncthreads();

  // ---- this rank's row of the count matrix: own points below each splitter, from the own histograms
  // below[j] = # own keys in B bins < split_bin[j]  (j = 1..R-1), below[0] = 0, below[R] = own n
  __shared__ long long below[kMaxRanks + 1];
  for (int j = 1; j < R; ++j) {
    const unsigned sb = split_bin[j];
    const int a = (int)(sb >> (kBBits - kABits)), bin = (int)(sb & (kBPerA - 1));
    long long part = 0;
    for (int i = t; i < a; i += blockDim.x) part += p.histA[p.rank][i];
    if (t < bin) part += p.histB[p.rank][(size_t)a * kBPerA + t];
    const long long tot = block_scan_incl(part, warp_sums);
    if (t == blockDim.x - 1) below[j] = tot;
    __syncthreads();
  }
  if (t == 0) {
    below[0] = 0;
    below[R] = p.n_local;
    for (int d = 0; d < R; ++d) p.row[d] = below[d + 1] - below[d];
  }
}

// M[s][d] (gathered rows) -> where this rank's runs start inside every owner's buffers, and the sizes the host needs.
__global__ void rd_plan_kernel(const long long* __restrict__ M, int R, int rank, RdPlan* __restrict__ plan) {
  if (threadIdx.x != 0) return;
  long long max_recv = 0;
  for (int d = 0; d < R; ++d) {
    long long before = 0, total = 0;
    for (int s = 0; s < R; ++s) {
      if (s < rank) before += M[(size_t)s * R + d];
      total += M[(size_t)s * R + d];
    }
    plan->recv_offset[d] = before;
    if (d == rank) plan->my_recv = total;
    max_recv = total > max_recv ? total : max_recv;
    // the way back: owner o gets from sender s what o sent to s; this rank's run in o's landing zone starts
    // after the runs of the senders before it
    long long ret = 0;
    for (int s = 0; s < rank; ++s) ret += M[(size_t)d * R + s];
    plan->ret_offset[d] = ret;
  }
  plan->max_recv = max_recv;
  for (int d = 0; d < R; ++d) { plan->cursor[d] = 0ull; plan->ret_cursor[d] = 0ull; }
}

constexpr int kScatterThreads = 256, kScatterPer = 4, kScatterTile = kScatterThreads * kScatterPer;

__device__ __forceinline__ int dest_of(uint64_t key, const uint64_t* split, int R) {
  int d = 0;
  for (int j = 0; j < R - 1; ++j) d += key >= split[j] ? 1 : 0;   // splitters are sorted
  return d;
}

// Shared by both directions: a block's tile of items is partitioned by destination in shared memory; one
// atomic per destination reserves the block's range in the destination's buffer; consecutive lanes then
// write consecutive slots of a run.
template <class Item>
__device__ __forceinline__ void tile_scatter(const Item* items, const int* dest, int count_in_tile, int R,
                                             Item* const* dst_base, const long long* dst_offset,
                                             unsigned long long* cursor, Item* stage, int* s_count, int* s_start,
                                             long long* s_base) {
  const int t = threadIdx.x;
  if (t < kMaxRanks) s_count[t] = 0;
  __syncthreads();
  // warp-aggregated: lanes with the same destination share one shared-memory atomic (a tile has <= 16
  // destinations, so per-item atomics would pile up on a handful of addresses)
  int slot[kScatterPer];
  const unsigned lt = (1u << (t & 31)) - 1u;
#pragma unroll
  for (int j = 0; j < kScatterPer; ++j) {
    const int i = t + j * kScatterThreads;
    const int dj = i < count_in_tile ? dest[j] : -1;
    const unsigned peers = __match_any_sync(kFullMask, dj);
    int base = 0;
    const int leader = __ffs(peers) - 1;
    if ((t & 31) == leader && dj >= 0) base = atomicAdd(&s_count[dj], __popc(peers));
    base = __shfl_sync(kFullMask, base, leader);
    slot[j] = base + __popc(peers & lt);
  }
  __syncthreads();
  if (t == 0) {
    int acc = 0;
    for (int d = 0; d < R; ++d) { s_start[d] = acc; acc += s_count[d]; }
    s_start[R] = acc;
  }
  if (t < R && s_count[t] > 0) s_base[t] = dst_offset[t] + (long long)atomicAdd(&cursor[t], (unsigned long long)s_count[t]);
  __syncthreads();
#pragma unroll
  for (int j = 0; j < kScatterPer; ++j) {
    const int i = t + j * kScatterThreads;
    if (i < count_in_tile) stage[s_start[dest[j]] + slot[j]] = items[j];
  }
  __syncthreads();
  for (int i = t; i < count_in_tile; i += kScatterThreads) {
    int d = 0;
    while (i >= s_start[d + 1]) ++d;
    dst_base[d][s_base[d] + (i - s_start[d])] = stage[i];
  }
  __syncthreads();
}

__global__ void __launch_bounds__(kScatterThreads)
rd_scatter_kernel(const RdScatterParams p) {
  __shared__ float4 stage[kScatterTile];
  __shared__ uint64_t split[kMaxRanks];
  __shared__ int s_count[kMaxRanks], s_start[kMaxRanks + 1];
  __shared__ long long s_base[kMaxRanks];
  __shared__ float4* dst_base[kMaxRanks];
  __shared__ long long dst_offset[kMaxRanks];
  const int t = threadIdx.x, R = p.R;
  if (t < R - 1) split[t] = p.split_keys[t];
  if (t < R) { dst_base[t] = reinterpret_cast<float4*>(p.recv[t]); dst_offset[t] = p.plan->recv_offset[t]; }
  const KeyFrame f = key_frame(p.info_all, R);
  __syncthreads();
  const int64_t tiles = (p.n + kScatterTile - 1) / kScatterTile;
  for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const int64_t first = tile * kScatterTile;
    const int cnt = (int)((p.n - first) < kScatterTile ? (p.n - first) : kScatterTile);
    float4 items[kScatterPer];
    int dest[kScatterPer];
#pragma unroll
    for (int j = 0; j < kScatterPer; ++j) {
      const int i = t + j * kScatterThreads;
      dest[j] = 0;
      if (i < cnt) {
        const int64_t g = first + i;
        const float x = p.xyz[3 * g], y = p.xyz[3 * g + 1], z = p.xyz[3 * g + 2];
        items[j] = make_float4(x, y, z, __uint_as_float((uint32_t)(p.global_begin + g)));
        dest[j] = dest_of(morton63(x, y, z, f.lox, f.loy, f.loz, f.ivx, f.ivy, f.ivz), split, R);
      }
    }
    tile_scatter<float4>(items, dest, cnt, R, dst_base, dst_offset, p.plan->cursor, stage, s_count, s_start, s_base);
  }
}

__global__ void __launch_bounds__(kScatterThreads)
rd_return_kernel(const RdReturnParams p) {
  __shared__ uint2 stage[kScatterTile];
  __shared__ long long begins[kMaxRanks + 1];
  __shared__ int s_count[kMaxRanks], s_start[kMaxRanks + 1];
  __shared__ long long s_base[kMaxRanks];
  __shared__ uint2* dst_base[kMaxRanks];
  __shared__ long long dst_offset[kMaxRanks];
  const int t = threadIdx.x, R = p.R;
  if (t <= R) begins[t] = p.begins[t];
  if (t < R) { dst_base[t] = reinterpret_cast<uint2*>(p.dst[t]); dst_offset[t] = p.plan->ret_offset[t]; }
  __syncthreads();
  const int64_t tiles = (p.n + kScatterTile - 1) / kScatterTile;
  for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const int64_t first = tile * kScatterTile;
    const int cnt = (int)((p.n - first) < kScatterTile ? (p.n - first) : kScatterTile);
    uint2 items[kScatterPer];
    int dest[kScatterPer];
#pragma unroll
    for (int j = 0; j < kScatterPer; ++j) {
      const int i = t + j * kScatterThreads;
      dest[j] = 0;
      if (i < cnt) {
        const int64_t g = first + i;
        const uint32_t id = __float_as_uint(p.pts[g].w);
        items[j] = make_uint2(id, __float_as_uint(p.res[g]));
        int o = 0;
        for (int r = 1; r < R; ++r) o += (long long)id >= begins[r] ? 1 : 0;
        dest[j] = o;
      }
    }
    tile_scatter<uint2>(items, dest, cnt, R, dst_base, dst_offset, p.plan->ret_cursor, stage, s_count, s_start, s_base);
  }
}

// Thread s decomposes rank s's key interval, in units of B bins (21 bits = 7 octree levels), into maximal
// aligned cells and turns each into a box in world coordinates.
__global__ void rd_territory_kernel(const uint64_t* __restrict__ split_keys, const unsigned* __restrict__ info_all, int R,
                                    Box* __restrict__ cells, int* __restrict__ counts) {
  const int s = threadIdx.x;
  if (s >= R) return;
  const KeyFrame f = key_frame(info_all, R);
  const float ex = f.ivx > 0.f ? 1.f / f.ivx : 0.f, ey = f.ivy > 0.f ? 1.f / f.ivy : 0.f, ez = f.ivz > 0.f ? 1.f / f.ivz : 0.f;
  const unsigned end_all = 1u << kBBits;
  unsigned a = s == 0 ? 0u : (unsigned)(split_keys[s - 1] >> (kKeyBits - kBBits));
  const unsigned b = s == R - 1 ? end_all : (unsigned)(split_keys[s] >> (kKeyBits - kBBits));
  Box* out = cells + (size_t)s * kTerritoryCells;
  int n = 0;
  const int depth_max = kBBits / 3;                       // 7
  while (a < b && n < kTerritoryCells) {
    int j = 0;                                            // largest j: a is a multiple of 8^j and a + 8^j <= b
    while (j < depth_max && (a & ((1u << (3 * (j + 1))) - 1u)) == 0u && a + (1u << (3 * (j + 1))) <= b) ++j;
    const int d = depth_max - j;                          // octree depth of the cell
    const unsigned prefix = a >> (3 * j);                 // 3 d bits: x y z interleaved, x highest
    unsigned ix = 0, iy = 0, iz = 0;
    for (int l = 0; l < d; ++l) {
      const unsigned tri = (prefix >> (3 * (d - 1 - l))) & 7u;
      ix = (ix << 1) | ((tri >> 2) & 1u);
      iy = (iy << 1) | ((tri >> 1) & 1u);
      iz = (iz << 1) | (tri & 1u);
    }
    // quantised coordinate q (21 bits) belongs to the cell iff q >> (21 - d) == i; q = floor(t * (2^21 - 1)),
    // t = (x - lo) / extent. Two fine cells of slack on either side cover the rounding of that arithmetic.
    const float scale = (float)(1u << (21 - d)) / 2097151.f;
    const float slack = 2.f / 2097151.f;
    Box c;
    c.lox = f.lox + ((float)ix * scale - slack) * ex;  c.hix = f.lox + ((float)(ix + 1) * scale + slack) * ex;
    c.loy = f.loy + ((float)iy * scale - slack) * ey;  c.hiy = f.loy + ((float)(iy + 1) * scale + slack) * ey;
    c.loz = f.loz + ((float)iz * scale - slack) * ez;  c.hiz = f.loz + ((float)(iz + 1) * scale + slack) * ez;
    // the outermost cells also own everything the clamping of t maps onto them
    if (ix == 0) c.lox = -INFINITY;
    if (iy == 0) c.loy = -INFINITY;
    if (iz == 0) c.loz = -INFINITY;
    if (ix + 1 == (1u << d)) c.hix = INFINITY;
    if (iy + 1 == (1u << d)) c.hiy = INFINITY;
    if (iz + 1 == (1u << d)) c.hiz = INFINITY;
    c.cnt_bits = 0.f; c.pad = 0.f;
    out[n++] = c;
    a += 1u << (3 * j);
  }
  if (a < b) {                                            // cannot happen (<= 98 cells); stay conservative if it does
    Box c;
    c.lox = c.loy = c.loz = -INFINITY; c.hix = c.hiy = c.hiz = INFINITY; c.cnt_bits = 0.f; c.pad = 0.f;
    out[0] = c;
    n = 1;
  }
  counts[s] = n;
}

__global__ void rd_pack_info_kernel(const unsigned* __restrict__ bounds_ord, long long n, long long begin,
                                    unsigned* __restrict__ out) {
  const int t = threadIdx.x;
  if (t < 6) out[t] = bounds_ord[t];
  if (t == 6) out[6] = (unsigned)((unsigned long long)n & 0xffffffffull);
  if (t == 7) out[7] = (unsigned)((unsigned long long)n >> 32);
  if (t == 8) out[8] = (unsigned)((unsigned long long)begin & 0xffffffffull);
  if (t == 9) out[9] = (unsigned)((unsigned long long)begin >> 32);
}

}  // namespace

size_t rd_hist_bytes() { return ((size_t)kABins + ((size_t)1 << kBBits)) * sizeof(unsigned); }
size_t rd_histB_offset() { return (size_t)kABins * sizeof(unsigned); }

int rd_pack_info_launch(const unsigned* d_bounds_ord, int64_t n, int64_t begin, unsigned* d_out, cudaStream_t stream) {
  rd_pack_info_kernel<<<1, 32, 0, stream>>>(d_bounds_ord, (long long)n, (long long)begin, d_out);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

int rd_territory_launch(const uint64_t* d_split_keys, const unsigned* d_info_all, int R, Box* d_cells, int* d_counts,
                        cudaStream_t stream) {
  rd_territory_kernel<<<1, 32, 0, stream>>>(d_split_keys, d_info_all, R, d_cells, d_counts);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

int rd_hist_launch(const float* d_xyz, int64_t n, const unsigned* d_info_all, int R, unsigned* histA, unsigned* histB,
                   int num_sms, cudaStream_t stream) {
  KNN_CUDA_TRY(cudaMemsetAsync(histA, 0, rd_hist_bytes(), stream));     // A and B are adjacent
  if (n <= 0) return 0;
  int64_t blocks = (n + 255) / 256;
  if (blocks > (int64_t)num_sms * 8) blocks = (int64_t)num_sms * 8;
  rd_hist_kernel<<<(int)blocks, 256, 0, stream>>>(d_xyz, n, d_info_all, R, histA, histB);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

int rd_split_launch(const RdSplitParams& p, cudaStream_t stream) {
  rd_split_kernel<<<1, 1024, 0, stream>>>(p);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

int rd_plan_launch(const long long* d_M, int R, int rank, RdPlan* d_plan, cudaStream_t stream) {
  rd_plan_kernel<<<1, 32, 0, stream>>>(d_M, R, rank, d_plan);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

int rd_scatter_launch(const RdScatterParams& p, int num_sms, cudaStream_t stream) {
  if (p.n <= 0) return 0;
  int64_t blocks = (p.n + kScatterTile - 1) / kScatterTile;
  if (blocks > (int64_t)num_sms * 8) blocks = (int64_t)num_sms * 8;
  rd_scatter_kernel<<<(int)blocks, kScatterThreads, 0, stream>>>(p);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

int rd_return_launch(const RdReturnParams& p, int num_sms, cudaStream_t stream) {
  if (p.n <= 0) return 0;
  int64_t blocks = (p.n + kScatterTile - 1) / kScatterTile;
  if (blocks > (int64_t)num_sms * 8) blocks = (int64_t)num_sms * 8;
  rd_return_kernel<<<(int)blocks, kScatterThreads, 0, stream>>>(p);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

}  // namespace b200knn
