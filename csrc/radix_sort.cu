// radix_sort.cu — hand-written stable LSD radix sort of (key, uint32 value) pairs for sm_100a.
//
// Replaces the reference's thrust::sort / cub merge sort that the default k-d tree builder
// runs log2(N) times (submodules/cudaKDTree/cukd/builder_thrust.h:486-488,508-510): here the
// whole spatial order is produced by ONE sort of 63-bit Morton keys (8 digit passes).
//
// Per 8-bit digit pass, three kernels:
//   rs_count   : per-tile digit histogram (smem atomics)  -> counts[digit][tile], totals[digit]
//   rs_scan    : one block per digit: exclusive scan over tiles, offset by the digit's base
//   rs_scatter : stable in-tile ranking with __match_any_sync, tile reordered in shared memory so
//                each digit run leaves the SM as contiguous stores, then scattered to its slot
// Keys are read twice per pass (count + scatter); the build is <5 % of a kNN run, so the simpler
// non-look-back structure is kept (see DESIGN.md "radix sort").
#include "common.cuh"
#include "kernels.h"

namespace b200knn {

namespace {

constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;   // 4096 keys per tile
constexpr int RS_RADIX = 256;

template <typename K>
__device__ __forceinline__ unsigned digit_of(K key, int shift) {
  return (unsigned)(key >> shift) & 0xffu;
}

template <typename K>
__global__ void __launch_bounds__(RS_THREADS)
rs_count(const K* __restrict__ keys, uint32_t n, int shift, int num_tiles,
         uint32_t* __restrict__ counts /*[256][num_tiles]*/, uint32_t* __restrict__ totals /*[256]*/) {
  __shared__ uint32_t hist[RS_RADIX];
  const int tile = blockIdx.x;
  hist[threadIdx.x] = 0;
  __syncthreads();
  const uint32_t base = (uint32_t)tile * RS_TILE;
#pragma unroll
  for (int i = 0; i < RS_ITEMS; ++i) {
    const uint32_t idx = base + i * RS_THREADS + threadIdx.x;
    if (idx < n) atomicAdd(&hist[digit_of(keys[idx], shift)], 1u);
  }
  __syncthreads();
  const uint32_t c = hist[threadIdx.x];
  counts[(size_t)threadIdx.x * num_tiles + tile] = c;
  if (c) atomicAdd(&totals[threadIdx.x], c);
}

// One block per digit. Exclusive scan of counts[digit][0..num_tiles) plus the digit's global base.
__global__ void __launch_bounds__(1024)
rs_scan(uint32_t* __restrict__ counts, const uint32_t* __restrict__ totals, int num_tiles) {
  __shared__ uint32_t warp_sums[32];
  __shared__ uint32_t carry;
  const int digit = blockIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    uint32_t b = 0;
    for (int d = 0; d < digit; ++d) b += totals[d];
    carry = b;
  }
  __syncthreads();
  uint32_t* row = counts + (size_t)digit * num_tiles;
  for (int start = 0; start < num_tiles; start += 1024) {
    const int i = start + threadIdx.x;
    const uint32_t v = (i < num_tiles) ? row[i] : 0u;
    uint32_t x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(kFullMask, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) warp_sums[warp] = x;
    __syncthreads();
    if (warp == 0) {
      uint32_t w = warp_sums[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(kFullMask, w, o);
        if (lane >= o) w += y;
      }
      warp_sums[lane] = w;   // inclusive over warps
    }
    __syncthreads();
    const uint32_t warp_excl = warp ? warp_sums[warp - 1] : 0u;
    const uint32_t c = carry;
    if (i < num_tiles) row[i] = c + warp_excl + x - v;
    __syncthreads();
    if (threadIdx.x == 0) carry = c + warp_sums[31];
    __syncthreads();
  }
}

template <typename K>
__global__ void __launch_bounds__(RS_THREADS)
rs_scatter(const K* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
           K* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
           uint32_t n, int shift, int num_tiles, const uint32_t* __restrict__ offsets /*[256][tiles]*/) {
  extern __shared__ __align__(16) unsigned char rs_dyn_smem[];
  K* s_keys = reinterpret_cast<K*>(rs_dyn_smem);
  uint32_t* s_vals = reinterpret_cast<uint32_t*>(rs_dyn_smem + sizeof(K) * RS_TILE);
  __shared__ uint32_t warp_hist[RS_WARPS][RS_RADIX];   // per-warp digit counts -> per-warp bases
  __shared__ uint32_t tile_excl[RS_RADIX];             // exclusive start of each digit inside the tile
  __shared__ uint32_t glob_base[RS_RADIX];             // global slot of the digit's first key of this tile

  const int tile = blockIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t tile_base = (uint32_t)tile * RS_TILE;
  const uint32_t warp_base = tile_base + warp * (32 * RS_ITEMS);

#pragma unroll
  for (int w = 0; w < RS_WARPS; ++w) warp_hist[w][threadIdx.x] = 0;
  __syncthreads();

  K key[RS_ITEMS];
  uint32_t val[RS_ITEMS];
  uint32_t rank[RS_ITEMS];   // rank among keys of the same digit within this warp
#pragma unroll
  for (int i = 0; i < RS_ITEMS; ++i) {
    const uint32_t idx = warp_base + i * 32 + lane;
    const bool ok = idx < n;
    key[i] = ok ? keys_in[idx] : (K)~(K)0;
    val[i] = ok ? vals_in[idx] : 0u;
  }
  const unsigned lt_mask = (1u << lane) - 1u;
#pragma unroll
  for (int i = 0; i < RS_ITEMS; ++i) {
    const uint32_t idx = warp_base + i * 32 + lane;
    const bool ok = idx < n;
    // Out-of-range slots take digit 255 AND sit at the very end of the tile (last warp, last
    // items), so stable ranking places them after every real key of digit 255.
    const unsigned d = ok ? digit_of(key[i], shift) : 255u;
    const unsigned peers = __match_any_sync(kFullMask, d);
    const uint32_t prev = warp_hist[warp][d];
    __syncwarp();
    if ((peers & lt_mask) == 0) warp_hist[warp][d] = prev + __popc(peers);
    __syncwarp();
    rank[i] = prev + __popc(peers & lt_mask);
  }
  __syncthreads();

  // Thread d: turn per-warp counts of digit d into per-warp exclusive bases; get tile total.
  {
    const int d = threadIdx.x;
    uint32_t run = 0;
#pragma unroll
    for (int w = 0; w < RS_WARPS; ++w) {
      const uint32_t c = warp_hist[w][d];
      warp_hist[w][d] = run;
      run += c;
    }
    // exclusive scan of tile totals across the 256 digits
    uint32_t x = run;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(kFullMask, x, o);
      if (lane >= o) x += y;
    }
    __shared__ uint32_t wsum[RS_WARPS];
    if (lane == 31) wsum[warp] = x;
    __syncthreads();
    uint32_t wbase = 0;
#pragma unroll
    for (int w = 0; w < RS_WARPS; ++w) if (w < warp) wbase += wsum[w];
    tile_excl[d] = wbase + x - run;
    glob_base[d] = offsets[(size_t)d * num_tiles + tile];
  }
  __syncthreads();

  // Reorder the tile in shared memory.
#pragma unroll
  for (int i = 0; i < RS_ITEMS; ++i) {
    const uint32_t idx = warp_base + i * 32 + lane;
    const unsigned d = (idx < n) ? digit_of(key[i], shift) : 255u;
    const uint32_t pos = tile_excl[d] + warp_hist[warp][d] + rank[i];
    s_keys[pos] = key[i];
    s_vals[pos] = val[i];
  }
  __syncthreads();

  const uint32_t valid = (n - tile_base < (uint32_t)RS_TILE) ? (n - tile_base) : (uint32_t)RS_TILE;
#pragma unroll
  for (int i = 0; i < RS_ITEMS; ++i) {
    const uint32_t pos = i * RS_THREADS + threadIdx.x;
    if (pos < valid) {
      const K k = s_keys[pos];
      const unsigned d = digit_of(k, shift);
      const uint32_t dst = glob_base[d] + (pos - tile_excl[d]);
      keys_out[dst] = k;
      vals_out[dst] = s_vals[pos];
    }
  }
}

template <typename K>
int radix_sort_impl(K* keys, uint32_t* vals, K* keys_tmp, uint32_t* vals_tmp, uint32_t* scratch,
                    int64_t n, int begin_bit, int end_bit, cudaStream_t stream, int* result_in_tmp) {
  *result_in_tmp = 0;
  if (n <= 0) return 0;
  KNN_REQUIRE(n < (int64_t)0x7fffffff, "radix_sort: n must be < 2^31 per GPU");
  const int num_tiles = (int)div_up(n, RS_TILE);
  uint32_t* totals = scratch;                 // [256]
  uint32_t* counts = scratch + RS_RADIX;      // [256][num_tiles]
  K* kin = keys; K* kout = keys_tmp;
  uint32_t* vin = vals; uint32_t* vout = vals_tmp;
  int flips = 0;
  const int dyn_smem = (int)((sizeof(K) + sizeof(uint32_t)) * RS_TILE);
  KNN_CUDA_TRY(cudaFuncSetAttribute(rs_scatter<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_smem));
  for (int shift = begin_bit; shift < end_bit; shift += 8) {
    KNN_CUDA_TRY(cudaMemsetAsync(totals, 0, RS_RADIX * sizeof(uint32_t), stream));
    rs_count<K><<<num_tiles, RS_THREADS, 0, stream>>>(kin, (uint32_t)n, shift, num_tiles, counts, totals);
    note_launch();
    rs_scan<<<RS_RADIX, 1024, 0, stream>>>(counts, totals, num_tiles);
    note_launch();
    rs_scatter<K><<<num_tiles, RS_THREADS, dyn_smem, stream>>>(kin, vin, kout, vout, (uint32_t)n, shift,
                                                       num_tiles, counts);
    note_launch();
    K* tk = kin; kin = kout; kout = tk;
    uint32_t* tv = vin; vin = vout; vout = tv;
    ++flips;
  }
  KNN_CUDA_TRY(cudaGetLastError());
  *result_in_tmp = flips & 1;
  return 0;
}

}  // namespace

size_t radix_sort_scratch_bytes(int64_t n) {
  const int64_t tiles = div_up(n > 0 ? n : 1, RS_TILE);
  return (size_t)(RS_RADIX + RS_RADIX * tiles) * sizeof(uint32_t);
}

int radix_sort_u64(uint64_t* keys, uint32_t* vals, uint64_t* keys_tmp, uint32_t* vals_tmp,
                   uint32_t* scratch, int64_t n, int begin_bit, int end_bit, cudaStream_t stream,
                   int* result_in_tmp) {
  return radix_sort_impl<uint64_t>(keys, vals, keys_tmp, vals_tmp, scratch, n, begin_bit, end_bit,
                                   stream, result_in_tmp);
}

int radix_sort_u32(uint32_t* keys, uint32_t* vals, uint32_t* keys_tmp, uint32_t* vals_tmp,
                   uint32_t* scratch, int64_t n, int begin_bit, int end_bit, cudaStream_t stream,
                   int* result_in_tmp) {
  return radix_sort_impl<uint32_t>(keys, vals, keys_tmp, vals_tmp, scratch, n, begin_bit, end_bit,
                                   stream, result_in_tmp);
}

}  // namespace b200knn
