// build.cu — spatial index construction kernels (bounds, Morton keys, gather, box hierarchy).
//
// Role of cukd::buildTree (submodules/cudaKDTree/cukd/builder.h:108-151 -> builder_thrust.h:414-512)
// in the reference, which runs one full comparison sort per tree level. Here the order comes from a
// single radix sort of 63-bit Morton keys; the "tree" is the sorted point array plus a 32-ary
// hierarchy of bounding boxes built bottom-up by warp reductions. Bounds use the order-preserving
// uint encoding + atomicMin/Max (the reference does this on the CPU: prePartitionedDataVariant.cu:230-232).
#include "common.cuh"
#include "kernels.h"

namespace b200knn {

namespace {

__global__ void k_bounds_init(unsigned* b) {
  if (threadIdx.x < 3) b[threadIdx.x] = 0xffffffffu;       // running min (ordered encoding)
  else if (threadIdx.x < 6) b[threadIdx.x] = 0u;            // running max
}

__global__ void __launch_bounds__(256)
k_bounds(const float* __restrict__ xyz, const PointLayout lay, int64_t n, unsigned* __restrict__ b) {
  float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      const float v = a < lay.dims ? xyz[(size_t)lay.stride * i + a] : 0.f;      // planar points: z = 0
      lo[a] = fminf(lo[a], v);
      hi[a] = fmaxf(hi[a], v);
    }
  }
#pragma unroll
  for (int a = 0; a < 3; ++a) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo[a] = fminf(lo[a], __shfl_xor_sync(kFullMask, lo[a], o));
      hi[a] = fmaxf(hi[a], __shfl_xor_sync(kFullMask, hi[a], o));
    }
  }
  if (lane_id() == 0) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      if (lo[a] <= hi[a]) {
        atomicMin(&b[a], f2ord(lo[a]));
        atomicMax(&b[3 + a], f2ord(hi[a]));
      }
    }
  }
}

__global__ void k_bounds_decode(const unsigned* __restrict__ b, float* __restrict__ out) {
  const int t = threadIdx.x;
  if (t < 6) out[t] = ord2f(b[t]);
}

__global__ void __launch_bounds__(256)
k_morton(const float* __restrict__ xyz, const PointLayout lay, int64_t n, const float* __restrict__ bounds6,
         uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float lox = bounds6[0], loy = bounds6[1], loz = bounds6[2];
  const float ex = bounds6[3] - lox, ey = bounds6[4] - loy, ez = bounds6[5] - loz;
  const float ivx = ex > 0.f ? 1.f / ex : 0.f;
  const float ivy = ey > 0.f ? 1.f / ey : 0.f;
  const float ivz = ez > 0.f ? 1.f / ez : 0.f;
  const size_t o = (size_t)lay.stride * i;
  keys[i] = morton63(xyz[o], xyz[o + 1], lay.dims >= 3 ? xyz[o + 2] : 0.f, lox, loy, loz, ivx, ivy, ivz);
  vals[i] = (uint32_t)i;
}

__global__ void __launch_bounds__(256)
k_gather(const float* __restrict__ xyz, const PointLayout lay, const uint32_t* __restrict__ order,
         const uint32_t* __restrict__ ids, uint32_t id_offset, int64_t n, Float4* __restrict__ sorted) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t src = order[i];
  float4 p;
  const size_t o = (size_t)lay.stride * src;
  p.x = xyz[o];
  p.y = xyz[o + 1];
  p.z = lay.dims >= 3 ? xyz[o + 2] : 0.f;
  // the id travels in .w of the input (redistributed points); else explicit ids or position
  p.w = lay.w_is_id ? xyz[o + 3] : __uint_as_float(ids ? ids[src] : src + id_offset);
  reinterpret_cast<float4*>(sorted)[i] = p;
}

__device__ __forceinline__ void warp_minmax3(float& lx, float& ly, float& lz, float& hx, float& hy, float& hz) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lx = fminf(lx, __shfl_xor_sync(kFullMask, lx, o));
    ly = fminf(ly, __shfl_xor_sync(kFullMask, ly, o));
    lz = fminf(lz, __shfl_xor_sync(kFullMask, lz, o));
    hx = fmaxf(hx, __shfl_xor_sync(kFullMask, hx, o));
    hy = fmaxf(hy, __shfl_xor_sync(kFullMask, hy, o));
    hz = fmaxf(hz, __shfl_xor_sync(kFullMask, hz, o));
  }
}

// One warp per leaf (32 consecutive sorted points).
__global__ void __launch_bounds__(256)
k_leaf_boxes(const Float4* __restrict__ pts, int64_t n, int num_leaves, Box* __restrict__ out) {
  const int leaf = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (leaf >= num_leaves) return;
  const int lane = lane_id();
  const int64_t i = (int64_t)leaf * kGroup + lane;
  float lx = INFINITY, ly = INFINITY, lz = INFINITY, hx = -INFINITY, hy = -INFINITY, hz = -INFINITY;
  if (i < n) {
    const float4 p = reinterpret_cast<const float4*>(pts)[i];
    lx = hx = p.x; ly = hy = p.y; lz = hz = p.z;
  }
  warp_minmax3(lx, ly, lz, hx, hy, hz);
  if (lane == 0) {
    const int64_t rem = n - (int64_t)leaf * kGroup;
    const int cnt = rem < kGroup ? (int)rem : kGroup;
    float4 a = make_float4(lx, ly, lz, __int_as_float(cnt));
    float4 b = make_float4(hx, hy, hz, 0.f);
    reinterpret_cast<float4*>(out)[2 * (size_t)leaf] = a;
    reinterpret_cast<float4*>(out)[2 * (size_t)leaf + 1] = b;
  }
}

// One warp per parent: union of up to 32 child boxes.
__global__ void __launch_bounds__(256)
k_upper_boxes(const Box* __restrict__ child, int child_count, int parent_count, Box* __restrict__ out) {
  const int parent = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (parent >= parent_count) return;
  const int lane = lane_id();
  const int c = parent * kFanout + lane;
  float lx = INFINITY, ly = INFINITY, lz = INFINITY, hx = -INFINITY, hy = -INFINITY, hz = -INFINITY;
  if (c < child_count) {
    const float4 a = reinterpret_cast<const float4*>(child)[2 * (size_t)c];
    const float4 b = reinterpret_cast<const float4*>(child)[2 * (size_t)c + 1];
    lx = a.x; ly = a.y; lz = a.z; hx = b.x; hy = b.y; hz = b.z;
  }
  warp_minmax3(lx, ly, lz, hx, hy, hz);
  if (lane == 0) {
    const int rem = child_count - parent * kFanout;
    const int cnt = rem < kFanout ? rem : kFanout;
    reinterpret_cast<float4*>(out)[2 * (size_t)parent] = make_float4(lx, ly, lz, __int_as_float(cnt));
    reinterpret_cast<float4*>(out)[2 * (size_t)parent + 1] = make_float4(hx, hy, hz, 0.f);
  }
}

// ---- k-d refinement of Morton runs --------------------------------------------------------------
// A run = 1024 consecutive Morton-sorted points = the 32 leaves under one level-1 node. Consecutive
// Z-curve points make ragged leaves (elongated, sometimes straddling a curve jump); inside a run we
// therefore re-order by 5 levels of exact median splits along the widest axis of each segment, which
// turns the run into a small balanced k-d tree whose 32-point leaves are compact boxes. One CTA per
// run, everything in shared memory: per level a segment-bbox reduction, then a bitonic sort of
// 64-bit keys (segment | ordered coordinate | slot). This plays the role of the reference's per-level
// sort + updateTags (cukd/builder_thrust.h:175-235,486-499) but only below the 1024-point level.
constexpr int RF_RUN = 1024;
constexpr int RF_THREADS = 512;
constexpr int RF_LEVELS = 5;     // 1024 -> 32

__global__ void __launch_bounds__(RF_THREADS)
k_refine_runs(Float4* __restrict__ pts, int64_t n) {
  __shared__ float4 P[RF_RUN];
  __shared__ unsigned long long keys[RF_RUN];
  __shared__ unsigned segbox[32][6];
  const int tid = threadIdx.x, lane = tid & 31;
  const int64_t base = (int64_t)blockIdx.x * RF_RUN;
  const int nrun = (int)((n - base) < RF_RUN ? (n - base) : RF_RUN);
  float4* gp = reinterpret_cast<float4*>(pts) + base;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int t = tid + h * RF_THREADS;
    P[t] = (t < nrun) ? gp[t] : make_float4(INFINITY, INFINITY, INFINITY, 0.f);
  }
  __syncthreads();

  for (int level = 0; level < RF_LEVELS; ++level) {
    const int seg_shift = 10 - level;            // segment size = 1 << seg_shift
    const int nseg = 1 << level;
    if (tid < nseg * 6) segbox[tid / 6][tid % 6] = (tid % 6) < 3 ? 0xffffffffu : 0u;
    __syncthreads();
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int t = tid + h * RF_THREADS;
      const bool ok = t < nrun;                   // padding always stays at the tail of the run
      const float4 p = P[t];
      float lx = ok ? p.x : INFINITY, ly = ok ? p.y : INFINITY, lz = ok ? p.z : INFINITY;
      float hx = ok ? p.x : -INFINITY, hy = ok ? p.y : -INFINITY, hz = ok ? p.z : -INFINITY;
      warp_minmax3(lx, ly, lz, hx, hy, hz);       // a warp's 32 slots lie in one segment (size >= 64)
      if (lane == 0 && lx <= hx) {
        unsigned* b = segbox[t >> seg_shift];
        atomicMin(&b[0], f2ord(lx)); atomicMin(&b[1], f2ord(ly)); atomicMin(&b[2], f2ord(lz));
        atomicMax(&b[3], f2ord(hx)); atomicMax(&b[4], f2ord(hy)); atomicMax(&b[5], f2ord(hz));
      }
    }
    __syncthreads();
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int t = tid + h * RF_THREADS;
      const int seg = t >> seg_shift;
      const unsigned* b = segbox[seg];
      const float ex = ord2f(b[3]) - ord2f(b[0]), ey = ord2f(b[4]) - ord2f(b[1]), ez = ord2f(b[5]) - ord2f(b[2]);
      const int axis = (ex >= ey && ex >= ez) ? 0 : (ey >= ez ? 1 : 2);
      const float4 p = P[t];
      const float c = axis == 0 ? p.x : (axis == 1 ? p.y : p.z);
      const unsigned oc = (t < nrun) ? f2ord(c) : 0xffffffffu;
      keys[t] = ((unsigned long long)seg << 42) | ((unsigned long long)oc << 10) | (unsigned long long)t;
    }
    __syncthreads();
    // Bitonic sort of every segment. Compare-exchange steps whose partner distance is >= 64 go through
    // shared memory (one __syncthreads each); all steps with distance <= 32 of a merge run in registers:
    // a warp owns 64 consecutive keys (lane holds e0 = 64w + lane and e1 = e0 + 32), distance 32 is the
    // thread's own pair, distances 16..1 are shfl_xor partners. 35 barriers per run instead of 185.
    const int seg_size = 1 << seg_shift;
    const int e0 = ((tid >> 5) << 6) + lane, e1 = e0 + 32;
    auto cx = [](unsigned long long& x, unsigned long long y, bool keep_min) {
      const bool y_smaller = y < x;
      if (y_smaller == keep_min) x = y;
    };
    auto in_warp = [&](int size_from, int size_to) {
      // merges of sizes size_from..size_to, restricted to distances <= 32 (size_from > 64: only the tail)
      unsigned long long a = keys[e0], b = keys[e1];
      for (int size = size_from; size <= size_to; size <<= 1) {
        const bool asc0 = (size == seg_size) || ((e0 & size) == 0);
        const bool asc1 = (size == seg_size) || ((e1 & size) == 0);
        int stride = size >> 1;
        if (stride > 32) stride = 32;
        if (stride == 32) {                      // partner of e0 is e1: both in this thread, e0 is the lower index
          const unsigned long long lo = a < b ? a : b, hi = a < b ? b : a;
          a = asc0 ? lo : hi;
          b = asc0 ? hi : lo;                    // (e0 & size) == (e1 & size) for size >= 64
          stride = 16;
        }
        for (; stride > 0; stride >>= 1) {
          const unsigned long long pa = __shfl_xor_sync(0xffffffffu, a, stride);
          const unsigned long long pb = __shfl_xor_sync(0xffffffffu, b, stride);
          const bool lower = (lane & stride) == 0;          // this lane holds the lower index of the pair
          cx(a, pa, lower == asc0);
          cx(b, pb, lower == asc1);
        }
      }
      keys[e0] = a;
      keys[e1] = b;
    };
    in_warp(2, 64 < seg_size ? 64 : seg_size);
    __syncthreads();
    for (int size = 128; size <= seg_size; size <<= 1) {
      for (int stride = size >> 1; stride >= 64; stride >>= 1) {
        const int i = 2 * tid - (tid & (stride - 1));
        const int j = i + stride;
        const bool asc = (size == seg_size) || ((i & size) == 0);
        const unsigned long long a = keys[i], b = keys[j];
        if ((a > b) == asc) { keys[i] = b; keys[j] = a; }
        __syncthreads();
      }
      in_warp(size, size);
      __syncthreads();
    }
    float4 moved[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) moved[h] = P[(int)(keys[tid + h * RF_THREADS] & 1023ull)];
    __syncthreads();
#pragma unroll
    for (int h = 0; h < 2; ++h) P[tid + h * RF_THREADS] = moved[h];
    __syncthreads();
  }
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int t = tid + h * RF_THREADS;
    if (t < nrun) gp[t] = P[t];
  }
}

__global__ void __launch_bounds__(256)
k_scatter_results(const Float4* __restrict__ pts, const float* __restrict__ res, int64_t n,
                  uint32_t id_offset, float* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t id = __float_as_uint(pts[i].w) - id_offset;
  out[id] = res[i];
}

inline int grid_for(int64_t n, int block) { return (int)div_up(n, block); }

}  // namespace

int bounds_init(unsigned* d_bounds_ord, cudaStream_t stream) {
  k_bounds_init<<<1, 32, 0, stream>>>(d_bounds_ord);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

int bounds_accumulate(const float* d_xyz, PointLayout lay, int64_t n, unsigned* d_bounds_ord, cudaStream_t stream) {
  if (n <= 0) return 0;
  int64_t blocks = div_up(n, 256 * 8);
  if (blocks > 148 * 16) blocks = 148 * 16;
  k_bounds<<<(int)blocks, 256, 0, stream>>>(d_xyz, lay, n, d_bounds_ord);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

int bounds_decode(const unsigned* d_bounds_ord, float* d_bounds6, cudaStream_t stream) {
  k_bounds_decode<<<1, 32, 0, stream>>>(d_bounds_ord, d_bounds6);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

int morton_keys(const float* d_xyz, PointLayout lay, int64_t n, const float* d_bounds6, uint64_t* keys,
                uint32_t* vals, cudaStream_t stream) {
  if (n <= 0) return 0;
  k_morton<<<grid_for(n, 256), 256, 0, stream>>>(d_xyz, lay, n, d_bounds6, keys, vals);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

// Set *flag if any cell of the key's leading (64 - shift) bits holds more than 1024 points: in sorted order that is
// the same prefix 1024 places apart.
__global__ void __launch_bounds__(256)
k_heavy_cells(const uint64_t* __restrict__ keys, int64_t n, int shift, unsigned* __restrict__ flag) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i + 1024 < n && (keys[i] >> shift) == (keys[i + 1024] >> shift)) *flag = 1u;
}

int heavy_cells(const uint64_t* sorted_keys, int64_t n, int shift, unsigned* d_flag, cudaStream_t stream) {
  KNN_CUDA_TRY(cudaMemsetAsync(d_flag, 0, sizeof(unsigned), stream));
  if (n <= 1024) return 0;
  k_heavy_cells<<<grid_for(n - 1024, 256), 256, 0, stream>>>(sorted_keys, n, shift, d_flag);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

int gather_points(const float* d_xyz, PointLayout lay, const uint32_t* order, const uint32_t* ids, uint32_t id_offset,
                  int64_t n, Float4* sorted, cudaStream_t stream) {
  if (n <= 0) return 0;
  k_gather<<<grid_for(n, 256), 256, 0, stream>>>(d_xyz, lay, order, ids, id_offset, n, sorted);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

int refine_runs(Float4* sorted, int64_t n, cudaStream_t stream) {
  if (n <= kGroup) return 0;
  k_refine_runs<<<(int)div_up(n, RF_RUN), RF_THREADS, 0, stream>>>(sorted, n);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

int build_leaf_boxes(const Float4* sorted, int64_t n, Box* level0, cudaStream_t stream) {
  if (n <= 0) return 0;
  const int num_leaves = (int)div_up(n, kGroup);
  k_leaf_boxes<<<grid_for(num_leaves, 8), 256, 0, stream>>>(sorted, n, num_leaves, level0);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

int build_upper_boxes(const Box* child, int child_count, Box* parent, cudaStream_t stream) {
  const int parent_count = (int)div_up(child_count, kFanout);
  k_upper_boxes<<<grid_for(parent_count, 8), 256, 0, stream>>>(child, child_count, parent_count, parent);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

int scatter_results(const Float4* sorted_pts, const float* sorted_result, int64_t n,
                    uint32_t id_offset, float* out, cudaStream_t stream) {
  if (n <= 0) return 0;
  k_scatter_results<<<grid_for(n, 256), 256, 0, stream>>>(sorted_pts, sorted_result, n, id_offset, out);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

}  // namespace b200knn
