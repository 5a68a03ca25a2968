// brute_tile.cu — SURVEY.md section 8, row 10: leaf tiles on the 5th-generation tensor cores, measured.
//
// The question: the query kernel spends a quarter of its instructions computing 32 x 32 squared distances
// per (group, leaf) on the CUDA cores (packed fp32x2) only to turn them into "is it below this query's
// k-th distance" bits. Can tcgen05 do that part cheaper? A tensor-core distance can never be the reported
// value (tf32 operands, re-associated |q|^2 + |p|^2 - 2 q.p), but it can be a CONSERVATIVE FILTER: with an
// explicit error bound E, "d2_tc - E < threshold" keeps every true candidate, and only the survivors need
// the exact fp32 expression.
//
// This file is that experiment as a stand-alone micro-kernel pair over the same inputs:
//   brute_tile_cuda   the product's phase 1: warp = 32 queries, leaf broadcast from shared memory,
//                     FADD2/FMUL2/FFMA2, exact; counts candidates below each query's threshold.
//   brute_tile_tc     CTA = 128 queries. A = 128 x 8 tf32 rows (-2q', |q'|^2 hi, lo, 1, 1, 0), B = 128 x 8
//                     rows (p', 1, 1, |p'|^2 hi, lo, 0) — q' = q - c, p' = p - c re-centred in fp32 by the
//                     prep kernel so that the products are small; both arrive by TMA
//                     (cp.async.bulk.tensor.2d, two 16-byte-wide boxes per tile = the K-major no-swizzle
//                     core-matrix layout), one tcgen05.mma kind::tf32 (M 128, N 128, K 8) per 4 leaves writes
//                     D = |q'|^2 + |p'|^2 - 2 q'.p' into TMEM, tcgen05.commit -> mbarrier, four warps read
//                     their 32 lanes back with tcgen05.ld and count entries below threshold + E.
// tools/brute_tile_bench.py times both on the same box and checks that the tensor-core filter never
// drops a true candidate; tests/test_gpu_brute_tile.py keeps that check in the GPU test tier. The verdict
// (profiles/README.md) decides whether the product kernel gets a tensor-core leaf path.
//
// No counterpart in the reference (all of its kernels are scalar CUDA-core code).
#include <cuda.h>
#include "common.cuh"
#include "kernels.h"
#include "traverse.cuh"
#include "../include/b200knn.h"

namespace b200knn {
namespace {

constexpr int kTileQ = 128;          // queries per CTA (MMA M)
constexpr int kTileP = 128;          // points per MMA (N): 4 leaves; 128 TMEM columns per CTA, so 4 CTAs share an SM
constexpr int kK = 8;                // tf32 MMA K
constexpr float kErrRel = 1.f / 256.f;   // 2^-8: two truncated tf32 operands per product (2^-10 each), doubled for margin

// ---------------------------------------------------------------- operand preparation (fp32, exact re-centring)
__global__ void __launch_bounds__(256)
bt_prep_kernel(const float* __restrict__ xyz, int64_t n, float cx, float cy, float cz, int is_query,
               float* __restrict__ rows8, float* __restrict__ norm) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float x = xyz[3 * i] - cx, y = xyz[3 * i + 1] - cy, z = xyz[3 * i + 2] - cz;
  const float s = __fmaf_rn(z, z, __fmaf_rn(x, x, __fmul_rn(y, y)));
  // hi = the part of s a tf32 operand keeps (10 explicit mantissa bits), lo = the rest (again truncated by the MMA,
  // leaving an error of 2^-20 s)
  const float hi = __uint_as_float(__float_as_uint(s) & 0xffffe000u);
  const float lo = s - hi;
  float4 a, b;
  if (is_query) { a = make_float4(-2.f * x, -2.f * y, -2.f * z, hi); b = make_float4(lo, 1.f, 1.f, 0.f); }
  else          { a = make_float4(x, y, z, 1.f);                    b = make_float4(1.f, hi, lo, 0.f); }
  reinterpret_cast<float4*>(rows8)[2 * i] = a;
  reinterpret_cast<float4*>(rows8)[2 * i + 1] = b;
  norm[i] = sqrtf(s);
}

// ---------------------------------------------------------------- CUDA-core tile (the product's phase 1)
__global__ void __launch_bounds__(128)
bt_cuda_kernel(const float* __restrict__ q_xyz, int64_t nq, const float* __restrict__ p_xyz, int np,
               const float* __restrict__ thr, unsigned* __restrict__ counts, unsigned* __restrict__ masks) {
  __shared__ float soa[4][96];
  const int w = threadIdx.x >> 5, lane = lane_id();
  const int64_t group = (int64_t)blockIdx.x * 4 + w;
  const int64_t qi = group * 32 + lane;
  const bool valid = qi < nq;
  const float qx = valid ? q_xyz[3 * qi] : 0.f, qy = valid ? q_xyz[3 * qi + 1] : 0.f, qz = valid ? q_xyz[3 * qi + 2] : 0.f;
  const float top0 = valid ? thr[qi] : -INFINITY;
  const unsigned long long qx2 = pack2(qx, qx), qy2 = pack2(qy, qy), qz2 = pack2(qz, qz);
  const unsigned long long* sx = reinterpret_cast<const unsigned long long*>(soa[w]);
  const unsigned long long* sy = reinterpret_cast<const unsigned long long*>(soa[w] + 32);
  const unsigned long long* sz = reinterpret_cast<const unsigned long long*>(soa[w] + 64);
  unsigned count = 0;
  for (int leaf = 0; leaf * 32 < np; ++leaf) {
    const int pi = leaf * 32 + lane;
    soa[w][lane] = pi < np ? p_xyz[3 * pi] : INFINITY;
    soa[w][32 + lane] = pi < np ? p_xyz[3 * pi + 1] : INFINITY;
    soa[w][64 + lane] = pi < np ? p_xyz[3 * pi + 2] : INFINITY;
    __syncwarp();
    unsigned mask = 0;
#pragma unroll
    for (int jj = 0; jj < 16; ++jj) {
      const unsigned long long dx = sub2(qx2, sx[jj]), dy = sub2(qy2, sy[jj]), dz = sub2(qz2, sz[jj]);
      const unsigned long long d2 = fma2(dz, dz, fma2(dx, dx, mul2(dy, dy)));
      mask |= (__uint_as_float((unsigned)(d2 & 0xffffffffull)) < top0) ? (1u << (2 * jj)) : 0u;
      mask |= (__uint_as_float((unsigned)(d2 >> 32)) < top0) ? (2u << (2 * jj)) : 0u;
    }
    count += __popc(mask);
    if (masks != nullptr && valid) masks[qi * ((np + 31) / 32) + leaf] = mask;
    __syncwarp();
  }
  if (valid) counts[qi] = count;
}

// ---------------------------------------------------------------- tcgen05 tile
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// Bounded wait: a mis-programmed pipeline must end in a trap, never in a hung GPU.
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  unsigned done = 0;
  for (unsigned spins = 0; !done; ++spins) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    if (spins > (1u << 24)) __trap();
  }
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, int c0, int c1, unsigned long long* bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
               ::"r"(smem_u32(smem_dst)), "l"(map), "r"(c0), "r"(c1), "r"(smem_u32(bar)) : "memory");
}
// K-major, no swizzle: 8-row x 16-byte core matrices, 128 bytes each; SBO between 8-row groups, LBO between the two
// 16-byte halves of K (cute::UMMA::SmemDescriptor, version 1).
__device__ __forceinline__ unsigned long long umma_desc(const void* smem, unsigned lbo_bytes, unsigned sbo_bytes) {
  unsigned long long d = 0;
  d |= (unsigned long long)((smem_u32(smem) & 0x3ffffu) >> 4);
  d |= (unsigned long long)((lbo_bytes >> 4) & 0x3fffu) << 16;
  d |= (unsigned long long)((sbo_bytes >> 4) & 0x3fffu) << 32;
  d |= 1ull << 46;                       // descriptor version (sm_100)
  return d;
}

__global__ void __launch_bounds__(128)
bt_tc_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_p, int64_t nq, int np,
             const float* __restrict__ thr, const float* __restrict__ q_norm, float p_norm_max,
             unsigned* __restrict__ counts, unsigned* __restrict__ masks) {
  __shared__ __align__(128) float a_tile[2][kTileQ][4];      // K halves of A: [k half][row][4 floats]
  __shared__ __align__(128) float b_tile[2][kTileP][4];
  __shared__ __align__(8) unsigned long long bar_full, bar_mma;
  __shared__ unsigned tmem_base_slot;
  const int warp = threadIdx.x >> 5, lane = lane_id();

  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_slot)), "n"(kTileP));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  if (threadIdx.x == 0) {
    mbar_init(&bar_full, 1);
    mbar_init(&bar_mma, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const unsigned tmem_base = tmem_base_slot;

  // instruction descriptor: D fp32, A and B tf32, both K-major, N = kTileP, M = 128 (cute::UMMA::InstrDescriptor)
  const unsigned idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((unsigned)(kTileP >> 3) << 17) | ((unsigned)(kTileQ >> 4) << 24);
  const unsigned long long a_desc = umma_desc(a_tile, kTileQ * 16, 128);
  const unsigned long long b_desc = umma_desc(b_tile, kTileP * 16, 128);

  unsigned phase_full = 0, phase_mma = 0;
  const int64_t tiles = (nq + kTileQ - 1) / kTileQ;
  const int batches = (np + kTileP - 1) / kTileP;
  for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const int64_t qi = tile * kTileQ + threadIdx.x;
    const bool valid = qi < nq;
    // conservative threshold: true d2 < top  =>  d2_tc < top + E, E = 2^-8 * 2|q'||p'| + 2^-18 (|q'|^2 + |p'|^2)
    const float qn = valid ? q_norm[qi] : 0.f;
    const float bound = valid ? thr[qi] + kErrRel * 2.f * qn * p_norm_max + (qn * qn + p_norm_max * p_norm_max) * (1.f / 262144.f)
                              : -INFINITY;
    unsigned count = 0;
    for (int b = 0; b < batches; ++b) {
      if (threadIdx.x == 0) {
        const unsigned bytes = (b == 0 ? (unsigned)sizeof(a_tile) : 0u) + (unsigned)sizeof(b_tile);
        mbar_expect_tx(&bar_full, bytes);
        if (b == 0) {
          tma_load_2d(&a_tile[0][0][0], &map_q, 0, (int)(tile * kTileQ), &bar_full);
          tma_load_2d(&a_tile[1][0][0], &map_q, 4, (int)(tile * kTileQ), &bar_full);
        }
        tma_load_2d(&b_tile[0][0][0], &map_p, 0, b * kTileP, &bar_full);
        tma_load_2d(&b_tile[1][0][0], &map_p, 4, b * kTileP, &bar_full);
        mbar_wait(&bar_full, phase_full);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        // one thread issues the MMA for the whole CTA; D = A * B^T, no accumulation
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                     ::"r"(tmem_base), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(0u) : "memory");
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                     ::"r"(smem_u32(&bar_mma)) : "memory");
      }
      phase_full ^= 1u;
      mbar_wait(&bar_mma, phase_mma);
      phase_mma ^= 1u;
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      __syncwarp();                       // tcgen05.ld is .sync.aligned: the warp's lanes left their wait loops separately
      // warp w owns TMEM lanes 32w..32w+31 = its 32 queries; chunks of 32 columns = leaves
      const int cols_here = np - b * kTileP < kTileP ? np - b * kTileP : kTileP;
#pragma unroll 1
      for (int c = 0; c < kTileP; c += 32) {
        unsigned v[32];
        const unsigned taddr = tmem_base + ((unsigned)(warp * 32) << 16) + (unsigned)c;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                     "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                     "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                       "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                       "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                       "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                     : "r"(taddr));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        unsigned mask = 0;
#pragma unroll
        for (int j = 0; j < 32; ++j) mask |= (__uint_as_float(v[j]) < bound) ? (1u << j) : 0u;
        const int live = cols_here - c;                       // columns of this chunk that are real points
        mask &= live >= 32 ? 0xffffffffu : (live <= 0 ? 0u : ((1u << live) - 1u));
        count += __popc(mask);
        if (masks != nullptr && valid && c < cols_here) masks[qi * ((np + 31) / 32) + (b * kTileP + c) / 32] = mask;
      }
      // nobody may still be reading TMEM or the B tile when the next batch overwrites them
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncthreads();
    }
    if (valid) counts[qi] = count;
  }
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kTileP));
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int make_rows8_map(CUtensorMap* map, float* rows8, int64_t rows, int box_rows) {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  KNN_CUDA_TRY(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
  KNN_REQUIRE(fn != nullptr && qres == cudaDriverEntryPointSuccess, "brute_tile: cuTensorMapEncodeTiled is not available");
  const cuuint64_t dims[2] = {(cuuint64_t)kK, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)kK * sizeof(float)};
  const cuuint32_t box[2] = {4u, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1u, 1u};
  const CUresult r = reinterpret_cast<EncodeTiledFn>(fn)(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, rows8, dims, strides, box, estr,
                                                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                                         CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  KNN_REQUIRE(r == CUDA_SUCCESS, "brute_tile: cuTensorMapEncodeTiled failed");
  return 0;
}

}  // namespace
}  // namespace b200knn

using namespace b200knn;

extern "C" {

// rows8: [n][8] floats, norm: [n] floats (both device, caller-allocated). Centre = what both sides are re-centred on.
int b200knn_brute_tile_prep(const float* d_xyz, int64_t n, const float centre[3], int is_query, float* d_rows8,
                            float* d_norm, void* stream) {
  KNN_REQUIRE(d_xyz && d_rows8 && d_norm && centre && n > 0, "brute_tile_prep: bad arguments");
  bt_prep_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(d_xyz, n, centre[0], centre[1], centre[2],
                                                                               is_query, d_rows8, d_norm);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

int b200knn_brute_tile_cuda(const float* d_q_xyz, int64_t nq, const float* d_p_xyz, int np, const float* d_thr,
                            unsigned* d_counts, unsigned* d_masks_or_null, void* stream) {
  KNN_REQUIRE(d_q_xyz && d_p_xyz && d_thr && d_counts && nq > 0 && np > 0, "brute_tile_cuda: bad arguments");
  bt_cuda_kernel<<<(unsigned)((nq + 127) / 128), 128, 0, (cudaStream_t)stream>>>(d_q_xyz, nq, d_p_xyz, np, d_thr, d_counts, d_masks_or_null);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

// d_q_rows8 / d_p_rows8 from b200knn_brute_tile_prep; p_norm_max = max of the points' norms (error bound).
int b200knn_brute_tile_tc(float* d_q_rows8, const float* d_q_norm, int64_t nq, float* d_p_rows8, int np, float p_norm_max,
                          const float* d_thr, unsigned* d_counts, unsigned* d_masks_or_null, int num_ctas, void* stream) {
  KNN_REQUIRE(d_q_rows8 && d_q_norm && d_p_rows8 && d_thr && d_counts && nq > 0 && np > 0, "brute_tile_tc: bad arguments");
  CUtensorMap map_q, map_p;
  int rc;
  if ((rc = make_rows8_map(&map_q, d_q_rows8, nq, kTileQ))) return rc;
  if ((rc = make_rows8_map(&map_p, d_p_rows8, np, kTileP))) return rc;
  const int64_t tiles = (nq + kTileQ - 1) / kTileQ;
  const int grid = (int)(tiles < num_ctas ? tiles : num_ctas);
  bt_tc_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(map_q, map_p, nq, np, d_thr, d_q_norm, p_norm_max, d_counts, d_masks_or_null);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

}  // extern "C"
