// knn_math.h — arithmetic shared by device kernels, the CPU mirror engine and the oracle.
//
// Everything that decides *which value ends up in the output file* lives here so that
// host and device evaluate the same expression, operation for operation.
//
// Parity notes (reference = /root/reference):
//   * canonical squared distance  = fmaf(dz,dz, fmaf(dx,dx, dy*dy)),  d = query - point.
//     This is what nvcc 12.9 emits for the reference's dot(d,d)
//     (submodules/cudaKDTree/cukd/cukd-math.h:109-110,210-213; SASS in SURVEY.md A.1).
//   * box lower bounds use the same expression on per-axis gaps, which makes them
//     conservative under round-to-nearest (every step is monotone in each |d_axis|),
//     so culling never changes the k-th distance (cf. the plane test in
//     cukd/traverse-stack-free.h:80-87 and box distance in cukd/box.h:96-108).
//   * candidate list = bounded max-heap of floats only. The reference keeps
//     (dist<<32|id) uint64 entries (cukd/knn.h:311-389) but only ever outputs the
//     k-th distance, for which the id tie-breaker is irrelevant.
#pragma once
#include <stdint.h>
#include <math.h>

#if defined(__CUDACC__)
#define KNN_HD __host__ __device__ __forceinline__
#else
#define KNN_HD inline
#endif

namespace b200knn {

constexpr int kGroup = 32;          // queries per group == points per leaf == warp width
constexpr int kFanout = 32;         // children per inner node of the wide tree
constexpr int kMaxLevels = 6;       // 32^6 leaves * 32 points: far beyond 2^31 points per GPU
constexpr int kHeapArity = 4;       // d-ary max-heap: half the dependent-load depth of a binary heap

struct alignas(16) Float4 { float x, y, z, w; };
struct alignas(16) Box { float lox, loy, loz, cnt_bits; float hix, hiy, hiz, pad; };  // 32 bytes

// The one distance expression. Explicit fmaf: host builds use -ffp-contract=off, device
// code cannot re-associate explicit fma intrinsics.
KNN_HD float dist2(float qx, float qy, float qz, float px, float py, float pz) {
  const float dx = qx - px, dy = qy - py, dz = qz - pz;
#if defined(__CUDA_ARCH__)
  return __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
#else
  return fmaf(dz, dz, fmaf(dx, dx, dy * dy));
#endif
}

KNN_HD float gap3(float gx, float gy, float gz) {
#if defined(__CUDA_ARCH__)
  return __fmaf_rn(gz, gz, __fmaf_rn(gx, gx, __fmul_rn(gy, gy)));
#else
  return fmaf(gz, gz, fmaf(gx, gx, gy * gy));
#endif
}

// Lower bound of dist2(q, p) over all p inside [lo,hi].
KNN_HD float point_box_lb(float qx, float qy, float qz,
                          float lox, float loy, float loz, float hix, float hiy, float hiz) {
  const float gx = fmaxf(0.f, fmaxf(lox - qx, qx - hix));
  const float gy = fmaxf(0.f, fmaxf(loy - qy, qy - hiy));
  const float gz = fmaxf(0.f, fmaxf(loz - qz, qz - hiz));
  return gap3(gx, gy, gz);
}

// Lower bound of dist2(q, p) over all q in box A and p in box B.
KNN_HD float box_box_lb(float alx, float aly, float alz, float ahx, float ahy, float ahz,
                        float blx, float bly, float blz, float bhx, float bhy, float bhz) {
  const float gx = fmaxf(0.f, fmaxf(blx - ahx, alx - bhx));
  const float gy = fmaxf(0.f, fmaxf(bly - ahy, aly - bhy));
  const float gz = fmaxf(0.f, fmaxf(blz - ahz, alz - bhz));
  return gap3(gx, gy, gz);
}

// ---- 63-bit Morton code (21 bits per axis). Ordering only: never affects results. ----
KNN_HD uint64_t spread21(uint64_t v) {
  v &= 0x1fffffull;
  v = (v | (v << 32)) & 0x1f00000000ffffull;
  v = (v | (v << 16)) & 0x1f0000ff0000ffull;
  v = (v | (v << 8))  & 0x100f00f00f00f00full;
  v = (v | (v << 4))  & 0x10c30c30c30c30c3ull;
  v = (v | (v << 2))  & 0x1249249249249249ull;
  return v;
}

KNN_HD uint32_t quantize21(float v, float lo, float inv_extent) {
  float t = (v - lo) * inv_extent;            // in [0,1] for in-box points
  t = fminf(fmaxf(t, 0.f), 1.f);
  float s = t * 2097151.f;                    // 2^21 - 1
  uint32_t q = (uint32_t)s;
  return q > 2097151u ? 2097151u : q;
}

KNN_HD uint64_t morton63(float x, float y, float z,
                         float lox, float loy, float loz, float ivx, float ivy, float ivz) {
  return (spread21(quantize21(x, lox, ivx)) << 2) |
         (spread21(quantize21(y, loy, ivy)) << 1) |
          spread21(quantize21(z, loz, ivz));
}

// ---- bounded max-heap of k floats, strided storage (stride 32 on the GPU: one column per lane).
// replace_top: precondition e < h[0]. Returns the new top.
template <typename Ptr>
KNN_HD float heap_replace_top(Ptr h, int stride, int k, float e) {
  int i = 0;
  float newtop = e;
  bool first = true;
  for (;;) {
    const int c0 = kHeapArity * i + 1;
    if (c0 >= k) break;
    float best = h[c0 * stride];
    int bi = c0;
#pragma unroll
    for (int t = 1; t < kHeapArity; ++t) {
      const int c = c0 + t;
      if (c < k) {
        const float v = h[c * stride];
        if (v > best) { best = v; bi = c; }
      }
    }
    if (!(best > e)) break;
    h[i * stride] = best;
    if (first) { newtop = best; first = false; }
    i = bi;
  }
  h[i * stride] = e;
  return newtop;
}

// Final value written to the output file (reference: extractFinalResult, unorderedDataVariant.cu:89-103).
KNN_HD float finalize_kth(float top_d2) {
  return isinf(top_d2) ? top_d2 : sqrtf(top_d2);
}

// Sentinel the heap is filled with: cutoff radius squared (+inf when -r is absent),
// cf. FlexHeapCandidateList ctor, cukd/knn.h:329-339.
KNN_HD float sentinel_d2(float max_radius) {
  return isinf(max_radius) ? max_radius : max_radius * max_radius;
}

inline int64_t div_up(int64_t a, int64_t b) { return (a + b - 1) / b; }

}  // namespace b200knn
