// verify.cu — an independent check of k-th-neighbour distances at ANY size: no tree, no heap, no sort.
//
// r is the distance to the k-th nearest point of q  <=>  #{p : dist(q,p) < r} < k <= #{p : dist(q,p) <= r}
// (with the library's own fp32 distance expression and sqrt). So a handful of sampled queries can be
// verified against ALL points with one counting pass — 64 queries x 256 Mi points is a few milliseconds
// of GPU time, where the CPU oracle would need minutes. Role of the reference's verifyKNN
// (cudaKDTree/testing/floatN-knn-and-fcp.cu:292-335) for sizes no brute-force k-selection can reach;
// across ranks the counts simply add up.
#include "common.cuh"
#include "kernels.h"
#include "../include/b200knn.h"

namespace b200knn {
namespace {

constexpr int kVerifyMaxQueries = 64;

__global__ void __launch_bounds__(256)
count_within_kernel(const float* __restrict__ xyz, int64_t n, const float* __restrict__ queries, int nq,
                    const float* __restrict__ radius, unsigned long long* __restrict__ less,
                    unsigned long long* __restrict__ less_equal) {
  __shared__ float q[kVerifyMaxQueries][4];
  __shared__ unsigned long long cnt[2][kVerifyMaxQueries];
  const int lane = lane_id();
  for (int j = threadIdx.x; j < nq; j += blockDim.x) {
    q[j][0] = queries[3 * j]; q[j][1] = queries[3 * j + 1]; q[j][2] = queries[3 * j + 2]; q[j][3] = radius[j];
    cnt[0][j] = cnt[1][j] = 0ull;
  }
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t rounds = (n + stride - 1) / stride;            // every thread of a warp takes part in every vote
  for (int64_t it = 0; it < rounds; ++it) {
    const int64_t i = it * stride + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool have = i < n;
    float px = 0.f, py = 0.f, pz = 0.f;
    if (have) { px = xyz[3 * i]; py = xyz[3 * i + 1]; pz = xyz[3 * i + 2]; }
    for (int j = 0; j < nq; ++j) {
      const float d = finalize_kth(dist2(q[j][0], q[j][1], q[j][2], px, py, pz));
      const unsigned lt = __ballot_sync(kFullMask, have && d < q[j][3]);
      const unsigned le = __ballot_sync(kFullMask, have && d <= q[j][3]);
      if (lane == 0) {
        if (lt) atomicAdd(&cnt[0][j], (unsigned long long)__popc(lt));
        if (le) atomicAdd(&cnt[1][j], (unsigned long long)__popc(le));
      }
    }
  }
  __syncthreads();
  for (int j = threadIdx.x; j < nq; j += blockDim.x) {
    if (cnt[0][j]) atomicAdd(less + j, cnt[0][j]);
    if (cnt[1][j]) atomicAdd(less_equal + j, cnt[1][j]);
  }
}

}  // namespace
}  // namespace b200knn

using namespace b200knn;

extern "C" int b200knn_count_within(const float* d_xyz, int64_t n, const float* d_queries, int nq, const float* d_radius,
                                    uint64_t* d_less, uint64_t* d_less_equal, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  KNN_REQUIRE(n >= 0 && (n == 0 || d_xyz), "count_within: bad point buffer");
  KNN_REQUIRE(nq >= 1 && nq <= kVerifyMaxQueries, "count_within: 1..64 queries per call");
  KNN_REQUIRE(d_queries && d_radius && d_less && d_less_equal, "count_within: NULL argument");
  KNN_CUDA_TRY(cudaMemsetAsync(d_less, 0, (size_t)nq * sizeof(uint64_t), stream));
  KNN_CUDA_TRY(cudaMemsetAsync(d_less_equal, 0, (size_t)nq * sizeof(uint64_t), stream));
  if (n == 0) return 0;
  int device = 0, sms = 148;
  KNN_CUDA_TRY(cudaGetDevice(&device));
  KNN_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
  int64_t blocks = (n + 255) / 256;
  if (blocks > (int64_t)sms * 8) blocks = (int64_t)sms * 8;
  count_within_kernel<<<(int)blocks, 256, 0, stream>>>(d_xyz, n, d_queries, nq, d_radius,
                                                       reinterpret_cast<unsigned long long*>(d_less),
                                                       reinterpret_cast<unsigned long long*>(d_less_equal));
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}
