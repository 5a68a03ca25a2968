/* sample_knn.c — the C ABI from a plain C program: k-th-neighbour distances of a random cloud on one
 * GPU, checked against the library's own brute-force oracle on a few queries.
 *
 *   make examples && build/sample_knn [numPoints] [k]
 *
 * Counterpart of the reference library's sample program (submodules/cudaKDTree/sample.cu), for the
 * operation the reference's two executables are built around (runQuery + extractFinalResult,
 * /root/reference/unorderedDataVariant.cu:75-103). */
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "b200knn.h"

#define CHECK_CUDA(call)                                                              \
  do {                                                                                \
    cudaError_t e_ = (call);                                                          \
    if (e_ != cudaSuccess) {                                                          \
      fprintf(stderr, "%s failed: %s\n", #call, cudaGetErrorString(e_));              \
      return 1;                                                                       \
    }                                                                                 \
  } while (0)

#define CHECK_KNN(call)                                                               \
  do {                                                                                \
    if ((call) != 0) {                                                                \
      fprintf(stderr, "%s failed: %s\n", #call, b200knn_last_error());                \
      return 1;                                                                       \
    }                                                                                 \
  } while (0)

int main(int argc, char** argv) {
  const long long n = argc > 1 ? atoll(argv[1]) : 1000000;
  const int k = argc > 2 ? atoi(argv[2]) : 100;
  if (n < 1 || k < 1) {
    fprintf(stderr, "usage: %s [numPoints >= 1] [k >= 1]\n", argv[0]);
    return 1;
  }

  float* h_xyz = (float*)malloc((size_t)n * 3 * sizeof(float));
  float* h_out = (float*)malloc((size_t)n * sizeof(float));
  if (!h_xyz || !h_out) return 1;
  srand(42);
  for (long long i = 0; i < 3 * n; ++i) h_xyz[i] = (float)rand() / (float)RAND_MAX;

  float *d_xyz = NULL, *d_out = NULL;
  cudaStream_t stream;
  cudaEvent_t e0, e1, e2;
  CHECK_CUDA(cudaStreamCreate(&stream));
  CHECK_CUDA(cudaEventCreate(&e0));
  CHECK_CUDA(cudaEventCreate(&e1));
  CHECK_CUDA(cudaEventCreate(&e2));
  CHECK_CUDA(cudaMalloc((void**)&d_xyz, (size_t)n * 3 * sizeof(float)));
  CHECK_CUDA(cudaMalloc((void**)&d_out, (size_t)n * sizeof(float)));
  CHECK_CUDA(cudaMemcpyAsync(d_xyz, h_xyz, (size_t)n * 3 * sizeof(float), cudaMemcpyHostToDevice, stream));

  b200knn_index* index = NULL;
  for (int rep = 0; rep < 2; ++rep) { /* first pass warms up the allocator and the kernels */
    CHECK_CUDA(cudaEventRecord(e0, stream));
    CHECK_KNN(b200knn_index_create(d_xyz, n, NULL, 0, stream, &index));
    CHECK_CUDA(cudaEventRecord(e1, stream));
    CHECK_KNN(b200knn_knn_self(index, k, INFINITY, d_out, stream));
    CHECK_CUDA(cudaEventRecord(e2, stream));
    CHECK_CUDA(cudaStreamSynchronize(stream));
    if (rep == 0) b200knn_index_destroy(index);
  }
  float ms_build = 0.f, ms_query = 0.f;
  CHECK_CUDA(cudaEventElapsedTime(&ms_build, e0, e1));
  CHECK_CUDA(cudaEventElapsedTime(&ms_query, e1, e2));
  CHECK_CUDA(cudaMemcpy(h_out, d_out, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost));
  printf("%lld points, k = %d: index %.2f ms (%d levels), all-kNN %.2f ms = %.1f M queries/s\n", n, k, ms_build,
         b200knn_index_num_levels(index), ms_query, 1e-3 * (double)n / ms_query);

  /* the oracle: all n distances per query on the host, same fp32 expression, must agree bit for bit */
  const int checks = n < 64 ? (int)n : 64;
  float* want = (float*)malloc((size_t)checks * sizeof(float));
  float* queries = (float*)malloc((size_t)checks * 3 * sizeof(float));
  for (int i = 0; i < checks; ++i) memcpy(queries + 3 * i, h_xyz + 3 * ((n / checks) * i), 3 * sizeof(float));
  CHECK_KNN(b200knn_cpu_bruteforce(h_xyz, n, queries, checks, k, INFINITY, want, 0));
  int bad = 0;
  for (int i = 0; i < checks; ++i) bad += memcmp(&want[i], &h_out[(n / checks) * i], sizeof(float)) != 0;
  printf("checked %d queries against the brute-force oracle: %s\n", checks, bad ? "MISMATCH" : "bit-identical");

  b200knn_index_destroy(index);
  cudaFree(d_xyz);
  cudaFree(d_out);
  free(h_xyz); free(h_out); free(want); free(queries);
  return bad != 0;
}
