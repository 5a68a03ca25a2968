// backend.cpp — GPU and CPU compute paths shared by the two executables (see knn_apps.hpp).
#include <cuda_runtime.h>
#include <algorithm>
#include "knn_apps.hpp"

namespace knnapp {

#define APP_CUDA(expr)                                                                           \
  do {                                                                                           \
    cudaError_t e_ = (expr);                                                                     \
    if (e_ != cudaSuccess) throw std::runtime_error(std::string(#expr) + ": " + cudaGetErrorString(e_)); \
  } while (0)

bool gpu_available() {
  const char* b = getenv("B200KNN_BACKEND");
  if (b && std::string(b) == "cpu") return false;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return false; }
  return n > 0;
}

void select_gpu(const Bootstrap& bs, const Options& o) {
  int n = 1;
  APP_CUDA(cudaGetDeviceCount(&n));
  const int dev = o.gpus_per_node > 0 ? bs.rank() % o.gpus_per_node : bs.rank() % n;
  if (o.gpus_per_node > 0)
    std::cout << "#" << bs.rank() << "/" << bs.size() << " setting active GPU #" << dev << std::endl;
  APP_CUDA(cudaSetDevice(dev));
}

std::vector<float> run_gpu(Bootstrap& bs, const std::vector<float>& pts, int64_t global_begin, const Options& o,
                           int mode) {
  const int64_t n = (int64_t)pts.size() / 3;
  float *d_xyz = nullptr, *d_out = nullptr;
  APP_CUDA(cudaMalloc(&d_xyz, std::max<size_t>(pts.size(), 3) * sizeof(float)));
  APP_CUDA(cudaMalloc(&d_out, std::max<int64_t>(n, 1) * sizeof(float)));
  if (n) APP_CUDA(cudaMemcpy(d_xyz, pts.data(), pts.size() * sizeof(float), cudaMemcpyHostToDevice));
  b200knn_dist* eng = nullptr;
  if (b200knn_dist_create(bs.c_struct(), &eng)) throw std::runtime_error(b200knn_last_error());
  b200knn_dist_report rep;
  const int rc = b200knn_dist_knn(eng, d_xyz, n, global_begin, o.k, o.max_radius, mode, 0, d_out, 0, &rep);
  if (rc) throw std::runtime_error(std::string("b200knn_dist_knn: ") + b200knn_last_error());
  std::vector<float> out((size_t)n);
  if (n) APP_CUDA(cudaMemcpy(out.data(), d_out, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost));
  if (bs.rank() == 0)
    std::cout << "timing(rank 0): total " << rep.ms_total << " ms, redistribute " << rep.ms_redistribute << ", build "
              << rep.ms_build << ", query " << rep.ms_query << ", return " << rep.ms_return << "; groups that visited "
              << "another rank: " << rep.pending_groups << std::endl;
  b200knn_dist_destroy(eng);
  cudaFree(d_xyz);
  cudaFree(d_out);
  return out;
}

std::vector<float> run_cpu(Bootstrap& bs, const std::vector<float>& pts, const Options& o) {
  const int R = bs.size();
  const int64_t n = (int64_t)pts.size() / 3;
  std::vector<int64_t> counts(R);
  bs.allgather(&n, counts.data(), sizeof(int64_t));
  int64_t max_n = 0, total = 0;
  for (int64_t c : counts) { max_n = std::max(max_n, c); total += c; }
  std::vector<float> padded((size_t)std::max<int64_t>(max_n, 1) * 3, 0.f), gathered((size_t)std::max<int64_t>(max_n, 1) * 3 * R);
  std::copy(pts.begin(), pts.end(), padded.begin());
  bs.allgather(padded.data(), gathered.data(), (int64_t)padded.size() * sizeof(float));
  std::vector<float> all;
  all.reserve((size_t)total * 3);
  for (int r = 0; r < R; ++r)
    all.insert(all.end(), gathered.begin() + (size_t)r * padded.size(), gathered.begin() + (size_t)r * padded.size() + counts[r] * 3);
  std::vector<float> out((size_t)n);
  if (n && b200knn_cpu_knn(all.data(), total, pts.data(), n, o.k, o.max_radius, out.data(), 0))
    throw std::runtime_error(b200knn_last_error());
  return out;
}

}  // namespace knnapp
