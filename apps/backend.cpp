// backend.cpp — GPU and CPU compute paths shared by the two executables (see knn_apps.hpp).
#include <cuda_runtime.h>
#include <algorithm>
#include <fcntl.h>
#include <string.h>
#include <unistd.h>
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

// File -> device -> results, for inputs of any size: the rank's slice is read in 64 MB pieces into two pinned
// buffers and each piece is on its way to the device while the next one is being read; results come back the
// same way. (The reference reads its slice into pageable memory and uploads it with one blocking cudaMemcpy,
// /root/reference/unorderedDataVariant.cu:41-63,157-160; its results reach the file through managed memory, :207-237.)
std::vector<float> run_gpu_from_file(Bootstrap& bs, const std::string& path, int64_t begin, int64_t n,
                                     const Options& o, int mode) {
  constexpr size_t kPiece = (size_t)64 << 20;
  float *d_xyz = nullptr, *d_out = nullptr;
  char* pin[2] = {nullptr, nullptr};
  cudaStream_t copy_stream;
  cudaEvent_t done[2];
  APP_CUDA(cudaMalloc(&d_xyz, (size_t)std::max<int64_t>(n, 1) * 12));
  APP_CUDA(cudaMalloc(&d_out, (size_t)std::max<int64_t>(n, 1) * 4));
  APP_CUDA(cudaStreamCreateWithFlags(&copy_stream, cudaStreamNonBlocking));
  for (int i = 0; i < 2; ++i) {
    APP_CUDA(cudaHostAlloc(&pin[i], kPiece, cudaHostAllocDefault));
    APP_CUDA(cudaEventCreateWithFlags(&done[i], cudaEventDisableTiming));
  }
  if (n > 0) {
    const int fd = open(path.c_str(), O_RDONLY);
    if (fd < 0) throw std::runtime_error("cannot open " + path);
    const size_t want = (size_t)n * 12;
    int which = 0;
    for (size_t off = 0; off < want; off += kPiece, which ^= 1) {
      const size_t len = std::min(kPiece, want - off);
      APP_CUDA(cudaEventSynchronize(done[which]));          // the copy that last used this buffer has finished
      size_t got = 0;
      while (got < len) {
        const ssize_t r = pread(fd, pin[which] + got, len - got, (off_t)(begin * 12 + off + got));
        if (r <= 0) { close(fd); throw std::runtime_error("short read on " + path); }
        got += (size_t)r;
      }
      APP_CUDA(cudaMemcpyAsync((char*)d_xyz + off, pin[which], len, cudaMemcpyHostToDevice, copy_stream));
      APP_CUDA(cudaEventRecord(done[which], copy_stream));
    }
    close(fd);
    APP_CUDA(cudaStreamSynchronize(copy_stream));
  }
  b200knn_dist* eng = nullptr;
  if (b200knn_dist_create(bs.c_struct(), &eng)) throw std::runtime_error(b200knn_last_error());
  b200knn_dist_report rep;
  const int rc = b200knn_dist_knn(eng, d_xyz, n, begin, o.k, o.max_radius, mode, 0, d_out, 0, &rep);
  if (rc) throw std::runtime_error(std::string("b200knn_dist_knn: ") + b200knn_last_error());
  std::vector<float> out((size_t)n);
  {
    const size_t want = (size_t)n * 4;
    size_t prev_off = 0, prev_len = 0;
    int which = 0;
    for (size_t off = 0; off < want; off += kPiece, which ^= 1) {
      const size_t len = std::min(kPiece, want - off);
      APP_CUDA(cudaMemcpyAsync(pin[which], (const char*)d_out + off, len, cudaMemcpyDeviceToHost, copy_stream));
      APP_CUDA(cudaEventRecord(done[which], copy_stream));
      if (prev_len) {                                        // unpack the previous piece while this one travels
        APP_CUDA(cudaEventSynchronize(done[which ^ 1]));
        memcpy((char*)out.data() + prev_off, pin[which ^ 1], prev_len);
      }
      prev_off = off;
      prev_len = len;
    }
    if (prev_len) {
      APP_CUDA(cudaEventSynchronize(done[which ^ 1]));
      memcpy((char*)out.data() + prev_off, pin[which ^ 1], prev_len);
    }
  }
  if (bs.rank() == 0)
    std::cout << "timing(rank 0): total " << rep.ms_total << " ms, redistribute " << rep.ms_redistribute << ", build "
              << rep.ms_build << ", query " << rep.ms_query << ", return " << rep.ms_return << "; groups that visited "
              << "another rank: " << rep.pending_groups << std::endl;
  b200knn_dist_destroy(eng);
  for (int i = 0; i < 2; ++i) { cudaFreeHost(pin[i]); cudaEventDestroy(done[i]); }
  cudaStreamDestroy(copy_stream);
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
