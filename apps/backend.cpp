// backend.cpp — GPU and CPU compute paths shared by the two executables (see knn_apps.hpp).
#include <cuda_runtime.h>
#include <algorithm>
#include <fcntl.h>
#include <string.h>
#include <unistd.h>
#include <thread>
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

// One file range moved by a few threads: a read from /dev/shm or the page cache is a memcpy inside the kernel and
// one core does 2-3 GB/s of that; four do close to four times as much (3.2 GB of input: 1.5 s -> 0.4 s).
namespace {
template <class Op>
void split_io(size_t len, Op op) {
  unsigned t = std::thread::hardware_concurrency();
  t = t >= 8 ? 4 : (t >= 4 ? 2 : 1);
  if (len < ((size_t)8 << 20) * t) t = 1;
  if (t == 1) { op(0, len); return; }
  std::vector<std::thread> th;
  std::vector<std::string> err(t);
  for (unsigned i = 0; i < t; ++i) {
    const size_t lo = len * i / t / 4096 * 4096, hi = i + 1 == t ? len : len * (i + 1) / t / 4096 * 4096;
    th.emplace_back([&, i, lo, hi] { try { op(lo, hi); } catch (const std::exception& e) { err[i] = e.what(); } });
  }
  for (auto& x : th) x.join();
  for (auto& e : err) if (!e.empty()) throw std::runtime_error(e);
}
}  // namespace

void parallel_pread(int fd, char* dst, size_t len, int64_t off, const std::string& what) {
  split_io(len, [&](size_t lo, size_t hi) {
    while (lo < hi) {
      const ssize_t r = pread(fd, dst + lo, hi - lo, (off_t)(off + (int64_t)lo));
      if (r <= 0) throw std::runtime_error("short read on " + what);
      lo += (size_t)r;
    }
  });
}

void parallel_pwrite(int fd, const char* src, size_t len, int64_t off, const std::string& what) {
  split_io(len, [&](size_t lo, size_t hi) {
    while (lo < hi) {
      const ssize_t w = pwrite(fd, src + lo, hi - lo, (off_t)(off + (int64_t)lo));
      if (w <= 0) throw std::runtime_error("short write on " + what);
      lo += (size_t)w;
    }
  });
}

// File -> device -> results, for inputs of any size: the rank's slice is read in 64 MB pieces into two pinned
// buffers and each piece is on its way to the device while the next one is being read; results come back the
// same way. (The reference reads its slice into pageable memory and uploads it with one blocking cudaMemcpy,
// /root/reference/unorderedDataVariant.cu:41-63,157-160; its results reach the file through managed memory, :207-237.)
std::vector<float> run_gpu_from_file(Bootstrap& bs, const std::string& path, int64_t begin, int64_t n,
                                     const Options& o, int mode, const std::string& out_path, int64_t total) {
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
      try { parallel_pread(fd, pin[which], len, begin * 12 + (int64_t)off, path); }
      catch (...) { close(fd); throw; }
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
  // Results: pieces come back through the same two pinned buffers; the previous piece is consumed while the next one
  // travels. With an output file they go straight from the pinned buffer into the file at this rank's offset (rank 0
  // creates the file at its final size first: the same bytes as the reference's rank-ordered append,
  // /root/reference/unorderedDataVariant.cu:229-237); without one they are collected and returned.
  const bool to_file = !out_path.empty();
  std::vector<float> out(to_file ? 0 : (size_t)n);
  int ofd = -1;
  if (to_file) {
    if (bs.rank() == 0) {
      std::cout << "done all queries..." << std::endl;
      ofd = open(out_path.c_str(), O_WRONLY | O_CREAT | O_TRUNC, 0644);
      if (ofd < 0) throw std::runtime_error("cannot open output " + out_path);
      if (ftruncate(ofd, (off_t)(total * 4)) != 0) { close(ofd); throw std::runtime_error("ftruncate failed"); }
    }
    bs.barrier();
    if (bs.rank() != 0) {
      ofd = open(out_path.c_str(), O_WRONLY);
      if (ofd < 0) throw std::runtime_error("cannot open output " + out_path);
    }
  }
  {
    const size_t want = (size_t)n * 4;
    size_t prev_off = 0, prev_len = 0;
    int which = 0;
    auto consume = [&](int buf, size_t off, size_t len) {
      APP_CUDA(cudaEventSynchronize(done[buf]));
      if (to_file) parallel_pwrite(ofd, pin[buf], len, begin * 4 + (int64_t)off, out_path);
      else memcpy((char*)out.data() + off, pin[buf], len);
    };
    for (size_t off = 0; off < want; off += kPiece, which ^= 1) {
      const size_t len = std::min(kPiece, want - off);
      APP_CUDA(cudaMemcpyAsync(pin[which], (const char*)d_out + off, len, cudaMemcpyDeviceToHost, copy_stream));
      APP_CUDA(cudaEventRecord(done[which], copy_stream));
      if (prev_len) consume(which ^ 1, prev_off, prev_len);
      prev_off = off;
      prev_len = len;
    }
    if (prev_len) consume(which ^ 1, prev_off, prev_len);
  }
  if (to_file) {
    if (close(ofd) != 0) throw std::runtime_error("close failed on " + out_path);
    bs.barrier();
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
