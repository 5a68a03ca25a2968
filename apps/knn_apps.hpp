// knn_apps.hpp — what the two executables share: command line, file formats, process bootstrap.
//
// Behavioural contract = the reference's:
//   * flags -o <path|prefix> (required), -k <int >= 1> (required), -r <maxRadius>, -g <gpusPerNode>;
//     any argument not starting with '-' is the input path (last one wins); unknown flag, missing
//     input/output/k -> usage on stderr, exit code 1   (/root/reference/unorderedDataVariant.cu:114-135)
//   * input: raw float32 triples, count = fileSize/12; rank r of R reads [count*r/R, count*(r+1)/R)
//     (/root/reference/unorderedDataVariant.cu:41-63)
//   * output: float32 per point, input order.
// Divergences (documented, deliberate): ranks write their slice with pwrite at its offset instead of
// a barrier-chained append (same bytes); without -g the GPU is rank % visible-device-count instead
// of always device 0; extra flags --ring / --prepartitioned-mode select the multi-GPU strategy; a file
// list whose last line lacks '\n' is accepted (the reference silently drops that line).
//
// Process model: `knnrun -n R exe args...` forks R copies with KNNRUN_RANK / KNNRUN_SIZE /
// KNNRUN_SOCK; ranks talk to the launcher over one unix stream socket each for two collectives
// (allgather of equal-sized blobs, barrier). Started without knnrun an executable is rank 0 of 1.
// Role of MPI_Init / MPI_Comm_rank / MPI_Comm_size / MPI_Allgather / MPI_Barrier in the reference.
#pragma once
#include <errno.h>
#include <fcntl.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/stat.h>
#include <sys/un.h>
#include <unistd.h>
#include <fstream>
#include <iostream>
#include <limits>
#include <stdexcept>
#include <string>
#include <vector>
#include "b200knn.h"

namespace knnapp {

struct Options {
  float max_radius = std::numeric_limits<float>::infinity();
  int k = 0;
  int gpus_per_node = 0;
  std::string input;
  std::string output;
  int mode = -1;   // -1: program default
};

[[noreturn]] inline void usage(const char* prog, const std::string& error) {
  std::cerr << "Error: " << error << "\n\n";
  std::cerr << prog << " -k <k> [-r <maxRadius>] [-g <gpusPerNode>] <input> -o <output>\n";
  exit(error.empty() ? 0 : 1);
}

inline Options parse_args(int ac, char** av, const char* what_input, const char* what_output) {
  Options o;
  auto need_value = [&](int& i, const std::string& flag) -> const char* {
    if (i + 1 >= ac) usage(av[0], "missing value after '" + flag + "'");
    return av[++i];
  };
  for (int i = 1; i < ac; ++i) {
    const std::string arg = av[i];
    if (arg == "-o") o.output = need_value(i, arg);
    else if (arg.empty() || arg[0] != '-') o.input = arg;
    else if (arg == "-r") o.max_radius = (float)atof(need_value(i, arg));
    else if (arg == "-g") o.gpus_per_node = atoi(need_value(i, arg));
    else if (arg == "-k") o.k = atoi(need_value(i, arg));
    else if (arg == "--ring") o.mode = B200KNN_MODE_RING;
    else if (arg == "--redistribute") o.mode = B200KNN_MODE_REDISTRIBUTE;
    else if (arg == "--in-place") o.mode = B200KNN_MODE_PREPARTITIONED;
    else usage(av[0], "unknown cmdline arg '" + arg + "'");
  }
  if (o.input.empty()) usage(av[0], std::string("no ") + what_input + " specified");
  if (o.output.empty()) usage(av[0], std::string("no ") + what_output + " specified");
  if (o.k < 1) usage(av[0], "no k specified, or invalid k value");
  return o;
}

// ---- files ------------------------------------------------------------------------------------
inline int64_t count_points(const std::string& path) {
  struct stat st;
  if (stat(path.c_str(), &st) != 0) return 0;      // the reference treats a missing file as empty
  return (int64_t)(st.st_size / 12);
}

// Rank's slice of a raw float3 file. begin/total are reported like readFilePortion does.
inline std::vector<float> read_points(const std::string& path, int rank, int size, int64_t* begin_out,
                                      int64_t* total_out) {
  const int64_t total = count_points(path);
  const int64_t begin = total * rank / size, end = total * (rank + 1) / size;
  if (begin_out) *begin_out = begin;
  if (total_out) *total_out = total;
  std::vector<float> pts((size_t)(end - begin) * 3);
  if (end > begin) {
    const int fd = open(path.c_str(), O_RDONLY);
    if (fd < 0) throw std::runtime_error("cannot open " + path);
    size_t done = 0;
    const size_t want = pts.size() * sizeof(float);
    while (done < want) {
      const ssize_t r = pread(fd, (char*)pts.data() + done, want - done, (off_t)(begin * 12 + done));
      if (r <= 0) { close(fd); throw std::runtime_error("short read on " + path); }
      done += (size_t)r;
    }
    close(fd);
  }
  return pts;
}

inline std::vector<std::string> read_file_list(const std::string& path) {
  std::vector<std::string> names;
  std::ifstream in(path);
  std::string line;
  while (std::getline(in, line)) {
    while (!line.empty() && (line.back() == '\r' || line.back() == ' ')) line.pop_back();
    if (!line.empty()) names.push_back(line);
  }
  return names;
}

inline void write_at(const std::string& path, const float* values, int64_t count, int64_t begin, bool create,
                     int64_t total) {
  const int fd = open(path.c_str(), create ? (O_WRONLY | O_CREAT | O_TRUNC) : O_WRONLY, 0644);
  if (fd < 0) throw std::runtime_error("cannot open output " + path);
  if (create && total >= 0 && ftruncate(fd, (off_t)(total * 4)) != 0) { close(fd); throw std::runtime_error("ftruncate failed"); }
  size_t done = 0;
  const size_t want = (size_t)count * sizeof(float);
  while (done < want) {
    const ssize_t w = pwrite(fd, (const char*)values + done, want - done, (off_t)(begin * 4 + done));
    if (w <= 0) { close(fd); throw std::runtime_error("short write on " + path); }
    done += (size_t)w;
  }
  close(fd);
}

inline std::string per_rank_name(const std::string& prefix, int rank) {
  char suffix[64];
  snprintf(suffix, sizeof suffix, "_%06d.float", rank);
  return prefix + suffix;
}

// ---- bootstrap over the knnrun socket -----------------------------------------------------------
struct Wire { uint32_t op; uint32_t rank; uint64_t bytes; };   // op 1 = allgather, 2 = bye

inline void io_all(int fd, void* buf, size_t n, bool writing) {
  size_t done = 0;
  while (done < n) {
    const ssize_t r = writing ? ::write(fd, (char*)buf + done, n - done) : ::read(fd, (char*)buf + done, n - done);
    if (r <= 0) {
      if (r < 0 && errno == EINTR) continue;
      throw std::runtime_error("knnrun socket closed");
    }
    done += (size_t)r;
  }
}

class Bootstrap {
 public:
  Bootstrap() {
    const char* r = getenv("KNNRUN_RANK");
    const char* s = getenv("KNNRUN_SIZE");
    const char* p = getenv("KNNRUN_SOCK");
    rank_ = r ? atoi(r) : 0;
    size_ = s ? atoi(s) : 1;
    if (size_ > 1) {
      if (!p) throw std::runtime_error("KNNRUN_SOCK not set (start through knnrun)");
      fd_ = socket(AF_UNIX, SOCK_STREAM, 0);
      sockaddr_un addr;
      memset(&addr, 0, sizeof addr);
      addr.sun_family = AF_UNIX;
      strncpy(addr.sun_path, p, sizeof(addr.sun_path) - 1);
      if (connect(fd_, (sockaddr*)&addr, sizeof addr) != 0) throw std::runtime_error("cannot connect to knnrun");
    }
    c_.ctx = this;
    c_.rank = rank_;
    c_.world = size_;
    c_.allgather = &Bootstrap::c_allgather;
    c_.barrier = &Bootstrap::c_barrier;
  }
  ~Bootstrap() {
    if (fd_ >= 0) {
      Wire w{2, (uint32_t)rank_, 0};
      try { io_all(fd_, &w, sizeof w, true); } catch (...) {}
      close(fd_);
    }
  }
  int rank() const { return rank_; }
  int size() const { return size_; }
  const b200knn_bootstrap* c_struct() const { return &c_; }

  void allgather(const void* send, void* recv, int64_t bytes) {
    if (size_ == 1) { memcpy(recv, send, (size_t)bytes); return; }
    Wire w{1, (uint32_t)rank_, (uint64_t)bytes};
    io_all(fd_, &w, sizeof w, true);
    if (bytes) io_all(fd_, const_cast<void*>(send), (size_t)bytes, true);
    if (bytes) io_all(fd_, recv, (size_t)bytes * size_, false);
    else { char ack; io_all(fd_, &ack, 1, false); }
  }
  void barrier() { allgather(nullptr, nullptr, 0); }

 private:
  static int c_allgather(void* ctx, const void* send, void* recv, int64_t bytes) {
    try { static_cast<Bootstrap*>(ctx)->allgather(send, recv, bytes); return 0; } catch (...) { return 1; }
  }
  static int c_barrier(void* ctx) {
    try { static_cast<Bootstrap*>(ctx)->barrier(); return 0; } catch (...) { return 1; }
  }
  int rank_ = 0, size_ = 1, fd_ = -1;
  b200knn_bootstrap c_{};
};

// ---- compute back ends ----------------------------------------------------------------------------
bool gpu_available();
// GPU path: points -> device, b200knn_dist_knn, results -> host.
std::vector<float> run_gpu(Bootstrap& bs, const std::vector<float>& pts, int64_t global_begin, const Options& o,
                           int mode);
// The same, reading the rank's slice [begin, begin + n) of a raw float3 file straight into device memory through
// pinned double buffers (no pageable copy of the input). With `out_path` the results go from the pinned buffers
// straight into that file at float offset `begin` (a file of `total` floats that rank 0 creates; collective: two
// barriers) and the returned vector is empty; "done all queries..." is printed before the file is written.
std::vector<float> run_gpu_from_file(Bootstrap& bs, const std::string& path, int64_t begin, int64_t n,
                                     const Options& o, int mode, const std::string& out_path = std::string(),
                                     int64_t total = 0);
void parallel_pread(int fd, char* dst, size_t len, int64_t off, const std::string& what);
void parallel_pwrite(int fd, const char* src, size_t len, int64_t off, const std::string& what);
// CPU path (B200KNN_BACKEND=cpu or no CUDA device): every rank gathers all points through the
// bootstrap and answers its own queries with the host mirror engine. For plumbing tests.
std::vector<float> run_cpu(Bootstrap& bs, const std::vector<float>& pts, const Options& o);
void select_gpu(const Bootstrap& bs, const Options& o);

}  // namespace knnapp
