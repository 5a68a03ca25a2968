// mpi_shim.cpp — file-based implementation of baseline/mpi_shim/mpi.h (see there).
// Ranks are plain processes started by baseline/run_reference.py with SHIM_RANK / SHIM_SIZE / SHIM_DIR.
// It also records wall-clock marks (first MPI_Comm_size, every MPI_Barrier, MPI_Finalize) into
// $SHIM_DIR/timing_<rank> because the reference has no timers of its own.
#include "mpi.h"
#include <cuda_runtime.h>
#include <fcntl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>
#include <string>
#include <vector>

namespace {
int g_rank = 0, g_size = 1;
std::string g_dir;
std::vector<long> g_send_seq, g_recv_seq;
long g_coll_epoch = 0;
double g_t_first = -1, g_t_first_barrier = -1, g_t_last_barrier = -1;

struct Req { bool recv; void* buf; size_t bytes; int peer; long seq; bool done; };
std::vector<Req> g_reqs;

double now() { timespec ts; clock_gettime(CLOCK_REALTIME, &ts); return ts.tv_sec + 1e-9 * ts.tv_nsec; }
size_t type_size(MPI_Datatype t) { return t == MPI_BYTE ? 1 : 4; }
bool is_device(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeDevice;
}
void write_file(const std::string& name, const void* data, size_t bytes) {
  const std::string tmp = name + ".tmp";
  FILE* f = fopen(tmp.c_str(), "wb");
  if (!f) { perror("shim fopen"); abort(); }
  if (bytes && fwrite(data, 1, bytes, f) != bytes) { perror("shim fwrite"); abort(); }
  fclose(f);
  rename(tmp.c_str(), name.c_str());
}
void read_file_blocking(const std::string& name, void* data, size_t bytes, bool remove_after) {
  struct stat st;
  long spins = 0;
  while (stat(name.c_str(), &st) != 0) { usleep(spins++ < 1000 ? 50 : 500); }
  FILE* f = fopen(name.c_str(), "rb");
  if (!f) { perror("shim fopen r"); abort(); }
  if (bytes && fread(data, 1, bytes, f) != bytes) { fprintf(stderr, "shim: short read %s\n", name.c_str()); abort(); }
  fclose(f);
  if (remove_after) unlink(name.c_str());
}
std::string msg_name(int src, int dst, long seq) {
  char b[256]; snprintf(b, sizeof b, "%s/m_%d_%d_%ld", g_dir.c_str(), src, dst, seq); return b;
}
void allgather_bytes(const void* send, void* recv, size_t bytes) {
  const long e = g_coll_epoch++;
  char b[256];
  snprintf(b, sizeof b, "%s/c_%ld_%d", g_dir.c_str(), e, g_rank);
  write_file(b, send, bytes);
  for (int r = 0; r < g_size; ++r) {
    snprintf(b, sizeof b, "%s/c_%ld_%d", g_dir.c_str(), e, r);
    read_file_blocking(b, (char*)recv + (size_t)r * bytes, bytes, false);
  }
}
void dump_timing(double t_final) {
  char b[256]; snprintf(b, sizeof b, "%s/timing_%d", g_dir.c_str(), g_rank);
  char txt[256];
  int n = snprintf(txt, sizeof txt, "%.6f %.6f %.6f %.6f\n", g_t_first, g_t_first_barrier, g_t_last_barrier, t_final);
  write_file(b, txt, (size_t)n);
}
}  // namespace

extern "C" {
int MPI_Init(int*, char***) {
  const char* r = getenv("SHIM_RANK"); const char* s = getenv("SHIM_SIZE"); const char* d = getenv("SHIM_DIR");
  g_rank = r ? atoi(r) : 0; g_size = s ? atoi(s) : 1; g_dir = d ? d : "/tmp";
  g_send_seq.assign(g_size, 0); g_recv_seq.assign(g_size, 0);
  return MPI_SUCCESS;
}
int MPI_Finalize(void) { dump_timing(now()); return MPI_SUCCESS; }
int MPI_Comm_rank(MPI_Comm, int* rank) { *rank = g_rank; return MPI_SUCCESS; }
int MPI_Comm_size(MPI_Comm, int* size) { if (g_t_first < 0) g_t_first = now(); *size = g_size; return MPI_SUCCESS; }
int MPI_Isend(const void* buf, int count, MPI_Datatype type, int dest, int, MPI_Comm, MPI_Request* req) {
  const size_t bytes = (size_t)count * type_size(type);
  const std::string name = msg_name(g_rank, dest, g_send_seq[dest]++);
  if (bytes && is_device(buf)) {
    std::vector<char> host(bytes);
    if (cudaMemcpy(host.data(), buf, bytes, cudaMemcpyDeviceToHost) != cudaSuccess) { fprintf(stderr, "shim: D2H failed\n"); abort(); }
    write_file(name, host.data(), bytes);
  } else {
    write_file(name, buf, bytes);
  }
  g_reqs.push_back({false, nullptr, bytes, dest, 0, true});
  *req = (int)g_reqs.size() - 1;
  return MPI_SUCCESS;
}
int MPI_Irecv(void* buf, int count, MPI_Datatype type, int source, int, MPI_Comm, MPI_Request* req) {
  g_reqs.push_back({true, buf, (size_t)count * type_size(type), source, g_recv_seq[source]++, false});
  *req = (int)g_reqs.size() - 1;
  return MPI_SUCCESS;
}
int MPI_Waitall(int count, MPI_Request* reqs, MPI_Status*) {
  for (int i = 0; i < count; ++i) {
    Req& q = g_reqs[reqs[i]];
    if (q.done) continue;
    const std::string name = msg_name(q.peer, g_rank, q.seq);
    if (q.bytes && is_device(q.buf)) {
      std::vector<char> host(q.bytes);
      read_file_blocking(name, host.data(), q.bytes, true);
      if (cudaMemcpy(q.buf, host.data(), q.bytes, cudaMemcpyHostToDevice) != cudaSuccess) { fprintf(stderr, "shim: H2D failed\n"); abort(); }
    } else {
      read_file_blocking(name, q.buf, q.bytes, true);
    }
    q.done = true;
  }
  return MPI_SUCCESS;
}
int MPI_Barrier(MPI_Comm) {
  char dummy = 0;
  std::vector<char> all(g_size);
  allgather_bytes(&dummy, all.data(), 1);
  const double t = now();
  if (g_t_first_barrier < 0) g_t_first_barrier = t;
  g_t_last_barrier = t;
  return MPI_SUCCESS;
}
int MPI_Allreduce(const void* send, void* recv, int count, MPI_Datatype type, MPI_Op, MPI_Comm) {
  const size_t bytes = (size_t)count * type_size(type);
  std::vector<char> all(bytes * g_size);
  allgather_bytes(send, all.data(), bytes);
  for (int i = 0; i < count; ++i) {
    if (type == MPI_INT) { int m = ((int*)all.data())[i]; for (int r = 1; r < g_size; ++r) { int v = ((int*)(all.data() + r * bytes))[i]; if (v > m) m = v; } ((int*)recv)[i] = m; }
    else { float m = ((float*)all.data())[i]; for (int r = 1; r < g_size; ++r) { float v = ((float*)(all.data() + r * bytes))[i]; if (v > m) m = v; } ((float*)recv)[i] = m; }
  }
  return MPI_SUCCESS;
}
int MPI_Allgather(const void* send, int scount, MPI_Datatype stype, void* recv, int, MPI_Datatype, MPI_Comm) {
  allgather_bytes(send, recv, (size_t)scount * type_size(stype));
  return MPI_SUCCESS;
}
}
