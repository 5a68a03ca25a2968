// mpi_shim.cpp — implementation of baseline/mpi_shim/mpi.h (see there): just enough MPI to run the
// UNMODIFIED reference on an image that has no MPI installation.
// Ranks are plain processes started by baseline/run_reference.py with SHIM_RANK / SHIM_SIZE / SHIM_DIR.
//
// Transport, chosen so that it does not distort the reference's timing:
//   * device buffers (the k-d trees the reference rotates between ranks): rendezvous over CUDA IPC —
//     the sender publishes an IPC handle of its buffer, the receiver maps it and pulls with one
//     device-to-device cudaMemcpy (NVLink between GPUs, on-device for two ranks of one GPU), then
//     acknowledges; the send request completes on that acknowledgement. This is what a CUDA-aware
//     MPI does at best. (NCCL is not usable here: the reference needs several ranks per GPU.)
//   * host buffers (4..24-byte counts, boxes, votes): small files in $SHIM_DIR (/dev/shm).
// The shim also records wall-clock marks (first MPI_Comm_size, every MPI_Barrier, MPI_Finalize) into
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
#include <map>
#include <string>
#include <vector>

namespace {
int g_rank = 0, g_size = 1;
std::string g_dir;
std::vector<long> g_send_seq, g_recv_seq;
long g_coll_epoch = 0;
double g_t_first = -1, g_t_first_barrier = -1, g_t_last_barrier = -1;

struct Meta { int ipc; int pad; unsigned long long bytes; cudaIpcMemHandle_t handle; };
struct Req { bool recv; void* buf; size_t bytes; int peer; long seq; bool done; bool wait_ack; };
std::vector<Req> g_reqs;
std::map<const void*, cudaIpcMemHandle_t> g_my_handles;
std::map<std::string, void*> g_opened;   // handle bytes -> mapped pointer

double now() { timespec ts; clock_gettime(CLOCK_REALTIME, &ts); return ts.tv_sec + 1e-9 * ts.tv_nsec; }
size_t type_size(MPI_Datatype t) { return t == MPI_BYTE ? 1 : 4; }
void die(const char* what) { fprintf(stderr, "mpi_shim[%d]: %s\n", g_rank, what); abort(); }
bool is_device(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeDevice;
}
void write_file(const std::string& name, const void* a, size_t na, const void* b = nullptr, size_t nb = 0) {
  const std::string tmp = name + ".tmp";
  FILE* f = fopen(tmp.c_str(), "wb");
  if (!f) die("cannot create message file");
  if (na && fwrite(a, 1, na, f) != na) die("short write");
  if (nb && fwrite(b, 1, nb, f) != nb) die("short write");
  fclose(f);
  rename(tmp.c_str(), name.c_str());
}
void wait_for(const std::string& name) {
  struct stat st;
  long spins = 0;
  while (stat(name.c_str(), &st) != 0) usleep(spins++ < 2000 ? 20 : 200);
}
void read_file_blocking(const std::string& name, void* a, size_t na, void* b, size_t nb, bool remove_after) {
  wait_for(name);
  FILE* f = fopen(name.c_str(), "rb");
  if (!f) die("cannot open message file");
  if (na && fread(a, 1, na, f) != na) die("short read");
  if (nb && fread(b, 1, nb, f) != nb) die("short read (payload)");
  fclose(f);
  if (remove_after) unlink(name.c_str());
}
std::string msg_name(const char* kind, int src, int dst, long seq) {
  char b[256]; snprintf(b, sizeof b, "%s/%s_%d_%d_%ld", g_dir.c_str(), kind, src, dst, seq); return b;
}
void allgather_bytes(const void* send, void* recv, size_t bytes) {
  const long e = g_coll_epoch++;
  char b[256];
  snprintf(b, sizeof b, "%s/c_%ld_%d", g_dir.c_str(), e, g_rank);
  write_file(b, send, bytes);
  for (int r = 0; r < g_size; ++r) {
    snprintf(b, sizeof b, "%s/c_%ld_%d", g_dir.c_str(), e, r);
    read_file_blocking(b, (char*)recv + (size_t)r * bytes, bytes, nullptr, 0, false);
  }
}
void dump_timing(double t_final) {
  char b[256]; snprintf(b, sizeof b, "%s/timing_%d", g_dir.c_str(), g_rank);
  char txt[256];
  const int n = snprintf(txt, sizeof txt, "%.6f %.6f %.6f %.6f\n", g_t_first, g_t_first_barrier, g_t_last_barrier, t_final);
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
  const long seq = g_send_seq[dest]++;
  Meta m;
  memset(&m, 0, sizeof m);
  m.bytes = bytes;
  bool wait_ack = false;
  if (bytes && is_device(buf)) {
    auto it = g_my_handles.find(buf);
    if (it == g_my_handles.end()) {
      cudaIpcMemHandle_t h;
      if (cudaIpcGetMemHandle(&h, const_cast<void*>(buf)) != cudaSuccess) die("cudaIpcGetMemHandle failed");
      it = g_my_handles.emplace(buf, h).first;
    }
    if (cudaDeviceSynchronize() != cudaSuccess) die("device sync before send failed");
    m.ipc = 1;
    m.handle = it->second;
    write_file(msg_name("m", g_rank, dest, seq), &m, sizeof m);
    wait_ack = true;
  } else {
    write_file(msg_name("m", g_rank, dest, seq), &m, sizeof m, buf, bytes);
  }
  g_reqs.push_back({false, nullptr, bytes, dest, seq, !wait_ack, wait_ack});
  *req = (int)g_reqs.size() - 1;
  return MPI_SUCCESS;
}
int MPI_Irecv(void* buf, int count, MPI_Datatype type, int source, int, MPI_Comm, MPI_Request* req) {
  g_reqs.push_back({true, buf, (size_t)count * type_size(type), source, g_recv_seq[source]++, false, false});
  *req = (int)g_reqs.size() - 1;
  return MPI_SUCCESS;
}
int MPI_Waitall(int count, MPI_Request* reqs, MPI_Status*) {
  // receives first (they only need the peer's Isend to have been posted), then send acknowledgements
  for (int pass = 0; pass < 2; ++pass) {
    for (int i = 0; i < count; ++i) {
      Req& q = g_reqs[reqs[i]];
      if (q.done || q.recv != (pass == 0)) continue;
      if (q.recv) {
        const std::string name = msg_name("m", q.peer, g_rank, q.seq);
        wait_for(name);
        FILE* f = fopen(name.c_str(), "rb");
        Meta m;
        if (!f || fread(&m, 1, sizeof m, f) != sizeof m) die("bad message header");
        if (m.bytes > q.bytes) die("message longer than the receive buffer");
        if (m.ipc) {
          fclose(f);
          const std::string key((const char*)&m.handle, sizeof m.handle);
          auto it = g_opened.find(key);
          if (it == g_opened.end()) {
            void* p = nullptr;
            if (cudaIpcOpenMemHandle(&p, m.handle, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) die("cudaIpcOpenMemHandle failed");
            it = g_opened.emplace(key, p).first;
          }
          if (cudaMemcpy(q.buf, it->second, m.bytes, cudaMemcpyDefault) != cudaSuccess) die("peer copy failed");
          char ok = 1;
          write_file(msg_name("a", q.peer, g_rank, q.seq), &ok, 1);
        } else {
          if (m.bytes) {
            if (is_device(q.buf)) {
              std::vector<char> host(m.bytes);
              if (fread(host.data(), 1, m.bytes, f) != m.bytes) die("short payload");
              if (cudaMemcpy(q.buf, host.data(), m.bytes, cudaMemcpyHostToDevice) != cudaSuccess) die("H2D failed");
            } else if (fread(q.buf, 1, m.bytes, f) != m.bytes) {
              die("short payload");
            }
          }
          fclose(f);
        }
        unlink(name.c_str());
      } else if (q.wait_ack) {
        const std::string ack = msg_name("a", g_rank, q.peer, q.seq);
        wait_for(ack);
        unlink(ack.c_str());
      }
      q.done = true;
    }
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
    if (type == MPI_INT) {
      int m = ((int*)all.data())[i];
      for (int r = 1; r < g_size; ++r) { const int v = ((int*)(all.data() + r * bytes))[i]; if (v > m) m = v; }
      ((int*)recv)[i] = m;
    } else {
      float m = ((float*)all.data())[i];
      for (int r = 1; r < g_size; ++r) { const float v = ((float*)(all.data() + r * bytes))[i]; if (v > m) m = v; }
      ((float*)recv)[i] = m;
    }
  }
  return MPI_SUCCESS;
}
int MPI_Allgather(const void* send, int scount, MPI_Datatype stype, void* recv, int, MPI_Datatype, MPI_Comm) {
  allgather_bytes(send, recv, (size_t)scount * type_size(stype));
  return MPI_SUCCESS;
}
}
