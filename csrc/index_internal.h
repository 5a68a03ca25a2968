// index_internal.h — layout of the opaque b200knn_index, shared by api.cu and dist.cu.
#pragma once
#include "../include/b200knn.h"
#include "kernels.h"

using namespace b200knn;

namespace b200knn {
int frontier_capacity();   // api.cu: 128, or B200KNN_QUEUE_CAP (a test knob)
int index_sort_bits();     // api.cu: 40, or B200KNN_SORT_BITS

// A few words from device memory to the host WITHOUT a copy engine: a one-warp kernel stores them into a slot of
// page-locked host memory that the GPU has mapped. An in-stream cudaMemcpyAsync of 24 bytes looks harmless, but it
// queues on the device-to-host copy engine behind whatever a caller's other stream is downloading — a 1 GB result
// array of the previous step holds the next step's bounds (and with them its whole build) back by 19 ms.
// begin() enqueues on `stream`; after the caller has synchronised that stream, end() copies slot -> h_dst and
// frees the slot (the destructor frees it on error paths). Falls back to cudaMemcpyAsync when no slot is free.
class SmallRead {
 public:
  SmallRead() = default;
  SmallRead(const SmallRead&) = delete;
  SmallRead& operator=(const SmallRead&) = delete;
  ~SmallRead() { release(); }
  int begin(const void* d_src, size_t bytes, void* h_dst, cudaStream_t stream);
  void end();
 private:
  void release();
  int slot_ = -1;
  void* dst_ = nullptr;
  size_t bytes_ = 0;
};
}  // namespace b200knn

struct b200knn_index {
  int device = 0;
  int sort_bits = 64;                      // leading key bits the build sorted by (40, or 64 after the heavy-cell fallback)
  int dims = 3;                            // 2: the index was built from [n][2] points (z = 0) and takes [nq][2] queries
  int num_sms = 148;
  int64_t n = 0;
  uint32_t id_offset = 0;
  Float4* pts = nullptr;
  Box* boxes[kMaxLevels] = {};
  int count[kMaxLevels] = {};
  int levels = 0;
  unsigned* d_bounds_ord = nullptr;
  float* d_bounds6 = nullptr;
  float h_bounds[6] = {};
  float* result_sorted = nullptr;          // [groups*32]
  unsigned long long* task_counter = nullptr;
  unsigned long long* stats = nullptr;     // [4]
  bool stats_on = false;
  cudaStream_t stream = 0;                 // stream the buffers were allocated on / last used on

  TreeView view() const {
    TreeView t;
    t.pts = pts;
    for (int l = 0; l < kMaxLevels; ++l) { t.boxes[l] = boxes[l]; t.count[l] = count[l]; }
    t.levels = levels;
    t.n = (int)n;
    t.queue_cap = frontier_capacity();
    t.cull_scale = 1.f;
    t.node_budget = -1;
    return t;
  }
  int64_t groups() const { return div_up(n, kGroup); }
};

