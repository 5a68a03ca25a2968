// index_internal.h — layout of the opaque b200knn_index, shared by api.cu and dist.cu.
#pragma once
#include "../include/b200knn.h"
#include "kernels.h"

using namespace b200knn;

namespace b200knn { int frontier_capacity(); }   // api.cu: 128, or B200KNN_QUEUE_CAP (a test knob)

struct b200knn_index {
  int device = 0;
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

