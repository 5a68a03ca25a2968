// noalloc.cu — b200knn_knn_self_noalloc: the all-kNN pipeline for small inputs with every byte of
// working memory taken from one fixed device array that is part of the library image.
//
// Who this is for. Some users of this library run where asking the CUDA driver (or any allocator
// built on it) for device memory is off the table:
//   * inside a host application that claimed the whole GPU up front for its own arena — an inference
//     server or a simulation code with a fixed memory budget — where a late cudaMalloc either fails
//     or steals from a budget somebody else has already planned;
//   * in control or acquisition loops that must not block: cudaMalloc/cudaFree may take driver locks
//     and synchronise the device for an unbounded time;
//   * in plug-ins called from a framework callback whose contract forbids allocator calls after
//     initialisation.
// The normal path (b200knn_index_create + b200knn_knn_self) allocates the sorted copy of the points,
// the box tree, sort buffers and counters on every call. This variant places all of that inside
// g_scratch below and calls no cudaMalloc, cudaMallocAsync, cudaFree, cudaFreeAsync or any other
// allocating routine. The caller still owns and passes the input points and the output array.
//
// Limits and rules
//   * n <= B200KNN_NOALLOC_MAX_POINTS (262144). Larger inputs are rejected with an error, never
//     truncated. k must fit the shared-memory candidate lists (k <= 1729); the normal path's fallback for
//     larger k keeps its lists in a global scratch buffer that it allocates, so it is refused here — before
//     anything is launched.
//   * The array (about 11.9 MB) is held by every GPU that loads the library, until the process exits.
//   * The first call on each device looks up the array's device address with cudaGetSymbolAddress
//     (the name g_scratch in host code is only a stand-in and must never be handed to a kernel).
//     That first call does the look-up, so make it before any stream capture is started.
//   * All calls on a device share the one array and it keeps whatever the previous call left there.
//     Everything that has to start from zero (bounds accumulators, digit totals, the task counter) is
//     cleared inside each call, on the caller's stream. Two calls whose GPU work overlaps in time —
//     different streams, or different threads — race on the array and produce garbage: keep calls on
//     one stream, or order them yourself.
//
// This entry point has no counterpart in the reference.
#include <string.h>
#include <atomic>
#include "common.cuh"
#include "kernels.h"
#include "../include/b200knn.h"

namespace {

using namespace b200knn;

constexpr int64_t kMaxPoints = B200KNN_NOALLOC_MAX_POINTS;
constexpr size_t a256(size_t v) { return (v + 255) / 256 * 256; }

// Fixed carve-up of the scratch array (offsets in bytes, all 256-byte aligned).
constexpr int64_t kMaxLeaves = kMaxPoints / kGroup;                       // 8192
constexpr int64_t kLvl1 = (kMaxLeaves + kFanout - 1) / kFanout;           // 256
constexpr int64_t kLvl2 = (kLvl1 + kFanout - 1) / kFanout;                // 8
constexpr int64_t kSortTiles = (kMaxPoints + 4095) / 4096;                // radix sort tile = 4096 keys
constexpr size_t kOffBoundsOrd = 0;
constexpr size_t kOffBounds6 = kOffBoundsOrd + a256(6 * sizeof(unsigned));
constexpr size_t kOffCounter = kOffBounds6 + a256(6 * sizeof(float));
constexpr size_t kOffKeys = kOffCounter + a256(sizeof(unsigned long long));
constexpr size_t kOffKeysTmp = kOffKeys + a256(kMaxPoints * sizeof(uint64_t));
constexpr size_t kOffVals = kOffKeysTmp + a256(kMaxPoints * sizeof(uint64_t));
constexpr size_t kOffValsTmp = kOffVals + a256(kMaxPoints * sizeof(uint32_t));
constexpr size_t kOffSort = kOffValsTmp + a256(kMaxPoints * sizeof(uint32_t));
constexpr size_t kOffPts = kOffSort + a256((256 + 256 * kSortTiles) * sizeof(uint32_t));
constexpr size_t kOffResult = kOffPts + a256(kMaxPoints * sizeof(Float4));
constexpr size_t kOffBox0 = kOffResult + a256(kMaxPoints * sizeof(float));
constexpr size_t kOffBox1 = kOffBox0 + a256(kMaxLeaves * sizeof(Box));
constexpr size_t kOffBox2 = kOffBox1 + a256(kLvl1 * sizeof(Box));
constexpr size_t kScratchBytes = kOffBox2 + a256(kLvl2 * sizeof(Box));

static __device__ __align__(256) unsigned char g_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
// device address of g_scratch, per device. Several host threads may make their first call at once: they all
// look up the same address, so the race is benign, and the slot is atomic so that it is also well defined.
std::atomic<unsigned char*> g_scratch_addr[kMaxDevices];

// Set-up, once per device, with that device current.
int scratch_address(unsigned char** out) {
  int dev = 0;
  KNN_CUDA_TRY(cudaGetDevice(&dev));
  KNN_REQUIRE(dev >= 0 && dev < kMaxDevices, "knn_self_noalloc: device ordinal out of range");
  unsigned char* known = g_scratch_addr[dev].load(std::memory_order_acquire);
  if (known == nullptr) {
    void* ptr = nullptr;
    KNN_CUDA_TRY(cudaGetSymbolAddress(&ptr, g_scratch));
    KNN_REQUIRE(ptr != nullptr, "knn_self_noalloc: cudaGetSymbolAddress returned a null address");
    known = static_cast<unsigned char*>(ptr);
    g_scratch_addr[dev].store(known, std::memory_order_release);
  }
  *out = known;
  return 0;
}

}  // namespace

extern "C" {

int64_t b200knn_noalloc_max_points(void) { return kMaxPoints; }
int64_t b200knn_noalloc_scratch_bytes(void) { return (int64_t)kScratchBytes; }

int b200knn_knn_self_noalloc(const float* d_xyz, int64_t n, int k, float max_radius, float* d_out,
                             void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  KNN_REQUIRE(n >= 0, "knn_self_noalloc: n must be >= 0");
  KNN_REQUIRE(n <= kMaxPoints,
              "knn_self_noalloc: n exceeds B200KNN_NOALLOC_MAX_POINTS (262144); use b200knn_index_create + "
              "b200knn_knn_self for larger inputs");
  KNN_REQUIRE(k >= 1, "knn_self_noalloc: k must be >= 1");
  KNN_REQUIRE(ring_pick_warps_per_cta(k) >= 1,
              "knn_self_noalloc: k does not fit the shared-memory candidate lists (k <= 1729); larger k needs a "
              "scratch allocation, use b200knn_index_create + b200knn_knn_self");
  if (n == 0) return 0;
  KNN_REQUIRE(d_xyz != nullptr && d_out != nullptr, "knn_self_noalloc: NULL point or output buffer");

  unsigned char* base = nullptr;
  int rc;
  if ((rc = scratch_address(&base))) return rc;
  unsigned* bounds_ord = reinterpret_cast<unsigned*>(base + kOffBoundsOrd);
  float* bounds6 = reinterpret_cast<float*>(base + kOffBounds6);
  unsigned long long* counter = reinterpret_cast<unsigned long long*>(base + kOffCounter);
  uint64_t* keys = reinterpret_cast<uint64_t*>(base + kOffKeys);
  uint64_t* keys_tmp = reinterpret_cast<uint64_t*>(base + kOffKeysTmp);
  uint32_t* vals = reinterpret_cast<uint32_t*>(base + kOffVals);
  uint32_t* vals_tmp = reinterpret_cast<uint32_t*>(base + kOffValsTmp);
  uint32_t* sort_scratch = reinterpret_cast<uint32_t*>(base + kOffSort);
  Float4* pts = reinterpret_cast<Float4*>(base + kOffPts);
  float* result_sorted = reinterpret_cast<float*>(base + kOffResult);
  Box* box_level[3] = {reinterpret_cast<Box*>(base + kOffBox0), reinterpret_cast<Box*>(base + kOffBox1),
                       reinterpret_cast<Box*>(base + kOffBox2)};

  // ---- index build, same kernels as b200knn_index_create. The accumulators that must start from a
  // known state are reset here on every call: bounds (bounds_init), digit totals (inside the sort),
  // task counter (inside ring_launch).
  if ((rc = bounds_init(bounds_ord, stream))) return rc;
  if ((rc = bounds_accumulate(d_xyz, 3, n, bounds_ord, stream))) return rc;
  if ((rc = bounds_decode(bounds_ord, bounds6, stream))) return rc;
  if ((rc = morton_keys(d_xyz, 3, n, bounds6, keys, vals, stream))) return rc;
  int in_tmp = 0;
  if ((rc = radix_sort_u64(keys, vals, keys_tmp, vals_tmp, sort_scratch, n, 0, 64, stream, &in_tmp))) return rc;
  if ((rc = gather_points(d_xyz, 3, in_tmp ? vals_tmp : vals, nullptr, 0, n, pts, stream))) return rc;
  if ((rc = refine_runs(pts, n, stream))) return rc;

  TreeView tree;
  memset(&tree, 0, sizeof(tree));
  tree.pts = pts;
  tree.n = (int)n;
  int cnt = (int)div_up(n, kGroup);
  int levels = 0;
  for (;;) {
    KNN_REQUIRE(levels < 3, "knn_self_noalloc: internal error (tree deeper than the scratch layout)");
    tree.boxes[levels] = box_level[levels];
    tree.count[levels] = cnt;
    ++levels;
    if (cnt <= kFanout) break;
    cnt = (int)div_up(cnt, kFanout);
  }
  tree.levels = levels;
  tree.queue_cap = 128;
  tree.cull_scale = 1.f;
  tree.node_budget = -1;
  if ((rc = build_leaf_boxes(pts, n, box_level[0], stream))) return rc;
  for (int l = 1; l < levels; ++l)
    if ((rc = build_upper_boxes(box_level[l - 1], tree.count[l - 1], box_level[l], stream))) return rc;

  // ---- query: the same persistent kernel as b200knn_knn_self (single rank, one step)
  int dev = 0, num_sms = 0;
  KNN_CUDA_TRY(cudaGetDevice(&dev));
  KNN_CUDA_TRY(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev));
  RingParams p;
  memset(&p, 0, sizeof(p));
  p.tree = tree;
  p.k = k;
  p.sentinel = sentinel_d2(max_radius);
  p.R = 1;
  p.rank = 0;
  p.groups_per_chunk = tree.count[0];
  p.num_chunks = 1;
  p.step_begin = 0;
  p.step_end = 1;
  p.use_flags = 0;
  p.home_queries = pts;
  p.home_n = n;
  p.home_is_tree = 1;
  p.result_sorted = result_sorted;
  p.task_counter = counter;
  p.stats = nullptr;
  if ((rc = ring_launch(p, num_sms, stream))) return rc;
  return scatter_results(pts, result_sorted, n, 0, d_out, stream);
}

}  // extern "C"
