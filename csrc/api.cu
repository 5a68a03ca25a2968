// api.cu — C ABI glue: index lifetime, single-GPU queries, fused-ring object (peer mapping via CUDA IPC).
// Host-side only; all kernels live in radix_sort.cu / build.cu / query.cu.
//
// Role in the reference: the body of main() between "points are on the host" and "results are on
// the host" (/root/reference/unorderedDataVariant.cu:153-211), minus MPI, plus what MPI hid
// (peer mapping, flow control).
#include <string.h>
#include <atomic>
#include <mutex>
#include <vector>
#include "../include/b200knn.h"
#include "common.cuh"
#include "kernels.h"
#include "index_internal.h"

namespace b200knn {
static thread_local std::string g_last_error;
void set_last_error(const std::string& msg) { g_last_error = msg; }
const char* last_error_cstr() { return g_last_error.c_str(); }
static std::atomic<unsigned long long> g_launches{0};
void note_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

// ---- SmallRead (index_internal.h): mapped page-locked mailbox slots ------------------------------------
namespace {
constexpr int kMailSlots = 64;
constexpr size_t kMailSlotBytes = 2048;
std::mutex g_mail_mutex;
char* g_mail_host = nullptr;       // kMailSlots * kMailSlotBytes, cudaHostAllocMapped | Portable, never freed
char* g_mail_dev = nullptr;
bool g_mail_failed = false;
unsigned long long g_mail_used = 0;

__global__ void small_read_kernel(const unsigned* __restrict__ src, unsigned* __restrict__ dst, int words) {
  for (int i = threadIdx.x; i < words; i += blockDim.x) dst[i] = src[i];
  __threadfence_system();
}
}  // namespace

int SmallRead::begin(const void* d_src, size_t bytes, void* h_dst, cudaStream_t stream) {
  release();
  dst_ = h_dst;
  bytes_ = bytes;
  if (bytes == 0) return 0;
  if (bytes <= kMailSlotBytes && bytes % 4 == 0) {
    std::lock_guard<std::mutex> lock(g_mail_mutex);
    if (g_mail_host == nullptr && !g_mail_failed) {
      void *h = nullptr, *dv = nullptr;
      if (cudaHostAlloc(&h, kMailSlots * kMailSlotBytes, cudaHostAllocMapped | cudaHostAllocPortable) == cudaSuccess &&
          cudaHostGetDevicePointer(&dv, h, 0) == cudaSuccess) {
        g_mail_host = static_cast<char*>(h);
        g_mail_dev = static_cast<char*>(dv);
      } else {
        cudaGetLastError();
        g_mail_failed = true;
      }
    }
    if (g_mail_host != nullptr)
      for (int i = 0; i < kMailSlots; ++i)
        if (!(g_mail_used >> i & 1ull)) { g_mail_used |= 1ull << i; slot_ = i; break; }
  }
  if (slot_ >= 0) {
    small_read_kernel<<<1, 128, 0, stream>>>(static_cast<const unsigned*>(d_src),
                                             reinterpret_cast<unsigned*>(g_mail_dev + (size_t)slot_ * kMailSlotBytes),
                                             (int)(bytes / 4));
    note_launch();
    KNN_CUDA_TRY(cudaGetLastError());
  } else {
    KNN_CUDA_TRY(cudaMemcpyAsync(h_dst, d_src, bytes, cudaMemcpyDeviceToHost, stream));
  }
  return 0;
}

void SmallRead::end() {
  if (slot_ >= 0) memcpy(dst_, g_mail_host + (size_t)slot_ * kMailSlotBytes, bytes_);
  release();
}

void SmallRead::release() {
  if (slot_ < 0) return;
  std::lock_guard<std::mutex> lock(g_mail_mutex);
  g_mail_used &= ~(1ull << slot_);
  slot_ = -1;
}
}  // namespace b200knn

namespace {

struct RingLayout {
  int64_t heap[2], pos[2], cnt[2], flag_in[2], ack_in[2], result, bytes;
};

struct IpcBlob {
  cudaIpcMemHandle_t handle;   // 64 bytes
  RingLayout layout;
  int32_t rank, world, k, groups_per_chunk;
  int64_t pid;
};
static_assert(sizeof(IpcBlob) <= B200KNN_IPC_BLOB_BYTES, "blob too large");

inline int64_t align_up(int64_t v, int64_t a) { return (v + a - 1) / a * a; }

RingLayout make_layout(int k, int Gc, int64_t result_floats) {
  RingLayout L;
  int64_t off = 0;
  for (int b = 0; b < 2; ++b) { L.heap[b] = off; off = align_up(off + (int64_t)Gc * k * 32 * 4, 256); }
  for (int b = 0; b < 2; ++b) { L.pos[b] = off; off = align_up(off + (int64_t)Gc * 32 * 16, 256); }
  for (int b = 0; b < 2; ++b) { L.cnt[b] = off; off = align_up(off + (int64_t)Gc * 4, 256); }
  for (int b = 0; b < 2; ++b) { L.flag_in[b] = off; off = align_up(off + (int64_t)Gc * 4, 256); }
  for (int b = 0; b < 2; ++b) { L.ack_in[b] = off; off = align_up(off + (int64_t)Gc * 4, 256); }
  L.result = off; off = align_up(off + result_floats * 4, 256);
  L.bytes = off;
  return L;
}

}  // namespace

struct b200knn_ring {
  b200knn_index* idx = nullptr;
  int rank = 0, world = 1, k = 0, Gc = 0, C = 0;
  char* base = nullptr;
  RingLayout mine{};
  char* next_base = nullptr; RingLayout next_layout{};
  char* prev_base = nullptr; RingLayout prev_layout{};
  bool opened_next = false, opened_prev = false;
  unsigned seq_base = 1;
};

namespace b200knn {
// Stream-ordered allocations are served from the device's default pool; keep freed blocks cached
// instead of returning them to the driver at every synchronisation (index build/destroy per kNN
// call would otherwise pay a multi-GB cudaMalloc/cudaFree each time).
void keep_pool_memory(int device) {
  static std::atomic<unsigned> done{0};
  const unsigned bit = 1u << (device & 31);
  if (done.load() & bit) return;
  cudaMemPool_t pool;
  if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
    unsigned long long keep = ~0ull;
    cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
  }
  done.fetch_or(bit);
}

// Leading bits of the Morton key that the index build sorts by (see index_create_impl): 40, or B200KNN_SORT_BITS
// (8..64, rounded up to whole 8-bit digits).
int index_sort_bits() {
  static const int bits = [] {
    const char* e = getenv("B200KNN_SORT_BITS");
    int b = e ? atoi(e) : 40;
    if (b < 8) b = 8;
    if (b > 64) b = 64;
    return (b + 7) / 8 * 8;
  }();
  return bits;
}

// Convergent fill of fresh heaps (traverse.cuh::fill_leaf): on by default (measured 58.4 -> 54.5 ms for
// 16 M points, k = 100); B200KNN_FILL=0 switches back to sifting every candidate.
// Frontier entries per warp: all 128, unless B200KNN_QUEUE_CAP (8..128) shrinks it so that tests reach the
// purge and depth-first fallback paths with small inputs.
int frontier_capacity() {
  static const int cap = [] {
    const char* e = getenv("B200KNN_QUEUE_CAP");
    const int v = e ? atoi(e) : 128;
    return v < 8 ? 8 : (v > 128 ? 128 : v);
  }();
  return cap;
}

bool convergent_fill_enabled() {
  static const bool on = [] {
    const char* e = getenv("B200KNN_FILL");
    return e ? (e[0] != '0') : true;
  }();
  return on;
}

}  // namespace b200knn

extern "C" {

const char* b200knn_last_error(void) { return last_error_cstr(); }
int b200knn_version(void) { return 100; }
uint64_t b200knn_launch_count(void) { return g_launches.load(); }

static int index_create_impl(const float* d_xyz, PointLayout stride, int64_t n, const uint32_t* d_ids, uint32_t id_offset,
                             cudaStream_t stream, b200knn_index** out);

int b200knn_index_create(const float* d_xyz, int64_t n, const uint32_t* d_ids, uint32_t id_offset,
                         void* stream_, b200knn_index** out) {
  return index_create_impl(d_xyz, layout_of(3), n, d_ids, id_offset, (cudaStream_t)stream_, out);
}

int b200knn_index_create_xyzw(const float* d_xyzw, int64_t n, uint32_t id_offset, void* stream_,
                              b200knn_index** out) {
  return index_create_impl(d_xyzw, layout_of(4), n, nullptr, id_offset, (cudaStream_t)stream_, out);
}

int b200knn_index_create_xy(const float* d_xy, int64_t n, const uint32_t* d_ids, uint32_t id_offset, void* stream_,
                            b200knn_index** out) {
  const int rc = index_create_impl(d_xy, layout_of(2), n, d_ids, id_offset, (cudaStream_t)stream_, out);
  if (out && *out) (*out)->dims = 2;
  return rc;
}

int b200knn_index_create_records(const void* d_records, int64_t n, size_t record_bytes, size_t position_offset,
                                 int dims, const uint32_t* d_ids, uint32_t id_offset, void* stream_,
                                 b200knn_index** out) {
  KNN_REQUIRE(dims == 2 || dims == 3, "index_create_records: dims must be 2 or 3");
  KNN_REQUIRE(record_bytes % 4 == 0 && position_offset % 4 == 0, "index_create_records: record size and position offset must be multiples of 4 bytes");
  KNN_REQUIRE(position_offset + (size_t)dims * 4 <= record_bytes, "index_create_records: the position does not fit inside a record");
  KNN_REQUIRE(record_bytes / 4 <= 0x7fffffff, "index_create_records: record too large");
  KNN_REQUIRE(n == 0 || (reinterpret_cast<uintptr_t>(d_records) & 3u) == 0, "index_create_records: records must be 4-byte aligned");
  const float* first = reinterpret_cast<const float*>(static_cast<const char*>(d_records) + position_offset);
  const int rc = index_create_impl(first, PointLayout{(int)(record_bytes / 4), dims, 0}, n, d_ids, id_offset,
                                   (cudaStream_t)stream_, out);
  if (out && *out) (*out)->dims = dims;
  return rc;
}

static int index_create_impl(const float* d_xyz, PointLayout stride, int64_t n, const uint32_t* d_ids, uint32_t id_offset,
                             cudaStream_t stream, b200knn_index** out) {
  NvtxRange range("b200knn:index_build");
  KNN_REQUIRE(out != nullptr, "index_create: out is NULL");
  KNN_REQUIRE(n >= 0 && n < (int64_t)0x7fffffff - 64, "index_create: n must be in [0, 2^31-64) per GPU");
  b200knn_index* idx = new b200knn_index();
  *out = idx;
  idx->n = n;
  idx->id_offset = id_offset;
  idx->stream = stream;
  KNN_CUDA_TRY(cudaGetDevice(&idx->device));
  keep_pool_memory(idx->device);
  KNN_CUDA_TRY(cudaDeviceGetAttribute(&idx->num_sms, cudaDevAttrMultiProcessorCount, idx->device));
  KNN_CUDA_TRY(cudaMallocAsync(&idx->d_bounds_ord, 6 * sizeof(unsigned), stream));
  KNN_CUDA_TRY(cudaMallocAsync(&idx->d_bounds6, 6 * sizeof(float), stream));
  KNN_CUDA_TRY(cudaMallocAsync(&idx->task_counter, sizeof(unsigned long long), stream));
  KNN_CUDA_TRY(cudaMallocAsync(&idx->stats, 16 * sizeof(unsigned long long), stream));
  KNN_CUDA_TRY(cudaMemsetAsync(idx->stats, 0, 16 * sizeof(unsigned long long), stream));
  int rc;
  if ((rc = bounds_init(idx->d_bounds_ord, stream))) return rc;
  if ((rc = bounds_accumulate(d_xyz, stride, n, idx->d_bounds_ord, stream))) return rc;
  if ((rc = bounds_decode(idx->d_bounds_ord, idx->d_bounds6, stream))) return rc;
  SmallRead bounds_read;
  if ((rc = bounds_read.begin(idx->d_bounds6, 6 * sizeof(float), idx->h_bounds, stream))) return rc;
  if (n == 0) {
    KNN_CUDA_TRY(cudaStreamSynchronize(stream));
    bounds_read.end();
    return 0;
  }
  const int64_t groups = idx->groups();
  const int64_t padded = groups * kGroup;
  KNN_CUDA_TRY(cudaMallocAsync(&idx->pts, padded * sizeof(Float4), stream));
  KNN_CUDA_TRY(cudaMallocAsync(&idx->result_sorted, padded * sizeof(float), stream));
  // level sizes
  int cnt = (int)groups;
  int levels = 0;
  for (;;) {
    KNN_REQUIRE(levels < kMaxLevels, "index_create: too many tree levels");
    idx->count[levels] = cnt;
    KNN_CUDA_TRY(cudaMallocAsync(&idx->boxes[levels], (size_t)cnt * sizeof(Box), stream));
    ++levels;
    if (cnt <= kFanout) break;
    cnt = (int)div_up(cnt, kFanout);
  }
  idx->levels = levels;

  // temporaries (stream-ordered)
  StreamBuf<uint64_t> keys(stream), keys_tmp(stream);
  StreamBuf<uint32_t> vals(stream), vals_tmp(stream), scratch(stream);
  KNN_CUDA_TRY(keys.alloc_bytes(n * sizeof(uint64_t)));
  KNN_CUDA_TRY(keys_tmp.alloc_bytes(n * sizeof(uint64_t)));
  KNN_CUDA_TRY(vals.alloc_bytes(n * sizeof(uint32_t)));
  KNN_CUDA_TRY(vals_tmp.alloc_bytes(n * sizeof(uint32_t)));
  KNN_CUDA_TRY(scratch.alloc_bytes(radix_sort_scratch_bytes(n)));
  StreamBuf<unsigned> heavy(stream);
  KNN_CUDA_TRY(heavy.alloc_bytes(sizeof(unsigned)));
  // The order only has to be spatial down to the 1024-point runs that refine_runs re-sorts anyway, so the sort
  // looks at the leading kIndexSortBits of the 63-bit key: 5 digit passes instead of 8. That is enough unless more
  // than 1024 points share one such cell (extent / 10321 per axis) — checked on the device, read back with the
  // synchronisation below, and answered by building again on all 64 bits (B200KNN_SORT_BITS=64: always).
  for (int attempt = 0; attempt < 2; ++attempt) {
    const int low_bit = attempt == 0 ? 64 - index_sort_bits() : 0;
    if ((rc = morton_keys(d_xyz, stride, n, idx->d_bounds6, keys, vals, stream))) return rc;
    int in_tmp = 0;
    if ((rc = radix_sort_u64(keys, vals, keys_tmp, vals_tmp, scratch, n, low_bit, 64, stream, &in_tmp))) return rc;
    const uint32_t* order = in_tmp ? vals_tmp.get() : vals.get();
    unsigned h_heavy = 0;
    SmallRead heavy_read;
    if (low_bit > 0) {
      if ((rc = heavy_cells(in_tmp ? keys_tmp.get() : keys.get(), n, low_bit, heavy, stream))) return rc;
      if ((rc = heavy_read.begin(heavy, sizeof(unsigned), &h_heavy, stream))) return rc;
    }
    if ((rc = gather_points(d_xyz, stride, order, d_ids, id_offset, n, idx->pts, stream))) return rc;
    if ((rc = refine_runs(idx->pts, n, stream))) return rc;
    if ((rc = build_leaf_boxes(idx->pts, n, idx->boxes[0], stream))) return rc;
    for (int l = 1; l < levels; ++l)
      if ((rc = build_upper_boxes(idx->boxes[l - 1], idx->count[l - 1], idx->boxes[l], stream))) return rc;
    KNN_CUDA_TRY(cudaStreamSynchronize(stream));
    heavy_read.end();
    if (attempt == 0) bounds_read.end();
    idx->sort_bits = 64 - low_bit;
    if (low_bit == 0 || h_heavy == 0) break;
  }
  keys.release();
  keys_tmp.release();
  vals.release();
  vals_tmp.release();
  scratch.release();
  heavy.release();
  return 0;
}

void b200knn_index_destroy(b200knn_index* idx) {
  if (!idx) return;
  cudaStreamSynchronize(idx->stream);   // buffers go back to the pool only after their last use
  if (idx->pts) cudaFreeAsync(idx->pts, idx->stream);
  for (int l = 0; l < kMaxLevels; ++l) if (idx->boxes[l]) cudaFreeAsync(idx->boxes[l], idx->stream);
  if (idx->d_bounds_ord) cudaFreeAsync(idx->d_bounds_ord, idx->stream);
  if (idx->d_bounds6) cudaFreeAsync(idx->d_bounds6, idx->stream);
  if (idx->result_sorted) cudaFreeAsync(idx->result_sorted, idx->stream);
  if (idx->task_counter) cudaFreeAsync(idx->task_counter, idx->stream);
  if (idx->stats) cudaFreeAsync(idx->stats, idx->stream);
  delete idx;
}

int64_t b200knn_index_num_points(const b200knn_index* idx) { return idx ? idx->n : 0; }
int b200knn_index_num_levels(const b200knn_index* idx) { return idx ? idx->levels : 0; }
int b200knn_index_sort_bits(const b200knn_index* idx) { return idx ? idx->sort_bits : 0; }

int b200knn_index_bounds(const b200knn_index* idx, float* host_bounds6) {
  KNN_REQUIRE(idx && host_bounds6, "index_bounds: NULL argument");
  memcpy(host_bounds6, idx->h_bounds, sizeof(idx->h_bounds));
  return 0;
}

int b200knn_index_copy_sorted(const b200knn_index* idx, float* host_xyzw) {
  KNN_REQUIRE(idx && host_xyzw, "index_copy_sorted: NULL argument");
  if (idx->n == 0) return 0;
  KNN_CUDA_TRY(cudaMemcpy(host_xyzw, idx->pts, idx->n * sizeof(Float4), cudaMemcpyDeviceToHost));
  return 0;
}

int b200knn_index_copy_boxes(const b200knn_index* idx, int level, float* host_boxes8, int* count) {
  KNN_REQUIRE(idx && count, "index_copy_boxes: NULL argument");
  KNN_REQUIRE(level >= 0 && level < idx->levels, "index_copy_boxes: level out of range");
  *count = idx->count[level];
  if (host_boxes8)
    KNN_CUDA_TRY(cudaMemcpy(host_boxes8, idx->boxes[level], (size_t)idx->count[level] * sizeof(Box),
                            cudaMemcpyDeviceToHost));
  return 0;
}

static void fill_local_only(RingParams& p, b200knn_index* idx, int k, float max_radius) {
  memset(&p, 0, sizeof(p));
  p.fill = convergent_fill_enabled() ? 1 : 0;
  p.tree = idx->view();
  p.k = k;
  p.sentinel = sentinel_d2(max_radius);
  p.R = 1;
  p.rank = 0;
  p.num_chunks = 1;
  p.step_begin = 0;
  p.step_end = 1;
  p.use_flags = 0;
  p.seq_base = 0;
  p.task_counter = idx->task_counter;
  p.stats = idx->stats_on ? idx->stats : nullptr;
  p.allow_global_heap = 1;
}

int b200knn_knn_self(b200knn_index* idx, int k, float max_radius, float* d_out, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  NvtxRange range("b200knn:knn_self");
  KNN_REQUIRE(idx && d_out, "knn_self: NULL argument");
  KNN_REQUIRE(k >= 1, "knn_self: k must be >= 1");
  if (idx->n == 0) return 0;
  RingParams p;
  fill_local_only(p, idx, k, max_radius);
  p.groups_per_chunk = (int)idx->groups();
  p.home_queries = idx->pts;
  p.home_n = idx->n;
  p.home_is_tree = 1;
  p.result_sorted = idx->result_sorted;
  if (idx->stats_on) KNN_CUDA_TRY(cudaMemsetAsync(idx->stats, 0, 16 * sizeof(unsigned long long), stream));
  int rc;
  if ((rc = ring_launch(p, idx->num_sms, stream))) return rc;
  return scatter_results(idx->pts, idx->result_sorted, idx->n, idx->id_offset, d_out, stream);
}

int b200knn_knn_query(b200knn_index* idx, const float* d_queries, int64_t nq, int k, float max_radius,
                      float* d_out, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  KNN_REQUIRE(idx && d_out, "knn_query: NULL argument");
  KNN_REQUIRE(k >= 1, "knn_query: k must be >= 1");
  KNN_REQUIRE(nq >= 0 && nq < (int64_t)0x7fffffff - 64, "knn_query: nq out of range");
  if (nq == 0) return 0;
  const int64_t groups = div_up(nq, kGroup);
  StreamBuf<uint64_t> keys(stream), keys_tmp(stream);
  StreamBuf<uint32_t> vals(stream), vals_tmp(stream), scratch(stream);
  StreamBuf<Float4> qsorted(stream);
  StreamBuf<float> rsorted(stream);
  KNN_CUDA_TRY(keys.alloc_bytes(nq * sizeof(uint64_t)));
  KNN_CUDA_TRY(keys_tmp.alloc_bytes(nq * sizeof(uint64_t)));
  KNN_CUDA_TRY(vals.alloc_bytes(nq * sizeof(uint32_t)));
  KNN_CUDA_TRY(vals_tmp.alloc_bytes(nq * sizeof(uint32_t)));
  KNN_CUDA_TRY(scratch.alloc_bytes(radix_sort_scratch_bytes(nq)));
  KNN_CUDA_TRY(qsorted.alloc_bytes(groups * kGroup * sizeof(Float4)));
  KNN_CUDA_TRY(rsorted.alloc_bytes(groups * kGroup * sizeof(float)));
  int rc;
  // Order queries along the index's own Morton grid (queries outside its box clamp to the faces).
  if ((rc = morton_keys(d_queries, idx->dims == 2 ? 2 : 3, nq, idx->d_bounds6, keys, vals, stream))) return rc;
  int in_tmp = 0;
  if ((rc = radix_sort_u64(keys, vals, keys_tmp, vals_tmp, scratch, nq, 64 - index_sort_bits(), 64, stream, &in_tmp))) return rc;
  if ((rc = gather_points(d_queries, idx->dims == 2 ? 2 : 3, in_tmp ? vals_tmp.get() : vals.get(), nullptr, 0, nq, qsorted, stream))) return rc;
  if ((rc = refine_runs(qsorted, nq, stream))) return rc;   // compact query groups, same as leaves
  RingParams p;
  fill_local_only(p, idx, k, max_radius);
  p.groups_per_chunk = (int)groups;
  p.home_queries = qsorted;
  p.home_n = nq;
  p.result_sorted = rsorted;
  if (idx->stats_on) KNN_CUDA_TRY(cudaMemsetAsync(idx->stats, 0, 16 * sizeof(unsigned long long), stream));
  if ((rc = ring_launch(p, idx->num_sms, stream))) return rc;
  if ((rc = scatter_results(qsorted, rsorted, nq, 0, d_out, stream))) return rc;
  keys.release();
  keys_tmp.release();
  vals.release();
  vals_tmp.release();
  scratch.release();
  qsorted.release();
  rsorted.release();
  KNN_CUDA_TRY(cudaStreamSynchronize(stream));
  return 0;
}

// k nearest neighbours with ids: pass 1 = the k-th-distance kernel with raw output, pass 2 = collect
// (neighbors.cu). d_queries == NULL: every indexed point is a query (rows in the caller's point order;
// the index must have been created without explicit ids, so that id - id_offset is the position).
int b200knn_knn_neighbors(b200knn_index* idx, const float* d_queries, int64_t nq, int k, float max_radius,
                          int32_t* d_out_ids, float* d_out_dist, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  NvtxRange range("b200knn:knn_neighbors");
  KNN_REQUIRE(idx && d_out_ids && d_out_dist, "knn_neighbors: NULL argument");
  KNN_REQUIRE(k >= 1, "knn_neighbors: k must be >= 1");
  const bool self = d_queries == nullptr;
  if (self) nq = idx->n;
  KNN_REQUIRE(nq >= 0 && nq < (int64_t)0x7fffffff - 64, "knn_neighbors: nq out of range");
  if (nq == 0) return 0;
  const int64_t groups = div_up(nq, kGroup);
  StreamBuf<uint64_t> keys(stream), keys_tmp(stream);
  StreamBuf<uint32_t> vals(stream), vals_tmp(stream), scratch(stream);
  StreamBuf<Float4> qsorted(stream);
  StreamBuf<float> limits(stream);
  int rc;
  KNN_CUDA_TRY(limits.alloc_bytes(groups * kGroup * sizeof(float)));
  if (!self) {
    KNN_CUDA_TRY(keys.alloc_bytes(nq * sizeof(uint64_t)));
    KNN_CUDA_TRY(keys_tmp.alloc_bytes(nq * sizeof(uint64_t)));
    KNN_CUDA_TRY(vals.alloc_bytes(nq * sizeof(uint32_t)));
    KNN_CUDA_TRY(vals_tmp.alloc_bytes(nq * sizeof(uint32_t)));
    KNN_CUDA_TRY(scratch.alloc_bytes(radix_sort_scratch_bytes(nq)));
    KNN_CUDA_TRY(qsorted.alloc_bytes(groups * kGroup * sizeof(Float4)));
    if ((rc = morton_keys(d_queries, idx->dims == 2 ? 2 : 3, nq, idx->d_bounds6, keys, vals, stream))) return rc;
    int in_tmp = 0;
    if ((rc = radix_sort_u64(keys, vals, keys_tmp, vals_tmp, scratch, nq, 64 - index_sort_bits(), 64, stream, &in_tmp))) return rc;
    if ((rc = gather_points(d_queries, idx->dims == 2 ? 2 : 3, in_tmp ? vals_tmp.get() : vals.get(), nullptr, 0, nq, qsorted, stream))) return rc;
    if ((rc = refine_runs(qsorted, nq, stream))) return rc;
  }
  RingParams p;
  fill_local_only(p, idx, k, max_radius);
  p.groups_per_chunk = (int)groups;
  p.home_queries = self ? idx->pts : qsorted;
  p.home_n = nq;
  p.home_is_tree = self ? 1 : 0;
  p.result_sorted = limits;
  p.raw_result = 1;
  if ((rc = ring_launch(p, idx->num_sms, stream))) return rc;
  CollectParams c;
  memset(&c, 0, sizeof(c));
  c.tree = idx->view();
  c.k = k;
  c.sentinel = p.sentinel;
  c.queries = p.home_queries;
  c.nq = nq;
  c.row_offset = self ? idx->id_offset : 0u;
  c.home_is_tree = p.home_is_tree;
  c.limit_sorted = limits;
  c.out_ids = d_out_ids;
  c.out_dist = d_out_dist;
  c.task_counter = idx->task_counter;
  if ((rc = collect_launch(c, idx->num_sms, stream))) return rc;
  limits.release();
  if (!self) {
    keys.release();
    keys_tmp.release();
    vals.release();
    vals_tmp.release();
    scratch.release();
    qsorted.release();
  }
  KNN_CUDA_TRY(cudaStreamSynchronize(stream));
  return 0;
}

// Find-closest-point: k = 1 traversal that also reports WHICH point is closest.
// Role of cukd::stackFree::fcp (cudaKDTree/cukd/fcp.h:210-254).
int b200knn_fcp_query(b200knn_index* idx, const float* d_queries, int64_t nq, float max_radius, int32_t* d_out_ids,
                      float* d_out_dist, void* stream_) {
  b200knn_fcp_params prm;
  prm.max_radius = max_radius;
  prm.eps = 0.f;
  prm.far_node_inspect_budget = -1;
  return b200knn_fcp_query_ex(idx, d_queries, nq, &prm, d_out_ids, d_out_dist, stream_);
}

// The same with the reference's FcpSearchParams (cukd/fcp.h:35-52): cutOffRadius, eps, far_node_inspect_budget.
int b200knn_fcp_query_ex(b200knn_index* idx, const float* d_queries, int64_t nq, const b200knn_fcp_params* prm,
                         int32_t* d_out_ids, float* d_out_dist, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  KNN_REQUIRE(prm, "fcp_query_ex: NULL parameters");
  KNN_REQUIRE(prm->eps >= 0.f, "fcp_query_ex: eps must be >= 0");
  const float max_radius = prm->max_radius;
  KNN_REQUIRE(idx && d_out_ids, "fcp_query: NULL argument");
  KNN_REQUIRE(nq >= 0 && nq < (int64_t)0x7fffffff - 64, "fcp_query: nq out of range");
  if (nq == 0) return 0;
  const int64_t groups = div_up(nq, kGroup);
  StreamBuf<uint64_t> keys(stream), keys_tmp(stream);
  StreamBuf<uint32_t> vals(stream), vals_tmp(stream), scratch(stream), ids_sorted(stream);
  StreamBuf<Float4> qsorted(stream);
  StreamBuf<float> rsorted(stream);
  KNN_CUDA_TRY(keys.alloc_bytes(nq * sizeof(uint64_t)));
  KNN_CUDA_TRY(keys_tmp.alloc_bytes(nq * sizeof(uint64_t)));
  KNN_CUDA_TRY(vals.alloc_bytes(nq * sizeof(uint32_t)));
  KNN_CUDA_TRY(vals_tmp.alloc_bytes(nq * sizeof(uint32_t)));
  KNN_CUDA_TRY(scratch.alloc_bytes(radix_sort_scratch_bytes(nq)));
  KNN_CUDA_TRY(qsorted.alloc_bytes(groups * kGroup * sizeof(Float4)));
  KNN_CUDA_TRY(rsorted.alloc_bytes(groups * kGroup * sizeof(float)));
  KNN_CUDA_TRY(ids_sorted.alloc_bytes(groups * kGroup * sizeof(uint32_t)));
  int rc;
  if ((rc = morton_keys(d_queries, idx->dims == 2 ? 2 : 3, nq, idx->d_bounds6, keys, vals, stream))) return rc;
  int in_tmp = 0;
  if ((rc = radix_sort_u64(keys, vals, keys_tmp, vals_tmp, scratch, nq, 64 - index_sort_bits(), 64, stream, &in_tmp))) return rc;
  if ((rc = gather_points(d_queries, idx->dims == 2 ? 2 : 3, in_tmp ? vals_tmp.get() : vals.get(), nullptr, 0, nq, qsorted, stream))) return rc;
  if ((rc = refine_runs(qsorted, nq, stream))) return rc;
  RingParams p;
  fill_local_only(p, idx, 1, max_radius);
  p.groups_per_chunk = (int)groups;
  p.home_queries = qsorted;
  p.home_n = nq;
  p.result_sorted = rsorted;
  p.result_ids = ids_sorted;
  p.tree.cull_scale = 1.f + prm->eps;
  p.tree.node_budget = prm->far_node_inspect_budget;
  if ((rc = ring_launch(p, idx->num_sms, stream))) return rc;
  // ids are 32-bit payloads: the float scatter moves them unchanged
  if ((rc = scatter_results(qsorted, reinterpret_cast<const float*>(ids_sorted.get()), nq, 0,
                            reinterpret_cast<float*>(d_out_ids), stream))) return rc;
  if (d_out_dist && (rc = scatter_results(qsorted, rsorted, nq, 0, d_out_dist, stream))) return rc;
  keys.release();
  keys_tmp.release();
  vals.release();
  vals_tmp.release();
  scratch.release();
  qsorted.release();
  rsorted.release();
  ids_sorted.release();
  KNN_CUDA_TRY(cudaStreamSynchronize(stream));
  return 0;
}

int b200knn_index_enable_stats(b200knn_index* idx, int on) {
  KNN_REQUIRE(idx, "index_enable_stats: NULL index");
  idx->stats_on = on != 0;
  return 0;
}

int b200knn_index_stats(b200knn_index* idx, uint64_t out8[8]) {
  KNN_REQUIRE(idx && out8, "index_stats: NULL argument");
  unsigned long long h[8];
  KNN_CUDA_TRY(cudaMemcpy(h, idx->stats, sizeof(h), cudaMemcpyDeviceToHost));
  for (int i = 0; i < 8; ++i) out8[i] = h[i];
  return 0;
}

int b200knn_mempool_usage(int64_t* reserved_now, int64_t* used_high, int reset_high) {
  int dev = 0;
  KNN_CUDA_TRY(cudaGetDevice(&dev));
  cudaMemPool_t pool;
  KNN_CUDA_TRY(cudaDeviceGetDefaultMemPool(&pool, dev));
  uint64_t v = 0;
  if (reserved_now) { KNN_CUDA_TRY(cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReservedMemCurrent, &v)); *reserved_now = (int64_t)v; }
  if (used_high) { KNN_CUDA_TRY(cudaMemPoolGetAttribute(pool, cudaMemPoolAttrUsedMemHigh, &v)); *used_high = (int64_t)v; }
  if (reset_high) { v = 0; KNN_CUDA_TRY(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrUsedMemHigh, &v)); }
  return 0;
}

int b200knn_index_stats_slowest(b200knn_index* idx, uint64_t out8[8]) {
  KNN_REQUIRE(idx && out8, "index_stats_slowest: NULL argument");
  unsigned long long h[8];
  KNN_CUDA_TRY(cudaMemcpy(h, idx->stats + 8, sizeof(h), cudaMemcpyDeviceToHost));
  for (int i = 0; i < 8; ++i) out8[i] = h[i];
  return 0;
}

int b200knn_index_unsort(b200knn_index* idx, const float* d_sorted_result, float* d_out, void* stream_) {
  KNN_REQUIRE(idx && d_sorted_result && d_out, "index_unsort: NULL argument");
  return scatter_results(idx->pts, d_sorted_result, idx->n, idx->id_offset, d_out, (cudaStream_t)stream_);
}

int b200knn_run_steps(b200knn_index* idx, const b200knn_step_args* a, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  KNN_REQUIRE(idx && a, "run_steps: NULL argument");
  KNN_REQUIRE(a->world >= 1 && a->world <= B200KNN_MAX_RANKS, "run_steps: world out of range");
  KNN_REQUIRE(a->groups_per_chunk >= 1 && a->num_chunks >= 1, "run_steps: bad chunking");
  KNN_REQUIRE(a->step_begin >= 0 && a->step_end <= a->num_chunks * a->world && a->step_begin <= a->step_end,
              "run_steps: bad step range");
  RingParams p;
  fill_local_only(p, idx, a->k, a->max_radius);
  p.R = a->world;
  p.rank = a->rank;
  p.groups_per_chunk = a->groups_per_chunk;
  p.num_chunks = a->num_chunks;
  p.step_begin = a->step_begin;
  p.step_end = a->step_end;
  p.home_queries = idx->pts;
  p.home_n = idx->n;
  p.home_is_tree = 1;
  for (int b = 0; b < 2; ++b) {
    p.inbox_heap[b] = a->inbox_heap[b];
    p.inbox_pos[b] = reinterpret_cast<const Float4*>(a->inbox_pos[b]);
    p.inbox_cnt[b] = a->inbox_cnt[b];
    p.next_heap[b] = a->outbox_heap[b];
    p.next_pos[b] = reinterpret_cast<Float4*>(a->outbox_pos[b]);
    p.next_cnt[b] = a->outbox_cnt[b];
  }
  p.result_sorted = a->result_sorted;
  return ring_launch(p, idx->num_sms, stream);
}

// ------------------------------------------------------------------------------------- fused ring

int b200knn_ring_create(b200knn_index* idx, int rank, int world, int k, int64_t max_points_any_rank,
                        int groups_per_chunk, b200knn_ring** out) {
  KNN_REQUIRE(idx && out, "ring_create: NULL argument");
  KNN_REQUIRE(world >= 1 && world <= B200KNN_MAX_RANKS, "ring_create: world must be in [1,16]");
  KNN_REQUIRE(rank >= 0 && rank < world, "ring_create: rank out of range");
  KNN_REQUIRE(k >= 1, "ring_create: k must be >= 1");
  KNN_REQUIRE(max_points_any_rank >= idx->n, "ring_create: max_points_any_rank < local point count");
  b200knn_ring* r = new b200knn_ring();
  *out = r;
  r->idx = idx;
  r->rank = rank;
  r->world = world;
  r->k = k;
  const int64_t max_groups = div_up(max_points_any_rank > 0 ? max_points_any_rank : 1, kGroup);
  int64_t Gc = groups_per_chunk > 0 ? groups_per_chunk : 65536;
  if (Gc > max_groups) Gc = max_groups;
  r->Gc = (int)Gc;
  r->C = (int)div_up(max_groups, Gc);
  r->mine = make_layout(k, world > 1 ? r->Gc : 1, max_groups * kGroup);
  KNN_CUDA_TRY(cudaMalloc(&r->base, r->mine.bytes));
  KNN_CUDA_TRY(cudaMemset(r->base, 0, r->mine.bytes));
  KNN_CUDA_TRY(cudaDeviceSynchronize());   // zeroed before the handle is exported or a kernel on another stream polls a flag
  return 0;
}

// Unmap the peers' buffers. Call on every rank, then fence (barrier), then ring_destroy: an exporter
// must not free memory that an importer still has mapped.
int b200knn_ring_disconnect(b200knn_ring* r) {
  KNN_REQUIRE(r, "ring_disconnect: NULL ring");
  if (r->opened_next) cudaIpcCloseMemHandle(r->next_base);
  if (r->opened_prev) cudaIpcCloseMemHandle(r->prev_base);
  r->opened_next = r->opened_prev = false;
  r->next_base = r->prev_base = nullptr;
  return 0;
}

void b200knn_ring_destroy(b200knn_ring* r) {
  if (!r) return;
  b200knn_ring_disconnect(r);
  cudaFree(r->base);
  delete r;
}

int64_t b200knn_ring_comm_bytes(const b200knn_ring* r) { return r ? r->mine.bytes : 0; }

int b200knn_ring_export(b200knn_ring* r, void* blob_) {
  KNN_REQUIRE(r && blob_, "ring_export: NULL argument");
  IpcBlob blob;
  memset(&blob, 0, sizeof(blob));
  if (r->world > 1) KNN_CUDA_TRY(cudaIpcGetMemHandle(&blob.handle, r->base));
  blob.layout = r->mine;
  blob.rank = r->rank;
  blob.world = r->world;
  blob.k = r->k;
  blob.groups_per_chunk = r->Gc;
  blob.pid = 0;
  memset(blob_, 0, B200KNN_IPC_BLOB_BYTES);
  memcpy(blob_, &blob, sizeof(blob));
  return 0;
}

int b200knn_ring_connect(b200knn_ring* r, const void* all_blobs) {
  KNN_REQUIRE(r && all_blobs, "ring_connect: NULL argument");
  if (r->world == 1) return 0;
  const char* bytes = static_cast<const char*>(all_blobs);
  auto blob_of = [&](int rank) {
    IpcBlob b;
    memcpy(&b, bytes + (size_t)rank * B200KNN_IPC_BLOB_BYTES, sizeof(b));
    return b;
  };
  for (int q = 0; q < r->world; ++q) {
    const IpcBlob b = blob_of(q);
    KNN_REQUIRE(b.rank == q && b.world == r->world, "ring_connect: blobs are not in rank order");
    KNN_REQUIRE(b.k == r->k && b.groups_per_chunk == r->Gc, "ring_connect: ranks disagree on k or chunking");
  }
  const int next = (r->rank + 1) % r->world, prev = (r->rank + r->world - 1) % r->world;
  IpcBlob nb = blob_of(next);
  void* ptr = nullptr;
  KNN_CUDA_TRY(cudaIpcOpenMemHandle(&ptr, nb.handle, cudaIpcMemLazyEnablePeerAccess));
  r->next_base = static_cast<char*>(ptr);
  r->next_layout = nb.layout;
  r->opened_next = true;
  if (prev == next) {
    r->prev_base = r->next_base;
    r->prev_layout = nb.layout;
  } else {
    IpcBlob pb = blob_of(prev);
    KNN_CUDA_TRY(cudaIpcOpenMemHandle(&ptr, pb.handle, cudaIpcMemLazyEnablePeerAccess));
    r->prev_base = static_cast<char*>(ptr);
    r->prev_layout = pb.layout;
    r->opened_prev = true;
  }
  return 0;
}

int b200knn_ring_run(b200knn_ring* r, float max_radius, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  KNN_REQUIRE(r, "ring_run: NULL ring");
  KNN_REQUIRE(r->world == 1 || r->next_base, "ring_run: ring_connect has not been called");
  b200knn_index* idx = r->idx;
  RingParams p;
  fill_local_only(p, idx, r->k, max_radius);
  p.R = r->world;
  p.rank = r->rank;
  p.groups_per_chunk = r->Gc;
  p.num_chunks = r->C;
  p.step_begin = 0;
  p.step_end = r->C * r->world;
  p.use_flags = r->world > 1;
  p.seq_base = r->seq_base;
  p.home_queries = idx->pts;
  p.home_n = idx->n;
  p.home_is_tree = 1;
  char* nb = r->world > 1 ? r->next_base : r->base;
  char* pb = r->world > 1 ? r->prev_base : r->base;
  const RingLayout& NL = r->world > 1 ? r->next_layout : r->mine;
  const RingLayout& PL = r->world > 1 ? r->prev_layout : r->mine;
  for (int b = 0; b < 2; ++b) {
    p.inbox_heap[b] = reinterpret_cast<const float*>(r->base + r->mine.heap[b]);
    p.inbox_pos[b] = reinterpret_cast<const Float4*>(r->base + r->mine.pos[b]);
    p.inbox_cnt[b] = reinterpret_cast<const int*>(r->base + r->mine.cnt[b]);
    p.flag_in[b] = reinterpret_cast<const unsigned*>(r->base + r->mine.flag_in[b]);
    p.ack_in[b] = reinterpret_cast<const unsigned*>(r->base + r->mine.ack_in[b]);
    p.next_heap[b] = reinterpret_cast<float*>(nb + NL.heap[b]);
    p.next_pos[b] = reinterpret_cast<Float4*>(nb + NL.pos[b]);
    p.next_cnt[b] = reinterpret_cast<int*>(nb + NL.cnt[b]);
    p.next_flag[b] = reinterpret_cast<unsigned*>(nb + NL.flag_in[b]);
    p.prev_ack[b] = reinterpret_cast<unsigned*>(pb + PL.ack_in[b]);
  }
  // last hop of a group runs on the rank before its home: results go to the NEXT rank's array
  p.result_sorted = reinterpret_cast<float*>(nb + NL.result);
  r->seq_base += (unsigned)(r->C * r->world + 8);
  if (idx->stats_on) KNN_CUDA_TRY(cudaMemsetAsync(idx->stats, 0, 16 * sizeof(unsigned long long), stream));
  return ring_launch(p, idx->num_sms, stream);
}

int b200knn_ring_results(b200knn_ring* r, float* d_out, void* stream_) {
  KNN_REQUIRE(r && d_out, "ring_results: NULL argument");
  return scatter_results(r->idx->pts, reinterpret_cast<const float*>(r->base + r->mine.result), r->idx->n,
                         r->idx->id_offset, d_out, (cudaStream_t)stream_);
}

}  // extern "C"
