// kernels.h — internal C++ interfaces between the .cu translation units and api.cpp.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include "knn_math.h"

namespace b200knn {

// ---------------- radix sort (radix_sort.cu) ----------------
size_t radix_sort_scratch_bytes(int64_t n);
// Sorts bits [begin_bit, end_bit) in 8-bit digits. Ping-pongs between (keys,vals) and the tmp
// pair; *result_in_tmp tells where the sorted sequence ended up.
int radix_sort_u64(uint64_t* keys, uint32_t* vals, uint64_t* keys_tmp, uint32_t* vals_tmp,
                   uint32_t* scratch, int64_t n, int begin_bit, int end_bit, cudaStream_t stream,
                   int* result_in_tmp);
int radix_sort_u32(uint32_t* keys, uint32_t* vals, uint32_t* keys_tmp, uint32_t* vals_tmp,
                   uint32_t* scratch, int64_t n, int begin_bit, int end_bit, cudaStream_t stream,
                   int* result_in_tmp);

// ---------------- tree (build.cu) ----------------
// Wide (32-ary) bounding-box tree over Morton-sorted points. Level 0 = leaf boxes (32 points
// each), level l+1 groups 32 consecutive boxes of level l. Implicit topology: no pointers.
struct TreeView {
  const Float4* pts;            // sorted points; .w carries the caller's point id (as bits)
  const Box* boxes[kMaxLevels]; // boxes[l][i]
  int count[kMaxLevels];        // number of boxes on level l
  int levels;                   // number of levels in use (top level has <= 32 boxes)
  int n;                        // number of points
  int queue_cap;                // frontier entries a warp may use (<= 128; smaller only to test the overflow paths)
  // Approximate search, role of FcpSearchParams::eps / far_node_inspect_budget and of the traversal's eps
  // (cukd/fcp.h:35-52, cukd/traverse-stack-free.h:32,85). Exact search: 1.0f and -1.
  float cull_scale;             // 1 + eps: a node is skipped when lower_bound * (1 + eps) >= current k-th distance
  int node_budget;              // heap kernels (find-closest-point): after the seed neighbourhood, at most this many
                                // further nodes (inner or leaf) are inspected per query group; < 0 = no limit
};

struct Bounds { float lo[3]; float hi[3]; };

// d_bounds_ord: 6 unsigned (order-preserving encoding), must be initialised by bounds_init.
int bounds_init(unsigned* d_bounds_ord, cudaStream_t stream);
// Where the coordinates of point i live: `dims` (2 or 3) floats at d_xyz + stride * i; a missing z is 0.
// w_is_id: the float after z carries the point's id (points that arrive from another rank).
struct PointLayout { int stride, dims, w_is_id; };
KNN_HD PointLayout layout_of(int stride) { return PointLayout{stride, stride >= 3 ? 3 : 2, stride == 4 ? 1 : 0}; }
int bounds_accumulate(const float* d_xyz, PointLayout lay, int64_t n, unsigned* d_bounds_ord, cudaStream_t stream);
inline int bounds_accumulate(const float* d_xyz, int stride, int64_t n, unsigned* d_bounds_ord, cudaStream_t stream) {
  return bounds_accumulate(d_xyz, layout_of(stride), n, d_bounds_ord, stream);
}
int bounds_decode(const unsigned* d_bounds_ord, float* d_bounds6, cudaStream_t stream);
// `stride` = floats per input point: 3 (xyz) or 4 (xyzw, id bits in w).
// keys[i] = morton63(point i) relative to *d_bounds6 (device-resident lo[3],hi[3]); vals[i] = i.
int morton_keys(const float* d_xyz, PointLayout lay, int64_t n, const float* d_bounds6, uint64_t* keys,
                uint32_t* vals, cudaStream_t stream);
inline int morton_keys(const float* d_xyz, int stride, int64_t n, const float* d_bounds6, uint64_t* keys,
                       uint32_t* vals, cudaStream_t stream) {
  return morton_keys(d_xyz, layout_of(stride), n, d_bounds6, keys, vals, stream);
}
// sorted[i] = (xyz[order[i]], id = ids ? ids[order[i]] : order[i] + id_offset)
int gather_points(const float* d_xyz, PointLayout lay, const uint32_t* order, const uint32_t* ids, uint32_t id_offset,
                  int64_t n, Float4* sorted, cudaStream_t stream);
inline int gather_points(const float* d_xyz, int stride, const uint32_t* order, const uint32_t* ids, uint32_t id_offset,
                         int64_t n, Float4* sorted, cudaStream_t stream) {
  return gather_points(d_xyz, layout_of(stride), order, ids, id_offset, n, sorted, stream);
}
// *d_flag = 1 if more than 1024 points share the leading (64 - shift) bits of their key (keys sorted), else 0.
int heavy_cells(const uint64_t* sorted_keys, int64_t n, int shift, unsigned* d_flag, cudaStream_t stream);
// Re-order every 1024-point Morton run into a balanced k-d tree with compact 32-point leaves.
int refine_runs(Float4* sorted, int64_t n, cudaStream_t stream);
int build_leaf_boxes(const Float4* sorted, int64_t n, Box* level0, cudaStream_t stream);
int build_upper_boxes(const Box* child, int child_count, Box* parent, cudaStream_t stream);

// ---------------- query / ring (query.cu) ----------------
constexpr int kMaxRanks = 16;

// One persistent kernel does everything between "points are sorted" and "k-th distances exist":
// it walks tasks (step s, group g); step s = chunk*R + hop. hop 0 takes fresh queries from the
// local tree's own points, hops 1..R-1 take groups another rank pushed into the local inbox,
// the last hop writes k-th distances to the group's home rank. R == 1 degenerates to a plain
// single-GPU all-kNN kernel with no flags touched.
struct RingParams {
  TreeView tree;
  int k;
  float sentinel;              // max_radius^2 or +inf
  int R, rank;
  int groups_per_chunk;        // Gc: identical on all ranks
  int num_chunks;              // C: identical on all ranks
  int step_begin, step_end;    // run steps [step_begin, step_end) (fused ring: 0 .. C*R)
  int use_flags;               // 1: in-kernel peer handshake; 0: host orders the steps (baseline transport)
  unsigned seq_base;
  // queries of hop 0: the local tree's own points, or an external sorted query set
  const Float4* home_queries;  // tree.pts for self-kNN
  int64_t home_n;
  int home_is_tree;            // 1 if home_queries are exactly the tree's own sorted points
  // local inbox (written by the previous rank)
  const float* inbox_heap[2];  // [Gc][k][32]
  const Float4* inbox_pos[2];  // [Gc][32]
  const int* inbox_cnt[2];     // [Gc]
  const unsigned* flag_in[2];  // [Gc]  data-ready sequence numbers (written by prev)
  const unsigned* ack_in[2];   // [Gc]  slot-consumed sequence numbers (written by next)
  // peer pointers
  float* next_heap[2];
  Float4* next_pos[2];
  int* next_cnt[2];
  unsigned* next_flag[2];      // next rank's flag_in
  unsigned* prev_ack[2];       // prev rank's ack_in
  float* result_sorted;        // home rank's result array for the LAST hop (next rank's, or own if R==1)
  unsigned* result_ids;        // optional (single rank, k == 1): id of the closest point, 0xffffffff if none
  unsigned long long* task_counter;   // local, zeroed before launch
  unsigned long long* stats;   // optional [8]: leaves visited, candidates tested, candidates accepted, inner nodes,
                               //               DFS fallbacks, frontier purges, list flushes
  int allow_global_heap;       // 1: k beyond the shared-memory limit may use a global-memory heap (allocates)
  float* global_heaps;         // set by ring_launch on that fallback path
  int fill;                    // heap kernels only. 1: fresh heaps take their first floor(k/32) leaves by plain appends + heapify
  int raw_result;              // 1: result_sorted receives the heap top itself (squared distance / sentinel)
};
bool convergent_fill_enabled();   // api.cu: host-side switch (B200KNN_FILL=0/1), read once

// Warps per CTA the query kernel would run with for this k; 0 = k does not fit in shared memory (the
// global-memory heap fallback, which allocates, is the only way then).
int ring_pick_warps_per_cta(int k);
int ring_launch(const RingParams& p, int num_sms, cudaStream_t stream);

// ---------------- neighbour lists (neighbors.cu) ----------------
// Second pass of the k-nearest-NEIGHBOURS query: with every query's exact k-th squared distance known,
// collect the ids and distances of all points inside it (fixed-radius traversal, no heap).
struct CollectParams {
  TreeView tree;
  int k;
  float sentinel;              // max_radius^2 or +inf: a limit equal to it means "fewer than k points in range"
  const Float4* queries;       // sorted queries, .w = row of the output (as bits) + row_offset
  int64_t nq;
  uint32_t row_offset;
  int home_is_tree;            // queries are exactly the tree's own sorted points
  const float* limit_sorted;   // per sorted query: heap top of the first pass
  int32_t* out_ids;            // [rows][k]
  float* out_dist;             // [rows][k]
  unsigned long long* task_counter;
};
int collect_launch(const CollectParams& p, int num_sms, cudaStream_t stream);

// ---------------- culled exchange / redistribution (pv.cu) ----------------
// A slot carries one group: heap [k][32] floats, positions [32] float4, meta [4] int32.
KNN_HD size_t pv_slot_bytes(int k) { return (size_t)k * 128 + 512 + 16; }

// Every rank's upper tree levels (level 1 and above: one box per 1024 points and coarser), replicated on every
// rank by peer stores (pv.cu, tops_publish_kernel). "Does this group's ball reach rank s?" is answered by
// walking rank s's published levels: the walk only descends where the ball goes, so it is as fine as the
// data is dense — a fixed number of cover boxes per rank is not (a rank that owns a piece of a dense
// cluster AND far-flung sparse points spends its boxes on the latter, and one coarse box over its piece
// of the cluster then attracts every group of the neighbour that owns the rest of it).
struct TopHead {
  int levels;                   // levels of that rank's tree (0: the rank is empty)
  int base_level;               // published levels: base_level .. levels-1 (base_level = 1, or 0 for a single-level tree)
  int count[kMaxLevels];        // boxes per level
  int offset[kMaxLevels];       // where a published level starts inside the rank's table
  float box[6];                 // overall bounding box (quick reject)
};
struct TopsView {
  const Box* boxes;             // [R][cap]
  const TopHead* heads;         // [R]
  int64_t cap;
};
struct PvMainParams {
  TreeView tree;
  int k;
  float sentinel;
  int phase;                   // 0 = LOCAL (home groups, fresh heaps), 1 = SERVE (visiting groups from the inbox)
  int fill;                    // as RingParams::fill, LOCAL phase only
  float* global_heaps;         // set by pv_main_launch when k exceeds the shared-memory heap
  int R, rank;
  TopsView tops;               // every rank's published tree levels
  const Box* territory;        // optional [R][kTerritoryCells]: the octree cells that make up each rank's share of
  const int* territory_count;  // key space after a redistribution (redist.cu); NULL: ranks own arbitrary point sets
  // LOCAL
  const Float4* home_queries;
  int64_t home_n;
  int home_is_tree;
  int64_t group_begin;         // first home group of this chunk
  int group_count;             // groups in this chunk
  float* result_sorted;
  char* pending;               // local pending slots
  int* pending_count;
  // SERVE
  const char* inbox;           // local inbox slots (filled by a peer's pv_push)
  const int* inbox_count;
  char* src_pending;           // the visiting groups' home GPU's pending slots (peer pointer)
  unsigned long long* task_counter;
  unsigned long long* stats;
  // One-round exchange (phases LOCAL with fused != 0, SERVE_FRESH, MERGE; sorted-list kernels only):
  // LOCAL posts a REQUEST (32 queries with their current k-th distances, 528 bytes) to every rank a pending group
  // reaches; every rank serves all requests it received in one pass, starting each from an empty list cut off at the
  // lanes' k-th distances, and sends back only the candidates it found (RETURN record); home merges the returns.
  int fused;
  int64_t cap_req;              // requests per (source, destination) region
  int64_t cap_ret;              // return records a rank can receive
  char* req_inbox[kMaxRanks];   // rank s's request area [R][cap_req] (peer pointers); this rank writes region [rank]
  int* req_out_count;           // [R] local: requests posted to each rank so far
  const char* my_req_inbox;     // SERVE_FRESH: this rank's request area
  int req_in[kMaxRanks];        // SERVE_FRESH: requests received from each rank (known to every host from the vote)
  char* ret_area[kMaxRanks];    // rank h's return area [cap_ret] (peer pointers)
  int* ret_count[kMaxRanks];    // rank h's return counter (incremented over NVLink by the servers)
  const char* my_ret_area;      // MERGE
  const int* my_ret_count;
};
KNN_HD size_t pv_request_bytes() { return 512 + 16; }
KNN_HD size_t pv_return_bytes(int k) { return (size_t)k * 128 + 16; }
int pv_main_launch(const PvMainParams& p, int max_tasks, int num_sms, cudaStream_t stream);
int pv_push_launch(const char* pending, const int* pending_count, int k, const TopsView& tops, const Box* territory,
                   const int* territory_count, int dst, char* peer_inbox, int* out_count, int num_sms, cudaStream_t stream);
struct TopsPublishParams {
  TreeView tree;
  float overall[6];             // this rank's bounding box
  int R, rank;
  int64_t cap;
  Box* tables[kMaxRanks];       // every rank's [R][cap] table (peer pointers)
  TopHead* heads[kMaxRanks];    // every rank's [R] heads
};
int64_t tops_capacity(int64_t max_groups);      // boxes a rank with at most max_groups leaves publishes (upper bound)
int tops_publish_launch(const TopsPublishParams& p, int num_sms, cudaStream_t stream);
bool pv_one_round_capable(int k);     // the sorted-list kernels serve this k (one-round exchange possible)
int pv_fused_launch(const PvMainParams& p, int phase, int max_tasks, int num_sms, cudaStream_t stream);   // 2 = SERVE_FRESH, 3 = MERGE
int pv_publish_launch(const int* out_count, int* peer_inbox_count, cudaStream_t stream);
int pv_final_launch(const char* pending, const int* pending_count, int k, float* result_sorted, int num_sms,
                    cudaStream_t stream);

struct DevBarrierParams {
  unsigned* peer_flags[kMaxRanks];   // base of every rank's flag array (peer pointers); slot [rank] is ours there
  const unsigned* local_flags;       // this rank's flag array: slot [r] is written by rank r
  unsigned seq;
  int rank, R;
};
int dev_barrier_launch(const DevBarrierParams& p, cudaStream_t stream);

// Small device-side collectives over the symmetric heap: every rank pushes up to
// kCollWords 32-bit words into a mailbox on every peer and raises a sequence flag there, then reads
// its own mailboxes. Replaces MPI_Allgather / MPI_Allreduce on host staging buffers
// (/root/reference/prePartitionedDataVariant.cu:253-262, 309-313) for values that are produced and
// consumed on the device; the host never sees the payload unless it asks for it.
constexpr int kCollWords = 32;
enum CollOp { kCollGather = 0, kCollMaxI64 = 1, kCollSumI64 = 2, kCollMinI64 = 3 };
struct DevCollParams {
  unsigned* peer_flags[kMaxRanks];   // every rank's flag array of this parity; slot [rank] is ours there
  unsigned* peer_data[kMaxRanks];    // every rank's mailboxes of this parity; [rank * kCollWords ..] is ours
  const unsigned* local_flags;       // slot [r] is raised by rank r
  const unsigned* local_data;        // mailbox [r] is filled by rank r
  const unsigned* send;              // `words` 32-bit words (device)
  unsigned* recv;                    // gather: R * words words; reductions: words words (int64 pairs)
  int words, op;
  unsigned seq;
  int rank, R;
};
int dev_collective_launch(const DevCollParams& p, cudaStream_t stream);

int scatter_pairs_launch(const void* pairs, int64_t n, int64_t begin, float* out, cudaStream_t stream);

// ---------------- sort-free redistribution (redist.cu) ----------------
constexpr int kRdInfoWords = 10;   // per rank: bounds (6 order-preserving uints), n (2 words), global begin (2 words)
struct RdPlan {                    // device-resident; the host reads my_recv / max_recv once
  long long recv_offset[kMaxRanks];      // where this rank's run starts inside owner d's receive buffer (points)
  long long ret_offset[kMaxRanks];       // where this rank's run starts inside reader o's landing zone (pairs)
  long long my_recv, max_recv;
  unsigned long long cursor[kMaxRanks];      // running positions inside those runs
  unsigned long long ret_cursor[kMaxRanks];
};
struct RdSplitParams {
  const unsigned* histA[kMaxRanks];      // every rank's histograms (peer pointers)
  const unsigned* histB[kMaxRanks];
  int R, rank;
  long long n_local;
  uint64_t* split_keys;                  // out [R-1]: owner d takes keys in [split[d-1], split[d])
  long long* row;                        // out [R]: points this rank sends to each owner
};
struct RdScatterParams {
  const float* xyz;
  int64_t n, global_begin;
  int R;
  const unsigned* info_all;              // [R][kRdInfoWords]
  const uint64_t* split_keys;
  RdPlan* plan;
  Float4* recv[kMaxRanks];               // owners' receive buffers (peer pointers)
};
struct RdReturnParams {
  const Float4* pts;                     // local sorted points, .w = global id
  const float* res;                      // their k-th distances
  int64_t n;
  int R;
  const long long* begins;               // [R+1] global id ranges of the readers
  RdPlan* plan;
  void* dst[kMaxRanks];                  // readers' landing zones of (id, value) pairs (peer pointers)
};
// After a redistribution rank s owns exactly the points whose Morton key lies in [split[s-1], split[s]): a union
// of at most kTerritoryCells octree cells. rd_territory writes them as (slightly enlarged) boxes, identically on
// every rank and without communication. A group whose ball misses all of rank s's cells cannot have a neighbour
// there, whatever s's bounding boxes say — and the boxes of a Morton-ordered tree do say silly things: a node
// that straddles a jump of the Z-curve has a box that covers other ranks' whole territory.
constexpr int kTerritoryCells = 128;
int rd_territory_launch(const uint64_t* d_split_keys, const unsigned* d_info_all, int R, Box* d_cells, int* d_counts,
                        cudaStream_t stream);
size_t rd_hist_bytes();                  // histogram A followed by histogram B
size_t rd_histB_offset();
int rd_pack_info_launch(const unsigned* d_bounds_ord, int64_t n, int64_t begin, unsigned* d_out, cudaStream_t stream);
int rd_hist_launch(const float* d_xyz, int64_t n, const unsigned* d_info_all, int R, unsigned* histA, unsigned* histB,
                   int num_sms, cudaStream_t stream);
int rd_split_launch(const RdSplitParams& p, cudaStream_t stream);
int rd_plan_launch(const long long* d_M, int R, int rank, RdPlan* d_plan, cudaStream_t stream);
int rd_scatter_launch(const RdScatterParams& p, int num_sms, cudaStream_t stream);
int rd_return_launch(const RdReturnParams& p, int num_sms, cudaStream_t stream);

// out[id(i)] = sorted_result[i]  (id taken from pts[i].w, minus id_offset)
int scatter_results(const Float4* sorted_pts, const float* sorted_result, int64_t n,
                    uint32_t id_offset, float* out, cudaStream_t stream);

}  // namespace b200knn
