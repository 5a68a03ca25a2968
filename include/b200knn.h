/* b200knn.h — C ABI of libb200knn (Blackwell-native large-scale k-nearest-neighbour distances).
 *
 * Everything here is callable from C, C++ (the cudaMpiKNN_* executables) and Python (ctypes; see
 * mpi_cuda_largescaleknn_b200/ops/_lib.py). All functions return 0 on success, non-zero on error;
 * b200knn_last_error() then holds a human-readable message (thread-local).
 *
 * Conventions
 *   - points / queries: interleaved float32 x,y,z (the reference's raw "float3" layout,
 *     /root/reference/unorderedDataVariant.cu:41-63), in DEVICE memory unless the name says cpu.
 *   - result: one float32 per query = distance to the k-th nearest point, the query's own entry
 *     included when it is part of the indexed set; +inf if fewer than k points exist; with a finite
 *     max_radius r, sqrt(r*r) if fewer than k points lie inside r
 *     (/root/reference/unorderedDataVariant.cu:89-103).
 *   - `stream` is a cudaStream_t passed as void* (0 = default stream). Calls enqueue work and return;
 *     they are asynchronous unless stated otherwise.
 */
#ifndef B200KNN_H
#define B200KNN_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200KNN_MAX_RANKS 16
#define B200KNN_IPC_BLOB_BYTES 256

typedef struct b200knn_index b200knn_index;
typedef struct b200knn_ring b200knn_ring;

const char* b200knn_last_error(void);
int b200knn_version(void);
/* Number of kernels this library has launched in the calling process (monotonic). */
uint64_t b200knn_launch_count(void);

/* ---- spatial index over one GPU's points -------------------------------------------------------
 * Role of cukd::buildTree (cudaKDTree/cukd/builder.h:108-151). One radix sort of Morton keys + a
 * 32-ary box hierarchy. `ids` (device, uint32, optional) are caller-side identifiers stored with the
 * points; when NULL, id = position + id_offset. Synchronises `stream` before returning (allocation
 * sizes are computed on the host). */
int b200knn_index_create(const float* d_xyz, int64_t n, const uint32_t* d_ids, uint32_t id_offset,
                         void* stream, b200knn_index** out);
/* Same, from x,y,z,w records (16 bytes per point) whose w holds the point's uint32 id as raw bits —
 * the layout redistributed points arrive in. */
int b200knn_index_create_xyzw(const float* d_xyzw, int64_t n, uint32_t id_offset, void* stream,
                              b200knn_index** out);

/* Planar points: d_xy is [n][2] floats (the reference library's float2 trees, cukd/cukd-math.h:61-85,298-331). The
 * index treats them as z = 0 — fmaf(0,0, fmaf(dx,dx, dy*dy)) is bit for bit the 2-D squared distance — and every
 * query entry point of this index then takes [nq][2] queries. */
int b200knn_index_create_xy(const float* d_xy, int64_t n, const uint32_t* d_ids_or_null, uint32_t id_offset,
                            void* stream, b200knn_index** out);
/* Points embedded in caller records — the reference library's data traits (a struct with a position and any
 * payload, cukd/data.h:56-112; testing/testPayloadSampleFromReadme.cu): record i starts at
 * d_records + i * record_bytes and its position is `dims` (2 or 3) floats at byte offset position_offset. Nothing
 * is copied besides the positions; the ids that fcp / neighbors return are record indices (+ id_offset, or
 * d_ids[i]), so the payload of a hit is records[id]. Sizes and offsets must be multiples of 4 bytes. Queries are
 * plain [nq][dims] floats. */
int b200knn_index_create_records(const void* d_records, int64_t n, size_t record_bytes, size_t position_offset,
                                 int dims, const uint32_t* d_ids_or_null, uint32_t id_offset, void* stream,
                                 b200knn_index** out);
void b200knn_index_destroy(b200knn_index* idx);
int64_t b200knn_index_num_points(const b200knn_index* idx);
int b200knn_index_num_levels(const b200knn_index* idx);
/* Leading bits of the Morton key the build sorted by: 40 normally, 64 when some 40-bit cell held more than 1024 points
 * (the build then runs a second time on the full key) or B200KNN_SORT_BITS=64 is set. */
int b200knn_index_sort_bits(const b200knn_index* idx);
/* host float[6]: lo.xyz, hi.xyz of the indexed points (synchronous). */
int b200knn_index_bounds(const b200knn_index* idx, float* host_bounds6);
/* Debug/test helpers: copy the sorted points (x,y,z,id-bits as float4) or one level of boxes to
 * host memory (synchronous). */
int b200knn_index_copy_sorted(const b200knn_index* idx, float* host_xyzw);
int b200knn_index_copy_boxes(const b200knn_index* idx, int level, float* host_boxes8, int* count);

/* ---- single-GPU queries ------------------------------------------------------------------------
 * knn_self: every indexed point is a query (the reference's use case); d_out[i] belongs to the i-th
 * point handed to index_create. Role of runQuery + extractFinalResult
 * (/root/reference/unorderedDataVariant.cu:75-103). */
int b200knn_knn_self(b200knn_index* idx, int k, float max_radius, float* d_out, void* stream);
/* knn_query: arbitrary query points against the index (cukd::stackFree::knn, cukd/knn.h:617-631).
 * Synchronises `stream` (temporary buffers). */
int b200knn_knn_query(b200knn_index* idx, const float* d_queries, int64_t nq, int k, float max_radius,
                      float* d_out, void* stream);
/* fcp_query: find-closest-point. For every query the id of the nearest indexed point (the id given to
 * index_create: position + id_offset, or ids[]) and optionally its distance; -1 / +inf (or max_radius)
 * if no point lies within max_radius. Ties are broken arbitrarily. Role of cukd::stackFree::fcp
 * (cudaKDTree/cukd/fcp.h:210-254). Synchronises `stream`. */
int b200knn_fcp_query(b200knn_index* idx, const float* d_queries, int64_t nq, float max_radius,
                      int32_t* d_out_ids, float* d_out_dist_or_null, void* stream);

/* Find-closest-point with the reference's search parameters (cukd/fcp.h:35-52, FcpSearchParams):
 *   max_radius               cutOffRadius: points farther than this are ignored (+inf: none)
 *   eps                      approximate search: a subtree is skipped when its lower bound times (1 + eps) is not
 *                            below the best distance so far, so the reported distance is at most (1 + eps)^(1/2)
 *                            times the true one (0 = exact; role of the eps of cukd/traverse-stack-free.h:32,85)
 *   far_node_inspect_budget  after the leaves next to the query, at most this many further tree nodes are
 *                            inspected per group of 32 queries (< 0 = no limit; the reference counts "far"
 *                            subtrees of its binary tree, here a node is a 32-ary box-tree node or a leaf) */
typedef struct b200knn_fcp_params {
  float max_radius;
  float eps;
  int32_t far_node_inspect_budget;
} b200knn_fcp_params;
int b200knn_fcp_query_ex(b200knn_index* idx, const float* d_queries, int64_t nq, const b200knn_fcp_params* params,
                         int32_t* d_out_ids, float* d_out_dist, void* stream);
/* knn_neighbors: the k nearest indexed points of every query — their ids (as fcp_query reports them)
 * and distances — in d_out_ids / d_out_dist, both [nq][k] row-major. Rows are NOT ordered by distance
 * (like the reference's heap-ordered candidate lists, cukd/knn.h:311-326); slots beyond the points
 * available within max_radius hold id -1 / distance +inf. Which of several points at exactly the k-th
 * distance are reported is unspecified. d_queries == NULL: the indexed points themselves are the
 * queries (nq is ignored; rows in the order the points were handed to index_create, which must have
 * been called without explicit ids). Two passes: the k-th-distance kernel, then a fixed-radius
 * collection; no (distance, id) heap is kept. Synchronises `stream`. */
int b200knn_knn_neighbors(b200knn_index* idx, const float* d_queries_or_null, int64_t nq, int k, float max_radius,
                          int32_t* d_out_ids, float* d_out_dist, void* stream);
/* count_within: for up to 64 queries, how many of the n points lie strictly closer than / no farther than
 * radius[j] (distances evaluated exactly as the query kernels do, square root included). A result r of a
 * k-th-neighbour query is right if and only if less[j] < k <= less_equal[j], so this verifies sampled results
 * against ALL points in one pass, at sizes where no brute-force selection is affordable; counts of several
 * ranks add up. d_less / d_less_equal: device, nq uint64 each (zeroed by the call). Role of verifyKNN
 * (cudaKDTree/testing/floatN-knn-and-fcp.cu:292-335). */
int b200knn_count_within(const float* d_xyz, int64_t n, const float* d_queries, int nq, const float* d_radius,
                         uint64_t* d_less, uint64_t* d_less_equal, void* stream);
/* knn_self_noalloc: same result as index_create + knn_self for n <= B200KNN_NOALLOC_MAX_POINTS, but
 * the library allocates nothing: the sorted points, the box tree, the sort buffers and the counters
 * all live in one fixed device array inside the library (about 11.9 MB per GPU that loads it, held
 * until exit). Meant for callers that may not allocate device memory once they are running — a host
 * application that has handed the whole GPU to its own arena, a loop that cannot tolerate the
 * device-wide stalls cudaMalloc/cudaFree can cause, a plug-in whose host forbids allocator calls
 * after start-up. The caller passes its own input and output buffers. Larger n is an error. The
 * first call on a device resolves the array's address (cudaGetSymbolAddress), so make that call
 * before starting any stream capture. All calls share the array and it keeps the previous call's
 * contents; whatever must start at zero is cleared by each call, and calls whose GPU work overlaps
 * (two streams, two threads) race — keep them on one stream. Enqueues on `stream` and returns.
 * (No counterpart in the reference.) */
#define B200KNN_NOALLOC_MAX_POINTS 262144
int b200knn_knn_self_noalloc(const float* d_xyz, int64_t n, int k, float max_radius, float* d_out,
                             void* stream);
int64_t b200knn_noalloc_max_points(void);
int64_t b200knn_noalloc_scratch_bytes(void);
/* Traversal counters of the most recent query call on this index (synchronous):
 * [0] leaves visited by groups, [1] candidate distance evaluations per lane, [2] heap insertions,
 * [3] inner nodes opened, [4] frontier overflows handled depth-first, [5] frontier purges. Role of CUKD_ENABLE_STATS (cukd/common.h:349-361). */
int b200knn_index_stats(b200knn_index* idx, uint64_t out8[8]);
int b200knn_index_enable_stats(b200knn_index* idx, int on);
/* Diagnostics of the last query with stats enabled: out8[7] of b200knn_index_stats is the slowest query group as
   (kilo-cycles << 32) | group index; this returns what that one group did: leaves swept, inner nodes opened,
   list flushes, depth-first fallbacks, frontier purges, candidates accepted by its first lane, 0, 0.
   (No counterpart in the reference; its CUKD_STATS counts traversal steps only, cukd/common.h:349-361.) */
int b200knn_index_stats_slowest(b200knn_index* idx, uint64_t out8[8]);
/* Leaf tiles on the tensor cores, as a measured experiment (csrc/brute_tile.cu; no
   counterpart in the reference): for nq queries and np points, how many points lie below each query's threshold
   (squared distance), and optionally which (masks: [nq][ceil(np/32)] words, bit j of word l = point 32 l + j).
   _cuda is the product's exact CUDA-core sweep; _tc is TMA + tcgen05.mma kind::tf32 + TMEM with a conservative error
   bound: its masks are supersets of the exact ones. _prep turns xyz into the [n][8] operand rows (re-centred on
   `centre`) and norms that _tc takes. */
int b200knn_brute_tile_prep(const float* d_xyz, int64_t n, const float centre[3], int is_query, float* d_rows8,
                            float* d_norm, void* stream);
int b200knn_brute_tile_cuda(const float* d_q_xyz, int64_t nq, const float* d_p_xyz, int np, const float* d_thr,
                            unsigned* d_counts, unsigned* d_masks_or_null, void* stream);
int b200knn_brute_tile_tc(float* d_q_rows8, const float* d_q_norm, int64_t nq, float* d_p_rows8, int np, float p_norm_max,
                          const float* d_thr, unsigned* d_counts, unsigned* d_masks_or_null, int num_ctas, void* stream);
/* The stream-ordered pool this library allocates from (the current device's default pool): bytes reserved right
   now and the high-water mark of bytes in use; reset_high != 0 restarts the high-water mark at zero afterwards.
   Lets a caller (and tests/test_zz_no_alloc.py) see whether a call allocated. Role of the reference's
   GpuMemoryResource bookkeeping, cukd/helpers.h:31-71, which has no such query. */
int b200knn_mempool_usage(int64_t* reserved_now, int64_t* used_high, int reset_high);

/* ---- multi-GPU fused ring ----------------------------------------------------------------------
 * One process per GPU. Query groups and their k-candidate heaps travel rank -> rank+1; every GPU's
 * tree stays home. The kernel writes into the next GPU's inbox over NVLink and signals with
 * system-scope release/acquire flags; there is no host or NCCL call between hops.
 * Role of the MPI ring in /root/reference/unorderedDataVariant.cu:173-205.
 *
 * Call order on every rank:
 *   ring_create -> ring_export -> (allgather the blobs by any means) -> ring_connect -> barrier
 *   -> ring_run -> stream sync -> barrier -> ring_results
 * `max_points_any_rank` must be identical on all ranks. `groups_per_chunk` (0 = default) bounds the
 * in-flight candidate memory: 2 * groups_per_chunk * (k*128 + 516) bytes per GPU. */
int b200knn_ring_create(b200knn_index* idx, int rank, int world, int k, int64_t max_points_any_rank,
                        int groups_per_chunk, b200knn_ring** out);
/* Tear-down on every rank: ring_disconnect -> barrier -> ring_destroy (peers must unmap before the
 * owner frees). ring_destroy alone is fine for world == 1. */
int b200knn_ring_disconnect(b200knn_ring* ring);
void b200knn_ring_destroy(b200knn_ring* ring);
int b200knn_ring_export(b200knn_ring* ring, void* blob /* B200KNN_IPC_BLOB_BYTES */);
int b200knn_ring_connect(b200knn_ring* ring, const void* all_blobs /* world * B200KNN_IPC_BLOB_BYTES */);
int b200knn_ring_run(b200knn_ring* ring, float max_radius, void* stream);
int b200knn_ring_results(b200knn_ring* ring, float* d_out, void* stream);
int64_t b200knn_ring_comm_bytes(const b200knn_ring* ring);

/* Low-level: run a range of ring steps with caller-provided buffers and NO in-kernel handshake
 * (use_flags = 0). This is the baseline transport path: the host moves outbox -> peer inbox with
 * NCCL between launches. Buffer shapes: heap [groups_per_chunk][k][32] float, pos
 * [groups_per_chunk][32] float4, cnt [groups_per_chunk] int32. */
typedef struct b200knn_step_args {
  int k, world, rank;
  int groups_per_chunk, num_chunks;
  int step_begin, step_end;
  float max_radius;
  const float* inbox_heap[2];
  const float* inbox_pos[2];
  const int32_t* inbox_cnt[2];
  float* outbox_heap[2];
  float* outbox_pos[2];
  int32_t* outbox_cnt[2];
  float* result_sorted;   /* receives k-th distances of groups finishing their last hop here */
} b200knn_step_args;
int b200knn_run_steps(b200knn_index* idx, const b200knn_step_args* args, void* stream);
/* out[id - id_offset] = sorted_result[i] for the index's own points (undo the spatial order). */
int b200knn_index_unsort(b200knn_index* idx, const float* d_sorted_result, float* d_out, void* stream);

/* ---- distributed driver ------------------------------------------------------------------------
 * The body of the reference's two main() functions, minus file I/O: every rank calls dist_knn with
 * the points it holds and gets the k-th-neighbour distance over the UNION of all ranks' points.
 *
 * The library never talks to a launcher or a network itself. The caller supplies a bootstrap: two
 * blocking host-side collectives over small buffers (torch.distributed in Python, the unix-socket
 * rendezvous of apps/knnrun in C++). They are used for set-up (CUDA IPC handles, boxes, counts) and
 * as round fences; all bulk data moves by kernels over NVLink-mapped peer memory.
 * Role of MPI_Init/Allgather/Allreduce/Barrier in the reference
 * (/root/reference/prePartitionedDataVariant.cu:254-319). */
typedef struct b200knn_bootstrap {
  void* ctx;
  int rank, world;
  /* recv holds world * bytes_per_rank bytes, in rank order. Return 0 on success. */
  int (*allgather)(void* ctx, const void* send, void* recv, int64_t bytes_per_rank);
  int (*barrier)(void* ctx);
} b200knn_bootstrap;

typedef struct b200knn_dist b200knn_dist;

enum {
  B200KNN_MODE_RING = 0,            /* all-pairs fused ring; any input (unorderedDataVariant.cu:173-205) */
  B200KNN_MODE_PREPARTITIONED = 1,  /* keep points where they are; bbox-culled per-group exchange
                                       (prePartitionedDataVariant.cu:304-357) */
  B200KNN_MODE_REDISTRIBUTE = 2,    /* scatter points to spatial owners first, then as mode 1, then
                                       return results to the reading rank (no counterpart in the
                                       reference; makes unordered input scale with GPU count) */
  B200KNN_MODE_AUTO = 3             /* per-rank inputs: mode 1 if the ranks' bounding boxes are nearly
                                       disjoint (pairwise overlap < 2 % of a box), else mode 2 on the
                                       concatenation of the inputs (global_begin is computed) */
};

typedef struct b200knn_dist_report {
  double ms_total, ms_redistribute, ms_build, ms_query, ms_return;   /* host wall clock, this rank */
  int64_t pending_groups;           /* home groups that had to visit at least one other rank */
  int64_t points_after_redistribute;
  int32_t chunks, rounds;
  double ms_local_phase, ms_host_collectives, ms_rounds;   /* breakdown of ms_query in modes 1 and 2 */
  int32_t auto_chose_redistribute;  /* mode 3 only: 1 if the inputs overlapped and were redistributed */
  int32_t reserved;
} b200knn_dist_report;

/* Collective over all ranks of the bootstrap. The engine keeps a symmetric, IPC-mapped heap that
 * grows on demand and is reused across dist_knn calls. */
int b200knn_dist_create(const b200knn_bootstrap* bs, b200knn_dist** out);
void b200knn_dist_destroy(b200knn_dist* d);
int64_t b200knn_dist_heap_bytes(const b200knn_dist* d);
/* d_out[i] = k-th-neighbour distance of local point i. `global_begin` = index of local point 0 in the
 * global order (mode 2 only: ranks must hold consecutive slices, as the reference's file split does).
 * groups_per_chunk = 0 picks the default. Synchronous: returns when d_out is complete. */
int b200knn_dist_knn(b200knn_dist* d, const float* d_xyz, int64_t n_local, int64_t global_begin, int k,
                     float max_radius, int mode, int groups_per_chunk, float* d_out, void* stream,
                     b200knn_dist_report* report_or_null);

/* Small device-side collectives over the engine's symmetric heap: payload and result live in device
 * memory, the exchange is one kernel enqueued on `stream` (peer stores + sequence flags), the host
 * does not wait. Role of MPI_Allgather / MPI_Allreduce on counts, boxes and radii
 * (prePartitionedDataVariant.cu:253-262, 309-313) for values produced and consumed on the GPU.
 * Collective: every rank calls with the same sizes, in the same order, one GPU per rank.
 *   allgather_dev: d_recv gets world * nbytes bytes in rank order; nbytes a multiple of 4, <= 128.
 *   allreduce_dev: element-wise over `count` <= 16 int64 values. */
#define B200KNN_DEV_COLLECTIVE_MAX_BYTES 128
enum { B200KNN_OP_MAX = 0, B200KNN_OP_SUM = 1, B200KNN_OP_MIN = 2 };
int b200knn_dist_allgather_dev(b200knn_dist* d, const void* d_send, int nbytes, void* d_recv, void* stream);
int b200knn_dist_allreduce_dev(b200knn_dist* d, const int64_t* d_send, int count, int op, int64_t* d_recv,
                               void* stream);

/* ---- CPU side (no GPU needed): oracle and mirror engine ----------------------------------------
 * cpu_bruteforce: exact O(nq*n) k-th distance with the canonical fp32 expression
 *   fmaf(dz,dz, fmaf(dx,dx, dy*dy)), d = query - point  (SURVEY.md fact 7); the test oracle.
 *   Role of verifyKNN (cudaKDTree/testing/floatN-knn-and-fcp.cu:292-335).
 * cpu_knn: the same index + group traversal as the GPU path, on the host (role of
 *   buildTree_host, cukd/builder_thrust.h:514-601); used as CPU backend and for cross-checks. */
int b200knn_cpu_bruteforce(const float* xyz, int64_t n, const float* queries, int64_t nq, int k,
                           float max_radius, float* out, int num_threads);
int b200knn_cpu_knn(const float* xyz, int64_t n, const float* queries_or_null, int64_t nq, int k,
                    float max_radius, float* out, int num_threads);
/* cpu_neighbors: brute-force oracle of knn_neighbors; rows sorted by (distance, id). */
int b200knn_cpu_neighbors(const float* xyz, int64_t n, const float* queries, int64_t nq, int k, float max_radius,
                          int32_t* out_ids, float* out_dist, int num_threads);

/* Traversal counters accumulated by cpu_knn calls (same meaning as b200knn_index_stats). */
int b200knn_cpu_stats(uint64_t out4[4], int reset);

#ifdef __cplusplus
}
#endif
#endif /* B200KNN_H */
