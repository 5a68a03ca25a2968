// dist.cu — the distributed driver (host side): one process per GPU, peer memory mapped with CUDA
// IPC into a symmetric heap, three strategies over the same kernels.
//
// Role of main() in the reference's two programs (/root/reference/unorderedDataVariant.cu:137-211,
// /root/reference/prePartitionedDataVariant.cu:211-364) with MPI replaced by
//   * a caller-supplied bootstrap (allgather + barrier of small host buffers: torch.distributed in
//     Python, a unix-socket rendezvous in the C++ launcher) for set-up and round fences, and
//   * in-kernel loads/stores on NVLink-mapped peer memory for every byte of bulk data.
//
// Modes
//   RING            all-pairs ring, query groups + heaps circulate (query.cu); valid for any input
//   PREPARTITIONED  every rank keeps the points it was given; per-group bounding-box culling decides
//                   which groups visit which ranks (pv.cu)
//   REDISTRIBUTE    unordered input is first scattered to spatial owners (global Morton order, sample
//                   splitters, all-to-all by peer stores), then PREPARTITIONED runs on compact,
//                   nearly disjoint regions, then results are scattered back to the reading rank.
//                   This is what makes N fixed / more GPUs faster (the ring does R times the work).
#include <string.h>
#include <algorithm>
#include <chrono>
#include <vector>
#include "common.cuh"
#include "index_internal.h"

namespace {

struct HeapBlob {
  cudaIpcMemHandle_t handle;
  int64_t bytes;
  int32_t rank, world;
};

inline size_t align256(size_t v) { return (v + 255) / 256 * 256; }

double now_ms() {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// Opt-in step-by-step timing of the distributed pipeline (B200KNN_TRACE_PHASES=1): every mark
// synchronises the stream and prints the time since the previous mark to stderr, so it perturbs the run
// it measures — a diagnosis tool, not a benchmark. Without the variable a mark is one predictable branch.
struct PhaseTrace {
  bool on;
  int rank;
  cudaStream_t stream;
  double last;
  PhaseTrace(int rank_, cudaStream_t s) : rank(rank_), stream(s) {
    static const bool enabled = getenv("B200KNN_TRACE_PHASES") != nullptr;
    on = enabled;
    last = on ? now_ms() : 0.0;
  }
  void mark(const char* what) {
    if (!on) return;
    cudaStreamSynchronize(stream);
    const double t = now_ms();
    fprintf(stderr, "[b200knn phase] rank %d %-28s %8.3f ms\n", rank, what, t - last);
    last = t;
  }
};

}  // namespace

struct b200knn_dist {
  b200knn_bootstrap bs{};
  int rank = 0, world = 1, device = 0, num_sms = 148;
  char* base = nullptr;
  size_t bytes = 0;
  char* peer[B200KNN_MAX_RANKS] = {};
  size_t bump = 0;
  unsigned barrier_seq = 0;     // device barrier sequence number; identical on all ranks
  unsigned coll_seq = 0;        // device collective sequence number; identical on all ranks
  unsigned long long* task_counter = nullptr;

  int allgather(const void* send, void* recv, int64_t nbytes) {
    if (world == 1) { memcpy(recv, send, (size_t)nbytes); return 0; }
    return bs.allgather(bs.ctx, send, recv, nbytes);
  }
  int barrier() { return world == 1 ? 0 : bs.barrier(bs.ctx); }
  size_t alloc(size_t n) { const size_t o = bump; bump = align256(bump + n); return o; }
};

namespace {

// Collective. Every rank first unmaps its peers' heaps, then (after a fence) frees its own: an
// exporter must not free memory that an importer still has mapped.
int close_heap(b200knn_dist* d) {
  for (int r = 0; r < d->world; ++r) {
    if (r != d->rank && d->peer[r]) cudaIpcCloseMemHandle(d->peer[r]);
    d->peer[r] = nullptr;
  }
  if (d->world > 1 && d->base) d->barrier();
  if (d->base) cudaFree(d->base);
  d->base = nullptr;
  d->bytes = 0;
  return 0;
}

// Collective: every rank must call with the same `need`.
int ensure_heap(b200knn_dist* d, size_t need) {
  need = align256(need);
  if (need <= d->bytes) return 0;
  KNN_CUDA_TRY(cudaDeviceSynchronize());
  if (d->barrier()) { set_last_error("bootstrap barrier failed"); return -3; }
  close_heap(d);
  const size_t grow = need + need / 8;
  KNN_CUDA_TRY(cudaMalloc(&d->base, grow));
  // the zeroing runs on the legacy default stream, the kernels that use the flags run on the caller's
  // (possibly non-blocking) stream and on other GPUs: make it complete before anybody can see the heap
  KNN_CUDA_TRY(cudaMemset(d->base, 0, grow));
  KNN_CUDA_TRY(cudaDeviceSynchronize());
  d->bytes = grow;
  d->peer[d->rank] = d->base;
  if (d->world > 1) {
    HeapBlob mine;
    memset(&mine, 0, sizeof(mine));
    KNN_CUDA_TRY(cudaIpcGetMemHandle(&mine.handle, d->base));
    mine.bytes = (int64_t)grow;
    mine.rank = d->rank;
    mine.world = d->world;
    std::vector<HeapBlob> all(d->world);
    if (d->allgather(&mine, all.data(), sizeof(HeapBlob))) { set_last_error("bootstrap allgather failed"); return -3; }
    for (int r = 0; r < d->world; ++r) {
      if (r == d->rank) continue;
      KNN_REQUIRE(all[r].rank == r && all[r].bytes == (int64_t)grow, "ranks disagree on the symmetric heap size");
      void* p = nullptr;
      KNN_CUDA_TRY(cudaIpcOpenMemHandle(&p, all[r].handle, cudaIpcMemLazyEnablePeerAccess));
      d->peer[r] = static_cast<char*>(p);
    }
    if (d->barrier()) { set_last_error("bootstrap barrier failed"); return -3; }
  }
  return 0;
}

// The first bytes of the symmetric heap persist across calls; everything after them is re-carved by
// every call:
//   [0, 256)       barrier flags, one unsigned per peer
//   [256, 512)     collective flags, [parity][peer]
//   [512, 4608)    collective mailboxes, [parity][peer][kCollWords]
constexpr size_t kCollFlagsOff = 256;
constexpr size_t kCollDataOff = 512;
constexpr size_t kHeapReserved = kCollDataOff + 2 * (size_t)kMaxRanks * kCollWords * sizeof(unsigned);
static_assert(2 * kMaxRanks * sizeof(unsigned) <= kCollDataOff - kCollFlagsOff, "collective flags overflow their area");
static_assert(kHeapReserved % 256 == 0, "reserved area keeps the 256-byte alignment of what follows");

// Exchange slots (pending + inbox) exist only when there is somebody to exchange with.
inline int64_t pv_slots(int Gc, int R) { return R > 1 ? Gc : 1; }

// The one-round exchange (requests out, candidates back; pv.cu SERVE_FRESH / MERGE) runs whenever the lists are
// sorted columns and a chunk's requests fit the areas below; otherwise that chunk takes the R - 1 sequential rounds.
inline bool pv_one_round(int k, int R) { return R > 1 && pv_one_round_capable(k) && getenv("B200KNN_SEQ_EXCHANGE") == nullptr; }
inline int64_t pv_cap_req(int Gc) {                                             // per (source, destination) pair
  if (const char* e = getenv("B200KNN_REQ_CAP")) return std::max<int64_t>(1, atoll(e));   // tests: force the fallback
  return std::max<int64_t>(256, Gc / 4);
}
inline int64_t pv_cap_ret(int Gc) { return std::max<int64_t>(256, Gc / 2); }    // per home rank

size_t pv_heap_need(int k, int Gc, int64_t max_groups, int R) {
  size_t need = kHeapReserved + 2 * align256((size_t)pv_slots(Gc, R) * pv_slot_bytes(k)) + 6 * 256 +
                align256((size_t)R * tops_capacity(max_groups) * sizeof(Box)) + align256((size_t)kMaxRanks * sizeof(TopHead));
  if (pv_one_round(k, R))
    need += align256((size_t)R * pv_cap_req(Gc) * pv_request_bytes()) + align256((size_t)pv_cap_ret(Gc) * pv_return_bytes(k)) + 256;
  return need;
}

// Device-side barrier over all ranks, enqueued on `stream` (replaces MPI_Barrier,
// /root/reference/prePartitionedDataVariant.cu:305). Work enqueued before it on every rank has
// completed — including its stores into peer memory — before work enqueued after it starts anywhere.
int device_barrier(b200knn_dist* d, cudaStream_t stream) {
  if (d->world == 1) return 0;
  DevBarrierParams bp;
  memset(&bp, 0, sizeof(bp));
  bp.R = d->world;
  bp.rank = d->rank;
  bp.seq = ++d->barrier_seq;
  bp.local_flags = reinterpret_cast<const unsigned*>(d->base);
  for (int r = 0; r < d->world; ++r) bp.peer_flags[r] = reinterpret_cast<unsigned*>(d->peer[r]);
  return dev_barrier_launch(bp, stream);
}

// All-gather (op = kCollGather: recv gets R * words words) or int64 all-reduce (recv gets words / 2
// elements) of device-resident values, enqueued on `stream`. Collective; same arguments on every rank.
int device_collective(b200knn_dist* d, const void* d_send, int words, int op, void* d_recv, cudaStream_t stream) {
  DevCollParams cp;
  memset(&cp, 0, sizeof(cp));
  cp.R = d->world;
  cp.rank = d->rank;
  cp.words = words;
  cp.op = op;
  cp.seq = ++d->coll_seq;
  const size_t par = cp.seq & 1u;
  const size_t foff = kCollFlagsOff + par * kMaxRanks * sizeof(unsigned);
  const size_t doff = kCollDataOff + par * (size_t)kMaxRanks * kCollWords * sizeof(unsigned);
  for (int r = 0; r < d->world; ++r) {
    cp.peer_flags[r] = reinterpret_cast<unsigned*>(d->peer[r] + foff);
    cp.peer_data[r] = reinterpret_cast<unsigned*>(d->peer[r] + doff);
  }
  cp.local_flags = reinterpret_cast<const unsigned*>(d->base + foff);
  cp.local_data = reinterpret_cast<const unsigned*>(d->base + doff);
  cp.send = static_cast<const unsigned*>(d_send);
  cp.recv = static_cast<unsigned*>(d_recv);
  return dev_collective_launch(cp, stream);
}

constexpr int kSplitShift = 32;   // redistribution: keys are compared by their upper 64 - kSplitShift bits only

struct PvPlan { int Gc; int C; int64_t max_groups; };

// Destroys a local index on every path out of a function.
struct IndexGuard {
  b200knn_index* p = nullptr;
  ~IndexGuard() { b200knn_index_destroy(p); }
};

PvPlan pv_plan(int64_t max_n, int k, int groups_per_chunk) {
  PvPlan pl;
  pl.max_groups = div_up(max_n > 0 ? max_n : 1, kGroup);
  // default: as many home groups per chunk as a 32 GB slot budget (pending + inbox) allows
  int64_t Gc = groups_per_chunk > 0 ? groups_per_chunk : (int64_t)((16ull << 30) / pv_slot_bytes(k));
  if (Gc > pl.max_groups) Gc = pl.max_groups;
  pl.Gc = (int)Gc;
  pl.C = (int)div_up(pl.max_groups, Gc);
  return pl;
}

// The culled exchange over an existing local index. Heap must already be large enough; sub-allocations
// are taken from d->bump (symmetric: same sequence on every rank). Results land in idx->result_sorted.
int run_pv(b200knn_dist* d, b200knn_index* idx, int k, float max_radius, const PvPlan& pl, cudaStream_t stream,
           b200knn_dist_report* rep, const Box* territory = nullptr, const int* territory_count = nullptr) {
  NvtxRange range("b200knn:culled_exchange");
  const int R = d->world, me = d->rank;
  PhaseTrace trace(me, stream);
  // everyone's upper tree levels, replicated on every rank by peer stores (pv.cu, tops_publish_kernel); a
  // device barrier makes the tables complete. No host collective.
  const int64_t tops_cap = tops_capacity(pl.max_groups);
  const size_t off_tops = d->alloc((size_t)R * tops_cap * sizeof(Box));
  const size_t off_heads = d->alloc((size_t)kMaxRanks * sizeof(TopHead));
  KNN_REQUIRE(d->bump <= d->bytes, "symmetric heap too small (internal sizing error)");
  TopsView tops;
  tops.boxes = reinterpret_cast<const Box*>(d->base + off_tops);
  tops.heads = reinterpret_cast<const TopHead*>(d->base + off_heads);
  tops.cap = tops_cap;
  if (R > 1) {
    TopsPublishParams tp;
    memset(&tp, 0, sizeof(tp));
    tp.tree = idx->view();
    memcpy(tp.overall, idx->h_bounds, sizeof(tp.overall));
    tp.R = R;
    tp.rank = me;
    tp.cap = tops_cap;
    for (int r = 0; r < R; ++r) {
      tp.tables[r] = reinterpret_cast<Box*>(d->peer[r] + off_tops);
      tp.heads[r] = reinterpret_cast<TopHead*>(d->peer[r] + off_heads);
    }
    int rcb;
    if ((rcb = tops_publish_launch(tp, idx->num_sms, stream))) return rcb;
    if ((rcb = device_barrier(d, stream))) return rcb;
    trace.mark("exchange: publish tree tops + barrier");
  }

  const size_t slot = pv_slot_bytes(k);
  const size_t off_pending = d->alloc((size_t)pv_slots(pl.Gc, R) * slot);
  const size_t off_inbox = d->alloc((size_t)pv_slots(pl.Gc, R) * slot);
  const size_t off_pcount = d->alloc((1 + kMaxRanks) * sizeof(int));   // pending count, then requests posted per rank
  const size_t off_icount = d->alloc(4);
  const size_t off_ocount = d->alloc(4);
  const bool one_round = pv_one_round(k, R);
  const int64_t cap_req = pv_cap_req(pl.Gc), cap_ret = pv_cap_ret(pl.Gc);
  size_t off_req = 0, off_ret = 0, off_retcount = 0;
  if (one_round) {
    off_req = d->alloc((size_t)R * cap_req * pv_request_bytes());
    off_ret = d->alloc((size_t)cap_ret * pv_return_bytes(k));
    off_retcount = d->alloc(4);
  }
  KNN_REQUIRE(d->bump <= d->bytes, "symmetric heap too small (internal sizing error)");

  PvMainParams p;
  memset(&p, 0, sizeof(p));
  p.tree = idx->view();
  p.k = k;
  p.sentinel = sentinel_d2(max_radius);
  p.R = R;
  p.rank = me;
  p.tops = tops;
  p.territory = territory;
  p.territory_count = territory_count;
  p.home_queries = idx->pts;
  p.home_n = idx->n;
  p.home_is_tree = 1;
  p.fill = convergent_fill_enabled() ? 1 : 0;
  p.result_sorted = idx->result_sorted;
  p.pending = d->base + off_pending;
  p.pending_count = reinterpret_cast<int*>(d->base + off_pcount);
  p.inbox = d->base + off_inbox;
  p.inbox_count = reinterpret_cast<const int*>(d->base + off_icount);
  p.task_counter = d->task_counter;
  p.stats = idx->stats_on ? idx->stats : nullptr;
  if (one_round) {
    p.fused = 1;
    p.cap_req = cap_req;
    p.cap_ret = cap_ret;
    p.req_out_count = p.pending_count + 1;
    p.my_req_inbox = d->base + off_req;
    p.my_ret_area = d->base + off_ret;
    p.my_ret_count = reinterpret_cast<const int*>(d->base + off_retcount);
    for (int r = 0; r < R; ++r) {
      p.req_inbox[r] = d->peer[r] + off_req;
      p.ret_area[r] = d->peer[r] + off_ret;
      p.ret_count[r] = reinterpret_cast<int*>(d->peer[r] + off_retcount);
    }
  }
  const int vote_words = one_round ? 1 + R : 1;
  int* out_count = reinterpret_cast<int*>(d->base + off_ocount);
  const int64_t my_groups = idx->groups();
  if (idx->stats_on) KNN_CUDA_TRY(cudaMemsetAsync(idx->stats, 0, 16 * sizeof(unsigned long long), stream));

  // the per-chunk vote "does anybody have groups that must travel?" is a device-side all-gather of
  // the pending counters (B200KNN_HOST_VOTE=1: through the bootstrap, as a cross-check)
  const bool host_vote = getenv("B200KNN_HOST_VOTE") != nullptr;
  StreamBuf<int> d_votes(stream);
  if (R > 1 && !host_vote) KNN_CUDA_TRY(d_votes.alloc_bytes((size_t)R * vote_words * sizeof(int)));

  int rc;
  for (int c = 0; c < pl.C; ++c) {
    KNN_CUDA_TRY(cudaMemsetAsync(p.pending_count, 0, (size_t)vote_words * sizeof(int), stream));
    // nobody adds to this counter before the vote below has seen every rank finish its local phase
    if (one_round) KNN_CUDA_TRY(cudaMemsetAsync(d->base + off_retcount, 0, sizeof(int), stream));
    p.phase = 0;
    p.group_begin = (int64_t)c * pl.Gc;
    int64_t gc = my_groups - p.group_begin;
    gc = gc < 0 ? 0 : (gc > pl.Gc ? pl.Gc : gc);
    p.group_count = (int)gc;
    const double tl0 = now_ms();
    if ((rc = pv_main_launch(p, p.group_count, idx->num_sms, stream))) return rc;
    if (R == 1) continue;
    // any work for the exchange at all? (one tiny allgather per chunk saves 2(R-1) fences when not)
    // votes[r * vote_words]: rank r's pending groups; [+ 1 + s]: requests rank r posted to rank s
    std::vector<int> votes((size_t)R * vote_words), mine(vote_words);
    double tl1;
    if (host_vote) {
      SmallRead vote_read;
      if ((rc = vote_read.begin(p.pending_count, (size_t)vote_words * sizeof(int), mine.data(), stream))) return rc;
      KNN_CUDA_TRY(cudaStreamSynchronize(stream));
      vote_read.end();
      tl1 = now_ms();
      if (d->allgather(mine.data(), votes.data(), (int64_t)vote_words * sizeof(int))) { set_last_error("bootstrap allgather failed"); return -3; }
      // the device-side vote doubles as the fence that makes the posted requests visible; this one does not
      if (one_round && (rc = device_barrier(d, stream))) return rc;
    } else {
      if ((rc = device_collective(d, p.pending_count, vote_words, kCollGather, d_votes, stream))) return rc;
      SmallRead vote_read;
      if ((rc = vote_read.begin(d_votes, votes.size() * sizeof(int), votes.data(), stream))) return rc;
      KNN_CUDA_TRY(cudaStreamSynchronize(stream));
      vote_read.end();
      tl1 = now_ms();
    }
    const int my_pending = votes[(size_t)me * vote_words];
    if (rep) rep->ms_local_phase += tl1 - tl0;
    trace.mark("exchange: local phase + vote");
    int64_t total_pending = 0;
    for (int r = 0; r < R; ++r) total_pending += votes[(size_t)r * vote_words];
    if (rep) { rep->pending_groups += my_pending; rep->chunks += 1; }
    if (rep) rep->ms_host_collectives += now_ms() - tl1;
    if (total_pending == 0) continue;
    const double tr0 = now_ms();
    // one round if every rank's requests fitted their areas (every host sees the same counts, so all agree)
    bool fits = one_round;
    for (int r = 0; r < R && fits; ++r) {
      int64_t sent = 0;
      for (int s = 0; s < R; ++s) {
        const int v = votes[(size_t)r * vote_words + 1 + s];
        if (v > cap_req) fits = false;
        sent += v;
      }
      if (sent > cap_ret) fits = false;       // at most one return record per request
    }
    if (fits) {
      int64_t to_serve = 0, sent = 0;
      for (int s = 0; s < R; ++s) {
        p.req_in[s] = votes[(size_t)s * vote_words + 1 + me];
        to_serve += p.req_in[s];
        sent += votes[(size_t)me * vote_words + 1 + s];
      }
      if ((rc = pv_fused_launch(p, 2, (int)to_serve, idx->num_sms, stream))) return rc;
      if ((rc = device_barrier(d, stream))) return rc;
      if ((rc = pv_fused_launch(p, 3, (int)sent, idx->num_sms, stream))) return rc;
      if ((rc = pv_final_launch(p.pending, p.pending_count, k, idx->result_sorted, idx->num_sms, stream))) return rc;
      if (rep) { rep->rounds += 1; KNN_CUDA_TRY(cudaStreamSynchronize(stream)); rep->ms_rounds += now_ms() - tr0; }
      trace.mark("exchange: one round (serve, merge, final)");
      continue;
    }
    for (int j = 1; j < R; ++j) {
      const int dst = (me + j) % R, src = (me - j + R) % R;
      if ((rc = pv_push_launch(p.pending, p.pending_count, k, tops, territory, territory_count, dst,
                               d->peer[dst] + off_inbox, out_count, idx->num_sms, stream))) return rc;
      if ((rc = pv_publish_launch(out_count, reinterpret_cast<int*>(d->peer[dst] + off_icount), stream))) return rc;
      // round fences are device-side sequence-flag barriers, enqueued on the stream: the host does
      // not synchronise between push, serve and the next round
      if ((rc = device_barrier(d, stream))) return rc;
      p.phase = 1;
      p.src_pending = d->peer[src] + off_pending;
      if ((rc = pv_main_launch(p, pl.Gc, idx->num_sms, stream))) return rc;
      if ((rc = device_barrier(d, stream))) return rc;
      if (rep) rep->rounds += 1;
    }
    if ((rc = pv_final_launch(p.pending, p.pending_count, k, idx->result_sorted, idx->num_sms, stream))) return rc;
    if (rep) { KNN_CUDA_TRY(cudaStreamSynchronize(stream)); rep->ms_rounds += now_ms() - tr0; }
    trace.mark("exchange: rounds + final");
  }
  if (d_votes) d_votes.release();
  return 0;
}

int knn_prepartitioned(b200knn_dist* d, const float* d_xyz, int64_t n_local, int k, float max_radius,
                       int groups_per_chunk, float* d_out, cudaStream_t stream, b200knn_dist_report* rep) {
  int rc;
  int64_t max_n = 0;
  {
    std::vector<int64_t> all(d->world);
    if (d->allgather(&n_local, all.data(), sizeof(int64_t))) { set_last_error("bootstrap allgather failed"); return -3; }
    for (int64_t v : all) max_n = std::max(max_n, v);
  }
  const PvPlan pl = pv_plan(max_n, k, groups_per_chunk);
  if ((rc = ensure_heap(d, pv_heap_need(k, pl.Gc, pl.max_groups, d->world)))) return rc;
  d->bump = kHeapReserved;
  const double t0 = now_ms();
  b200knn_index* idx = nullptr;
  if ((rc = b200knn_index_create(d_xyz, n_local, nullptr, 0, stream, &idx))) { b200knn_index_destroy(idx); return rc; }
  const double t1 = now_ms();
  rc = run_pv(d, idx, k, max_radius, pl, stream, rep);
  if (!rc) rc = scatter_results(idx->pts, idx->result_sorted, idx->n, 0, d_out, stream);
  if (!rc) { cudaError_t e = cudaStreamSynchronize(stream); if (e != cudaSuccess) { set_last_error(cudaGetErrorString(e)); rc = (int)e; } }
  if (rep) { rep->ms_build += t1 - t0; rep->ms_query += now_ms() - t1; }
  b200knn_index_destroy(idx);
  return rc;
}

// Elapsed GPU time between phases of a pipeline that the host no longer synchronises with: events recorded
// on the stream, read after the final synchronisation.
struct PhaseEvents {
  static constexpr int kMax = 6;
  cudaEvent_t ev[kMax] = {};
  int n = 0;
  bool on;
  explicit PhaseEvents(bool enabled) : on(enabled) {}
  ~PhaseEvents() { for (int i = 0; i < n; ++i) cudaEventDestroy(ev[i]); }
  void mark(cudaStream_t s) {
    if (!on || n >= kMax) return;
    if (cudaEventCreate(&ev[n]) != cudaSuccess) { on = false; return; }
    cudaEventRecord(ev[n++], s);
  }
  double ms(int from, int to) const {
    float t = 0.f;
    if (!on || to >= n || cudaEventElapsedTime(&t, ev[from], ev[to]) != cudaSuccess) return 0.0;
    return t;
  }
};

int knn_redistribute(b200knn_dist* d, const float* d_xyz, int64_t n_local, int64_t global_begin, int k,
                     float max_radius, int groups_per_chunk, float* d_out, cudaStream_t stream,
                     b200knn_dist_report* rep) {
  NvtxRange range("b200knn:redistribute_pipeline");
  const int R = d->world, me = d->rank;
  int rc;
  PhaseTrace trace(me, stream);
  PhaseEvents pe(rep != nullptr);
  KNN_REQUIRE(n_local < (int64_t)0x7fffffff - 64, "redistribute: more than 2^31 points on one rank");

  // ---- the histograms live in the symmetric heap (peers read them): make sure it exists
  const size_t hist_bytes = align256(rd_hist_bytes());
  if ((rc = ensure_heap(d, kHeapReserved + hist_bytes + 4 * 256))) return rc;
  d->bump = kHeapReserved;
  const size_t off_hist = d->alloc(hist_bytes);
  pe.mark(stream);

  // ---- who has what: bounds, count and first global index of every rank, gathered on the device
  StreamBuf<unsigned> d_bord(stream), d_info(stream), d_info_all(stream);
  StreamBuf<uint64_t> d_split(stream);
  StreamBuf<long long> d_row(stream), d_M(stream), d_begins(stream);
  StreamBuf<RdPlan> d_plan(stream);
  StreamBuf<Box> d_cells(stream);
  StreamBuf<int> d_cell_counts(stream);
  KNN_CUDA_TRY(d_cells.alloc_bytes((size_t)kMaxRanks * kTerritoryCells * sizeof(Box)));
  KNN_CUDA_TRY(d_cell_counts.alloc_bytes(kMaxRanks * sizeof(int)));
  KNN_CUDA_TRY(d_bord.alloc_bytes(6 * sizeof(unsigned)));
  KNN_CUDA_TRY(d_info.alloc_bytes(kCollWords * sizeof(unsigned)));
  KNN_CUDA_TRY(d_info_all.alloc_bytes((size_t)kMaxRanks * kCollWords * sizeof(unsigned)));
  KNN_CUDA_TRY(d_split.alloc_bytes(kMaxRanks * sizeof(uint64_t)));
  KNN_CUDA_TRY(d_row.alloc_bytes(kMaxRanks * sizeof(long long)));
  KNN_CUDA_TRY(d_M.alloc_bytes((size_t)kMaxRanks * kMaxRanks * sizeof(long long)));
  KNN_CUDA_TRY(d_begins.alloc_bytes((kMaxRanks + 1) * sizeof(long long)));
  KNN_CUDA_TRY(d_plan.alloc_bytes(sizeof(RdPlan)));
  if ((rc = bounds_init(d_bord, stream))) return rc;
  if ((rc = bounds_accumulate(d_xyz, 3, n_local, d_bord, stream))) return rc;
  if ((rc = rd_pack_info_launch(d_bord, n_local, global_begin, d_info, stream))) return rc;
  if (R > 1) {
    if ((rc = device_collective(d, d_info, kRdInfoWords, kCollGather, d_info_all, stream))) return rc;
  } else {
    KNN_CUDA_TRY(cudaMemcpyAsync(d_info_all, d_info, kRdInfoWords * sizeof(unsigned), cudaMemcpyDeviceToDevice, stream));
  }

  // ---- histograms of the leading Morton bits -> splitters and this rank's row of the count matrix
  unsigned* histA = reinterpret_cast<unsigned*>(d->base + off_hist);
  unsigned* histB = reinterpret_cast<unsigned*>(d->base + off_hist + rd_histB_offset());
  if ((rc = rd_hist_launch(d_xyz, n_local, d_info_all, R, histA, histB, d->num_sms, stream))) return rc;
  if ((rc = device_barrier(d, stream))) return rc;
  RdSplitParams sp;
  memset(&sp, 0, sizeof(sp));
  for (int r = 0; r < R; ++r) {
    sp.histA[r] = reinterpret_cast<const unsigned*>(d->peer[r] + off_hist);
    sp.histB[r] = reinterpret_cast<const unsigned*>(d->peer[r] + off_hist + rd_histB_offset());
  }
  sp.R = R;
  sp.rank = me;
  sp.n_local = n_local;
  sp.split_keys = d_split;
  sp.row = d_row;
  if ((rc = rd_split_launch(sp, stream))) return rc;
  if (R > 1) {
    if ((rc = device_collective(d, d_row, 2 * R, kCollGather, d_M, stream))) return rc;
  } else {
    KNN_CUDA_TRY(cudaMemcpyAsync(d_M, d_row, sizeof(long long), cudaMemcpyDeviceToDevice, stream));
  }
  if ((rc = rd_plan_launch(d_M, R, me, d_plan, stream))) return rc;
  if ((rc = rd_territory_launch(d_split, d_info_all, R, d_cells, d_cell_counts, stream))) return rc;

  // ---- the one host synchronisation: sizes
  RdPlan plan;
  std::vector<unsigned> info((size_t)R * kRdInfoWords);
  {
    SmallRead plan_read, info_read;      // no copy engine: see index_internal.h
    if ((rc = plan_read.begin(d_plan, sizeof(RdPlan), &plan, stream))) return rc;
    if ((rc = info_read.begin(d_info_all, info.size() * sizeof(unsigned), info.data(), stream))) return rc;
    KNN_CUDA_TRY(cudaStreamSynchronize(stream));
    plan_read.end();
    info_read.end();
  }
  trace.mark("redistribute: bounds, histograms, splitters, plan");
  std::vector<long long> begins(R + 1);
  int64_t max_local = 0, total = 0;
  for (int r = 0; r < R; ++r) {
    const unsigned* w = info.data() + (size_t)r * kRdInfoWords;
    const int64_t n_r = (int64_t)(((uint64_t)w[7] << 32) | w[6]), b_r = (int64_t)(((uint64_t)w[9] << 32) | w[8]);
    KNN_REQUIRE(r == 0 || b_r == begins[0] + total,
                "redistribute: ranks must hold consecutive slices of the global point order");
    begins[r] = b_r;
    max_local = std::max(max_local, n_r);
    total += n_r;
  }
  begins[R] = begins[0] + total;
  KNN_REQUIRE(begins[R] <= (int64_t)0xffffffffll, "redistribute: global point ids must fit 32 bits");
  const int64_t max_recv = plan.max_recv, my_recv = plan.my_recv;
  KNN_REQUIRE(max_recv < (int64_t)0x7fffffff - 64, "redistribute: a rank would receive more than 2^31 points");

  // ---- symmetric heap: histograms | receive buffer | result landing zone | exchange slots
  const PvPlan pl = pv_plan(max_recv, k, groups_per_chunk);
  const size_t need = hist_bytes + align256((size_t)std::max<int64_t>(max_recv, 1) * sizeof(Float4)) +
                      align256((size_t)std::max<int64_t>(max_local, 1) * 8) + pv_heap_need(k, pl.Gc, pl.max_groups, R);
  if ((rc = ensure_heap(d, need))) return rc;
  d->bump = kHeapReserved;
  (void)d->alloc(hist_bytes);
  const size_t off_recv = d->alloc((size_t)std::max<int64_t>(max_recv, 1) * sizeof(Float4));
  const size_t off_out = d->alloc((size_t)std::max<int64_t>(max_local, 1) * 8);   // (id, value) pairs land here

  RdScatterParams rp;
  memset(&rp, 0, sizeof(rp));
  rp.xyz = d_xyz;
  rp.n = n_local;
  rp.global_begin = global_begin;
  rp.R = R;
  rp.info_all = d_info_all;
  rp.split_keys = d_split;
  rp.plan = d_plan;
  for (int dst = 0; dst < R; ++dst) rp.recv[dst] = reinterpret_cast<Float4*>(d->peer[dst] + off_recv);
  if ((rc = rd_scatter_launch(rp, d->num_sms, stream))) return rc;
  // everybody's points have landed (and are visible) before anybody builds on them
  if ((rc = device_barrier(d, stream))) return rc;
  pe.mark(stream);
  trace.mark("redistribute: scatter + barrier");

  // ---- local index over the received (spatially compact) points, then the culled exchange
  IndexGuard index_guard;
  b200knn_index*& idx = index_guard.p;
  rc = b200knn_index_create_xyzw(reinterpret_cast<const float*>(d->base + off_recv), my_recv, 0, stream, &idx);
  pe.mark(stream);
  trace.mark("index build");
  if (!rc) rc = run_pv(d, idx, k, max_radius, pl, stream, rep, d_cells, d_cell_counts);
  pe.mark(stream);
  trace.mark("culled exchange (total)");

  // ---- results go back to the rank that read the point: (id, value) pairs partitioned by reader, each
  // reader's run pushed as contiguous pairs into its landing zone, the reader scatters locally into d_out
  if (!rc) {
    KNN_CUDA_TRY(cudaMemcpyAsync(d_begins, begins.data(), (size_t)(R + 1) * sizeof(long long), cudaMemcpyHostToDevice, stream));
    RdReturnParams qp;
    memset(&qp, 0, sizeof(qp));
    qp.pts = idx->pts;
    qp.res = idx->result_sorted;
    qp.n = idx->n;
    qp.R = R;
    qp.begins = d_begins;
    qp.plan = d_plan;
    for (int o = 0; o < R; ++o) qp.dst[o] = d->peer[o] + off_out;
    rc = rd_return_launch(qp, d->num_sms, stream);
  }
  trace.mark("return: push pairs");
  if (!rc) rc = device_barrier(d, stream);
  trace.mark("return: barrier");
  if (!rc && n_local > 0) rc = scatter_pairs_launch(d->base + off_out, n_local, global_begin, d_out, stream);
  pe.mark(stream);
  // No fence is needed before the next call: a peer writes into this rank's buffers only behind a device
  // barrier or collective of that next call, which this rank joins after everything above (stream order).
  if (!rc) {
    cudaError_t e = cudaStreamSynchronize(stream);
    if (e != cudaSuccess) { set_last_error(cudaGetErrorString(e)); rc = (int)e; }
  }
  trace.mark("return: local scatter");
  if (rep) {
    rep->ms_redistribute += pe.ms(0, 1);
    rep->ms_build += pe.ms(1, 2);
    rep->ms_query += pe.ms(2, 3);
    rep->ms_return += pe.ms(3, 4);
    rep->points_after_redistribute = my_recv;
  }
  return rc;
}

int knn_ring(b200knn_dist* d, const float* d_xyz, int64_t n_local, int k, float max_radius, int groups_per_chunk,
             float* d_out, cudaStream_t stream, b200knn_dist_report* rep) {
  int rc;
  int64_t max_n = 0;
  {
    std::vector<int64_t> all(d->world);
    if (d->allgather(&n_local, all.data(), sizeof(int64_t))) { set_last_error("bootstrap allgather failed"); return -3; }
    for (int64_t v : all) max_n = std::max(max_n, v);
  }
  const double t0 = now_ms();
  b200knn_index* idx = nullptr;
  b200knn_ring* ring = nullptr;
  rc = b200knn_index_create(d_xyz, n_local, nullptr, 0, stream, &idx);
  const double t1 = now_ms();
  if (!rc) rc = b200knn_ring_create(idx, d->rank, d->world, k, max_n, groups_per_chunk, &ring);
  std::vector<char> mine(B200KNN_IPC_BLOB_BYTES), all((size_t)B200KNN_IPC_BLOB_BYTES * d->world);
  if (!rc) rc = b200knn_ring_export(ring, mine.data());
  if (!rc && d->allgather(mine.data(), all.data(), B200KNN_IPC_BLOB_BYTES)) { set_last_error("bootstrap allgather failed"); rc = -3; }
  if (!rc) rc = b200knn_ring_connect(ring, all.data());
  if (!rc && d->barrier()) { set_last_error("bootstrap barrier failed"); rc = -3; }
  if (!rc) rc = b200knn_ring_run(ring, max_radius, stream);
  if (!rc) { cudaError_t e = cudaStreamSynchronize(stream); if (e != cudaSuccess) { set_last_error(cudaGetErrorString(e)); rc = (int)e; } }
  if (!rc && d->barrier()) { set_last_error("bootstrap barrier failed"); rc = -3; }
  if (!rc) rc = b200knn_ring_results(ring, d_out, stream);
  if (!rc) { cudaError_t e = cudaStreamSynchronize(stream); if (e != cudaSuccess) { set_last_error(cudaGetErrorString(e)); rc = (int)e; } }
  if (!rc && d->barrier()) { set_last_error("bootstrap barrier failed"); rc = -3; }
  if (rep) { rep->ms_build += t1 - t0; rep->ms_query += now_ms() - t1; }
  if (ring) {
    b200knn_ring_disconnect(ring);
    d->barrier();
  }
  b200knn_ring_destroy(ring);
  b200knn_index_destroy(idx);
  return rc;
}

}  // namespace

extern "C" {

int b200knn_dist_create(const b200knn_bootstrap* bs, b200knn_dist** out) {
  KNN_REQUIRE(bs && out, "dist_create: NULL argument");
  KNN_REQUIRE(bs->world >= 1 && bs->world <= B200KNN_MAX_RANKS, "dist_create: world must be in [1,16]");
  KNN_REQUIRE(bs->rank >= 0 && bs->rank < bs->world, "dist_create: rank out of range");
  KNN_REQUIRE(bs->world == 1 || (bs->allgather && bs->barrier), "dist_create: bootstrap callbacks missing");
  b200knn_dist* d = new b200knn_dist();
  *out = d;
  d->bs = *bs;
  d->rank = bs->rank;
  d->world = bs->world;
  KNN_CUDA_TRY(cudaGetDevice(&d->device));
  keep_pool_memory(d->device);
  KNN_CUDA_TRY(cudaDeviceGetAttribute(&d->num_sms, cudaDevAttrMultiProcessorCount, d->device));
  KNN_CUDA_TRY(cudaMalloc(&d->task_counter, sizeof(unsigned long long)));
  return 0;
}

void b200knn_dist_destroy(b200knn_dist* d) {
  if (!d) return;
  cudaDeviceSynchronize();
  if (d->world > 1) d->barrier();
  close_heap(d);
  cudaFree(d->task_counter);
  delete d;
}

int64_t b200knn_dist_heap_bytes(const b200knn_dist* d) { return d ? (int64_t)d->bytes : 0; }

int b200knn_dist_allgather_dev(b200knn_dist* d, const void* d_send, int nbytes, void* d_recv, void* stream) {
  KNN_REQUIRE(d && d_send && d_recv, "dist_allgather_dev: NULL argument");
  KNN_REQUIRE(nbytes >= 4 && nbytes % 4 == 0 && nbytes <= B200KNN_DEV_COLLECTIVE_MAX_BYTES,
              "dist_allgather_dev: nbytes must be a multiple of 4 in [4, 128]");
  if (d->world == 1) {
    KNN_CUDA_TRY(cudaMemcpyAsync(d_recv, d_send, (size_t)nbytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return 0;
  }
  int rc;
  if ((rc = ensure_heap(d, kHeapReserved))) return rc;
  return device_collective(d, d_send, nbytes / 4, kCollGather, d_recv, (cudaStream_t)stream);
}

int b200knn_dist_allreduce_dev(b200knn_dist* d, const int64_t* d_send, int count, int op, int64_t* d_recv, void* stream) {
  KNN_REQUIRE(d && d_send && d_recv, "dist_allreduce_dev: NULL argument");
  KNN_REQUIRE(count >= 1 && count * 8 <= B200KNN_DEV_COLLECTIVE_MAX_BYTES, "dist_allreduce_dev: count must be in [1, 16]");
  KNN_REQUIRE(op == B200KNN_OP_MAX || op == B200KNN_OP_SUM || op == B200KNN_OP_MIN, "dist_allreduce_dev: unknown op");
  if (d->world == 1) {
    KNN_CUDA_TRY(cudaMemcpyAsync(d_recv, d_send, (size_t)count * 8, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return 0;
  }
  int rc;
  if ((rc = ensure_heap(d, kHeapReserved))) return rc;
  const int kop = op == B200KNN_OP_MAX ? kCollMaxI64 : (op == B200KNN_OP_SUM ? kCollSumI64 : kCollMinI64);
  return device_collective(d, d_send, 2 * count, kop, d_recv, (cudaStream_t)stream);
}

int b200knn_dist_knn(b200knn_dist* d, const float* d_xyz, int64_t n_local, int64_t global_begin, int k,
                     float max_radius, int mode, int groups_per_chunk, float* d_out, void* stream_,
                     b200knn_dist_report* report) {
  cudaStream_t stream = (cudaStream_t)stream_;
  KNN_REQUIRE(d, "dist_knn: NULL engine");
  KNN_REQUIRE(k >= 1, "dist_knn: k must be >= 1");
  KNN_REQUIRE(n_local >= 0 && (n_local == 0 || (d_xyz && d_out)), "dist_knn: bad point buffers");
  if (report) memset(report, 0, sizeof(*report));
  const double t0 = now_ms();
  int rc;
  switch (mode) {
    case B200KNN_MODE_RING:
      rc = knn_ring(d, d_xyz, n_local, k, max_radius, groups_per_chunk, d_out, stream, report);
      break;
    case B200KNN_MODE_PREPARTITIONED:
      rc = knn_prepartitioned(d, d_xyz, n_local, k, max_radius, groups_per_chunk, d_out, stream, report);
      break;
    case B200KNN_MODE_REDISTRIBUTE:
      rc = knn_redistribute(d, d_xyz, n_local, global_begin, k, max_radius, groups_per_chunk, d_out, stream, report);
      break;
    case B200KNN_MODE_AUTO: {
      // Per-rank inputs: keep the points in place when the ranks' boxes are (nearly) disjoint, otherwise
      // treat the concatenation of the inputs as one unordered cloud and redistribute. Box culling cannot
      // separate clouds that interpenetrate (e.g. Gaussian blobs with overlapping tails): every group
      // would visit every rank, which costs far more than one all-to-all of the points.
      struct Info { int64_t n; float box[6]; };
      Info mine{};
      mine.n = n_local;
      StreamBuf<unsigned> d_bord(stream);
      StreamBuf<float> d_box6(stream);
      KNN_CUDA_TRY(d_bord.alloc_bytes(6 * sizeof(unsigned)));
      KNN_CUDA_TRY(d_box6.alloc_bytes(6 * sizeof(float)));
      if ((rc = bounds_init(d_bord, stream))) return rc;
      if ((rc = bounds_accumulate(d_xyz, 3, n_local, d_bord, stream))) return rc;
      if ((rc = bounds_decode(d_bord, d_box6, stream))) return rc;
      SmallRead box_read;
      if ((rc = box_read.begin(d_box6, 6 * sizeof(float), mine.box, stream))) return rc;
      d_bord.release();
      d_box6.release();
      KNN_CUDA_TRY(cudaStreamSynchronize(stream));
      box_read.end();
      std::vector<Info> info(d->world);
      if (d->allgather(&mine, info.data(), sizeof(Info))) { set_last_error("bootstrap allgather failed"); return -3; }
      double worst = 0.0;
      int64_t begin = 0;
      for (int a = 0; a < d->world; ++a) {
        if (a < d->rank) begin += info[a].n;
        if (info[a].n == 0) continue;
        double va = 1.0;
        for (int c = 0; c < 3; ++c) va *= std::max(1e-30, (double)info[a].box[3 + c] - info[a].box[c]);
        for (int b = 0; b < d->world; ++b) {
          if (b == a || info[b].n == 0) continue;
          double vi = 1.0;
          for (int c = 0; c < 3; ++c)
            vi *= std::max(0.0, (double)std::min(info[a].box[3 + c], info[b].box[3 + c]) - std::max(info[a].box[c], info[b].box[c]));
          worst = std::max(worst, vi / va);
        }
      }
      const bool in_place = worst < 0.02;
      if (report) report->auto_chose_redistribute = in_place ? 0 : 1;
      rc = in_place ? knn_prepartitioned(d, d_xyz, n_local, k, max_radius, groups_per_chunk, d_out, stream, report)
                    : knn_redistribute(d, d_xyz, n_local, begin, k, max_radius, groups_per_chunk, d_out, stream, report);
      break;
    }
    default:
      set_last_error("dist_knn: unknown mode");
      rc = -2;
  }
  if (report) report->ms_total = now_ms() - t0;
  return rc;
}

}  // extern "C"
