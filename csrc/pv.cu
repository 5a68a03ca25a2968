// pv.cu — kernels of the bounding-box-culled exchange (pre-partitioned variant and the redistributed
// fast path of the unordered variant), plus point redistribution and result return.
//
// Reference behaviour being replaced (/root/reference/prePartitionedDataVariant.cu:304-357): every
// round each rank pulls ONE whole peer tree, chosen by one scalar radius per rank against one box
// per peer, and re-runs all of its queries against it. Here the trees never move and culling is per
// query group:
//   pv_main (phase LOCAL)  all home groups against the home tree; a group whose ball still reaches
//                          another rank's box keeps its heaps in a "pending" slot
//   pv_push                per round, pending groups that reach the round's destination box are
//                          compacted and written straight into that GPU's inbox (peer stores)
//   pv_main (phase SERVE)  the destination traverses its tree for the visiting groups and writes the
//                          improved heaps straight back into the home GPU's pending slots (peer
//                          stores) — the merge-back
//   pv_final               k-th distances of pending groups
// Rounds are separated by device-side barriers (dev_barrier_kernel); all data movement is done by these kernels
// over NVLink-mapped peer memory, never by a collective library.
//
// redistribute / push_pairs: the all-to-all scatter of points to their spatial owner and of the
// results back to the rank that read the point — both as contiguous in-kernel peer stores (new
// capability; the reference never redistributes).
#include "traverse.cuh"

namespace b200knn {

namespace {

__device__ __forceinline__ float* slot_heap(char* slot) { return reinterpret_cast<float*>(slot); }
__device__ __forceinline__ const float* slot_heap(const char* slot) { return reinterpret_cast<const float*>(slot); }
__device__ __forceinline__ float4* slot_pos(char* slot, int k) {
  return reinterpret_cast<float4*>(slot + (size_t)k * 128);
}
__device__ __forceinline__ const float4* slot_pos(const char* slot, int k) {
  return reinterpret_cast<const float4*>(slot + (size_t)k * 128);
}
__device__ __forceinline__ int* slot_meta(char* slot, int k) {
  return reinterpret_cast<int*>(slot + (size_t)k * 128 + 512);
}
__device__ __forceinline__ const int* slot_meta(const char* slot, int k) {
  return reinterpret_cast<const int*>(slot + (size_t)k * 128 + 512);
}

// Does the group's ball reach rank s's points, as far as s's published tree levels can tell? The warp walks
// them depth-first (lane = child box, group-level test with the group's box and largest radius) and settles
// the matter at the lowest published level with the per-query test. `stack` holds kTopsStack node ids; if it
// overflows the answer is a conservative yes.
constexpr int kTopsStack = 128;
__device__ __forceinline__ bool reaches_rank(const TopsView& tv, int s, float qx, float qy, float qz, float top,
                                             bool valid, float glx, float gly, float glz, float ghx, float ghy,
                                             float ghz, float maxR2, unsigned* stack) {
  const TopHead* h = tv.heads + s;
  const int levels = __ldg(&h->levels);
  if (levels == 0) return false;
  if (!(box_box_lb(glx, gly, glz, ghx, ghy, ghz, __ldg(&h->box[0]), __ldg(&h->box[1]), __ldg(&h->box[2]),
                   __ldg(&h->box[3]), __ldg(&h->box[4]), __ldg(&h->box[5])) < maxR2))
    return false;
  const int base_level = __ldg(&h->base_level);
  const Box* boxes = tv.boxes + (size_t)s * tv.cap;
  const int lane = lane_id();
  const unsigned lt = (1u << lane) - 1u;
  int sp = 0;
  unsigned node = (unsigned)levels << kNodeIdxBits;      // virtual root: its children are the top level
  for (;;) {
    const int L = (int)(node >> kNodeIdxBits), idx = (int)(node & ((1u << kNodeIdxBits) - 1u));
    const int cl = L - 1;
    const int cbase = L == levels ? 0 : idx * kFanout;
    const int rem = __ldg(&h->count[cl]) - cbase;
    const int ccount = rem < kFanout ? rem : kFanout;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
    if (lane < ccount) {
      const float4* bp = reinterpret_cast<const float4*>(boxes + __ldg(&h->offset[cl]) + cbase + lane);
      a = __ldg(bp);
      b = __ldg(bp + 1);
    }
    const bool pass = lane < ccount && box_box_lb(glx, gly, glz, ghx, ghy, ghz, a.x, a.y, a.z, b.x, b.y, b.z) < maxR2;
    unsigned pm = __ballot_sync(kFullMask, pass);
    if (cl == base_level) {
      while (pm != 0u) {
        const int bi = __ffs(pm) - 1;
        pm &= pm - 1;
        const float lb = point_box_lb(qx, qy, qz, __shfl_sync(kFullMask, a.x, bi), __shfl_sync(kFullMask, a.y, bi),
                                      __shfl_sync(kFullMask, a.z, bi), __shfl_sync(kFullMask, b.x, bi),
                                      __shfl_sync(kFullMask, b.y, bi), __shfl_sync(kFullMask, b.z, bi));
        if (__any_sync(kFullMask, valid && (lb < top))) return true;
      }
    } else if (pm != 0u) {
      const int np = __popc(pm);
      if (sp + np > kTopsStack) return true;
      if (pass) stack[sp + __popc(pm & lt)] = ((unsigned)cl << kNodeIdxBits) | (unsigned)(cbase + lane);
      sp += np;
      __syncwarp();
    }
    if (sp == 0) return false;
    node = stack[--sp];
    __syncwarp();
  }
}

// Does the group's ball touch rank s's share of key space (redist.cu, rd_territory)? lane = cell.
__device__ __forceinline__ bool reaches_territory(const Box* __restrict__ territory, const int* __restrict__ territory_count,
                                                  int s, float qx, float qy, float qz, float top, bool valid, float glx,
                                                  float gly, float glz, float ghx, float ghy, float ghz, float maxR2) {
  const int n = __ldg(territory_count + s);
  const Box* cells = territory + (size_t)s * kTerritoryCells;
  const int lane = lane_id();
  for (int base = 0; base < n; base += 32) {
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
    bool hit = false;
    if (base + lane < n) {
      const float4* bp = reinterpret_cast<const float4*>(cells + base + lane);
      a = __ldg(bp);
      b = __ldg(bp + 1);
      hit = box_box_lb(glx, gly, glz, ghx, ghy, ghz, a.x, a.y, a.z, b.x, b.y, b.z) < maxR2;
    }
    unsigned hm = __ballot_sync(kFullMask, hit);
    while (hm != 0u) {
      const int bi = __ffs(hm) - 1;
      hm &= hm - 1;
      const float lb = point_box_lb(qx, qy, qz, __shfl_sync(kFullMask, a.x, bi), __shfl_sync(kFullMask, a.y, bi),
                                    __shfl_sync(kFullMask, a.z, bi), __shfl_sync(kFullMask, b.x, bi),
                                    __shfl_sync(kFullMask, b.y, bi), __shfl_sync(kFullMask, b.z, bi));
      if (__any_sync(kFullMask, valid && (lb < top))) return true;
    }
  }
  return false;
}

// SORTED / GLOBAL_HEAP: as in query.cu — sorted lists + append buffers in shared memory for 2 <= k <= ~1700,
// the heap for k == 1 and (in a global scratch block per resident warp) for k beyond the shared-memory limit.
template <bool GLOBAL_HEAP, bool SORTED>
__global__ void __launch_bounds__(Q_MAX_WARPS * 32, 1)
pv_main_kernel_t(const PvMainParams p) {
  extern __shared__ __align__(16) float q_smem[];
  const int lane = lane_id();
  const int k = p.k;
  GroupCtx g;
  if constexpr (GLOBAL_HEAP) {
    const size_t warp_global = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    group_ctx_init(g, q_smem, threadIdx.x >> 5, lane, k, p.global_heaps + warp_global * (size_t)heap_padded(k) * 32);
  } else {
    group_ctx_init(g, q_smem, threadIdx.x >> 5, lane, k, nullptr, SORTED);
  }
  const int kp = g.kp;
  const size_t slot_bytes = pv_slot_bytes(k);
  const int total = p.phase == 0 ? p.group_count : ldcg_i(p.inbox_count);

  for (;;) {
    unsigned long long t = 0;
    if (lane == 0) t = atomicAdd(p.task_counter, 1ull);
    t = __shfl_sync(kFullMask, t, 0);
    if (t >= (unsigned long long)total) break;

    int cnt, home_slot = 0;
    int64_t home_group = 0;
    if (p.phase == 0) {
      home_group = p.group_begin + (int64_t)t;
      const int64_t rem = p.home_n - home_group * kGroup;
      cnt = rem <= 0 ? 0 : (rem < kGroup ? (int)rem : kGroup);
      if (cnt > 0) {
        const int src = lane < cnt ? lane : cnt - 1;
        const float4 q = __ldg(reinterpret_cast<const float4*>(p.home_queries) + home_group * kGroup + src);
        g.qx = q.x; g.qy = q.y; g.qz = q.z;
        const float sent = p.sentinel;
        for (int i = 0; i < k; ++i) g.heap[i * 32] = sent;
        g.top = sent;
        g.last_id = 0xffffffffu;
        g.live = 0;
      }
    } else {
      const char* slot = p.inbox + (size_t)t * slot_bytes;
      const int* meta = slot_meta(slot, k);
      home_slot = ldcg_i(meta + 0);
      cnt = ldcg_i(meta + 1);
      const float4 q = ldcg_f4(slot_pos(slot, k) + lane);
      g.qx = q.x; g.qy = q.y; g.qz = q.z;
      const float* src = slot_heap(slot) + lane;
      for (int i = 0; i < k; ++i) g.heap[i * 32] = ldcg_f(src + (size_t)i * 32);
      g.top = g.heap[0];
      g.live = k;       // an arriving list may be full
    }
    g.valid = lane < cnt;

    if (cnt > 0 && p.tree.n > 0) {
      for (int i = k; i < kp; ++i) g.heap[i * 32] = -INFINITY;
      const int own_seed = (p.phase == 0 && p.home_is_tree)
                               ? (p.tree.levels >= 2 ? (int)(home_group / kFanout) : -1) : -2;
      traverse_group_knn<SORTED>(p.tree, g, cnt, own_seed, k, (p.phase == 0 && p.fill) ? k / kGroup : 0);
    }

    if (p.phase == 0) {
      if (cnt > 0) {
        if (g.valid) p.result_sorted[home_group * kGroup + lane] = finalize_kth(g.top);
        // does the group's ball still reach another rank's region? lane = one of that rank's boxes
        group_bounds(g);
        // which ranks? (the sequential exchange only needs to know whether there is one)
        unsigned reach = 0u;
        for (int s = 0; s < p.R && (p.fused || reach == 0u); ++s) {
          if (s == p.rank) continue;
          if (p.territory != nullptr &&
              !reaches_territory(p.territory, p.territory_count, s, g.qx, g.qy, g.qz, g.top, g.valid, g.glx, g.gly, g.glz,
                                 g.ghx, g.ghy, g.ghz, g.maxR2))
            continue;
          if (reaches_rank(p.tops, s, g.qx, g.qy, g.qz, g.top, g.valid, g.glx, g.gly, g.glz, g.ghx, g.ghy, g.ghz,
                           g.maxR2, g.qnode))
            reach |= 1u << s;
        }
        const bool need = reach != 0u;
        if (need) {
          int slot_idx = 0;
          if (lane == 0) slot_idx = atomicAdd(p.pending_count, 1);
          slot_idx = __shfl_sync(kFullMask, slot_idx, 0);
          char* slot = p.pending + (size_t)slot_idx * slot_bytes;
          float* dst = slot_heap(slot) + lane;
          for (int i = 0; i < k; ++i) dst[(size_t)i * 32] = g.heap[i * 32];
          slot_pos(slot, k)[lane] = make_float4(g.qx, g.qy, g.qz, 0.f);
          if (lane == 0) {
            int* meta = slot_meta(slot, k);
            meta[0] = (int)(home_group & 0xffffffffll);
            meta[1] = cnt;
            meta[2] = (int)(home_group >> 32);
            meta[3] = 0;            // MERGE's lock word
          }
          // one-round exchange: a request to every rank the ball reaches — queries + their k-th distances
          for (unsigned todo = p.fused ? reach : 0u; todo != 0u; todo &= todo - 1) {
            const int dst = __ffs(todo) - 1;
            int r_idx = 0;
            if (lane == 0) r_idx = atomicAdd(p.req_out_count + dst, 1);
            r_idx = __shfl_sync(kFullMask, r_idx, 0);
            if (r_idx < p.cap_req) {
              char* req = p.req_inbox[dst] + ((size_t)p.rank * p.cap_req + r_idx) * pv_request_bytes();
              reinterpret_cast<float4*>(req)[lane] = make_float4(g.qx, g.qy, g.qz, g.top);
              if (lane == 0) {
                int* rmeta = reinterpret_cast<int*>(req + 512);
                rmeta[0] = slot_idx; rmeta[1] = cnt; rmeta[2] = 0; rmeta[3] = 0;
              }
            }
          }
        }
      }
    } else if (cnt > 0) {
      // merge-back: improved heaps go straight into the home GPU's pending slot
      float* dst = slot_heap(p.src_pending + (size_t)home_slot * slot_bytes) + lane;
      for (int i = 0; i < k; ++i) dst[(size_t)i * 32] = g.heap[i * 32];
    }
    __syncwarp();
  }
  flush_stats(g, p.stats);
}

// ---- one-round exchange: SERVE_FRESH (PHASE 2) and MERGE (PHASE 3) --------------------------------------
// SERVE_FRESH: every request received from any rank, in one pass (one tail instead of R - 1): the 32 queries start
// from an EMPTY list whose sentinel is each lane's k-th distance at home — nothing at or beyond it can matter —
// and walk the local tree; what they find (rows [k - found, k) of the sorted lists) goes back to the home GPU as a
// RETURN record, the slot reserved by one atomic on the home GPU's counter over NVLink. A request that finds
// nothing sends nothing.
// MERGE: home appends every returned value that still beats the lane's k-th distance to the lane's buffer and
// flushes — the same code path as a leaf. Several servers may answer the same group; a spin lock in the pending
// slot's meta word serialises them (one to three returns per group, the holder never waits for anything).
template <int PHASE>
__global__ void __launch_bounds__(Q_MAX_WARPS * 32, 1)
pv_fused_kernel_t(const PvMainParams p) {
  extern __shared__ __align__(16) float q_smem[];
  const int lane = lane_id();
  const int k = p.k;
  GroupCtx g;
  group_ctx_init(g, q_smem, threadIdx.x >> 5, lane, k, nullptr, true);
  const size_t slot_bytes = pv_slot_bytes(k);
  int total = 0;
  auto capped = [](int v, int64_t cap) { return v < (int)cap ? v : (int)cap; };
  if (PHASE == 2) { for (int s = 0; s < p.R; ++s) total += capped(p.req_in[s], p.cap_req); }
  else total = capped(ldcg_i(p.my_ret_count), p.cap_ret);

  for (;;) {
    unsigned long long t = 0;
    if (lane == 0) t = atomicAdd(p.task_counter, 1ull);
    t = __shfl_sync(kFullMask, t, 0);
    if (t >= (unsigned long long)total) break;

    if (PHASE == 2) {
      int src = 0, idx = (int)t;
      for (;; ++src) {                                    // which source's region does task t fall into
        const int c = capped(p.req_in[src], p.cap_req);
        if (idx < c) break;
        idx -= c;
      }
      const char* req = p.my_req_inbox + ((size_t)src * p.cap_req + idx) * pv_request_bytes();
      const float4 q = ldcg_f4(reinterpret_cast<const float4*>(req) + lane);
      const int home_slot = ldcg_i(reinterpret_cast<const int*>(req + 512));
      const int cnt = ldcg_i(reinterpret_cast<const int*>(req + 512) + 1);
      g.qx = q.x; g.qy = q.y; g.qz = q.z;
      g.valid = lane < cnt;
      // idle lanes: a list of zeros takes nothing and, unlike -inf, cannot be mistaken for the column's terminator
      const float thr = g.valid ? q.w : 0.f;
      for (int i = 0; i < k; ++i) g.heap[i * 32] = thr;
      g.heap[k * 32] = -INFINITY;
      g.top = thr;
      g.live = 0;
      if (cnt > 0 && p.tree.n > 0) traverse_group_knn<true>(p.tree, g, cnt, -2, k, 0);
      const int found = (int)__reduce_max_sync(kFullMask, (unsigned)(g.valid ? g.live : 0));
      if (found > 0) {
        int r_idx = 0;
        if (lane == 0) r_idx = atomicAdd_system(p.ret_count[src], 1);
        r_idx = __shfl_sync(kFullMask, r_idx, 0);
        if (r_idx < p.cap_ret) {
          char* rec = p.ret_area[src] + (size_t)r_idx * pv_return_bytes(k);
          float* dst = reinterpret_cast<float*>(rec) + lane;
          for (int i = 0; i < found; ++i) dst[(size_t)i * 32] = g.heap[(k - found + i) * 32];
          if (lane == 0) {
            int* meta = reinterpret_cast<int*>(rec + (size_t)k * 128);
            meta[0] = home_slot; meta[1] = found; meta[2] = p.rank; meta[3] = 0;
          }
        }
      }
    } else {
      const char* rec = p.my_ret_area + (size_t)t * pv_return_bytes(k);
      const int* rmeta = reinterpret_cast<const int*>(rec + (size_t)k * 128);
      const int home_slot = ldcg_i(rmeta), found = ldcg_i(rmeta + 1);
      char* slot = p.pending + (size_t)home_slot * slot_bytes;
      int* lock = slot_meta(slot, k) + 3;
      if (lane == 0) {
        unsigned ns = 32;
        while (atomicCAS(lock, 0, 1) != 0) { __nanosleep(ns); if (ns < 1024) ns <<= 1; }
        __threadfence();
      }
      __syncwarp();
      const int cnt = slot_meta(slot, k)[1];
      g.valid = lane < cnt;
      float* list = slot_heap(slot) + lane;
      for (int i = 0; i < k; ++i) g.heap[i * 32] = __ldcg(list + (size_t)i * 32);
      g.heap[k * 32] = -INFINITY;
      g.top = g.heap[0];
      g.live = k;
      g.cnt = 0;
      g.maxR2 = 0.f;
      const float* src = reinterpret_cast<const float*>(rec) + lane;
      const float top0 = g.valid ? g.top : -INFINITY;
      for (int i = 0; i < found; ++i) {
        const float v = ldcg_f(src + (size_t)i * 32);
        if (v < top0) { g.buf[g.cnt * 32] = v; ++g.cnt; }      // sentinel rows of the server's list equal top0: dropped
        if ((i & 31) == 31 || i == found - 1) {
          __syncwarp();
          if (__any_sync(kFullMask, g.cnt > 0)) sorted_flush(g, k);
        }
      }
      for (int i = 0; i < k; ++i) list[(size_t)i * 32] = g.heap[i * 32];
      __threadfence();                       // every lane's column is out before lane 0 opens the lock
      __syncwarp();
      if (lane == 0) atomicExch(lock, 0);
    }
    __syncwarp();
  }
  flush_stats(g, p.stats);
}

// One warp per pending slot: if the group's ball (with the radius it has by now) still reaches the
// destination rank, copy the whole slot into the next free slot of the destination GPU's inbox.
__global__ void __launch_bounds__(256)
pv_push_kernel(const char* __restrict__ pending, const int* __restrict__ pending_count, int k, const TopsView tops,
               const Box* __restrict__ territory, const int* __restrict__ territory_count, int dst_rank,
               char* __restrict__ peer_inbox, int* __restrict__ out_count) {
  __shared__ unsigned stacks[8][kTopsStack];
  const int lane = lane_id();
  const int warps = (gridDim.x * blockDim.x) >> 5;
  const size_t slot_bytes = pv_slot_bytes(k);
  const int total = *pending_count;
  for (int s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; s < total; s += warps) {
    const char* slot = pending + (size_t)s * slot_bytes;
    const int cnt = slot_meta(slot, k)[1];
    const float4 q = slot_pos(slot, k)[lane];
    const bool valid = lane < cnt;
    float lx = valid ? q.x : INFINITY, ly = valid ? q.y : INFINITY, lz = valid ? q.z : INFINITY;
    float hx = valid ? q.x : -INFINITY, hy = valid ? q.y : -INFINITY, hz = valid ? q.z : -INFINITY;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lx = fminf(lx, __shfl_xor_sync(kFullMask, lx, o));
      ly = fminf(ly, __shfl_xor_sync(kFullMask, ly, o));
      lz = fminf(lz, __shfl_xor_sync(kFullMask, lz, o));
      hx = fmaxf(hx, __shfl_xor_sync(kFullMask, hx, o));
      hy = fmaxf(hy, __shfl_xor_sync(kFullMask, hy, o));
      hz = fmaxf(hz, __shfl_xor_sync(kFullMask, hz, o));
    }
    const float top = slot_heap(slot)[lane];
    const float maxR2 = __uint_as_float(__reduce_max_sync(kFullMask, valid ? __float_as_uint(top) : 0u));
    if (territory != nullptr &&
        !reaches_territory(territory, territory_count, dst_rank, q.x, q.y, q.z, top, valid, lx, ly, lz, hx, hy, hz, maxR2))
      continue;
    if (!reaches_rank(tops, dst_rank, q.x, q.y, q.z, top, valid, lx, ly, lz, hx, hy, hz, maxR2, stacks[threadIdx.x >> 5]))
      continue;
    int o = 0;
    if (lane == 0) o = atomicAdd(out_count, 1);
    o = __shfl_sync(kFullMask, o, 0);
    char* dst = peer_inbox + (size_t)o * slot_bytes;
    // heap rows + positions: k*128 + 512 bytes = (k*8 + 32) float4 per warp pass of 32
    const float4* src4 = reinterpret_cast<const float4*>(slot);
    float4* dst4 = reinterpret_cast<float4*>(dst);
    const int n4 = k * 8 + 32;
    for (int i = lane; i < n4; i += 32) dst4[i] = src4[i];
    if (lane == 0) {
      int* meta = slot_meta(dst, k);
      meta[0] = s;        // home slot: where the merge-back goes
      meta[1] = cnt;
      meta[2] = 0;
      meta[3] = 0;
    }
  }
}

// Cross-GPU barrier as a stream-ordered kernel: lane r tells rank r "I am at sequence seq" and waits
// until rank r has said the same. Runs as one warp; only ever waits on OTHER GPUs.
__global__ void dev_barrier_kernel(const DevBarrierParams p) {
  const int lane = threadIdx.x;
  if (lane < p.R) {
    __threadfence_system();
    st_release_sys(p.peer_flags[lane] + p.rank, p.seq);
    wait_flag_ge(p.local_flags + lane, p.seq);
  }
}

// All-gather / all-reduce of a few words per rank as one stream-ordered kernel. Warp w talks to rank w:
// it pushes this rank's words into mailbox [rank] over there, raises flag [rank] there, then waits
// for rank w's flag here and picks up rank w's words. Mailboxes are double-buffered by the parity of
// the sequence number (the host alternates the pointers): a rank can be at most one collective
// ahead of the slowest one, because finishing collective s needs everybody's contribution to s.
__global__ void dev_collective_kernel(const DevCollParams p) {
  __shared__ unsigned gathered[kMaxRanks][kCollWords];
  const int w = threadIdx.x >> 5, lane = lane_id();
  if (w < p.R) {
    if (lane < p.words) {
      volatile unsigned* box = p.peer_data[w] + (size_t)p.rank * kCollWords;
      box[lane] = p.send[lane];
    }
    __threadfence_system();
    __syncwarp();
    if (lane == 0) {
      st_release_sys(p.peer_flags[w] + p.rank, p.seq);
      wait_flag_ge(p.local_flags + w, p.seq);
    }
    __syncwarp();
    unsigned v = 0;
    if (lane < p.words) v = __ldcg(p.local_data + (size_t)w * kCollWords + lane);
    if (p.op == kCollGather) {
      if (lane < p.words) p.recv[(size_t)w * p.words + lane] = v;
    } else {
      gathered[w][lane] = v;
    }
  }
  if (p.op == kCollGather) return;
  __syncthreads();
  const int e = threadIdx.x;                       // one thread per int64 element
  if (e < p.words / 2) {
    auto at = [&](int r) {
      return (long long)(((unsigned long long)gathered[r][2 * e + 1] << 32) | gathered[r][2 * e]);
    };
    long long acc = at(0);
    for (int r = 1; r < p.R; ++r) {
      const long long x = at(r);
      acc = p.op == kCollSumI64 ? acc + x : (p.op == kCollMaxI64 ? (x > acc ? x : acc) : (x < acc ? x : acc));
    }
    reinterpret_cast<long long*>(p.recv)[e] = acc;
  }
}

__global__ void pv_publish_kernel(const int* __restrict__ out_count, int* __restrict__ peer_inbox_count) {
  if (threadIdx.x == 0) *peer_inbox_count = *out_count;
}

__global__ void __launch_bounds__(256)
pv_final_kernel(const char* __restrict__ pending, const int* __restrict__ pending_count, int k,
                float* __restrict__ result_sorted) {
  const int lane = lane_id();
  const int warps = (gridDim.x * blockDim.x) >> 5;
  const size_t slot_bytes = pv_slot_bytes(k);
  const int total = *pending_count;
  for (int s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; s < total; s += warps) {
    const char* slot = pending + (size_t)s * slot_bytes;
    const int* meta = slot_meta(slot, k);
    const int64_t home_group = ((int64_t)meta[2] << 32) | (uint32_t)meta[0];
    if (lane < meta[1]) result_sorted[home_group * kGroup + lane] = finalize_kth(slot_heap(slot)[lane]);
  }
}

// ---- result return ------------------------------------------------------------------------------
// The pairs arrive in no particular order, so this is a random permutation of 4-byte values: written in
// one go, every store would be a read-modify-write of a 32-byte DRAM sector (measured 5.2 ms for 134 M
// pairs). Instead the output is covered in windows small enough to stay in L2; every window re-reads the
// pairs (sequential, cheap) and writes only its own ids, which merge in L2 before they go to DRAM.
__global__ void __launch_bounds__(256)
scatter_pairs_kernel(const uint2* __restrict__ pairs, int64_t n, int64_t begin, float* __restrict__ out,
                     int64_t window_lo, int64_t window_hi) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint2 pr = __ldcs(pairs + i);
    const int64_t o = (int64_t)pr.x - begin;
    if (o >= window_lo && o < window_hi) out[o] = __uint_as_float(pr.y);
  }
}

// ---- publishing the upper tree levels ---------------------------------------------------------------------
__global__ void __launch_bounds__(256)
tops_publish_kernel(const TopsPublishParams p) {
  const TreeView& tr = p.tree;
  const int base_level = tr.levels > 1 ? 1 : 0;
  int64_t off = 0;
  for (int l = base_level; l < tr.levels; ++l) {
    const float4* src = reinterpret_cast<const float4*>(tr.boxes[l]);
    const int64_t n4 = 2 * (int64_t)tr.count[l];
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
      const float4 v = src[i];
      for (int r = 0; r < p.R; ++r)
        reinterpret_cast<float4*>(p.tables[r] + (size_t)p.rank * p.cap + off)[i] = v;
    }
    off += tr.count[l];
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    TopHead h;
    h.levels = tr.n > 0 ? tr.levels : 0;
    h.base_level = base_level;
    int64_t o = 0;
    for (int l = 0; l < kMaxLevels; ++l) {
      h.count[l] = l < tr.levels ? tr.count[l] : 0;
      h.offset[l] = (l >= base_level && l < tr.levels) ? (int)o : 0;
      if (l >= base_level && l < tr.levels) o += tr.count[l];
    }
    for (int a = 0; a < 6; ++a) h.box[a] = p.overall[a];
    for (int r = 0; r < p.R; ++r) p.heads[r][p.rank] = h;
  }
}

inline int blocks_for(int64_t n, int block) { return (int)div_up(n, block); }

}  // namespace

namespace {
template <bool SORTED>
int pv_launch_shared(PvMainParams& p, int max_tasks, int num_sms, cudaStream_t stream) {
  auto kernel = pv_main_kernel_t<false, SORTED>;
  const int warps = group_pick_warps(p.k, (const void*)kernel, SORTED);
  KNN_REQUIRE(warps >= 1, "pv kernel: k does not fit in shared memory");
  const size_t smem = group_smem_bytes(p.k, warps, SORTED);
  KNN_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  KNN_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, warps * 32, smem));
  KNN_REQUIRE(occ >= 1, "pv kernel does not fit on an SM");
  if (max_tasks <= 0) return 0;
  const long long want = ((long long)max_tasks + warps - 1) / warps;
  const long long cap = (long long)num_sms * occ;
  const int grid = (int)(want < cap ? want : cap);
  KNN_CUDA_TRY(cudaMemsetAsync(p.task_counter, 0, sizeof(unsigned long long), stream));
  kernel<<<grid, warps * 32, smem, stream>>>(p);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}
}  // namespace

int pv_main_launch(const PvMainParams& p_in, int max_tasks, int num_sms, cudaStream_t stream) {
  PvMainParams p = p_in;
  KNN_REQUIRE(p.k >= 1, "k must be >= 1");
  if (use_sorted_list(p.k)) return pv_launch_shared<true>(p, max_tasks, num_sms, stream);
  if (p.k == 1) return pv_launch_shared<false>(p, max_tasks, num_sms, stream);
  // k does not fit in shared memory: heaps go to a stream-ordered global scratch buffer (slower)
  if (max_tasks <= 0) return 0;
  const int gw = 8;
  const size_t gsmem = (size_t)gw * warp_smem_floats_noheap() * sizeof(float);
  int gocc = 0;
  KNN_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&gocc, pv_main_kernel_t<true, false>, gw * 32, gsmem));
  KNN_REQUIRE(gocc >= 1, "pv kernel does not fit on an SM");
  if (gocc > 2) gocc = 2;
  const long long gwant = ((long long)max_tasks + gw - 1) / gw, gcap = (long long)num_sms * gocc;
  const int ggrid = (int)(gwant < gcap ? gwant : gcap);
  float* heaps = nullptr;
  KNN_CUDA_TRY(cudaMallocAsync(&heaps, (size_t)ggrid * gw * heap_padded(p.k) * 32 * sizeof(float), stream));
  p.global_heaps = heaps;
  KNN_CUDA_TRY(cudaMemsetAsync(p.task_counter, 0, sizeof(unsigned long long), stream));
  pv_main_kernel_t<true, false><<<ggrid, gw * 32, gsmem, stream>>>(p);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  KNN_CUDA_TRY(cudaFreeAsync(heaps, stream));
  return 0;
}

bool pv_one_round_capable(int k) { return use_sorted_list(k); }

int pv_fused_launch(const PvMainParams& p_in, int phase, int max_tasks, int num_sms, cudaStream_t stream) {
  PvMainParams p = p_in;
  KNN_REQUIRE(use_sorted_list(p.k), "one-round exchange needs the sorted-list kernels");
  KNN_REQUIRE(phase == 2 || phase == 3, "pv_fused_launch: phase must be 2 or 3");
  auto kernel = phase == 2 ? pv_fused_kernel_t<2> : pv_fused_kernel_t<3>;
  const int warps = group_pick_warps(p.k, (const void*)kernel, true);
  KNN_REQUIRE(warps >= 1, "pv fused kernel: k does not fit in shared memory");
  const size_t smem = group_smem_bytes(p.k, warps, true);
  KNN_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  KNN_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, warps * 32, smem));
  KNN_REQUIRE(occ >= 1, "pv fused kernel does not fit on an SM");
  if (max_tasks <= 0) return 0;
  const long long want = ((long long)max_tasks + warps - 1) / warps, cap = (long long)num_sms * occ;
  KNN_CUDA_TRY(cudaMemsetAsync(p.task_counter, 0, sizeof(unsigned long long), stream));
  kernel<<<(int)(want < cap ? want : cap), warps * 32, smem, stream>>>(p);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

int pv_push_launch(const char* pending, const int* pending_count, int k, const TopsView& tops, const Box* territory,
                   const int* territory_count, int dst, char* peer_inbox, int* out_count, int num_sms, cudaStream_t stream) {
  KNN_CUDA_TRY(cudaMemsetAsync(out_count, 0, sizeof(int), stream));
  pv_push_kernel<<<num_sms * 8, 256, 0, stream>>>(pending, pending_count, k, tops, territory, territory_count, dst,
                                                 peer_inbox, out_count);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

int64_t tops_capacity(int64_t max_groups) {
  int64_t c = max_groups > 1 ? max_groups : 1, total = 0;
  if (c <= kFanout) return c + 64;                       // single-level tree: the leaf level is what gets published
  while (c > kFanout) { c = div_up(c, kFanout); total += c; }
  return total + 64;
}

int tops_publish_launch(const TopsPublishParams& p, int num_sms, cudaStream_t stream) {
  tops_publish_kernel<<<num_sms * 2, 256, 0, stream>>>(p);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

int dev_barrier_launch(const DevBarrierParams& p, cudaStream_t stream) {
  dev_barrier_kernel<<<1, 32, 0, stream>>>(p);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

int dev_collective_launch(const DevCollParams& p, cudaStream_t stream) {
  KNN_REQUIRE(p.R >= 1 && p.R <= kMaxRanks, "dev_collective: bad world size");
  KNN_REQUIRE(p.words >= 1 && p.words <= kCollWords, "dev_collective: payload must be 1..32 words");
  KNN_REQUIRE(p.op == kCollGather || p.words % 2 == 0, "dev_collective: reductions work on int64 elements");
  dev_collective_kernel<<<1, 32 * p.R, 0, stream>>>(p);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

int pv_publish_launch(const int* out_count, int* peer_inbox_count, cudaStream_t stream) {
  pv_publish_kernel<<<1, 32, 0, stream>>>(out_count, peer_inbox_count);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

int pv_final_launch(const char* pending, const int* pending_count, int k, float* result_sorted, int num_sms,
                    cudaStream_t stream) {
  pv_final_kernel<<<num_sms * 4, 256, 0, stream>>>(pending, pending_count, k, result_sorted);
  note_launch();
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}

int scatter_pairs_launch(const void* pairs, int64_t n, int64_t begin, float* out, cudaStream_t stream) {
  if (n <= 0) return 0;
  const int64_t window = (int64_t)20 << 20;             // 20 Mi results = 80 MB of output per window (L2: 126 MB)
  int64_t blocks = blocks_for(n, 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  for (int64_t lo = 0; lo < n; lo += window) {
    scatter_pairs_kernel<<<(int)blocks, 256, 0, stream>>>(static_cast<const uint2*>(pairs), n, begin, out, lo,
                                                        lo + window < n ? lo + window : n);
    note_launch();
  }
  KNN_CUDA_TRY(cudaGetLastError());
  return 0;
}


}  // namespace b200knn
