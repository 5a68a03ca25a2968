// simt_host.h — run this library's CUDA kernels on the CPU, for tests on machines without a GPU.
//
// The kernels are compiled UNCHANGED as host C++ (tests/simt/preprocess.py only rewrites the
// `kernel<<<grid, block, smem, stream>>>(args)` launch syntax and `extern __shared__` declarations).
// Execution model:
//   * one OS thread per emulated GPU (a "rank" of the multi-GPU tests; peer memory is plain shared
//     memory, CUDA IPC handles are pointers); within it blocks of a grid run one after the other;
//   * the threads of a block are fibers (ucontext) inside ONE operating-system thread, resumed
//     round-robin; a fiber runs until it reaches a warp- or block-level collective (__shfl*_sync,
//     __ballot_sync, __reduce_*_sync, __match_any_sync, __syncwarp, __syncthreads) and yields there
//     until the rest of its warp / block has arrived;
//   * `__shared__` variables are statics (shared by all fibers, stale between blocks like the real
//     thing), dynamic shared memory is one buffer per launch, "device" memory is host memory, streams
//     are synchronous.
// What it checks: the logic of the kernels — indexing, warp-collective protocols, heap and traversal
// code, launch geometry, and the flag protocols between ranks as far as a sequentially consistent
// x86 host exercises them. What it cannot check: the GPU memory model, performance.
#pragma once
#include <time.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <functional>
#include <vector>

#define __CUDACC__ 1
#define B200KNN_SIMT_HOST 1
#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __restrict__
#define __launch_bounds__(...)
#define __align__(n) __attribute__((aligned(n)))
#define __shared__ static thread_local      // one OS thread per emulated GPU (rank): its blocks run one at a time

// ---------------------------------------------------------------- vector types
struct alignas(16) float4 { float x, y, z, w; };
struct alignas(8) float2 { float x, y; };
struct alignas(8) uint2 { unsigned x, y; };
struct alignas(16) uint4 { unsigned x, y, z, w; };
struct alignas(16) int4 { int x, y, z, w; };
struct dim3 { unsigned x = 1, y = 1, z = 1; dim3() {} dim3(unsigned a, unsigned b = 1, unsigned c = 1) : x(a), y(b), z(c) {} };
inline float4 make_float4(float x, float y, float z, float w) { return float4{x, y, z, w}; }
inline uint2 make_uint2(unsigned x, unsigned y) { return uint2{x, y}; }
inline uint4 make_uint4(unsigned x, unsigned y, unsigned z, unsigned w) { return uint4{x, y, z, w}; }
inline int4 make_int4(int x, int y, int z, int w) { return int4{x, y, z, w}; }

namespace simt {

struct Barrier { int expected = 0, arrived = 0; unsigned gen = 0; };
struct Warp {
  Barrier bar;
  unsigned live = 0;            // lanes that have not returned from the kernel yet
  uint64_t slot[2][32];
};

// tests/simt/simt_runtime.cpp
void launch(const char* name, dim3 grid, dim3 block, size_t smem_bytes, const std::function<void()>& body);
void arrive_and_wait(Barrier& b);       // generation barrier among the cooperative fibers of a warp / block
Barrier& block_barrier();
Warp& my_warp();
int my_lane();
void* dyn_smem();
void yield_if_fiber();                  // spin loops on another rank's flags must let their own block run

// Deposit one 64-bit word per lane, wait for the warp, return the buffer all lanes wrote into. Buffers
// alternate by barrier generation: a lane can only reach collective c+2 after everybody has read c.
inline const uint64_t* exchange(uint64_t v) {
  Warp& w = my_warp();
  uint64_t* buf = w.slot[w.bar.gen & 1u];
  buf[my_lane()] = v;
  arrive_and_wait(w.bar);
  return buf;
}

template <class T> inline uint64_t to_bits(T v) { uint64_t u = 0; static_assert(sizeof(T) <= 8, "shuffle type"); memcpy(&u, &v, sizeof(T)); return u; }
template <class T> inline T from_bits(uint64_t u) { T v; memcpy(&v, &u, sizeof(T)); return v; }

}  // namespace simt

// Built-in variables: one OS thread, so plain globals that the scheduler rewrites before every resume.
extern thread_local dim3 threadIdx, blockIdx, blockDim, gridDim;

// ---------------------------------------------------------------- warp / block collectives
inline void __syncwarp(unsigned = 0xffffffffu) { simt::arrive_and_wait(simt::my_warp().bar); }
inline void __syncthreads() { simt::arrive_and_wait(simt::block_barrier()); }
inline void __threadfence() { __atomic_thread_fence(__ATOMIC_SEQ_CST); }
inline void __threadfence_block() {}
inline void __threadfence_system() { __atomic_thread_fence(__ATOMIC_SEQ_CST); }

template <class T> inline T __shfl_sync(unsigned, T v, int src, int = 32) {
  const uint64_t* b = simt::exchange(simt::to_bits(v));
  return simt::from_bits<T>(b[src & 31]);
}
template <class T> inline T __shfl_xor_sync(unsigned, T v, int lane_mask, int = 32) {
  const int lane = simt::my_lane();
  const uint64_t* b = simt::exchange(simt::to_bits(v));
  return simt::from_bits<T>(b[(lane ^ lane_mask) & 31]);
}
template <class T> inline T __shfl_up_sync(unsigned, T v, unsigned delta, int = 32) {
  const int lane = simt::my_lane();
  const uint64_t* b = simt::exchange(simt::to_bits(v));
  return simt::from_bits<T>(b[lane >= (int)delta ? lane - (int)delta : lane]);
}
template <class T> inline T __shfl_down_sync(unsigned, T v, unsigned delta, int = 32) {
  const int lane = simt::my_lane();
  const uint64_t* b = simt::exchange(simt::to_bits(v));
  return simt::from_bits<T>(b[lane + (int)delta < 32 ? lane + (int)delta : lane]);
}
inline unsigned __ballot_sync(unsigned, int pred) {
  simt::Warp& w = simt::my_warp();
  const unsigned voters = w.live;           // lanes taking part in THIS vote (they are all still running)
  const uint64_t* b = simt::exchange(pred ? 1u : 0u);
  unsigned m = 0;
  for (int i = 0; i < 32; ++i) m |= (b[i] ? 1u : 0u) << i;
  return m & voters;
}
inline int __any_sync(unsigned m, int pred) { return __ballot_sync(m, pred) != 0u; }
inline int __all_sync(unsigned m, int pred) { return __ballot_sync(m, !pred) == 0u; }
inline unsigned __activemask() { return 0xffffffffu; }
inline unsigned __reduce_min_sync(unsigned, unsigned v) {
  const uint64_t* b = simt::exchange(v);
  unsigned r = 0xffffffffu;
  for (int i = 0; i < 32; ++i) r = (unsigned)b[i] < r ? (unsigned)b[i] : r;
  return r;
}
inline unsigned __reduce_max_sync(unsigned, unsigned v) {
  const uint64_t* b = simt::exchange(v);
  unsigned r = 0;
  for (int i = 0; i < 32; ++i) r = (unsigned)b[i] > r ? (unsigned)b[i] : r;
  return r;
}
inline unsigned __reduce_add_sync(unsigned, unsigned v) {
  const uint64_t* b = simt::exchange(v);
  unsigned r = 0;
  for (int i = 0; i < 32; ++i) r += (unsigned)b[i];
  return r;
}
inline unsigned __reduce_or_sync(unsigned, unsigned v) {
  const uint64_t* b = simt::exchange(v);
  unsigned r = 0;
  for (int i = 0; i < 32; ++i) r |= (unsigned)b[i];
  return r;
}
template <class T> inline unsigned __match_any_sync(unsigned, T v) {
  const uint64_t mine = simt::to_bits(v);
  const uint64_t* b = simt::exchange(mine);
  unsigned m = 0;
  for (int i = 0; i < 32; ++i) m |= (b[i] == mine ? 1u : 0u) << i;
  return m;
}

// ---------------------------------------------------------------- scalar intrinsics
inline unsigned __float_as_uint(float f) { unsigned u; memcpy(&u, &f, 4); return u; }
inline int __float_as_int(float f) { int u; memcpy(&u, &f, 4); return u; }
inline float __uint_as_float(unsigned u) { float f; memcpy(&f, &u, 4); return f; }
inline float __int_as_float(int u) { float f; memcpy(&f, &u, 4); return f; }
inline float __fmaf_rn(float a, float b, float c) { return fmaf(a, b, c); }
inline float __fmul_rn(float a, float b) { return a * b; }
inline float __fadd_rn(float a, float b) { return a + b; }
inline float __fsub_rn(float a, float b) { return a - b; }
inline int __ffs(unsigned v) { return __builtin_ffs((int)v); }
inline int __popc(unsigned v) { return __builtin_popcount(v); }
inline int __clz(unsigned v) { return v ? __builtin_clz(v) : 32; }
inline int __clzll(unsigned long long v) { return v ? __builtin_clzll(v) : 64; }
template <class T> inline T __ldg(const T* p) { return *p; }
template <class T> inline T __ldcg(const T* p) { return *p; }
template <class T> inline T __ldcs(const T* p) { return *p; }
inline void __nanosleep(unsigned) { simt::yield_if_fiber(); }
inline long long clock64() { static thread_local long long c = 0; return c += 1000; }
inline void __trap() { fprintf(stderr, "simt: __trap()\n"); abort(); }

// Atomics are real atomics: another rank's thread may touch the same word (peer memory).
template <class T> inline T atomicAdd(T* p, T v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
inline unsigned long long atomicAdd(unsigned long long* p, unsigned long long v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
template <class T> inline T atomicAdd_system(T* p, T v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
template <class T> inline T atomicExch(T* p, T v) { return __atomic_exchange_n(p, v, __ATOMIC_SEQ_CST); }
template <class T> inline T atomicCAS(T* p, T cmp, T v) { __atomic_compare_exchange_n(p, &cmp, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST); return cmp; }
template <class T> inline T atomicOr(T* p, T v) { return __atomic_fetch_or(p, v, __ATOMIC_SEQ_CST); }
template <class T> inline T atomicMin(T* p, T v) { T o = __atomic_load_n(p, __ATOMIC_SEQ_CST); while (v < o && !__atomic_compare_exchange_n(p, &o, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) {} return o; }
template <class T> inline T atomicMax(T* p, T v) { T o = __atomic_load_n(p, __ATOMIC_SEQ_CST); while (v > o && !__atomic_compare_exchange_n(p, &o, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) {} return o; }

// ---------------------------------------------------------------- runtime API (synchronous, host memory)
typedef int cudaError_t;
constexpr cudaError_t cudaSuccess = 0;
constexpr cudaError_t cudaErrorNotSupported = 801;
typedef struct simt_stream* cudaStream_t;
typedef struct simt_event* cudaEvent_t;
typedef void* cudaMemPool_t;
enum cudaMemcpyKind { cudaMemcpyHostToHost, cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice, cudaMemcpyDefault };
enum cudaDeviceAttr { cudaDevAttrMultiProcessorCount = 16 };
enum cudaFuncAttribute { cudaFuncAttributeMaxDynamicSharedMemorySize = 8 };
enum cudaMemPoolAttr { cudaMemPoolAttrReleaseThreshold = 4, cudaMemPoolAttrReservedMemCurrent = 5, cudaMemPoolAttrUsedMemHigh = 8 };
constexpr unsigned cudaIpcMemLazyEnablePeerAccess = 1;
struct cudaIpcMemHandle_t { char reserved[64]; };
struct cudaFuncAttributes { int numRegs = 0; size_t sharedSizeBytes = 0; int maxThreadsPerBlock = 1024; };

inline const char* cudaGetErrorString(cudaError_t e) { return e == cudaSuccess ? "no error" : "simt: operation not available on the host emulation"; }
inline cudaError_t cudaGetLastError() { return cudaSuccess; }
// exactly n bytes (256-byte aligned like the real allocator): under AddressSanitizer an access one byte past a
// "device" buffer is reported
inline cudaError_t cudaMalloc(void** p, size_t n) { *p = nullptr; return posix_memalign(p, 256, n ? n : 1) == 0 ? cudaSuccess : 2; }
template <class T> inline cudaError_t cudaMalloc(T** p, size_t n) { return cudaMalloc((void**)p, n); }
inline cudaError_t cudaFree(void* p) { free(p); return cudaSuccess; }
constexpr unsigned cudaHostAllocDefault = 0, cudaHostAllocPortable = 1, cudaHostAllocMapped = 2;
inline cudaError_t cudaHostAlloc(void** p, size_t n, unsigned) { *p = nullptr; return posix_memalign(p, 256, n ? n : 1) == 0 ? cudaSuccess : 2; }
inline cudaError_t cudaHostGetDevicePointer(void** d, void* h, unsigned) { *d = h; return cudaSuccess; }
template <class T> inline cudaError_t cudaMallocAsync(T** p, size_t n, cudaStream_t) { return cudaMalloc((void**)p, n); }
inline cudaError_t cudaFreeAsync(void* p, cudaStream_t) { free(p); return cudaSuccess; }
inline cudaError_t cudaMemcpy(void* d, const void* s, size_t n, cudaMemcpyKind) { if (n) memmove(d, s, n); return cudaSuccess; }
inline cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t = nullptr) { if (n) memmove(d, s, n); return cudaSuccess; }
inline cudaError_t cudaMemset(void* d, int v, size_t n) { if (n) memset(d, v, n); return cudaSuccess; }
inline cudaError_t cudaMemsetAsync(void* d, int v, size_t n, cudaStream_t = nullptr) { if (n) memset(d, v, n); return cudaSuccess; }
inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
// events: everything runs synchronously here, so an event is a host timestamp
struct simt_event { double ms; };
inline cudaError_t cudaEventCreate(cudaEvent_t* e) { *e = new simt_event{0.0}; return cudaSuccess; }
inline cudaError_t cudaEventDestroy(cudaEvent_t e) { delete e; return cudaSuccess; }
inline cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t = nullptr) {
  timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); e->ms = ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6; return cudaSuccess;
}
inline cudaError_t cudaEventElapsedTime(float* out, cudaEvent_t a, cudaEvent_t b) { *out = (float)(b->ms - a->ms); return cudaSuccess; }
inline cudaError_t cudaDeviceSynchronize() { return cudaSuccess; }
inline cudaError_t cudaGetDevice(int* d) { *d = 0; return cudaSuccess; }
inline cudaError_t cudaSetDevice(int) { return cudaSuccess; }
inline cudaError_t cudaDeviceGetAttribute(int* v, cudaDeviceAttr, int) { *v = 2; return cudaSuccess; }   // "2 SMs": small grids
inline cudaError_t cudaDeviceGetDefaultMemPool(cudaMemPool_t* p, int) { *p = nullptr; return cudaSuccess; }
inline cudaError_t cudaMemPoolSetAttribute(cudaMemPool_t, cudaMemPoolAttr, void*) { return cudaSuccess; }
inline cudaError_t cudaMemPoolGetAttribute(cudaMemPool_t, cudaMemPoolAttr, void* v) { *static_cast<uint64_t*>(v) = 0; return cudaSuccess; }
template <class F> inline cudaError_t cudaFuncSetAttribute(F, cudaFuncAttribute, int) { return cudaSuccess; }
template <class F> inline cudaError_t cudaFuncGetAttributes(cudaFuncAttributes* a, F) { *a = cudaFuncAttributes(); return cudaSuccess; }
template <class F> inline cudaError_t cudaOccupancyMaxActiveBlocksPerMultiprocessor(int* n, F, int, size_t) { *n = 1; return cudaSuccess; }
template <class T> inline cudaError_t cudaGetSymbolAddress(void** p, T& symbol) { *p = (void*)&symbol; return cudaSuccess; }
// "Peer memory": ranks are threads of one process, so a handle is the pointer itself.
inline cudaError_t cudaIpcGetMemHandle(cudaIpcMemHandle_t* h, void* p) { memset(h, 0, sizeof(*h)); memcpy(h->reserved, &p, sizeof(p)); return cudaSuccess; }
inline cudaError_t cudaIpcOpenMemHandle(void** p, cudaIpcMemHandle_t h, unsigned) { memcpy(p, h.reserved, sizeof(*p)); return cudaSuccess; }
inline cudaError_t cudaIpcCloseMemHandle(void*) { return cudaSuccess; }
