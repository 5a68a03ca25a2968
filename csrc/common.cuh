// common.cuh — error handling, small device helpers, system-scope flag primitives.
// Role of cukd/common.h:276-343 (CUKD_CUDA_CHECK/CALL) in the reference; re-designed:
// library code never throws or exits, it records an error string and returns a code.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include "knn_math.h"

namespace b200knn {

void set_last_error(const std::string& msg);   // api.cpp
const char* last_error_cstr();
void keep_pool_memory(int device);               // api.cu: raise the default mempool release threshold
void note_launch();                             // api.cu: counts kernel launches (bench evidence)

#define KNN_CUDA_TRY(expr)                                                               \
  do {                                                                                   \
    cudaError_t _e = (expr);                                                             \
    if (_e != cudaSuccess) {                                                             \
      ::b200knn::set_last_error(std::string(#expr) + " failed: " +                       \
                                cudaGetErrorString(_e) + " (" + __FILE__ + ":" +         \
                                std::to_string(__LINE__) + ")");                         \
      return (int)_e == 0 ? -1 : (int)_e;                                                \
    }                                                                                    \
  } while (0)

#define KNN_REQUIRE(cond, msg)                                                           \
  do {                                                                                   \
    if (!(cond)) {                                                                       \
      ::b200knn::set_last_error(std::string(msg) + " (" + __FILE__ + ":" +               \
                                std::to_string(__LINE__) + ")");                         \
      return -2;                                                                         \
    }                                                                                    \
  } while (0)

// Stream-ordered device allocation that goes back to the pool on EVERY path out of a function (library code
// returns error codes from the middle of long pipelines; without this each early return leaked its temporaries).
// release() frees early, at the point where the pipeline is done with the buffer.
template <class T>
struct StreamBuf {
  T* p = nullptr;
  cudaStream_t stream;
  explicit StreamBuf(cudaStream_t s) : stream(s) {}
  StreamBuf(const StreamBuf&) = delete;
  StreamBuf& operator=(const StreamBuf&) = delete;
  ~StreamBuf() { release(); }
  cudaError_t alloc_bytes(size_t bytes) {
    release();
    return cudaMallocAsync(&p, bytes, stream);
  }
  void release() {
    if (p) cudaFreeAsync(p, stream);
    p = nullptr;
  }
  T* get() const { return p; }
  operator T*() const { return p; }
};

}  // namespace b200knn
#include <nvtx3/nvToolsExt.h>
namespace b200knn {
// Named range visible in Nsight Systems / ncu (role of the reference's getCurrentTime() brackets,
// cukd/common.h:219-237, which exist only in its tests; the MPI programs have no timers at all).
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
};

#if defined(__CUDACC__)

constexpr unsigned kFullMask = 0xffffffffu;

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

// Order-preserving float <-> uint mapping (for atomicMin/Max on floats of either sign).
// Same idea as the reference's AtomicBox encoding (cukd/spatial-kdtree.h:102-187).
__device__ __forceinline__ unsigned f2ord(float f) {
  unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float ord2f(unsigned u) {
  u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
#if defined(__CUDA_ARCH__)
  return __uint_as_float(u);
#else
  float f; memcpy(&f, &u, 4); return f;
#endif
}

// ---- system-scope release/acquire flags (visible across NVLink peers) ----
#if defined(B200KNN_SIMT_HOST) && defined(B200KNN_SIMT_BROKEN_FLAGS)
// Self-test of the race detector (make simt-tsan-selftest): flags without release/acquire must be reported.
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) { *(volatile unsigned*)p = v; }
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) { return *(volatile const unsigned*)p; }
__device__ __forceinline__ unsigned ld_relaxed_sys(const unsigned* p) { return *(volatile const unsigned*)p; }
#elif defined(B200KNN_SIMT_HOST)   // host emulation of the kernels (tests/simt): ranks are threads, flags are C++ atomics
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) { __atomic_store_n(p, v, __ATOMIC_RELEASE); }
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) { return __atomic_load_n(p, __ATOMIC_ACQUIRE); }
__device__ __forceinline__ unsigned ld_relaxed_sys(const unsigned* p) { return __atomic_load_n(p, __ATOMIC_RELAXED); }
#else
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned ld_relaxed_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
#endif
// Spin until *p >= want (sequence numbers only grow). Called by one lane.
// Failure detection: a peer that died (or never launched) would otherwise leave this GPU spinning
// forever. After ~60 s of waiting on one flag the kernel traps, which surfaces on the host as a
// launch failure from the next synchronisation instead of a hang.
__device__ __forceinline__ void wait_flag_ge(const unsigned* p, unsigned want) {
  unsigned ns = 32;
  long long start = 0;
  unsigned spins = 0;
  while ((int)(ld_acquire_sys(p) - want) < 0) {
    __nanosleep(ns);
    if (ns < 1024) ns <<= 1;
    if ((++spins & 0xfffu) == 0u) {
      const long long now = clock64();
      if (start == 0) start = now;
      else if (now - start > 120000000000ll) __trap();     // ~60 s at 2 GHz
    }
  }
}

// L1-bypassing loads for buffers a peer GPU writes into (L2 is the coherence point).
__device__ __forceinline__ float ldcg_f(const float* p) { return __ldcg(p); }
__device__ __forceinline__ float4 ldcg_f4(const float4* p) { return __ldcg(p); }
__device__ __forceinline__ int ldcg_i(const int* p) { return __ldcg(p); }

#endif  // __CUDACC__

}  // namespace b200knn
