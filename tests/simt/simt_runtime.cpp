// simt_runtime.cpp — scheduler of the host emulation (see simt_host.h): blocks run one after the other,
// the threads of a block are fibers in one OS thread, switched by a dozen instructions of x86-64
// assembly (no system call, unlike swapcontext, which makes a launch of 256 blocks x 1024 threads cost
// milliseconds instead of seconds).
#include <execinfo.h>
#include <signal.h>
#include <unistd.h>
#include "simt_host.h"

// ThreadSanitizer build (make simt-tsan). By default it is compiled without function entry/exit
// instrumentation (--param tsan-instrument-func-entry-exit=0): the sanitizer then keeps no shadow call
// stack that the fiber switches could upset, all fibers of a rank are simply the rank's OS thread, and what
// is left for it to report are unordered accesses BETWEEN ranks, i.e. holes in the flag / barrier protocols.
// With -DSIMT_TSAN_FIBERS the fibers are announced through the sanitizer's fiber interface instead (full
// stacks in the reports, an order of magnitude slower).
#if defined(__SANITIZE_THREAD__) && defined(SIMT_TSAN_FIBERS)
extern "C" {
void* __tsan_get_current_fiber(void);
void* __tsan_create_fiber(unsigned flags);
void __tsan_destroy_fiber(void* fiber);
void __tsan_switch_to_fiber(void* fiber, unsigned flags);
}
#define SIMT_TSAN_SWITCH(f) __tsan_switch_to_fiber((f), 0)
#else
#define SIMT_TSAN_SWITCH(f) ((void)0)
#endif

#if !defined(__x86_64__)
#error "the fiber switch of the host emulation is written for x86-64"
#endif

// void simt_switch(void** save_sp, void* load_sp): push the callee-saved registers, store the stack
// pointer, adopt the other fiber's stack pointer, pop its registers, return into it.
asm(R"(
.text
.globl simt_switch
.type simt_switch,@function
simt_switch:
    pushq %rbp
    pushq %rbx
    pushq %r12
    pushq %r13
    pushq %r14
    pushq %r15
    movq %rsp, (%rdi)
    movq %rsi, %rsp
    popq %r15
    popq %r14
    popq %r13
    popq %r12
    popq %rbx
    popq %rbp
    ret
.size simt_switch,.-simt_switch
)");
extern "C" void simt_switch(void** save_sp, void* load_sp);

thread_local dim3 threadIdx, blockIdx, blockDim, gridDim;

namespace simt {
namespace {

constexpr size_t kFiberStack = 96 * 1024;

struct Fiber {
  void* sp = nullptr;
  void* tsan = nullptr;
  bool done = false;
};

struct State {
  void* sched_sp = nullptr;
  void* sched_tsan = nullptr;
  std::vector<void*> tsan_pool;          // sanitizer fibers are expensive to create: one pool per rank thread, reused
  std::vector<Fiber> fibers;
  std::vector<Warp> warps;
  Barrier block_bar;
  std::vector<char> stacks;
  char* dyn = nullptr;                   // dynamic shared memory of the current launch, exactly as large as requested
  int current = -1;
  unsigned long long progress = 0;       // barrier releases + fiber exits: no change over a whole round = deadlock
  const std::function<void()>* body = nullptr;
};
State& state() { static thread_local State s; return s; }

void yield() {
  State& s = state();
  SIMT_TSAN_SWITCH(s.sched_tsan);
  simt_switch(&s.fibers[s.current].sp, s.sched_sp);
}


void fiber_main() {
  State& s = state();
  (*s.body)();
  Fiber& f = s.fibers[s.current];
  f.done = true;
  ++s.progress;
  // A finished lane no longer takes part in barriers (CUDA: exited threads are not waited for) and votes
  // "false" from then on. Its exchange slots are left alone: slower lanes of the same collective may not
  // have read them yet.
  Warp& w = s.warps[s.current >> 5];
  w.live &= ~(1u << (s.current & 31));
  --w.bar.expected;
  if (w.bar.expected > 0 && w.bar.arrived >= w.bar.expected) { w.bar.arrived = 0; ++w.bar.gen; ++s.progress; }
  --s.block_bar.expected;
  if (s.block_bar.expected > 0 && s.block_bar.arrived >= s.block_bar.expected) { s.block_bar.arrived = 0; ++s.block_bar.gen; ++s.progress; }
  SIMT_TSAN_SWITCH(s.sched_tsan);
  simt_switch(&f.sp, s.sched_sp);
  abort();                               // a finished fiber is never resumed
}

void* fresh_stack(char* base) {
  // top of the stack, as simt_switch expects to find it: six register slots, then the address `ret` jumps to.
  // The return-address slot is 16-byte aligned so that fiber_main starts with the ABI's stack alignment.
  uintptr_t top = ((uintptr_t)base + kFiberStack) & ~(uintptr_t)15;
  void** ret_slot = (void**)(top - 16);
  *ret_slot = (void*)&fiber_main;
  void** regs = ret_slot - 6;
  for (int i = 0; i < 6; ++i) regs[i] = nullptr;
  return regs;
}

void on_alarm(int) {
  void* frames[48];
  const int n = backtrace(frames, 48);
  fprintf(stderr, "simt: launch still running, thread %u of block %u is here:\n", threadIdx.x, blockIdx.x);
  backtrace_symbols_fd(frames, n, 2);
  State& s = state();
  for (size_t w = 0; w < s.warps.size(); ++w) {
    int live = 0;
    for (int l = 0; l < 32 && w * 32 + l < s.fibers.size(); ++l) live += !s.fibers[w * 32 + l].done;
    fprintf(stderr, "  warp %zu: %d live lanes, barrier expects %d, arrived %d, generation %u\n", w, live,
            s.warps[w].bar.expected, s.warps[w].bar.arrived, s.warps[w].bar.gen);
  }
  fprintf(stderr, "  block barrier expects %d, arrived %d; progress counter %llu\n", s.block_bar.expected,
          s.block_bar.arrived, s.progress);
  _exit(3);
}

}  // namespace

void arrive_and_wait(Barrier& b) {
  const unsigned my = b.gen;
  if (++b.arrived >= b.expected) {
    b.arrived = 0;
    ++b.gen;
    ++state().progress;
    return;
  }
  while (b.gen == my) yield();
}

// Called from spin loops that wait for ANOTHER rank (its thread makes the progress): yield to the rest of
// this block and tell the deadlock detector that waiting here is legitimate.
void yield_if_fiber() {
  State& s = state();
  if (s.current < 0) return;
  ++s.progress;
  yield();
}
Barrier& block_barrier() { return state().block_bar; }
Warp& my_warp() { State& s = state(); return s.warps[s.current >> 5]; }
int my_lane() { return state().current & 31; }

void* dyn_smem() { return state().dyn; }

// SIMT_DEBUG=<seconds>: print every launch, and if one launch runs longer than that, say where it is.
void launch(const char* name, dim3 grid, dim3 block, size_t smem_bytes, const std::function<void()>& body) {
  State& s = state();
  const int nthreads = (int)(block.x * block.y * block.z);
  static const bool debug = getenv("SIMT_DEBUG") != nullptr;
  if (debug) {
    fprintf(stderr, "simt: %s<<<%u, %u, %zu>>>\n", name, grid.x, block.x, smem_bytes);
    const int secs = atoi(getenv("SIMT_DEBUG"));
    if (secs > 0) { signal(SIGALRM, on_alarm); alarm((unsigned)secs); }
  }
  if (nthreads <= 0 || grid.x == 0) return;
  if (block.y != 1 || block.z != 1 || grid.y != 1 || grid.z != 1) { fprintf(stderr, "simt: only 1-D launches\n"); abort(); }
  if (s.current >= 0) { fprintf(stderr, "simt: kernels cannot launch kernels\n"); abort(); }
  if (s.stacks.size() < (size_t)nthreads * kFiberStack) s.stacks.resize((size_t)nthreads * kFiberStack);
  free(s.dyn);
  s.dyn = nullptr;
  if (posix_memalign((void**)&s.dyn, 128, smem_bytes ? smem_bytes : 1)) abort();
  memset(s.dyn, 0, smem_bytes ? smem_bytes : 1);
  s.body = &body;
  gridDim = grid;
  blockDim = block;
  for (unsigned b = 0; b < grid.x; ++b) {
    blockIdx = dim3(b);
    s.fibers.assign(nthreads, Fiber());
    s.warps.assign((nthreads + 31) / 32, Warp());
    for (int t = 0; t < nthreads; ++t) { ++s.warps[t >> 5].bar.expected; s.warps[t >> 5].live |= 1u << (t & 31); }
    s.block_bar = Barrier();
    s.block_bar.expected = nthreads;
    for (int t = 0; t < nthreads; ++t) s.fibers[t].sp = fresh_stack(s.stacks.data() + (size_t)t * kFiberStack);
#if defined(__SANITIZE_THREAD__) && defined(SIMT_TSAN_FIBERS)
    s.sched_tsan = __tsan_get_current_fiber();
    while ((int)s.tsan_pool.size() < nthreads) s.tsan_pool.push_back(__tsan_create_fiber(0));
    for (int t = 0; t < nthreads; ++t) s.fibers[t].tsan = s.tsan_pool[t];
#endif
    int live = nthreads, stuck = 0;
    while (live > 0) {
      const unsigned long long before = s.progress;
      for (int t = 0; t < nthreads; ++t) {
        Fiber& f = s.fibers[t];
        if (f.done) continue;
        s.current = t;
        threadIdx = dim3((unsigned)t);
        SIMT_TSAN_SWITCH(f.tsan);
        simt_switch(&s.sched_sp, f.sp);
        if (f.done) --live;
      }
      // a whole round in which no barrier was released and no fiber finished: every fiber is parked at a
      // collective that can never complete (lanes of one warp took different paths to it)
      stuck = (s.progress == before) ? stuck + 1 : 0;
      if (stuck > 2) { fprintf(stderr, "simt: %s, block %u: deadlock at a collective that not all lanes reach\n", name, b); abort(); }
    }

  }
  s.current = -1;
  s.body = nullptr;
  if (debug) alarm(0);
}

}  // namespace simt
