// cuda_emu.h -- TEST INFRASTRUCTURE ONLY: a small SIMT emulator for checking kernel LOGIC on a CPU.
//
// The product's kernels live in devspace_b200/csrc/*.cuh. This header lets g++ compile those same
// headers (with -DDSG_EMU) and run a grid one block at a time, each CUDA thread as a fiber (a hand-written context
// switch on x86-64 -- swapcontext costs two signal-mask system calls per switch -- ucontext elsewhere):
//   * __syncthreads / __syncwarp / warp collectives (__shfl*, __ballot, __match_any, __reduce_*) are
//     rendezvous points; a fiber that reaches one yields until its peers have arrived, so code that
//     forgets a barrier sees a legal (non-lockstep) interleaving and fails here rather than on the GPU;
//   * atomics are plain read-modify-writes (one OS thread runs everything);
//   * `__shared__` becomes `static` (one block runs at a time), dynamic shared memory is emu::dyn_smem().
// It is compiled only by tests/ (tests/emu_harness.cc); libdsgpack.so never contains or calls it, and
// nothing here is a fallback for the CUDA path.
#pragma once

#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <ucontext.h>

#include <algorithm>
#include <functional>
#include <stdexcept>
#include <vector>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __shared__ static
#define __restrict__
#define __launch_bounds__(...)
#define __align__(n) alignas(n)
#define __noinline__

struct dim3 {
  unsigned x, y, z;
  dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};
struct uint4 {
  unsigned x, y, z, w;
};
struct uint2 {
  unsigned x, y;
};
inline uint4 make_uint4(unsigned a, unsigned b, unsigned c, unsigned d) { return uint4{a, b, c, d}; }
inline uint2 make_uint2(unsigned a, unsigned b) { return uint2{a, b}; }

namespace emu {

#if defined(__x86_64__) && !defined(EMU_USE_UCONTEXT)
#define EMU_ASM_SWITCH 1
// void emu_switch_asm(void **save_sp, void *new_sp): callee-saved registers on the old stack, stack pointers swapped
extern "C" void emu_switch_asm(void **save_sp, void *new_sp);
asm(R"(
.text
.globl emu_switch_asm
.type emu_switch_asm,@function
emu_switch_asm:
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
.size emu_switch_asm, .-emu_switch_asm
)");
#endif

struct Fiber {
#if defined(EMU_ASM_SWITCH)
  void *sp = nullptr;
#else
  ucontext_t ctx;
#endif
  char *stack = nullptr;
  bool done = false;
  bool at_barrier = false;   // parked in __syncthreads: not scheduled until the barrier generation changes
  unsigned bar_gen_seen = 0;
  dim3 tid;
};

struct WarpState {
  unsigned arrived_mask = 0;
  unsigned gen = 0;
  uint64_t vals[32];
  int args[32];  // per-lane operand of a collective (the source lane of an indexed shuffle)
  uint64_t res[2][32];
  unsigned res_mask[2];
};

struct BlockState {
  Fiber *fibers = nullptr;
  std::vector<WarpState> warps;
  unsigned nthreads = 0;
  unsigned live = 0;
  unsigned bar_arrived = 0;
  unsigned bar_gen = 0;
  int current = -1;
  unsigned nb_arrived[16] = {0};
  unsigned nb_gen[16] = {0};
#if defined(EMU_ASM_SWITCH)
  void *sched_sp = nullptr;
#else
  ucontext_t sched;
#endif
  std::function<void()> body;
  std::vector<unsigned char> dyn;
};

inline BlockState *&cur_block() {
  static BlockState *b = nullptr;
  return b;
}
inline unsigned char *dyn_smem() { return cur_block()->dyn.data(); }

inline void yield() {
  BlockState *b = cur_block();
#if defined(EMU_ASM_SWITCH)
  emu_switch_asm(&b->fibers[b->current].sp, b->sched_sp);
#else
  swapcontext(&b->fibers[b->current].ctx, &b->sched);
#endif
}

}  // namespace emu

// CUDA built-in variables (one OS thread => plain globals, set on every fiber switch)
extern dim3 threadIdx, blockIdx, blockDim, gridDim;

namespace emu {

inline void fiber_entry() {
  BlockState *b = cur_block();
  b->body();
  b->fibers[b->current].done = true;
  b->live--;
#if defined(EMU_ASM_SWITCH)
  emu_switch_asm(&b->fibers[b->current].sp, b->sched_sp);  // never resumed
  abort();
#else
  swapcontext(&b->fibers[b->current].ctx, &b->sched);
#endif
}

// Run one grid. `body` is called once per thread and should call the kernel with its arguments.
inline void launch(dim3 grid, dim3 block, size_t dyn_bytes, const std::function<void()> &body,
                   size_t stack_bytes = 96 * 1024) {
  const unsigned nthreads = block.x * block.y * block.z;
  gridDim = grid;
  blockDim = block;
  std::vector<Fiber> fibers(nthreads);
  std::vector<char> stacks((size_t)nthreads * stack_bytes);  // reused by every block of the grid
  for (unsigned t = 0; t < nthreads; t++) fibers[t].stack = stacks.data() + (size_t)t * stack_bytes;
  for (unsigned bz = 0; bz < grid.z; bz++)
    for (unsigned by = 0; by < grid.y; by++)
      for (unsigned bx = 0; bx < grid.x; bx++) {
        BlockState bs;
        bs.nthreads = nthreads;
        bs.live = nthreads;
        bs.body = body;
        bs.dyn.assign(dyn_bytes + 256, 0xCD);  // poison: kernels must not assume zeroed shared memory
        bs.fibers = fibers.data();
        bs.warps.resize((nthreads + 31) / 32);
        cur_block() = &bs;
        for (unsigned t = 0; t < nthreads; t++) {
          Fiber &f = fibers[t];
          f.done = false;
          f.at_barrier = false;
          f.tid = dim3(t % block.x, (t / block.x) % block.y, t / (block.x * block.y));
#if defined(EMU_ASM_SWITCH)
          {  // a stack that emu_switch_asm "returns" from into fiber_entry: six zeroed registers, the entry address
            // in a 16-byte aligned slot (so that the stack is aligned as behind a call), one slot of padding above
            uintptr_t top = (reinterpret_cast<uintptr_t>(f.stack) + stack_bytes) & ~uintptr_t(15);
            void **slot = reinterpret_cast<void **>(top - 16);
            slot[0] = reinterpret_cast<void *>(&fiber_entry);
            slot[1] = nullptr;
            for (int k = 1; k <= 6; k++) slot[-k] = nullptr;
            f.sp = slot - 6;
          }
#else
          getcontext(&f.ctx);
          f.ctx.uc_stack.ss_sp = f.stack;
          f.ctx.uc_stack.ss_size = stack_bytes;
          f.ctx.uc_link = &bs.sched;
          makecontext(&f.ctx, (void (*)())fiber_entry, 0);
#endif
        }
        blockIdx = dim3(bx, by, bz);
        unsigned long spins = 0;
        // EMU_SCHED=reverse|random runs the fibers in another order: code that silently relies on threads
        // executing in thread-id order (e.g. atomics "arriving" in stream order) then fails here too
        static const char *sched_env = getenv("EMU_SCHED");
        std::vector<unsigned> order(nthreads);
        for (unsigned t = 0; t < nthreads; t++) order[t] = t;
        if (sched_env && sched_env[0] == 'r' && sched_env[1] == 'e') std::reverse(order.begin(), order.end());
        unsigned lcg = 12345u + bx;
        while (bs.live > 0) {
          unsigned ran = 0;
          if (sched_env && sched_env[0] == 'r' && sched_env[1] == 'a')
            for (unsigned t = nthreads - 1; t > 0; t--) {
              lcg = lcg * 1664525u + 1013904223u;
              std::swap(order[t], order[(lcg >> 8) % (t + 1)]);
            }
          for (unsigned oi = 0; oi < nthreads; oi++) {
            const unsigned t = order[oi];
            Fiber &f = fibers[t];
            if (f.done) continue;
            if (f.at_barrier && f.bar_gen_seen == bs.bar_gen) continue;  // still parked
            bs.current = (int)t;
            threadIdx = f.tid;
#if defined(EMU_ASM_SWITCH)
            emu_switch_asm(&bs.sched_sp, f.sp);
#else
            swapcontext(&bs.sched, &f.ctx);
#endif
            ran++;
          }
          if (ran == 0) {
            // everybody alive is parked at the barrier: threads that exited do not take part
            if (bs.bar_arrived >= bs.live && bs.live > 0) {
              bs.bar_arrived = 0;
              bs.bar_gen++;
            } else {
              throw std::runtime_error("emu: deadlock (barrier never satisfied)");
            }
          }
          if (++spins > 2000000000ul) throw std::runtime_error("emu: livelock");
        }
        cur_block() = nullptr;
      }
}

// bar.sync id, count: `count` threads of the block rendezvous on barrier `id` (1..15)
inline void named_barrier_sync(int id, int count) {
  BlockState *b = cur_block();
  const unsigned my_gen = b->nb_gen[id];
  if (++b->nb_arrived[id] >= (unsigned)count) {
    b->nb_arrived[id] = 0;
    b->nb_gen[id]++;
    return;
  }
  while (b->nb_gen[id] == my_gen) yield();
}

// bar.arrive id, count: counts the thread in without waiting
inline void named_barrier_arrive(int id, int count) {
  BlockState *b = cur_block();
  if (++b->nb_arrived[id] >= (unsigned)count) {
    b->nb_arrived[id] = 0;
    b->nb_gen[id]++;
  }
}

inline unsigned lane_id() { return (threadIdx.x + threadIdx.y * blockDim.x + threadIdx.z * blockDim.x * blockDim.y) & 31; }
inline unsigned warp_id() { return (threadIdx.x + threadIdx.y * blockDim.x + threadIdx.z * blockDim.x * blockDim.y) >> 5; }

// Generic warp rendezvous: every lane in `mask` deposits `v`; `combine(vals, mask, lane)` is evaluated
// per lane once all have arrived.
template <typename F>
inline uint64_t warp_collective(unsigned mask, uint64_t v, F combine, int arg = 0) {
  BlockState *b = cur_block();
  WarpState &w = b->warps[warp_id()];
  const unsigned lane = lane_id();
  w.args[lane] = arg;
  // lanes of a partial last warp that do not exist never arrive: clip the mask
  unsigned first = warp_id() * 32;
  unsigned exist = b->nthreads - first >= 32 ? 0xffffffffu : ((1u << (b->nthreads - first)) - 1u);
  mask &= exist;
  if (!(mask & (1u << lane))) throw std::runtime_error("emu: lane not in its own sync mask");
  const unsigned my_gen = w.gen;
  w.vals[lane] = v;
  w.arrived_mask |= 1u << lane;
  if ((w.arrived_mask & mask) == mask) {
    const unsigned slot = my_gen & 1;
    for (unsigned l = 0; l < 32; l++)
      if (mask & (1u << l)) w.res[slot][l] = combine(w.vals, mask, l);
    w.res_mask[slot] = mask;
    w.arrived_mask &= ~mask;
    w.gen++;
  } else {
    while (w.gen == my_gen) yield();
  }
  return w.res[my_gen & 1][lane];
}

}  // namespace emu

// ---- synchronisation -------------------------------------------------------------------------------
inline void __syncthreads() {
  emu::BlockState *b = emu::cur_block();
  emu::Fiber &f = b->fibers[b->current];
  const unsigned my_gen = b->bar_gen;
  b->bar_arrived++;
  if (b->bar_arrived >= b->live) {  // last arriver releases everybody (exited threads do not take part)
    b->bar_arrived = 0;
    b->bar_gen++;
    return;
  }
  f.at_barrier = true;
  f.bar_gen_seen = my_gen;
  while (b->bar_gen == my_gen) emu::yield();  // parked: the scheduler skips this fiber until the release
  f.at_barrier = false;
}
inline void __syncwarp(unsigned mask = 0xffffffffu) {
  emu::warp_collective(mask, 0, [](const uint64_t *, unsigned, unsigned) { return (uint64_t)0; });
}
inline long long clock64() { return 0; }
inline void __threadfence() {}
inline void __threadfence_block() {}

// ---- warp collectives --------------------------------------------------------------------------------
inline unsigned __ballot_sync(unsigned mask, int pred) {
  return (unsigned)emu::warp_collective(mask, pred ? 1 : 0, [](const uint64_t *v, unsigned m, unsigned) {
    unsigned r = 0;
    for (unsigned l = 0; l < 32; l++)
      if ((m & (1u << l)) && v[l]) r |= 1u << l;
    return (uint64_t)r;
  });
}
inline int __any_sync(unsigned mask, int pred) { return __ballot_sync(mask, pred) != 0; }
inline int __all_sync(unsigned mask, int pred) {
  unsigned m = mask;
  return __ballot_sync(mask, pred) == (m & __ballot_sync(mask, 1));
}
template <typename T>
inline T emu_shfl(unsigned mask, T var, int mode, int arg, int width) {
  static_assert(sizeof(T) <= 8, "shfl of <= 64-bit values only");
  uint64_t bits = 0;
  memcpy(&bits, &var, sizeof(T));
  // the lambda runs in the fiber of the LAST lane to arrive, for every lane: a per-lane operand (the source lane of
  // an indexed shuffle may differ from lane to lane) is read from the warp state, not from the capture
  const int *args = emu::cur_block()->warps[emu::warp_id()].args;
  uint64_t r = emu::warp_collective(mask, bits, [=](const uint64_t *v, unsigned m, unsigned lane) {
    int base = (int)lane & ~(width - 1);
    int src;
    const int arg = args[lane];
    switch (mode) {
      case 0: src = base + (arg & (width - 1)); break;                       // idx
      case 1: src = (int)lane - arg; if (src < base) src = (int)lane; break;  // up
      case 2: src = (int)lane + arg; if (src >= base + width) src = (int)lane; break;  // down
      default: src = (int)lane ^ arg; if (src >= base + width) src = (int)lane; break;  // xor
    }
    if (src < 0 || src > 31 || !(m & (1u << src))) return v[lane];  // undefined on HW: return own value
    return v[src];
  }, arg);
  T out;
  memcpy(&out, &r, sizeof(T));
  return out;
}
template <typename T> inline T __shfl_sync(unsigned m, T v, int src, int w = 32) { return emu_shfl(m, v, 0, src, w); }
template <typename T> inline T __shfl_up_sync(unsigned m, T v, unsigned d, int w = 32) { return emu_shfl(m, v, 1, (int)d, w); }
template <typename T> inline T __shfl_down_sync(unsigned m, T v, unsigned d, int w = 32) { return emu_shfl(m, v, 2, (int)d, w); }
template <typename T> inline T __shfl_xor_sync(unsigned m, T v, int x, int w = 32) { return emu_shfl(m, v, 3, x, w); }

inline unsigned __match_any_sync(unsigned mask, uint64_t value) {
  return (unsigned)emu::warp_collective(mask, value, [](const uint64_t *v, unsigned m, unsigned lane) {
    unsigned r = 0;
    for (unsigned l = 0; l < 32; l++)
      if ((m & (1u << l)) && v[l] == v[lane]) r |= 1u << l;
    return (uint64_t)r;
  });
}
#define EMU_REDUCE(name, init, op)                                                                   \
  inline unsigned name(unsigned mask, unsigned value) {                                              \
    return (unsigned)emu::warp_collective(mask, value, [](const uint64_t *v, unsigned m, unsigned) { \
      unsigned r = init;                                                                             \
      for (unsigned l = 0; l < 32; l++)                                                              \
        if (m & (1u << l)) r = op(r, (unsigned)v[l]);                                                \
      return (uint64_t)r;                                                                            \
    });                                                                                              \
  }
inline unsigned emu_or(unsigned a, unsigned b) { return a | b; }
inline unsigned emu_and(unsigned a, unsigned b) { return a & b; }
inline unsigned emu_add(unsigned a, unsigned b) { return a + b; }
inline unsigned emu_max(unsigned a, unsigned b) { return a > b ? a : b; }
inline unsigned emu_min(unsigned a, unsigned b) { return a < b ? a : b; }
EMU_REDUCE(__reduce_or_sync, 0u, emu_or)
EMU_REDUCE(__reduce_and_sync, 0xffffffffu, emu_and)
EMU_REDUCE(__reduce_add_sync, 0u, emu_add)
EMU_REDUCE(__reduce_max_sync, 0u, emu_max)
EMU_REDUCE(__reduce_min_sync, 0xffffffffu, emu_min)
inline unsigned __activemask() { return 0xffffffffu; }

// ---- atomics (single OS thread: plain RMW) --------------------------------------------------------------
template <typename T> inline T atomicAdd(T *p, T v) { T o = *p; *p = o + v; return o; }
template <typename T> inline T atomicSub(T *p, T v) { T o = *p; *p = o - v; return o; }
template <typename T> inline T atomicExch(T *p, T v) { T o = *p; *p = v; return o; }
template <typename T> inline T atomicMin(T *p, T v) { T o = *p; if (v < o) *p = v; return o; }
template <typename T> inline T atomicMax(T *p, T v) { T o = *p; if (v > o) *p = v; return o; }
template <typename T> inline T atomicOr(T *p, T v) { T o = *p; *p = o | v; return o; }
template <typename T> inline T atomicAnd(T *p, T v) { T o = *p; *p = o & v; return o; }
template <typename T> inline T atomicXor(T *p, T v) { T o = *p; *p = o ^ v; return o; }
template <typename T> inline T atomicCAS(T *p, T cmp, T v) { T o = *p; if (o == cmp) *p = v; return o; }

// ---- integer intrinsics ------------------------------------------------------------------------------------
inline int __popc(unsigned x) { return __builtin_popcount(x); }
inline int __popcll(unsigned long long x) { return __builtin_popcountll(x); }
inline int __clz(int x) { return x ? __builtin_clz((unsigned)x) : 32; }
inline int __clzll(long long x) { return x ? __builtin_clzll((unsigned long long)x) : 64; }
inline int __ffs(int x) { return __builtin_ffs(x); }
inline int __ffsll(long long x) { return __builtin_ffsll(x); }
inline unsigned __brev(unsigned x) {
  unsigned r = 0;
  for (int i = 0; i < 32; i++) r |= ((x >> i) & 1u) << (31 - i);
  return r;
}
inline unsigned __funnelshift_r(unsigned lo, unsigned hi, unsigned s) {
  uint64_t v = ((uint64_t)hi << 32) | lo;
  return (unsigned)(v >> (s & 31));
}
inline unsigned __funnelshift_l(unsigned lo, unsigned hi, unsigned s) {
  uint64_t v = ((uint64_t)hi << 32) | lo;
  return (unsigned)((v << (s & 31)) >> 32);
}
inline unsigned __byte_perm(unsigned a, unsigned b, unsigned sel) {
  uint64_t v = ((uint64_t)b << 32) | a;
  unsigned r = 0;
  for (int i = 0; i < 4; i++) {
    unsigned s = (sel >> (4 * i)) & 0xf;
    unsigned byte = (unsigned)(v >> (8 * (s & 7))) & 0xff;
    if (s & 8) byte = (byte & 0x80) ? 0xff : 0x00;
    r |= byte << (8 * i);
  }
  return r;
}
template <typename T> inline T __ldg(const T *p) { return *p; }
using std::max;
using std::min;
