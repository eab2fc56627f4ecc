// Host-side fiber emulator of the CUDA execution model.  TEST INFRASTRUCTURE ONLY.
//
// The build container has nvcc but no GPU.  To debug the indexing and the math of
// the CUDA-core kernels before spending GPU minutes, tests/emu/build_emu.py compiles
// the SAME kernel sources (msgat_b200/csrc/*.cu) with g++ -DMSGAT_EMU against this
// header: every CUDA thread of a block becomes a ucontext fiber, __syncthreads and
// warp shuffles are scheduler yield points, blocks run one after another on one OS
// thread.  Shared memory is poisoned with NaNs per block so reads of unwritten smem
// show up.  Nothing under msgat_b200/ loads the resulting library; only
// tests/test_emu_*.py do (marked "not gpu").  It is not a product path and never a
// fallback: msgat_b200/_lib.py refuses to run without a CUDA device.
#pragma once
#include <ucontext.h>

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <vector>

struct uint3 { unsigned x, y, z; };
struct dim3 {
  unsigned x, y, z;
  dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};
struct alignas(16) float4 { float x, y, z, w; };
struct alignas(8) float2 { float x, y; };
struct alignas(16) int4 { int x, y, z, w; };
static inline float4 make_float4(float a, float b, float c, float d) { return float4{a, b, c, d}; }
static inline float2 make_float2(float a, float b) { return float2{a, b}; }

typedef void* cudaStream_t;
typedef int cudaError_t;
#define cudaSuccess 0
static inline cudaError_t cudaGetLastError() { return 0; }
static inline cudaError_t cudaPeekAtLastError() { return 0; }
static inline const char* cudaGetErrorString(cudaError_t) { return "emu"; }
static inline cudaError_t cudaMemsetAsync(void* p, int v, size_t n, cudaStream_t) { memset(p, v, n); return 0; }

namespace emu {

struct Ctx {
  uint3 tid, bid;
  dim3 bdim, gdim;
  int lin;  // linear thread id in block
  unsigned char* smem;
  int nb_id = 0, nb_need = 0;  // named barrier the fiber is parked at (state AT_NAMED)
};

enum { RUN = 0, AT_BLOCK = 1, AT_WARP = 2, DONE = 3, SPIN = 4, AT_NAMED = 5 };

struct Fiber {
  ucontext_t uc;
  char* stack = nullptr;
  int state = DONE;
  Ctx ctx;
};

struct Sched {
  // tcgen05 / mbarrier emulation (tc_common.cuh under MSGAT_EMU): tensor memory of the running block and an event
  // counter that every barrier arrival bumps, so that a round of pure spinning without any event is a deadlock
  std::vector<float> tmem;          // [128 lanes][512 columns]
  unsigned long long events = 0;
  int named_id[16] = {0}, named_need[16] = {0};
  ucontext_t main_uc;
  std::vector<Fiber> fibers;
  std::vector<uint64_t> slots;  // one per thread, warp shuffle exchange
  Fiber* running = nullptr;
  std::function<void()>* body = nullptr;
};

inline Sched& sched() { static Sched s; return s; }
inline Ctx*& cur() { static Ctx* c = nullptr; return c; }

inline void yield_to_sched(int st) {
  Sched& s = sched();
  Fiber* f = s.running;
  f->state = st;
  swapcontext(&f->uc, &s.main_uc);
}

inline void syncthreads() { yield_to_sched(AT_BLOCK); }
inline void spin_yield() { yield_to_sched(SPIN); }
// bar.sync id, nthreads: released when `nthreads` fibers of the block wait on the same id
inline void named_sync(int id, int nthreads) { cur()->nb_id = id; cur()->nb_need = nthreads; yield_to_sched(AT_NAMED); }
inline void event() { sched().events++; }
inline void syncwarp() { yield_to_sched(AT_WARP); }

inline uint64_t shfl_idx(uint64_t v, int src) {
  Sched& s = sched();
  int lin = cur()->lin;
  s.slots[lin] = v;
  syncwarp();
  uint64_t r = s.slots[(lin & ~31) | (src & 31)];
  syncwarp();
  return r;
}

static void fiber_entry() {
  Sched& s = sched();
  (*s.body)();
  s.running->state = DONE;
  swapcontext(&s.running->uc, &s.main_uc);
}

inline void run_block(dim3 grid, dim3 block, uint3 bid, size_t smem_bytes, std::function<void()>& body) {
  Sched& s = sched();
  const int nt = block.x * block.y * block.z;
  const size_t STACK = 256 * 1024;
  if ((int)s.fibers.size() < nt) s.fibers.resize(nt);
  s.slots.assign(nt, 0);
  std::vector<unsigned char> smem(smem_bytes + 64);
  memset(smem.data(), 0xFF, smem.size());  // NaN poison
  unsigned char* sbase = (unsigned char*)(((uintptr_t)smem.data() + 15) & ~(uintptr_t)15);
  s.body = &body;
  for (int t = 0; t < nt; ++t) {
    Fiber& f = s.fibers[t];
    if (!f.stack) f.stack = (char*)malloc(STACK);
    getcontext(&f.uc);
    f.uc.uc_stack.ss_sp = f.stack;
    f.uc.uc_stack.ss_size = STACK;
    f.uc.uc_link = nullptr;
    makecontext(&f.uc, (void (*)())fiber_entry, 0);
    f.state = RUN;
    f.ctx.tid = uint3{(unsigned)(t % block.x), (unsigned)((t / block.x) % block.y), (unsigned)(t / (block.x * block.y))};
    f.ctx.bid = bid;
    f.ctx.bdim = block;
    f.ctx.gdim = grid;
    f.ctx.lin = t;
    f.ctx.smem = sbase;
  }
  s.tmem.assign(128 * 512, 0.f);
  int idle_rounds = 0;
  for (;;) {
    bool progressed = false;
    int alive = 0;
    const unsigned long long ev0 = s.events;
    for (int t = 0; t < nt; ++t) {
      Fiber& f = s.fibers[t];
      if (f.state == RUN || f.state == SPIN) {
        const bool spinning = f.state == SPIN;
        f.state = RUN;
        s.running = &f;
        cur() = &f.ctx;
        swapcontext(&s.main_uc, &f.uc);
        if (!(spinning && f.state == SPIN)) progressed = true;   // a fiber that only re-polls made no progress
      }
      if (f.state != DONE) ++alive;
    }
    if (s.events != ev0) progressed = true;
    if (alive == 0) break;
    bool released = false;
    // block-wide barrier: every live thread must have arrived
    int at_block = 0;
    for (int t = 0; t < nt; ++t) at_block += s.fibers[t].state == AT_BLOCK;
    if (at_block == alive) {
      for (int t = 0; t < nt; ++t) if (s.fibers[t].state == AT_BLOCK) s.fibers[t].state = RUN;
      released = true;
    }
    // named barriers
    for (int id = 1; id < 16; ++id) {
      int cnt = 0, need = 0;
      for (int t = 0; t < nt; ++t) if (s.fibers[t].state == AT_NAMED && s.fibers[t].ctx.nb_id == id) { ++cnt; need = s.fibers[t].ctx.nb_need; }
      if (cnt && cnt == need) {
        for (int t = 0; t < nt; ++t) if (s.fibers[t].state == AT_NAMED && s.fibers[t].ctx.nb_id == id) s.fibers[t].state = RUN;
        released = true;
      }
    }
    // warp-level sync points
    for (int w = 0; w * 32 < nt; ++w) {
      int live = 0, atw = 0;
      for (int t = w * 32; t < nt && t < w * 32 + 32; ++t) {
        live += s.fibers[t].state != DONE;
        atw += s.fibers[t].state == AT_WARP;
      }
      if (live && atw == live) {
        for (int t = w * 32; t < nt && t < w * 32 + 32; ++t) if (s.fibers[t].state == AT_WARP) s.fibers[t].state = RUN;
        released = true;
      }
    }
    if (!progressed && !released) {
      if (++idle_rounds < 4) continue;   // spinners get a few more polls before the verdict
      fprintf(stderr, "emu: deadlock in block (%u,%u,%u): divergent barrier or mbarrier wait that can never complete\n", bid.x, bid.y, bid.z);
      for (int t = 0; t < nt; ++t) if (s.fibers[t].state != DONE) fprintf(stderr, "  thread %d state %d\n", t, s.fibers[t].state);
      abort();
    }
    idle_rounds = 0;
  }
}

template <class F>
inline void launch(dim3 grid, dim3 block, size_t smem, F&& f) {
  std::function<void()> body = f;
  for (unsigned z = 0; z < grid.z; ++z)
    for (unsigned y = 0; y < grid.y; ++y)
      for (unsigned x = 0; x < grid.x; ++x) run_block(grid, block, uint3{x, y, z}, smem, body);
}

}  // namespace emu

#define threadIdx (emu::cur()->tid)
#define blockIdx (emu::cur()->bid)
#define blockDim (emu::cur()->bdim)
#define gridDim (emu::cur()->gdim)

static inline void __syncthreads() { emu::syncthreads(); }
static inline void __syncwarp(unsigned = 0xffffffffu) { emu::syncwarp(); }
static inline float __shfl_xor_sync(unsigned, float v, int m) {
  uint32_t u; memcpy(&u, &v, 4);
  uint64_t r = emu::shfl_idx(u, (emu::cur()->lin & 31) ^ m);
  u = (uint32_t)r; memcpy(&v, &u, 4); return v;
}
static inline double __shfl_xor_sync(unsigned, double v, int m) {
  uint64_t u; memcpy(&u, &v, 8);
  uint64_t r = emu::shfl_idx(u, (emu::cur()->lin & 31) ^ m);
  memcpy(&v, &r, 8); return v;
}
static inline float __shfl_sync(unsigned, float v, int src) {
  uint32_t u; memcpy(&u, &v, 4);
  uint64_t r = emu::shfl_idx(u, src);
  u = (uint32_t)r; memcpy(&v, &u, 4); return v;
}
static inline float __shfl_down_sync(unsigned, float v, int d) {
  uint32_t u; memcpy(&u, &v, 4);
  int lane = emu::cur()->lin & 31;
  uint64_t r = emu::shfl_idx(u, lane + d < 32 ? lane + d : lane);
  u = (uint32_t)r; memcpy(&v, &u, 4); return v;
}
static inline float atomicAdd(float* p, float v) { float o = *p; *p = o + v; return o; }
static inline int atomicAdd(int* p, int v) { int o = *p; *p = o + v; return o; }
static inline unsigned atomicAdd(unsigned* p, unsigned v) { unsigned o = *p; *p = o + v; return o; }
static inline void __threadfence() {}
template <class T> static inline T __ldg(const T* p) { return *p; }
#define __expf(x) expf(x)
static inline float rsqrtf(float x) { return 1.0f / sqrtf(x); }
static inline float __fdividef(float a, float b) { return a / b; }
