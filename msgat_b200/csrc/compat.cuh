// compat.cuh -- one spelling of launch / shared-memory / qualifiers for two builds:
//   * the product build: nvcc -gencode arch=compute_100a,code=sm_100a (libmsgat_sm100.so)
//   * the debug build used by tests/emu only: g++ -DMSGAT_EMU with tests/emu/emu_runtime.h,
//     which runs the same kernel bodies as host fibers (the build container has no GPU).
// The product never contains or loads the emulated build.
#pragma once
#include <stdint.h>

#ifdef MSGAT_EMU
#include "emu_runtime.h"
#define MSGAT_KERNEL static void
#define MSGAT_DEV static inline
#define MSGAT_LB(n, m)
#define MSGAT_RESTRICT __restrict__
#define MSGAT_DYN_SMEM(name) unsigned char* name = emu::cur()->smem
#define MSGAT_LAUNCH(kern, grid, block, smem, stream, ...) \
  emu::launch((grid), (block), (smem), [=]() { kern(__VA_ARGS__); })
#define MSGAT_UNROLL
#define MSGAT_UNROLL_N(n)
#define MSGAT_PIN(x)
static inline void msgat_prof_tag(const char*, int, int) {}
static inline int msgat_current_device() { return 0; }
static inline int msgat_sm_count() { return 148; }
#else
#include <cuda_runtime.h>
#include <atomic>
#define MSGAT_KERNEL static __global__ void
#define MSGAT_DEV __device__ __forceinline__
#define MSGAT_LB(n, m) __launch_bounds__(n, m)
#define MSGAT_RESTRICT __restrict__
#define MSGAT_DYN_SMEM(name) extern __shared__ __align__(1024) unsigned char name[]
#define MSGAT_MAX_SMEM 232448
int msgat_prof_begin(const char* name, cudaStream_t s);  // counts the launch; >= 0 when event timing is on
void msgat_prof_end(int slot, cudaStream_t s);
void msgat_prof_tag(const char* fmt, int a, int b);
int msgat_current_device();   // ordinal of the calling thread's current device
int msgat_sm_count();         // multiprocessors of the current device (148 on a B200), cached per device
// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device attribute of the function: one bit per device
// ordinal and call site (a process may drive several GPUs, e.g. under nn.DataParallel, main.py:53-54).
#define MSGAT_LAUNCH(kern, grid, block, smem, stream, ...)                                        \
  do {                                                                                            \
    static std::atomic<unsigned long long> _attr{0ull};                                           \
    const unsigned long long _bit = 1ull << (msgat_current_device() & 63);                        \
    if (!(_attr.load(std::memory_order_relaxed) & _bit)) {                                        \
      cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, MSGAT_MAX_SMEM);    \
      _attr.fetch_or(_bit, std::memory_order_relaxed);                                            \
    }                                                                                             \
    int _slot = msgat_prof_begin(#kern, (cudaStream_t)(stream));                                  \
    kern<<<(grid), (block), (smem), (cudaStream_t)(stream)>>>(__VA_ARGS__);                       \
    if (_slot >= 0) msgat_prof_end(_slot, (cudaStream_t)(stream));                                \
  } while (0)
#define MSGAT_UNROLL _Pragma("unroll")
#define MSGAT_STR_(x) #x
#define MSGAT_UNROLL_N(n) _Pragma(MSGAT_STR_(unroll n))
// keep a kernel-parameter value (e.g. dil[i], an indexed constant-bank load) in a register instead of
// letting the compiler reload it inside the inner loops
#define MSGAT_PIN(x) asm volatile("" : "+r"(x))
#endif

#define MSGAT_T12 12

MSGAT_DEV float warp_sum(float v) {
  v += __shfl_xor_sync(0xffffffffu, v, 16);
  v += __shfl_xor_sync(0xffffffffu, v, 8);
  v += __shfl_xor_sync(0xffffffffu, v, 4);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  return v;
}
MSGAT_DEV float warp_max(float v) {
  v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 16));
  v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 8));
  v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 4));
  v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 2));
  v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 1));
  return v;
}

// 16-byte asynchronous global -> shared copy (LDGSTS): the data never passes through registers, so a thread
// can keep the next rows of its stream in flight at no register cost.  live = false writes zeros.
#ifdef MSGAT_EMU
MSGAT_DEV void cp_async16(float* dst, const float* src, bool live) {
  for (int i = 0; i < 4; ++i) dst[i] = live ? src[i] : 0.f;
}
MSGAT_DEV void cp_async_commit() {}
template <int N> MSGAT_DEV void cp_async_wait() {}
#else
MSGAT_DEV void cp_async16(float* dst, const float* src, bool live) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst);
  const int n = live ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(src), "r"(n) : "memory");
}
MSGAT_DEV void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> MSGAT_DEV void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
#endif

// exp(x) for x <= 0 in the dense softmax passes: one multiply + MUFU.EX2 (ex2.approx, 2^-22 relative) instead of the
// ~10-instruction expf.  The product x * log2(e) is rounded once, so the relative error grows as |x| * 6e-8: the
// columns that matter to a softmax row have |x| of a few units, the ones with large |x| contribute ~0.
#ifdef MSGAT_EMU
MSGAT_DEV float exp_neg(float x) { return exp2f(x * 1.4426950408889634f); }
#else
MSGAT_DEV float exp_neg(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x * 1.4426950408889634f));
  return y;
}
#endif

// 1 / sqrt(v) of the LayerNorm rows to ~1 ulp: MUFU.RSQ (2 ulp, a relative error that is COMMON to the 12 values of a
// row and to its whole gradient, so it does not average out in the cancelling sums behind bias / LayerNorm-affine
// gradients) followed by one Newton step.  The reference's ATen LayerNorm divides by a correctly rounded sqrt.
#ifdef MSGAT_EMU
MSGAT_DEV float rsqrt_ln(float v) { return 1.0f / sqrtf(v); }
#else
MSGAT_DEV float rsqrt_ln(float v) {
  const float y = rsqrtf(v);
  return y * fmaf(-0.5f * v * y, y, 1.5f);
}
#endif

// two fp32 FMAs in one instruction (FFMA2, sm_100): halves the issue slots of the FMA-dense inner loops
#ifdef MSGAT_EMU
MSGAT_DEV float2 ffma2(float2 a, float2 b, float2 c) { return make_float2(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y)); }
MSGAT_DEV float2 fmul2(float2 a, float2 b) { return make_float2(a.x * b.x, a.y * b.y); }
#else
MSGAT_DEV float2 ffma2(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
MSGAT_DEV float2 fmul2(float2 a, float2 b) { return __fmul2_rn(a, b); }
#endif
MSGAT_DEV float2 lo2(const float4& v) { return make_float2(v.x, v.y); }
MSGAT_DEV float2 hi2(const float4& v) { return make_float2(v.z, v.w); }

MSGAT_DEV float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
MSGAT_DEV void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
MSGAT_DEV float4 f4zero() { return make_float4(0.f, 0.f, 0.f, 0.f); }
MSGAT_DEV void fma4(float4& a, float w, const float4& x) {
  a.x = fmaf(w, x.x, a.x); a.y = fmaf(w, x.y, a.y); a.z = fmaf(w, x.z, a.z); a.w = fmaf(w, x.w, a.w);
}
MSGAT_DEV float dot4(const float4& a, const float4& b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
// value of a 12-step row at t-d for t = t0..t0+3, given the quad at t0 and the quad before it
// (zeros when t0 == 0): the causal "past" tap of Conv2d(k=[1,2], dilation d) + Chomp(d).
// Two aligned float4 loads and a select on the sub-quad shift replace four predicated scalar loads:
// rows are 16-byte aligned, t0 is a multiple of 4, d is uniform over the block.
MSGAT_DEV float4 quad_shifted(const float4& A, const float4& B, int r) {
  if (r == 0) return A;
  if (r == 1) return make_float4(A.y, A.z, A.w, B.x);
  if (r == 2) return make_float4(A.z, A.w, B.x, B.y);
  return make_float4(A.w, B.x, B.y, B.z);
}
MSGAT_DEV float4 quad_past(const float* row, int t0, int d) {
  const int t = t0 - d, q0 = t & ~3, r = t & 3;   // floor to a quad boundary (also for negative t)
  const float4 A = (q0 >= 0) ? ld4(row + q0) : f4zero();
  const float4 B = (r != 0 && q0 + 4 >= 0) ? ld4(row + q0 + 4) : f4zero();   // q0 + 4 <= t0 < 12
  return quad_shifted(A, B, r);
}
// value at t+d (transpose of the past tap), zero beyond the 12-step window
MSGAT_DEV float4 quad_future(const float* row, int t0, int d) {
  const int t = t0 + d, q0 = t & ~3, r = t & 3;
  const float4 A = (q0 < MSGAT_T12) ? ld4(row + q0) : f4zero();
  const float4 B = (r != 0 && q0 + 4 < MSGAT_T12) ? ld4(row + q0 + 4) : f4zero();
  return quad_shifted(A, B, r);
}
