// tc_common.cuh -- sm_100a tensor-core plumbing shared by the tc_*.cu kernels: tcgen05.mma
// (kind::tf32, accumulators in TMEM), TMEM allocation and tcgen05.ld, mbarriers, and the one
// shared-memory operand layout every MS-GAT tile uses.
//
// Operand layout ("SW128 tile").  A tile holds R rows (channels) of W consecutive positions
// (W a multiple of 32, a position = one (node, t) pair: the (n, t) plane of [B,C,N,12] is
// contiguous in HBM, msgat.py:121).  It is stored as W/32 column blocks; a column block is
// R/8 atoms of 8 rows x 128 bytes with the 128-byte swizzle (16-byte chunk index ^= row & 7):
//     byte(r, p) = (p / 32) * CB + (r / 8) * 1024 + (r % 8) * 128 + ((((p % 32) / 4) ^ (r % 8)) * 16) + (p % 4) * 4
// with CB = (R / 8) * 1024.  The same bytes serve tcgen05.mma in both roles it needs:
//   * as an MN-major operand (positions = M or N, rows = K): the channel mixes
//     D[pos, o] = sum_c X[c, pos] W[c, o]                      (LBO = CB, SBO = 1024)
//   * as a K-major operand (rows = M or N, positions = K): the weight gradients
//     dW[o, c] = sum_pos G[o, pos] X[c, pos]                    (SBO = 1024, 32 B per k-step)
#pragma once
#ifdef MSGAT_EMU
// ---------------------------------------------------------------------------------------------------------------
// Host-fiber emulation of the primitives below (tests/emu only; the product build never sees this branch): mbarriers
// as counters in the emulated shared memory, tensor memory as a per-block array, tcgen05.mma executed synchronously
// at issue by decoding the SAME descriptors the hardware gets (operand values truncated to TF32).  It checks the
// kernels' indexing, pipeline protocol and arithmetic; it cannot check asynchronous hazards or memory-model fences.
#include <stdint.h>
#include "compat.cuh"
#define TC_DEV static inline
TC_DEV uint32_t smem_u32(const void* p) { return (uint32_t)((const unsigned char*)p - emu::cur()->smem); }
TC_DEV unsigned char* emu_smem_at(uint32_t off) { return emu::cur()->smem + off; }
TC_DEV uint32_t sw128_off(int r, int p, uint32_t cb) {
  return (uint32_t)(p >> 5) * cb + (uint32_t)(r >> 3) * 1024u + (uint32_t)(r & 7) * 128u +
         ((((uint32_t)(p & 31) >> 2) ^ (uint32_t)(r & 7)) << 4) + ((uint32_t)(p & 3) << 2);
}
TC_DEV uint32_t core_off(int r, int p, uint32_t gs) {
  return (uint32_t)(r >> 3) * gs + (uint32_t)(p >> 2) * 128u + (uint32_t)(r & 7) * 16u + ((uint32_t)(p & 3) << 2);
}
TC_DEV uint32_t sw128b32_off(int r, int p, uint32_t cb) {
  return (uint32_t)(p >> 5) * cb + (uint32_t)(r >> 2) * 512u + (uint32_t)(r & 3) * 128u +
         ((((uint32_t)(p & 31) >> 3) ^ (uint32_t)(r & 3)) << 5) + ((uint32_t)(p & 7) << 2);
}
struct EmuBar { uint32_t count, pending; uint32_t phase; int32_t tx; };   // lives in 16 bytes of emulated shared memory
TC_DEV EmuBar* emu_bar(uint32_t bar) { return reinterpret_cast<EmuBar*>(emu_smem_at(bar)); }
TC_DEV void mbar_init(uint32_t bar, uint32_t count) { EmuBar* b = emu_bar(bar); b->count = b->pending = count; b->phase = 0; b->tx = 0; }
TC_DEV void emu_bar_check(EmuBar* b) { if (b->pending == 0 && b->tx == 0) { b->pending = b->count; b->phase ^= 1u; } emu::event(); }
TC_DEV void mbar_arrive(uint32_t bar) {
  EmuBar* b = emu_bar(bar);
  if (b->pending == 0) { fprintf(stderr, "emu: arrive on an uninitialised mbarrier\n"); abort(); }
  --b->pending;
  emu_bar_check(b);
}
// arrive and announce `bytes` of asynchronous copies that will complete on this barrier
TC_DEV void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  EmuBar* b = emu_bar(bar);
  if (b->pending == 0) { fprintf(stderr, "emu: arrive on an uninitialised mbarrier\n"); abort(); }
  b->tx += (int32_t)bytes; --b->pending;
  emu_bar_check(b);
}
// cp.async.bulk global -> shared (1-D TMA), completion counted in bytes on the barrier
TC_DEV void bulk_g2s(void* dst_smem, const void* src, uint32_t bytes, uint32_t bar) {
  if ((bytes & 15u) || ((uintptr_t)src & 15u) || (smem_u32(dst_smem) & 15u)) { fprintf(stderr, "emu: cp.async.bulk needs 16-byte aligned operands and size\n"); abort(); }
  memcpy(dst_smem, src, bytes);
  EmuBar* b = emu_bar(bar);
  b->tx -= (int32_t)bytes;
  emu_bar_check(b);
}
TC_DEV void mbar_fence_init() {}
TC_DEV void mbar_wait(uint32_t bar, uint32_t parity) {
  while (emu_bar(bar)->phase == (parity & 1u)) emu::spin_yield();
}
TC_DEV void named_bar_sync(int id, int nthreads) { emu::named_sync(id, nthreads); }
TC_DEV void tc_fence_before_sync() {}
TC_DEV void tc_fence_after_sync() {}
TC_DEV void fence_smem_to_async() {}
TC_DEV void tmem_alloc(uint32_t slot_smem, uint32_t) { *reinterpret_cast<uint32_t*>(emu_smem_at(slot_smem)) = 0u; __syncwarp(); }
TC_DEV void tmem_dealloc(uint32_t, uint32_t) {}
TC_DEV uint32_t tmem_addr(uint32_t base, int lane, int col) { return base + ((uint32_t)lane << 16) + (uint32_t)col; }
TC_DEV uint32_t warp_uniform(uint32_t v) { return v; }
TC_DEV float* emu_tmem(uint32_t addr) {
  const int lane_base = (int)(addr >> 16), col = (int)(addr & 0xffffu), warp = emu::cur()->lin >> 5, lane = emu::cur()->lin & 31;
  if (lane_base != 32 * (warp & 3)) { fprintf(stderr, "emu: warp %d reads TMEM lanes %d.. (allowed: %d..)\n", warp, lane_base, 32 * (warp & 3)); abort(); }
  if (col < 0 || col + 8 > 512) { fprintf(stderr, "emu: TMEM column %d out of range\n", col); abort(); }
  return emu::sched().tmem.data() + (size_t)(lane_base + lane) * 512 + col;
}
TC_DEV void tmem_ld8(uint32_t addr, float (&v)[8]) { const float* p = emu_tmem(addr); for (int i = 0; i < 8; ++i) v[i] = p[i]; }
TC_DEV void tmem_ld8_nowait(uint32_t addr, uint32_t (&r)[8]) { const float* p = emu_tmem(addr); memcpy(r, p, 32); }
TC_DEV void tmem_ld_wait() {}
TC_DEV float __uint_as_float(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
TC_DEV uint32_t __float_as_uint(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
TC_DEV uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout_type = 2) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)layout_type << 61;
  return d;
}
TC_DEV uint32_t umma_idesc_tf32(int M, int N, int a_mn_major, int b_mn_major) {
  uint32_t d = 0;
  d |= 1u << 4; d |= 2u << 7; d |= 2u << 10;
  d |= (uint32_t)(a_mn_major & 1) << 15;
  d |= (uint32_t)(b_mn_major & 1) << 16;
  d |= (uint32_t)(N >> 3) << 17;
  d |= (uint32_t)(M >> 4) << 24;
  return d;
}
TC_DEV float emu_tf32(float x) { uint32_t u; memcpy(&u, &x, 4); u &= 0xffffe000u; memcpy(&x, &u, 4); return x; }
// element (row = M or N index, k) of an operand described by `desc`; one instruction covers k = 0..7
TC_DEV float emu_operand(uint64_t desc, int mn_major, int idx, int k) {
  const uint32_t start = (uint32_t)(desc & 0x3FFFu) << 4, lbo = (uint32_t)((desc >> 16) & 0x3FFFu) << 4,
                 sbo = (uint32_t)((desc >> 32) & 0x3FFFu) << 4, layout = (uint32_t)(desc >> 61);
  uint32_t off;
  if (mn_major) {
    if (layout != 1 || sbo != 512u) { fprintf(stderr, "emu: MN-major TF32 operands need layout 1 (128B swizzle, 32B atoms), SBO 512\n"); abort(); }
    off = sw128b32_off(k, idx, lbo);
  } else {
    if (layout != 0 || lbo != 128u) { fprintf(stderr, "emu: K-major operands are emulated for the unswizzled core-matrix layout only\n"); abort(); }
    off = core_off(idx, k, sbo);
  }
  float v;
  memcpy(&v, emu_smem_at(start + off), 4);
  return emu_tf32(v);
}
TC_DEV void umma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  const int M = (int)((idesc >> 24) & 0x1fu) << 4, N = (int)((idesc >> 17) & 0x3fu) << 3, amn = (idesc >> 15) & 1, bmn = (idesc >> 16) & 1;
  if (M != 128 || N < 16 || N > 256 || (N & 15)) { fprintf(stderr, "emu: tcgen05.mma shape M=%d N=%d not supported\n", M, N); abort(); }
  const int col = (int)(d_tmem & 0xffffu);
  if (col + N > 512) { fprintf(stderr, "emu: accumulator columns %d..%d beyond TMEM\n", col, col + N); abort(); }
  float* T = emu::sched().tmem.data();
  for (int m = 0; m < M; ++m)
    for (int n = 0; n < N; ++n) {
      float s = accumulate ? T[(size_t)m * 512 + col + n] : 0.f;
      for (int k = 0; k < 8; ++k) s += emu_operand(adesc, amn, m, k) * emu_operand(bdesc, bmn, n, k);
      T[(size_t)m * 512 + col + n] = s;
    }
}
TC_DEV void umma_commit(uint32_t bar) { mbar_arrive(bar); }
// converged-warp variants: the emulated warp's lane 0 is the elected lane
TC_DEV void umma_tf32_e(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  if ((threadIdx.x & 31) == 0) umma_tf32(d_tmem, adesc, bdesc, idesc, accumulate);
}
TC_DEV void umma_commit_e(uint32_t bar) { if ((threadIdx.x & 31) == 0) umma_commit(bar); }
#else

#include <cuda_runtime.h>
#include <stdint.h>

#define TC_DEV __device__ __forceinline__

TC_DEV uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// byte offset of element (row r, position p) inside a SW128 tile whose column blocks are cb bytes apart
TC_DEV uint32_t sw128_off(int r, int p, uint32_t cb) {
  return (uint32_t)(p >> 5) * cb + (uint32_t)(r >> 3) * 1024u + (uint32_t)(r & 7) * 128u +
         ((((uint32_t)(p & 31) >> 2) ^ (uint32_t)(r & 7)) << 4) + ((uint32_t)(p & 3) << 2);
}

// Unswizzled "core matrix" tile: 8 rows x 4 positions (128 contiguous bytes) per core matrix,
// core matrices of one 8-row group contiguous along the positions, row groups gs bytes apart.
TC_DEV uint32_t core_off(int r, int p, uint32_t gs) {
  return (uint32_t)(r >> 3) * gs + (uint32_t)(p >> 2) * 128u + (uint32_t)(r & 7) * 16u + ((uint32_t)(p & 3) << 2);
}
// 128-byte swizzle with 32-byte atoms (the swizzled MN-major layout of 32-bit operands):
// 4-row atoms of 128 bytes per row, 32-byte chunk index ^= row & 3.
TC_DEV uint32_t sw128b32_off(int r, int p, uint32_t cb) {
  return (uint32_t)(p >> 5) * cb + (uint32_t)(r >> 2) * 512u + (uint32_t)(r & 3) * 128u +
         ((((uint32_t)(p & 31) >> 3) ^ (uint32_t)(r & 3)) << 5) + ((uint32_t)(p & 7) << 2);
}

// ---- mbarrier ----------------------------------------------------------------------------
TC_DEV void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
TC_DEV void mbar_arrive(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }
TC_DEV void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
// Waits for the phase of the given parity to complete.  try_wait suspends the warp in hardware up to the time hint,
// so a waiting warp costs almost no issue slots (a plain test-and-branch loop of the 8 epilogue warps of the
// pipelined kernels executed 180 M instructions per launch); a wait that can never complete traps after 2^20
// expirations (<= 4 us each) instead of hanging the device.  The whole loop is one PTX block: 3 instructions per expiry.
// The branches are NOT marked .uni: the lanes of a warp poll separately and may see the phase flip in different iterations.
TC_DEV void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p, q;\n\t"
      ".reg .u32 n;\n\t"
      "mov.u32 n, 0;\n\t"
      "MBAR_WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"
      "@p bra MBAR_WAIT_DONE;\n\t"
      "add.u32 n, n, 1;\n\t"
      "setp.lt.u32 q, n, 1048576;\n\t"
      "@q bra MBAR_WAIT_LOOP;\n\t"
      "trap;\n\t"
      "MBAR_WAIT_DONE:\n\t"
      "}"
      ::"r"(bar), "r"(parity), "r"(4000u)
      : "memory");
}

// arrive and announce `bytes` of asynchronous copies that will complete on this barrier
TC_DEV void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// cp.async.bulk global -> shared (1-D TMA, UBLKCP): `bytes` (multiple of 16) from 16-byte aligned `src`; the copy
// engine signals the mbarrier with the byte count when the data has landed, no thread waits on the load itself
TC_DEV void bulk_g2s(void* dst_smem, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst_smem)), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

// bar.sync id, nthreads: a barrier among a subset of the CTA's warps (nthreads a multiple of 32, id 1..15)
TC_DEV void named_bar_sync(int id, int nthreads) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory"); }

// ---- fences ------------------------------------------------------------------------------
TC_DEV void tc_fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
TC_DEV void tc_fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// generic-proxy writes to shared memory -> visible to the async proxy (tcgen05.mma operand reads)
TC_DEV void fence_smem_to_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- TMEM --------------------------------------------------------------------------------
// One full warp allocates `cols` (power of two >= 32) columns; the base address lands in *slot.
TC_DEV void tmem_alloc(uint32_t slot_smem, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot_smem), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
TC_DEV void tmem_dealloc(uint32_t base, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(base), "r"(cols) : "memory");
}
// 32 lanes x 32 bit, 8 / 16 consecutive columns per thread; warp w reads lanes 32 * (w % 4) ..
TC_DEV void tmem_ld8(uint32_t addr, float (&v)[8]) {
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(addr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}
TC_DEV void tmem_ld8_nowait(uint32_t addr, uint32_t (&r)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(addr));
}
TC_DEV void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
TC_DEV uint32_t tmem_addr(uint32_t base, int lane, int col) { return base + ((uint32_t)lane << 16) + (uint32_t)col; }
// A value every lane of the (converged, full) warp holds, handed back through the uniform datapath (REDUX writes a
// uniform register).  tcgen05.mma takes its TMEM address from a uniform register; for an address the compiler cannot
// prove uniform (the TMEM base is read from shared memory) it wraps EVERY MMA in an ELECT / R2UR.BROADCAST / BRA.U.ANY
// loop, ~12 dependent uniform-datapath instructions of ~10 cycles each: the MMA-issuing thread of k_l2_out spent 130
// cycles per MMA on it and was the pipeline's bottleneck (ncu: 76 % of its samples in the issue code).
TC_DEV uint32_t warp_uniform(uint32_t v) { return __reduce_or_sync(0xffffffffu, v); }

// ---- descriptors -------------------------------------------------------------------------
// Shared-memory matrix descriptor, 128-byte swizzle (layout type 2), version 1 (sm_100).
TC_DEV uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout_type = 2) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)layout_type << 61;  // 0 none, 1 128B with 32B atoms, 2 128B, 4 64B, 6 32B
  return d;
}
// Instruction descriptor: TF32 x TF32 -> FP32, M x N, per-operand major-ness (0 = K, 1 = MN).
TC_DEV uint32_t umma_idesc_tf32(int M, int N, int a_mn_major, int b_mn_major) {
  uint32_t d = 0;
  d |= 1u << 4;                         // D format: F32
  d |= 2u << 7;                         // A format: TF32
  d |= 2u << 10;                        // B format: TF32
  d |= (uint32_t)(a_mn_major & 1) << 15;
  d |= (uint32_t)(b_mn_major & 1) << 16;
  d |= (uint32_t)(N >> 3) << 17;
  d |= (uint32_t)(M >> 4) << 24;
  return d;
}
// D[tmem] (+)= A[smem] * B[smem]; issued by ONE thread.
TC_DEV void umma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// All MMAs issued so far by this thread arrive on the mbarrier when they complete
// (implies tcgen05.fence::before_thread_sync).
TC_DEV void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

// The same two instructions for a CONVERGED warp: every lane runs the issue loop (waits, descriptor arithmetic in the
// uniform datapath) and one elected lane executes the tcgen05 instruction.  Inside an `if (lane == 0)` region the
// compiler cannot know that a single lane is active and wraps every MMA in an ELECT / BRA.U.ANY loop.
TC_DEV void umma_tf32_e(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, e;\n\t"
      "elect.sync _|e, 0xffffffff;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "@e tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
TC_DEV void umma_commit_e(uint32_t bar) {
  asm volatile(
      "{\n\t.reg .pred e;\n\t"
      "elect.sync _|e, 0xffffffff;\n\t"
      "@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(bar) : "memory");
}

#endif  // MSGAT_EMU
