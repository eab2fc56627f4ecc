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
// Spins on the phase parity; a wait that can never complete traps instead of hanging the device.
TC_DEV void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok = 0, spins = 0;
  while (true) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    if (ok) break;
    if (++spins > (1u << 26)) __trap();
  }
}

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
