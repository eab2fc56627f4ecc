// l2_common.cuh -- the warp-specialised, mbarrier-pipelined skeleton shared by the layer-2 (C % 8 == 0) MEAM kernels.
//
// Every channel mix of a layer-2 MEAM is a GEMM over positions p = (node, t) with a contraction over <= 96
// channels (msgat.py:27,71-72,94,116).  The kernels here run them on tcgen05 (kind::tf32, accumulators in TMEM)
// with three roles per CTA that only meet at mbarriers:
//   producer warp     cp.async.bulk (1-D TMA) copies of the raw fp32 rows of a K-chunk (one contiguous run of
//                     8 nodes x 48 bytes per channel) into a ring of raw chunks, completion counted on mbarriers:
//                     tens of KB are in flight per SM without a register or a waiting thread;
//   converter warps   raw rows of 12 floats -> (LayerNorm / temporal mix in registers) -> error-compensated TF32
//                     operands (hi, lo) in a second ring of shared-memory K-chunks;
//   one MMA thread    waits for a chunk, issues its tcgen05.mma group, commits the chunk back to the converters and,
//                     at the end of a tile, the accumulators to the epilogue (two accumulator sets alternate);
//   epilogue warps    TMEM -> registers -> bias / ReLU / sparse-neighbour terms -> coalesced global stores.
//
// Tile = 8 nodes on the 128 TMEM lanes: lane = 16 * node + t, t < 12 live (a warp quarter = 2 nodes), so that the
// rows of 12 of a node never straddle a warp.  Operands are MN-major (positions contiguous): the 128-byte swizzle
// with 32-byte atoms, the one layout tcgen05 accepts for 32-bit MN-major operands (measured, tc_probe.cu).
//
// Precision.  fp32-grade mode splits every operand v into hi = tf32(v), lo = tf32(v - hi) and issues
// hi*hi + hi*lo + lo*hi.  The tensor core TRUNCATES when it adds an MMA's products into the fp32
// accumulator, so the error grows with the number of accumulations at full magnitude: hi*hi goes to a MAIN
// accumulator (one rounding per k-step), the three correction terms (2^-11 of it) to a second, SMALL one, and the
// epilogue adds the two in fp32 (measured on the B200: forward rel-L2 4.6e-7 -> 2.4e-7, dX 5.3e-6 -> 1.8e-6).
// TF32 mode issues hi*hi only (2e-3 bar).
#pragma once
#include "kernels.h"
#include "tc_common.cuh"
#include "tile_ops.cuh"

#define L2_NT 8            // nodes per tile
#define L2_EPI_WARPS 8     // warps 0..7: TMEM lane quarter warp % 4, column groups of parity warp / 4
#define L2_MMA_WARP 8
#define L2_CONV_WARP0 9
#define L2_CONV_WARPS 6    // 192 converter threads = 24 rows x 8 nodes per K-chunk
#define L2_CONV_GROUPS 2   // converter groups: group g converts the chunks q with q % 2 == g.  A converter warp is a chain
                           // of dependent steps per chunk (wait, load, LayerNorm, split, store, fence, arrive: ~2 k cycles
                           // measured); two groups in flight double the chunk rate
#define L2_PROD_WARP (L2_CONV_WARP0 + L2_CONV_WARPS * L2_CONV_GROUPS)   // one more warp: issues the bulk copies (1-D TMA) of the raw rows
#define L2_PROD_WARPS 2    // a bulk copy takes per-lane operands through the uniform datapath, so a warp issues its rows one
                           // after the other (R2UR loop in the SASS).  In k_l2_mix, whose converters are light, ONE producer
                           // warp issuing the 24 copies of a chunk was the longest stage (ncu: the producer never idle, every
                           // other role waiting): 477 us with one producer warp, 350 us with two or three.  k_l2_out is bound
                           // by its converters and gains nothing (measured 1.12 / 1.15 / 1.19 ms with 1 / 2 / 3); it keeps one
                           // producer warp and the higher register cap of a 704-thread block.
#define L2_THREADS_NP(np) ((L2_PROD_WARP + (np)) * 32)
#define L2_THREADS L2_THREADS_NP(L2_PROD_WARPS)
#define L2_BAR_BYTES 16    // an mbarrier is 8 bytes; the host-fiber emulation keeps 16

// round to TF32, nearest with ties away from zero in magnitude: two integer instructions (cvt.rna.tf32.f32 is lowered
// to the same add-and-mask plus an Inf / NaN guard the activations here never need)
MSGAT_DEV float l2_tf32(float x) { return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u); }
// v -> hi tile and lo tile at the same byte offset
MSGAT_DEV void l2_split4(unsigned char* hi, unsigned char* lo, uint32_t off, const float4& v) {
  const float4 h = make_float4(l2_tf32(v.x), l2_tf32(v.y), l2_tf32(v.z), l2_tf32(v.w));
  st4(reinterpret_cast<float*>(hi + off), h);
  st4(reinterpret_cast<float*>(lo + off), make_float4(l2_tf32(v.x - h.x), l2_tf32(v.y - h.y), l2_tf32(v.z - h.z), l2_tf32(v.w - h.w)));
}
MSGAT_DEV void l2_split1(unsigned char* hi, unsigned char* lo, uint32_t off, float v) {
  const float h = l2_tf32(v);
  *reinterpret_cast<float*>(hi + off) = h;
  *reinterpret_cast<float*>(lo + off) = l2_tf32(v - h);
}
// One k-step (8 contraction rows) of an error-compensated product: main (+)= Ah Bh, small (+)= Ah Bl + Al Bh.
// ka / kb: k-step index inside the A / B operand (8 rows = 1024 bytes = 64 units of the descriptor's address field).
MSGAT_DEV void l2_mma(uint32_t d_main, uint32_t d_small, uint64_t ah, uint64_t al, uint64_t bh, uint64_t bl, uint32_t ka, uint32_t kb,
                      uint32_t idesc, uint32_t acc_main, uint32_t acc_small, int fp32_grade) {
  const uint64_t oa = (uint64_t)(ka * 64u), ob = (uint64_t)(kb * 64u);
  umma_tf32(d_main, ah + oa, bh + ob, idesc, acc_main);
  if (fp32_grade) {
    // lo*lo (2^-22 of the product) is not issued: adding it changed no measured error (round-2 experiment), and every
    // MMA re-reads its 4 KB + N * 32 B of operands from shared memory, the pipe these kernels are bound by
    umma_tf32(d_small, ah + oa, bl + ob, idesc, acc_small);
    umma_tf32(d_small, al + oa, bh + ob, idesc, 1u);
  }
}
// 8 consecutive accumulator columns of this thread's lane: main (+ small)
MSGAT_DEV void l2_ld8(uint32_t addr_main, uint32_t addr_small, int fp32_grade, float (&v)[8]) {
  uint32_t r0[8], r1[8];
  tmem_ld8_nowait(addr_main, r0);
  if (fp32_grade) tmem_ld8_nowait(addr_small, r1);
  tmem_ld_wait();
  MSGAT_UNROLL
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r0[i]) + (fp32_grade ? __uint_as_float(r1[i]) : 0.f);
}
// three accumulators: two full-magnitude ones (kept apart so that neither sees more than one GEMM's chain of
// truncating accumulations) and the small one
MSGAT_DEV void l2_ld8x3(uint32_t addr_a, uint32_t addr_b, uint32_t addr_small, int fp32_grade, float (&v)[8]) {
  uint32_t r0[8], r1[8], r2[8];
  tmem_ld8_nowait(addr_a, r0);
  tmem_ld8_nowait(addr_b, r1);
  if (fp32_grade) tmem_ld8_nowait(addr_small, r2);
  tmem_ld_wait();
  MSGAT_UNROLL
  for (int i = 0; i < 8; ++i) v[i] = (__uint_as_float(r0[i]) + __uint_as_float(r1[i])) + (fp32_grade ? __uint_as_float(r2[i]) : 0.f);
}
// rows of a K-chunk: 24 when the channel count allows it (192 converter threads = 24 rows x 8 nodes), else 16 or 8
static inline int l2_chunk_rows(int C) { return C % 24 == 0 ? 24 : (C % 16 == 0 ? 16 : 8); }

// A converter thread's place in a chunk: a warp = 4 consecutive rows x 8 nodes (row in the low lane bits, so that the
// four lanes of a node hit the four swizzle positions of a 128-byte line)
struct L2Conv {
  int grp, cid, kl, nl;   // converter group, thread index inside it, row inside the chunk, node inside the tile
};
MSGAT_DEV L2Conv l2_conv_thread() {
  L2Conv c;
  c.cid = (int)threadIdx.x - L2_CONV_WARP0 * 32;
  c.grp = c.cid / (L2_CONV_WARPS * 32);
  c.cid -= c.grp * L2_CONV_WARPS * 32;
  c.kl = (c.cid & 3) + 4 * (c.cid >> 5);
  c.nl = (c.cid >> 2) & 7;
  return c;
}
// Byte offsets of the three quads of the row (k, nl) inside an operand chunk (positions 16 * nl + t): they depend on
// the thread only, so a converter computes them once.
struct L2RowOff { uint32_t o0, o4, o8; };
MSGAT_DEV L2RowOff l2_row_off(uint32_t cb, int k, int nl) {
  L2RowOff r;
  r.o0 = sw128b32_off(k, 16 * nl, cb); r.o4 = sw128b32_off(k, 16 * nl + 4, cb); r.o8 = sw128b32_off(k, 16 * nl + 8, cb);
  return r;
}
// store one row of 12 into an operand chunk (hi / lo)
MSGAT_DEV void l2_store_row(unsigned char* hi, unsigned char* lo, const L2RowOff& ro, bool odd_node, const float4& a, const float4& b, const float4& d) {
  // the two nodes of a quarter-warp hit the same four 16-byte bank groups with their first quad: odd nodes start
  // with the second quad (the other half of the 32-byte swizzle atom), which makes two of the three stores conflict-free
  if (odd_node) { l2_split4(hi, lo, ro.o4, b); l2_split4(hi, lo, ro.o0, a); }
  else { l2_split4(hi, lo, ro.o0, a); l2_split4(hi, lo, ro.o4, b); }
  l2_split4(hi, lo, ro.o8, d);
}
// Sum of 12 values over the 8 lanes that differ in lane bits 2..4 (the 8 nodes of a tile), recursive halving: 11
// shuffles instead of 36.  On return lane (bits 2,3,4 = h2,h3,h4) holds in r0 the total of element 6 h2 + 3 h3 + (h4 ? 2 : 0)
// and, when h4 == 0, in r1 the total of the element after it.
MSGAT_DEV void l2_reduce12_nodes(const float (&v)[12], int lane, float& r0, float& r1) {
  const bool h2 = lane & 4, h3 = lane & 8, h4 = lane & 16;
  float k6[6];
  MSGAT_UNROLL
  for (int i = 0; i < 6; ++i) {
    const float send = h2 ? v[i] : v[6 + i], keep = h2 ? v[6 + i] : v[i];
    k6[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
  }
  float k3[3];
  MSGAT_UNROLL
  for (int i = 0; i < 3; ++i) {
    const float send = h3 ? k6[i] : k6[3 + i], keep = h3 ? k6[3 + i] : k6[i];
    k3[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
  }
  // lanes with h4 == 0 keep elements 0 and 1, lanes with h4 == 1 keep element 2
  const float s0 = h4 ? k3[0] : k3[2], s1 = h4 ? k3[1] : 0.f;
  const float g0 = __shfl_xor_sync(0xffffffffu, s0, 16), g1 = __shfl_xor_sync(0xffffffffu, s1, 16);
  r0 = h4 ? k3[2] + g0 : k3[0] + g0;
  r1 = h4 ? 0.f : k3[1] + g1;
}
// LayerNorm of one row in three quads (nn.LayerNorm([12]), msgat.py:115,122); same operation order as tile_load_ln
MSGAT_DEV void l2_row_ln(float4& a, float4& b, float4& d, const float* LW) {
  const float mu = (a.x + a.y + a.z + a.w + b.x + b.y + b.z + b.w + d.x + d.y + d.z + d.w) * (1.0f / 12.0f);
  a.x -= mu; a.y -= mu; a.z -= mu; a.w -= mu; b.x -= mu; b.y -= mu; b.z -= mu; b.w -= mu;
  d.x -= mu; d.y -= mu; d.z -= mu; d.w -= mu;
  const float var = (dot4(a, a) + dot4(b, b) + dot4(d, d)) * (1.0f / 12.0f);
  const float rs = rsqrt_ln(var + LN_EPS);
  a.x = a.x * rs * LW[0] + LW[12]; a.y = a.y * rs * LW[1] + LW[13]; a.z = a.z * rs * LW[2] + LW[14]; a.w = a.w * rs * LW[3] + LW[15];
  b.x = b.x * rs * LW[4] + LW[16]; b.y = b.y * rs * LW[5] + LW[17]; b.z = b.z * rs * LW[6] + LW[18]; b.w = b.w * rs * LW[7] + LW[19];
  d.x = d.x * rs * LW[8] + LW[20]; d.y = d.y * rs * LW[9] + LW[21]; d.z = d.z * rs * LW[10] + LW[22]; d.w = d.w * rs * LW[11] + LW[23];
}
