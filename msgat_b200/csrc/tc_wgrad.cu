// tc_wgrad.cu -- every weight gradient of a MEAM as tcgen05 (kind::tf32) GEMMs whose
// contraction runs over the positions (node, t):
//   dWres[o,c]   = sum_p G[o,p]   X[c,p]          residual 1x1 conv            (msgat.py:116,131)
//   dMc[o,c]     = sum_p G_c[o,p] H[c,p]          folded channel attention     (attention.py:94, msgat.py:94), per sample
//   dW1now/past  = sum_p dA_now/past[k,p] H[c,p]  first conv behind the temporal attention (msgat.py:71-72, attention.py:66)
//   dWs[o,c]     = sum_p GZ[o,p]  H[c,p]          GACN.W commuted in front of the aggregation (msgat.py:27)
//   dW2now/past  = sum_p G_t[o,p] Y1[k,p] / Y1[k,p-d2]   second conv
// with G the masked upstream gradient [cacn | tacn | gacn], H = LayerNorm_T(X), Y1 the first
// conv's output (saved by the forward), GZ the transposed sparse aggregation of G_s, and
// (both computed and written to HBM by k_bwd_main, which has them in shared memory anyway)
//   dY1[k,t] = sum_o W2now[o,k] G_t[o,t] + W2past[o,k] G_t[o,t+d2],
//   dA_now[k,i] = sum_t Pt[t,i] dY1[k,t],  dA_past[k,i] = sum_{t>=d1} Pt[t-d1,i] dY1[k,t].
//
// Error-compensated TF32 ("3xTF32"): every operand v is staged twice, hi = rna_tf32(v) and
// lo = v - hi (exact in fp32), and each product is issued as hi*hi + hi*lo + lo*hi with fp32
// accumulation in TMEM; the dropped lo*lo term is 2^-24 relative.  That gives fp32-grade
// gradients from the tensor cores, so the kernel serves the fp32 path too.
//
// Positions are contiguous in HBM, so every operand is K-major.  A tile is 2 nodes = 24
// positions in the unswizzled core-matrix layout (8 rows x 4 positions = 128 contiguous bytes,
// tc_common.cuh:core_off), double-buffered so that staging tile i+1 (global loads, LayerNorm,
// hi/lo split) overlaps the MMAs of tile i; accumulators stay in TMEM across all tiles of the CTA and are
// flushed once as per-CTA partials (k_reduce sums them in a fixed order).
#include <stdlib.h>
#include "kernels.h"
#include "tc_common.cuh"
#include "tile_ops.cuh"

#define TCW_NT 2
#define TCW_P (TCW_NT * 12)
#define TCW_P4 (TCW_P / 4)
#define TCW_GS (TCW_P * 32)  // bytes of one 8-row group of a tile
#define TCW_THREADS 512      // staging (producer) threads; warp 16 issues the MMAs
#define TCW_ALL (TCW_THREADS + 32)
#define TCW_STAGES 2         // operand tiles are double-buffered: staging tile i+1 overlaps the MMAs of tile i
#define TCW_SLACK (16 * TCW_GS)  // an M=128 operand descriptor reads 16 row groups from its base

struct TcWgradArgs {   // grouped: blockIdx.y = component; inputs x_sm / g_sm apart, saved slices sv_sm, scratch slices sc_sm apart
  const float* x; int64_t x_sb, x_sc, x_sm, g_sm, sv_sm, sc_sm;
  const float* G;   // masked dout [B,O,N,12]
  int C, Cp, N, O, b, ntile, J, B, d2, has_ln;
  const float *ln_w[MSGAT_MAX_GROUP], *ln_b[MSGAT_MAX_GROUP];
  const float *Y1, *tcA;
  float *p_res, *p_ws, *p_c1past, *p_c1now, *p_c2past, *p_c2now;  // partial bases (block 0), row-major
  int64_t wg_stride;                                             // floats between consecutive blocks
  float* p_mc; int64_t mc_stride;
};

static inline int rup8(int v) { return (v + 7) & ~7; }

static inline size_t tcw_tiles_bytes(int C, int O, int b) {
  const int Cp = rup8(C);
  return (size_t)(rup8(O) + 2 * Cp + 3 * b + 2 * b) / 8 * TCW_GS;  // one precision part (hi or lo)
}

size_t tc_wgrad_smem(int C, int O, int b) {
  size_t s = TCW_STAGES * 2 * tcw_tiles_bytes(C, O, b) + TCW_SLACK;
  s += (24 + 12) * 4;  // ln, four mbarriers + TMEM slot
  return s + 1024;
}

// round to TF32, nearest with ties away from zero: the add-and-mask cvt.rna.tf32.f32 is lowered to, without its Inf / NaN guard
// (two integer instructions instead of four; bit-identical for finite values)
TC_DEV float tf32_rna(float x) { return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u); }
// v -> hi tile and lo tile at the same offset
TC_DEV void st_split(unsigned char* hi, unsigned char* lo, uint32_t off, float4 v) {
  float4 h = make_float4(tf32_rna(v.x), tf32_rna(v.y), tf32_rna(v.z), tf32_rna(v.w));
  st4(reinterpret_cast<float*>(hi + off), h);
  // lo is rounded (not truncated by the tensor core) to TF32: residual 2^-23 |v|, unbiased
  st4(reinterpret_cast<float*>(lo + off), make_float4(tf32_rna(v.x - h.x), tf32_rna(v.y - h.y), tf32_rna(v.z - h.z), tf32_rna(v.w - h.w)));
}

__global__ void __launch_bounds__(TCW_ALL, 1) k_tc_wgrad(TcWgradArgs a) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // aligned inside the shared window: the pointer stays a shared-memory pointer (LDS / STS, not generic LD / ST)
  const int C = a.C, Cp = a.Cp, O = a.O, b = a.b, tid = threadIdx.x, warp = tid >> 5, m = blockIdx.y;
  const int64_t svo = (int64_t)m * a.sv_sm, sco = (int64_t)m * a.sc_sm;
  const int Op = (O + 7) & ~7;
  const uint32_t GS = TCW_GS;
  const uint32_t szG = (uint32_t)(Op / 8) * GS, szX = (uint32_t)(2 * Cp / 8) * GS, szA = (uint32_t)(3 * b / 8) * GS,
                 szY = (uint32_t)(2 * b / 8) * GS;
  const uint32_t part = szG + szX + szA + szY;  // hi part, then the lo part in the same order
  // stage s at smem + s * 2 * part: [O][P] G, [2Cp][P] X rows 0..Cp | H rows Cp..2Cp,
  // [3b][P] dA_now | dA_past | GZ, [2b][P] Y1 | shift_d2(Y1); hi part then lo part
  float* LW = reinterpret_cast<float*>(smem + TCW_STAGES * 2 * part + TCW_SLACK);
  uint64_t* bars = reinterpret_cast<uint64_t*>(LW + 24);   // stage drained: the MMAs that read it retired
  uint64_t* full = reinterpret_cast<uint64_t*>(LW + 28);   // stage staged: one arrival per producer warp
  uint32_t& tmem_slot = *reinterpret_cast<uint32_t*>(LW + 32);

  // zero all operand tiles once: pad rows (C..Cp) and the slack must be finite
  for (int i = tid; i < (int)((TCW_STAGES * 2 * part + TCW_SLACK) / 16); i += TCW_ALL)
    reinterpret_cast<float4*>(smem)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (tid < 12) { LW[tid] = a.has_ln ? a.ln_w[m][tid] : 1.f; LW[12 + tid] = a.has_ln ? a.ln_b[m][tid] : 0.f; }
  if (tid == 0) {
    mbar_init(smem_u32(&bars[0]), 1); mbar_init(smem_u32(&bars[1]), 1);
    mbar_init(smem_u32(&full[0]), TCW_THREADS / 32); mbar_init(smem_u32(&full[1]), TCW_THREADS / 32);
    mbar_fence_init();
  }
  if (warp == 0) tmem_alloc(smem_u32(&tmem_slot), 512);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tm = warp_uniform(tmem_slot);   // uniform register: no per-MMA R2UR loop in the issue thread (tc_common.cuh)
  const int N1 = (2 * Cp + 15) & ~15, N2 = (Cp + 15) & ~15, N3 = 2 * b;
  const uint32_t D1 = tm, D2 = tm + 160, D3 = tm + 256;
  const uint32_t id1 = umma_idesc_tf32(128, N1, 0, 0), id2 = umma_idesc_tf32(128, N2, 0, 0), id3 = umma_idesc_tf32(128, N3, 0, 0);
  // A CTA serves one component and walks the logical blocks (sample bi, slot j), lb = bi * J + j = blockIdx.x, + gridDim.x, ...:
  // block (bi, j) owns the tiles j, j + J, ... of its sample, accumulates them in TMEM and flushes its own slice of partials,
  // exactly as when every logical block was a CTA.  The stage / barrier sequence (itg) runs on across the blocks.
  const int nlb = a.J * a.B;

  if (warp == TCW_THREADS / 32) {
    // ---- MMA warp: one thread waits for a staged tile, issues its 27 MMAs and commits them to the
    // stage's "drained" barrier.  Nothing else waits on this thread, so the producers run ahead.
    if ((tid & 31) == 0) {
      const uint32_t s0 = smem_u32(smem);
      const uint32_t offH = (uint32_t)(Cp / 8) * GS, offT = (uint32_t)(b / 8) * GS;
      // K-major, no swizzle: LBO = next 4 positions (128 B), SBO = next 8 rows (GS); stage 0, k = 0
      const uint64_t gh0 = umma_desc(s0, 128u, GS, 0), gl0 = umma_desc(s0 + part, 128u, GS, 0);
      const uint64_t xh0 = umma_desc(s0 + szG, 128u, GS, 0), xl0 = umma_desc(s0 + part + szG, 128u, GS, 0);
      const uint64_t ah0 = umma_desc(s0 + szG + szX, 128u, GS, 0), al0 = umma_desc(s0 + part + szG + szX, 128u, GS, 0);
      const uint64_t hh0 = umma_desc(s0 + szG + offH, 128u, GS, 0), hl0 = umma_desc(s0 + part + szG + offH, 128u, GS, 0);
      const uint64_t th0 = umma_desc(s0 + offT, 128u, GS, 0), tl0 = umma_desc(s0 + part + offT, 128u, GS, 0);
      const uint64_t yh0 = umma_desc(s0 + szG + szX + szA, 128u, GS, 0), yl0 = umma_desc(s0 + part + szG + szX + szA, 128u, GS, 0);
      int itg = 0;
      for (int lb = blockIdx.x; lb < nlb; lb += gridDim.x) {
      const int ntl = (a.ntile - (lb % a.J) + a.J - 1) / a.J;
      for (int it = 0; it < ntl; ++it, ++itg) {
        const int st = itg & 1;
        mbar_wait(smem_u32(&full[st]), (uint32_t)((itg >> 1) & 1));
        tc_fence_after_sync();
        for (int k8 = 0; k8 < TCW_P / 8; ++k8) {
          const uint64_t d = (uint64_t)(((uint32_t)st * 2u * part + (uint32_t)k8 * 256u) >> 4);   // start-address field, 16-byte units
          const uint32_t acc = (it == 0 && k8 == 0) ? 0u : 1u;
          umma_tf32(D1, gh0 + d, xh0 + d, id1, acc); umma_tf32(D1, gh0 + d, xl0 + d, id1, 1u); umma_tf32(D1, gl0 + d, xh0 + d, id1, 1u);
          umma_tf32(D2, ah0 + d, hh0 + d, id2, acc); umma_tf32(D2, ah0 + d, hl0 + d, id2, 1u); umma_tf32(D2, al0 + d, hh0 + d, id2, 1u);
          umma_tf32(D3, th0 + d, yh0 + d, id3, acc); umma_tf32(D3, th0 + d, yl0 + d, id3, 1u); umma_tf32(D3, tl0 + d, yh0 + d, id3, 1u);
        }
        umma_commit(smem_u32(&bars[st]));
      }
      }
    }
  }
  int it = 0;   // tiles staged so far by this CTA (all logical blocks)
  if (warp < TCW_THREADS / 32)
  for (int lb = blockIdx.x; lb < nlb; lb += gridDim.x) {
  const int bi = lb / a.J, jslot = lb - bi * a.J;
  const int ntl = (a.ntile - jslot + a.J - 1) / a.J;   // tiles of this logical block (>= 1)
  const float* xb = a.x + (int64_t)m * a.x_sm + bi * a.x_sb;
  const float* Gb = a.G + (int64_t)m * a.g_sm + (int64_t)bi * O * a.N * 12;
  const float* Yb = a.Y1 + svo + (int64_t)bi * b * a.N * 12;
  const float* Ab = a.tcA + sco + (int64_t)bi * 3 * b * a.N * 12;
  {

  // Work assignment, fixed for the whole kernel, every item one float4 (4 positions of one row):
  //  * G and dA rows: flat items on warps 8..15 (row-within-group in the low 3 bits, so 8
  //    consecutive threads fill one 128-byte core matrix and a warp four consecutive ones);
  //  * X (+LayerNorm) and Y1 (+causal shift) rows: warp items of (8 rows) x (one node), lane =
  //    row + 8 * quad with 24 lanes live; the LayerNorm statistics and the shifted taps move
  //    between the three quads of a row by shuffles.  Items 0..15 go to warps 0..15, items
  //    16..23 to warps 0..7, which evens out the instruction count per warp (the slowest warp
  //    sets the time of a tile: everything meets at the barrier below).
  const int lane = tid & 31;
  const int ql = lane >> 3, rl = lane & 7;
  uint32_t g_off[2], w_off[2], w_off2[2];
  int g_go[2], g_nl[2], w_nl[2], w_kind[2];
  int64_t w_go[2];
  bool g_on[2], a_on[2];
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const int gi = tid - 256 + 256 * k;
    const int r = (gi & 7) + 8 * (gi / (8 * TCW_P4)), p4 = ((gi >> 3) % TCW_P4) * 4;
    g_on[k] = tid >= 256 && gi < Op * TCW_P4 && r < O;
    a_on[k] = tid >= 256 && gi < 3 * b * TCW_P4;
    g_off[k] = core_off(r, p4, GS); g_go[k] = r * a.N * 12 + p4; g_nl[k] = p4 / 12;
    if (tid >= 256 && gi < Op * TCW_P4 && r >= O) g_on[k] = false;   // pad rows O..Op stay zero
    const int nX = (Cp / 8) * TCW_NT, nY = (b / 8) * TCW_NT;
    const int j = warp + 16 * k;
    w_kind[k] = 0; w_off[k] = w_off2[k] = 0; w_go[k] = 0; w_nl[k] = 0;
    if ((k == 0 || warp < 8) && ql < 3) {
      if (j < nX) {
        const int c = (j / TCW_NT) * 8 + rl, nl = j % TCW_NT;
        if (c < C) {
          w_kind[k] = 1; w_nl[k] = nl; w_go[k] = (int64_t)c * a.x_sc + nl * 12 + 4 * ql;
          w_off[k] = szG + core_off(c, nl * 12 + 4 * ql, GS); w_off2[k] = szG + core_off(Cp + c, nl * 12 + 4 * ql, GS);
        }
      } else if (j < nX + nY) {
        const int o = ((j - nX) / TCW_NT) * 8 + rl, nl = (j - nX) % TCW_NT;
        w_kind[k] = 2; w_nl[k] = nl; w_go[k] = ((int64_t)o * a.N + nl) * 12 + 4 * ql;
        w_off[k] = szG + szX + szA + core_off(o, nl * 12 + 4 * ql, GS);
        w_off2[k] = szG + szX + szA + core_off(b + o, nl * 12 + 4 * ql, GS);
      }
    }
  }
  const float4 lw4 = a.has_ln && ql < 3 ? ld4(a.ln_w[m] + 4 * ql) : make_float4(1.f, 1.f, 1.f, 1.f);
  const float4 lb4 = a.has_ln && ql < 3 ? ld4(a.ln_b[m] + 4 * ql) : f4zero();
  const int qd = a.d2 >> 2, rd = a.d2 & 3;   // causal shift = qd whole quads + rd positions
  // The global loads of tile i+1 are issued before tile i is staged (register prefetch): their
  // latency hides behind the staging arithmetic and the wait for the stage to drain.
  float4 pg[2], pa[2], pw[2];
  auto prefetch = [&](int tile) {
    const int n0 = tile * TCW_NT;
    const bool in = tile < a.ntile;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      pg[k] = pa[k] = pw[k] = f4zero();
      const bool gin = in && n0 + g_nl[k] < a.N;
      if (g_on[k] && gin) pg[k] = ld4(Gb + g_go[k] + (int64_t)n0 * 12);
      if (a_on[k] && gin) pa[k] = ld4(Ab + g_go[k] + (int64_t)n0 * 12);
      if (w_kind[k] && in && n0 + w_nl[k] < a.N) pw[k] = ld4((w_kind[k] == 1 ? xb : Yb) + w_go[k] + (int64_t)n0 * 12);
    }
  };
  prefetch(jslot);
  for (int tile = jslot; tile < a.ntile; tile += a.J, ++it) {
    const int n0 = tile * TCW_NT, st = it & 1;
    unsigned char* SB = smem + (size_t)st * 2 * part;   // stage base: G | X,H | dA,GZ | Y1,shift
    unsigned char* TG = SB;
    unsigned char* TX = TG + szG;
    unsigned char* TA = TX + szX;
    unsigned char* TY = TA + szA;
    const float4 cg[2] = {pg[0], pg[1]}, ca[2] = {pa[0], pa[1]}, cw[2] = {pw[0], pw[1]};
    prefetch(tile + a.J);
    // the MMAs that read this stage two tiles ago must have retired before it is overwritten
    if (it >= TCW_STAGES) mbar_wait(smem_u32(&bars[st]), (uint32_t)(((it >> 1) - 1) & 1));
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      // ---- X rows and their LayerNorm (msgat.py:122); Y1 rows and their causal shift.
      // Whole warps take this branch or skip it (the item index depends on the warp only).
      if ((k == 0 || warp < 8) && warp + 16 * k < (Cp / 8 + b / 8) * TCW_NT) {
        const bool isx = warp + 16 * k < (Cp / 8) * TCW_NT;
        const float4 v = cw[k];
        if (w_kind[k]) st_split(SB, SB + part, w_off[k], v);
        float4 o2;
        if (isx) {
          float sum = v.x + v.y + v.z + v.w;
          sum += __shfl_xor_sync(0xffffffffu, sum, 8); sum += __shfl_xor_sync(0xffffffffu, sum, 16);
          const float mu = sum * (1.0f / 12.0f);
          float4 d = make_float4(v.x - mu, v.y - mu, v.z - mu, v.w - mu);
          float ss = ql < 3 ? dot4(d, d) : 0.f;
          ss += __shfl_xor_sync(0xffffffffu, ss, 8); ss += __shfl_xor_sync(0xffffffffu, ss, 16);
          const float rs = rsqrt_ln(ss * (1.0f / 12.0f) + LN_EPS);
          o2 = v;
          if (a.has_ln && n0 + w_nl[k] < a.N)
            o2 = make_float4(d.x * rs * lw4.x + lb4.x, d.y * rs * lw4.y + lb4.y, d.z * rs * lw4.z + lb4.z, d.w * rs * lw4.w + lb4.w);
        } else {
          // shifted quad = y[4 ql + e - d2]: from quad (ql - qd) and, if rd > 0, the one before it
          const int sb_ = lane - 8 * qd, sa_ = sb_ - 8;
          float4 B, A;
          B.x = __shfl_sync(0xffffffffu, v.x, sb_ & 31); B.y = __shfl_sync(0xffffffffu, v.y, sb_ & 31);
          B.z = __shfl_sync(0xffffffffu, v.z, sb_ & 31); B.w = __shfl_sync(0xffffffffu, v.w, sb_ & 31);
          A.x = __shfl_sync(0xffffffffu, v.x, sa_ & 31); A.y = __shfl_sync(0xffffffffu, v.y, sa_ & 31);
          A.z = __shfl_sync(0xffffffffu, v.z, sa_ & 31); A.w = __shfl_sync(0xffffffffu, v.w, sa_ & 31);
          if (sb_ < 0) B = f4zero();
          if (sa_ < 0) A = f4zero();
          o2 = rd == 0 ? B : rd == 1 ? make_float4(A.w, B.x, B.y, B.z) : rd == 2 ? make_float4(A.z, A.w, B.x, B.y)
                                                                             : make_float4(A.y, A.z, A.w, B.x);
        }
        if (w_kind[k]) st_split(SB, SB + part, w_off2[k], o2);
      }
      // ---- G rows; dA_now | dA_past | GZ rows, written by k_bwd_main (meam_bwd.cu) for this kernel
      if (g_on[k]) st_split(TG, TG + part, g_off[k], cg[k]);
      if (a_on[k]) st_split(TA, TA + part, g_off[k], ca[k]);
    }
    fence_smem_to_async();
    __syncwarp();
    if (lane == 0) mbar_arrive(smem_u32(&full[st]));
  }
  }
  // the last commit of the block covers every MMA issued before it
  mbar_wait(smem_u32(&bars[(it - 1) & 1]), (uint32_t)(((it - 1) >> 1) & 1));
  tc_fence_after_sync();
  // ---- flush: TMEM lane = output row, columns = input channel
  {
    const int blk = lb;
    const int row = (warp & 3) * 32 + (tid & 31);
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    const int quarter = warp >> 2;  // the four warpgroups split the column groups
    float* wres = a.p_res + sco + (int64_t)blk * a.wg_stride;
    float* wmc = a.p_mc + sco + (int64_t)blk * a.mc_stride;
    int cg = 0;
    for (int c0 = 0; c0 < N1; c0 += 8, ++cg) {
      float v[8];
      if ((cg & 3) != quarter) continue;
      tmem_ld8(D1 + lane_base + c0, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int c = c0 + j;
        if (c < C && row < O) wres[row * C + c] = v[j];
        if (c >= Cp && c < Cp + C && row < b) wmc[row * C + c - Cp] = v[j];
      }
    }
    float* w1n = a.p_c1now + sco + (int64_t)blk * a.wg_stride;
    float* w1p = a.p_c1past + sco + (int64_t)blk * a.wg_stride;
    float* wws = a.p_ws + sco + (int64_t)blk * a.wg_stride;
    cg = 0;
    for (int c0 = 0; c0 < N2; c0 += 8, ++cg) {
      float v[8];
      if ((cg & 3) != quarter) continue;
      tmem_ld8(D2 + lane_base + c0, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int c = c0 + j;
        if (c >= C) continue;
        if (row < b) w1n[row * C + c] = v[j];
        else if (row < 2 * b) w1p[(row - b) * C + c] = v[j];
        else if (row < 3 * b) wws[(row - 2 * b) * C + c] = v[j];
      }
    }
    float* w2n = a.p_c2now + sco + (int64_t)blk * a.wg_stride;
    float* w2p = a.p_c2past + sco + (int64_t)blk * a.wg_stride;
    cg = 0;
    for (int c0 = 0; c0 < N3; c0 += 8, ++cg) {
      float v[8];
      if ((cg & 3) != quarter) continue;
      tmem_ld8(D3 + lane_base + c0, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int c = c0 + j;
        if (row >= b) continue;
        if (c < b) w2n[row * b + c] = v[j];
        else w2p[row * b + c - b] = v[j];
      }
    }
  }
  tc_fence_before_sync();   // the flush's TMEM reads are ordered before this warp's next arrival on a `full` barrier, after
                            // which the MMA thread overwrites the accumulators with the next block's first product
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tm, 512);
}

bool tc_wgrad_supported(const Plan& p) {
  const char* off = getenv("MSGAT_TC_DISABLE");  // test switch: bit 1 = weight-gradient kernel
  if (off && (atoi(off) & 2)) return false;
  const int full = MSGAT_F_RES | MSGAT_F_CMIX | MSGAT_F_SMIX;
  const int b = p.Oc;
  return (p.flags & full) == full && p.n_conv == 2 && p.Oc == p.Ot && p.Ot == p.Os &&
         b % 8 == 0 && b >= 8 && b <= 24 && p.C >= 1 && p.C <= 72 && p.O <= 72 &&
         p.dil[0] >= 1 && p.dil[0] < 12 && p.dil[1] >= 1 && p.dil[1] < 12;
}

int tc_wgrad_blocks_per_sample(const Plan& p) {
  const int ntile = (p.N + TCW_NT - 1) / TCW_NT;
  return tc_ctas_per_sample(p.B, ntile, 2, "MSGAT_TW_WAVES");
}

int launch_tc_wgrad(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, const float* x, const float* G, int64_t g_sm,
                    const float* saved, const float* tcA, float* p_res, float* p_ws, float* p_c1past, float* p_c1now, float* p_c2past,
                    float* p_c2now, int64_t wg_stride, float* p_mc, int64_t mc_stride, void* stream) {
  TcWgradArgs a;
  a.x = x; a.x_sb = p.x_sb; a.x_sc = p.x_sc; a.x_sm = p.x_sm; a.g_sm = g_sm; a.sv_sm = p.saved_total; a.sc_sm = p.scratch_sm; a.G = G;
  a.C = p.C; a.Cp = rup8(p.C); a.N = p.N; a.O = p.O; a.b = p.Oc; a.ntile = (p.N + TCW_NT - 1) / TCW_NT;
  a.J = tc_wgrad_blocks_per_sample(p); a.B = p.B;
  a.d2 = p.dil[1]; a.has_ln = (p.flags & MSGAT_F_LN) ? 1 : 0;
  for (int m = 0; m < MSGAT_MAX_GROUP; ++m) { a.ln_w[m] = prm[m < p.M ? m : 0].ln_w; a.ln_b[m] = prm[m < p.M ? m : 0].ln_b; }
  a.Y1 = saved + p.o_Y1; a.tcA = tcA;
  a.p_res = p_res; a.p_ws = p_ws; a.p_c1past = p_c1past; a.p_c1now = p_c1now; a.p_c2past = p_c2past; a.p_c2now = p_c2now;
  a.wg_stride = wg_stride; a.p_mc = p_mc; a.mc_stride = mc_stride;
  // one CTA per SM, every CTA on ONE component: sm_count / M CTAs per component walk its J * B logical blocks
  int G1 = msgat_sm_count() / p.M;
  if (G1 < 1) G1 = 1;
  if (G1 > a.J * p.B) G1 = a.J * p.B;
  MSGAT_LAUNCH(k_tc_wgrad, dim3(G1, p.M), dim3(TCW_ALL), tc_wgrad_smem(p.C, p.O, p.Oc), stream, a);
  return 0;
}
