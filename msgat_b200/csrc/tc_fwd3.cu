// tc_fwd3.cu -- the MEAM apply pass of tc_fwd.cu with error-compensated TF32 ("3xTF32"):
// every GEMM operand v is staged as hi = rna_tf32(v) and lo = v - hi and every product is
// issued as hi*hi + hi*lo + lo*hi into the same TMEM accumulator, which gives fp32-grade
// results from tcgen05 -- this kernel serves the default fp32 path (MSGAT_PREC_FP32) for the
// layer-2 MEAM shape.  Same math and same commuting of the temporal mix as k_tc_apply:
//   D1[p, 0:O]   = X^T  Wres^T                       (msgat.py:131)
//   D2[p, 0:b]   = H^T (W1x1 Pc)^T                   (attention.py:94, msgat.py:94)
//   D2[p, b:3b]  = H^T [Wnow1 ; Wpast1]^T            first-conv taps before the temporal mix
//   y1           = Pt a_now + Pt(t-d1) a_past + b1   rows of 12 on CUDA cores (attention.py:66, msgat.py:71-72)
//   D3[p, 0:b]   = [y1 ; shift_d2(y1)]^T [Wnow2 ; Wpast2]^T
//   out          = relu([D2[:, :b] + ca_b | D3 + b2 | sparse aggregation of Z] + D1 + res_b)
// The doubled operands halve the tile: 8 nodes = 96 positions = one M = 128 block; X and H
// go through the same pair of shared-memory tiles one after the other (x rows stay in
// registers in between).  224 TMEM columns, 200 KB shared memory.
#include <stdlib.h>
#include "kernels.h"
#include "tc_common.cuh"
#include "tile_ops.cuh"

#define TF3_NT 8
#define TF3_P 96
#define TF3_THREADS 512
#define TF3_PITCH 100

struct TcApply3Args {
  const float* x; int64_t x_sb, x_sc;
  int C, N, O, b, ntile, J, d1, d2, nnz;
  const float *ln_w, *ln_b, *res_w, *res_b, *ca_b, *c1_w, *c1_b, *c2_w, *c2_b;
  const float *Mc, *Pt, *Z, *pw;
  const int32_t *row_ptr, *col_idx;
  float *out, *Y1, *A2;
};

static size_t tc_apply3_smem(int C, int b) {
  size_t s = 0;
  s += 2 * (size_t)3 * C * 128;            // AH, AL (X then H; later y1 hi/lo)
  s += 4 * (size_t)3 * C * 128;            // WRH, WRL, WHH, WHL
  s += 2 * (size_t)2 * b * 128;            // W2H, W2L
  s += (size_t)2 * b * TF3_PITCH * 4;      // staging of a_now / a_past
  s += (144 * 2 + 96 + 32 * 3 + 24 + 4) * 4;
  return s + 1024;
}

// round to TF32, nearest with ties away from zero: the add-and-mask cvt.rna.tf32.f32 is lowered to, without its Inf / NaN guard
// (two integer instructions instead of four; bit-identical for finite values)
TC_DEV float tf3_rna(float x) { return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u); }
TC_DEV void tf3_split4(unsigned char* hi, unsigned char* lo, uint32_t off, float4 v) {
  float4 h = make_float4(tf3_rna(v.x), tf3_rna(v.y), tf3_rna(v.z), tf3_rna(v.w));
  st4(reinterpret_cast<float*>(hi + off), h);
  // lo is rounded (not truncated by the tensor core) to TF32: residual 2^-23 |v|, unbiased
  st4(reinterpret_cast<float*>(lo + off), make_float4(tf3_rna(v.x - h.x), tf3_rna(v.y - h.y), tf3_rna(v.z - h.z), tf3_rna(v.w - h.w)));
}
TC_DEV void tf3_split1(unsigned char* hi, unsigned char* lo, uint32_t off, float v) {
  const float h = tf3_rna(v);
  *reinterpret_cast<float*>(hi + off) = h;
  *reinterpret_cast<float*>(lo + off) = tf3_rna(v - h);
}
// acc (+)= Ah Bh + Ah Bl + Al Bh for one k-step; descriptors: MN-major, 128 B swizzle with 32 B atoms
// The descriptors of k-step 0 are built once; k-step ks (8 contraction rows = 1024 B) adds ks * 64 to the
// start-address field (16-byte units).
TC_DEV void mma3(uint32_t d, uint64_t dah, uint64_t dal, uint64_t dbh, uint64_t dbl, uint32_t ks, uint32_t idesc, uint32_t acc) {
  const uint64_t o = (uint64_t)(ks * 64u);
  // The tensor core adds each MMA's products into the fp32 accumulator with truncation, not round-to-nearest: the
  // error grows with the NUMBER of accumulations at full magnitude.  The hi*hi terms therefore go to one accumulator
  // (one rounding per k-step) and the three correction terms, 2^-11 of it, to a second one 256 columns further;
  // the epilogue adds the two in fp32.
  umma_tf32(d, dah + o, dbh + o, idesc, acc);
  umma_tf32(d + 256u, dah + o, dbl + o, idesc, acc);
  umma_tf32(d + 256u, dal + o, dbh + o, idesc, 1u);
  umma_tf32(d + 256u, dal + o, dbl + o, idesc, 1u);
}
TC_DEV void tmem_ld8_sum(uint32_t addr, float (&v)[8]) {
  uint32_t r0[8], r1[8];
  tmem_ld8_nowait(addr, r0);
  tmem_ld8_nowait(addr + 256u, r1);
  tmem_ld_wait();
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r0[i]) + __uint_as_float(r1[i]);
}

__global__ void __launch_bounds__(TF3_THREADS, 1) k_tc_apply3(TcApply3Args a) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // aligned inside the shared window: the pointer stays a shared-memory pointer (LDS / STS, not generic LD / ST)
  const int C = a.C, b = a.b, O = a.O, tid = threadIdx.x, warp = tid >> 5, bi = blockIdx.y;
  const uint32_t cbX = (uint32_t)C * 128u, cbY = (uint32_t)(2 * b) * 128u;
  unsigned char* AH = smem;
  unsigned char* AL = AH + 3 * cbX;
  unsigned char* WRH = AL + 3 * cbX;
  unsigned char* WRL = WRH + 3 * cbX;
  unsigned char* WHH = WRL + 3 * cbX;
  unsigned char* WHL = WHH + 3 * cbX;
  unsigned char* W2H = WHL + 3 * cbX;
  unsigned char* W2L = W2H + cbY;
  float* STG = reinterpret_cast<float*>(W2L + cbY);   // [2b][TF3_PITCH]
  float* Pt = STG + 2 * b * TF3_PITCH;
  float* Ptd = Pt + 144;
  float* resB = Ptd + 144;   // [96]
  float* caB = resB + 96;    // [32]
  float* c1B = caB + 32;
  float* c2B = c1B + 32;
  float* LW = c2B + 32;      // [24]
  uint64_t& bar = *reinterpret_cast<uint64_t*>(LW + 24);
  uint32_t& tmem_slot = *reinterpret_cast<uint32_t*>(LW + 26);
  unsigned char* YH = AH;             // [2b][96] y1 | shift(y1), aliases the A tiles once D2 is complete
  unsigned char* YL = AH + 3 * cbY;

  // ---- one-time staging: weights as B operands [K][N] (N contiguous = MN-major), hi / lo, zero padded
  for (int i = tid; i < C * 96; i += TF3_THREADS) {
    const int k = i / 96, n = i - k * 96;
    const uint32_t off = sw128b32_off(k, n, cbX);
    tf3_split1(WRH, WRL, off, n < O ? a.res_w[n * C + k] : 0.f);
    float w = 0.f;
    if (n < b) w = a.Mc[((int64_t)bi * b + n) * C + k];                 // (W1x1 Pc)[n][k] of this sample
    else if (n < 2 * b) w = a.c1_w[((n - b) * C + k) * 2 + 1];          // tap 1 multiplies x[t]
    else if (n < 3 * b) w = a.c1_w[((n - 2 * b) * C + k) * 2 + 0];      // tap 0 multiplies x[t - d]
    tf3_split1(WHH, WHL, off, w);
  }
  for (int i = tid; i < 2 * b * 32; i += TF3_THREADS) {
    const int k = i / 32, n = i - k * 32;
    float w = 0.f;
    if (n < b) w = k < b ? a.c2_w[(n * b + k) * 2 + 1] : a.c2_w[(n * b + (k - b)) * 2 + 0];
    tf3_split1(W2H, W2L, sw128b32_off(k, n, cbY), w);
  }
  for (int i = tid; i < 144; i += TF3_THREADS) {
    const int t = i / 12, j = i - t * 12;
    Pt[i] = a.Pt[(int64_t)bi * 144 + i];
    Ptd[i] = t >= a.d1 ? a.Pt[(int64_t)bi * 144 + (t - a.d1) * 12 + j] : 0.f;
  }
  for (int i = tid; i < 96; i += TF3_THREADS) resB[i] = i < O ? a.res_b[i] : 0.f;
  for (int i = tid; i < 32; i += TF3_THREADS) {
    caB[i] = i < b ? a.ca_b[i] : 0.f; c1B[i] = i < b ? a.c1_b[i] : 0.f; c2B[i] = i < b ? a.c2_b[i] : 0.f;
  }
  if (tid < 12) { LW[tid] = a.ln_w[tid]; LW[12 + tid] = a.ln_b[tid]; }
  if (tid == 0) { mbar_init(smem_u32(&bar), 1); mbar_fence_init(); }
  if (warp == 0) tmem_alloc(smem_u32(&tmem_slot), 512);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tm = tmem_slot;
  const uint32_t D1 = tm, D2 = tm + 96, D3 = tm + 192;
  const uint64_t dAH = umma_desc(smem_u32(AH), cbX, 512u, 1), dAL = umma_desc(smem_u32(AL), cbX, 512u, 1);
  const uint64_t dWRH = umma_desc(smem_u32(WRH), cbX, 512u, 1), dWRL = umma_desc(smem_u32(WRL), cbX, 512u, 1);
  const uint64_t dWHH = umma_desc(smem_u32(WHH), cbX, 512u, 1), dWHL = umma_desc(smem_u32(WHL), cbX, 512u, 1);
  const uint64_t dW2H = umma_desc(smem_u32(W2H), cbY, 512u, 1), dW2L = umma_desc(smem_u32(W2L), cbY, 512u, 1);
  const uint64_t dYH = umma_desc(smem_u32(YH), cbY, 512u, 1), dYL = umma_desc(smem_u32(YL), cbY, 512u, 1);
  const uint32_t idesc80 = umma_idesc_tf32(128, 80, 1, 1), idesc32 = umma_idesc_tf32(128, 32, 1, 1);
  uint32_t phase = 0;
  const int l = tid & 127, quarter = warp >> 2;    // l = TMEM lane = position; the 4 warpgroups split the column groups
  const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
  const float* xb = a.x + bi * a.x_sb;
  const float* pwb = a.pw + (int64_t)bi * a.nnz;
  const float* Zb = a.Z + (int64_t)bi * b * a.N * 12;

  // the one column group (if any) of this thread's quarter that holds spatial channels
  int sp_c0 = -1;
  { int g = 0; for (int c0 = 0; c0 < O; c0 += 8, ++g) if ((g & 3) == quarter && c0 >= 2 * b) sp_c0 = c0; }
  for (int tile = blockIdx.x; tile < a.ntile; tile += a.J) {
    const int n0 = tile * TF3_NT;
    // L2 prefetch of the next tile's x rows (8 nodes x 48 B = 384 B per channel), one 128-byte line per thread
    if (tile + a.J < a.ntile) {
      for (int i = tid; i < C * 4; i += TF3_THREADS) {
        int64_t off = (int64_t)(n0 + a.J * TF3_NT) * 12 + (i & 3) * 32;
        const int64_t last = (int64_t)a.N * 12 - 1;
        if (off > last) off = last;
        asm volatile("prefetch.global.L2 [%0];" ::"l"(xb + (i >> 2) * a.x_sc + off));
      }
    }
    // ---- 1. x rows -> registers and the A tiles (hi / lo)
    float4 xr[2][3];
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const int r = tid + q * TF3_THREADS;
      xr[q][0] = xr[q][1] = xr[q][2] = f4zero();
      if (r < C * TF3_NT) {
        const int c = r / TF3_NT, nl = r - c * TF3_NT, n = n0 + nl, pp = nl * 12;
        if (n < a.N) { const float* s = xb + c * a.x_sc + (int64_t)n * 12; xr[q][0] = ld4(s); xr[q][1] = ld4(s + 4); xr[q][2] = ld4(s + 8); }
        tf3_split4(AH, AL, sw128b32_off(c, pp, cbX), xr[q][0]);
        tf3_split4(AH, AL, sw128b32_off(c, pp + 4, cbX), xr[q][1]);
        tf3_split4(AH, AL, sw128b32_off(c, pp + 8, cbX), xr[q][2]);
      }
    }
    fence_smem_to_async();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after_sync();
      for (int ks = 0; ks < C / 8; ++ks) mma3(D1, dAH, dAL, dWRH, dWRL, ks, idesc80, ks > 0);
      umma_commit(smem_u32(&bar));
    }
    // ---- while the tensor core runs: sparse spatial aggregation of this thread's 8 channels
    // (attention.py:36 at the adjacency non-zeros); the values wait in registers for the epilogue
    float sp[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) sp[j] = 0.f;
    if (sp_c0 >= 0) {
      const int nl = l / 12, n = n0 + nl, t = l - nl * 12;
      if (l < TF3_P && n < a.N) {
        const int64_t cs = (int64_t)a.N * 12;
        const int e1 = a.row_ptr[n + 1];
        for (int e = a.row_ptr[n]; e < e1; ++e) {
          const float w = pwb[e];
          const float* z = Zb + ((int64_t)(sp_c0 - 2 * b) * a.N + a.col_idx[e]) * 12 + t;
#pragma unroll
          for (int j = 0; j < 8; ++j) sp[j] = fmaf(w, z[j * cs], sp[j]);
        }
      }
    }
    mbar_wait(smem_u32(&bar), phase); phase ^= 1;
    tc_fence_after_sync();
    // ---- 2. LayerNorm over t of the same rows (msgat.py:122) -> the A tiles again
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const int r = tid + q * TF3_THREADS;
      if (r < C * TF3_NT) {
        const int c = r / TF3_NT, nl = r - c * TF3_NT, n = n0 + nl, pp = nl * 12;
        float4 u = xr[q][0], v = xr[q][1], w = xr[q][2];
        if (n < a.N) {
          float mu = (u.x + u.y + u.z + u.w + v.x + v.y + v.z + v.w + w.x + w.y + w.z + w.w) * (1.0f / 12.0f);
          u.x -= mu; u.y -= mu; u.z -= mu; u.w -= mu; v.x -= mu; v.y -= mu; v.z -= mu; v.w -= mu;
          w.x -= mu; w.y -= mu; w.z -= mu; w.w -= mu;
          const float rs = rsqrt_ln((dot4(u, u) + dot4(v, v) + dot4(w, w)) * (1.0f / 12.0f) + LN_EPS);
          u.x = u.x * rs * LW[0] + LW[12]; u.y = u.y * rs * LW[1] + LW[13]; u.z = u.z * rs * LW[2] + LW[14]; u.w = u.w * rs * LW[3] + LW[15];
          v.x = v.x * rs * LW[4] + LW[16]; v.y = v.y * rs * LW[5] + LW[17]; v.z = v.z * rs * LW[6] + LW[18]; v.w = v.w * rs * LW[7] + LW[19];
          w.x = w.x * rs * LW[8] + LW[20]; w.y = w.y * rs * LW[9] + LW[21]; w.z = w.z * rs * LW[10] + LW[22]; w.w = w.w * rs * LW[11] + LW[23];
        }
        tf3_split4(AH, AL, sw128b32_off(c, pp, cbX), u);
        tf3_split4(AH, AL, sw128b32_off(c, pp + 4, cbX), v);
        tf3_split4(AH, AL, sw128b32_off(c, pp + 8, cbX), w);
      }
    }
    fence_smem_to_async();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after_sync();
      for (int ks = 0; ks < C / 8; ++ks) mma3(D2, dAH, dAL, dWHH, dWHL, ks, idesc80, ks > 0);
      umma_commit(smem_u32(&bar));
    }
    mbar_wait(smem_u32(&bar), phase); phase ^= 1;
    tc_fence_after_sync();
    // ---- 3. a_now / a_past (columns b .. 3b of D2) -> staging rows [2b][TF3_PITCH]
    {
      int g = 0;
      for (int c0 = 0; c0 < 2 * b; c0 += 8, ++g) {
        if ((g & 3) != quarter) continue;   // warp-uniform
        float v[8];
        tmem_ld8_sum(D2 + lane_base + b + c0, v);
        if (l < TF3_P) {
#pragma unroll
          for (int j = 0; j < 8; ++j) STG[(c0 + j) * TF3_PITCH + l] = v[j];
        }
      }
    }
    __syncthreads();
    // ---- 4. temporal attention + first conv on the mixed rows (attention.py:66, msgat.py:71-72)
    for (int r = tid; r < b * TF3_NT; r += TF3_THREADS) {
      const int o = r / TF3_NT, nl = r - o * TF3_NT;
      const float* an = STG + o * TF3_PITCH + nl * 12;
      const float* ap = STG + (b + o) * TF3_PITCH + nl * 12;
      float vn[12], vp[12], y[12];
#pragma unroll
      for (int i = 0; i < 12; ++i) { vn[i] = an[i]; vp[i] = ap[i]; }
      const float bias = c1B[o];
      if (n0 + nl < a.N) {   // taps before the temporal mix: dPt of the backward contracts them with dY1
        float* an2 = a.A2 + (((int64_t)bi * 2 * b + o) * a.N + n0 + nl) * 12;
        float* ap2 = a.A2 + (((int64_t)bi * 2 * b + b + o) * a.N + n0 + nl) * 12;
        st4(an2, make_float4(vn[0], vn[1], vn[2], vn[3])); st4(an2 + 4, make_float4(vn[4], vn[5], vn[6], vn[7]));
        st4(an2 + 8, make_float4(vn[8], vn[9], vn[10], vn[11]));
        st4(ap2, make_float4(vp[0], vp[1], vp[2], vp[3])); st4(ap2 + 4, make_float4(vp[4], vp[5], vp[6], vp[7]));
        st4(ap2 + 8, make_float4(vp[8], vp[9], vp[10], vp[11]));
      }
#pragma unroll
      for (int t = 0; t < 12; ++t) {
        float s = bias;
#pragma unroll
        for (int i = 0; i < 12; ++i) s = fmaf(Pt[t * 12 + i], vn[i], s);
#pragma unroll
        for (int i = 0; i < 12; ++i) s = fmaf(Ptd[t * 12 + i], vp[i], s);
        y[t] = s;
      }
      const int pp = nl * 12;
      if (n0 + nl < a.N) {
        float* yd = a.Y1 + (((int64_t)bi * b + o) * a.N + n0 + nl) * 12;
        st4(yd, make_float4(y[0], y[1], y[2], y[3])); st4(yd + 4, make_float4(y[4], y[5], y[6], y[7]));
        st4(yd + 8, make_float4(y[8], y[9], y[10], y[11]));
      }
      // shifted copy ys[t] = y[t - d2]: a switch over the uniform dilation keeps every index static
      float ys[12];
#pragma unroll
      for (int t = 0; t < 12; ++t) ys[t] = 0.f;
#define TF3_SHIFT(D) case D: _Pragma("unroll") for (int t = D; t < 12; ++t) ys[t] = y[t - D]; break;
      switch (a.d2) {
        TF3_SHIFT(1) TF3_SHIFT(2) TF3_SHIFT(3) TF3_SHIFT(4) TF3_SHIFT(5) TF3_SHIFT(6) TF3_SHIFT(7) TF3_SHIFT(8) TF3_SHIFT(9)
        TF3_SHIFT(10) TF3_SHIFT(11)
        default: break;
      }
#undef TF3_SHIFT
#pragma unroll
      for (int t0 = 0; t0 < 12; t0 += 4) {
        tf3_split4(YH, YL, sw128b32_off(o, pp + t0, cbY), make_float4(y[t0], y[t0 + 1], y[t0 + 2], y[t0 + 3]));
        tf3_split4(YH, YL, sw128b32_off(b + o, pp + t0, cbY), make_float4(ys[t0], ys[t0 + 1], ys[t0 + 2], ys[t0 + 3]));
      }
    }
    fence_smem_to_async();
    __syncthreads();
    // ---- 5. D3 = [y1 ; shift(y1)]^T [Wnow2 ; Wpast2]^T
    if (tid == 0) {
      tc_fence_after_sync();
      for (int ks = 0; ks < 2 * b / 8; ++ks) mma3(D3, dYH, dYL, dW2H, dW2L, ks, idesc32, ks > 0);
      umma_commit(smem_u32(&bar));
    }
    mbar_wait(smem_u32(&bar), phase); phase ^= 1;
    tc_fence_after_sync();
    // ---- 6. epilogue: one thread per (position, column-group quarter), channels in groups of 8
    {
      const int nl = l / 12, n = n0 + nl, t = l - nl * 12;
      const bool live = l < TF3_P && n < a.N;
      float* ob = a.out + (int64_t)bi * O * a.N * 12 + (int64_t)n * 12 + t;
      const int64_t cs = (int64_t)a.N * 12;
      int g = 0;
      for (int c0 = 0; c0 < O; c0 += 8, ++g) {
        if ((g & 3) != quarter) continue;   // warp-uniform
        float r[8], v[8];
        tmem_ld8_sum(D1 + lane_base + c0, r);
        if (c0 < b) {
          tmem_ld8_sum(D2 + lane_base + c0, v);
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] += caB[c0 + j];
        } else if (c0 < 2 * b) {
          tmem_ld8_sum(D3 + lane_base + (c0 - b), v);
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] += c2B[c0 - b + j];
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = sp[j];   // gathered inside the first MMA window (c0 == sp_c0)
        }
        if (live) {
#pragma unroll
          for (int j = 0; j < 8; ++j) ob[(c0 + j) * cs] = fmaxf(v[j] + r[j] + resB[c0 + j], 0.f);
        }
      }
    }
    tc_fence_before_sync();
    __syncthreads();
  }
  if (warp == 0) tmem_dealloc(tm, 512);
}

bool tc_apply3_supported(const Plan& p) {
  const char* off = getenv("MSGAT_TC_DISABLE");  // test switch: bit 0 = apply kernels
  if (off && (atoi(off) & 1)) return false;
  const int full = MSGAT_F_LN | MSGAT_F_RES | MSGAT_F_RELU | MSGAT_F_CMIX | MSGAT_F_SMIX;
  const int b = p.Oc;
  return p.precision == MSGAT_PREC_FP32 && (p.flags & full) == full && p.n_conv == 2 && p.s_pre && p.Oc == p.Ot &&
         p.Ot == p.Os && b % 8 == 0 && b >= 8 && b <= 24 && p.C % 8 == 0 && p.C >= 8 && p.C <= 72 && p.O <= 72 &&
         p.dil[0] >= 1 && p.dil[0] < 12 && p.dil[1] >= 1 && p.dil[1] < 12;
}

int launch_tc_apply3(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, const float* x, float* out,
                     float* saved, void* stream) {
  TcApply3Args a;
  a.x = x; a.x_sb = p.x_sb; a.x_sc = p.x_sc;
  a.C = p.C; a.N = p.N; a.O = p.O; a.b = p.Oc; a.ntile = (p.N + TF3_NT - 1) / TF3_NT;
  a.d1 = p.dil[0]; a.d2 = p.dil[1]; a.nnz = p.nnz;
  a.ln_w = prm->ln_w; a.ln_b = prm->ln_b; a.res_w = prm->res_w; a.res_b = prm->res_b; a.ca_b = prm->ca_b;
  a.c1_w = prm->conv_w[0]; a.c1_b = prm->conv_b[0]; a.c2_w = prm->conv_w[1]; a.c2_b = prm->conv_b[1];
  a.Mc = saved + p.o_Mc; a.Pt = saved + p.o_Pt; a.Z = saved + p.o_Z; a.pw = saved + p.o_pw;
  a.row_ptr = adj->row_ptr; a.col_idx = adj->col_idx; a.out = out; a.Y1 = saved + p.o_Y1; a.A2 = saved + p.o_A2;
  const int J = tc_ctas_per_sample(p.B, a.ntile, 2, "MSGAT_TF_WAVES");  // all tiles of a CTA in one sample (per-sample Mc / Pt staged once)
  a.J = J;
  MSGAT_LAUNCH(k_tc_apply3, dim3(J, p.B), dim3(TF3_THREADS), tc_apply3_smem(p.C, p.Oc), stream, a);
  return 0;
}
