// tile_ops.cuh -- shared-memory tile primitives of the MEAM kernels.
//
// A tile is NT consecutive nodes of one sample: P = NT*12 positions per channel, stored
// as rows [channel][PS] with PS = P + 4 floats (row pitch = 16 B more than a multiple of
// 128 B, which staggers the banks of consecutive rows for the weight-gradient reads).
// The (n, t) plane is contiguous in HBM ([B,C,N,12], msgat.py:121), so a channel row of a
// tile is one contiguous run of P floats: every global access here is a 128-bit access of
// consecutive threads.
#pragma once
#include "compat.cuh"

#define LN_EPS 1e-5f

// Load rows [C][NT*12] of x for nodes n0.. into dstX (raw) and, if dstH, LayerNorm_T of
// them into dstH (nn.LayerNorm([12]), msgat.py:115,122: biased variance, eps 1e-5).
// Nodes beyond N are zero in both.  Optionally keeps (mean, rstd) per row in stats[2*row].
MSGAT_DEV void tile_load_ln(const float* MSGAT_RESTRICT xb, int64_t x_sc, int C, int N, int n0, int NT, int PS,
                            float* dstX, float* dstH, const float* lnw, const float* lnb, bool do_ln,
                            float* stats) {
  const int rows = C * NT;
  for (int r = threadIdx.x; r < rows; r += blockDim.x) {
    const int c = r / NT, nl = r - c * NT, n = n0 + nl;
    float4 a = f4zero(), b = f4zero(), d = f4zero();
    if (n < N) {
      const float* src = xb + c * x_sc + (int64_t)n * 12;
      a = ld4(src); b = ld4(src + 4); d = ld4(src + 8);
    }
    if (dstX) {
      float* px = dstX + c * PS + nl * 12;
      st4(px, a); st4(px + 4, b); st4(px + 8, d);
    }
    if (dstH) {
      float* ph = dstH + c * PS + nl * 12;
      if (do_ln && n < N) {
        float mu = (a.x + a.y + a.z + a.w + b.x + b.y + b.z + b.w + d.x + d.y + d.z + d.w) * (1.0f / 12.0f);
        a.x -= mu; a.y -= mu; a.z -= mu; a.w -= mu; b.x -= mu; b.y -= mu; b.z -= mu; b.w -= mu;
        d.x -= mu; d.y -= mu; d.z -= mu; d.w -= mu;
        float var = (dot4(a, a) + dot4(b, b) + dot4(d, d)) * (1.0f / 12.0f);
        float rs = rsqrt_ln(var + LN_EPS);
        if (stats) { stats[2 * r] = mu; stats[2 * r + 1] = rs; }
        a.x = a.x * rs * lnw[0] + lnb[0]; a.y = a.y * rs * lnw[1] + lnb[1];
        a.z = a.z * rs * lnw[2] + lnb[2]; a.w = a.w * rs * lnw[3] + lnb[3];
        b.x = b.x * rs * lnw[4] + lnb[4]; b.y = b.y * rs * lnw[5] + lnb[5];
        b.z = b.z * rs * lnw[6] + lnb[6]; b.w = b.w * rs * lnw[7] + lnb[7];
        d.x = d.x * rs * lnw[8] + lnb[8]; d.y = d.y * rs * lnw[9] + lnb[9];
        d.z = d.z * rs * lnw[10] + lnb[10]; d.w = d.w * rs * lnw[11] + lnb[11];
      } else if (stats) {
        stats[2 * r] = 0.f; stats[2 * r + 1] = 0.f;
      }
      st4(ph, a); st4(ph + 4, b); st4(ph + 8, d);
    }
  }
}

// Stage W [Om][K] (row-major, global) transposed into smem Wt [K][WS], WS = rup4(Om),
// zero padded; `gstride` = elements between consecutive k of one row (1 for a plain
// matrix, 2 for one tap of a Conv2d [O,Cin,1,2] weight), `goff` = tap offset.
MSGAT_DEV void stage_wT(float* Wt, const float* MSGAT_RESTRICT W, int Om, int K, int WS, int gstride, int goff) {
  for (int i = threadIdx.x; i < K * WS; i += blockDim.x) {
    int k = i / WS, o = i - k * WS;
    Wt[i] = (o < Om) ? W[((int64_t)o * K + k) * gstride + goff] : 0.f;
  }
}
// Stage W [Om][K] as is (rows = reduction index of a W^T x product) into smem [Om][KS].
MSGAT_DEV void stage_w(float* Ws, const float* MSGAT_RESTRICT W, int Om, int K, int KS, int gstride, int goff) {
  for (int i = threadIdx.x; i < Om * KS; i += blockDim.x) {
    int o = i / KS, k = i - o * KS;
    Ws[i] = (k < K) ? W[((int64_t)o * K + k) * gstride + goff] : 0.f;
  }
}
MSGAT_DEV void stage_vec(float* dst, const float* MSGAT_RESTRICT src, int n, int npad) {
  for (int i = threadIdx.x; i < npad; i += blockDim.x) dst[i] = (i < n && src) ? src[i] : 0.f;
}

// acc[j] += sum_k Wt[k][o0+j] * src[k][nl*12 + t0 .. +3]      (1x1 channel mix)
MSGAT_DEV void mix_item(float4 acc[4], const float* src, int PS, int K, const float* Wt, int WS, int o0, int nl, int t0) {
  const float* s = src + nl * 12 + t0;
  const float* w = Wt + o0;
#pragma unroll 4
  for (int k = 0; k < K; ++k) {
    float4 x = ld4(s + k * PS);
    float4 ww = ld4(w + k * WS);
    fma4(acc[0], ww.x, x); fma4(acc[1], ww.y, x); fma4(acc[2], ww.z, x); fma4(acc[3], ww.w, x);
  }
}
// causal two-tap conv: acc[j] += sum_k Wnow[k][o0+j] x[k][t] + Wpast[k][o0+j] x[k][t-d]
MSGAT_DEV void mix_item_taps(float4 acc[4], const float* src, int PS, int K, const float* Wnow, const float* Wpast,
                             int WS, int o0, int nl, int t0, int d) {
  MSGAT_PIN(d);
#pragma unroll 4
  for (int k = 0; k < K; ++k) {
    const float* row = src + k * PS + nl * 12;
    float4 x = ld4(row + t0);
    float4 xp = quad_past(row, t0, d);
    float4 wa = ld4(Wnow + k * WS + o0);
    float4 wb = ld4(Wpast + k * WS + o0);
    fma4(acc[0], wa.x, x); fma4(acc[1], wa.y, x); fma4(acc[2], wa.z, x); fma4(acc[3], wa.w, x);
    fma4(acc[0], wb.x, xp); fma4(acc[1], wb.y, xp); fma4(acc[2], wb.z, xp); fma4(acc[3], wb.w, xp);
  }
}
// transpose of the two-tap conv: acc[j] += sum_o Wnow[o][c0+j] g[o][t] + Wpast[o][c0+j] g[o][t+d]
// (weights staged untransposed, rows = o)
MSGAT_DEV void mix_item_taps_T(float4 acc[4], const float* src, int PS, int K, const float* Wnow, const float* Wpast,
                               int WS, int o0, int nl, int t0, int d) {
  MSGAT_PIN(d);
#pragma unroll 4
  for (int k = 0; k < K; ++k) {
    const float* row = src + k * PS + nl * 12;
    float4 x = ld4(row + t0);
    float4 xf = quad_future(row, t0, d);
    float4 wa = ld4(Wnow + k * WS + o0);
    float4 wb = ld4(Wpast + k * WS + o0);
    fma4(acc[0], wa.x, x); fma4(acc[1], wa.y, x); fma4(acc[2], wa.z, x); fma4(acc[3], wa.w, x);
    fma4(acc[0], wb.x, xf); fma4(acc[1], wb.y, xf); fma4(acc[2], wb.z, xf); fma4(acc[3], wb.w, xf);
  }
}

// Temporal attention apply on a tile, row-wise: dst[c][n][t] = sum_i Pt[t][i] src[c][n][i]
// (attention.py:66).  With transpose=true: dst[c][n][i] (+)= sum_t Pt[t][i] src[c][n][t].
MSGAT_DEV void tile_temporal(float* dst, const float* src, int PS, int rows_c, int NT, const float* Pt, bool transpose,
                             bool accumulate) {
  const int rows = rows_c * NT;
  for (int r = threadIdx.x; r < rows; r += blockDim.x) {
    const int c = r / NT, nl = r - c * NT;
    const float* s = src + c * PS + nl * 12;
    float v[12], o[12];
    float4 a = ld4(s), b = ld4(s + 4), d = ld4(s + 8);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    v[8] = d.x; v[9] = d.y; v[10] = d.z; v[11] = d.w;
    MSGAT_UNROLL
    for (int t = 0; t < 12; ++t) {
      float acc = 0.f;
      MSGAT_UNROLL
      for (int i = 0; i < 12; ++i) acc = fmaf(transpose ? Pt[i * 12 + t] : Pt[t * 12 + i], v[i], acc);
      o[t] = acc;
    }
    float* q = dst + c * PS + nl * 12;
    if (accumulate) {
      MSGAT_UNROLL
      for (int t = 0; t < 12; ++t) o[t] += q[t];
    }
    st4(q, make_float4(o[0], o[1], o[2], o[3]));
    st4(q + 4, make_float4(o[4], o[5], o[6], o[7]));
    st4(q + 8, make_float4(o[8], o[9], o[10], o[11]));
  }
}
