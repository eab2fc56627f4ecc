// tc_fwd.cu -- tensor-core (tcgen05, kind::tf32) version of the MEAM apply pass for the
// layer-2 shape (C a multiple of 8, full MEAM flags, two dilated convs, W commuted in front of
// the spatial aggregation).  Same inputs and outputs as k_apply<> in meam_fwd.cu.
//
// Every channel mix of the MEAM is a GEMM with the positions (node, t) as M and the channels as
// K (msgat.py:94,116,71-72,27); the temporal attention and the causal shifts act on t only and
// commute with them (attention.py:66), so they are applied to the b = O/3 mixed channels instead
// of the C input channels:
//   D1[p, 0:O]   = X^T  Wres^T                       residual 1x1 conv            (msgat.py:131)
//   D2[p, 0:b]   = H^T (W1x1 Pc)^T                   channel attention + 1x1 conv (attention.py:94, msgat.py:94)
//   D2[p, b:3b]  = H^T [Wnow1 ; Wpast1]^T            first conv taps before the temporal mix
//   y1[o,n,t]    = sum_i Pt[t,i] a_now[o,n,i] + Pt[t-d1,i] a_past[o,n,i] + bias1   (CUDA cores, rows of 12)
//   D3[p, 0:b]   = [y1 ; shift_d2(y1)]^T [Wnow2 ; Wpast2]^T
//   out          = relu([D2[:, :b] + ca_b | D3 + bias2 | sparse aggregation of Z] + D1 + res_b)
// A tile is 16 nodes = 192 positions = two M = 128 blocks (the second one half used); operands
// sit in shared memory in the 128-byte-swizzle / 32-byte-atom layout MN-major TF32 needs
// (tc_common.cuh), accumulators in TMEM (448 of 512 columns).
#include <stdlib.h>
#include "kernels.h"
#include "tc_common.cuh"
#include "tile_ops.cuh"

#define TCF_NT 16
#define TCF_P 192
#define TCF_THREADS 256

struct TcApplyArgs {
  const float* x; int64_t x_sb, x_sc;
  int C, N, O, b, ntile, J, d1, d2, nnz;
  const float *ln_w, *ln_b, *res_w, *res_b, *ca_b, *c1_w, *c1_b, *c2_w, *c2_b;
  const float *Mc, *Pt, *Z, *pw;
  const int32_t *row_ptr, *col_idx;
  float* out;
  float* Y1;  // [B,b,N,12] first-conv output, kept for the weight gradients
  float* A2;  // [B,2b,N,12] first-conv taps before the temporal mix, kept for dPt (k_tc_dgrad)
};

// bytes of one operand tile with R rows and W positions
static inline size_t tc_tile_bytes(int R, int W) { return (size_t)(W / 32) * R * 128; }

size_t tc_apply_smem(int C, int b) {
  size_t s = 0;
  s += 2 * tc_tile_bytes(C, TCF_P);      // XS, HS
  s += 2 * tc_tile_bytes(C, 96);         // WR, WH
  s += tc_tile_bytes(2 * b, 32);         // W2
  s += (144 * 2 + 96 + 32 * 3 + 24 + 4) * 4; // Pt, Ptd, res_b, ca_b/c1_b/c2_b, ln, mbarrier + TMEM slot
  return s + 1024;                       // alignment slack
}

__global__ void __launch_bounds__(TCF_THREADS, 1) k_tc_apply(TcApplyArgs a) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // aligned inside the shared window: the pointer stays a shared-memory pointer (LDS / STS, not generic LD / ST)
  const int C = a.C, b = a.b, O = a.O, tid = threadIdx.x, warp = tid >> 5, bi = blockIdx.y;
  const uint32_t cbX = (uint32_t)C * 128u, cbY = (uint32_t)(2 * b) * 128u;
  unsigned char* XS = smem;
  unsigned char* HS = XS + 6 * cbX;
  unsigned char* WR = HS + 6 * cbX;
  unsigned char* WH = WR + 3 * cbX;
  unsigned char* W2 = WH + 3 * cbX;
  float* Pt = reinterpret_cast<float*>(W2 + cbY);
  float* Ptd = Pt + 144;
  float* resB = Ptd + 144;   // [96]
  float* caB = resB + 96;    // [32]
  float* c1B = caB + 32;
  float* c2B = c1B + 32;
  float* LW = c2B + 32;      // [24]
  uint64_t& bar = *reinterpret_cast<uint64_t*>(LW + 24);   // (dynamic: static shared memory would cut into the opt-in maximum)
  uint32_t& tmem_slot = *reinterpret_cast<uint32_t*>(LW + 26);
  float* STG = reinterpret_cast<float*>(XS);  // [2b][196] staging, aliases XS once the first MMAs are done
  unsigned char* YS = HS;                     // [2b][192] operand tile, aliases HS

  // ---- one-time staging: weights as B operands [K][N] (N contiguous = MN-major), zero padded
  for (int i = tid; i < C * 96; i += TCF_THREADS) {
    const int k = i / 96, n = i - k * 96;
    *reinterpret_cast<float*>(WR + sw128b32_off(k, n, cbX)) = n < O ? a.res_w[n * C + k] : 0.f;
    float w = 0.f;
    if (n < b) w = a.Mc[((int64_t)bi * b + n) * C + k];                 // (W1x1 Pc)[n][k] of this sample
    else if (n < 2 * b) w = a.c1_w[((n - b) * C + k) * 2 + 1];          // tap 1 multiplies x[t]
    else if (n < 3 * b) w = a.c1_w[((n - 2 * b) * C + k) * 2 + 0];      // tap 0 multiplies x[t - d]
    *reinterpret_cast<float*>(WH + sw128b32_off(k, n, cbX)) = w;
  }
  for (int i = tid; i < 2 * b * 32; i += TCF_THREADS) {
    const int k = i / 32, n = i - k * 32;
    float w = 0.f;
    if (n < b) w = k < b ? a.c2_w[(n * b + k) * 2 + 1] : a.c2_w[(n * b + (k - b)) * 2 + 0];
    *reinterpret_cast<float*>(W2 + sw128b32_off(k, n, cbY)) = w;
  }
  for (int i = tid; i < 144; i += TCF_THREADS) {
    const int t = i / 12, j = i - t * 12;
    Pt[i] = a.Pt[(int64_t)bi * 144 + i];
    Ptd[i] = t >= a.d1 ? a.Pt[(int64_t)bi * 144 + (t - a.d1) * 12 + j] : 0.f;
  }
  for (int i = tid; i < 96; i += TCF_THREADS) resB[i] = i < O ? a.res_b[i] : 0.f;
  for (int i = tid; i < 32; i += TCF_THREADS) {
    caB[i] = i < b ? a.ca_b[i] : 0.f; c1B[i] = i < b ? a.c1_b[i] : 0.f; c2B[i] = i < b ? a.c2_b[i] : 0.f;
  }
  if (tid < 12) { LW[tid] = a.ln_w[tid]; LW[12 + tid] = a.ln_b[tid]; }
  if (tid == 0) { mbar_init(smem_u32(&bar), 1); mbar_fence_init(); }
  if (warp == 0) tmem_alloc(smem_u32(&tmem_slot), 512);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tm = tmem_slot;
  const uint32_t D1 = tm, D2 = tm + 192, D3 = tm + 384;  // per M block: +96, +96, +32
  const uint32_t sXS = smem_u32(XS), sHS = smem_u32(HS), sWR = smem_u32(WR), sWH = smem_u32(WH), sW2 = smem_u32(W2);
  const uint32_t idesc80 = umma_idesc_tf32(128, 80, 1, 1), idesc32 = umma_idesc_tf32(128, 32, 1, 1);
  uint32_t phase = 0;
  const int wg = tid >> 7, l = tid & 127;          // warpgroup wg owns M block wg; l = TMEM lane
  const int p = wg * 128 + l;                      // position inside the tile
  const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
  const float* xb = a.x + bi * a.x_sb;
  const float* pwb = a.pw + (int64_t)bi * a.nnz;
  const float* Zb = a.Z + (int64_t)bi * b * a.N * 12;

  for (int tile = blockIdx.x; tile < a.ntile; tile += a.J) {
    const int n0 = tile * TCF_NT;
    // ---- 1. x tile -> XS (raw) and HS (LayerNorm over t, msgat.py:122), both as A operands
    for (int r = tid; r < C * TCF_NT; r += TCF_THREADS) {
      const int c = r / TCF_NT, nl = r - c * TCF_NT, n = n0 + nl;
      float4 u = f4zero(), v = f4zero(), w = f4zero();
      if (n < a.N) { const float* s = xb + c * a.x_sc + (int64_t)n * 12; u = ld4(s); v = ld4(s + 4); w = ld4(s + 8); }
      const int pp = nl * 12;
      st4(reinterpret_cast<float*>(XS + sw128b32_off(c, pp, cbX)), u);
      st4(reinterpret_cast<float*>(XS + sw128b32_off(c, pp + 4, cbX)), v);
      st4(reinterpret_cast<float*>(XS + sw128b32_off(c, pp + 8, cbX)), w);
      if (n < a.N) {
        float mu = (u.x + u.y + u.z + u.w + v.x + v.y + v.z + v.w + w.x + w.y + w.z + w.w) * (1.0f / 12.0f);
        u.x -= mu; u.y -= mu; u.z -= mu; u.w -= mu; v.x -= mu; v.y -= mu; v.z -= mu; v.w -= mu;
        w.x -= mu; w.y -= mu; w.z -= mu; w.w -= mu;
        const float rs = rsqrt_ln((dot4(u, u) + dot4(v, v) + dot4(w, w)) * (1.0f / 12.0f) + LN_EPS);
        u.x = u.x * rs * LW[0] + LW[12]; u.y = u.y * rs * LW[1] + LW[13]; u.z = u.z * rs * LW[2] + LW[14]; u.w = u.w * rs * LW[3] + LW[15];
        v.x = v.x * rs * LW[4] + LW[16]; v.y = v.y * rs * LW[5] + LW[17]; v.z = v.z * rs * LW[6] + LW[18]; v.w = v.w * rs * LW[7] + LW[19];
        w.x = w.x * rs * LW[8] + LW[20]; w.y = w.y * rs * LW[9] + LW[21]; w.z = w.z * rs * LW[10] + LW[22]; w.w = w.w * rs * LW[11] + LW[23];
      }
      st4(reinterpret_cast<float*>(HS + sw128b32_off(c, pp, cbX)), u);
      st4(reinterpret_cast<float*>(HS + sw128b32_off(c, pp + 4, cbX)), v);
      st4(reinterpret_cast<float*>(HS + sw128b32_off(c, pp + 8, cbX)), w);
    }
    fence_smem_to_async();
    __syncthreads();
    // ---- 2. D1 = X^T Wres^T, D2 = H^T [Mc | Wnow1 | Wpast1]^T
    if (tid == 0) {
      tc_fence_after_sync();
      for (int mb = 0; mb < 2; ++mb)
        for (int ks = 0; ks < C / 8; ++ks) {
          const uint32_t ao = (uint32_t)mb * 4u * cbX + (uint32_t)ks * 1024u;
          umma_tf32(D1 + mb * 96, umma_desc(sXS + ao, cbX, 512u, 1), umma_desc(sWR + ks * 1024u, cbX, 512u, 1), idesc80, ks > 0);
          umma_tf32(D2 + mb * 96, umma_desc(sHS + ao, cbX, 512u, 1), umma_desc(sWH + ks * 1024u, cbX, 512u, 1), idesc80, ks > 0);
        }
      umma_commit(smem_u32(&bar));
    }
    mbar_wait(smem_u32(&bar), phase); phase ^= 1;
    tc_fence_after_sync();
    // ---- 3. a_now / a_past (columns b .. 3b of D2) -> staging rows [2b][196]
    if (p < TCF_P) {
      for (int c0 = 0; c0 < 2 * b; c0 += 8) {
        float v[8];
        tmem_ld8(D2 + wg * 96 + lane_base + b + c0, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) STG[(c0 + j) * 196 + p] = v[j];
      }
    } else {  // keep the warp-collective tcgen05.ld uniform
      for (int c0 = 0; c0 < 2 * b; c0 += 8) { float v[8]; tmem_ld8(D2 + wg * 96 + lane_base + b + c0, v); }
    }
    __syncthreads();
    // ---- 4. temporal attention + first conv on the mixed rows (attention.py:66, msgat.py:71-72)
    for (int r = tid; r < b * TCF_NT; r += TCF_THREADS) {
      const int o = r / TCF_NT, nl = r - o * TCF_NT;
      const float* an = STG + o * 196 + nl * 12;
      const float* ap = STG + (b + o) * 196 + nl * 12;
      float vn[12], vp[12], y[12];
#pragma unroll
      for (int i = 0; i < 12; ++i) { vn[i] = an[i]; vp[i] = ap[i]; }
      const float bias = c1B[o];
      if (n0 + nl < a.N) {
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
#pragma unroll
      for (int t0 = 0; t0 < 12; t0 += 4) {
        st4(reinterpret_cast<float*>(YS + sw128b32_off(o, pp + t0, cbY)), make_float4(y[t0], y[t0 + 1], y[t0 + 2], y[t0 + 3]));
        float4 s;
        s.x = t0 >= a.d2 ? y[t0 - a.d2 < 0 ? 0 : t0 - a.d2] : 0.f;
        s.y = t0 + 1 >= a.d2 ? y[t0 + 1 - a.d2 < 0 ? 0 : t0 + 1 - a.d2] : 0.f;
        s.z = t0 + 2 >= a.d2 ? y[t0 + 2 - a.d2 < 0 ? 0 : t0 + 2 - a.d2] : 0.f;
        s.w = t0 + 3 >= a.d2 ? y[t0 + 3 - a.d2 < 0 ? 0 : t0 + 3 - a.d2] : 0.f;
        st4(reinterpret_cast<float*>(YS + sw128b32_off(b + o, pp + t0, cbY)), s);
      }
    }
    fence_smem_to_async();
    __syncthreads();
    // ---- 5. D3 = [y1 ; shift(y1)]^T [Wnow2 ; Wpast2]^T
    if (tid == 0) {
      tc_fence_after_sync();
      for (int mb = 0; mb < 2; ++mb)
        for (int ks = 0; ks < 2 * b / 8; ++ks)
          umma_tf32(D3 + mb * 32, umma_desc(smem_u32(YS) + (uint32_t)mb * 4u * cbY + (uint32_t)ks * 1024u, cbY, 512u, 1),
                    umma_desc(sW2 + ks * 1024u, cbY, 512u, 1), idesc32, ks > 0);
      umma_commit(smem_u32(&bar));
    }
    mbar_wait(smem_u32(&bar), phase); phase ^= 1;
    tc_fence_after_sync();
    // ---- 6. epilogue: one thread per position, channels in groups of 8
    {
      const int nl = p / 12, n = n0 + nl, t = p - nl * 12;
      const bool live = p < TCF_P && n < a.N;
      float* ob = a.out + (int64_t)bi * O * a.N * 12 + (int64_t)n * 12 + t;
      const int64_t cs = (int64_t)a.N * 12;
      int e0 = 0, e1 = 0;
      if (live) { e0 = a.row_ptr[n]; e1 = a.row_ptr[n + 1]; }
      for (int c0 = 0; c0 < O; c0 += 8) {
        float r[8], v[8];
        tmem_ld8(D1 + wg * 96 + lane_base + c0, r);
        if (c0 < b) {
          tmem_ld8(D2 + wg * 96 + lane_base + c0, v);
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] += caB[c0 + j];
        } else if (c0 < 2 * b) {
          tmem_ld8(D3 + wg * 32 + lane_base + (c0 - b), v);
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] += c2B[c0 - b + j];
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = 0.f;
          for (int e = e0; e < e1; ++e) {  // attention.py:36 at the adjacency non-zeros
            const float w = pwb[e];
            const float* z = Zb + ((int64_t)(c0 - 2 * b) * a.N + a.col_idx[e]) * 12 + t;
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = fmaf(w, z[j * cs], v[j]);
          }
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

bool tc_apply_supported(const Plan& p) {
  const char* off = getenv("MSGAT_TC_DISABLE");  // test switch: bit 0 = apply kernel
  if (off && (atoi(off) & 1)) return false;
  const int full = MSGAT_F_LN | MSGAT_F_RES | MSGAT_F_RELU | MSGAT_F_CMIX | MSGAT_F_SMIX;
  const int b = p.Oc;
  return p.precision == MSGAT_PREC_TF32 && (p.flags & full) == full && p.n_conv == 2 && p.s_pre && p.Oc == p.Ot &&
         p.Ot == p.Os && b % 8 == 0 && b >= 8 && b <= 24 && p.C % 8 == 0 && p.C >= 8 && p.C <= 72 && p.O <= 72 &&
         p.dil[0] >= 1 && p.dil[0] < 12 && p.dil[1] >= 1 && p.dil[1] < 12;
}

int launch_tc_apply(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, const float* x, float* out,
                    float* saved, void* stream) {
  TcApplyArgs a;
  a.x = x; a.x_sb = p.x_sb; a.x_sc = p.x_sc;
  a.C = p.C; a.N = p.N; a.O = p.O; a.b = p.Oc; a.ntile = (p.N + TCF_NT - 1) / TCF_NT;
  a.d1 = p.dil[0]; a.d2 = p.dil[1]; a.nnz = p.nnz;
  a.ln_w = prm->ln_w; a.ln_b = prm->ln_b; a.res_w = prm->res_w; a.res_b = prm->res_b; a.ca_b = prm->ca_b;
  a.c1_w = prm->conv_w[0]; a.c1_b = prm->conv_b[0]; a.c2_w = prm->conv_w[1]; a.c2_b = prm->conv_b[1];
  a.Mc = saved + p.o_Mc; a.Pt = saved + p.o_Pt; a.Z = saved + p.o_Z; a.pw = saved + p.o_pw;
  a.row_ptr = adj->row_ptr; a.col_idx = adj->col_idx; a.out = out; a.Y1 = saved + p.o_Y1; a.A2 = saved + p.o_A2;
  const int J = tc_ctas_per_sample(p.B, a.ntile, 2, "MSGAT_TF_WAVES");  // one CTA per SM (512 TMEM columns each), all tiles of a CTA in one sample
  a.J = J;
  const size_t smem = tc_apply_smem(p.C, p.Oc);
  MSGAT_LAUNCH(k_tc_apply, dim3(J, p.B), dim3(TCF_THREADS), smem, stream, a);
  return 0;
}
