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
// Positions are contiguous in HBM, so every operand is K-major.  A tile is 4 nodes = 48
// positions in the unswizzled core-matrix layout (8 rows x 4 positions = 128 contiguous bytes,
// tc_common.cuh:core_off); accumulators stay in TMEM across all tiles of the CTA and are
// flushed once as per-CTA partials (k_reduce sums them in a fixed order).
#include <stdlib.h>
#include "kernels.h"
#include "tc_common.cuh"
#include "tile_ops.cuh"

#define TCW_NT 4
#define TCW_P 48
#define TCW_GS (TCW_P * 32)  // bytes of one 8-row group of a tile
#define TCW_THREADS 512
#define TCW_SLACK (28 * 1024)

struct TcWgradArgs {
  const float* x; int64_t x_sb, x_sc;
  const float* G;   // masked dout [B,O,N,12]
  int C, Cp, N, O, b, ntile, J, d2, has_ln;
  const float *ln_w, *ln_b, *Y1, *tcA;
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
  size_t s = 2 * tcw_tiles_bytes(C, O, b) + TCW_SLACK;
  s += (24 + 8) * 4;   // ln, mbarrier + TMEM slot
  return s + 1024;
}

TC_DEV float tf32_rna(float x) {
  uint32_t u;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x));
  return __uint_as_float(u);
}
// v -> hi tile and lo tile at the same offset
TC_DEV void st_split(unsigned char* hi, unsigned char* lo, uint32_t off, float4 v) {
  float4 h = make_float4(tf32_rna(v.x), tf32_rna(v.y), tf32_rna(v.z), tf32_rna(v.w));
  st4(reinterpret_cast<float*>(hi + off), h);
  // lo is rounded (not truncated by the tensor core) to TF32: residual 2^-23 |v|, unbiased
  st4(reinterpret_cast<float*>(lo + off), make_float4(tf32_rna(v.x - h.x), tf32_rna(v.y - h.y), tf32_rna(v.z - h.z), tf32_rna(v.w - h.w)));
}

__global__ void __launch_bounds__(TCW_THREADS, 1) k_tc_wgrad(TcWgradArgs a) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int C = a.C, Cp = a.Cp, O = a.O, b = a.b, tid = threadIdx.x, warp = tid >> 5, bi = blockIdx.y;
  const int Op = (O + 7) & ~7;
  const uint32_t GS = TCW_GS;
  const uint32_t szG = (uint32_t)(Op / 8) * GS, szX = (uint32_t)(2 * Cp / 8) * GS, szA = (uint32_t)(3 * b / 8) * GS,
                 szY = (uint32_t)(2 * b / 8) * GS;
  const uint32_t part = szG + szX + szA + szY;  // hi part, then the lo part in the same order
  unsigned char* TG = smem;             // [O][48]            G
  unsigned char* TX = TG + szG;         // [2Cp][48]          X rows 0..Cp, H rows Cp..2Cp
  unsigned char* TA = TX + szX;         // [3b][48]           dA_now | dA_past | GZ
  unsigned char* TY = TA + szA;         // [2b][48]           Y1 | shift_d2(Y1)
  float* LW = reinterpret_cast<float*>(smem + 2 * part + TCW_SLACK);
  uint64_t& bar = *reinterpret_cast<uint64_t*>(LW + 24);
  uint32_t& tmem_slot = *reinterpret_cast<uint32_t*>(LW + 26);

  // zero all operand tiles once: pad rows (C..Cp) and the slack must be finite
  for (int i = tid; i < (int)((2 * part + TCW_SLACK) / 16); i += TCW_THREADS)
    reinterpret_cast<float4*>(smem)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (tid < 12) { LW[tid] = a.has_ln ? a.ln_w[tid] : 1.f; LW[12 + tid] = a.has_ln ? a.ln_b[tid] : 0.f; }
  if (tid == 0) { mbar_init(smem_u32(&bar), 1); mbar_fence_init(); }
  if (warp == 0) tmem_alloc(smem_u32(&tmem_slot), 512);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tm = tmem_slot;
  const int N1 = (2 * Cp + 15) & ~15, N2 = (Cp + 15) & ~15, N3 = 2 * b;
  const uint32_t D1 = tm, D2 = tm + 160, D3 = tm + 256;
  const uint32_t id1 = umma_idesc_tf32(128, N1, 0, 0), id2 = umma_idesc_tf32(128, N2, 0, 0), id3 = umma_idesc_tf32(128, N3, 0, 0);
  const uint32_t sTG = smem_u32(TG), sTX = smem_u32(TX), sTA = smem_u32(TA), sTY = smem_u32(TY);
  const float* xb = a.x + bi * a.x_sb;
  const float* Gb = a.G + (int64_t)bi * O * a.N * 12;
  const float* Yb = a.Y1 + (int64_t)bi * b * a.N * 12;
  const float* Ab = a.tcA + (int64_t)bi * 3 * b * a.N * 12;
  uint32_t phase = 0;
  int first = 1;

  // Work items put the row-within-group index in the low 3 bits: 8 consecutive threads fill one
  // 128-byte core matrix, a warp four consecutive ones -> conflict-free shared-memory stores.
  for (int tile = blockIdx.x; tile < a.ntile; tile += a.J) {
    const int n0 = tile * TCW_NT;
    // ---- G tile
    for (int i = tid; i < Op * 12; i += TCW_THREADS) {
      const int r = (i & 7) + 8 * (i / 96), p4 = ((i >> 3) % 12) * 4, n = n0 + p4 / 12;
      float4 v = f4zero();
      if (r < O && n < a.N) v = ld4(Gb + ((int64_t)r * a.N + n0) * 12 + p4);
      st_split(TG, TG + part, core_off(r, p4, GS), v);
    }
    // ---- X rows and their LayerNorm (msgat.py:122)
    for (int i = tid; i < Cp * TCW_NT; i += TCW_THREADS) {
      const int c = (i & 7) + 8 * (i / (8 * TCW_NT)), nl = (i >> 3) % TCW_NT, n = n0 + nl;
      if (c >= C) continue;
      float4 u = f4zero(), v = f4zero(), w = f4zero();
      if (n < a.N) { const float* s = xb + c * a.x_sc + (int64_t)n * 12; u = ld4(s); v = ld4(s + 4); w = ld4(s + 8); }
      const int pp = nl * 12;
      st_split(TX, TX + part, core_off(c, pp, GS), u);
      st_split(TX, TX + part, core_off(c, pp + 4, GS), v);
      st_split(TX, TX + part, core_off(c, pp + 8, GS), w);
      if (n < a.N && a.has_ln) {
        float mu = (u.x + u.y + u.z + u.w + v.x + v.y + v.z + v.w + w.x + w.y + w.z + w.w) * (1.0f / 12.0f);
        u.x -= mu; u.y -= mu; u.z -= mu; u.w -= mu; v.x -= mu; v.y -= mu; v.z -= mu; v.w -= mu;
        w.x -= mu; w.y -= mu; w.z -= mu; w.w -= mu;
        const float rs = rsqrtf((dot4(u, u) + dot4(v, v) + dot4(w, w)) * (1.0f / 12.0f) + LN_EPS);
        u.x = u.x * rs * LW[0] + LW[12]; u.y = u.y * rs * LW[1] + LW[13]; u.z = u.z * rs * LW[2] + LW[14]; u.w = u.w * rs * LW[3] + LW[15];
        v.x = v.x * rs * LW[4] + LW[16]; v.y = v.y * rs * LW[5] + LW[17]; v.z = v.z * rs * LW[6] + LW[18]; v.w = v.w * rs * LW[7] + LW[19];
        w.x = w.x * rs * LW[8] + LW[20]; w.y = w.y * rs * LW[9] + LW[21]; w.z = w.z * rs * LW[10] + LW[22]; w.w = w.w * rs * LW[11] + LW[23];
      }
      st_split(TX, TX + part, core_off(Cp + c, pp, GS), u);
      st_split(TX, TX + part, core_off(Cp + c, pp + 4, GS), v);
      st_split(TX, TX + part, core_off(Cp + c, pp + 8, GS), w);
    }
    // ---- Y1 rows and their causal shift
    for (int i = tid; i < b * TCW_NT; i += TCW_THREADS) {
      const int o = (i & 7) + 8 * (i / (8 * TCW_NT)), nl = (i >> 3) % TCW_NT, n = n0 + nl;
      float y[12];
#pragma unroll
      for (int t = 0; t < 12; ++t) y[t] = 0.f;
      if (n < a.N) {
        const float* s = Yb + ((int64_t)o * a.N + n) * 12;
        const float4 u = ld4(s), v = ld4(s + 4), w = ld4(s + 8);
        y[0] = u.x; y[1] = u.y; y[2] = u.z; y[3] = u.w; y[4] = v.x; y[5] = v.y; y[6] = v.z; y[7] = v.w;
        y[8] = w.x; y[9] = w.y; y[10] = w.z; y[11] = w.w;
      }
      const int pp = nl * 12;
#pragma unroll
      for (int t0 = 0; t0 < 12; t0 += 4) {
        st_split(TY, TY + part, core_off(o, pp + t0, GS), make_float4(y[t0], y[t0 + 1], y[t0 + 2], y[t0 + 3]));
        float4 s;
        s.x = t0 >= a.d2 ? y[t0 - a.d2 < 0 ? 0 : t0 - a.d2] : 0.f;
        s.y = t0 + 1 >= a.d2 ? y[t0 + 1 - a.d2 < 0 ? 0 : t0 + 1 - a.d2] : 0.f;
        s.z = t0 + 2 >= a.d2 ? y[t0 + 2 - a.d2 < 0 ? 0 : t0 + 2 - a.d2] : 0.f;
        s.w = t0 + 3 >= a.d2 ? y[t0 + 3 - a.d2 < 0 ? 0 : t0 + 3 - a.d2] : 0.f;
        st_split(TY, TY + part, core_off(b + o, pp + t0, GS), s);
      }
    }
    // ---- dA_now | dA_past | GZ rows, written by k_bwd_main (meam_bwd.cu) for this kernel
    for (int i = tid; i < 3 * b * 12; i += TCW_THREADS) {
      const int r = (i & 7) + 8 * (i / 96), p4 = ((i >> 3) % 12) * 4, n = n0 + p4 / 12;
      float4 v = f4zero();
      if (n < a.N) v = ld4(Ab + ((int64_t)r * a.N + n0) * 12 + p4);
      st_split(TA, TA + part, core_off(r, p4, GS), v);
    }
    fence_smem_to_async();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after_sync();
      const uint32_t offH = (uint32_t)(Cp / 8) * GS, offT = (uint32_t)(b / 8) * GS;
      for (int k8 = 0; k8 < TCW_P / 8; ++k8) {
        const uint32_t ko = (uint32_t)k8 * 256u;
        const uint32_t acc = (first && k8 == 0) ? 0u : 1u;
        // K-major, no swizzle: LBO = next 4 positions (128 B), SBO = next 8 rows (GS)
        const uint64_t gh = umma_desc(sTG + ko, 128u, GS, 0), gl = umma_desc(sTG + part + ko, 128u, GS, 0);
        const uint64_t xh = umma_desc(sTX + ko, 128u, GS, 0), xl = umma_desc(sTX + part + ko, 128u, GS, 0);
        umma_tf32(D1, gh, xh, id1, acc); umma_tf32(D1, gh, xl, id1, 1u); umma_tf32(D1, gl, xh, id1, 1u);
        const uint64_t ah = umma_desc(sTA + ko, 128u, GS, 0), al = umma_desc(sTA + part + ko, 128u, GS, 0);
        const uint64_t hh = umma_desc(sTX + offH + ko, 128u, GS, 0), hl = umma_desc(sTX + part + offH + ko, 128u, GS, 0);
        umma_tf32(D2, ah, hh, id2, acc); umma_tf32(D2, ah, hl, id2, 1u); umma_tf32(D2, al, hh, id2, 1u);
        const uint64_t th = umma_desc(sTG + offT + ko, 128u, GS, 0), tl = umma_desc(sTG + part + offT + ko, 128u, GS, 0);
        const uint64_t yh = umma_desc(sTY + ko, 128u, GS, 0), yl = umma_desc(sTY + part + ko, 128u, GS, 0);
        umma_tf32(D3, th, yh, id3, acc); umma_tf32(D3, th, yl, id3, 1u); umma_tf32(D3, tl, yh, id3, 1u);
      }
      umma_commit(smem_u32(&bar));
    }
    first = 0;
    mbar_wait(smem_u32(&bar), phase); phase ^= 1;
    tc_fence_after_sync();
  }
  // ---- flush: TMEM lane = output row, columns = input channel
  {
    const int blk = bi * a.J + blockIdx.x;
    const int row = (warp & 3) * 32 + (tid & 31);
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    const int quarter = warp >> 2;  // the four warpgroups split the column groups
    float* wres = a.p_res + (int64_t)blk * a.wg_stride;
    float* wmc = a.p_mc + (int64_t)blk * a.mc_stride;
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
    float* w1n = a.p_c1now + (int64_t)blk * a.wg_stride;
    float* w1p = a.p_c1past + (int64_t)blk * a.wg_stride;
    float* wws = a.p_ws + (int64_t)blk * a.wg_stride;
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
    float* w2n = a.p_c2now + (int64_t)blk * a.wg_stride;
    float* w2p = a.p_c2past + (int64_t)blk * a.wg_stride;
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
  int J = 148 / p.B;
  if (J > ntile) J = ntile;
  return J < 1 ? 1 : J;
}

int launch_tc_wgrad(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, const float* x, const float* G,
                    const float* saved, const float* tcA, float* p_res, float* p_ws, float* p_c1past, float* p_c1now, float* p_c2past,
                    float* p_c2now, int64_t wg_stride, float* p_mc, int64_t mc_stride, void* stream) {
  TcWgradArgs a;
  a.x = x; a.x_sb = p.x_sb; a.x_sc = p.x_sc; a.G = G;
  a.C = p.C; a.Cp = rup8(p.C); a.N = p.N; a.O = p.O; a.b = p.Oc; a.ntile = (p.N + TCW_NT - 1) / TCW_NT;
  a.J = tc_wgrad_blocks_per_sample(p);
  a.d2 = p.dil[1]; a.has_ln = (p.flags & MSGAT_F_LN) ? 1 : 0;
  a.ln_w = prm->ln_w; a.ln_b = prm->ln_b; a.Y1 = saved + p.o_Y1; a.tcA = tcA;
  a.p_res = p_res; a.p_ws = p_ws; a.p_c1past = p_c1past; a.p_c1now = p_c1now; a.p_c2past = p_c2past; a.p_c2now = p_c2now;
  a.wg_stride = wg_stride; a.p_mc = p_mc; a.mc_stride = mc_stride;
  MSGAT_LAUNCH(k_tc_wgrad, dim3(a.J, p.B), dim3(TCW_THREADS), tc_wgrad_smem(p.C, p.O, p.Oc), stream, a);
  return 0;
}
