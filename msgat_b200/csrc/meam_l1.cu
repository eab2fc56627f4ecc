// meam_l1.cu -- purpose-built forward of a FIRST-layer MEAM (msgat.py:121-131 with C = 1..4 input channels: the
// dataset channels of data/meta.yaml, 1 for PEMSD3/7, 3 for PEMSD4/8).
//
// With so few input channels the layer is a pure streaming kernel: one 12-float row per (channel, node) comes in,
// O = 3b rows go out, and every channel mix is a rank-C update.  Per node:
//   h   = LayerNorm_T(x)                                          (msgat.py:122)
//   channel:  out[o]      = ca_b[o] + sum_c Mc[o][c] h[c]          (attention.py:94 folded with msgat.py:94; Mc = W1x1 Pc)
//   temporal: out[b + o]  = bias2[o][t] + sum_c Wa[o][c] ot[c][t] + Wb[o][c] ot[c][t-d1] + Wc[o][c] ot[c][t-d2]
//                                        + Wd[o][c] ot[c][t-d1-d2],   ot = Pt h   (attention.py:66)
//             the two dilated causal convs of TACN (msgat.py:68-74, no activation in between) composed into one
//             four-tap filter: Wa = W2now W1now, Wb = W2now W1past, Wc = W2past W1now, Wd = W2past W1past,
//             bias2[o][t] = b2[o] + sum_k W2now[o][k] b1[k] + [t >= d2] sum_k W2past[o][k] b1[k]
//             (the second conv pads its INPUT with zeros, so the first conv's bias does not reach t < d2 through the past tap)
//   spatial:  out[2b + o] = sum_c Ws[o][c] agg[c],  agg[c] = sum_e pw[e] h[c][col e]     (attention.py:36, msgat.py:27)
//   all:      out = relu(out + res_b + Wres x)                      (msgat.py:131)
// A warp owns 10 consecutive nodes: lane = 3 * node + quad holds positions t0 = 4 * quad .. t0 + 3 of its node
// (30 lanes live), keeps the node's inputs in registers and walks the O output channels; every store instruction of
// the warp is one contiguous run of 480 bytes of a channel plane.  No shared-memory tile, no block barrier inside the
// node loop; the per-sample matrices (Mc, Pt) and the component's weights are staged once per block.
#include "kernels.h"
#include "tile_ops.cuh"

#define L1_THREADS 128
#define L1_WNODES 10   // nodes per warp pass
#define RUP4(v) (((v) + 3) & ~3)

struct L1FwdArgs {
  const float* x; int64_t x_sb, x_sc, x_sm, sv_sm, out_sm;
  int N, O, b, d1, d2, nnz, ngroups, J;
  msgat_meam_params prm[MSGAT_MAX_GROUP];
  const float *Mc, *Pt, *Z, *pw;        // saved (component 0)
  const int32_t *row_ptr, *col_idx;
  float* out;
};

// floats of shared memory: Wres [O][C] | resb [O] | Mc [b][C] | cab [b] | W4 [b][C][4] | bias2 [b][12] | Ws [b][C] | PtP [24][12]
static inline int l1_fwd_smem_floats(int C, int O, int b) {
  return rup4(O * C) + rup4(O) + rup4(b * C) + rup4(b) + b * C * 4 + b * 12 + rup4(b * C) + 24 * 12;
}

// The composed four-tap filter of the TACN chain, in fp64 (one rounding per entry).
MSGAT_DEV void l1_compose_taps(float* W4, float* bias2, const float* c1w, const float* c1b, const float* c2w, const float* c2b,
                               int C, int b, int d2) {
  for (int i = threadIdx.x; i < b * C; i += blockDim.x) {
    const int o = i / C, c = i - o * C;
    double wa = 0.0, wb = 0.0, wc = 0.0, wd = 0.0;
    for (int k = 0; k < b; ++k) {
      const double w2n = (double)c2w[(o * b + k) * 2 + 1], w2p = (double)c2w[(o * b + k) * 2 + 0];
      const double w1n = (double)c1w[(k * C + c) * 2 + 1], w1p = (double)c1w[(k * C + c) * 2 + 0];
      wa += w2n * w1n; wb += w2n * w1p; wc += w2p * w1n; wd += w2p * w1p;
    }
    W4[i * 4 + 0] = (float)wa; W4[i * 4 + 1] = (float)wb; W4[i * 4 + 2] = (float)wc; W4[i * 4 + 3] = (float)wd;
  }
  if (!bias2) return;
  for (int o = threadIdx.x; o < b; o += blockDim.x) {
    double sn = 0.0, sp = 0.0;
    for (int k = 0; k < b; ++k) { sn += (double)c2w[(o * b + k) * 2 + 1] * (double)c1b[k]; sp += (double)c2w[(o * b + k) * 2 + 0] * (double)c1b[k]; }
    for (int t = 0; t < 12; ++t) bias2[o * 12 + t] = (float)((double)c2b[o] + sn + (t >= d2 ? sp : 0.0));
  }
}

// LayerNorm of one row held in three quads (same operation order as tile_load_ln)
MSGAT_DEV void l1_row_ln(float4& a, float4& b, float4& d, const float* lnw, const float* lnb) {
  float mu = (a.x + a.y + a.z + a.w + b.x + b.y + b.z + b.w + d.x + d.y + d.z + d.w) * (1.0f / 12.0f);
  a.x -= mu; a.y -= mu; a.z -= mu; a.w -= mu; b.x -= mu; b.y -= mu; b.z -= mu; b.w -= mu;
  d.x -= mu; d.y -= mu; d.z -= mu; d.w -= mu;
  const float var = (dot4(a, a) + dot4(b, b) + dot4(d, d)) * (1.0f / 12.0f);
  const float rs = rsqrt_ln(var + LN_EPS);
  a.x = a.x * rs * lnw[0] + lnb[0]; a.y = a.y * rs * lnw[1] + lnb[1]; a.z = a.z * rs * lnw[2] + lnb[2]; a.w = a.w * rs * lnw[3] + lnb[3];
  b.x = b.x * rs * lnw[4] + lnb[4]; b.y = b.y * rs * lnw[5] + lnb[5]; b.z = b.z * rs * lnw[6] + lnb[6]; b.w = b.w * rs * lnw[7] + lnb[7];
  d.x = d.x * rs * lnw[8] + lnb[8]; d.y = d.y * rs * lnw[9] + lnb[9]; d.z = d.z * rs * lnw[10] + lnb[10]; d.w = d.w * rs * lnw[11] + lnb[11];
}
MSGAT_DEV float4 l1_pick(const float4& a, const float4& b, const float4& d, int q) { return q == 0 ? a : (q == 1 ? b : d); }
MSGAT_DEV float l1_dot12(const float* row, const float4& a, const float4& b, const float4& d) {
  const float4 r0 = ld4(row), r1 = ld4(row + 4), r2 = ld4(row + 8);
  return dot4(r0, a) + dot4(r1, b) + dot4(r2, d);
}
// the four shifted temporal-attention rows of one channel at this lane's quad: o4[s][j] = (Pt h)[t0 + j - shift_s]
MSGAT_DEV void l1_taps(float4 (&o4)[4], const float* PtP, int t0, int d1, int d2, const float4& a, const float4& b, const float4& d) {
  const int sh[4] = {0, d1, d2, d1 + d2};
  MSGAT_UNROLL
  for (int s = 0; s < 4; ++s) {
    const float* r = PtP + (12 + t0 - sh[s]) * 12;   // rows 0..11 of PtP are zero: t - shift < 0 contributes nothing
    o4[s].x = l1_dot12(r, a, b, d); o4[s].y = l1_dot12(r + 12, a, b, d);
    o4[s].z = l1_dot12(r + 24, a, b, d); o4[s].w = l1_dot12(r + 36, a, b, d);
  }
}

template <int C>
MSGAT_KERNEL MSGAT_LB(L1_THREADS, (C >= 3 ? 3 : 4)) k_l1_fwd(L1FwdArgs a) {
  MSGAT_DYN_SMEM(smem_raw);
  const int b = a.b, O = a.O, bi = blockIdx.y, m = blockIdx.z;
  const msgat_meam_params& PM = a.prm[m];
  float* sp = reinterpret_cast<float*>(smem_raw);
  float* Wres = sp; sp += RUP4(O * C);
  float* resb = sp; sp += RUP4(O);
  float* Mc = sp; sp += RUP4(b * C);
  float* cab = sp; sp += RUP4(b);
  float* W4 = sp; sp += b * C * 4;
  float* bias2 = sp; sp += b * 12;
  float* Ws = sp; sp += RUP4(b * C);
  float* PtP = sp;
  float* LW = PtP + 24 * 12;   // 24 floats, inside the launch's allocation (see l1_fwd_smem_floats + 24)
  const int64_t mo = (int64_t)m * a.sv_sm;
  for (int i = threadIdx.x; i < O * C; i += blockDim.x) Wres[i] = PM.res_w[i];
  for (int i = threadIdx.x; i < O; i += blockDim.x) resb[i] = PM.res_b[i];
  for (int i = threadIdx.x; i < b * C; i += blockDim.x) { Mc[i] = a.Mc[mo + (int64_t)bi * b * C + i]; Ws[i] = PM.sa_W[i]; }
  for (int i = threadIdx.x; i < b; i += blockDim.x) cab[i] = PM.ca_b[i];
  for (int i = threadIdx.x; i < 24 * 12; i += blockDim.x) PtP[i] = i < 144 ? 0.f : a.Pt[mo + (int64_t)bi * 144 + i - 144];
  if (threadIdx.x < 12) { LW[threadIdx.x] = PM.ln_w[threadIdx.x]; LW[12 + threadIdx.x] = PM.ln_b[threadIdx.x]; }
  l1_compose_taps(W4, bias2, PM.conv_w[0], PM.conv_b[0], PM.conv_w[1], PM.conv_b[1], C, b, a.d2);
  __syncthreads();

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const int nl = lane / 3, q = lane - nl * 3, t0 = q * 4;
  const float* xb = a.x + (int64_t)m * a.x_sm + (int64_t)bi * a.x_sb;
  const float* Zb = a.Z + mo + (int64_t)bi * C * a.N * 12;
  const float* pwb = a.pw + mo + (int64_t)bi * a.nnz;
  float* ob = a.out + (int64_t)m * a.out_sm + (int64_t)bi * O * a.N * 12;
  const int64_t cs = (int64_t)a.N * 12;
  for (int g = blockIdx.x * nwarp + warp; g < a.ngroups; g += a.J * nwarp) {
    const int n = g * L1_WNODES + nl;
    if (lane >= 3 * L1_WNODES || n >= a.N) continue;   // no warp-level primitive below: lanes run independently
    float4 x4[C], h4[C], ot[C][4], ag[C];
    MSGAT_UNROLL
    for (int c = 0; c < C; ++c) {
      const float* src = xb + c * a.x_sc + (int64_t)n * 12;
      float4 ra = ld4(src), rb = ld4(src + 4), rd = ld4(src + 8);
      x4[c] = l1_pick(ra, rb, rd, q);
      l1_row_ln(ra, rb, rd, LW, LW + 12);
      h4[c] = l1_pick(ra, rb, rd, q);
      l1_taps(ot[c], PtP, t0, a.d1, a.d2, ra, rb, rd);
      ag[c] = f4zero();
    }
    {
      const int e1 = a.row_ptr[n + 1];
      for (int e = a.row_ptr[n]; e < e1; ++e) {
        const float w = pwb[e];
        const float* zr = Zb + (int64_t)a.col_idx[e] * 12 + t0;
        MSGAT_UNROLL
        for (int c = 0; c < C; ++c) fma4(ag[c], w, ld4(zr + c * cs));
      }
    }
    float* on = ob + (int64_t)n * 12 + t0;
    // channel branch
    for (int o = 0; o < b; ++o) {
      const float bb = cab[o] + resb[o];
      float4 v = make_float4(bb, bb, bb, bb);
      MSGAT_UNROLL
      for (int c = 0; c < C; ++c) { fma4(v, Mc[o * C + c], h4[c]); fma4(v, Wres[o * C + c], x4[c]); }
      v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f);
      st4(on + o * cs, v);
    }
    // temporal branch
    for (int o = 0; o < b; ++o) {
      float4 v = ld4(bias2 + o * 12 + t0);
      const float rb2 = resb[b + o];
      v.x += rb2; v.y += rb2; v.z += rb2; v.w += rb2;
      MSGAT_UNROLL
      for (int c = 0; c < C; ++c) {
        const float4 w = ld4(W4 + (o * C + c) * 4);
        fma4(v, w.x, ot[c][0]); fma4(v, w.y, ot[c][1]); fma4(v, w.z, ot[c][2]); fma4(v, w.w, ot[c][3]);
        fma4(v, Wres[(b + o) * C + c], x4[c]);
      }
      v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f);
      st4(on + (b + o) * cs, v);
    }
    // spatial branch
    for (int o = 0; o < b; ++o) {
      const float bb = resb[2 * b + o];
      float4 v = make_float4(bb, bb, bb, bb);
      MSGAT_UNROLL
      for (int c = 0; c < C; ++c) { fma4(v, Ws[o * C + c], ag[c]); fma4(v, Wres[(2 * b + o) * C + c], x4[c]); }
      v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f);
      st4(on + (2 * b + o) * cs, v);
    }
  }
}

bool l1_fwd_supported(const Plan& p) {
  const char* off = getenv("MSGAT_L1_DISABLE");  // test switch: bit 0 = forward kernel
  if (off && (atoi(off) & 1)) return false;
  const int full = MSGAT_F_LN | MSGAT_F_RES | MSGAT_F_RELU | MSGAT_F_CMIX | MSGAT_F_SMIX;
  return (p.flags & full) == full && p.n_conv == 2 && p.C >= 1 && p.C <= 4 && !p.s_pre && p.Oc == p.Ot && p.Ot == p.Os &&
         p.Oc >= 4 && p.Oc <= 32 && p.dil[0] >= 1 && p.dil[1] >= 1 && p.dil[0] + p.dil[1] < 12;
}

int launch_l1_fwd(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, const float* x, float* out,
                  float* saved, void* stream) {
  L1FwdArgs a;
  a.x = x; a.x_sb = p.x_sb; a.x_sc = p.x_sc; a.x_sm = p.x_sm; a.sv_sm = p.saved_total; a.out_sm = (int64_t)p.B * p.O * p.N * 12;
  a.N = p.N; a.O = p.O; a.b = p.Oc; a.d1 = p.dil[0]; a.d2 = p.dil[1]; a.nnz = p.nnz;
  a.ngroups = (p.N + L1_WNODES - 1) / L1_WNODES;
  for (int m = 0; m < p.M; ++m) a.prm[m] = prm[m];
  a.Mc = saved + p.o_Mc; a.Pt = saved + p.o_Pt; a.Z = saved + p.o_Z; a.pw = saved + p.o_pw;
  a.row_ptr = adj->row_ptr; a.col_idx = adj->col_idx; a.out = out;
  // blocks per sample: about two waves of four resident blocks per SM over all samples of the group
  const int nw = L1_THREADS / 32, need = (a.ngroups + nw - 1) / nw;
  int J = (2 * 4 * msgat_sm_count()) / (p.B * p.M);
  if (J < 1) J = 1;
  if (J > need) J = need;
  a.J = J;
  const size_t smem = (size_t)(l1_fwd_smem_floats(p.C, p.O, p.Oc) + 24) * 4;
  switch (p.C) {
    case 1: MSGAT_LAUNCH(k_l1_fwd<1>, dim3(J, p.B, p.M), dim3(L1_THREADS), smem, stream, a); break;
    case 2: MSGAT_LAUNCH(k_l1_fwd<2>, dim3(J, p.B, p.M), dim3(L1_THREADS), smem, stream, a); break;
    case 3: MSGAT_LAUNCH(k_l1_fwd<3>, dim3(J, p.B, p.M), dim3(L1_THREADS), smem, stream, a); break;
    default: MSGAT_LAUNCH(k_l1_fwd<4>, dim3(J, p.B, p.M), dim3(L1_THREADS), smem, stream, a); break;
  }
  return 0;
}

// =======================================================================================
// Backward of the first-layer MEAM.  One streaming pass over the masked upstream gradient G (O rows per node):
//   data gradients per node (thread = (node, quad), G rows from a shared-memory tile):
//     DXR[c] = sum_o Wres[o][c] G[o]                       residual                      msgat.py:116,131
//     dh[c]  = sum_o Mc[o][c] G_c[o]                       channel branch               attention.py:94
//            + Pt^T dot[c],  dot[c][t] = sum_o Wa G_t[o][t] + Wb G_t[o][t+d1] + Wc G_t[o][t+d2] + Wd G_t[o][t+d1+d2]
//                                                          the composed four-tap filter transposed (see k_l1_fwd)
//     DA[c]  = sum_o Ws[o][c] G_s[o]                       d(aggregated h); its transposed aggregation and the
//                                                          alpha terms are added by k_bwd_fix (first-layer mode)
//     dSnz[e] = pw[e] <DA[n], h[col e]>, delta[n] = sum_e dSnz[e]      spatial softmax backward  attention.py:34-36
//     dPt[t][i] += dot[c][t] h[c][i]                       temporal softmax backward    attention.py:66
//   weight gradients: every one of them is a linear map of the feature correlations
//     Q[o][f] = sum_{n,t} G[o][n][t] F_f[n][t],  F = { x[c], h[c], agg[c], ot[c][t-s] (4 shifts), 1, 1[t >= d2] }
//   because the conv chain's intermediate y1 = b1 + W1now ot + W1past ot(t-d1) lives in the span of those
//   features (k_l1_wfin spells the maps out).  Q is accumulated with thread = (output row, feature set) from
//   the same shared-memory tile: O (5C+2) products per position instead of the 2 b^2 of a direct dW2.
#define L1B_NT 16
#define L1B_P (L1B_NT * 12)
#define L1B_PS (L1B_P + 4)
#define L1B_THREADS 128

struct L1BwdArgs {
  const float* x; int64_t x_sb, x_sc, x_sm, sv_sm, sc_sm, g_sm;
  const float* G;
  int N, O, b, d1, d2, nnz, ntile, J, QSZ;
  msgat_meam_params prm[MSGAT_MAX_GROUP];
  const float *Mc, *Pt, *Z, *pw;
  const int32_t *row_ptr, *col_idx;
  float *DH, *DXR, *DA, *delta, *dSnz, *dMc_part, *Qpart;
  float* dPt_part;   // (as doubles)
};

static inline int l1_qsz(int C, int O, int b) { return rup4(O * C + O + b * (4 * C + 1) + b * C); }
static inline int l1_bwd_smem_floats(int C, int O, int b) {
  return O * L1B_PS + (7 * C + 1) * L1B_PS + L1B_NT * 4 * C * 12 + 2 * L1B_NT * C * 12 + L1B_NT * 4 +
         rup4(O * C) + rup4(b * C) + b * C * 4 + rup4(b * C) + 24 * 12 + 144 + 24 + 2 * (5 * C + 2) * 32;
}

template <int C>
MSGAT_KERNEL MSGAT_LB(L1B_THREADS, (C >= 3 ? 2 : 3)) k_l1_bwd(L1BwdArgs a) {
  constexpr int PS = L1B_PS, NT = L1B_NT, NF = 5 * C + 2;
  MSGAT_DYN_SMEM(smem_raw);
  const int b = a.b, O = a.O, bi = blockIdx.y, m = blockIdx.z, tid = threadIdx.x, T = blockDim.x;
  const msgat_meam_params& PM = a.prm[m];
  float* sp = reinterpret_cast<float*>(smem_raw);
  float* GT = sp; sp += O * PS;                  // G tile [O][PS]
  float* FT = sp; sp += (7 * C + 1) * PS;        // features: x | h | agg | ot(4 shifts) | 1[t >= d2]
  float* EX = sp; sp += NT * 4 * C * 12;         // the four unshifted tap sums of every node
  float* DAs = sp; sp += NT * C * 12;            // d(aggregated h) rows
  float* DOT = sp; sp += NT * C * 12;            // d(Pt h) rows
  float* PD = sp; sp += NT * 4;                  // per-lane parts of delta
  float* Wres = sp; sp += RUP4(O * C);
  float* Mc = sp; sp += RUP4(b * C);
  float* W4 = sp; sp += b * C * 4;
  float* Ws = sp; sp += RUP4(b * C);
  float* PtP = sp; sp += 24 * 12;
  float* LW = sp; sp += 24;
  float* HB = sp;                                // [2][NF][32] temporal half sums (end of kernel)
  const int64_t mo = (int64_t)m * a.sv_sm, so = (int64_t)m * a.sc_sm;
  for (int i = tid; i < O * C; i += T) Wres[i] = PM.res_w[i];
  for (int i = tid; i < b * C; i += T) { Mc[i] = a.Mc[mo + (int64_t)bi * b * C + i]; Ws[i] = PM.sa_W[i]; }
  for (int i = tid; i < 24 * 12; i += T) PtP[i] = i < 144 ? 0.f : a.Pt[mo + (int64_t)bi * 144 + i - 144];
  if (tid < 12) { LW[tid] = PM.ln_w[tid]; LW[12 + tid] = PM.ln_b[tid]; }
  for (int p = tid; p < L1B_P; p += T) FT[7 * C * PS + p] = (p % 12) >= a.d2 ? 1.f : 0.f;
  l1_compose_taps(W4, nullptr, PM.conv_w[0], PM.conv_b[0], PM.conv_w[1], PM.conv_b[1], C, b, a.d2);
  __syncthreads();

  const float* xb = a.x + (int64_t)m * a.x_sm + (int64_t)bi * a.x_sb;
  const float* Gb = a.G + (int64_t)m * a.g_sm + (int64_t)bi * O * a.N * 12;
  const float* Zb = a.Z + mo + (int64_t)bi * C * a.N * 12;
  const float* pwb = a.pw + mo + (int64_t)bi * a.nnz;
  const int64_t cs = (int64_t)a.N * 12;
  const int64_t rowb = so + (int64_t)bi * C * cs;   // [B][C][N][12] scratch tensors
  const bool rowthr = tid < NT * 3;
  const int nl = tid / 3, q = tid - nl * 3, t0 = q * 4;
  const int lane = tid & 31, warp = tid >> 5;
  float ptacc[12][4];
  MSGAT_UNROLL
  for (int t = 0; t < 12; ++t) { ptacc[t][0] = ptacc[t][1] = ptacc[t][2] = ptacc[t][3] = 0.f; }
  float acc[NF];
  MSGAT_UNROLL
  for (int i = 0; i < NF; ++i) acc[i] = 0.f;

  for (int tile = blockIdx.x; tile < a.ntile; tile += a.J) {
    const int n0 = tile * NT;
    // ---- the G tile: O rows of NT nodes (zero beyond N)
    for (int i = tid; i < O * NT * 3; i += T) {
      const int o = i / (NT * 3), r = i - o * (NT * 3), nn = n0 + r / 3;
      cp_async16(GT + o * PS + r * 4, Gb + ((int64_t)o * a.N + (nn < a.N ? nn : 0)) * 12 + (r % 3) * 4, nn < a.N);
    }
    cp_async_commit();
    const int n = n0 + nl;
    const bool live = rowthr && n < a.N;
    // the node's x rows and its edge range are requested while the G tile is in flight (requested behind the barrier they were
    // a second exposed round trip per tile: 14 % of the kernel's samples), and the next tile's G rows are pulled into L2
    float4 xr[C][3];
    int e_lo = 0, e_hi = 0;
    MSGAT_UNROLL
    for (int c = 0; c < C; ++c) {
      xr[c][0] = xr[c][1] = xr[c][2] = f4zero();
      if (live) {
        const float* src = xb + c * a.x_sc + (int64_t)n * 12;
        xr[c][0] = ld4(src); xr[c][1] = ld4(src + 4); xr[c][2] = ld4(src + 8);
      }
    }
    if (live) { e_lo = a.row_ptr[n]; e_hi = a.row_ptr[n + 1]; }
    // ... and so is the node's aggregated row (edge range -> column / weight -> Z rows: three dependent round trips that need
    // nothing from the G tile)
    float4 ag[C];
    MSGAT_UNROLL
    for (int c = 0; c < C; ++c) ag[c] = f4zero();
    constexpr bool AGG_EARLY = C <= 2;   // measured: C = 1 273 -> 261 us; C = 3 327 -> 351 us (its 12 more live registers spill)
    if (AGG_EARLY)
      for (int e = e_lo; e < e_hi; ++e) {
        const float w = pwb[e];
        const float* zr = Zb + (int64_t)a.col_idx[e] * 12 + t0;
        MSGAT_UNROLL
        for (int c = 0; c < C; ++c) fma4(ag[c], w, ld4(zr + c * cs));
      }
#ifndef MSGAT_EMU
    if (tile + a.J < a.ntile) {   // O rows of NT * 48 contiguous bytes: one prefetch per 128-byte line
      const int n1 = (tile + a.J) * NT;
      constexpr int LPR = (NT * 48 + 127) / 128 + 1;   // lines a row's run can touch
      for (int i = tid; i < O * LPR; i += T) {
        const int o = i / LPR, l = i - o * LPR;
        const float* pa = Gb + ((int64_t)o * a.N + n1) * 12 + l * 32;
        if (n1 + (l * 32) / 12 < a.N) asm volatile("prefetch.global.L2 [%0];" ::"l"(pa));
      }
    }
#endif
    cp_async_wait<0>();
    __syncthreads();
    float4 h4[C], dhc[C];
    if (rowthr) {
      float4 x4[C], ot[C][4];
      MSGAT_UNROLL
      for (int c = 0; c < C; ++c) {
        float4 ra = xr[c][0], rb = xr[c][1], rd = xr[c][2];
        x4[c] = h4[c] = f4zero();
        ot[c][0] = ot[c][1] = ot[c][2] = ot[c][3] = f4zero();
        if (live) {
          x4[c] = l1_pick(ra, rb, rd, q);
          l1_row_ln(ra, rb, rd, LW, LW + 12);
          h4[c] = l1_pick(ra, rb, rd, q);
          l1_taps(ot[c], PtP, t0, a.d1, a.d2, ra, rb, rd);
        }
      }
      if (!AGG_EARLY)
        for (int e = e_lo; e < e_hi; ++e) {
          const float w = pwb[e];
          const float* zr = Zb + (int64_t)a.col_idx[e] * 12 + t0;
          MSGAT_UNROLL
          for (int c = 0; c < C; ++c) fma4(ag[c], w, ld4(zr + c * cs));
        }
      const int po = nl * 12 + t0;
      MSGAT_UNROLL
      for (int c = 0; c < C; ++c) {
        st4(FT + c * PS + po, x4[c]); st4(FT + (C + c) * PS + po, h4[c]); st4(FT + (2 * C + c) * PS + po, ag[c]);
        MSGAT_UNROLL
        for (int s = 0; s < 4; ++s) st4(FT + ((3 + s) * C + c) * PS + po, ot[c][s]);
      }
      // ---- walk the O rows of G: transposed channel mixes
      float4 dxr[C], E[C][4], da[C];
      MSGAT_UNROLL
      for (int c = 0; c < C; ++c) { dxr[c] = dhc[c] = da[c] = f4zero(); E[c][0] = E[c][1] = E[c][2] = E[c][3] = f4zero(); }
      const float* gp = GT + po;
      for (int o = 0; o < b; ++o) {
        const float4 g = ld4(gp + o * PS);
        MSGAT_UNROLL
        for (int c = 0; c < C; ++c) { fma4(dxr[c], Wres[o * C + c], g); fma4(dhc[c], Mc[o * C + c], g); }
      }
      for (int o = 0; o < b; ++o) {
        const float4 g = ld4(gp + (b + o) * PS);
        MSGAT_UNROLL
        for (int c = 0; c < C; ++c) {
          const float4 w = ld4(W4 + (o * C + c) * 4);
          fma4(dxr[c], Wres[(b + o) * C + c], g);
          fma4(E[c][0], w.x, g); fma4(E[c][1], w.y, g); fma4(E[c][2], w.z, g); fma4(E[c][3], w.w, g);
        }
      }
      for (int o = 0; o < b; ++o) {
        const float4 g = ld4(gp + (2 * b + o) * PS);
        MSGAT_UNROLL
        for (int c = 0; c < C; ++c) { fma4(dxr[c], Wres[(2 * b + o) * C + c], g); fma4(da[c], Ws[o * C + c], g); }
      }
      MSGAT_UNROLL
      for (int c = 0; c < C; ++c) {
        MSGAT_UNROLL
        for (int s = 0; s < 4; ++s) st4(EX + ((nl * 4 + s) * C + c) * 12 + t0, E[c][s]);
        st4(DAs + (nl * C + c) * 12 + t0, da[c]);
        if (live) {
          st4(a.DXR + rowb + ((int64_t)c * a.N + n) * 12 + t0, dxr[c]);
          st4(a.DA + rowb + ((int64_t)c * a.N + n) * 12 + t0, da[c]);
        }
      }
    }
    __syncthreads();
    if (rowthr) {   // dot[c][t] = Ea[t] + Eb[t + d1] + Ec[t + d2] + Ed[t + d1 + d2]   (zero past the window)
      const int sh[4] = {0, a.d1, a.d2, a.d1 + a.d2};
      MSGAT_UNROLL
      for (int c = 0; c < C; ++c) {
        float v[4] = {0.f, 0.f, 0.f, 0.f};
        MSGAT_UNROLL
        for (int s = 0; s < 4; ++s) {
          const float* e = EX + ((nl * 4 + s) * C + c) * 12;
          MSGAT_UNROLL
          for (int j = 0; j < 4; ++j) { const int t = t0 + j + sh[s]; if (t < 12) v[j] += e[t]; }
        }
        st4(DOT + (nl * C + c) * 12 + t0, make_float4(v[0], v[1], v[2], v[3]));
      }
    }
    __syncthreads();
    float pd = 0.f;
    if (live) {
      MSGAT_UNROLL
      for (int c = 0; c < C; ++c) {
        const float* dr = DOT + (nl * C + c) * 12;
        float4 dh = dhc[c];
        MSGAT_UNROLL
        for (int t = 0; t < 12; ++t) {
          const float dv = dr[t];
          fma4(dh, dv, ld4(PtP + (12 + t) * 12 + t0));          // dh[i] += Pt[t][i] dot[t]
          ptacc[t][0] = fmaf(dv, h4[c].x, ptacc[t][0]); ptacc[t][1] = fmaf(dv, h4[c].y, ptacc[t][1]);
          ptacc[t][2] = fmaf(dv, h4[c].z, ptacc[t][2]); ptacc[t][3] = fmaf(dv, h4[c].w, ptacc[t][3]);
        }
        st4(a.DH + rowb + ((int64_t)c * a.N + n) * 12 + t0, dh);
      }
      // dS at the adjacency non-zeros of this row, edges e = q, q + 3, ... of the row
      const int e1 = e_hi;
      for (int e = e_lo + q; e < e1; e += 3) {
        const float* zr = Zb + (int64_t)a.col_idx[e] * 12;
        float s = 0.f;
        MSGAT_UNROLL
        for (int c = 0; c < C; ++c) {
          const float* d = DAs + (nl * C + c) * 12;
          s += dot4(ld4(d), ld4(zr + c * cs)) + dot4(ld4(d + 4), ld4(zr + c * cs + 4)) + dot4(ld4(d + 8), ld4(zr + c * cs + 8));
        }
        const float dsn = pwb[e] * s;
        a.dSnz[so + (int64_t)bi * a.nnz + e] = dsn;
        pd += dsn;
      }
    }
    if (rowthr) PD[nl * 4 + q] = pd;
    // ---- feature correlations: warp 0 channel rows, warp 1 spatial rows, warps 2 and 3 temporal rows (half the
    // positions each).  Tile-local sums first, then one add into the long-running accumulators.
    if (lane < b) {
      float loc[NF];
      MSGAT_UNROLL
      for (int i = 0; i < NF; ++i) loc[i] = 0.f;
      if (warp < 2) {
        const float* gr = GT + (warp == 0 ? lane : 2 * b + lane) * PS;
        const float* f2 = FT + (warp == 0 ? C : 2 * C) * PS;   // h rows (channel) or agg rows (spatial)
        for (int p4 = 0; p4 < NT * 3; ++p4) {
          const float4 g = ld4(gr + p4 * 4);
          MSGAT_UNROLL
          for (int c = 0; c < C; ++c) {
            loc[c] += dot4(g, ld4(FT + c * PS + p4 * 4));
            loc[C + c] += dot4(g, ld4(f2 + c * PS + p4 * 4));
          }
          loc[2 * C] += (g.x + g.y) + (g.z + g.w);
        }
      } else {
        const float* gr = GT + (b + lane) * PS;
        const int lo = (warp - 2) * (NT * 3 / 2), hi = lo + NT * 3 / 2;
        for (int p4 = lo; p4 < hi; ++p4) {
          const float4 g = ld4(gr + p4 * 4);
          MSGAT_UNROLL
          for (int c = 0; c < C; ++c) {
            loc[c] += dot4(g, ld4(FT + c * PS + p4 * 4));
            MSGAT_UNROLL
            for (int s = 0; s < 4; ++s) loc[C + s * C + c] += dot4(g, ld4(FT + ((3 + s) * C + c) * PS + p4 * 4));
          }
          loc[5 * C] += (g.x + g.y) + (g.z + g.w);
          loc[5 * C + 1] += dot4(g, ld4(FT + 7 * C * PS + p4 * 4));
        }
      }
      MSGAT_UNROLL
      for (int i = 0; i < NF; ++i) acc[i] += loc[i];
    }
    __syncthreads();
    if (live && q == 0) a.delta[so + (int64_t)bi * a.N + n] = PD[nl * 4] + PD[nl * 4 + 1] + PD[nl * 4 + 2];
  }

  // ---- flush the per-block partials
  const int blk = bi * a.J + blockIdx.x;
  float* Q = a.Qpart + so + (int64_t)blk * a.QSZ;
  const int oQB = O * C, oQT = O * C + O, oQS = oQT + b * (4 * C + 1);
  if (warp == 3 && lane < b) {
    MSGAT_UNROLL
    for (int i = 0; i < NF; ++i) HB[i * 32 + lane] = acc[i];
  }
  __syncthreads();
  if (lane < b) {
    if (warp == 0) {
      float* dmc = a.dMc_part + so + (int64_t)blk * b * C;
      MSGAT_UNROLL
      for (int c = 0; c < C; ++c) { Q[lane * C + c] = acc[c]; dmc[lane * C + c] = acc[C + c]; }
      Q[oQB + lane] = acc[2 * C];
    } else if (warp == 1) {
      MSGAT_UNROLL
      for (int c = 0; c < C; ++c) { Q[(2 * b + lane) * C + c] = acc[c]; Q[oQS + lane * C + c] = acc[C + c]; }
      Q[oQB + 2 * b + lane] = acc[2 * C];
    } else if (warp == 2) {
      MSGAT_UNROLL
      for (int i = 0; i < NF; ++i) acc[i] += HB[i * 32 + lane];
      MSGAT_UNROLL
      for (int c = 0; c < C; ++c) {
        Q[(b + lane) * C + c] = acc[c];
        MSGAT_UNROLL
        for (int s = 0; s < 4; ++s) Q[oQT + lane * (4 * C + 1) + s * C + c] = acc[C + s * C + c];
      }
      Q[oQB + b + lane] = acc[5 * C];
      Q[oQT + lane * (4 * C + 1) + 4 * C] = acc[5 * C + 1];
    }
  }
  // dPt partial of the block: sum the row threads' 12 x 4 blocks over the nodes
  __syncthreads();
  if (rowthr) {
    MSGAT_UNROLL
    for (int t = 0; t < 12; ++t)
      MSGAT_UNROLL
      for (int j = 0; j < 4; ++j) GT[tid * 48 + t * 4 + j] = ptacc[t][j];
  }
  __syncthreads();
  for (int i = tid; i < 144; i += T) {
    const int t = i / 12, ii = i - t * 12, qq = ii >> 2, j = ii & 3;
    double s = 0.0;
    for (int k = 0; k < NT; ++k) s += (double)GT[(k * 3 + qq) * 48 + t * 4 + j];
    reinterpret_cast<double*>(a.dPt_part + so)[(int64_t)blk * 144 + i] = s;
  }
}

// Weight gradients of the first-layer MEAM from the feature correlations (one block per component).
struct L1WfinArgs {
  const float* Qpart; int64_t sc_sm;
  int C, O, b, nparts, QSZ;
  msgat_meam_params prm[MSGAT_MAX_GROUP];
  msgat_meam_grads g[MSGAT_MAX_GROUP];
};

MSGAT_KERNEL MSGAT_LB(256, 1) k_l1_wfin(L1WfinArgs a) {
  MSGAT_DYN_SMEM(smem_raw);
  double* Q = reinterpret_cast<double*>(smem_raw);
  const int m = blockIdx.x, C = a.C, O = a.O, b = a.b, tid = threadIdx.x, T = blockDim.x, nq = O * C + O + b * (4 * C + 1) + b * C;
  const msgat_meam_params& PM = a.prm[m];
  const msgat_meam_grads& GR = a.g[m];
  const float* src = a.Qpart + (int64_t)m * a.sc_sm;
  for (int i = tid; i < nq; i += T) {
    double s = 0.0;
    for (int p = 0; p < a.nparts; ++p) s += (double)src[(int64_t)p * a.QSZ + i];
    Q[i] = s;
  }
  __syncthreads();
  const double* QR = Q;
  const double* QB = Q + O * C;
  const double* QT = QB + O;
  const double* QS = QT + b * (4 * C + 1);
  const int FS = 4 * C + 1;
  const float* c1w = PM.conv_w[0];
  const float* c1b = PM.conv_b[0];
  const float* c2w = PM.conv_w[1];
  if (GR.res_w) for (int i = tid; i < O * C; i += T) GR.res_w[i] = (float)QR[i];
  if (GR.res_b) for (int i = tid; i < O; i += T) GR.res_b[i] = (float)QB[i];
  if (GR.ca_b) for (int i = tid; i < b; i += T) GR.ca_b[i] = (float)QB[i];
  if (GR.conv_b[1]) for (int i = tid; i < b; i += T) GR.conv_b[1][i] = (float)QB[b + i];
  if (GR.sa_W) for (int i = tid; i < b * C; i += T) GR.sa_W[i] = (float)QS[i];
  // second conv (msgat.py:71-72): dW2now[o][k] = sum_p G_t[o][p] y1[k][p], dW2past with y1 shifted by d2, where
  // y1[k] = b1[k] + sum_c W1now[k][c] ot[c] + W1past[k][c] ot[c](t - d1)
  if (GR.conv_w[1])
    for (int i = tid; i < b * b; i += T) {
      const int o = i / b, k = i - o * b;
      double sn = (double)c1b[k] * QB[b + o], spv = (double)c1b[k] * QT[o * FS + 4 * C];
      for (int c = 0; c < C; ++c) {
        const double w1n = (double)c1w[(k * C + c) * 2 + 1], w1p = (double)c1w[(k * C + c) * 2 + 0];
        sn += w1n * QT[o * FS + c] + w1p * QT[o * FS + C + c];
        spv += w1n * QT[o * FS + 2 * C + c] + w1p * QT[o * FS + 3 * C + c];
      }
      GR.conv_w[1][i * 2 + 1] = (float)sn;
      GR.conv_w[1][i * 2 + 0] = (float)spv;
    }
  // first conv: dy1[k][p] = sum_o W2now[o][k] G_t[o][p] + W2past[o][k] G_t[o][p + d2]
  if (GR.conv_b[0])
    for (int k = tid; k < b; k += T) {
      double s = 0.0;
      for (int o = 0; o < b; ++o) s += (double)c2w[(o * b + k) * 2 + 1] * QB[b + o] + (double)c2w[(o * b + k) * 2 + 0] * QT[o * FS + 4 * C];
      GR.conv_b[0][k] = (float)s;
    }
  if (GR.conv_w[0])
    for (int i = tid; i < b * C; i += T) {
      const int k = i / C, c = i - k * C;
      double sn = 0.0, spv = 0.0;
      for (int o = 0; o < b; ++o) {
        const double w2n = (double)c2w[(o * b + k) * 2 + 1], w2p = (double)c2w[(o * b + k) * 2 + 0];
        sn += w2n * QT[o * FS + c] + w2p * QT[o * FS + 2 * C + c];
        spv += w2n * QT[o * FS + C + c] + w2p * QT[o * FS + 3 * C + c];
      }
      GR.conv_w[0][i * 2 + 1] = (float)sn;
      GR.conv_w[0][i * 2 + 0] = (float)spv;
    }
}

bool l1_bwd_supported(const Plan& p) {
  const char* off = getenv("MSGAT_L1_DISABLE");  // test switch: bit 1 = backward kernels
  if (off && (atoi(off) & 2)) return false;
  const int full = MSGAT_F_LN | MSGAT_F_RES | MSGAT_F_RELU | MSGAT_F_CMIX | MSGAT_F_SMIX;
  return (p.flags & full) == full && p.n_conv == 2 && p.C >= 1 && p.C <= 4 && !p.s_pre && p.Oc == p.Ot && p.Ot == p.Os &&
         p.Oc >= 4 && p.Oc <= 32 && p.dil[0] >= 1 && p.dil[1] >= 1 && p.dil[0] + p.dil[1] < 12;
}

int l1_bwd_blocks_per_sample(const Plan& p) {
  const int ntile = (p.N + L1B_NT - 1) / L1B_NT;
  const int occ = p.C >= 3 ? 2 : 3;
  int J = (2 * occ * msgat_sm_count()) / (p.B * p.M);
  if (J < 1) J = 1;
  return J > ntile ? ntile : J;
}
int l1_bwd_qsz(const Plan& p) { return l1_qsz(p.C, p.O, p.Oc); }

int launch_l1_bwd(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, const float* x, const float* G, int64_t g_sm,
                  const float* saved, float* DH, float* DXR, float* DA, float* delta, float* dSnz, float* dMc_part,
                  float* dPt_part, float* Qpart, void* stream) {
  L1BwdArgs a;
  a.x = x; a.x_sb = p.x_sb; a.x_sc = p.x_sc; a.x_sm = p.x_sm; a.sv_sm = p.saved_total; a.sc_sm = p.scratch_sm; a.g_sm = g_sm;
  a.G = G; a.N = p.N; a.O = p.O; a.b = p.Oc; a.d1 = p.dil[0]; a.d2 = p.dil[1]; a.nnz = p.nnz;
  a.ntile = (p.N + L1B_NT - 1) / L1B_NT; a.J = l1_bwd_blocks_per_sample(p); a.QSZ = l1_bwd_qsz(p);
  for (int m = 0; m < p.M; ++m) a.prm[m] = prm[m];
  a.Mc = saved + p.o_Mc; a.Pt = saved + p.o_Pt; a.Z = saved + p.o_Z; a.pw = saved + p.o_pw;
  a.row_ptr = adj->row_ptr; a.col_idx = adj->col_idx;
  a.DH = DH; a.DXR = DXR; a.DA = DA; a.delta = delta; a.dSnz = dSnz; a.dMc_part = dMc_part; a.dPt_part = dPt_part; a.Qpart = Qpart;
  const size_t smem = (size_t)l1_bwd_smem_floats(p.C, p.O, p.Oc) * 4;
  const dim3 grid(a.J, p.B, p.M);
  switch (p.C) {
    case 1: MSGAT_LAUNCH(k_l1_bwd<1>, grid, dim3(L1B_THREADS), smem, stream, a); break;
    case 2: MSGAT_LAUNCH(k_l1_bwd<2>, grid, dim3(L1B_THREADS), smem, stream, a); break;
    case 3: MSGAT_LAUNCH(k_l1_bwd<3>, grid, dim3(L1B_THREADS), smem, stream, a); break;
    default: MSGAT_LAUNCH(k_l1_bwd<4>, grid, dim3(L1B_THREADS), smem, stream, a); break;
  }
  return 0;
}

int launch_l1_wfin(const Plan& p, const msgat_meam_params* prm, const msgat_meam_grads* g, const float* Qpart, int nparts, void* stream) {
  L1WfinArgs a;
  a.Qpart = Qpart; a.sc_sm = p.scratch_sm; a.C = p.C; a.O = p.O; a.b = p.Oc; a.nparts = nparts; a.QSZ = l1_bwd_qsz(p);
  for (int m = 0; m < p.M; ++m) { a.prm[m] = prm[m]; a.g[m] = g[m]; }
  MSGAT_LAUNCH(k_l1_wfin, dim3(p.M), dim3(256), (size_t)a.QSZ * sizeof(double), stream, a);
  return 0;
}
