// meam_fwd.cu -- forward of the MEAM-shaped operator (reference msgat.py:121-131 and the
// three attentions of attention.py) as four launches:
//   k_prep    reads x once: LayerNorm_T, the spatial query q_s, the temporal query q_t,
//             Z = W h (or h) for the spatial gather, per-tile partial sums of the two
//             per-sample reductions (channel query over N, temporal keys over N);
//   k_small   per sample: finish the reductions, C x C and 12 x 12 softmaxes, fold the
//             1x1 conv of CACN into the channel attention (Mc = W1 Pc);
//   k_sstats  (spatial.cu) softmax statistics of q Wg q^T over ALL N columns + the
//             attention weights at the adjacency non-zeros;
//   k_apply   reads x again: LN, residual 1x1, Mc h, temporal attention + dilated causal
//             conv chain, sparse spatial aggregation, concat, ReLU, one write of the output.
// The N x N attention matrix never exists in HBM.
#include <stdlib.h>
#include "kernels.h"
#include "tile_ops.cuh"

// ---------------------------------------------------------------------------------------
// Grouped launch: blockIdx.z = component m; every pointer below is component 0's, component m's buffers lie
// x_sm / sv_sm / sc_sm floats further (x, saved, scratch), its parameters are prm[m].
struct PrepArgs {
  const float* x; int64_t x_sb, x_sc, x_sm, sv_sm, sc_sm;
  int C, N, Oc, Ot, Os, Cz, s_pre, do_ln, ntile;
  msgat_meam_params prm[MSGAT_MAX_GROUP];
  float *qs, *qt, *Z;
  float *pqc, *pK;  // (as doubles) per-tile partial sums kept in fp64: they feed softmaxes whose scores reach O(100)
};

template <int NT>
MSGAT_KERNEL MSGAT_LB(256, 3) k_prep(PrepArgs a) {
  constexpr int P = NT * 12, PS = P + 4, PQ = NT * 3;
  MSGAT_DYN_SMEM(smem_raw);
  float* H = reinterpret_cast<float*>(smem_raw);  // [C][PS]
  float* QT = H + a.C * PS;                       // [P]
  float* LW = QT + P;                             // ln w,b [24]
  float* WsT = LW + 24;                           // [C][rup4(Os)] when s_pre
  const int WS = (a.Os + 3) & ~3;
  double* Wd = reinterpret_cast<double*>(WsT + (a.s_pre ? a.C * WS : 0));   // [20][NT + 1] Wt1 | Wt2 columns of the tile, as doubles
  double* QTd = Wd + 20 * (NT + 1);               // [P] the temporal query, as doubles
  const int tile = blockIdx.x, b = blockIdx.y, m = blockIdx.z, n0 = tile * NT;
  const msgat_meam_params& PM = a.prm[m];
  float* const qs_m = a.qs + m * a.sv_sm;
  float* const qt_m = a.qt + m * a.sv_sm;
  float* const Z_m = a.Z + m * a.sv_sm;
  double* const pqc_m = reinterpret_cast<double*>(a.pqc + m * a.sc_sm);
  double* const pK_m = reinterpret_cast<double*>(a.pK + m * a.sc_sm);
  const int lim = a.N - n0 < NT ? a.N - n0 : NT;   // live nodes of the tile
  if (threadIdx.x < 12) {
    LW[threadIdx.x] = a.do_ln ? PM.ln_w[threadIdx.x] : 1.f;
    LW[12 + threadIdx.x] = a.do_ln ? PM.ln_b[threadIdx.x] : 0.f;
  }
  // the tile's columns of Wt1 / Wt2 converted ONCE (the key partials below are 240 x NT fp64 products: converting both operands
  // per product was 2 x 240 x NT conversions on the quarter-rate pipe, 46 % of the kernel's samples with the loads behind them)
  if (a.Ot)
    for (int i = threadIdx.x; i < 20 * NT; i += blockDim.x) {
      const int r = i / NT, nl = i - r * NT;
      const float* W = (r < 10) ? PM.ta_Wt1 + (int64_t)r * a.N : PM.ta_Wt2 + (int64_t)(r - 10) * a.N;
      Wd[r * (NT + 1) + nl] = nl < lim ? (double)__ldg(W + n0 + nl) : 0.0;
    }
  if (a.s_pre) stage_wT(WsT, PM.sa_W, a.Os, a.C, WS, 1, 0);
  __syncthreads();
  tile_load_ln(a.x + m * a.x_sm + b * a.x_sb, a.x_sc, a.C, a.N, n0, NT, PS, nullptr, H, LW, LW + 12, a.do_ln, nullptr);
  __syncthreads();
  // spatial / temporal queries: q[n,t] = sum_c alpha[c] h[c,n,t]   (attention.py:33,63)
  for (int p = threadIdx.x; p < P; p += blockDim.x) {
    const int nl = p / 12, n = n0 + nl;
    float s = 0.f, t = 0.f;
    if (a.Os) for (int c = 0; c < a.C; ++c) s = fmaf(__ldg(PM.sa_alpha + c), H[c * PS + p], s);
    if (a.Ot) for (int c = 0; c < a.C; ++c) t = fmaf(__ldg(PM.ta_alpha + c), H[c * PS + p], t);
    QTd[p] = (double)t;
    if (n < a.N) {
      if (a.Os) qs_m[((int64_t)b * a.N + n0) * 12 + p] = s;
      if (a.Ot) qt_m[((int64_t)b * a.N + n0) * 12 + p] = t;
    }
  }
  // Z for the spatial aggregation: W h when that narrows the rows (msgat.py:27 commuted in
  // front of the aggregation: A (h W^T) == (A h) W^T), else h itself.
  if (a.Os) {
    const int nch = (a.Cz + 3) >> 2;
    for (int it = threadIdx.x; it < nch * PQ; it += blockDim.x) {
      const int ch = it / PQ, pq = it - ch * PQ, nl = pq / 3, t0 = (pq - nl * 3) * 4, o0 = ch * 4, n = n0 + nl;
      float4 acc[4] = {f4zero(), f4zero(), f4zero(), f4zero()};
      if (a.s_pre) {
        mix_item(acc, H, PS, a.C, WsT, WS, o0, nl, t0);
      } else {
        MSGAT_UNROLL
        for (int j = 0; j < 4; ++j)
          if (o0 + j < a.Cz) acc[j] = ld4(H + (o0 + j) * PS + nl * 12 + t0);
      }
      if (n < a.N)
        MSGAT_UNROLL
        for (int j = 0; j < 4; ++j)
          if (o0 + j < a.Cz) st4(Z_m + (((int64_t)b * a.Cz + o0 + j) * a.N + n) * 12 + t0, acc[j]);
    }
  }
  __syncthreads();   // QTd (and H, already synchronised) complete
  // per-tile partial of the temporal keys K1 = q_t Wt1^T, K2 = q_t Wt2^T  (attention.py:65): threads 0..239
  if (a.Ot && threadIdx.x < 240) {
    double* dst = pK_m + ((int64_t)b * a.ntile + tile) * 240;
    const int i = threadIdx.x, t = i / 20, r = i - t * 20;
    const double* w = Wd + r * (NT + 1);
    double s = 0.0;
    MSGAT_UNROLL_N(8)
    for (int nl = 0; nl < lim; ++nl) s += w[nl] * QTd[nl * 12 + t];
    dst[i] = s;
  }
  // per-tile partial of the channel query q_c[c,t] = sum_n alpha[n] h[c,n,t]  (attention.py:92).  Independent of the keys: the
  // items start on the 16 threads the key partials leave idle (a narrow input has 12 items: both run side by side; behind a
  // barrier of its own this phase was 20 % of the kernel's samples)
  if (a.Oc) {
    double* dst = pqc_m + ((int64_t)b * a.ntile + tile) * a.C * 12;
    for (int i = (threadIdx.x + blockDim.x - 240) % blockDim.x; i < a.C * 12; i += blockDim.x) {
      const int c = i / 12, t = i - c * 12;
      double s = 0.0;
      // same order of additions; without the per-node guard the loads of eight nodes are requested together
      MSGAT_UNROLL_N(8)
      for (int nl = 0; nl < lim; ++nl) s += (double)__ldg(PM.ca_alpha + n0 + nl) * (double)H[c * PS + nl * 12 + t];
      dst[i] = s;
    }
  }
}

// ---------------------------------------------------------------------------------------
struct SmallArgs {   // grouped like PrepArgs: blockIdx.y = component
  int C, Oc, Ot, cmix, ntile;
  int l2, N;           // l2: partials from k_l2_mix (fp32 q_c partials per 8-node tile; K from the saved q_t)
  const float* qt;     // saved q_t [B,N,12] (l2)
  int64_t sv_sm, sc_sm;
  const float *pqc, *pK;   // (as doubles)
  msgat_meam_params prm[MSGAT_MAX_GROUP];
  float *qc, *Pc, *Mc, *K, *Pt;
};

// All of this is O(C^2) work per sample; it runs in fp64 because the scores of the channel and
// temporal softmaxes are sums over all N nodes and reach O(100) with trained weights, where an
// fp32 score carries 1e-5 of relative error straight into the attention matrix.
MSGAT_KERNEL MSGAT_LB(1024, 1) k_small(SmallArgs a) {
  MSGAT_DYN_SMEM(smem_raw);
  const int C = a.C, b = blockIdx.x, m = blockIdx.y;
  const float* const Wc_m = a.prm[m].ca_Wc;
  const float* const caw_m = a.prm[m].ca_w;
  const double* const pqc_m = reinterpret_cast<const double*>(a.pqc + m * a.sc_sm);
  const double* const pK_m = reinterpret_cast<const double*>(a.pK + m * a.sc_sm);
  float* const qc_o = a.qc + m * a.sv_sm;
  float* const Pc_o = a.Pc + m * a.sv_sm;
  float* const Mc_o = a.Mc + m * a.sv_sm;
  float* const K_o = a.K + m * a.sv_sm;
  float* const Pt_o = a.Pt + m * a.sv_sm;
  double* qc = reinterpret_cast<double*>(smem_raw);  // [C][12]
  double* tmp = qc + C * 12;                         // [C][12]
  double* S = tmp + C * 12;                          // [C][C]
  double* K = S + C * C;                             // [12][20]
  double* St = K + 240;                              // [144]
  if (a.Oc) {
    const float* pqf = a.pqc + m * a.sc_sm;
    if (a.l2) {
      // independent loads, serial fp64 adds in tile order; a thread walks its (up to) two outputs together: sixteen loads in flight
      // instead of one L2 round trip per tile and output
      for (int i = threadIdx.x; i < C * 12; i += 2 * blockDim.x) {
        const int i1 = i + blockDim.x;
        const bool two = i1 < C * 12;
        const float* p0 = pqf + (int64_t)b * a.ntile * C * 12 + i;
        const float* p1 = p0 + (two ? blockDim.x : 0);
        double s0 = 0.0, s1 = 0.0;
        MSGAT_UNROLL_N(8)
        for (int tl = 0; tl < a.ntile; ++tl) { s0 += (double)p0[(int64_t)tl * C * 12]; s1 += (double)p1[(int64_t)tl * C * 12]; }
        qc[i] = s0;
        qc_o[(int64_t)b * C * 12 + i] = (float)s0;
        if (two) { qc[i1] = s1; qc_o[(int64_t)b * C * 12 + i1] = (float)s1; }
      }
    } else {
      for (int i = threadIdx.x; i < C * 12; i += blockDim.x) {
        double s = 0.0;
        for (int tl = 0; tl < a.ntile; ++tl) s += pqc_m[((int64_t)b * a.ntile + tl) * C * 12 + i];
        qc[i] = s;
        qc_o[(int64_t)b * C * 12 + i] = (float)s;
      }
    }
  }
  if (a.Ot && !a.l2) {
    for (int i = threadIdx.x; i < 240; i += blockDim.x) {
      double s = 0.0;
      for (int tl = 0; tl < a.ntile; ++tl) s += pK_m[((int64_t)b * a.ntile + tl) * 240 + i];
      K[i] = s;
      K_o[(int64_t)b * 240 + i] = (float)s;
    }
  }
  if (a.Ot && a.l2) {
    // temporal keys K1 = q_t Wt1^T, K2 = q_t Wt2^T over all N nodes (attention.py:65) from the saved query, fp64;
    // two halves of the nodes per output, combined through St (free until the scores are formed)
    const float* qtb = a.qt + m * a.sv_sm + (int64_t)b * a.N * 12;
    // The sample's q_t rows are staged in shared memory AS DOUBLES (converted once; the score matrix S, K, St and the partial
    // area behind them are free until the scores are formed), in chunks of as many nodes as fit.  Thread = (rank index r,
    // node group g): per node one weight W[r][n] (consecutive lanes read consecutive n) and the node's 12 staged values,
    // 12 fp64 FMAs.  The per-group sums meet in the same area.
    const int Rn = C * C + 240 + 144 + 8 * 240;             // doubles behind S (see the launch)
    double* R = S;
    int NG = (int)blockDim.x / 20;
    if (NG > Rn / 240) NG = Rn / 240;
    const int capn = Rn / 12;
    const int r = threadIdx.x / NG, g = threadIdx.x - r * NG;
    const bool on = threadIdx.x < 20 * NG;
    const float* W = on ? ((r < 10) ? a.prm[m].ta_Wt1 + (int64_t)r * a.N : a.prm[m].ta_Wt2 + (int64_t)(r - 10) * a.N) : nullptr;
    double acc[12];
    MSGAT_UNROLL
    for (int t = 0; t < 12; ++t) acc[t] = 0.0;
    for (int base = 0; base < a.N; base += capn) {
      const int cn = a.N - base < capn ? a.N - base : capn;
      __syncthreads();   // the chunk before has been consumed (first trip: qc is written, nobody reads R yet)
      for (int i = threadIdx.x; i < cn * 3; i += blockDim.x) {
        const float4 v = ld4(qtb + (int64_t)base * 12 + i * 4);
        R[i * 4 + 0] = (double)v.x; R[i * 4 + 1] = (double)v.y; R[i * 4 + 2] = (double)v.z; R[i * 4 + 3] = (double)v.w;
      }
      __syncthreads();
      if (on) {
        MSGAT_UNROLL_N(4)
        for (int n = g; n < cn; n += NG) {
          const double w = (double)__ldg(W + base + n);
          const double* q = R + n * 12;
          MSGAT_UNROLL
          for (int t = 0; t < 12; ++t) acc[t] += w * q[t];
        }
      }
    }
    __syncthreads();
    if (on) {
      MSGAT_UNROLL
      for (int t = 0; t < 12; ++t) R[g * 240 + t * 20 + r] = acc[t];
    }
    __syncthreads();
    double ksum = 0.0;
    if (threadIdx.x < 240) for (int k = 0; k < NG; ++k) ksum += R[k * 240 + threadIdx.x];
    __syncthreads();   // K lies inside R
    if (threadIdx.x < 240) {
      K[threadIdx.x] = ksum;
      K_o[(int64_t)b * 240 + threadIdx.x] = (float)ksum;
    }
  }
  __syncthreads();
  if (a.Oc) {
    // S = (qc Wc) qc^T, softmax over the last axis (attention.py:93)
    for (int i = threadIdx.x; i < C * 12; i += blockDim.x) {
      const int c = i / 12, t2 = i - c * 12;
      double s = 0.0;
      MSGAT_UNROLL
      for (int t = 0; t < 12; ++t) s += qc[c * 12 + t] * (double)__ldg(Wc_m + t * 12 + t2);
      tmp[i] = s;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < C * C; i += blockDim.x) {
      const int c = i / C, j = i - c * C;
      double s = 0.0;
      MSGAT_UNROLL
      for (int t = 0; t < 12; ++t) s += tmp[c * 12 + t] * qc[j * 12 + t];
      S[i] = s;
    }
    __syncthreads();
    // four threads per row (columns part, part + 4, ...), combined with two shuffles: the rows of a wide operator were 72
    // serial fp64 exponentials on 72 of the block's threads
    for (int base = 0; base < 4 * C; base += blockDim.x) {   // every thread takes every trip: the shuffles are warp-wide
      const int c4 = base + threadIdx.x, c = c4 >> 2, part = c4 & 3;
      const bool live = c4 < 4 * C;
      double mx = -INFINITY;
      if (live) for (int j = part; j < C; j += 4) mx = fmax(mx, S[c * C + j]);
      mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
      double l = 0.0;
      if (live) for (int j = part; j < C; j += 4) { const double e = exp(S[c * C + j] - mx); S[c * C + j] = e; l += e; }
      l += __shfl_xor_sync(0xffffffffu, l, 1);
      l += __shfl_xor_sync(0xffffffffu, l, 2);
      const double inv = 1.0 / l;
      if (live) for (int j = part; j < C; j += 4) S[c * C + j] *= inv;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < C * C; i += blockDim.x) Pc_o[(int64_t)b * C * C + i] = (float)S[i];
    // fold the 1x1 conv of CACN (msgat.py:94) into the attention: Mc = W1 Pc
    for (int i = threadIdx.x; i < a.Oc * C; i += blockDim.x) {
      const int o = i / C, j = i - o * C;
      double s;
      if (a.cmix) {
        s = 0.0;
        for (int c = 0; c < C; ++c) s += (double)__ldg(caw_m + o * C + c) * S[c * C + j];
      } else {
        s = S[o * C + j];
      }
      Mc_o[(int64_t)b * a.Oc * C + i] = (float)s;
    }
  }
  if (a.Ot) {
    // temporal scores K1 K2^T, softmax over the last axis (attention.py:65)
    for (int i = threadIdx.x; i < 144; i += blockDim.x) {
      const int t = i / 12, j = i - t * 12;
      double s = 0.0;
      for (int r = 0; r < 10; ++r) s += K[t * 20 + r] * K[j * 20 + 10 + r];
      St[i] = s;
    }
    __syncthreads();
    if (threadIdx.x < 12) {
      const int t = threadIdx.x;
      double m = -INFINITY;
      MSGAT_UNROLL
      for (int j = 0; j < 12; ++j) m = fmax(m, St[t * 12 + j]);
      double e[12], l = 0.0;
      MSGAT_UNROLL
      for (int j = 0; j < 12; ++j) { e[j] = exp(St[t * 12 + j] - m); l += e[j]; }
      const double inv = 1.0 / l;
      MSGAT_UNROLL
      for (int j = 0; j < 12; ++j) Pt_o[(int64_t)b * 144 + t * 12 + j] = (float)(e[j] * inv);
    }
  }
}

// ---------------------------------------------------------------------------------------
struct ApplyArgs {
  const float* x; int64_t x_sb, x_sc;
  int C, N, O, Oc, Ot, Os, Cz, s_pre, flags, n_conv, ntile, J;
  int dil[MSGAT_MAX_CONV];
  const float *ln_w, *ln_b, *res_w, *res_b, *ca_b, *sa_W;
  const float* conv_w[MSGAT_MAX_CONV];
  const float* conv_b[MSGAT_MAX_CONV];
  const float *Mc, *Pt, *Z, *pw;
  const int32_t *row_ptr, *col_idx;
  int nnz;
  float* out;
  float* Y1;  // optional [B,Ot,N,12]: output of the first conv of a two-conv chain
};

template <int NT>
MSGAT_KERNEL MSGAT_LB(512, 1) k_apply(ApplyArgs a) {
  constexpr int P = NT * 12, PS = P + 4, PQ = NT * 3;
  MSGAT_DYN_SMEM(smem_raw);
  const int C = a.C, b = blockIdx.y;
  const int WO = (a.O + 3) & ~3, WOc = (a.Oc + 3) & ~3, WOt = (a.Ot + 3) & ~3, WOs = (a.Os + 3) & ~3;
  const bool has_res = a.flags & MSGAT_F_RES, smix = a.flags & MSGAT_F_SMIX;
  // ---- carve shared memory (same order as apply_smem_floats in plan.h)
  float* sp = reinterpret_cast<float*>(smem_raw);
  float* X = sp; sp += C * PS;
  float* H = sp; sp += C * PS;
  float* OT = nullptr; if (a.Ot) { OT = sp; sp += C * PS; }
  float* YA = nullptr; if (a.n_conv >= 2) { YA = sp; sp += a.Ot * PS; }
  float* YB = nullptr; if (a.n_conv >= 3) { YB = sp; sp += a.Ot * PS; }
  float* OS = nullptr; if (a.Os && !a.s_pre) { OS = sp; sp += C * PS; }
  float *resT = nullptr, *resB = nullptr;
  if (has_res) { resT = sp; sp += C * WO; resB = sp; sp += WO; }
  float *McT = nullptr, *caB = nullptr;
  if (a.Oc) { McT = sp; sp += C * WOc; caB = sp; sp += WOc; }
  // conv i: [now taps K x WOt | past taps K x WOt | bias WOt] at cv0 + cv_off(i); offsets from one base
  // instead of pointer arrays, which would live in local memory and make every access through them generic
  float* cv0 = sp;
  for (int i = 0; i < a.n_conv; ++i) sp += 2 * (i == 0 ? C : a.Ot) * WOt + WOt;
  auto cv_off = [&](int i) { return i == 0 ? 0 : (2 * C * WOt + WOt) + (i - 1) * (2 * a.Ot * WOt + WOt); };
  float* WsT = nullptr; if (a.Os && smix && !a.s_pre) { WsT = sp; sp += C * WOs; }
  float* Pt = sp; sp += 144;
  float* LW = sp; sp += 24;
  // ---- stage the weights once per block (the block works on one sample only)
  if (has_res) { stage_wT(resT, a.res_w, a.O, C, WO, 1, 0); stage_vec(resB, a.res_b, a.O, WO); }
  if (a.Oc) {
    stage_wT(McT, a.Mc + (int64_t)b * a.Oc * C, a.Oc, C, WOc, 1, 0);
    stage_vec(caB, (a.flags & MSGAT_F_CMIX) ? a.ca_b : nullptr, a.Oc, WOc);
  }
  for (int i = 0; i < a.n_conv; ++i) {
    const int K = i == 0 ? C : a.Ot;
    float* cvA = cv0 + cv_off(i);
    stage_wT(cvA, a.conv_w[i], a.Ot, K, WOt, 2, 1);              // tap 1 multiplies x[t]
    stage_wT(cvA + K * WOt, a.conv_w[i], a.Ot, K, WOt, 2, 0);    // tap 0 multiplies x[t-d]
    stage_vec(cvA + 2 * K * WOt, a.conv_b[i], a.Ot, WOt);
  }
  if (WsT) stage_wT(WsT, a.sa_W, a.Os, C, WOs, 1, 0);
  if (a.Ot) for (int i = threadIdx.x; i < 144; i += blockDim.x) Pt[i] = a.Pt[(int64_t)b * 144 + i];
  const bool do_ln = a.flags & MSGAT_F_LN;
  if (threadIdx.x < 12) {
    LW[threadIdx.x] = do_ln ? a.ln_w[threadIdx.x] : 1.f;
    LW[12 + threadIdx.x] = do_ln ? a.ln_b[threadIdx.x] : 0.f;
  }
  const int chC = (a.Oc + 3) >> 2, chT = (a.Ot + 3) >> 2, chS = (a.Os + 3) >> 2;
  const float* pwb = a.pw + (int64_t)b * a.nnz;
  const float* Zb = a.Z + (int64_t)b * a.Cz * a.N * 12;

  for (int tile = blockIdx.x; tile < a.ntile; tile += a.J) {
    const int n0 = tile * NT;
    __syncthreads();  // previous tile fully consumed (also orders the weight staging)
    tile_load_ln(a.x + b * a.x_sb, a.x_sc, C, a.N, n0, NT, PS, X, H, LW, LW + 12, do_ln, nullptr);
    __syncthreads();
    // temporal attention apply (attention.py:66): OT[c][n][t] = sum_i Pt[t][i] H[c][n][i]
    if (a.Ot) tile_temporal(OT, H, PS, C, NT, Pt, false, false);
    // spatial aggregation of un-mixed rows into OS (few channels): sum over adjacency nnz
    if (OS) {
      for (int it = threadIdx.x; it < C * PQ; it += blockDim.x) {
        const int c = it / PQ, pq = it - c * PQ, nl = pq / 3, t0 = (pq - nl * 3) * 4, n = n0 + nl;
        float4 acc = f4zero();
        if (n < a.N) {
          const int e1 = a.row_ptr[n + 1];
          for (int e = a.row_ptr[n]; e < e1; ++e) {
            const float w = pwb[e];
            fma4(acc, w, ld4(Zb + ((int64_t)c * a.N + a.col_idx[e]) * 12 + t0));
          }
        }
        st4(OS + c * PS + nl * 12 + t0, acc);
      }
    }
    __syncthreads();
    // intermediate convs of the TACN chain (msgat.py:68-74): no activation in between
    const float* cin = OT;
    for (int i = 0; i + 1 < a.n_conv; ++i) {
      float* dst = (i & 1) ? YB : YA;
      const int K = i == 0 ? C : a.Ot;
      const float* cvA = cv0 + cv_off(i);
      for (int it = threadIdx.x; it < chT * PQ; it += blockDim.x) {
        const int ch = it / PQ, pq = it - ch * PQ, nl = pq / 3, t0 = (pq - nl * 3) * 4, o0 = ch * 4;
        float4 acc[4];
        MSGAT_UNROLL
        for (int j = 0; j < 4; ++j) { float bj = cvA[2 * K * WOt + o0 + j]; acc[j] = make_float4(bj, bj, bj, bj); }
        mix_item_taps(acc, cin, PS, K, cvA, cvA + K * WOt, WOt, o0, nl, t0, a.dil[i]);
        MSGAT_UNROLL
        for (int j = 0; j < 4; ++j)
          if (o0 + j < a.Ot) st4(dst + (o0 + j) * PS + nl * 12 + t0, acc[j]);
        if (a.Y1 && i == 0 && n0 + nl < a.N)
          MSGAT_UNROLL
          for (int j = 0; j < 4; ++j)
            if (o0 + j < a.Ot) st4(a.Y1 + (((int64_t)b * a.Ot + o0 + j) * a.N + n0 + nl) * 12 + t0, acc[j]);
      }
      __syncthreads();
      cin = dst;
    }
    // output items: 4 output channels x 4 time steps of one node each
    for (int it = threadIdx.x; it < (chC + chT + chS) * PQ; it += blockDim.x) {
      const int chg = it / PQ, pq = it - chg * PQ, nl = pq / 3, t0 = (pq - nl * 3) * 4, n = n0 + nl;
      float4 acc[4] = {f4zero(), f4zero(), f4zero(), f4zero()};
      int obase, width, o0;
      if (chg < chC) {  // channel branch: Mc[b] h + bias  (attention.py:94 + msgat.py:94)
        o0 = chg * 4; obase = 0; width = a.Oc;
        MSGAT_UNROLL
        for (int j = 0; j < 4; ++j) { float bj = caB[o0 + j]; acc[j] = make_float4(bj, bj, bj, bj); }
        mix_item(acc, H, PS, C, McT, WOc, o0, nl, t0);
      } else if (chg < chC + chT) {  // temporal branch: last conv of the chain (or OT itself)
        o0 = (chg - chC) * 4; obase = a.Oc; width = a.Ot;
        if (a.n_conv == 0) {
          MSGAT_UNROLL
          for (int j = 0; j < 4; ++j) if (o0 + j < width) acc[j] = ld4(OT + (o0 + j) * PS + nl * 12 + t0);
        } else {
          const int i = a.n_conv - 1, K = i == 0 ? C : a.Ot;
          const float* cvA = cv0 + cv_off(i);
          MSGAT_UNROLL
          for (int j = 0; j < 4; ++j) { float bj = cvA[2 * K * WOt + o0 + j]; acc[j] = make_float4(bj, bj, bj, bj); }
          mix_item_taps(acc, cin, PS, K, cvA, cvA + K * WOt, WOt, o0, nl, t0, a.dil[i]);
        }
      } else {  // spatial branch (attention.py:36 + msgat.py:27)
        o0 = (chg - chC - chT) * 4; obase = a.Oc + a.Ot; width = a.Os;
        if (a.s_pre) {
          if (n < a.N) {
            const int e1 = a.row_ptr[n + 1];
            for (int e = a.row_ptr[n]; e < e1; ++e) {
              const float w = pwb[e];
              const float* zr = Zb + (int64_t)a.col_idx[e] * 12 + t0;
              MSGAT_UNROLL
              for (int j = 0; j < 4; ++j)
                if (o0 + j < width) fma4(acc[j], w, ld4(zr + (int64_t)(o0 + j) * a.N * 12));
            }
          }
        } else if (smix) {
          mix_item(acc, OS, PS, C, WsT, WOs, o0, nl, t0);
        } else {
          MSGAT_UNROLL
          for (int j = 0; j < 4; ++j) if (o0 + j < width) acc[j] = ld4(OS + (o0 + j) * PS + nl * 12 + t0);
        }
      }
      if (has_res) {
        float4 r[4];
        MSGAT_UNROLL
        for (int j = 0; j < 4; ++j) { float bj = (o0 + j < width) ? resB[obase + o0 + j] : 0.f; r[j] = make_float4(bj, bj, bj, bj); }
        // residual rows obase+o0.. may be unaligned to 4 in resT when widths are not
        // multiples of 4 (bare attentions never have a residual), so index per row
        const float* s = X + nl * 12 + t0;
        for (int k = 0; k < C; ++k) {
          float4 xv = ld4(s + k * PS);
          const float* w = resT + k * WO + obase + o0;
          MSGAT_UNROLL
          for (int j = 0; j < 4; ++j) if (o0 + j < width) fma4(r[j], w[j], xv);
        }
        MSGAT_UNROLL
        for (int j = 0; j < 4; ++j) { acc[j].x += r[j].x; acc[j].y += r[j].y; acc[j].z += r[j].z; acc[j].w += r[j].w; }
      }
      if (n < a.N) {
        MSGAT_UNROLL
        for (int j = 0; j < 4; ++j) {
          if (o0 + j >= width) continue;
          float4 v = acc[j];
          if (a.flags & MSGAT_F_RELU) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
          st4(a.out + (((int64_t)b * a.O + obase + o0 + j) * a.N + n) * 12 + t0, v);
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------
static int pick_threads(int items, int cap) {
  int rounds = (items + cap - 1) / cap;
  int t = (items + rounds - 1) / rounds;
  t = (t + 31) & ~31;
  return t < 64 ? 64 : t;
}

static void apply_launch(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, const float* x, float* out,
                         float* saved, void* stream) {
  {
    ApplyArgs a;
    a.x = x; a.x_sb = p.x_sb; a.x_sc = p.x_sc;
    a.C = p.C; a.N = p.N; a.O = p.O; a.Oc = p.Oc; a.Ot = p.Ot; a.Os = p.Os; a.Cz = p.Cz; a.s_pre = p.s_pre;
    a.flags = p.flags; a.n_conv = p.n_conv; a.ntile = p.ntile_a;
    for (int i = 0; i < MSGAT_MAX_CONV; ++i) { a.dil[i] = p.dil[i]; a.conv_w[i] = prm->conv_w[i]; a.conv_b[i] = prm->conv_b[i]; }
    a.ln_w = prm->ln_w; a.ln_b = prm->ln_b; a.res_w = prm->res_w; a.res_b = prm->res_b; a.ca_b = prm->ca_b;
    a.sa_W = prm->sa_W;
    a.Mc = saved + p.o_Mc; a.Pt = saved + p.o_Pt; a.Z = saved + p.o_Z; a.pw = saved + p.o_pw;
    a.row_ptr = adj ? adj->row_ptr : nullptr; a.col_idx = adj ? adj->col_idx : nullptr; a.nnz = p.nnz;
    a.out = out;
    a.Y1 = (p.n_conv == 2) ? saved + p.o_Y1 : nullptr;
    const int chunks = ((p.Oc + 3) >> 2) + ((p.Ot + 3) >> 2) + ((p.Os + 3) >> 2);
    size_t smem = (size_t)apply_smem_floats(p, p.NTa) * 4;
    // Narrow operators (layer-1 MEAMs: a tile needs < 48 KB) run as 128-thread blocks, four or more per SM: the
    // kernel is a sequence of barrier-separated phases and independent blocks hide each other's barriers and
    // loads (measured at C = 1: 123 us with 448 threads x 1 block per SM, 98 us with 128 x 4).  Wide ones
    // are limited to one block per SM by their tile and keep up to 512 threads.  MSGAT_APPLY_CAP overrides.
    static int cap_env = -1;
    if (cap_env < 0) { const char* e = getenv("MSGAT_APPLY_CAP"); cap_env = e ? atoi(e) : 0; }
    const int cap = cap_env >= 64 ? cap_env : (smem <= 48 * 1024 ? 128 : 512);
    const int nthr = pick_threads(chunks * p.NTa * 3, cap);
    int occ = (int)(227 * 1024 / (smem + 1024));
    if (occ > 2048 / nthr) occ = 2048 / nthr;
    if (occ > 65536 / (nthr * 128)) occ = 65536 / (nthr * 128);
    if (occ < 1) occ = 1;
    int J = (2 * msgat_sm_count() * occ) / p.B;  // about two full waves of resident blocks
    if (J > p.ntile_a) J = p.ntile_a;
    if (J < 1) J = 1;
    a.J = J;
    switch (p.NTa) {
      case 16: MSGAT_LAUNCH(k_apply<16>, dim3(J, p.B), dim3(nthr), smem, stream, a); break;
      case 8: MSGAT_LAUNCH(k_apply<8>, dim3(J, p.B), dim3(nthr), smem, stream, a); break;
      case 4: MSGAT_LAUNCH(k_apply<4>, dim3(J, p.B), dim3(nthr), smem, stream, a); break;
      default: MSGAT_LAUNCH(k_apply<2>, dim3(J, p.B), dim3(nthr), smem, stream, a); break;
    }
  }
}

int meam_forward_impl(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, const float* x, float* out,
                      float* saved, float* scratch, void* stream) {
  const bool do_ln = p.flags & MSGAT_F_LN;
  const int M = p.M;
  const int64_t sv_sm = p.saved_total, sc_sm = p.scratch_sm, out_sm = (int64_t)p.B * p.O * p.N * 12;
  const bool l2 = l2_fwd_supported(p);
  if (l2) {
    int rc = launch_l2_mix(p, prm, x, saved, scratch, stream);
    if (rc) return rc;
  } else {   // one launch for all components
    PrepArgs a;
    a.x = x; a.x_sb = p.x_sb; a.x_sc = p.x_sc; a.x_sm = p.x_sm; a.sv_sm = sv_sm; a.sc_sm = sc_sm;
    a.C = p.C; a.N = p.N; a.Oc = p.Oc; a.Ot = p.Ot; a.Os = p.Os; a.Cz = p.Cz; a.s_pre = p.s_pre; a.do_ln = do_ln;
    a.ntile = p.ntile_p;
    for (int m = 0; m < M; ++m) a.prm[m] = prm[m];
    a.qs = saved + p.o_qs; a.qt = saved + p.o_qt; a.Z = saved + p.o_Z;
    a.pqc = scratch + p.o_pqc; a.pK = scratch + p.o_pK;
    const int NT = p.NTp;
    size_t smem = ((size_t)p.C * (NT * 12 + 4) + NT * 12 + 24 + (p.s_pre ? (size_t)p.C * rup4(p.Os) : 0)) * 4 +
                  (20 * (size_t)(NT + 1) + NT * 12) * 8;
    if (NT == 64) MSGAT_LAUNCH(k_prep<64>, dim3(p.ntile_p, p.B, M), dim3(256), smem, stream, a);
    else MSGAT_LAUNCH(k_prep<16>, dim3(p.ntile_p, p.B, M), dim3(256), smem, stream, a);
  }
  // the per-sample softmaxes (latency-bound, one block per sample) run on the side stream next to the dense score pass
  void* side = (p.Os && (p.Oc || p.Ot)) ? msgat_fork(stream) : stream;
  if (p.Oc || p.Ot) {
    SmallArgs a;
    a.C = p.C; a.Oc = p.Oc; a.Ot = p.Ot; a.cmix = (p.flags & MSGAT_F_CMIX) ? 1 : 0; a.ntile = l2 ? (p.N + 7) / 8 : p.ntile_p;
    a.l2 = l2 ? 1 : 0; a.N = p.N; a.qt = saved + p.o_qt;
    a.sv_sm = sv_sm; a.sc_sm = sc_sm;
    a.pqc = scratch + p.o_pqc; a.pK = scratch + p.o_pK;
    for (int m = 0; m < M; ++m) a.prm[m] = prm[m];
    a.qc = saved + p.o_qc; a.Pc = saved + p.o_Pc; a.Mc = saved + p.o_Mc; a.K = saved + p.o_K; a.Pt = saved + p.o_Pt;
    size_t smem = ((size_t)p.C * 24 + (size_t)p.C * p.C + 240 + 144 + (l2 ? 8 * 240 : 0)) * 8;
    const int smT = p.C >= 16 ? 512 : 256;   // one block per sample: the C x C phases of a wide operator use 512 threads (81 -> 69 us, 57 -> 46 us at C = 72)
    MSGAT_LAUNCH(k_small, dim3(p.B, M), dim3(smT), smem, side, a);
  }
  if (p.Os) launch_sstats(p, prm, adj, saved, stream);
  msgat_join(stream, side);
  if (l1_fwd_supported(p)) return launch_l1_fwd(p, prm, adj, x, out, saved, stream);   // first layer: one grouped launch
  if (l2 && !getenv("MSGAT_L2_OLD_APPLY")) return launch_l2_out(p, prm, adj, x, out, saved, stream);   // layer 2: one grouped launch
  // the apply pass: still one launch per component (the component's buffers by pointer offset)
  for (int m = 0; m < M; ++m) {
    const float* xm = x + m * p.x_sm;
    float* outm = out + m * out_sm;
    float* savedm = saved + m * sv_sm;
    const msgat_meam_params* prmm = prm + m;
#ifndef MSGAT_EMU
    if (tc_apply_supported(p)) { int rc = launch_tc_apply(p, prmm, adj, xm, outm, savedm, stream); if (rc) return rc; continue; }
    if (tc_apply3_supported(p)) { int rc = launch_tc_apply3(p, prmm, adj, xm, outm, savedm, stream); if (rc) return rc; continue; }
#endif
    apply_launch(p, prmm, adj, xm, outm, savedm, stream);
  }
  return 0;
}
