// meam_bwd.cu -- backward of the MEAM-shaped operator.  Nothing but x, the output (for the
// ReLU mask) and per-sample / per-node statistics was saved by the forward: LayerNorm, the
// temporal apply and the conv chain are recomputed per tile, the N x N attention is
// re-evaluated from (q, m, l) (north_star d: "recompute the scores instead of storing them").
//
// Launch sequence (see meam_backward_impl at the bottom):
//   k_relu_mask  G = dout * (out > 0)                      (skipped when the consumer pre-masked)
//   k_bwd_main   per tile: all data gradients that are local to a node, all weight gradients
//                (register-resident 4x4 blocks, one partial per thread block), the per-sample
//                reductions dMc, dPt and the sparse spatial terms dA, delta
//   k_scol       dense column pass of the spatial softmax backward (spatial.cu)
//   k_sfin       spatial: dU, dq_s, dWg partials
//   k_bwd_small  per sample: softmax backward of the C x C and 12 x 12 attentions
//   k_dwt        dWt1, dWt2
//   k_bwd_fix    second pass over x: the rank-structured terms alpha (x) dq that only exist once
//                the per-sample reductions are complete, pushed through the LayerNorm backward
//                (linear in its incoming gradient) and added to dx; alpha gradients
//   k_reduce     sums every partial into the caller's gradient tensors
#include <stdlib.h>
#include "kernels.h"
#include "tile_ops.cuh"

// Weight gradients: every 4x4 block of every weight matrix is owned by one (thread, register slot);
// when there are fewer blocks than slots each block is split into S position slices.

struct WJob {
  int start, nblk, nkb, Oa, Kb;
  float* out;          // partial base of block 0
  int64_t blk_stride;  // floats between partials of consecutive thread blocks
};
enum { JOB_RES = 0, JOB_MC = 1, JOB_WS = 2, JOB_CONV = 3, NJOBS = 3 + 2 * MSGAT_MAX_CONV };

struct BwdArgs {
  const float* x; int64_t x_sb, x_sc;
  const float* G;  // masked dout [B,O,N,12]
  int C, N, O, Oc, Ot, Os, Cz, s_pre, flags, n_conv, ntile, J, nnz;
  int dil[MSGAT_MAX_CONV];
  const float *ln_w, *ln_b, *res_w, *sa_W;
  const float* conv_w[MSGAT_MAX_CONV];
  const float* conv_b[MSGAT_MAX_CONV];
  const float *Mc, *Pt, *Z, *pw;
  const int32_t *row_ptr, *col_idx, *csc_ptr, *csc_row, *csc_eid;
  WJob jobs[NJOBS];
  float *bias_part, *ln_part, *delta, *dSnz, *dx;
  double* dPt_part;
  int nbias, arena_floats, S;
  float* tcA;  // [B,3b,N,12] dA_now | dA_past | GZ for the tensor-core weight-gradient kernel (null: not used)
  int skip_wgrad;  // weight gradients come from the tensor-core kernel (tc_wgrad.cu) instead
  int resident;    // the backward-chain weights are staged once per CTA, not once per tile and phase
};

// acc[r] (+)= A[oa..oa+3][p] . B[kb..kb+3][p] over the tile's positions, for the blocks of job jb
template <int RBT>
MSGAT_DEV void wgrad_accum(float (&acc)[RBT][16], const WJob& jb, const float* A, const float* Bm, int PS, int PQ, int tapd, int S) {
  MSGAT_UNROLL
  for (int r = 0; r < RBT; ++r) {
    const int vloc = (int)threadIdx.x + r * (int)blockDim.x - jb.start;
    if (vloc < 0 || vloc >= jb.nblk * S) continue;
    const int slice = vloc / jb.nblk, local = vloc - slice * jb.nblk;
    const int oa = (local / jb.nkb) * 4, kb = (local % jb.nkb) * 4;
    const float* ar[4];
    const float* br[4];
    MSGAT_UNROLL
    for (int i = 0; i < 4; ++i) {
      ar[i] = A + (oa + i < jb.Oa ? oa + i : jb.Oa - 1) * PS;
      br[i] = Bm + (kb + i < jb.Kb ? kb + i : jb.Kb - 1) * PS;
    }
    float tl[16];  // tile-local sums first, then one add into the long-running accumulator
    MSGAT_UNROLL
    for (int i = 0; i < 16; ++i) tl[i] = 0.f;
    for (int pq = slice; pq < PQ; pq += S) {
      const int nl = pq / 3, t0 = (pq - nl * 3) * 4, off = nl * 12;
      float4 a[4], b[4];
      MSGAT_UNROLL
      for (int i = 0; i < 4; ++i) a[i] = ld4(ar[i] + off + t0);
      MSGAT_UNROLL
      for (int i = 0; i < 4; ++i) b[i] = (tapd < 0) ? ld4(br[i] + off + t0) : quad_past(br[i] + off, t0, tapd);
      MSGAT_UNROLL
      for (int i = 0; i < 4; ++i) {
        MSGAT_UNROLL
        for (int j = 0; j < 4; ++j) tl[i * 4 + j] += dot4(a[i], b[j]);
      }
    }
    MSGAT_UNROLL
    for (int i = 0; i < 16; ++i) acc[r][i] += tl[i];
  }
}

template <int RBT>
MSGAT_DEV void wgrad_flush(const float (&acc)[RBT][16], const WJob* jobs, int blk, int S) {
  for (int r = 0; r < RBT; ++r) {
    const int id = (int)threadIdx.x + r * (int)blockDim.x;
    for (int q = 0; q < NJOBS; ++q) {
      const WJob& jb = jobs[q];
      const int vloc = id - jb.start;
      if (jb.nblk == 0 || vloc < 0 || vloc >= jb.nblk * S) continue;
      const int slice = vloc / jb.nblk, local = vloc - slice * jb.nblk;
      const int oa = (local / jb.nkb) * 4, kb = (local % jb.nkb) * 4;
      float* dst = jb.out + ((int64_t)blk * S + slice) * jb.blk_stride;
      MSGAT_UNROLL
      for (int i = 0; i < 4; ++i)
        MSGAT_UNROLL
        for (int j = 0; j < 4; ++j)
          if (oa + i < jb.Oa && kb + j < jb.Kb) dst[(oa + i) * jb.Kb + kb + j] = acc[r][i * 4 + j];
    }
  }
}

MSGAT_KERNEL k_relu_mask(const float4* dout, const float4* out, float4* G, int64_t n4) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float4 d = dout[i], o = out[i];
    d.x = o.x > 0.f ? d.x : 0.f; d.y = o.y > 0.f ? d.y : 0.f; d.z = o.z > 0.f ? d.z : 0.f; d.w = o.w > 0.f ? d.w : 0.f;
    G[i] = d;
  }
}

// ---------------------------------------------------------------------------------------
// shared memory (floats) of k_bwd_main; mirrored by the carve-up inside the kernel
static int bwd_arena_floats(const Plan& p, int nthreads) {
  int a = 0, v;
  if (p.flags & MSGAT_F_RES) { v = p.O * rup4(p.C); a = v > a ? v : a; }
  if (p.Oc) { v = p.Oc * rup4(p.C); a = v > a ? v : a; }
  for (int i = 0; i < p.n_conv; ++i) {
    const int K = i == 0 ? p.C : p.Ot;
    v = 2 * K * rup4(p.Ot) + rup4(p.Ot); a = v > a ? v : a;   // transposed taps + bias (recompute)
    v = 2 * p.Ot * rup4(K); a = v > a ? v : a;                // untransposed taps (conv^T)
  }
  if (p.Os && (p.flags & MSGAT_F_SMIX)) { v = p.Os * rup4(p.C); a = v > a ? v : a; }
  if (p.Ot) { v = (nthreads / 9) * 144; a = v > a ? v : a; }
  v = (nthreads / 32) * 24; a = v > a ? v : a;
  return rup4(a);
}
// Backward-chain weights small enough to stay in shared memory for the whole kernel (layer-1 MEAMs):
// Mc, both taps of every conv, Ws, Wres, in the [rows][rup4(K)] layout of stage_w.  0 = restage per tile.
static int bwd_resident_floats(const Plan& p) {
  const int RC = rup4(p.C);
  int f = 0;
  if (p.Oc) f += p.Oc * RC;
  for (int i = 0; i < p.n_conv; ++i) f += 2 * p.Ot * rup4(i == 0 ? p.C : p.Ot);
  if (p.Os && (p.flags & MSGAT_F_SMIX)) f += p.Os * RC;
  if (p.flags & MSGAT_F_RES) f += p.O * RC;
  return f <= 4096 ? f : 0;
}
static int64_t bwd_smem_floats(const Plan& p, int NT, int nthreads) {
  const int PS = NT * 12 + 4;
  int64_t f = 0;
  f += (int64_t)p.C * PS * 3;  // H, OT, DH
  f += (int64_t)p.O * PS;      // G
  if (p.n_conv >= 2) f += (int64_t)(p.n_conv - 1) * p.Ot * PS + 2 * (int64_t)p.Ot * PS;  // Y[i], DYa, DYb
  if (p.Os) f += (int64_t)p.Os * PS;                                                       // GZ
  if (p.Os && (p.flags & MSGAT_F_SMIX) && !p.s_pre) f += (int64_t)p.C * PS;                // DOS
  f += rup4(p.C * NT * 2) + 144 + 24;
  f += bwd_arena_floats(p, nthreads);
  f += bwd_resident_floats(p);
  return f;
}

template <int NT, int RBT>
MSGAT_KERNEL MSGAT_LB((RBT == 2 ? 512 : 384), 1) k_bwd_main(BwdArgs a) {
  constexpr int P = NT * 12, PS = P + 4, PQ = NT * 3;
  MSGAT_DYN_SMEM(smem_raw);
  const int C = a.C, b = blockIdx.y, T = blockDim.x, tid = threadIdx.x;
  const int RC = (C + 3) & ~3;
  const bool has_res = a.flags & MSGAT_F_RES, smix = a.flags & MSGAT_F_SMIX, do_ln = a.flags & MSGAT_F_LN;
  float* sp = reinterpret_cast<float*>(smem_raw);
  float* H = sp; sp += C * PS;
  float* OT = sp; sp += C * PS;
  float* DH = sp; sp += C * PS;
  float* G = sp; sp += a.O * PS;
  // conv intermediates Y[i] = Yb + i * Ot * PS and the two ping-pong gradient buffers DY[j] = DYb + j * Ot * PS:
  // offsets from one shared-memory base (pointer arrays would live in local memory and turn every access
  // through them into a generic load / store)
  float *Yb = nullptr, *DYb = nullptr;
  if (a.n_conv >= 2) {
    Yb = sp; sp += (a.n_conv - 1) * a.Ot * PS;
    DYb = sp; sp += 2 * a.Ot * PS;
  }
  float* GZ = nullptr; if (a.Os) { GZ = sp; sp += a.Os * PS; }
  float* DOS = nullptr; if (a.Os && smix && !a.s_pre) { DOS = sp; sp += C * PS; }
  float* ST = sp; sp += (C * NT * 2 + 3) & ~3;
  float* Pt = sp; sp += 144;
  float* LW = sp; sp += 24;
  float* AR = sp;  // arena
  float* RW = AR + a.arena_floats;  // resident weights (a.resident)
  float *rMc = nullptr, *rWc = nullptr, *rWs = nullptr, *rWres = nullptr;   // rWc: conv taps, [now | past] per conv
  const int RKt = (a.Ot + 3) & ~3;
  const int Ooff = a.Oc + a.Ot;
  const int chC = (C + 3) >> 2;
  const int blk = b * a.J + blockIdx.x;
  const float* Gb = a.G + (int64_t)b * a.O * a.N * 12;
  const float* pwb = a.pw + (int64_t)b * a.nnz;
  const float* Zb = a.Z + (int64_t)b * a.Cz * a.N * 12;

  if (a.Ot) for (int i = tid; i < 144; i += T) Pt[i] = a.Pt[(int64_t)b * 144 + i];
  if (a.resident) {   // stage every backward-chain weight once instead of once per tile and phase
    float* q = RW;
    if (a.Oc) { rMc = q; q += a.Oc * RC; stage_w(rMc, a.Mc + (int64_t)b * a.Oc * C, a.Oc, C, RC, 1, 0); }
    for (int i = 0; i < a.n_conv; ++i) {
      const int K = i == 0 ? C : a.Ot, RK = (K + 3) & ~3;
      if (i == 0) rWc = q;
      stage_w(q, a.conv_w[i], a.Ot, K, RK, 2, 1);
      stage_w(q + a.Ot * RK, a.conv_w[i], a.Ot, K, RK, 2, 0);
      q += 2 * a.Ot * RK;
    }
    if (a.Os && (a.flags & MSGAT_F_SMIX)) { rWs = q; q += a.Os * RC; stage_w(rWs, a.sa_W, a.Os, C, RC, 1, 0); }
    if (a.flags & MSGAT_F_RES) { rWres = q; stage_w(rWres, a.res_w, a.O, C, RC, 1, 0); }
  }
  if (tid < 12) { LW[tid] = do_ln ? a.ln_w[tid] : 1.f; LW[12 + tid] = do_ln ? a.ln_b[tid] : 0.f; }

  float acc[RBT][16];
  for (int r = 0; r < RBT; ++r) for (int i = 0; i < 16; ++i) acc[r][i] = 0.f;
  float bacc = 0.f, lnw_acc[12], lnb_acc[12];
  double pt_acc = 0.0;
  MSGAT_UNROLL
  for (int t = 0; t < 12; ++t) { lnw_acc[t] = 0.f; lnb_acc[t] = 0.f; }

  for (int tile = blockIdx.x; tile < a.ntile; tile += a.J) {
    const int n0 = tile * NT;
    __syncthreads();
    // ---- S0: x -> OT (raw) and H (LayerNorm, stats kept); masked dout tile -> G
    tile_load_ln(a.x + b * a.x_sb, a.x_sc, C, a.N, n0, NT, PS, OT, H, LW, LW + 12, do_ln, ST);
    for (int r = tid; r < a.O * NT; r += T) {
      const int o = r / NT, nl = r - o * NT, n = n0 + nl;
      float4 u = f4zero(), v = f4zero(), w = f4zero();
      if (n < a.N) { const float* s = Gb + ((int64_t)o * a.N + n) * 12; u = ld4(s); v = ld4(s + 4); w = ld4(s + 8); }
      float* q = G + o * PS + nl * 12;
      st4(q, u); st4(q + 4, v); st4(q + 8, w);
    }
    __syncthreads();
    // ---- S1: residual weight gradient (msgat.py:116,131) and the row sums behind three biases
    if (has_res && !a.skip_wgrad) wgrad_accum<RBT>(acc, a.jobs[JOB_RES], G, OT, PS, PQ, -1, a.S);
    if (tid < a.O) {
      const float* q = G + tid * PS;
      float s = 0.f;
      for (int i = 0; i < P; i += 4) { float4 v = ld4(q + i); s += (v.x + v.y) + (v.z + v.w); }
      bacc += s;
    }
    // ---- S2: channel branch: dh = Mc^T g_c ; dMc = g_c h^T
    if (a.Oc && !a.resident) stage_w(AR, a.Mc + (int64_t)b * a.Oc * C, a.Oc, C, RC, 1, 0);
    __syncthreads();
    for (int it = tid; it < chC * PQ; it += T) {
      const int ch = it / PQ, pq = it - ch * PQ, nl = pq / 3, t0 = (pq - nl * 3) * 4, c0 = ch * 4;
      float4 d[4] = {f4zero(), f4zero(), f4zero(), f4zero()};
      if (a.Oc) mix_item(d, G, PS, a.Oc, a.resident ? rMc : AR, RC, c0, nl, t0);
      MSGAT_UNROLL
      for (int j = 0; j < 4; ++j) if (c0 + j < C) st4(DH + (c0 + j) * PS + nl * 12 + t0, d[j]);
    }
    if (a.Oc && !a.skip_wgrad) wgrad_accum<RBT>(acc, a.jobs[JOB_MC], G, H, PS, PQ, -1, a.S);
    __syncthreads();
    // ---- S3: temporal branch
    if (a.Ot) {
      // forward temporal mix and intermediate conv outputs are inputs of weight gradients only
      if (!a.skip_wgrad) tile_temporal(OT, H, PS, C, NT, Pt, false, false);
      __syncthreads();
      const float* cin = OT;
      for (int i = 0; i + 1 < a.n_conv && !a.skip_wgrad; ++i) {
        const int K = i == 0 ? C : a.Ot, WOt = (a.Ot + 3) & ~3;
        float *cvA = AR, *cvP = AR + K * WOt, *cvB = AR + 2 * K * WOt;
        stage_wT(cvA, a.conv_w[i], a.Ot, K, WOt, 2, 1);
        stage_wT(cvP, a.conv_w[i], a.Ot, K, WOt, 2, 0);
        stage_vec(cvB, a.conv_b[i], a.Ot, WOt);
        __syncthreads();
        for (int it = tid; it < (WOt >> 2) * PQ; it += T) {
          const int ch = it / PQ, pq = it - ch * PQ, nl = pq / 3, t0 = (pq - nl * 3) * 4, o0 = ch * 4;
          float4 d[4];
          MSGAT_UNROLL
          for (int j = 0; j < 4; ++j) { float bj = cvB[o0 + j]; d[j] = make_float4(bj, bj, bj, bj); }
          mix_item_taps(d, cin, PS, K, cvA, cvP, WOt, o0, nl, t0, a.dil[i]);
          MSGAT_UNROLL
          for (int j = 0; j < 4; ++j) if (o0 + j < a.Ot) st4(Yb + i * a.Ot * PS + (o0 + j) * PS + nl * 12 + t0, d[j]);
        }
        __syncthreads();
        cin = Yb + i * a.Ot * PS;
      }
      // walk the chain backwards
      const float* dy = G + a.Oc * PS;
      for (int i = a.n_conv - 1; i >= 0; --i) {
        const int K = i == 0 ? C : a.Ot, RK = (K + 3) & ~3;
        const float* in = i == 0 ? OT : Yb + (i - 1) * a.Ot * PS;
        if (a.tcA && i == 0) {
          // dy = dY1 here.  The first conv sits behind the temporal mix, which commutes with it:
          // y1 = Pt a_now + Pt(t-d1) a_past with a = W1 h, so the weight gradients contract
          // dA_now[k,j] = sum_t Pt[t,j] dY1[k,t] and dA_past[k,j] = sum_{t>=d1} Pt[t-d1,j] dY1[k,t] with h.
          const int d1 = a.dil[0];
          for (int r = tid; r < a.Ot * NT; r += T) {
            const int k = r / NT, nl = r - k * NT, n = n0 + nl;
            if (n >= a.N) continue;
            const float* q = dy + k * PS + nl * 12;
            float v[12], an[12], ap[12];
            MSGAT_UNROLL
            for (int t = 0; t < 12; ++t) v[t] = q[t];
            MSGAT_UNROLL
            for (int j = 0; j < 12; ++j) {
              float s1 = 0.f, s2 = 0.f;
              MSGAT_UNROLL
              for (int t = 0; t < 12; ++t) s1 = fmaf(Pt[t * 12 + j], v[t], s1);
              MSGAT_UNROLL
              for (int t = 0; t < 12; ++t) s2 = t >= d1 ? fmaf(Pt[(t - d1 < 0 ? 0 : t - d1) * 12 + j], v[t], s2) : s2;
              an[j] = s1; ap[j] = s2;
            }
            float* o1 = a.tcA + (((int64_t)b * 3 * a.Ot + k) * a.N + n) * 12;
            float* o2 = a.tcA + (((int64_t)b * 3 * a.Ot + a.Ot + k) * a.N + n) * 12;
            st4(o1, make_float4(an[0], an[1], an[2], an[3])); st4(o1 + 4, make_float4(an[4], an[5], an[6], an[7]));
            st4(o1 + 8, make_float4(an[8], an[9], an[10], an[11]));
            st4(o2, make_float4(ap[0], ap[1], ap[2], ap[3])); st4(o2 + 4, make_float4(ap[4], ap[5], ap[6], ap[7]));
            st4(o2 + 8, make_float4(ap[8], ap[9], ap[10], ap[11]));
          }
        }
        if (!a.skip_wgrad) {
          wgrad_accum<RBT>(acc, a.jobs[JOB_CONV + 2 * i + 0], dy, in, PS, PQ, a.dil[i], a.S);
          wgrad_accum<RBT>(acc, a.jobs[JOB_CONV + 2 * i + 1], dy, in, PS, PQ, -1, a.S);
        }
        if (i + 1 < a.n_conv) {
          const int row = tid - a.O - i * a.Ot;
          if (row >= 0 && row < a.Ot) {
            const float* q = dy + row * PS;
            float s = 0.f;
            for (int k = 0; k < P; k += 4) { float4 v = ld4(q + k); s += (v.x + v.y) + (v.z + v.w); }
            bacc += s;
          }
        }
        // resident taps: conv 0 has RK = rup4(C) columns, the later ones rup4(Ot)
        float* Wn = a.resident ? rWc + (i == 0 ? 0 : 2 * a.Ot * RC + (i - 1) * 2 * a.Ot * RKt) : AR;
        float* Wp = Wn + a.Ot * RK;
        if (!a.resident) {
          stage_w(Wn, a.conv_w[i], a.Ot, K, RK, 2, 1);
          stage_w(Wp, a.conv_w[i], a.Ot, K, RK, 2, 0);
        }
        __syncthreads();
        float* dst = i == 0 ? OT : DYb + (i & 1) * a.Ot * PS;
        for (int it = tid; it < (RK >> 2) * PQ; it += T) {
          const int ch = it / PQ, pq = it - ch * PQ, nl = pq / 3, t0 = (pq - nl * 3) * 4, k0 = ch * 4;
          float4 d[4] = {f4zero(), f4zero(), f4zero(), f4zero()};
          mix_item_taps_T(d, dy, PS, a.Ot, Wn, Wp, RK, k0, nl, t0, a.dil[i]);
          MSGAT_UNROLL
          for (int j = 0; j < 4; ++j) if (k0 + j < K) st4(dst + (k0 + j) * PS + nl * 12 + t0, d[j]);
        }
        __syncthreads();
        dy = dst;
      }
      const float* dOt = dy;  // gradient of the temporal attention output, C rows
      // dPt[t][i] = sum_{c,n} dOt[c,n,t] h[c,n,i]   (attention.py:66 transposed)
      {
        const int ngr = T / 9;
        if (tid < ngr * 9) {
          const int blk9 = tid % 9, grp = tid / 9, tb = (blk9 / 3) * 4, ib = (blk9 % 3) * 4;
          double o16[16];
          MSGAT_UNROLL
          for (int i = 0; i < 16; ++i) o16[i] = 0.0;
          for (int r = grp; r < C * NT; r += ngr) {
            const int c = r / NT, nl = r - c * NT;
            const float4 g = ld4(dOt + c * PS + nl * 12 + tb), h = ld4(H + c * PS + nl * 12 + ib);
            o16[0] += g.x * h.x; o16[1] += g.x * h.y; o16[2] += g.x * h.z; o16[3] += g.x * h.w;
            o16[4] += g.y * h.x; o16[5] += g.y * h.y; o16[6] += g.y * h.z; o16[7] += g.y * h.w;
            o16[8] += g.z * h.x; o16[9] += g.z * h.y; o16[10] += g.z * h.z; o16[11] += g.z * h.w;
            o16[12] += g.w * h.x; o16[13] += g.w * h.y; o16[14] += g.w * h.z; o16[15] += g.w * h.w;
          }
          MSGAT_UNROLL
          for (int i = 0; i < 4; ++i)
            MSGAT_UNROLL
            for (int j = 0; j < 4; ++j) AR[grp * 144 + (tb + i) * 12 + ib + j] = (float)o16[i * 4 + j];
        }
        // dh[c,n,i] += sum_t Pt[t][i] dOt[c,n,t]
        tile_temporal(DH, dOt, PS, C, NT, Pt, true, true);
        __syncthreads();
        if (tid < 144) {
          double s = 0.0;
          for (int g2 = 0; g2 < ngr; ++g2) s += (double)AR[g2 * 144 + tid];
          pt_acc += s;
        }
        __syncthreads();
      }
    }
    // ---- S4: spatial branch
    if (a.Os) {
      const float* Gs = G + Ooff * PS;
      const int RCs = RC;
      const float* Wsm = a.resident ? rWs : AR;
      if (smix && !a.resident) stage_w(AR, a.sa_W, a.Os, C, RCs, 1, 0);
      // GZ[o][i] = sum_n A[n,i] g_s[o][n]  (transpose aggregation through the CSC view)
      const int chS = (a.Os + 3) >> 2;
      for (int it = tid; it < chS * PQ; it += T) {
        const int ch = it / PQ, pq = it - ch * PQ, nl = pq / 3, t0 = (pq - nl * 3) * 4, o0 = ch * 4, n = n0 + nl;
        float4 d[4] = {f4zero(), f4zero(), f4zero(), f4zero()};
        if (n < a.N) {
          const int k1 = a.csc_ptr[n + 1];
          for (int k = a.csc_ptr[n]; k < k1; ++k) {
            const float w = pwb[a.csc_eid[k]];
            const float* gr = Gb + ((int64_t)(Ooff + o0) * a.N + a.csc_row[k]) * 12 + t0;
            MSGAT_UNROLL
            for (int j = 0; j < 4; ++j) if (o0 + j < a.Os) fma4(d[j], w, ld4(gr + (int64_t)j * a.N * 12));
          }
        }
        MSGAT_UNROLL
        for (int j = 0; j < 4; ++j) if (o0 + j < a.Os) st4(GZ + (o0 + j) * PS + nl * 12 + t0, d[j]);
        if (a.tcA && n < a.N)
          MSGAT_UNROLL
          for (int j = 0; j < 4; ++j)
            if (o0 + j < a.Os) st4(a.tcA + (((int64_t)b * 3 * a.Os + 2 * a.Os + o0 + j) * a.N + n) * 12 + t0, d[j]);
      }
      __syncthreads();
      for (int it = tid; it < chC * PQ; it += T) {
        const int ch = it / PQ, pq = it - ch * PQ, nl = pq / 3, t0 = (pq - nl * 3) * 4, c0 = ch * 4;
        float4 d[4];
        MSGAT_UNROLL
        for (int j = 0; j < 4; ++j) d[j] = (c0 + j < C) ? ld4(DH + (c0 + j) * PS + nl * 12 + t0) : f4zero();
        if (smix) {
          mix_item(d, GZ, PS, a.Os, Wsm, RCs, c0, nl, t0);
        } else {
          MSGAT_UNROLL
          for (int j = 0; j < 4; ++j)
            if (c0 + j < C) { float4 z = ld4(GZ + (c0 + j) * PS + nl * 12 + t0); d[j].x += z.x; d[j].y += z.y; d[j].z += z.z; d[j].w += z.w; }
        }
        MSGAT_UNROLL
        for (int j = 0; j < 4; ++j) if (c0 + j < C) st4(DH + (c0 + j) * PS + nl * 12 + t0, d[j]);
        if (DOS) {  // d(aggregated h) = W^T g_s, the left operand of dA when h (not W h) was aggregated
          float4 e[4] = {f4zero(), f4zero(), f4zero(), f4zero()};
          mix_item(e, Gs, PS, a.Os, Wsm, RCs, c0, nl, t0);
          MSGAT_UNROLL
          for (int j = 0; j < 4; ++j) if (c0 + j < C) st4(DOS + (c0 + j) * PS + nl * 12 + t0, e[j]);
        }
      }
      if (smix && !a.skip_wgrad) wgrad_accum<RBT>(acc, a.jobs[JOB_WS], GZ, H, PS, PQ, -1, a.S);
      __syncthreads();
      // dA[n,i] at the adjacency non-zeros, delta[n] = sum_i A[n,i] dA[n,i]; one warp per row
      const float* GA = DOS ? DOS : Gs;
      const int warp = tid >> 5, lane = tid & 31, nwarp = T >> 5;
      for (int nl = warp; nl < NT; nl += nwarp) {
        const int n = n0 + nl;
        if (n >= a.N) continue;
        float delta = 0.f;
        const int e1 = a.row_ptr[n + 1];
        for (int e = a.row_ptr[n]; e < e1; ++e) {
          const float* zr = Zb + (int64_t)a.col_idx[e] * 12;
          float s = 0.f;
          for (int pr = lane; pr < a.Cz * 3; pr += 32) {
            const int cz = pr / 3, q4 = (pr - cz * 3) * 4;
            s += dot4(ld4(GA + cz * PS + nl * 12 + q4), ld4(zr + (int64_t)cz * a.N * 12 + q4));
          }
          s = warp_sum(s);
          const float dsn = pwb[e] * s;
          delta += dsn;
          if (lane == 0) a.dSnz[(int64_t)b * a.nnz + e] = dsn;
        }
        if (lane == 0) a.delta[(int64_t)b * a.N + n] = delta;
      }
    }
    __syncthreads();
    // ---- S5: LayerNorm backward on the rows of DH (in place) + affine gradient partials
    if (do_ln) {
      for (int r = tid; r < C * NT; r += T) {
        const int c = r / NT, nl = r - c * NT, n = n0 + nl;
        if (n >= a.N) continue;
        const float* src = a.x + b * a.x_sb + c * a.x_sc + (int64_t)n * 12;
        const float mu = ST[2 * r], rs = ST[2 * r + 1];
        float xh[12], dh[12];
        float* q = DH + c * PS + nl * 12;
        MSGAT_UNROLL
        for (int t = 0; t < 12; ++t) { xh[t] = (src[t] - mu) * rs; dh[t] = q[t]; }
        float m1 = 0.f, m2 = 0.f;
        MSGAT_UNROLL
        for (int t = 0; t < 12; ++t) {
          lnw_acc[t] += dh[t] * xh[t];
          lnb_acc[t] += dh[t];
          const float gg = dh[t] * LW[t];
          m1 += gg; m2 += gg * xh[t];
        }
        m1 *= (1.0f / 12.0f); m2 *= (1.0f / 12.0f);
        MSGAT_UNROLL
        for (int t = 0; t < 12; ++t) q[t] = rs * (dh[t] * LW[t] - m1 - xh[t] * m2);
      }
    }
    // ---- S6: dx = LN^T(dh) + Wres^T g
    if (has_res && !a.resident) stage_w(AR, a.res_w, a.O, C, RC, 1, 0);
    __syncthreads();
    for (int it = tid; it < chC * PQ; it += T) {
      const int ch = it / PQ, pq = it - ch * PQ, nl = pq / 3, t0 = (pq - nl * 3) * 4, c0 = ch * 4, n = n0 + nl;
      float4 d[4];
      MSGAT_UNROLL
      for (int j = 0; j < 4; ++j) d[j] = (c0 + j < C) ? ld4(DH + (c0 + j) * PS + nl * 12 + t0) : f4zero();
      if (has_res) mix_item(d, G, PS, a.O, a.resident ? rWres : AR, RC, c0, nl, t0);
      if (n < a.N) {
        MSGAT_UNROLL
        for (int j = 0; j < 4; ++j) {
          if (c0 + j >= C) continue;
          float4 v = d[j];
          if (a.flags & MSGAT_F_X_IS_RELU) {
            const float4 xv = ld4(a.x + b * a.x_sb + (c0 + j) * a.x_sc + (int64_t)n * 12 + t0);
            v.x = xv.x > 0.f ? v.x : 0.f; v.y = xv.y > 0.f ? v.y : 0.f; v.z = xv.z > 0.f ? v.z : 0.f; v.w = xv.w > 0.f ? v.w : 0.f;
          }
          st4(a.dx + (((int64_t)b * C + c0 + j) * a.N + n) * 12 + t0, v);
        }
      }
    }
  }
  // ---- flush the per-block partials
  __syncthreads();
  if (!a.skip_wgrad) wgrad_flush<RBT>(acc, a.jobs, blk, a.S);
  if (tid < a.nbias) a.bias_part[(int64_t)blk * a.nbias + tid] = bacc;
  if (a.Ot && tid < 144) a.dPt_part[(int64_t)blk * 144 + tid] = pt_acc;
  {
    const int warp = tid >> 5, lane = tid & 31, nwarp = T >> 5;
    MSGAT_UNROLL
    for (int t = 0; t < 12; ++t) {
      const float w = warp_sum(lnw_acc[t]), bb = warp_sum(lnb_acc[t]);
      if (lane == 0) { AR[warp * 24 + t] = w; AR[warp * 24 + 12 + t] = bb; }
    }
    __syncthreads();
    if (tid < 24) {
      float s = 0.f;
      for (int w = 0; w < nwarp; ++w) s += AR[w * 24 + tid];
      a.ln_part[(int64_t)blk * 24 + tid] = s;
    }
  }
}

// ---------------------------------------------------------------------------------------
// spatial finalisation, one thread per (b, n):
//   dU[n,:] = sum_e dSnz[e] q[i_e,:] - delta[n] E[n,:]
//   dq[n,:] = dq_dense[n,:] + sum_{k in col n} dSnz[eid_k] U[row_k,:] + dU[n,:] Wg^T
//   dWg    += q[n,:]^T dU[n,:]
struct SFinArgs {   // grouped: blockIdx.z = component (saved buffers sv_sm apart, scratch buffers sc_sm apart)
  int N, nnz;
  int64_t sv_sm, sc_sm;
  const float* Wg[MSGAT_MAX_GROUP];
  const float *qs, *delta, *dSnz, *dq_dense;
  const float* E;   // (as doubles)
  const int32_t *row_ptr, *col_idx, *csc_ptr, *csc_row, *csc_eid;
  float *dqs, *Wg_part;
};

MSGAT_KERNEL MSGAT_LB(128, 4) k_sfin(SFinArgs a) {
  MSGAT_DYN_SMEM(smem_raw);
  float* WG = reinterpret_cast<float*>(smem_raw);  // [144]
  float* QD = WG + 144;                            // [128][24]: q, dU of every node of the block
  const int b = blockIdx.y, n = blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t mo = (int64_t)blockIdx.z * a.sv_sm, so = (int64_t)blockIdx.z * a.sc_sm;
  const float* qb = a.qs + mo + (int64_t)b * a.N * 12;
  const float* ds = a.dSnz + so + (int64_t)b * a.nnz;
  const double* Eb = reinterpret_cast<const double*>(a.E + mo);
  const float* Wgm = a.Wg[blockIdx.z];
  for (int i = threadIdx.x; i < 144; i += blockDim.x) WG[i] = Wgm[i];
  __syncthreads();
  float q[12], dq[12];
  double dU[12];  // sum_e dSnz q_i and delta E nearly cancel
  MSGAT_UNROLL
  for (int t = 0; t < 12; ++t) { q[t] = 0.f; dU[t] = 0.0; dq[t] = 0.f; }
  if (n < a.N) {
    const int64_t g = (int64_t)b * a.N + n;
    // the edge ranges first: both edge loops below are chains of dependent gathers (range -> index / weight -> row)
    const int e_lo = a.row_ptr[n], e_hi = a.row_ptr[n + 1], k_lo = a.csc_ptr[n], k_hi = a.csc_ptr[n + 1];
    const double dl = (double)a.delta[so + g];
    {
      const float* qn = qb + (int64_t)n * 12;
      const float* dn = a.dq_dense + so + g * 12;
      const double* En = Eb + g * 12;
      const float4 q0 = ld4(qn), q1 = ld4(qn + 4), q2 = ld4(qn + 8), d0 = ld4(dn), d1 = ld4(dn + 4), d2 = ld4(dn + 8);
      q[0] = q0.x; q[1] = q0.y; q[2] = q0.z; q[3] = q0.w; q[4] = q1.x; q[5] = q1.y; q[6] = q1.z; q[7] = q1.w;
      q[8] = q2.x; q[9] = q2.y; q[10] = q2.z; q[11] = q2.w;
      dq[0] = d0.x; dq[1] = d0.y; dq[2] = d0.z; dq[3] = d0.w; dq[4] = d1.x; dq[5] = d1.y; dq[6] = d1.z; dq[7] = d1.w;
      dq[8] = d2.x; dq[9] = d2.y; dq[10] = d2.z; dq[11] = d2.w;
      MSGAT_UNROLL
      for (int t = 0; t < 12; ++t) dU[t] = -dl * En[t];
    }
    // four edges per step: their indices and weights, then their rows, are requested together (one pair of round trips per
    // four edges instead of per edge); a step's dead slots carry weight 0 and the node's own row
    for (int e = e_lo; e < e_hi; e += 4) {
      float w[4];
      float4 r[4][3];
      MSGAT_UNROLL
      for (int j = 0; j < 4; ++j) {
        const bool ok = e + j < e_hi;
        w[j] = ok ? ds[e + j] : 0.f;
        const float* qi = qb + (int64_t)(ok ? a.col_idx[e + j] : n) * 12;
        r[j][0] = ld4(qi); r[j][1] = ld4(qi + 4); r[j][2] = ld4(qi + 8);
      }
      MSGAT_UNROLL
      for (int j = 0; j < 4; ++j) {
        const double wd = (double)w[j];
        const float qi[12] = {r[j][0].x, r[j][0].y, r[j][0].z, r[j][0].w, r[j][1].x, r[j][1].y, r[j][1].z, r[j][1].w,
                              r[j][2].x, r[j][2].y, r[j][2].z, r[j][2].w};
        MSGAT_UNROLL
        for (int t = 0; t < 12; ++t) dU[t] = fma(wd, (double)qi[t], dU[t]);
      }
    }
    // sum_k dSnz[eid_k] U[row_k,:] with U = q Wg: the weighted rows are summed first, Wg is applied once
    float v[12];
    MSGAT_UNROLL
    for (int t = 0; t < 12; ++t) v[t] = 0.f;
    for (int k = k_lo; k < k_hi; k += 4) {
      float w[4];
      float4 r[4][3];
      MSGAT_UNROLL
      for (int j = 0; j < 4; ++j) {
        const bool ok = k + j < k_hi;
        w[j] = ok ? ds[a.csc_eid[k + j]] : 0.f;
        const float* qr = qb + (int64_t)(ok ? a.csc_row[k + j] : n) * 12;
        r[j][0] = ld4(qr); r[j][1] = ld4(qr + 4); r[j][2] = ld4(qr + 8);
      }
      MSGAT_UNROLL
      for (int j = 0; j < 4; ++j) {
        const float qr[12] = {r[j][0].x, r[j][0].y, r[j][0].z, r[j][0].w, r[j][1].x, r[j][1].y, r[j][1].z, r[j][1].w,
                              r[j][2].x, r[j][2].y, r[j][2].z, r[j][2].w};
        MSGAT_UNROLL
        for (int t = 0; t < 12; ++t) v[t] = fmaf(w[j], qr[t], v[t]);
      }
    }
    MSGAT_UNROLL
    for (int t = 0; t < 12; ++t) {
      float u = 0.f;
      MSGAT_UNROLL
      for (int t1 = 0; t1 < 12; ++t1) u = fmaf(v[t1], WG[t1 * 12 + t], u);
      dq[t] += u;
    }
    MSGAT_UNROLL
    for (int t = 0; t < 12; ++t) {
      double s = 0.0;
      MSGAT_UNROLL
      for (int t2 = 0; t2 < 12; ++t2) s = fma(dU[t2], (double)WG[t * 12 + t2], s);
      a.dqs[so + g * 12 + t] = (float)((double)dq[t] + s);
    }
  }
  MSGAT_UNROLL
  for (int t = 0; t < 12; ++t) { QD[threadIdx.x * 24 + t] = q[t]; QD[threadIdx.x * 24 + 12 + t] = (float)dU[t]; }
  __syncthreads();
  for (int i = threadIdx.x; i < 144; i += blockDim.x) {
    const int t = i / 12, t2 = i - t * 12;
    float s = 0.f;
    for (int k = 0; k < (int)blockDim.x; ++k) s = fmaf(QD[k * 24 + t], QD[k * 24 + 12 + t2], s);
    a.Wg_part[so + ((int64_t)b * gridDim.x + blockIdx.x) * 144 + i] = s;
  }
}

// ---------------------------------------------------------------------------------------
struct BSmallArgs {   // grouped: blockIdx.y = component
  int C, Oc, Ot, cmix, J, Jm;
  int64_t sv_sm, sc_sm;
  const float* Wc[MSGAT_MAX_GROUP];
  const float* ca_w[MSGAT_MAX_GROUP];
  const float *dMc_part, *Pc, *qc, *Pt, *K;
  const float* dPt_part;   // (as doubles)
  float *caw_part, *Wc_part, *dqc, *dK;
};

MSGAT_KERNEL MSGAT_LB(1024, 1) k_bwd_small(BSmallArgs a) {
  MSGAT_DYN_SMEM(smem_raw);
  const int C = a.C, b = blockIdx.x, T = blockDim.x, tid = threadIdx.x;
  const int64_t mo = (int64_t)blockIdx.y * a.sv_sm, so = (int64_t)blockIdx.y * a.sc_sm;
  const float* const Wc_m = a.Wc[blockIdx.y];
  const float* const caw_m = a.ca_w[blockIdx.y];
  const float* const g_qc = a.qc + mo;
  const float* const g_Pc = a.Pc + mo;
  const float* const g_Pt = a.Pt + mo;
  const float* const g_K = a.K + mo;
  const float* const g_dMc = a.dMc_part + so;
  const double* const g_dPt = reinterpret_cast<const double*>(a.dPt_part + so);
  float* const o_caw = a.caw_part + so;
  float* const o_Wc = a.Wc_part + so;
  float* const o_dqc = a.dqc + so;
  float* const o_dK = a.dK + so;
  float* sp = reinterpret_cast<float*>(smem_raw);
  float* qc = sp; sp += C * 12;
  float* A1 = sp; sp += C * 12;
  float* A2 = sp; sp += C * 12;
  float* T2 = sp; sp += C * 12;
  float* Pc = sp; sp += C * C;
  float* dS = sp; sp += C * C;
  float* dMc = sp; sp += a.Oc * C;
  float* TP = sp; sp += 144 * 2 + 240;  // Pt, dSt, K
  if (a.Oc) {
    for (int i = tid; i < C * 12; i += T) qc[i] = g_qc[(int64_t)b * C * 12 + i];
    for (int i = tid; i < C * C; i += T) Pc[i] = g_Pc[(int64_t)b * C * C + i];
    for (int i = tid; i < a.Oc * C; i += T) {
      double s = 0.0;
      MSGAT_UNROLL_N(8)   // the partials of eight blocks requested together (same order of additions)
      for (int j = 0; j < a.Jm; ++j) s += g_dMc[((int64_t)b * a.Jm + j) * a.Oc * C + i];
      dMc[i] = s;
    }
    __syncthreads();
    if (a.cmix) {  // dW1 = dMc Pc^T ; dPc = W1^T dMc   (msgat.py:94)
      for (int i = tid; i < a.Oc * C; i += T) {
        const int o = i / C, c = i - o * C;
        double s = 0.0;
        for (int k = 0; k < C; ++k) s += (double)(dMc[o * C + k]) * (double)(Pc[c * C + k]);
        o_caw[(int64_t)b * a.Oc * C + i] = s;
      }
      for (int i = tid; i < C * C; i += T) {
        const int c = i / C, k = i - c * C;
        double s = 0.0;
        for (int o = 0; o < a.Oc; ++o) s += (double)(__ldg(caw_m + o * C + c)) * (double)(dMc[o * C + k]);
        dS[i] = s;
      }
    } else {
      for (int i = tid; i < C * C; i += T) dS[i] = dMc[i];
    }
    __syncthreads();
    for (int c = tid; c < C; c += T) {  // softmax backward, row-wise
      double s = 0.0;
      for (int k = 0; k < C; ++k) s += (double)(Pc[c * C + k]) * (double)(dS[c * C + k]);
      for (int k = 0; k < C; ++k) dS[c * C + k] = (float)((double)Pc[c * C + k] * ((double)dS[c * C + k] - s));
    }
    for (int i = tid; i < C * 12; i += T) {
      const int c = i / 12, t = i - c * 12;
      double s1 = 0.0, s2 = 0.0;
      MSGAT_UNROLL
      for (int u = 0; u < 12; ++u) {
        s1 += (double)(__ldg(Wc_m + t * 12 + u)) * (double)(qc[c * 12 + u]);   // (qc Wc^T)[c][t]
        s2 += (double)(qc[c * 12 + u]) * (double)(__ldg(Wc_m + u * 12 + t));   // (qc Wc)[c][t]
      }
      A1[i] = s1; A2[i] = s2;
    }
    __syncthreads();
    for (int i = tid; i < C * 12; i += T) {
      const int c = i / 12, t = i - c * 12;
      double s = 0.0, s3 = 0.0;
      for (int k = 0; k < C; ++k) {
        s += (double)(dS[c * C + k]) * (double)(A1[k * 12 + t]);
        s += (double)(dS[k * C + c]) * (double)(A2[k * 12 + t]);
        s3 += (double)(dS[c * C + k]) * (double)(qc[k * 12 + t]);
      }
      o_dqc[(int64_t)b * C * 12 + i] = s;
      T2[i] = s3;
    }
    __syncthreads();
    for (int i = tid; i < 144; i += T) {
      const int t = i / 12, u = i - t * 12;
      double s = 0.0;
      for (int c = 0; c < C; ++c) s += (double)(qc[c * 12 + t]) * (double)(T2[c * 12 + u]);
      o_Wc[(int64_t)b * 144 + i] = s;
    }
  }
  if (a.Ot) {
    float *Pt = TP, *dSt = TP + 144, *K = TP + 288;
    __syncthreads();
    for (int i = tid; i < 144; i += T) {
      Pt[i] = g_Pt[(int64_t)b * 144 + i];
      double s = 0.0;
      MSGAT_UNROLL_N(8)
      for (int j = 0; j < a.J; ++j) s += g_dPt[((int64_t)b * a.J + j) * 144 + i];
      dSt[i] = s;
    }
    for (int i = tid; i < 240; i += T) K[i] = g_K[(int64_t)b * 240 + i];
    __syncthreads();
    if (tid < 12) {
      double s = 0.0;
      MSGAT_UNROLL
      for (int k = 0; k < 12; ++k) s += (double)(Pt[tid * 12 + k]) * (double)(dSt[tid * 12 + k]);
      MSGAT_UNROLL
      for (int k = 0; k < 12; ++k) dSt[tid * 12 + k] = (float)((double)Pt[tid * 12 + k] * ((double)dSt[tid * 12 + k] - s));
    }
    __syncthreads();
    for (int i = tid; i < 240; i += T) {
      const int t = i / 20, r = i - t * 20;
      double s = 0.0;
      if (r < 10) { for (int k = 0; k < 12; ++k) s += (double)(dSt[t * 12 + k]) * (double)(K[k * 20 + 10 + r]); }   // dK1 = dS K2
      else { for (int k = 0; k < 12; ++k) s += (double)(dSt[k * 12 + t]) * (double)(K[k * 20 + r - 10]); }         // dK2 = dS^T K1
      o_dK[(int64_t)b * 240 + i] = s;
    }
  }
}

// dWt1[r][n] = sum_{b,t} dK1[b,t,r] q_t[b,n,t]   (attention.py:65)
struct DwtArgs {   // grouped: blockIdx.y = component
  const float *dK, *qt;
  int64_t sv_sm, sc_sm;
  int B, N;
  float* dWt1[MSGAT_MAX_GROUP];
  float* dWt2[MSGAT_MAX_GROUP];
};
MSGAT_KERNEL k_dwt(DwtArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int B = a.B, N = a.N;
  const float* dK = a.dK + (int64_t)blockIdx.y * a.sc_sm;
  const float* qt = a.qt + (int64_t)blockIdx.y * a.sv_sm;
  float* dWt1 = a.dWt1[blockIdx.y];
  float* dWt2 = a.dWt2[blockIdx.y];
  if (i >= N * 20) return;
  const int n = i / 20, r = i - n * 20;
  float s = 0.f;
  MSGAT_UNROLL_N(4)   // the rows of four samples requested together (same order of additions)
  for (int b = 0; b < B; ++b)
    MSGAT_UNROLL
    for (int t = 0; t < 12; ++t) s = fmaf(dK[(int64_t)b * 240 + t * 20 + r], qt[((int64_t)b * N + n) * 12 + t], s);
  if (r < 10) { if (dWt1) dWt1[(int64_t)r * N + n] = s; }
  else { if (dWt2) dWt2[(int64_t)(r - 10) * N + n] = s; }
}

// ---------------------------------------------------------------------------------------
// second pass: dh2[c,n,t] = alpha_n[n] dqc[b,c,t] + alpha_t[c] dqt[b,t,n] + alpha_s[c] dqs[b,n,t]
// (the gradients of the three query reductions, attention.py:92,63,33), through the
// LayerNorm backward, added to dx; gradients of the three alphas.
struct FixArgs {   // grouped: blockIdx.z = component (x: x_sm, dx: dx_sm, scratch: sc_sm apart)
  const float* x; int64_t x_sb, x_sc, x_sm, dx_sm, sc_sm;
  int C, N, Oc, Ot, Os, flags, ntile, J;
  msgat_meam_params prm[MSGAT_MAX_GROUP];
  const float *dqc, *dK, *dqs;
  float *dx, *pa_n, *ln_part;
  float *pa_t, *pa_s;   // (as doubles)
  // First-layer mode (meam_l1.cu): the main pass left the pre-LayerNorm gradient DH, the residual gradient DXR and
  // d(aggregated h) DA in scratch ([B,C,N,12] each) instead of a partial dx; this kernel adds the transposed sparse
  // aggregation sum_n A[n,i] DA[n] (CSC view) and the alpha terms, runs the LayerNorm backward ONCE and writes dx.
  const float *DH, *DXR, *DA, *pw;
  const int32_t *csc_ptr, *csc_row, *csc_eid;
  int nnz;
  int64_t sv_sm;
};

template <int NT>
MSGAT_KERNEL MSGAT_LB(256, 2) k_bwd_fix(FixArgs a) {
  constexpr int P = NT * 12;
  MSGAT_DYN_SMEM(smem_raw);
  const int C = a.C, b = blockIdx.y, T = blockDim.x, tid = threadIdx.x;
  const bool do_ln = a.flags & MSGAT_F_LN;
  const msgat_meam_params& PM = a.prm[blockIdx.z];
  const int64_t so = (int64_t)blockIdx.z * a.sc_sm;
  const float* const xm = a.x + (int64_t)blockIdx.z * a.x_sm;
  float* const dxm = a.dx + (int64_t)blockIdx.z * a.dx_sm;
  const float* const g_dqc = a.dqc + so;
  const float* const g_dK = a.dK + so;
  const float* const g_dqs = a.dqs + so;
  const float* const g_DH = a.DH ? a.DH + so : nullptr;
  const float* const g_DXR = a.DH ? a.DXR + so : nullptr;
  const float* const g_DA = a.DH ? a.DA + so : nullptr;
  const float* const g_pw = a.DH ? a.pw + (int64_t)blockIdx.z * a.sv_sm + (int64_t)b * a.nnz : nullptr;
  float* sp = reinterpret_cast<float*>(smem_raw);
  float* DQC = sp; sp += C * 12;
  float* DQT = sp; sp += P;
  float* DQS = sp; sp += P;
  float* V1 = sp; sp += C * NT;
  double* V2 = reinterpret_cast<double*>(sp); sp += 2 * C * NT;  // offset so far is even: 8-byte aligned
  double* V3 = reinterpret_cast<double*>(sp); sp += 2 * C * NT;
  float* DK = sp; sp += 240;
  float* LW = sp; sp += 24;
  float* RED = sp;  // [T/32][24]
  if (a.Oc) for (int i = tid; i < C * 12; i += T) DQC[i] = g_dqc[(int64_t)b * C * 12 + i];
  if (a.Ot) for (int i = tid; i < 240; i += T) DK[i] = g_dK[(int64_t)b * 240 + i];
  if (tid < 12) { LW[tid] = do_ln ? PM.ln_w[tid] : 1.f; LW[12 + tid] = do_ln ? PM.ln_b[tid] : 0.f; }
  double at_acc = 0.0, as_acc = 0.0;
  float lnw_acc[12], lnb_acc[12];
  MSGAT_UNROLL
  for (int t = 0; t < 12; ++t) { lnw_acc[t] = 0.f; lnb_acc[t] = 0.f; }
  for (int tile = blockIdx.x; tile < a.ntile; tile += a.J) {
    const int n0 = tile * NT;
    __syncthreads();
    for (int p = tid; p < P; p += T) {
      const int nl = p / 12, t = p - nl * 12, n = n0 + nl;
      float s = 0.f, q = 0.f;
      if (n < a.N) {
        if (a.Ot)
          for (int r = 0; r < 10; ++r) {
            s = fmaf(DK[t * 20 + r], __ldg(PM.ta_Wt1 + (int64_t)r * a.N + n), s);
            s = fmaf(DK[t * 20 + 10 + r], __ldg(PM.ta_Wt2 + (int64_t)r * a.N + n), s);
          }
        if (a.Os) q = g_dqs[((int64_t)b * a.N + n) * 12 + t];
      }
      DQT[p] = s; DQS[p] = q;
    }
    __syncthreads();
    for (int r = tid; r < C * NT; r += T) {
      const int c = r / NT, nl = r - c * NT, n = n0 + nl;
      // the three alpha-gradient dot products of a row: 12 terms each in fp32 FMAs, the ROW sums then go to the fp64 accumulation
      // over rows, tiles and samples (per-term fp64 products cost four fp32 -> fp64 conversions per element on the quarter-rate
      // conversion pipe: 17 % of this kernel's samples; the error of a 12-term fp32 dot is random per row and averages out)
      float v1 = 0.f, v2 = 0.f, v3 = 0.f;
#ifndef MSGAT_EMU
      {   // the rows of the thread's NEXT item on their way into L1 while this one is computed
        const int r2 = r + T, c2 = r2 / NT, n2 = n0 + r2 - c2 * NT;
        if (r2 < C * NT && n2 < a.N) {
          asm volatile("prefetch.global.L1 [%0];" ::"l"(xm + b * a.x_sb + c2 * a.x_sc + (int64_t)n2 * 12));
          if (!g_DH) asm volatile("prefetch.global.L1 [%0];" ::"l"(dxm + (((int64_t)b * C + c2) * a.N + n2) * 12));
        }
      }
#endif
      if (n < a.N) {
        const float* src = xm + b * a.x_sb + c * a.x_sc + (int64_t)n * 12;
        float xv[12], xh[12], h[12], dh[12];
        float4 u = ld4(src), v = ld4(src + 4), w = ld4(src + 8);
        // the row of dx this row's result is added to: requested together with the x row (issued after the arithmetic it was
        // 23 % of the kernel's samples of exposed load latency)
        float* d = dxm + (((int64_t)b * C + c) * a.N + n) * 12;
        float4 d0 = f4zero(), d1 = f4zero(), d2 = f4zero();
        if (!g_DH) { d0 = ld4(d); d1 = ld4(d + 4); d2 = ld4(d + 8); }
        xv[0] = u.x; xv[1] = u.y; xv[2] = u.z; xv[3] = u.w; xv[4] = v.x; xv[5] = v.y; xv[6] = v.z; xv[7] = v.w;
        xv[8] = w.x; xv[9] = w.y; xv[10] = w.z; xv[11] = w.w;
        float mu = 0.f, rs = 1.f;
        if (do_ln) {
          MSGAT_UNROLL
          for (int t = 0; t < 12; ++t) mu += xv[t];
          mu *= (1.0f / 12.0f);
          float var = 0.f;
          MSGAT_UNROLL
          for (int t = 0; t < 12; ++t) { const float d = xv[t] - mu; var = fmaf(d, d, var); }
          rs = rsqrt_ln(var * (1.0f / 12.0f) + LN_EPS);
          MSGAT_UNROLL
          for (int t = 0; t < 12; ++t) { xh[t] = (xv[t] - mu) * rs; h[t] = xh[t] * LW[t] + LW[12 + t]; }
        } else {
          MSGAT_UNROLL
          for (int t = 0; t < 12; ++t) { xh[t] = 0.f; h[t] = xv[t]; }
        }
        const float an = a.Oc ? __ldg(PM.ca_alpha + n) : 0.f;
        const float att = a.Ot ? __ldg(PM.ta_alpha + c) : 0.f;
        const float as = a.Os ? __ldg(PM.sa_alpha + c) : 0.f;
        MSGAT_UNROLL
        for (int t = 0; t < 12; ++t) {
          const float dqc = a.Oc ? DQC[c * 12 + t] : 0.f, dqt = DQT[nl * 12 + t], dqs = DQS[nl * 12 + t];
          dh[t] = an * dqc + att * dqt + as * dqs;
          v1 = fmaf(dqc, h[t], v1); v2 = fmaf(dqt, h[t], v2); v3 = fmaf(dqs, h[t], v3);
        }
        const int64_t rowo = (((int64_t)b * C + c) * a.N + n) * 12;
        if (g_DH) {
          const float4 p0 = ld4(g_DH + rowo), p1 = ld4(g_DH + rowo + 4), p2 = ld4(g_DH + rowo + 8);
          float4 s0 = p0, s1 = p1, s2 = p2;
          const int k1 = a.csc_ptr[n + 1];
          for (int k = a.csc_ptr[n]; k < k1; ++k) {   // dh[c,i] += sum_n A[n,i] DA[c,n]   (attention.py:36 transposed)
            const float w = g_pw[a.csc_eid[k]];
            const float* dr = g_DA + (((int64_t)b * C + c) * a.N + a.csc_row[k]) * 12;
            fma4(s0, w, ld4(dr)); fma4(s1, w, ld4(dr + 4)); fma4(s2, w, ld4(dr + 8));
          }
          dh[0] += s0.x; dh[1] += s0.y; dh[2] += s0.z; dh[3] += s0.w; dh[4] += s1.x; dh[5] += s1.y; dh[6] += s1.z; dh[7] += s1.w;
          dh[8] += s2.x; dh[9] += s2.y; dh[10] += s2.z; dh[11] += s2.w;
        }
        float out[12];
        if (do_ln) {
          float m1 = 0.f, m2 = 0.f;
          MSGAT_UNROLL
          for (int t = 0; t < 12; ++t) {
            lnw_acc[t] += dh[t] * xh[t];
            lnb_acc[t] += dh[t];
            const float gg = dh[t] * LW[t];
            m1 += gg; m2 += gg * xh[t];
          }
          m1 *= (1.0f / 12.0f); m2 *= (1.0f / 12.0f);
          MSGAT_UNROLL
          for (int t = 0; t < 12; ++t) out[t] = rs * (dh[t] * LW[t] - m1 - xh[t] * m2);
        } else {
          MSGAT_UNROLL
          for (int t = 0; t < 12; ++t) out[t] = dh[t];
        }
        if (a.flags & MSGAT_F_X_IS_RELU)
          MSGAT_UNROLL
          for (int t = 0; t < 12; ++t) out[t] = xv[t] > 0.f ? out[t] : 0.f;
        if (g_DH) {   // first-layer mode: dx = LN^T(dh) + Wres^T G, written once (the residual is not behind the ReLU of x)
          d0 = ld4(g_DXR + rowo); d1 = ld4(g_DXR + rowo + 4); d2 = ld4(g_DXR + rowo + 8);
          if (a.flags & MSGAT_F_X_IS_RELU) {
            d0.x = xv[0] > 0.f ? d0.x : 0.f; d0.y = xv[1] > 0.f ? d0.y : 0.f; d0.z = xv[2] > 0.f ? d0.z : 0.f; d0.w = xv[3] > 0.f ? d0.w : 0.f;
            d1.x = xv[4] > 0.f ? d1.x : 0.f; d1.y = xv[5] > 0.f ? d1.y : 0.f; d1.z = xv[6] > 0.f ? d1.z : 0.f; d1.w = xv[7] > 0.f ? d1.w : 0.f;
            d2.x = xv[8] > 0.f ? d2.x : 0.f; d2.y = xv[9] > 0.f ? d2.y : 0.f; d2.z = xv[10] > 0.f ? d2.z : 0.f; d2.w = xv[11] > 0.f ? d2.w : 0.f;
          }
        }
        d0.x += out[0]; d0.y += out[1]; d0.z += out[2]; d0.w += out[3];
        d1.x += out[4]; d1.y += out[5]; d1.z += out[6]; d1.w += out[7];
        d2.x += out[8]; d2.y += out[9]; d2.z += out[10]; d2.w += out[11];
        st4(d, d0); st4(d + 4, d1); st4(d + 8, d2);
      }
      V1[r] = (float)v1; V2[r] = v2; V3[r] = v3;
    }
    __syncthreads();
    if (a.Oc) for (int nl = tid; nl < NT; nl += T) {
      if (n0 + nl < a.N) {
        float s = 0.f;
        for (int c = 0; c < C; ++c) s += V1[c * NT + nl];
        a.pa_n[so + (int64_t)b * a.N + n0 + nl] = s;
      }
    }
    for (int c = tid; c < C; c += T) {  // a thread owns channel c for the whole kernel (needs T >= C)
      double s2 = 0.0, s3 = 0.0;
      for (int nl = 0; nl < NT; ++nl) { s2 += V2[c * NT + nl]; s3 += V3[c * NT + nl]; }
      at_acc += s2; as_acc += s3;
    }
  }
  const int blk = b * a.J + blockIdx.x;
  if (tid < C) {
    reinterpret_cast<double*>(a.pa_t + so)[(int64_t)blk * C + tid] = at_acc;
    reinterpret_cast<double*>(a.pa_s + so)[(int64_t)blk * C + tid] = as_acc;
  }
  __syncthreads();
  const int warp = tid >> 5, lane = tid & 31, nwarp = T >> 5;
  MSGAT_UNROLL
  for (int t = 0; t < 12; ++t) {
    const float w = warp_sum(lnw_acc[t]), bb = warp_sum(lnb_acc[t]);
    if (lane == 0) { RED[warp * 24 + t] = w; RED[warp * 24 + 12 + t] = bb; }
  }
  __syncthreads();
  if (tid < 24) {
    float s = 0.f;
    for (int w = 0; w < nwarp; ++w) s += RED[w * 24 + tid];
    a.ln_part[so + (int64_t)blk * 24 + tid] = s;
  }
}

// ---------------------------------------------------------------------------------------
#include "reduce.cuh"

// ---------------------------------------------------------------------------------------
struct BwdLayout {
  int NT, T, RBT, S, J, nblk, nwblk, nbias, NT4, J4, nblk4, ntile4, nfin;
  int64_t WGT;  // floats of weight-gradient partial per thread block
  int64_t off_job[NJOBS];
  int64_t o_G, o_wg, o_bias, o_ln, o_dMc, o_dPt, o_delta, o_dSnz, o_dqd, o_dqs, o_dK, o_dqc, o_caw, o_Wc, o_Wg,
      o_pat, o_pas, o_pan, o_tcA, o_l1, o_q, total;
  int l1;   // first-layer kernels (meam_l1.cu) serve this shape
  WJob jobs[NJOBS];
};

// blocks per sample so that the grid is about two full waves of `slots` resident thread blocks
static int pick_J(int B, int ntile, int slots) {
  int J = (2 * slots) / B;
  if (J < 1) J = 1;
  return J > ntile ? ntile : J;
}

static int bwd_layout(Plan& p, BwdLayout& L) {
  // weight-gradient jobs as lists of 4x4 blocks
  for (int q = 0; q < NJOBS; ++q) { L.jobs[q] = WJob{0, 0, 1, 1, 1, nullptr, 0}; L.off_job[q] = 0; }
  int64_t woff = 0;
  int nw = 0;
  auto add = [&](int q, int Oa, int Kb, bool in_wg) {
    WJob& jb = L.jobs[q];
    jb.Oa = Oa; jb.Kb = Kb; jb.nkb = (Kb + 3) / 4; jb.nblk = ((Oa + 3) / 4) * jb.nkb;
    nw += jb.nblk;
    if (in_wg) { L.off_job[q] = woff; woff += rup4l((int64_t)Oa * Kb); }
  };
  if (p.flags & MSGAT_F_RES) add(JOB_RES, p.O, p.C, true);
  if (p.Oc) add(JOB_MC, p.Oc, p.C, false);
  if (p.Os && (p.flags & MSGAT_F_SMIX)) add(JOB_WS, p.Os, p.C, true);
  for (int i = 0; i < p.n_conv; ++i) {
    const int K = i == 0 ? p.C : p.Ot;
    add(JOB_CONV + 2 * i, p.Ot, K, true);
    add(JOB_CONV + 2 * i + 1, p.Ot, K, true);
  }
  L.nwblk = nw;
  L.WGT = woff;
  L.nbias = p.O + (p.n_conv > 1 ? (p.n_conv - 1) * p.Ot : 0);
  int T, RBT;
  if (nw <= 1024) { RBT = 2; T = nw > 512 ? 512 : 256; }
  else if (nw <= 1536) { RBT = 4; T = 384; }
  else return -30;
  if (T < ((L.nbias + 31) & ~31)) T = (L.nbias + 31) & ~31;
  if (T < ((p.C + 31) & ~31)) T = (p.C + 31) & ~31;
  if (T > (RBT == 2 ? 512 : 384)) return -30;
  L.T = T; L.RBT = RBT;
  int S = nw ? (RBT * T) / nw : 1;
  if (S < 1) S = 1;
  if (S > 8) S = 8;
  L.S = S;
  int start = 0;
  for (int q = 0; q < NJOBS; ++q) { L.jobs[q].start = start; start += L.jobs[q].nblk * S; }
  int NT = 16;
  while (NT > 4 && bwd_smem_floats(p, NT, T) * 4 > MSGAT_SMEM_BUDGET + 16 * 1024) NT >>= 1;
  if (bwd_smem_floats(p, NT, T) * 4 > MSGAT_SMEM_BUDGET + 16 * 1024) return -31;
  // a narrow kernel (256 threads) is latency-bound at one CTA per SM: halve the tile when that
  // lets a second CTA share the SM
  {
    const char* e = getenv("MSGAT_BWD_NT");
    if (e && atoi(e) >= 4 && atoi(e) <= NT) NT = atoi(e);
    else if (T <= 256 && NT > 4 && (bwd_smem_floats(p, NT, T) * 4 + 1024) * 2 > 227 * 1024 &&
             (bwd_smem_floats(p, NT / 2, T) * 4 + 1024) * 2 <= 227 * 1024) NT >>= 1;
  }
  L.NT = NT; p.NTb = NT; p.ntile_b = (p.N + NT - 1) / NT;
  {
    int64_t smem = bwd_smem_floats(p, NT, T) * 4 + 1024;
    int occ = (int)(227 * 1024 / smem);
    if (occ > 2048 / T) occ = 2048 / T;
    if (occ > 65536 / (T * 128)) occ = 65536 / (T * 128);
    if (occ < 1) occ = 1;
    L.J = pick_J(p.B, p.ntile_b, msgat_sm_count() * occ);
  }
#ifndef MSGAT_EMU
  if (tc_dgrad_supported(p)) L.J = tc_dgrad_blocks_per_sample(p);   // the tensor-core data-gradient kernel writes the per-block partials
#endif
  L.l1 = l1_bwd_supported(p) ? 1 : 0;
  if (L.l1) { L.J = l1_bwd_blocks_per_sample(p); L.S = 1; }
  L.nblk = L.J * p.B;
  // nodes per tile of k_bwd_fix: a tile is C x NT4 rows for 128 threads; a narrow (first-layer) input gets wider tiles so that
  // at least 64 rows are in flight per block (C = 1 with 16-node tiles kept 16 of 128 threads busy: 118 us for 17 MB)
  L.NT4 = 16;
  while (L.NT4 < 64 && p.C * L.NT4 < 64) L.NT4 <<= 1;
  { const char* e = getenv("MSGAT_FIX_NT"); if (e && (atoi(e) == 16 || atoi(e) == 32 || atoi(e) == 64)) L.NT4 = atoi(e); }
  L.ntile4 = (p.N + L.NT4 - 1) / L.NT4;
  L.J4 = pick_J(p.B, L.ntile4, msgat_sm_count() * 4); L.nblk4 = L.J4 * p.B;
  L.nfin = (p.N + 127) / 128;
  int64_t o = 0;
  const int64_t BN = (int64_t)p.B * p.N;
  const bool need_mask = (p.flags & MSGAT_F_RELU) && !(p.flags & MSGAT_F_DOUT_MASKED);
  L.o_G = o; o += need_mask ? rup4l(BN * 12 * p.O) : 0;
  int64_t nwp = (int64_t)L.nblk * L.S;  // weight-gradient partials: k_bwd_main's blocks or the tensor-core kernel's
#ifndef MSGAT_EMU
  { const int64_t t = (int64_t)p.B * tc_wgrad_blocks_per_sample(p); if (t > nwp) nwp = t; }
#endif
  L.o_wg = o; o += rup4l(nwp * L.WGT);
  L.o_bias = o; o += rup4l((int64_t)L.nblk * L.nbias);
  L.o_ln = o; o += rup4l((int64_t)(L.nblk + L.nblk4) * 24);
  L.o_dMc = o; o += rup4l(nwp * p.Oc * p.C);
  L.o_dPt = o; o += rup4l(2 * (int64_t)L.nblk * 144);  // fp64
  L.o_delta = o; o += rup4l(BN);
  L.o_dSnz = o; o += rup4l((int64_t)p.B * p.nnz);
  L.o_dqd = o; o += rup4l(BN * 12);
  L.o_dqs = o; o += rup4l(BN * 12);
  L.o_dK = o; o += rup4l((int64_t)p.B * 240);
  L.o_dqc = o; o += rup4l((int64_t)p.B * p.C * 12);
  L.o_caw = o; o += rup4l((int64_t)p.B * p.Oc * p.C);
  L.o_Wc = o; o += rup4l((int64_t)p.B * 144);
  L.o_Wg = o; o += rup4l((int64_t)p.B * L.nfin * 144);
  L.o_pat = o; o += rup4l(2 * (int64_t)L.nblk4 * p.C);  // fp64 partials of the alpha gradients
  L.o_pas = o; o += rup4l(2 * (int64_t)L.nblk4 * p.C);
  L.o_pan = o; o += rup4l(BN);
  L.o_tcA = o;
#ifndef MSGAT_EMU
  if (tc_wgrad_supported(p) && !L.l1) o += rup4l(BN * 12 * 3 * p.Ot);
#endif
  L.o_l1 = o; if (L.l1) o += 3 * rup4l(BN * 12 * p.C);                       // DH | DXR | DA
  L.o_q = o; if (L.l1) o += rup4l((int64_t)L.nblk * l1_bwd_qsz(p));          // feature-correlation partials
  L.total = o;
  return 0;
}

int64_t meam_backward_scratch(Plan& p) {
  BwdLayout L;
  int rc = bwd_layout(p, L);
  if (rc) return rc;
  return L.total;
}

// The per-sample and reduction kernels behind the main passes, each ONE launch for all components of a grouped call.
static int meam_backward_tail(Plan& p, BwdLayout& L, int Jw, const msgat_meam_params* prm, const msgat_adj* adj, const float* x,
                              const float* saved, float* scratch, float* dx, const msgat_meam_grads* g, void* stream) {
  const int M = p.M;
  const int64_t sv_sm = p.saved_total, sc_sm = p.scratch_sm, dx_sm = (int64_t)p.B * p.C * p.N * 12;
  // the softmax backward of the channel / temporal attention and dWt (per-sample, latency-bound) run on the side stream next to
  // the dense column pass of the spatial attention; both chains meet again in front of k_bwd_fix
  void* side = (p.Os && (p.Oc || p.Ot)) ? msgat_fork(stream) : stream;
  if (p.Os) {
    launch_scol(p, prm->sa_Wg, saved, scratch + L.o_delta, scratch + L.o_dqd, stream);
    SFinArgs a;
    a.N = p.N; a.nnz = p.nnz; a.sv_sm = sv_sm; a.sc_sm = sc_sm; a.qs = saved + p.o_qs; a.E = saved + p.o_E;
    for (int m = 0; m < M; ++m) a.Wg[m] = prm[m].sa_Wg;
    a.delta = scratch + L.o_delta; a.dSnz = scratch + L.o_dSnz; a.dq_dense = scratch + L.o_dqd;
    a.row_ptr = adj->row_ptr; a.col_idx = adj->col_idx; a.csc_ptr = adj->csc_ptr; a.csc_row = adj->csc_row; a.csc_eid = adj->csc_eid;
    a.dqs = scratch + L.o_dqs; a.Wg_part = scratch + L.o_Wg;
    MSGAT_LAUNCH(k_sfin, dim3(L.nfin, p.B, M), dim3(128), (144 + 128 * 24) * 4, stream, a);
  }
  if (p.Oc || p.Ot) {
    BSmallArgs a;
    a.C = p.C; a.Oc = p.Oc; a.Ot = p.Ot; a.cmix = (p.flags & MSGAT_F_CMIX) ? 1 : 0; a.J = L.J; a.Jm = Jw;
    a.sv_sm = sv_sm; a.sc_sm = sc_sm;
    a.dMc_part = scratch + L.o_dMc; a.dPt_part = scratch + L.o_dPt; a.Pc = saved + p.o_Pc; a.qc = saved + p.o_qc;
    for (int m = 0; m < M; ++m) { a.Wc[m] = prm[m].ca_Wc; a.ca_w[m] = prm[m].ca_w; }
    a.Pt = saved + p.o_Pt; a.K = saved + p.o_K;
    a.caw_part = scratch + L.o_caw; a.Wc_part = scratch + L.o_Wc; a.dqc = scratch + L.o_dqc; a.dK = scratch + L.o_dK;
    size_t smem = ((size_t)p.C * 48 + 2 * (size_t)p.C * p.C + (size_t)p.Oc * p.C + 528) * 4;
    const int smT = p.C >= 16 ? 512 : 256;   // one block per sample: the C x C phases of a wide operator use 512 threads (81 -> 69 us, 57 -> 46 us at C = 72)
    MSGAT_LAUNCH(k_bwd_small, dim3(p.B, M), dim3(smT), smem, side, a);
  }
  if (p.Ot && (g->ta_Wt1 || g->ta_Wt2)) {
    DwtArgs a;
    a.dK = scratch + L.o_dK; a.qt = saved + p.o_qt; a.sv_sm = sv_sm; a.sc_sm = sc_sm; a.B = p.B; a.N = p.N;
    for (int m = 0; m < M; ++m) { a.dWt1[m] = g[m].ta_Wt1; a.dWt2[m] = g[m].ta_Wt2; }
    MSGAT_LAUNCH(k_dwt, dim3((p.N * 20 + 255) / 256, M), dim3(256), 0, side, a);
  }
  msgat_join(stream, side);
  {
    FixArgs a;
    a.x = x; a.x_sb = p.x_sb; a.x_sc = p.x_sc; a.x_sm = p.x_sm; a.dx_sm = dx_sm; a.sc_sm = sc_sm;
    a.C = p.C; a.N = p.N; a.Oc = p.Oc; a.Ot = p.Ot; a.Os = p.Os; a.flags = p.flags; a.ntile = L.ntile4; a.J = L.J4;
    for (int m = 0; m < M; ++m) a.prm[m] = prm[m];
    a.dqc = scratch + L.o_dqc; a.dK = scratch + L.o_dK; a.dqs = scratch + L.o_dqs;
    a.dx = dx; a.pa_n = scratch + L.o_pan; a.pa_t = scratch + L.o_pat; a.pa_s = scratch + L.o_pas;
    a.DH = a.DXR = a.DA = a.pw = nullptr; a.csc_ptr = a.csc_row = a.csc_eid = nullptr; a.nnz = p.nnz; a.sv_sm = sv_sm;
    if (L.l1) {
      const int64_t sz = rup4l((int64_t)p.B * p.N * 12 * p.C);
      a.DH = scratch + L.o_l1; a.DXR = scratch + L.o_l1 + sz; a.DA = scratch + L.o_l1 + 2 * sz; a.pw = saved + p.o_pw;
      a.csc_ptr = adj->csc_ptr; a.csc_row = adj->csc_row; a.csc_eid = adj->csc_eid;
    }
    a.ln_part = scratch + L.o_ln + (int64_t)L.nblk * 24;
    size_t smem = ((size_t)p.C * 12 + 2 * 12 * (size_t)L.NT4 + 5 * (size_t)p.C * L.NT4 + 240 + 24 + 8 * 24) * 4;
    // 128-thread blocks, four per SM (a thread owns channel tid across the tiles: at least C threads): more
    // independent blocks hide the barriers between the phases of a tile (C = 1: 35 -> 25 us, C = 72: 118 -> 114 us)
    const int fixT = p.C > 128 ? ((p.C + 31) & ~31) : 128;
    if (L.NT4 == 64) MSGAT_LAUNCH(k_bwd_fix<64>, dim3(L.J4, p.B, M), dim3(fixT), smem, stream, a);
    else if (L.NT4 == 32) MSGAT_LAUNCH(k_bwd_fix<32>, dim3(L.J4, p.B, M), dim3(fixT), smem, stream, a);
    else MSGAT_LAUNCH(k_bwd_fix<16>, dim3(L.J4, p.B, M), dim3(fixT), smem, stream, a);
  }
  {
    ReduceArgs r;
    int nj = 0, tot = 0;
    r.sc_sm = sc_sm;
    // `field`: the gradient tensor of component 0 selects the job; every component's own pointer is taken from g[m]
    auto add = [&](float* msgat_meam_grads::* field, int conv, int which, const float* src, int n, int nparts, int64_t pstride,
                   int dmul, int doff, int f64 = 0) {
      auto get = [&](int m) -> float* {
        if (conv < 0) return g[m].*field;
        return which ? g[m].conv_b[conv] : g[m].conv_w[conv];
      };
      if (!get(0) || n <= 0 || nj >= MAX_RJOBS) return;
      RJob jb;
      for (int m = 0; m < MSGAT_MAX_GROUP; ++m) jb.dst[m] = m < M ? get(m) : nullptr;
      jb.src = src; jb.n = n; jb.nparts = nparts; jb.dmul = dmul; jb.doff = doff; jb.pstride = pstride; jb.f64 = f64; jb.tr = 0;
      r.j[nj] = jb;
      r.starts[nj] = tot; tot += n; ++nj;
    };
    typedef msgat_meam_grads GR;
    const float* wg = scratch + L.o_wg;
    const float* bp = scratch + L.o_bias;
    if ((p.flags & MSGAT_F_RES) && !L.l1) {
      add(&GR::res_w, -1, 0, wg + L.off_job[JOB_RES], p.O * p.C, p.B * Jw, L.WGT, 1, 0);
      add(&GR::res_b, -1, 0, bp, p.O, L.nblk, L.nbias, 1, 0);
    }
    if (p.Os && (p.flags & MSGAT_F_SMIX) && !L.l1) add(&GR::sa_W, -1, 0, wg + L.off_job[JOB_WS], p.Os * p.C, p.B * Jw, L.WGT, 1, 0);
    for (int i = 0; i < p.n_conv && !L.l1; ++i) {
      const int K = i == 0 ? p.C : p.Ot;
      add(nullptr, i, 0, wg + L.off_job[JOB_CONV + 2 * i], p.Ot * K, p.B * Jw, L.WGT, 2, 0);
      add(nullptr, i, 0, wg + L.off_job[JOB_CONV + 2 * i + 1], p.Ot * K, p.B * Jw, L.WGT, 2, 1);
      add(nullptr, i, 1, i + 1 == p.n_conv ? bp + p.Oc : bp + p.O + i * p.Ot, p.Ot, L.nblk, L.nbias, 1, 0);
    }
    if (p.flags & MSGAT_F_LN) {
      // first-layer mode: the LayerNorm backward runs once, in k_bwd_fix, whose partials follow the main pass's slots
      const float* lp = scratch + L.o_ln + (L.l1 ? (int64_t)L.nblk * 24 : 0);
      const int np = L.l1 ? L.nblk4 : L.nblk + L.nblk4;
      add(&GR::ln_w, -1, 0, lp, 12, np, 24, 1, 0);
      add(&GR::ln_b, -1, 0, lp + 12, 12, np, 24, 1, 0);
    }
    if (p.Oc) {
      if (p.flags & MSGAT_F_CMIX) {
        add(&GR::ca_w, -1, 0, scratch + L.o_caw, p.Oc * p.C, p.B, (int64_t)p.Oc * p.C, 1, 0);
        if (!L.l1) add(&GR::ca_b, -1, 0, bp, p.Oc, L.nblk, L.nbias, 1, 0);
      }
      add(&GR::ca_Wc, -1, 0, scratch + L.o_Wc, 144, p.B, 144, 1, 0);
      add(&GR::ca_alpha, -1, 0, scratch + L.o_pan, p.N, p.B, p.N, 1, 0);
    }
    if (p.Ot) add(&GR::ta_alpha, -1, 0, scratch + L.o_pat, p.C, L.nblk4, p.C, 1, 0, 1);
    if (p.Os) {
      add(&GR::sa_alpha, -1, 0, scratch + L.o_pas, p.C, L.nblk4, p.C, 1, 0, 1);
      add(&GR::sa_Wg, -1, 0, scratch + L.o_Wg, 144, p.B * L.nfin, 144, 1, 0);
    }
    r.starts[nj] = tot; r.njobs = nj;
    if (tot > 0) LAUNCH_REDUCE(tot, M, stream, r);
  }
  if (L.l1) launch_l1_wfin(p, prm, g, scratch + L.o_q, L.nblk, stream);
  return 0;
}

int meam_backward_impl(Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, const float* x, const float* out,
                       const float* dout, const float* saved, float* scratch, float* dx, const msgat_meam_grads* g,
                       void* stream) {
  BwdLayout L;
  int rc = bwd_layout(p, L);
  if (rc) { msgat_set_error("MEAM backward: unsupported shape (code %d)", rc); return rc; }
  const int M = p.M;
  const int64_t sv_sm = p.saved_total, sc_sm = p.scratch_sm, out_sm = (int64_t)p.B * p.O * p.N * 12,
                dx_sm = (int64_t)p.B * p.C * p.N * 12;
  const bool need_mask = (p.flags & MSGAT_F_RELU) && !(p.flags & MSGAT_F_DOUT_MASKED);
  bool tc_w = false, tc_d = false;
#ifndef MSGAT_EMU
  tc_w = tc_wgrad_supported(p);
  tc_d = tc_dgrad_supported(p);
#endif
  const int Jw = tc_w ? tc_wgrad_blocks_per_sample(p) : L.J * L.S;  // weight-gradient partials per sample
  if (need_mask && !out) { msgat_set_error("MEAM backward: forward output needed for the ReLU mask"); return -32; }
  if (L.l1) {   // first layer: one grouped streaming pass (meam_l1.cu), masked gradient from dout or from a masking pre-pass
    const float* G = dout;
    int64_t g_sm = out_sm;
    if (need_mask) {
      const int64_t n4 = (int64_t)p.B * p.O * p.N * 3;
      int blocks = (int)((n4 + 255) / 256);
      if (blocks > msgat_sm_count() * 8) blocks = msgat_sm_count() * 8;
      for (int m = 0; m < M; ++m)
        MSGAT_LAUNCH(k_relu_mask, dim3(blocks), dim3(256), 0, stream, (const float4*)(dout + m * out_sm),
                     (const float4*)(out + m * out_sm), (float4*)(scratch + m * sc_sm + L.o_G), n4);
      G = scratch + L.o_G; g_sm = sc_sm;
    }
    const int64_t sz = rup4l((int64_t)p.B * p.N * 12 * p.C);
    rc = launch_l1_bwd(p, prm, adj, x, G, g_sm, saved, scratch + L.o_l1, scratch + L.o_l1 + sz, scratch + L.o_l1 + 2 * sz,
                       scratch + L.o_delta, scratch + L.o_dSnz, scratch + L.o_dMc, scratch + L.o_dPt, scratch + L.o_q, stream);
    if (rc) return rc;
    return meam_backward_tail(p, L, L.J, prm, adj, x, saved, scratch, dx, g, stream);
  }
  if (need_mask) {
    const int64_t n4 = (int64_t)p.B * p.O * p.N * 3;
    int blocks = (int)((n4 + 255) / 256);
    if (blocks > msgat_sm_count() * 8) blocks = msgat_sm_count() * 8;
    for (int m = 0; m < M; ++m)
      MSGAT_LAUNCH(k_relu_mask, dim3(blocks), dim3(256), 0, stream, (const float4*)(dout + m * out_sm), (const float4*)(out + m * out_sm),
                   (float4*)(scratch + m * sc_sm + L.o_G), n4);
  }
#ifndef MSGAT_EMU
  if (tc_d) {   // the sparse stages of all components in one launch (l2_sparse_bwd.cu): GZ -> tcA rows [2b, 3b), dSnz, delta
    rc = launch_l2_gz(p, adj, need_mask ? scratch + L.o_G : dout, need_mask ? sc_sm : out_sm, saved, scratch + L.o_tcA,
                      scratch + L.o_delta, scratch + L.o_dSnz, stream);
    if (rc) return rc;
    // ... and the node-local data gradients of all components in one persistent launch (tc_dgrad.cu)
    rc = launch_tc_dgrad(p, prm, adj, x, need_mask ? scratch + L.o_G : dout, need_mask ? sc_sm : out_sm, saved, scratch + L.o_tcA, dx,
                         dx_sm, scratch + L.o_bias, L.nbias, scratch + L.o_ln, reinterpret_cast<double*>(scratch + L.o_dPt), stream);
    if (rc) return rc;
  }
#endif
  // ---- the remaining passes over the activations: one launch per component (buffers by pointer offset)
  for (int m = 0; m < M; ++m) {
    const float* xm = x + m * p.x_sm;
    const float* savedm = saved + m * sv_sm;
    float* scr = scratch + m * sc_sm;
    float* dxm = dx + m * dx_sm;
    const msgat_meam_params* prmm = prm + m;
    const float* G = need_mask ? scr + L.o_G : dout + m * out_sm;
    if (!tc_d) {
      BwdArgs a;
      a.x = xm; a.x_sb = p.x_sb; a.x_sc = p.x_sc; a.G = G;
      a.C = p.C; a.N = p.N; a.O = p.O; a.Oc = p.Oc; a.Ot = p.Ot; a.Os = p.Os; a.Cz = p.Cz; a.s_pre = p.s_pre;
      a.flags = p.flags; a.n_conv = p.n_conv; a.ntile = p.ntile_b; a.J = L.J; a.nnz = p.nnz;
      for (int i = 0; i < MSGAT_MAX_CONV; ++i) { a.dil[i] = p.dil[i]; a.conv_w[i] = prmm->conv_w[i]; a.conv_b[i] = prmm->conv_b[i]; }
      a.ln_w = prmm->ln_w; a.ln_b = prmm->ln_b; a.res_w = prmm->res_w; a.sa_W = prmm->sa_W;
      a.Mc = savedm + p.o_Mc; a.Pt = savedm + p.o_Pt; a.Z = savedm + p.o_Z; a.pw = savedm + p.o_pw;
      a.row_ptr = adj ? adj->row_ptr : nullptr; a.col_idx = adj ? adj->col_idx : nullptr;
      a.csc_ptr = adj ? adj->csc_ptr : nullptr; a.csc_row = adj ? adj->csc_row : nullptr; a.csc_eid = adj ? adj->csc_eid : nullptr;
      for (int q = 0; q < NJOBS; ++q) {
        a.jobs[q] = L.jobs[q];
        if (q == JOB_MC) { a.jobs[q].out = scr + L.o_dMc; a.jobs[q].blk_stride = (int64_t)p.Oc * p.C; }
        else { a.jobs[q].out = scr + L.o_wg + L.off_job[q]; a.jobs[q].blk_stride = L.WGT; }
      }
      a.bias_part = scr + L.o_bias; a.ln_part = scr + L.o_ln; a.dPt_part = reinterpret_cast<double*>(scr + L.o_dPt);
      a.delta = scr + L.o_delta; a.dSnz = scr + L.o_dSnz; a.dx = dxm;
      a.nbias = L.nbias; a.arena_floats = bwd_arena_floats(p, L.T);
      size_t smem = (size_t)bwd_smem_floats(p, L.NT, L.T) * 4;
      a.S = L.S;
      a.skip_wgrad = tc_w ? 1 : 0;
      a.resident = (tc_w && bwd_resident_floats(p) > 0) ? 1 : 0;
      a.tcA = tc_w ? scr + L.o_tcA : nullptr;
#define BWD_CASE(NTV, RBV) MSGAT_LAUNCH((k_bwd_main<NTV, RBV>), dim3(L.J, p.B), dim3(L.T), smem, stream, a)
      if (L.RBT == 2) {
        if (L.NT == 16) BWD_CASE(16, 2); else if (L.NT == 8) BWD_CASE(8, 2); else BWD_CASE(4, 2);
      } else {
        if (L.NT == 16) BWD_CASE(16, 4); else if (L.NT == 8) BWD_CASE(8, 4); else BWD_CASE(4, 4);
      }
#undef BWD_CASE
    }
  }
#ifndef MSGAT_EMU
  if (tc_w) {   // every weight gradient of all components in one persistent launch (tc_wgrad.cu); dA / GZ come from the passes above
    float* wg = scratch + L.o_wg;
    launch_tc_wgrad(p, prm, adj, x, need_mask ? scratch + L.o_G : dout, need_mask ? sc_sm : out_sm, saved, scratch + L.o_tcA,
                    wg + L.off_job[JOB_RES], wg + L.off_job[JOB_WS], wg + L.off_job[JOB_CONV + 0], wg + L.off_job[JOB_CONV + 1],
                    wg + L.off_job[JOB_CONV + 2], wg + L.off_job[JOB_CONV + 3], L.WGT, scratch + L.o_dMc, (int64_t)p.Oc * p.C, stream);
  }
#endif
  return meam_backward_tail(p, L, Jw, prm, adj, x, saved, scratch, dx, g, stream);
}
