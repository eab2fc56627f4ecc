// spatial.cu -- the N x N part of GraphAttention (attention.py:32-36) on the fp32 path.
//
// The reference materialises S = q Wg q^T, softmax(S) and softmax(S) * adj as [B,N,N]
// tensors.  Here the softmax statistics of every row (max m, denominator l) are taken over
// ALL N columns -- the adjacency mask is applied AFTER the softmax (attention.py:34,36),
// so off-graph columns still feed the denominator -- while the masked attention itself is
// evaluated only at the adjacency non-zeros (CSR): pw[b][e] = exp(S[n,i]-m[n])/l[n]*adj[n,i].
// The backward needs two more dense reductions of the same scores, E = softmax(S) q (row
// direction, fused here) and the column-direction pass k_scol below.
#include "kernels.h"
#include "compat.cuh"

#define SS_ROWS 64
#define SS_QUART 4
#define SS_CHUNK 256

struct SStatsArgs {   // grouped: blockIdx.z = component m, buffers sv_sm floats apart, Wg per component
  int N, nnz;
  int64_t sv_sm;
  const float* Wg[MSGAT_MAX_GROUP];
  const float *qs, *val;
  const int32_t *row_ptr, *col_idx;
  float *m, *U, *pw;
  float *l, *E;  // (as doubles) kept in fp64: the backward subtracts delta * E and delta / l * P from sums they nearly cancel
};

// Thread map: a block owns SS_ROWS = 64 rows; a warp is 32 rows x ONE column quarter, so the q row of a
// column is the same shared-memory address for every lane (a broadcast: one wavefront per 128-bit read
// instead of one per 8 lanes) and the 24 FMAs + exp of a pair run against registers only.  The four
// column quarters of a row are merged through shared memory in a fixed order.
MSGAT_KERNEL MSGAT_LB(256, 2) k_sstats(SStatsArgs a) {
  MSGAT_DYN_SMEM(smem_raw);
  float* Q = reinterpret_cast<float*>(smem_raw);  // [SS_CHUNK][12]
  float* WG = Q + SS_CHUNK * 12;                  // [144]
  double* MG = reinterpret_cast<double*>(WG + 144);  // [SS_QUART][SS_ROWS][14]: m, l, E[12]
  const int b = blockIdx.y, rl = threadIdx.x % SS_ROWS, cq = threadIdx.x / SS_ROWS;
  const int n = blockIdx.x * SS_ROWS + rl;
  const bool live = n < a.N;
  const int64_t mo = (int64_t)blockIdx.z * a.sv_sm;
  const float* qb = a.qs + mo + (int64_t)b * a.N * 12;
  const float* Wgm = a.Wg[blockIdx.z];
  for (int i = threadIdx.x; i < 144; i += blockDim.x) WG[i] = Wgm[i];
  __syncthreads();
  // U[n,:] = q[n,:] Wg   (k @ Wg, attention.py:34)
  float U[12];
  {
    float q[12];
    for (int t = 0; t < 12; ++t) q[t] = live ? qb[(int64_t)n * 12 + t] : 0.f;
    for (int t2 = 0; t2 < 12; ++t2) {
      float s = 0.f;
      for (int t = 0; t < 12; ++t) s = fmaf(q[t], WG[t * 12 + t2], s);
      U[t2] = s;
    }
  }
  float2 U2[6];
  for (int k = 0; k < 6; ++k) U2[k] = make_float2(U[2 * k], U[2 * k + 1]);
  float m = -INFINITY;
  double l = 0.0, E[12];
  for (int t = 0; t < 12; ++t) E[t] = 0.0;
  for (int j0 = 0; j0 < a.N; j0 += SS_CHUNK) {
    const int len = (a.N - j0 < SS_CHUNK) ? a.N - j0 : SS_CHUNK;
    __syncthreads();
    for (int i = threadIdx.x; i < len * 3; i += blockDim.x) st4(Q + i * 4, ld4(qb + (int64_t)j0 * 12 + i * 4));
    __syncthreads();
    // Four consecutive columns per step: the scores and the products p * q of a group are summed in fp32 (p
    // itself is an fp32 expf), the running sums stay in fp64 -- one conversion per group instead of one per pair.
    for (int j = cq * 4; j < len; j += 4 * SS_QUART) {
      float4 q[4][3];
      float sc4[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int ju = j + u;
        const float* qj = Q + (ju < len ? ju : j) * 12;
        q[u][0] = ld4(qj); q[u][1] = ld4(qj + 4); q[u][2] = ld4(qj + 8);
        float2 s2 = fmul2(U2[0], lo2(q[u][0]));
        s2 = ffma2(U2[1], hi2(q[u][0]), s2); s2 = ffma2(U2[2], lo2(q[u][1]), s2); s2 = ffma2(U2[3], hi2(q[u][1]), s2);
        s2 = ffma2(U2[4], lo2(q[u][2]), s2); s2 = ffma2(U2[5], hi2(q[u][2]), s2);
        const float sv = s2.x + s2.y;
        sc4[u] = ju < len ? sv : -INFINITY;
      }
      const float mg = fmaxf(fmaxf(sc4[0], sc4[1]), fmaxf(sc4[2], sc4[3]));
      if (mg > m) {
        const double c = (double)expf(m - mg);
        l *= c;
        for (int t = 0; t < 12; ++t) E[t] *= c;
        m = mg;
      }
      float lc = 0.f;
      float2 Ec2[6];
#pragma unroll
      for (int k = 0; k < 6; ++k) Ec2[k] = make_float2(0.f, 0.f);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float pf = exp_neg(sc4[u] - m);   // 0 for the padded columns
        const float2 p2 = make_float2(pf, pf);
        lc += pf;
        Ec2[0] = ffma2(p2, lo2(q[u][0]), Ec2[0]); Ec2[1] = ffma2(p2, hi2(q[u][0]), Ec2[1]);
        Ec2[2] = ffma2(p2, lo2(q[u][1]), Ec2[2]); Ec2[3] = ffma2(p2, hi2(q[u][1]), Ec2[3]);
        Ec2[4] = ffma2(p2, lo2(q[u][2]), Ec2[4]); Ec2[5] = ffma2(p2, hi2(q[u][2]), Ec2[5]);
      }
      l += (double)lc;
#pragma unroll
      for (int k = 0; k < 6; ++k) { E[2 * k] += (double)Ec2[k].x; E[2 * k + 1] += (double)Ec2[k].y; }
    }
  }
  // merge the column quarters of each row, quarter 0 first
  {
    double* mine = MG + ((int64_t)cq * SS_ROWS + rl) * 14;
    mine[0] = (double)m; mine[1] = l;
    for (int t = 0; t < 12; ++t) mine[2 + t] = E[t];
  }
  __syncthreads();
  if (cq != 0 || !live) return;
  float M = m;
  for (int k = 1; k < SS_QUART; ++k) M = fmaxf(M, (float)MG[((int64_t)k * SS_ROWS + rl) * 14]);
  {
    const double sc = (m == -INFINITY) ? 0.0 : (double)expf(m - M);
    l *= sc;
    for (int t = 0; t < 12; ++t) E[t] *= sc;
  }
  for (int k = 1; k < SS_QUART; ++k) {
    const double* o = MG + ((int64_t)k * SS_ROWS + rl) * 14;
    const float mk = (float)o[0];
    const double sc = (mk == -INFINITY) ? 0.0 : (double)expf(mk - M);
    l += o[1] * sc;
    for (int t = 0; t < 12; ++t) E[t] += o[2 + t] * sc;
  }
  const double invd = 1.0 / l;
  const float inv = (float)invd;
  a.m[mo + (int64_t)b * a.N + n] = M;
  reinterpret_cast<double*>(a.l + mo)[(int64_t)b * a.N + n] = l;
  for (int t = 0; t < 12; ++t) reinterpret_cast<double*>(a.E + mo)[((int64_t)b * a.N + n) * 12 + t] = E[t] * invd;
  for (int t = 0; t < 12; ++t) a.U[mo + ((int64_t)b * a.N + n) * 12 + t] = U[t];  // reused by k_scol
  // masked attention at the adjacency non-zeros of this row (attention.py:36: att * adjacency)
  const int e1 = a.row_ptr[n + 1];
  for (int e = a.row_ptr[n]; e < e1; ++e) {
    // the score of the pair in EXACTLY the arithmetic of the dense pass above (same FFMA2 pairing and order):
    // for the row's arg-max s - M is then 0.0 and not the rounding difference of two summation orders, which
    // at |s| of several hundred (peaked weights) would be 1e-5 of the dominant attention weight
    const float* qi = qb + (int64_t)a.col_idx[e] * 12;
    const float4 q0 = ld4(qi), q1 = ld4(qi + 4), q2 = ld4(qi + 8);
    float2 s2 = fmul2(U2[0], lo2(q0));
    s2 = ffma2(U2[1], hi2(q0), s2); s2 = ffma2(U2[2], lo2(q1), s2); s2 = ffma2(U2[3], hi2(q1), s2);
    s2 = ffma2(U2[4], lo2(q2), s2); s2 = ffma2(U2[5], hi2(q2), s2);
    const float s = s2.x + s2.y;
    a.pw[mo + (int64_t)b * a.nnz + e] = expf(s - M) * inv * a.val[e];
  }
}

int launch_sstats(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, float* saved, void* stream) {
  if (score_tc_supported(p)) return launch_sstats_tc(p, prm, adj, saved, stream);
  if (score_tc32_supported(p)) return launch_sstats_tc32(p, prm, adj, saved, stream);
  SStatsArgs a;
  a.N = p.N; a.nnz = p.nnz;
  a.sv_sm = p.saved_total;
  for (int m = 0; m < p.M; ++m) a.Wg[m] = prm[m].sa_Wg;
  a.qs = saved + p.o_qs; a.val = adj->val;
  a.row_ptr = adj->row_ptr; a.col_idx = adj->col_idx;
  a.m = saved + p.o_m; a.l = saved + p.o_l; a.E = saved + p.o_E;
  a.U = saved + p.o_U; a.pw = saved + p.o_pw;
  dim3 grid((p.N + SS_ROWS - 1) / SS_ROWS, p.B, p.M);
  size_t smem = (SS_CHUNK * 12 + 144) * sizeof(float) + (size_t)SS_QUART * SS_ROWS * 14 * sizeof(double);
  MSGAT_LAUNCH(k_sstats, grid, dim3(256), smem, stream, a);
  return 0;
}

// ---------------------------------------------------------------------------------------
// Backward, column direction.  With dS[n,i] = P[n,i] (adj[n,i] dA[n,i] - delta[n]) the dense
// part of dq[i,:] = sum_n dS[n,i] U[n,:] is  -sum_n P[n,i] delta[n] U[n,:]  over ALL rows n:
//   dq_dense[i,:] = -sum_n exp(q[i].U[n] - m[n]) * (delta[n]/l[n]) * U[n,:]
// One thread group per column i, rows streamed through shared memory as (U[n,:], m[n], w[n]).
struct SColArgs {   // grouped: blockIdx.z = component; saved buffers sv_sm apart, scratch buffers sc_sm apart
  int N;
  int64_t sv_sm, sc_sm;
  const float *qs, *U, *m, *delta;
  const float* l;   // (as doubles)
  float* dq;  // [B,N,12]  written (dense part)
};

MSGAT_KERNEL MSGAT_LB(256, 2) k_scol(SColArgs a) {
  MSGAT_DYN_SMEM(smem_raw);
  float* UR = reinterpret_cast<float*>(smem_raw);  // [SS_CHUNK][16]: U[12], m, w = delta / l; read as warp-wide broadcasts
  double* MG = reinterpret_cast<double*>(UR + SS_CHUNK * 16);  // [SS_QUART][SS_ROWS][12]
  const int b = blockIdx.y, rl = threadIdx.x % SS_ROWS, cq = threadIdx.x / SS_ROWS;   // same thread map as k_sstats
  const int i = blockIdx.x * SS_ROWS + rl;
  const bool live = i < a.N;
  const int64_t mo = (int64_t)blockIdx.z * a.sv_sm, so = (int64_t)blockIdx.z * a.sc_sm;
  const float* qb = a.qs + mo + (int64_t)b * a.N * 12;
  const float* Ub = a.U + mo + (int64_t)b * a.N * 12;
  const float* mb = a.m + mo;
  const double* lb = reinterpret_cast<const double*>(a.l + mo);
  const float* deltab = a.delta + so;
  float qi[12];
  double acc[12];  // the dense sum cancels against the sparse part of dq: keep it in fp64
  for (int t = 0; t < 12; ++t) { qi[t] = live ? qb[(int64_t)i * 12 + t] : 0.f; acc[t] = 0.0; }
  float2 qi2[6];
  for (int k = 0; k < 6; ++k) qi2[k] = make_float2(qi[2 * k], qi[2 * k + 1]);
  for (int n0 = 0; n0 < a.N; n0 += SS_CHUNK) {
    const int len = (a.N - n0 < SS_CHUNK) ? a.N - n0 : SS_CHUNK;
    __syncthreads();
    // stage the U rows of this chunk (written by k_sstats) with their (m, delta / l)
    for (int k = threadIdx.x; k < len * 3; k += blockDim.x) {
      const int r = k / 3, c4 = (k - r * 3) * 4;
      st4(UR + r * 16 + c4, ld4(Ub + (int64_t)(n0 + r) * 12 + c4));
    }
    for (int r = threadIdx.x; r < len; r += blockDim.x) {
      const int64_t g = (int64_t)b * a.N + n0 + r;
      UR[r * 16 + 12] = mb[g];
      UR[r * 16 + 13] = (float)((double)deltab[g] / lb[g]);
    }
    __syncthreads();
    // four consecutive rows per step, summed in fp32 inside the group and in fp64 across groups (see k_sstats)
    for (int r = cq * 4; r < len; r += 4 * SS_QUART) {
      float2 ac2[6];
#pragma unroll
      for (int k = 0; k < 6; ++k) ac2[k] = make_float2(0.f, 0.f);
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        const int rg = r + g;
        if (rg < len) {
          const float* u = UR + rg * 16;
          const float4 u0 = ld4(u), u1 = ld4(u + 4), u2 = ld4(u + 8), u3 = ld4(u + 12);
          float2 s2 = fmul2(qi2[0], lo2(u0));
          s2 = ffma2(qi2[1], hi2(u0), s2); s2 = ffma2(qi2[2], lo2(u1), s2); s2 = ffma2(qi2[3], hi2(u1), s2);
          s2 = ffma2(qi2[4], lo2(u2), s2); s2 = ffma2(qi2[5], hi2(u2), s2);
          const float pf = exp_neg((s2.x + s2.y) - u3.x) * u3.y;   // u3.y = (float)(delta / l)
          const float2 p2 = make_float2(pf, pf);
          ac2[0] = ffma2(p2, lo2(u0), ac2[0]); ac2[1] = ffma2(p2, hi2(u0), ac2[1]);
          ac2[2] = ffma2(p2, lo2(u1), ac2[2]); ac2[3] = ffma2(p2, hi2(u1), ac2[3]);
          ac2[4] = ffma2(p2, lo2(u2), ac2[4]); ac2[5] = ffma2(p2, hi2(u2), ac2[5]);
        }
      }
#pragma unroll
      for (int k = 0; k < 6; ++k) { acc[2 * k] += (double)ac2[k].x; acc[2 * k + 1] += (double)ac2[k].y; }
    }
  }
  {
    double* mine = MG + ((int64_t)cq * SS_ROWS + rl) * 12;
    for (int t = 0; t < 12; ++t) mine[t] = acc[t];
  }
  __syncthreads();
  if (cq != 0 || !live) return;
  for (int k = 1; k < SS_QUART; ++k) {
    const double* o = MG + ((int64_t)k * SS_ROWS + rl) * 12;
    for (int t = 0; t < 12; ++t) acc[t] += o[t];
  }
  for (int t = 0; t < 12; ++t) a.dq[so + ((int64_t)b * a.N + i) * 12 + t] = (float)(-acc[t]);
}

int launch_scol(const Plan& p, const float* Wg, const float* saved, const float* delta, float* dq, void* stream) {
  if (score_tc_supported(p)) return launch_scol_tc(p, saved, delta, dq, stream);
  if (score_tc32_supported(p)) return launch_scol_tc32(p, saved, delta, dq, stream);
  SColArgs a;
  a.N = p.N; a.sv_sm = p.saved_total; a.sc_sm = p.scratch_sm;
  a.qs = saved + p.o_qs; a.U = saved + p.o_U; a.m = saved + p.o_m; a.l = saved + p.o_l; a.delta = delta; a.dq = dq;
  dim3 grid((p.N + SS_ROWS - 1) / SS_ROWS, p.B, p.M);
  size_t smem = (SS_CHUNK * 16) * sizeof(float) + (size_t)SS_QUART * SS_ROWS * 12 * sizeof(double);
  MSGAT_LAUNCH(k_scol, grid, dim3(256), smem, stream, a);
  return 0;
}
