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

#define SS_ROWS 32
#define SS_LANES 8
#define SS_CHUNK 256

struct SStatsArgs {
  int N, nnz;
  const float *qs, *Wg, *val;
  const int32_t *row_ptr, *col_idx;
  float *m, *U, *pw;
  double *l, *E;  // kept in fp64: the backward subtracts delta * E and delta / l * P from sums they nearly cancel
};

MSGAT_KERNEL MSGAT_LB(256, 2) k_sstats(SStatsArgs a) {
  MSGAT_DYN_SMEM(smem_raw);
  float* Q = reinterpret_cast<float*>(smem_raw);  // [SS_CHUNK][12]
  float* WG = Q + SS_CHUNK * 12;                  // [144]
  const int b = blockIdx.y, rl = threadIdx.x / SS_LANES, cl = threadIdx.x % SS_LANES;
  const int n = blockIdx.x * SS_ROWS + rl;
  const bool live = n < a.N;
  const float* qb = a.qs + (int64_t)b * a.N * 12;
  for (int i = threadIdx.x; i < 144; i += blockDim.x) WG[i] = a.Wg[i];
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
  float m = -INFINITY;
  double l = 0.0, E[12];
  for (int t = 0; t < 12; ++t) E[t] = 0.0;
  for (int j0 = 0; j0 < a.N; j0 += SS_CHUNK) {
    const int len = (a.N - j0 < SS_CHUNK) ? a.N - j0 : SS_CHUNK;
    __syncthreads();
    for (int i = threadIdx.x; i < len * 3; i += blockDim.x) st4(Q + i * 4, ld4(qb + (int64_t)j0 * 12 + i * 4));
    __syncthreads();
    for (int j = cl; j < len; j += SS_LANES) {
      const float* qj = Q + j * 12;
      float4 q0 = ld4(qj), q1 = ld4(qj + 4), q2 = ld4(qj + 8);
      float s = U[0] * q0.x;
      s = fmaf(U[1], q0.y, s); s = fmaf(U[2], q0.z, s); s = fmaf(U[3], q0.w, s);
      s = fmaf(U[4], q1.x, s); s = fmaf(U[5], q1.y, s); s = fmaf(U[6], q1.z, s); s = fmaf(U[7], q1.w, s);
      s = fmaf(U[8], q2.x, s); s = fmaf(U[9], q2.y, s); s = fmaf(U[10], q2.z, s); s = fmaf(U[11], q2.w, s);
      if (s > m) {
        const double c = (double)expf(m - s);
        l *= c;
        for (int t = 0; t < 12; ++t) E[t] *= c;
        m = s;
      }
      const double p = (double)expf(s - m);
      l += p;
      E[0] = fma(p, (double)q0.x, E[0]); E[1] = fma(p, (double)q0.y, E[1]); E[2] = fma(p, (double)q0.z, E[2]);
      E[3] = fma(p, (double)q0.w, E[3]); E[4] = fma(p, (double)q1.x, E[4]); E[5] = fma(p, (double)q1.y, E[5]);
      E[6] = fma(p, (double)q1.z, E[6]); E[7] = fma(p, (double)q1.w, E[7]); E[8] = fma(p, (double)q2.x, E[8]);
      E[9] = fma(p, (double)q2.y, E[9]); E[10] = fma(p, (double)q2.z, E[10]); E[11] = fma(p, (double)q2.w, E[11]);
    }
  }
  // combine the 8 column lanes of each row (lanes of a row are adjacent in the warp)
  float M = m;
  M = fmaxf(M, __shfl_xor_sync(0xffffffffu, M, 1));
  M = fmaxf(M, __shfl_xor_sync(0xffffffffu, M, 2));
  M = fmaxf(M, __shfl_xor_sync(0xffffffffu, M, 4));
  const double sc = (m == -INFINITY) ? 0.0 : (double)expf(m - M);
  l *= sc;
  l += __shfl_xor_sync(0xffffffffu, l, 1);
  l += __shfl_xor_sync(0xffffffffu, l, 2);
  l += __shfl_xor_sync(0xffffffffu, l, 4);
  for (int t = 0; t < 12; ++t) {
    double e = E[t] * sc;
    e += __shfl_xor_sync(0xffffffffu, e, 1);
    e += __shfl_xor_sync(0xffffffffu, e, 2);
    e += __shfl_xor_sync(0xffffffffu, e, 4);
    E[t] = e;
  }
  if (!live) return;
  const double invd = 1.0 / l;
  const float inv = (float)invd;
  if (cl == 0) {
    a.m[(int64_t)b * a.N + n] = M;
    a.l[(int64_t)b * a.N + n] = l;
    for (int t = 0; t < 12; ++t) a.E[((int64_t)b * a.N + n) * 12 + t] = E[t] * invd;
    for (int t = 0; t < 12; ++t) a.U[((int64_t)b * a.N + n) * 12 + t] = U[t];  // reused by k_scol
  }
  // masked attention at the adjacency non-zeros of this row (attention.py:36: att * adjacency)
  const int e1 = a.row_ptr[n + 1];
  for (int e = a.row_ptr[n] + cl; e < e1; e += SS_LANES) {
    const float* qi = qb + (int64_t)a.col_idx[e] * 12;
    float s = 0.f;
    for (int t = 0; t < 12; ++t) s = fmaf(U[t], qi[t], s);
    a.pw[(int64_t)b * a.nnz + e] = expf(s - M) * inv * a.val[e];
  }
}

int launch_sstats(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, float* saved, void* stream) {
  SStatsArgs a;
  a.N = p.N; a.nnz = p.nnz;
  a.qs = saved + p.o_qs; a.Wg = prm->sa_Wg; a.val = adj->val;
  a.row_ptr = adj->row_ptr; a.col_idx = adj->col_idx;
  a.m = saved + p.o_m; a.l = reinterpret_cast<double*>(saved + p.o_l); a.E = reinterpret_cast<double*>(saved + p.o_E);
  a.U = saved + p.o_U; a.pw = saved + p.o_pw;
  dim3 grid((p.N + SS_ROWS - 1) / SS_ROWS, p.B);
  size_t smem = (SS_CHUNK * 12 + 144) * sizeof(float);
  MSGAT_LAUNCH(k_sstats, grid, dim3(256), smem, stream, a);
  return 0;
}

// ---------------------------------------------------------------------------------------
// Backward, column direction.  With dS[n,i] = P[n,i] (adj[n,i] dA[n,i] - delta[n]) the dense
// part of dq[i,:] = sum_n dS[n,i] U[n,:] is  -sum_n P[n,i] delta[n] U[n,:]  over ALL rows n:
//   dq_dense[i,:] = -sum_n exp(q[i].U[n] - m[n]) * (delta[n]/l[n]) * U[n,:]
// One thread group per column i, rows streamed through shared memory as (U[n,:], m[n], w[n]).
struct SColArgs {
  int N;
  const float *qs, *U, *m, *delta;
  const double* l;
  float* dq;  // [B,N,12]  written (dense part)
};

MSGAT_KERNEL MSGAT_LB(256, 2) k_scol(SColArgs a) {
  MSGAT_DYN_SMEM(smem_raw);
  float* UR = reinterpret_cast<float*>(smem_raw);  // [SS_CHUNK][20]: U[12], m, pad, w = delta / l as one fp64 in slots 14-15; the 80-byte pitch keeps the 8 row lanes on distinct banks
  const int b = blockIdx.y, rl = threadIdx.x / SS_LANES, cl = threadIdx.x % SS_LANES;
  const int i = blockIdx.x * SS_ROWS + rl;
  const bool live = i < a.N;
  const float* qb = a.qs + (int64_t)b * a.N * 12;
  const float* Ub = a.U + (int64_t)b * a.N * 12;
  float qi[12];
  double acc[12];  // the dense sum cancels against the sparse part of dq: keep it in fp64
  for (int t = 0; t < 12; ++t) { qi[t] = live ? qb[(int64_t)i * 12 + t] : 0.f; acc[t] = 0.0; }
  for (int n0 = 0; n0 < a.N; n0 += SS_CHUNK) {
    const int len = (a.N - n0 < SS_CHUNK) ? a.N - n0 : SS_CHUNK;
    __syncthreads();
    // stage the U rows of this chunk (written by k_sstats) with their (m, delta / l)
    for (int k = threadIdx.x; k < len * 3; k += blockDim.x) {
      const int r = k / 3, c4 = (k - r * 3) * 4;
      st4(UR + r * 20 + c4, ld4(Ub + (int64_t)(n0 + r) * 12 + c4));
    }
    for (int r = threadIdx.x; r < len; r += blockDim.x) {
      const int64_t g = (int64_t)b * a.N + n0 + r;
      UR[r * 20 + 12] = a.m[g];
      *reinterpret_cast<double*>(UR + r * 20 + 14) = (double)a.delta[g] / a.l[g];
    }
    __syncthreads();
    for (int r = cl; r < len; r += SS_LANES) {
      const float* u = UR + r * 20;
      float4 u0 = ld4(u), u1 = ld4(u + 4), u2 = ld4(u + 8), u3 = ld4(u + 12);
      float s = qi[0] * u0.x;
      s = fmaf(qi[1], u0.y, s); s = fmaf(qi[2], u0.z, s); s = fmaf(qi[3], u0.w, s);
      s = fmaf(qi[4], u1.x, s); s = fmaf(qi[5], u1.y, s); s = fmaf(qi[6], u1.z, s); s = fmaf(qi[7], u1.w, s);
      s = fmaf(qi[8], u2.x, s); s = fmaf(qi[9], u2.y, s); s = fmaf(qi[10], u2.z, s); s = fmaf(qi[11], u2.w, s);
      const double p = (double)expf(s - u3.x) * *reinterpret_cast<const double*>(u + 14);
      acc[0] = fma(p, (double)u0.x, acc[0]); acc[1] = fma(p, (double)u0.y, acc[1]);
      acc[2] = fma(p, (double)u0.z, acc[2]); acc[3] = fma(p, (double)u0.w, acc[3]);
      acc[4] = fma(p, (double)u1.x, acc[4]); acc[5] = fma(p, (double)u1.y, acc[5]);
      acc[6] = fma(p, (double)u1.z, acc[6]); acc[7] = fma(p, (double)u1.w, acc[7]);
      acc[8] = fma(p, (double)u2.x, acc[8]); acc[9] = fma(p, (double)u2.y, acc[9]);
      acc[10] = fma(p, (double)u2.z, acc[10]); acc[11] = fma(p, (double)u2.w, acc[11]);
    }
  }
  for (int t = 0; t < 12; ++t) {
    double e = acc[t];
    e += __shfl_xor_sync(0xffffffffu, e, 1);
    e += __shfl_xor_sync(0xffffffffu, e, 2);
    e += __shfl_xor_sync(0xffffffffu, e, 4);
    acc[t] = e;
  }
  if (live && cl == 0)
    for (int t = 0; t < 12; ++t) a.dq[((int64_t)b * a.N + i) * 12 + t] = (float)(-acc[t]);
}

int launch_scol(const Plan& p, const float* Wg, const float* saved, const float* delta, float* dq, void* stream) {
  SColArgs a;
  a.N = p.N; a.qs = saved + p.o_qs; a.U = saved + p.o_U; a.m = saved + p.o_m; a.l = reinterpret_cast<const double*>(saved + p.o_l); a.delta = delta; a.dq = dq;
  dim3 grid((p.N + SS_ROWS - 1) / SS_ROWS, p.B);
  size_t smem = (SS_CHUNK * 20) * sizeof(float);
  MSGAT_LAUNCH(k_scol, grid, dim3(256), smem, stream, a);
  return 0;
}
