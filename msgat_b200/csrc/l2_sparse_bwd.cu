// l2_sparse_bwd.cu -- the two sparse stages of a layer-2 MEAM's backward (attention.py:35-36 transposed), taken out of
// the tcgen05 data-gradient kernel and run as one grouped launch over all components:
//   GZ[o,i,t]  = sum_n A[n,i] G_s[o,n,t]          transposed aggregation through the CSC view, A = pw at the non-zeros
//   dSnz[e]    = A[n,i_e] sum_{o,t} G_s[o,n,t] Z[o,i_e,t]        dS at the adjacency non-zeros of row n
//   delta[n]   = sum_e dSnz[e]                                    its row sums (softmax backward)
// Inside k_tc_dgrad these were chains of dependent gathers (index -> weight -> rows) on a few warps of a 1-CTA/SM kernel,
// in the shadow of a short MMA window they outlasted several times over.  Here a warp owns one node of one (sample,
// component): the edge lists are read by the lanes in parallel and broadcast, the row gathers of different edges are
// independent, and ~60 such warps are resident per SM.  GZ goes to rows [2b, 3b) of the tcA buffer, where k_tc_dgrad
// (operand of its second GEMM) and k_tc_wgrad read it.
#include "kernels.h"
#include "compat.cuh"

#define GZ_WARPS 8

struct GzArgs {
  const float* G; int64_t g_sm;       // upstream gradient [B, 3b, N, 12] per component
  int N, b, nnz;
  int64_t sv_sm, sc_sm;
  const float *Z, *pw;                // saved
  const int32_t *row_ptr, *col_idx, *csc_ptr, *csc_row, *csc_eid;
  float *tcA, *delta, *dSnz;          // scratch
};

MSGAT_KERNEL MSGAT_LB(GZ_WARPS * 32, 4) k_l2_gz(GzArgs a) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, bi = blockIdx.y, m = blockIdx.z, b = a.b;
  const int n = blockIdx.x * GZ_WARPS + warp;
  if (n >= a.N) return;
  const int64_t cs = (int64_t)a.N * 12, mo = (int64_t)m * a.sv_sm, so = (int64_t)m * a.sc_sm;
  const float* Gs = a.G + (int64_t)m * a.g_sm + ((int64_t)bi * 3 * b + 2 * b) * cs;
  const float* pwb = a.pw + mo + (int64_t)bi * a.nnz;
  const float* Zb = a.Z + mo + (int64_t)bi * b * cs;
  float* gzo = a.tcA + so + ((int64_t)bi * 3 * b + 2 * b) * cs + (int64_t)n * 12;
  // items of a node: (channel o, quad q4) pairs, item = lane + 32 u
  int64_t ioff[3];
  bool ion[3];
#pragma unroll
  for (int u = 0; u < 3; ++u) {
    const int it = lane + 32 * u, o = it / 3, q4 = (it - o * 3) * 4;
    ion[u] = it < b * 3;
    ioff[u] = ion[u] ? (int64_t)o * cs + q4 : 0;
  }
  // Both stages are chains of dependent gathers (pointer -> index -> weight -> rows).  The heads of the second chain (the row's
  // range, its first 32 edges, the node's own G rows) are requested here, in front of the first stage's loop, so that the two
  // chains run side by side instead of one after the other (the stores of GZ in between keep the compiler from doing it).
  const int k_lo = a.csc_ptr[n], k1 = a.csc_ptr[n + 1];
  const int e_lo = a.row_ptr[n], e1 = a.row_ptr[n + 1];
  const int cnt0 = e1 - e_lo < 32 ? e1 - e_lo : 32;
  const int myc0 = lane < cnt0 ? a.col_idx[e_lo + lane] : 0;
  const float myw0 = lane < cnt0 ? pwb[e_lo + lane] : 0.f;
  float4 gs[3];
#pragma unroll
  for (int u = 0; u < 3; ++u) gs[u] = ion[u] ? ld4(Gs + (int64_t)n * 12 + ioff[u]) : f4zero();
  // ---- GZ: the node's column of A
  float4 gz[3] = {f4zero(), f4zero(), f4zero()};
  {
    for (int k0 = k_lo; k0 < k1; k0 += 32) {
      const int cnt = k1 - k0 < 32 ? k1 - k0 : 32;
      const int myr = lane < cnt ? a.csc_row[k0 + lane] : 0;
      const float myw = lane < cnt ? pwb[a.csc_eid[k0 + lane]] : 0.f;
#pragma unroll 4
      for (int j = 0; j < cnt; ++j) {
        const float w = __shfl_sync(0xffffffffu, myw, j);
        const float* gr = Gs + (int64_t)__shfl_sync(0xffffffffu, myr, j) * 12;
#pragma unroll
        for (int u = 0; u < 3; ++u)
          if (ion[u]) fma4(gz[u], w, ld4(gr + ioff[u]));
      }
    }
  }
#pragma unroll
  for (int u = 0; u < 3; ++u)
    if (ion[u]) st4(gzo + ioff[u], gz[u]);
  // ---- dS at the non-zeros of the node's row, and their sum
  float dl = 0.f;
  for (int e0 = e_lo; e0 < e1; e0 += 32) {
    const int cnt = e1 - e0 < 32 ? e1 - e0 : 32;
    const int myc = e0 == e_lo ? myc0 : (lane < cnt ? a.col_idx[e0 + lane] : 0);
    const float myw = e0 == e_lo ? myw0 : (lane < cnt ? pwb[e0 + lane] : 0.f);
    float mine = 0.f;
#pragma unroll 4
    for (int j = 0; j < cnt; ++j) {
      const float* zr = Zb + (int64_t)__shfl_sync(0xffffffffu, myc, j) * 12;
      float s = 0.f;
#pragma unroll
      for (int u = 0; u < 3; ++u)
        if (ion[u]) s += dot4(gs[u], ld4(zr + ioff[u]));
      s = warp_sum(s);
      if (lane == j) mine = myw * s;
    }
    if (lane < cnt) a.dSnz[so + (int64_t)bi * a.nnz + e0 + lane] = mine;
    dl += warp_sum(mine);   // lanes >= cnt hold 0
  }
  if (lane == 0) a.delta[so + (int64_t)bi * a.N + n] = dl;
}

int launch_l2_gz(const Plan& p, const msgat_adj* adj, const float* G, int64_t g_sm, const float* saved, float* tcA, float* delta,
                 float* dSnz, void* stream) {
  GzArgs a;
  a.G = G; a.g_sm = g_sm; a.N = p.N; a.b = p.Oc; a.nnz = p.nnz; a.sv_sm = p.saved_total; a.sc_sm = p.scratch_sm;
  a.Z = saved + p.o_Z; a.pw = saved + p.o_pw;
  a.row_ptr = adj->row_ptr; a.col_idx = adj->col_idx; a.csc_ptr = adj->csc_ptr; a.csc_row = adj->csc_row; a.csc_eid = adj->csc_eid;
  a.tcA = tcA; a.delta = delta; a.dSnz = dSnz;
  MSGAT_LAUNCH(k_l2_gz, dim3((p.N + GZ_WARPS - 1) / GZ_WARPS, p.B, p.M), dim3(GZ_WARPS * 32), 0, stream, a);
  return 0;
}
