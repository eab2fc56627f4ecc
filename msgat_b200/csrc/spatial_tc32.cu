// spatial_tc32.cu -- the N x N score passes of GraphAttention (attention.py:32-36) with the SCORES on tcgen05 at fp32
// grade, for the default MSGAT_PREC_FP32 path.
//
// k_sstats / k_scol (spatial.cu) spend a third of their instructions on the 12-deep dot product of every (row, column)
// pair: 3 shared-memory loads + 6 FFMA2 + a horizontal add per score.  Here the tensor core produces the scores:
//   forward  (k_score_tc32<0>):  S[n,i]   = U[n,:] . q[i,:]   (U = q Wg)      per 128 x 128 tile
//   backward (k_score_tc32<1>):  S^T[i,n] = q[i,:] . U[n,:]
// Every operand v is split THREE ways, v = v1 + v2 + v3 with v1 = tf32(v), v2 = tf32(v - v1), v3 = v - v1 - v2 (exact: 11 +
// 11 + 2 bits), and the six products of total order <= 2 are issued (the dropped ones are 2^-33 of |U||q|): all partial
// products are exact in fp32, so S differs from the exact dot product only by the accumulator roundings -- like the FFMA
// chain it replaces.  A two-way split is NOT enough here: with trained (peaked) weights |U||q| reaches several hundred and
// its 2^-22 goes straight into the exponent.  v1.v1 goes to a MAIN accumulator, the five corrections to a SMALL one
// (tcgen05 truncates on accumulation, see l2_common.cuh); the reader adds them.
// The forward and the backward issue the SAME products in the same order (operand roles swapped), and the scores at the
// adjacency non-zeros are picked out of the same accumulators, so m, pw and the backward's exp(S - m) describe one set of
// scores (for a row's arg-max S - m is exactly 0), as in the CUDA-core kernels.
// Everything after the scores is the arithmetic of k_sstats / k_scol: online maximum, exp by ex2.approx(x log2 e), the
// P q / P' U contraction in fp32 FFMA2 over 8 columns, running sums in fp64.
//
// A CTA owns 128 rows of one (sample, component); 256 threads: warp w reads TMEM lanes 32 (w % 4).., the two warps of a
// lane quarter split the 128 columns of a tile.  256 TMEM columns and ~80 KB of shared memory: two CTAs per SM overlap each
// other's MMA and exp phases.  K-major operands in the unswizzled core-matrix layout, K = 12 padded to 16.
#include <math.h>
#include <stdlib.h>
#include "kernels.h"
#include "tc_common.cuh"

#define S3_ROWS 128
#define S3_COLS 128
#define S3_THREADS 256
#define S3_PART 8192   // one split part of a 128 x 16 operand tile

struct S3Args {   // grouped: blockIdx.z = component
  int N, nnz;
  int64_t sv_sm, sc_sm;
  const float* Wg[MSGAT_MAX_GROUP];
  const float *qs, *val;
  const int32_t *row_ptr, *col_idx;
  float *m, *U, *pw;
  float *l, *E;            // (as doubles)
  const float* delta;      // backward
  float* dq;
};

MSGAT_DEV float s3_tf32(float x) { return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u); }
// the three parts of a quad, stored at the same offset of the three part tiles
MSGAT_DEV void s3_split_store(unsigned char* base, uint32_t off, const float4& v) {
  const float4 a = make_float4(s3_tf32(v.x), s3_tf32(v.y), s3_tf32(v.z), s3_tf32(v.w));
  const float4 r = make_float4(v.x - a.x, v.y - a.y, v.z - a.z, v.w - a.w);
  const float4 b = make_float4(s3_tf32(r.x), s3_tf32(r.y), s3_tf32(r.z), s3_tf32(r.w));
  st4(reinterpret_cast<float*>(base + off), a);
  st4(reinterpret_cast<float*>(base + S3_PART + off), b);
  st4(reinterpret_cast<float*>(base + 2 * S3_PART + off), make_float4(r.x - b.x, r.y - b.y, r.z - b.z, r.w - b.w));
}

static inline size_t score_tc32_smem() {
  return 6 * S3_PART + 2 * S3_COLS * 12 * 4 + 2 * S3_COLS * 8 + (size_t)S3_ROWS * 14 * 8 + 144 * 4 + 64 + 1024;
}

template <int BWD>
MSGAT_KERNEL MSGAT_LB(S3_THREADS, 2) k_score_tc32(S3Args a) {
  MSGAT_DYN_SMEM(smem_raw);
  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // aligned inside the shared window: the pointer stays a shared-memory pointer (LDS / STS, not generic LD / ST)
  unsigned char* RA = smem;                        // row block, 3 parts   (forward: U, backward: q)
  unsigned char* CB = RA + 3 * S3_PART;            // column tile, 3 parts (forward: q, backward: U)
  float* QF = reinterpret_cast<float*>(CB + 3 * S3_PART);      // [2][128][12] the column tile's rows in fp32 (second contraction)
  float2* MW = reinterpret_cast<float2*>(QF + 2 * S3_COLS * 12);   // [2][128] backward: (m, delta / l) of the tile's columns
  double* MG = reinterpret_cast<double*>(MW + 2 * S3_COLS);    // [128][14] the upper column half's (m, l, E[12]) / acc[12]
  float* WG = reinterpret_cast<float*>(MG + S3_ROWS * 14);     // [144]
  uint64_t* bar = reinterpret_cast<uint64_t*>(WG + 144);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 2);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int r = (warp & 3) * 32 + lane, half = warp >> 2, b = blockIdx.y, n0 = blockIdx.x * S3_ROWS;
  const int64_t mo = (int64_t)blockIdx.z * a.sv_sm, so = (int64_t)blockIdx.z * a.sc_sm;
  const float* qb = a.qs + mo + (int64_t)b * a.N * 12;
  float* Ub = a.U + mo + (int64_t)b * a.N * 12;
  const float* colsrc = BWD ? Ub : qb;
  const int ntile = (a.N + S3_COLS - 1) / S3_COLS;
  const uint32_t barS = smem_u32(bar);

  if (!BWD) for (int i = tid; i < 144; i += S3_THREADS) WG[i] = a.Wg[blockIdx.z][i];
  // zero the K padding (k = 12..15) of all six part tiles
  for (int i = tid; i < 6 * S3_ROWS; i += S3_THREADS)
    st4(reinterpret_cast<float*>(RA + (i / S3_ROWS) * S3_PART + core_off(i % S3_ROWS, 12, 512u)), f4zero());
  __syncthreads();
  // ---- the row block
  if (tid < S3_ROWS) {
    const int n = n0 + tid;
    float v[12];
    if (BWD) {
      for (int t = 0; t < 12; ++t) v[t] = n < a.N ? qb[(int64_t)n * 12 + t] : 0.f;
    } else {
      float q[12];
      for (int t = 0; t < 12; ++t) q[t] = n < a.N ? qb[(int64_t)n * 12 + t] : 0.f;
      for (int t2 = 0; t2 < 12; ++t2) {   // U[n,:] = q[n,:] Wg   (k @ Wg, attention.py:34)
        float s = 0.f;
        for (int t = 0; t < 12; ++t) s = fmaf(q[t], WG[t * 12 + t2], s);
        v[t2] = s;
      }
      if (n < a.N) for (int t = 0; t < 12; ++t) Ub[(int64_t)n * 12 + t] = v[t];   // reused by the backward
    }
    for (int p4 = 0; p4 < 12; p4 += 4) s3_split_store(RA, core_off(tid, p4, 512u), make_float4(v[p4], v[p4 + 1], v[p4 + 2], v[p4 + 3]));
  }
  if (tid == 0) { mbar_init(barS, 1); mbar_fence_init(); }
  if (warp == 0) tmem_alloc(smem_u32(tmem_slot), 256);

  // items of a column tile: 128 rows x 3 quads; item x -> (row x / 3, quad x % 3)
  float4 pre[2];
  float2 premw = make_float2(0.f, 0.f);
  auto prefetch = [&](int j) {
    const int j0 = j * S3_COLS;
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int x = tid + u * S3_THREADS, rr = x / 3, p4 = (x - rr * 3) * 4;
      pre[u] = (x < S3_COLS * 3 && j0 + rr < a.N) ? ld4(colsrc + (int64_t)(j0 + rr) * 12 + p4) : f4zero();
    }
    if (BWD && tid < S3_COLS) {
      const int n = j0 + tid;
      premw = make_float2(0.f, 0.f);   // padded columns: weight 0
      if (n < a.N) {
        const int64_t g = (int64_t)b * a.N + n;
        premw.x = a.m[mo + g];
        premw.y = (float)((double)a.delta[so + g] / reinterpret_cast<const double*>(a.l + mo)[g]);
      }
    }
  };
  auto store_tile = [&](int j) {   // tile j: operand parts (single buffer: the MMAs of tile j - 1 are complete) and fp32 rows (double buffer)
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int x = tid + u * S3_THREADS, rr = x / 3, p4 = (x - rr * 3) * 4;
      if (x < S3_COLS * 3) {
        s3_split_store(CB, core_off(rr, p4, 512u), pre[u]);
        st4(QF + (j & 1) * S3_COLS * 12 + x * 4, pre[u]);
      }
    }
    if (BWD && tid < S3_COLS) MW[(j & 1) * S3_COLS + tid] = premw;
  };
  prefetch(0);
  store_tile(0);
  fence_smem_to_async();
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tm = warp_uniform(*tmem_slot);
  const uint32_t tMain = tmem_addr(tm, (warp & 3) * 32, half * 64), tSmall = tMain + S3_COLS;
  const uint32_t idS = umma_idesc_tf32(128, S3_COLS, 0, 0);
  uint32_t phase = 0;
  auto issue_scores = [&]() {
    // (U part, q part): (1,1) -> main; (1,2) (2,1) (1,3) (2,2) (3,1) -> small, in this order in BOTH directions (the backward's
    // row block is q and its column tile U, so it reads the list with the roles swapped)
    const int pa[6] = {0, 0, 1, 0, 1, 2}, pb[6] = {0, 1, 0, 2, 1, 0};
    for (int t = 0; t < 6; ++t)
      for (int ks = 0; ks < 2; ++ks)
        umma_tf32(tm + (t == 0 ? 0u : (uint32_t)S3_COLS), umma_desc(smem_u32(RA) + (BWD ? pb[t] : pa[t]) * S3_PART + ks * 256u, 128u, 512u, 0),
                  umma_desc(smem_u32(CB) + (BWD ? pa[t] : pb[t]) * S3_PART + ks * 256u, 128u, 512u, 0), idS, (t == 0 || t == 1) ? (ks > 0) : 1u);
    umma_commit(barS);
  };

  const int n = n0 + r;
  const bool live = n < a.N;
  float m = -INFINITY;
  double l = 0.0, E[12];
#pragma unroll
  for (int t = 0; t < 12; ++t) E[t] = 0.0;
  // the row's adjacency entries in ascending column order: this thread serves the ones inside its own column ranges
  int e = 0, e1 = 0, nc = 0x7fffffff;
  if (!BWD && live) { e = a.row_ptr[n]; e1 = a.row_ptr[n + 1]; nc = e < e1 ? a.col_idx[e] : 0x7fffffff; }
  float* pwb = BWD ? nullptr : a.pw + mo + (int64_t)b * a.nnz;

  for (int j = 0; j < ntile; ++j) {
    if (tid == 0) issue_scores();
    __syncwarp();
    if (j + 1 < ntile) prefetch(j + 1);
    mbar_wait(barS, phase); phase ^= 1;
    tc_fence_after_sync();
    const int cbase0 = j * S3_COLS + half * 64;
    const float* qf = QF + (j & 1) * S3_COLS * 12 + half * 64 * 12;
    const float2* mw = MW + (j & 1) * S3_COLS + half * 64;
    // fp32 partial sums of TWO chunks (16 columns) go to the fp64 accumulators at once: the fp32 -> fp64 conversions run on the
    // quarter-rate conversion pipe, 13 per chunk next to its 8 exponentials
    float lc = 0.f;
    float2 Ec2[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) Ec2[k] = make_float2(0.f, 0.f);
#pragma unroll 1
    for (int c0 = 0; c0 < 64; c0 += 8) {
      if (cbase0 + c0 >= a.N) break;   // warp-uniform: the chunk lies beyond the last column (every chunk that is processed has a live
                                       // column, so the running maximum is finite before the first exponential)
      uint32_t v0[8], v1[8];
      tmem_ld8_nowait(tMain + c0, v0);
      tmem_ld8_nowait(tSmall + c0, v1);
      tmem_ld_wait();
      float s[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) s[i] = __uint_as_float(v0[i]) + __uint_as_float(v1[i]);
      const int cbase = cbase0 + c0;
      if (!BWD) {
        // raw scores of the row's adjacency entries in this chunk (normalised once m and l are final)
        if (nc < cbase + 8) {   // rare: ~3 entries per row of N columns
          while (nc < cbase) { ++e; nc = e < e1 ? a.col_idx[e] : 0x7fffffff; }   // entries of the other half's ranges
#pragma unroll
          for (int i = 0; i < 8; ++i)   // static register indices (a dynamic pick turns s[] into a local-memory array)
            if (nc == cbase + i) { pwb[e] = s[i]; ++e; nc = e < e1 ? a.col_idx[e] : 0x7fffffff; }
        }
        if (cbase + 8 > a.N) {   // warp-uniform: only the last chunk of a row has columns beyond N
#pragma unroll
          for (int i = 0; i < 8; ++i)
            if (cbase + i >= a.N) s[i] = -INFINITY;
        }
        const float mg = fmaxf(fmaxf(fmaxf(s[0], s[1]), fmaxf(s[2], s[3])), fmaxf(fmaxf(s[4], s[5]), fmaxf(s[6], s[7])));
        if (mg > m) {
          const float cf = expf(m - mg);
          const double c = (double)cf;
          l *= c;
#pragma unroll
          for (int t = 0; t < 12; ++t) E[t] *= c;
          lc *= cf;   // the pending fp32 partials of the chunk before
#pragma unroll
          for (int k = 0; k < 6; ++k) Ec2[k] = fmul2(Ec2[k], make_float2(cf, cf));
          m = mg;
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float pf = exp_neg(s[i] - m);   // 0 for the padded columns
          const float2 p2 = make_float2(pf, pf);
          const float* qj = qf + (c0 + i) * 12;
          const float4 q0 = ld4(qj), q1 = ld4(qj + 4), q2 = ld4(qj + 8);
          lc += pf;
          Ec2[0] = ffma2(p2, lo2(q0), Ec2[0]); Ec2[1] = ffma2(p2, hi2(q0), Ec2[1]);
          Ec2[2] = ffma2(p2, lo2(q1), Ec2[2]); Ec2[3] = ffma2(p2, hi2(q1), Ec2[3]);
          Ec2[4] = ffma2(p2, lo2(q2), Ec2[4]); Ec2[5] = ffma2(p2, hi2(q2), Ec2[5]);
        }
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float2 w = mw[c0 + i];
          const float pf = exp_neg(s[i] - w.x) * w.y;
          const float2 p2 = make_float2(pf, pf);
          const float* uj = qf + (c0 + i) * 12;
          const float4 u0 = ld4(uj), u1 = ld4(uj + 4), u2 = ld4(uj + 8);
          Ec2[0] = ffma2(p2, lo2(u0), Ec2[0]); Ec2[1] = ffma2(p2, hi2(u0), Ec2[1]);
          Ec2[2] = ffma2(p2, lo2(u1), Ec2[2]); Ec2[3] = ffma2(p2, hi2(u1), Ec2[3]);
          Ec2[4] = ffma2(p2, lo2(u2), Ec2[4]); Ec2[5] = ffma2(p2, hi2(u2), Ec2[5]);
        }
      }
      if (c0 & 8) {   // every second chunk (the last chunk of a tile's half is an odd one: 64 columns = 8 chunks)
        if (!BWD) { l += (double)lc; lc = 0.f; }
#pragma unroll
        for (int k = 0; k < 6; ++k) { E[2 * k] += (double)Ec2[k].x; E[2 * k + 1] += (double)Ec2[k].y; Ec2[k] = make_float2(0.f, 0.f); }
      }
    }
    // a tile that ended on an even chunk (ragged last tile) leaves partials behind
    if (!BWD) { l += (double)lc; lc = 0.f; }
#pragma unroll
    for (int k = 0; k < 6; ++k) { E[2 * k] += (double)Ec2[k].x; E[2 * k + 1] += (double)Ec2[k].y; Ec2[k] = make_float2(0.f, 0.f); }
    if (j + 1 < ntile) store_tile(j + 1);
    fence_smem_to_async();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
  }
  // ---- merge the two column halves of each row, lower half first
  if (half == 1) {
    double* mine = MG + r * 14;
    mine[0] = (double)m; mine[1] = l;
#pragma unroll
    for (int t = 0; t < 12; ++t) mine[2 + t] = E[t];
  }
  __syncthreads();   // also: the raw scores the upper half wrote to pw are visible to the lower half's thread of the row
  if (half == 0 && live) {
    const double* o = MG + r * 14;
    const int64_t g = (int64_t)b * a.N + n;
    if (BWD) {
      float* d = a.dq + so + g * 12;
#pragma unroll
      for (int t = 0; t < 12; ++t) d[t] = (float)(-(E[t] + o[2 + t]));
    } else {
      const float mk = (float)o[0];
      const float M = fmaxf(m, mk);
      const double s0 = (m == -INFINITY) ? 0.0 : (double)expf(m - M), s1 = (mk == -INFINITY) ? 0.0 : (double)expf(mk - M);
      l = l * s0 + o[1] * s1;
      const double invd = 1.0 / l;
      a.m[mo + g] = M;
      reinterpret_cast<double*>(a.l + mo)[g] = l;
#pragma unroll
      for (int t = 0; t < 12; ++t) reinterpret_cast<double*>(a.E + mo)[g * 12 + t] = (E[t] * s0 + o[2 + t] * s1) * invd;
      // masked attention at the adjacency non-zeros of this row (attention.py:36: att * adjacency)
      const float inv = (float)invd;
      const int ea = a.row_ptr[n], eb = a.row_ptr[n + 1];
      for (int q = ea; q < eb; ++q) pwb[q] = expf(pwb[q] - M) * inv * a.val[q];
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tm, 256);
}

bool score_tc32_supported(const Plan& p) {
  const char* off = getenv("MSGAT_TC_DISABLE");   // test switch: bit 4 = fp32-grade tensor-core score passes
  if (off && (atoi(off) & 16)) return false;
  return p.precision == MSGAT_PREC_FP32 && p.Os > 0;
}

static void s3_fill(S3Args& a, const Plan& p) {
  a.N = p.N; a.nnz = p.nnz; a.sv_sm = p.saved_total; a.sc_sm = p.scratch_sm;
  for (int m = 0; m < MSGAT_MAX_GROUP; ++m) a.Wg[m] = nullptr;
}

int launch_sstats_tc32(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, float* saved, void* stream) {
  S3Args a;
  s3_fill(a, p);
  for (int m = 0; m < p.M; ++m) a.Wg[m] = prm[m].sa_Wg;
  a.qs = saved + p.o_qs; a.val = adj->val; a.row_ptr = adj->row_ptr; a.col_idx = adj->col_idx;
  a.m = saved + p.o_m; a.l = saved + p.o_l; a.E = saved + p.o_E; a.U = saved + p.o_U; a.pw = saved + p.o_pw;
  a.delta = nullptr; a.dq = nullptr;
  MSGAT_LAUNCH(k_score_tc32<0>, dim3((p.N + S3_ROWS - 1) / S3_ROWS, p.B, p.M), dim3(S3_THREADS), score_tc32_smem(), stream, a);
  return 0;
}

int launch_scol_tc32(const Plan& p, const float* saved, const float* delta, float* dq, void* stream) {
  S3Args a;
  s3_fill(a, p);
  float* sv = const_cast<float*>(saved);
  a.qs = sv + p.o_qs; a.val = nullptr; a.row_ptr = nullptr; a.col_idx = nullptr;
  a.m = sv + p.o_m; a.l = sv + p.o_l; a.E = nullptr; a.U = sv + p.o_U; a.pw = nullptr;
  a.delta = delta; a.dq = dq;
  MSGAT_LAUNCH(k_score_tc32<1>, dim3((p.N + S3_ROWS - 1) / S3_ROWS, p.B, p.M), dim3(S3_THREADS), score_tc32_smem(), stream, a);
  return 0;
}
