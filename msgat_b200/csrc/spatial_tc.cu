// spatial_tc.cu -- the N x N score passes of GraphAttention (attention.py:32-36) on tcgen05, for MSGAT_PREC_TF32.
//
// The dense part of the spatial attention is two products per tile of the N x N score matrix, exactly the shape of
// an attention kernel with a head dimension of 12 (q = k, attention.py:34):
//   forward  (k_score_tc<0>, replaces k_sstats):   S = U q^T  (U = q Wg),  P = exp(S - m_row),  l = rowsum P,  E = P q / l
//   backward (k_score_tc<1>, replaces k_scol):     S^T = q U^T,  P' = exp(S^T - m_col) * (delta_col / l_col),  dq = -P' U
// Both products run on the tensor core (kind::tf32, accumulators in TMEM): the score tile S[128 x 128] is one MMA group
// with K = 16 (12 zero padded), P goes from registers to a shared-memory A tile and the second product accumulates
// E / dq [128 x 16] in TMEM over all column tiles.  The CUDA cores keep what is not a contraction: one FFMA + one
// MUFU.EX2 + the TF32 rounding per score.  The forward makes two sweeps over the columns (row maxima first; the
// score MMAs are cheap enough to repeat), so that no accumulator is ever rescaled.
//
// Operands are rounded to TF32 (nearest) when they are staged, so every product is exact in fp32 and the scores at the
// adjacency non-zeros -- evaluated on the CUDA cores from the same rounded operands -- agree with the tensor core's to
// the rounding of an fp32 sum: pw, m and l describe ONE softmax.  P is rounded before it is summed into l, so that E
// is a convex combination of q rows with exactly the weights P / l.
//
// A CTA owns 128 rows of one (sample, component); 256 threads: warp w reads TMEM lanes 32 (w % 4).., the two warps of a
// lane quarter split the 128 columns of a tile.  256 TMEM columns and ~96 KB of shared memory per CTA: two CTAs per
// SM overlap each other's MMA and exp phases.  K-major operands in the unswizzled core-matrix layout (tc_common.cuh).
#include <math.h>
#include <stdlib.h>
#include "kernels.h"
#include "tc_common.cuh"

#define ST_ROWS 128
#define ST_COLS 128
#define ST_THREADS 256
#define ST_PGS (4096 + 128)   // row-group stride of the P tile: 32 core matrices + 128 bytes, so that the four row groups
                              // a warp stores to fall on different banks
#define ST_LOG2E 1.4426950408889634f

struct STcArgs {   // grouped like SStatsArgs / SColArgs: blockIdx.z = component
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

MSGAT_DEV float st_tf32(float x) { return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u); }
MSGAT_DEV float4 st_tf32x4(const float4& v) { return make_float4(st_tf32(v.x), st_tf32(v.y), st_tf32(v.z), st_tf32(v.w)); }
#ifdef MSGAT_EMU
MSGAT_DEV float st_ex2(float x) { return exp2f(x); }
#else
MSGAT_DEV float st_ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
#endif

static inline size_t score_tc_smem() {
  return 3 * 8192 + 16 * ST_PGS + 2 * ST_COLS * 8 + 3 * ST_ROWS * 4 + 144 * 4 + 64 + 1024;
}

template <int BWD>
MSGAT_KERNEL MSGAT_LB(ST_THREADS, 2) k_score_tc(STcArgs a) {
  MSGAT_DYN_SMEM(smem_raw);
  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // aligned inside the shared window: the pointer stays a shared-memory pointer (LDS / STS, not generic LD / ST)
  unsigned char* RA = smem;                       // row block  [128][16]   (forward: U, backward: q), TF32
  unsigned char* CB = RA + 8192;                  // column tile [128][16]  (forward: q, backward: U)
  unsigned char* CT = CB + 8192;                  // its transpose [16][128]
  unsigned char* PA = CT + 8192;                  // P tile [128][128], row groups ST_PGS apart
  float2* MW = reinterpret_cast<float2*>(PA + 16 * ST_PGS);   // [2][128] backward: (m log2e, delta / l) of the tile's columns
  float* RM = reinterpret_cast<float*>(MW + 2 * ST_COLS);     // [2][128] row maxima of the two column halves
  float* RL = RM + 2 * ST_ROWS;                               // [128] row sums of the upper column half
  float* WG = RL + ST_ROWS;                                   // [144]
  uint64_t* bar = reinterpret_cast<uint64_t*>(WG + 144);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 2);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int r = (warp & 3) * 32 + lane, half = warp >> 2, b = blockIdx.y, n0 = blockIdx.x * ST_ROWS;
  const int64_t mo = (int64_t)blockIdx.z * a.sv_sm, so = (int64_t)blockIdx.z * a.sc_sm;
  const float* qb = a.qs + mo + (int64_t)b * a.N * 12;
  float* Ub = a.U + mo + (int64_t)b * a.N * 12;
  const float* rowsrc = BWD ? qb : nullptr;       // forward computes its row block (U) itself
  const float* colsrc = BWD ? Ub : qb;
  const int ntile = (a.N + ST_COLS - 1) / ST_COLS;
  const uint32_t barS = smem_u32(bar);

  if (!BWD) for (int i = tid; i < 144; i += ST_THREADS) WG[i] = a.Wg[blockIdx.z][i];
  // zero the K padding (k = 12..15) of the row block and the column tile, and the rows f = 12..15 of the transpose
  for (int i = tid; i < 2 * ST_ROWS; i += ST_THREADS) {
    unsigned char* t = (i < ST_ROWS ? RA : CB) + core_off(i & (ST_ROWS - 1), 12, 512u);
    st4(reinterpret_cast<float*>(t), f4zero());
  }
  for (int i = tid; i < 4 * ST_COLS / 4; i += ST_THREADS) st4(reinterpret_cast<float*>(CT + core_off(12 + i / (ST_COLS / 4), (i % (ST_COLS / 4)) * 4, 4096u)), f4zero());
  __syncthreads();
  // ---- the row block
  if (tid < ST_ROWS) {
    const int n = n0 + tid;
    float v[12];
    if (BWD) {
      for (int t = 0; t < 12; ++t) v[t] = n < a.N ? rowsrc[(int64_t)n * 12 + t] : 0.f;
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
    for (int p4 = 0; p4 < 12; p4 += 4)
      st4(reinterpret_cast<float*>(RA + core_off(tid, p4, 512u)), st_tf32x4(make_float4(v[p4], v[p4 + 1], v[p4 + 2], v[p4 + 3])));
  }
  if (tid == 0) { mbar_init(barS, 1); mbar_fence_init(); }
  if (warp == 0) tmem_alloc(smem_u32(tmem_slot), 256);

  // items of a column tile: 128 rows x 3 quads of the source rows; item x -> (row x / 3, quad x % 3)
  float4 pre[2];
  float2 premw = make_float2(0.f, 0.f);
  auto prefetch = [&](int j) {
    const int j0 = j * ST_COLS;
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int x = tid + u * ST_THREADS, rr = x / 3, p4 = (x - rr * 3) * 4;
      pre[u] = (x < ST_COLS * 3 && j0 + rr < a.N) ? ld4(colsrc + (int64_t)(j0 + rr) * 12 + p4) : f4zero();
    }
    if (BWD && tid < ST_COLS) {
      const int n = j0 + tid;
      premw = make_float2(0.f, 0.f);   // padded columns: exp(0 - 0) * 0
      if (n < a.N) {
        const int64_t g = (int64_t)b * a.N + n;
        premw.x = a.m[mo + g] * ST_LOG2E;
        premw.y = (float)((double)a.delta[so + g] / reinterpret_cast<const double*>(a.l + mo)[g]);
      }
    }
  };
  auto store_tile = [&](int j, bool transpose_current) {
    // the transpose of the CURRENT tile (operand of the second product) is read back from CB before CB takes tile j + 1;
    // every thread touches the same items in both steps, so no barrier is needed in between
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int x = tid + u * ST_THREADS, rr = x / 3, p4 = (x - rr * 3) * 4;
      if (x < ST_COLS * 3) {
        float* cb = reinterpret_cast<float*>(CB + core_off(rr, p4, 512u));
        if (transpose_current) {
          const float4 c = ld4(cb);
          *reinterpret_cast<float*>(CT + core_off(p4 + 0, rr, 4096u)) = c.x;
          *reinterpret_cast<float*>(CT + core_off(p4 + 1, rr, 4096u)) = c.y;
          *reinterpret_cast<float*>(CT + core_off(p4 + 2, rr, 4096u)) = c.z;
          *reinterpret_cast<float*>(CT + core_off(p4 + 3, rr, 4096u)) = c.w;
        }
        st4(cb, st_tf32x4(pre[u]));
      }
    }
    if (BWD && tid < ST_COLS) MW[((j + 1) & 1) * ST_COLS + tid] = premw;
  };
  (void)store_tile;
  prefetch(0);
  store_tile(-1, false);
  fence_smem_to_async();
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tm = *tmem_slot;
  const uint32_t tS = tmem_addr(tm, (warp & 3) * 32, half * 64), tE = tmem_addr(tm, (warp & 3) * 32, ST_COLS);
  const uint32_t idS = umma_idesc_tf32(128, ST_COLS, 0, 0), idE = umma_idesc_tf32(128, 16, 0, 0);
  uint32_t phase = 0;
  auto issue_scores = [&]() {
    for (int ks = 0; ks < 2; ++ks)
      umma_tf32(tm, umma_desc(smem_u32(RA) + ks * 256u, 128u, 512u, 0), umma_desc(smem_u32(CB) + ks * 256u, 128u, 512u, 0), idS, ks > 0);
    umma_commit(barS);
  };

  float mrow = -INFINITY;
  if (!BWD) {
    // ---- sweep 1: row maxima
    for (int j = 0; j < ntile; ++j) {
      if (tid == 0) issue_scores();
      __syncwarp();
      prefetch(j + 1 < ntile ? j + 1 : 0);   // after the last tile: tile 0 again, for sweep 2
      mbar_wait(barS, phase); phase ^= 1;
      tc_fence_after_sync();
      const int cmax = a.N - j * ST_COLS - half * 64;   // live columns of this thread's half
#pragma unroll
      for (int c0 = 0; c0 < 64; c0 += 16) {
        uint32_t v0[8], v1[8];
        tmem_ld8_nowait(tS + c0, v0);
        tmem_ld8_nowait(tS + c0 + 8, v1);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          if (c0 + i < cmax) mrow = fmaxf(mrow, __uint_as_float(v0[i]));
          if (c0 + 8 + i < cmax) mrow = fmaxf(mrow, __uint_as_float(v1[i]));
        }
      }
      store_tile(j, false);
      fence_smem_to_async();
      tc_fence_before_sync();
      __syncthreads();
      tc_fence_after_sync();
    }
    RM[half * ST_ROWS + r] = mrow;
    __syncthreads();
    mrow = fmaxf(RM[r], RM[ST_ROWS + r]);
  }
  const float mrow2 = mrow * ST_LOG2E;

  // ---- main sweep: P tile -> second product
  float lrow = 0.f;
  for (int j = 0; j < ntile; ++j) {
    if (tid == 0) issue_scores();
    __syncwarp();
    if (j + 1 < ntile) prefetch(j + 1);
    mbar_wait(barS, phase); phase ^= 1;   // the scores of tile j; the MMAs run in order, so the second product of tile j - 1 is done too
    tc_fence_after_sync();
    const int cmax = a.N - j * ST_COLS - half * 64;
    const float2* mw = MW + (j & 1) * ST_COLS + half * 64;
#pragma unroll
    for (int c0 = 0; c0 < 64; c0 += 16) {
      uint32_t v0[8], v1[8];
      tmem_ld8_nowait(tS + c0, v0);
      tmem_ld8_nowait(tS + c0 + 8, v1);
      tmem_ld_wait();
      float p[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const float s = __uint_as_float(i < 8 ? v0[i] : v1[i - 8]);
        float e;
        if (BWD) {
          const float2 w = mw[c0 + i];
          e = st_ex2(fmaf(s, ST_LOG2E, -w.x)) * w.y;
        } else {
          e = c0 + i < cmax ? st_ex2(fmaf(s, ST_LOG2E, -mrow2)) : 0.f;
        }
        p[i] = st_tf32(e);
        lrow += p[i];
      }
#pragma unroll
      for (int i = 0; i < 16; i += 4)
        st4(reinterpret_cast<float*>(PA + core_off(r, half * 64 + c0 + i, ST_PGS)), make_float4(p[i], p[i + 1], p[i + 2], p[i + 3]));
    }
    store_tile(j, true);
    fence_smem_to_async();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    if (tid == 0) {
      for (int ks = 0; ks < ST_COLS / 8; ++ks)
        umma_tf32(tm + ST_COLS, umma_desc(smem_u32(PA) + ks * 256u, 128u, ST_PGS, 0), umma_desc(smem_u32(CT) + ks * 256u, 128u, 4096u, 0), idE,
                  (j > 0 || ks > 0) ? 1u : 0u);
    }
    __syncwarp();
  }
  if (tid == 0) umma_commit(barS);
  __syncwarp();
  if (half == 1) RL[r] = lrow;
  mbar_wait(barS, phase);
  tc_fence_after_sync();
  __syncthreads();
  const int n = n0 + r;
  uint32_t v0[8], v1[8];
  if (half == 0) {   // warp-uniform: tcgen05.ld is a .sync.aligned instruction, every lane of the warp has to execute it
    tmem_ld8_nowait(tE, v0);
    tmem_ld8_nowait(tE + 8, v1);
    tmem_ld_wait();
  }
  if (half == 0 && n < a.N) {
    if (BWD) {
      float* d = a.dq + so + ((int64_t)b * a.N + n) * 12;
      for (int t = 0; t < 12; ++t) d[t] = -__uint_as_float(t < 8 ? v0[t] : v1[t - 8]);
    } else {
      const float l = lrow + RL[r];
      const double invd = 1.0 / (double)l;
      const int64_t g = (int64_t)b * a.N + n;
      a.m[mo + g] = mrow;
      reinterpret_cast<double*>(a.l + mo)[g] = (double)l;
      for (int t = 0; t < 12; ++t) reinterpret_cast<double*>(a.E + mo)[g * 12 + t] = (double)__uint_as_float(t < 8 ? v0[t] : v1[t - 8]) * invd;
      // masked attention at the adjacency non-zeros of this row (attention.py:36), from the same rounded operands
      float u[12];
      for (int p4 = 0; p4 < 12; p4 += 4) {
        const float4 c = ld4(reinterpret_cast<const float*>(RA + core_off(r, p4, 512u)));
        u[p4] = c.x; u[p4 + 1] = c.y; u[p4 + 2] = c.z; u[p4 + 3] = c.w;
      }
      const float inv = (float)invd;
      const int e1 = a.row_ptr[n + 1];
      for (int e = a.row_ptr[n]; e < e1; ++e) {
        const float* qi = qb + (int64_t)a.col_idx[e] * 12;
        float s = 0.f;
        for (int t = 0; t < 12; ++t) s = fmaf(u[t], st_tf32(qi[t]), s);
        a.pw[mo + (int64_t)b * a.nnz + e] = st_ex2(fmaf(s, ST_LOG2E, -mrow2)) * inv * a.val[e];
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tm, 256);
}

bool score_tc_supported(const Plan& p) {
  const char* off = getenv("MSGAT_TC_DISABLE");   // test switch: bit 3 = tensor-core score passes
  if (off && (atoi(off) & 8)) return false;
  return p.precision == MSGAT_PREC_TF32 && p.Os > 0;
}

int launch_sstats_tc(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, float* saved, void* stream) {
  STcArgs a;
  a.N = p.N; a.nnz = p.nnz; a.sv_sm = p.saved_total; a.sc_sm = 0;
  for (int m = 0; m < p.M; ++m) a.Wg[m] = prm[m].sa_Wg;
  a.qs = saved + p.o_qs; a.val = adj->val; a.row_ptr = adj->row_ptr; a.col_idx = adj->col_idx;
  a.m = saved + p.o_m; a.l = saved + p.o_l; a.E = saved + p.o_E; a.U = saved + p.o_U; a.pw = saved + p.o_pw;
  a.delta = nullptr; a.dq = nullptr;
  MSGAT_LAUNCH(k_score_tc<0>, dim3((p.N + ST_ROWS - 1) / ST_ROWS, p.B, p.M), dim3(ST_THREADS), score_tc_smem(), stream, a);
  return 0;
}

int launch_scol_tc(const Plan& p, const float* saved, const float* delta, float* dq, void* stream) {
  STcArgs a;
  a.N = p.N; a.nnz = p.nnz; a.sv_sm = p.saved_total; a.sc_sm = p.scratch_sm;
  for (int m = 0; m < MSGAT_MAX_GROUP; ++m) a.Wg[m] = nullptr;
  a.qs = const_cast<float*>(saved) + p.o_qs; a.val = nullptr; a.row_ptr = nullptr; a.col_idx = nullptr;
  a.m = const_cast<float*>(saved) + p.o_m; a.l = const_cast<float*>(saved) + p.o_l; a.E = nullptr;
  a.U = const_cast<float*>(saved) + p.o_U; a.pw = nullptr;
  a.delta = delta; a.dq = dq;
  MSGAT_LAUNCH(k_score_tc<1>, dim3((p.N + ST_ROWS - 1) / ST_ROWS, p.B, p.M), dim3(ST_THREADS), score_tc_smem(), stream, a);
  return 0;
}
