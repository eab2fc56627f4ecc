// tc_dgrad.cu -- the node-local data gradients of a layer-2 MEAM on tcgen05 (3xTF32, see
// tc_fwd3.cu), the backward counterpart of k_tc_apply3 and the replacement of k_bwd_main for
// that shape.  With G the masked upstream gradient [cacn | tacn | gacn] (O = 3b rows) and
// positions p = (node, t) on the TMEM lanes:
//   D_x[p, c]  = sum_o G[o,p] Wres[o,c]                               d(residual)          msgat.py:116,131
//   D_y[p, k'] = sum_o G_t[o,p] [W2now | W2past][o,k']                second conv, transposed  msgat.py:71-72
//   dY1[k,t]   = D_y[(n,t),k] + D_y[(n,t+d2), b+k]
//   dA_now[k,i] = sum_t Pt[t,i] dY1[k,t],  dA_past[k,i] = sum_{t>=d1} Pt[t-d1,i] dY1[k,t]    attention.py:66 transposed
//   GZ[o,i]    = sum_n A[n,i] G_s[o,n]                                sparse transposed aggregation  attention.py:36
//   D_h[p, c]  = sum_o G_c[o,p] Mc[o,c] + sum_j [dA_now;dA_past;GZ][j,p] [W1now;W1past;Ws][j,c]
//   dx         = LayerNorm_T^T(D_h) + D_x   (masked by x > 0 when x is a ReLU output)       msgat.py:122
// plus everything k_bwd_main hands on: dA_now/dA_past/GZ for k_tc_wgrad, bias / LayerNorm-affine /
// dPt partials, and the spatial attention's dS at the adjacency non-zeros with its row sums delta.
// dPt[t,i] = sum_{k,n} dY1[k,n,t] a_now[k,n,i] + dY1[k,n,t+d1] a_past[k,n,i] uses the first-conv
// taps a = W1 h kept by the forward (saved A2).
//
// Tile = 8 nodes = 96 positions (one M = 128 block).  The G tile and later the [dA|GZ] tile are
// the MN-major A operand (hi, lo); the four weight matrices are resident B operands.  Row-wise
// work (temporal mixes, LayerNorm backward) goes through shared-memory staging rows of 12.
#include <stdlib.h>
#include "kernels.h"
#include "tc_common.cuh"
#include "tile_ops.cuh"

#define TD_NT 8
#define TD_P 96
#define TD_THREADS 512
#define TD_PITCH 100
#define TD_NGR 32      // 9-thread groups of the dPt contraction
#define TD_ISSUER 320  // MMA-issuing thread: warp 10 has no work inside the two MMA windows, so nothing waits for the issue
#define TD_ISSUER2 352 // second issuing thread (warp 11) of the first window: D_x (27 of its 45 MMAs) on its own barrier

struct TcDgradArgs {   // grouped: blockIdx.y = component; inputs x_sm / g_sm apart, saved slices sv_sm, scratch slices sc_sm, dx dx_sm apart
  const float* x; int64_t x_sb, x_sc, x_sm, g_sm, sv_sm, sc_sm, dx_sm;
  const float* G;
  int C, N, O, b, ntile, J, B, d1, d2, nnz, x_is_relu, nbias;
  const float *ln_w[MSGAT_MAX_GROUP], *ln_b[MSGAT_MAX_GROUP], *res_w[MSGAT_MAX_GROUP], *c1_w[MSGAT_MAX_GROUP], *c2_w[MSGAT_MAX_GROUP],
      *sa_W[MSGAT_MAX_GROUP];
  const float *Mc, *Pt, *Z, *pw, *A2;
  const int32_t *row_ptr, *col_idx, *csc_ptr, *csc_row, *csc_eid;
  float *dx, *tcA, *delta, *dSnz, *bias_part, *ln_part;
  double* dPt_part;
};

static inline size_t td_pre_floats(int b) { const size_t a = (size_t)2 * b * TD_PITCH, c = (size_t)TD_NGR * 144; return a > c ? a : c; }

static size_t tc_dgrad_smem(int O, int b) {
  size_t s = 0;
  s += 2 * (size_t)3 * O * 128;            // GH, GL: G tile, then [dA_now | dA_past | GZ]; later the D_h staging rows
  s += 4 * (size_t)3 * O * 128;            // WXH, WXL (Wres), W3H, W3L ([W1now; W1past; Ws])
  s += 2 * (size_t)3 * b * 128;            // WMH, WML (Mc of the sample)
  s += 2 * (size_t)2 * b * 128;            // W2H, W2L
  s += td_pre_floats(b) * 4;               // PRE: D_y rows, later the dPt group partials
  s += (size_t)b * TD_PITCH * 4;           // DY: dY1 rows
  s += (144 * 2 + 24 + 16 + 16 * 24 + 8) * 4;
  return s + 1024;
}

TC_DEV float td_rna(float x) {
  uint32_t u;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x));
  return __uint_as_float(u);
}
TC_DEV void td_split4(unsigned char* hi, unsigned char* lo, uint32_t off, float4 v) {
  float4 h = make_float4(td_rna(v.x), td_rna(v.y), td_rna(v.z), td_rna(v.w));
  st4(reinterpret_cast<float*>(hi + off), h);
  st4(reinterpret_cast<float*>(lo + off), make_float4(td_rna(v.x - h.x), td_rna(v.y - h.y), td_rna(v.z - h.z), td_rna(v.w - h.w)));
}
TC_DEV void td_split1(unsigned char* hi, unsigned char* lo, uint32_t off, float v) {
  const float h = td_rna(v);
  *reinterpret_cast<float*>(hi + off) = h;
  *reinterpret_cast<float*>(lo + off) = td_rna(v - h);
}
// acc (+)= Ah Bh + Ah Bl + Al Bh for one k-step (lo*lo, 2^-22 of the product, changed no measured error and is not issued); MN-major operands, 128 B swizzle with 32 B atoms
// The descriptors of k-step 0 are built once; k-step ks (8 contraction rows = 1024 B) adds ks * 64 to
// the start-address field (16-byte units).
TC_DEV void td_mma3(uint32_t d, uint64_t dah, uint64_t dal, uint64_t dbh, uint64_t dbl, uint32_t ka, uint32_t kb, uint32_t idesc,
                    uint32_t acc) {
  const uint64_t oa = (uint64_t)(ka * 64u), ob = (uint64_t)(kb * 64u);
  // hi*hi into the main accumulator, the three correction terms (2^-11 of it) into a second one 256 columns further:
  // the tensor core truncates when it adds into the fp32 accumulator, so the error grows with the number of
  // accumulations at full magnitude (see tc_fwd3.cu); the readers add the two accumulators in fp32.
  umma_tf32(d, dah + oa, dbh + ob, idesc, acc);
  umma_tf32(d + 256u, dah + oa, dbl + ob, idesc, acc);
  umma_tf32(d + 256u, dal + oa, dbh + ob, idesc, 1u);
}
TC_DEV void td_ld8_sum(uint32_t addr, float (&v)[8]) {
  uint32_t r0[8], r1[8];
  tmem_ld8_nowait(addr, r0);
  tmem_ld8_nowait(addr + 256u, r1);
  tmem_ld_wait();
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r0[i]) + __uint_as_float(r1[i]);
}

__global__ void __launch_bounds__(TD_THREADS, 1) k_tc_dgrad(TcDgradArgs a) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // aligned inside the shared window: the pointer stays a shared-memory pointer (LDS / STS, not generic LD / ST)
  const int C = a.C, b = a.b, O = a.O, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, m = blockIdx.y;
  const int64_t svo = (int64_t)m * a.sv_sm, sco = (int64_t)m * a.sc_sm;
  const uint32_t cbO = (uint32_t)O * 128u, cbM = (uint32_t)b * 128u;
  unsigned char* GH = smem;
  unsigned char* GL = GH + 3 * cbO;
  unsigned char* WXH = GL + 3 * cbO;
  unsigned char* WXL = WXH + 3 * cbO;
  unsigned char* W3H = WXL + 3 * cbO;
  unsigned char* W3L = W3H + 3 * cbO;
  unsigned char* WMH = W3L + 3 * cbO;
  unsigned char* WML = WMH + 3 * cbM;
  unsigned char* W2H = WML + 3 * cbM;
  unsigned char* W2L = W2H + 2 * cbM;
  float* PRE = reinterpret_cast<float*>(W2L + 2 * cbM);   // [2b][TD_PITCH]
  float* DY = PRE + (2 * b * TD_PITCH > TD_NGR * 144 ? 2 * b * TD_PITCH : TD_NGR * 144);   // [b][TD_PITCH]
  float* Pt = DY + b * TD_PITCH;                          // [144]
  float* Ptd = Pt + 144;                                  // Pt shifted down by d1 rows
  float* LW = Ptd + 144;                                  // [24]
  float* DL = LW + 24;                                    // [16] delta halves
  float* RED = DL + 16;                                   // [16][24] final LayerNorm-affine reduction
  uint64_t& bar = *reinterpret_cast<uint64_t*>(RED + 16 * 24);
  uint32_t& tmem_slot = *reinterpret_cast<uint32_t*>(RED + 16 * 24 + 2);
  uint64_t& bar2 = *reinterpret_cast<uint64_t*>(RED + 16 * 24 + 4);
  float* STG = reinterpret_cast<float*>(GH);              // [C][TD_PITCH] D_h rows, aliases the A tile once it is consumed

  // ---- one-time staging: weights as B operands [K][N] (N contiguous), hi / lo, zero padded
  for (int i = tid; i < O * 96; i += TD_THREADS) {
    const int k = i / 96, n = i - k * 96;
    const uint32_t off = sw128b32_off(k, n, cbO);
    td_split1(WXH, WXL, off, n < C ? a.res_w[m][k * C + n] : 0.f);
    float w = 0.f;
    if (n < C) {
      if (k < b) w = a.c1_w[m][(k * C + n) * 2 + 1];
      else if (k < 2 * b) w = a.c1_w[m][((k - b) * C + n) * 2 + 0];
      else w = a.sa_W[m][(k - 2 * b) * C + n];
    }
    td_split1(W3H, W3L, off, w);
  }
  for (int i = tid; i < b * 64; i += TD_THREADS) {
    const int k = i / 64, n = i - k * 64;   // k = output channel o of the second conv (contraction), n = its input channel (+ b: past tap)
    float w = 0.f;
    if (n < b) w = a.c2_w[m][(k * b + n) * 2 + 1];
    else if (n < 2 * b) w = a.c2_w[m][(k * b + (n - b)) * 2 + 0];
    td_split1(W2H, W2L, sw128b32_off(k, n, cbM), w);
  }
  if (tid < 12) { LW[tid] = a.ln_w[m][tid]; LW[12 + tid] = a.ln_b[m][tid]; }
  if (tid == 0) { mbar_init(smem_u32(&bar), 1); mbar_init(smem_u32(&bar2), 1); mbar_fence_init(); }
  if (warp == 0) tmem_alloc(smem_u32(&tmem_slot), 512);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tm = warp_uniform(tmem_slot);   // uniform register: no per-MMA R2UR loop in the issue thread (tc_common.cuh)
  const uint32_t Dx = tm, Dh = tm + 96, Dy = tm + 192;
  const uint64_t dGH = umma_desc(smem_u32(GH), cbO, 512u, 1), dGL = umma_desc(smem_u32(GL), cbO, 512u, 1);
  const uint64_t dWXH = umma_desc(smem_u32(WXH), cbO, 512u, 1), dWXL = umma_desc(smem_u32(WXL), cbO, 512u, 1);
  const uint64_t dW3H = umma_desc(smem_u32(W3H), cbO, 512u, 1), dW3L = umma_desc(smem_u32(W3L), cbO, 512u, 1);
  const uint64_t dWMH = umma_desc(smem_u32(WMH), cbM, 512u, 1), dWML = umma_desc(smem_u32(WML), cbM, 512u, 1);
  const uint64_t dW2H = umma_desc(smem_u32(W2H), cbM, 512u, 1), dW2L = umma_desc(smem_u32(W2L), cbM, 512u, 1);
  const uint32_t idescC = umma_idesc_tf32(128, (C + 15) & ~15, 1, 1), idescY = umma_idesc_tf32(128, 2 * b, 1, 1);
  uint32_t phase = 0, phase2 = 0;
  const int l = tid & 127, quarter = warp >> 2;   // l = TMEM lane = position; the 4 warpgroups split the column groups
  const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;

  // GZ gather items on the top threads: 4 channels x 4 positions each
  const int gz_i = tid - (TD_THREADS - (b / 4) * TD_NT * 3);
  const bool gz_on = gz_i >= 0;
  const int gz_ch = gz_i / (TD_NT * 3), gz_pq = gz_i - gz_ch * (TD_NT * 3), gz_nl = gz_pq / 3, gz_t0 = (gz_pq - gz_nl * 3) * 4, gz_o0 = gz_ch * 4;
  uint32_t* MSK = reinterpret_cast<uint32_t*>(DY);   // [C][8] x > 0 bits of a row, over the dY1 rows once dPt has read them

  // A CTA serves one component (its weights are staged once) and walks the logical blocks (sample bi, slot j) lb = bi * J + j
  // with lb = blockIdx.x, blockIdx.x + gridDim.x, ...: block (bi, j) owns the tiles j, j + J, ... of its sample and its own
  // slice of the per-block partial sums, exactly as when every logical block was a CTA of its own launch.
  for (int blk = blockIdx.x; blk < a.J * a.B; blk += gridDim.x) {
  const int bi = blk / a.J, jslot = blk - bi * a.J;
  // per-sample operands: Mc as a B operand, the temporal attention matrix and its d1-shifted copy (the MMAs of the previous
  // block have completed: every tile ends with a wait on its second MMA group)
  for (int i = tid; i < b * 96; i += TD_THREADS) {
    const int k = i / 96, n = i - k * 96;
    td_split1(WMH, WML, sw128b32_off(k, n, cbM), n < C ? a.Mc[svo + ((int64_t)bi * b + k) * C + n] : 0.f);
  }
  for (int i = tid; i < 144; i += TD_THREADS) {
    const int t = i / 12, j = i - t * 12;
    Pt[i] = a.Pt[svo + (int64_t)bi * 144 + i];
    Ptd[i] = t >= a.d1 ? a.Pt[svo + (int64_t)bi * 144 + (t - a.d1) * 12 + j] : 0.f;
  }
  __syncthreads();
  const float* xb = a.x + (int64_t)m * a.x_sm + bi * a.x_sb;
  const float* Gb = a.G + (int64_t)m * a.g_sm + (int64_t)bi * O * a.N * 12;
  const float* A2b = a.A2 + svo + (int64_t)bi * 2 * b * a.N * 12;
  float* tcAb = a.tcA + sco + (int64_t)bi * 3 * b * a.N * 12;
  float* dxb = a.dx + (int64_t)m * a.dx_sm;
  float bsG[2] = {0.f, 0.f}, bsY = 0.f, lnw_acc[12], lnb_acc[12];
  double pt_acc = 0.0;
#pragma unroll
  for (int t = 0; t < 12; ++t) { lnw_acc[t] = 0.f; lnb_acc[t] = 0.f; }

  for (int tile = jslot; tile < a.ntile; tile += a.J) {
    const int n0 = tile * TD_NT;
    // the thread's first G row is requested before the prefetch instructions below queue up in front of it
    float4 g0a = f4zero(), g0b = f4zero(), g0c = f4zero(), g1a = f4zero(), g1b = f4zero(), g1c = f4zero();
    {
      const int o = tid / TD_NT, nl = tid - o * TD_NT, n = n0 + nl, o1 = o + TD_THREADS / TD_NT;
      if (n < a.N) {
        const float* s = Gb + ((int64_t)o * a.N + n) * 12; g0a = ld4(s); g0b = ld4(s + 4); g0c = ld4(s + 8);
        if (o1 < O) { const float* s1 = Gb + ((int64_t)o1 * a.N + n) * 12; g1a = ld4(s1); g1b = ld4(s1 + 4); g1c = ld4(s1 + 8); }
      }
    }
    // L2 prefetch of what this tile reads late and latency-bound: the saved first-conv taps (dPt, inside
    // the second MMA window) and the x rows of the LayerNorm backward.  One 128-byte line per thread.
    {
      const int nrow = 2 * b + C, seg = tid >> 2, part = tid & 3;   // a row segment = 8 nodes x 48 B = 384 B
      for (int sgi = seg; sgi < nrow; sgi += TD_THREADS / 4) {
        const float* base = sgi < 2 * b ? A2b + (int64_t)sgi * a.N * 12 : xb + (int64_t)(sgi - 2 * b) * a.x_sc;
        int64_t off = (int64_t)n0 * 12 + part * 32;                 // floats
        const int64_t last = (int64_t)a.N * 12 - 1;
        if (off > last) off = last;
        asm volatile("prefetch.global.L2 [%0];" ::"l"(base + off));
      }
    }
    // ... and of what the NEXT tile of this CTA reads first: its G rows and the GZ rows k_l2_gz left in tcA (the loads at the top
    // of a tile were exposed DRAM latency: 7 % of the kernel's samples)
    if (tile + a.J < a.ntile) {
      const int nrow = O + b, seg = tid >> 2, part = tid & 3;
      const int64_t n1 = (int64_t)(n0 + a.J * TD_NT) * 12;
      for (int sgi = seg; sgi < nrow; sgi += TD_THREADS / 4) {
        const float* base = sgi < O ? Gb + (int64_t)sgi * a.N * 12 : tcAb + (int64_t)(2 * b + sgi - O) * a.N * 12;
        int64_t off = n1 + part * 32;
        const int64_t last = (int64_t)a.N * 12 - 1;
        if (off > last) off = last;
        asm volatile("prefetch.global.L2 [%0];" ::"l"(base + off));
      }
    }
    // ---- 1. G rows -> the A tile (hi / lo); row sums for the bias gradients
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const int r = tid + q * TD_THREADS;
      if (r < O * TD_NT) {
        const int o = r / TD_NT, nl = r - o * TD_NT, n = n0 + nl, pp = nl * 12;
        const float4 u = q ? g1a : g0a, v = q ? g1b : g0b, w = q ? g1c : g0c;
        bsG[q] += ((u.x + u.y) + (u.z + u.w)) + ((v.x + v.y) + (v.z + v.w)) + ((w.x + w.y) + (w.z + w.w));
        td_split4(GH, GL, sw128b32_off(o, pp, cbO), u);
        td_split4(GH, GL, sw128b32_off(o, pp + 4, cbO), v);
        td_split4(GH, GL, sw128b32_off(o, pp + 8, cbO), w);
      }
    }
    fence_smem_to_async();
    __syncthreads();
    if (tid == TD_ISSUER) {
      tc_fence_after_sync();
      for (int ks = 0; ks < b / 8; ++ks) td_mma3(Dy, dGH, dGL, dW2H, dW2L, b / 8 + ks, ks, idescY, ks > 0);
      umma_commit(smem_u32(&bar));
    }
    if (tid == TD_ISSUER2) {   // the issue of an MMA costs ~40 cycles of dependent uniform-datapath work: two threads halve the window
      tc_fence_after_sync();
      for (int ks = 0; ks < b / 8; ++ks) td_mma3(Dh, dGH, dGL, dWMH, dWML, ks, ks, idescC, ks > 0);
      for (int ks = 0; ks < O / 8; ++ks) td_mma3(Dx, dGH, dGL, dWXH, dWXL, ks, ks, idescC, ks > 0);
      umma_commit(smem_u32(&bar2));
    }
    // ---- while the tensor core runs: the GZ rows of this tile, computed by k_l2_gz (transposed sparse aggregation
    // through the CSC view, attention.py:36), come in from the tcA buffer: 4 channels x 4 positions per item
    float4 gz[4] = {f4zero(), f4zero(), f4zero(), f4zero()};
    if (gz_on && n0 + gz_nl < a.N) {
#pragma unroll
      for (int j = 0; j < 4; ++j) gz[j] = ld4(tcAb + ((int64_t)(2 * b + gz_o0 + j) * a.N + n0 + gz_nl) * 12 + gz_t0);
    }
    mbar_wait(smem_u32(&bar), phase); phase ^= 1;
    tc_fence_after_sync();
    // ---- 2. D_y -> staging rows PRE[2b][.]
    {
      int g = 0;
      for (int c0 = 0; c0 < 2 * b; c0 += 8, ++g) {
        if ((g & 3) != quarter) continue;   // warp-uniform
        float v[8];
        td_ld8_sum(Dy + lane_base + c0, v);
        if (l < TD_P) {
#pragma unroll
          for (int j = 0; j < 8; ++j) PRE[(c0 + j) * TD_PITCH + l] = v[j];
        }
      }
    }
    __syncthreads();
    mbar_wait(smem_u32(&bar2), phase2); phase2 ^= 1;   // the D_x products read the G tile that phase 3 overwrites
    tc_fence_after_sync();
    // ---- 3. rows of 12: dY1, dA_now / dA_past (two threads per row), GZ gather -> the A tile again
    if (tid < 2 * b * TD_NT) {
      const int item = tid >> 1, half = tid & 1, k = item / TD_NT, nl = item - k * TD_NT, n = n0 + nl, pp = nl * 12;
      const float* pn = PRE + k * TD_PITCH + pp;
      const float* pq = PRE + (b + k) * TD_PITCH + pp;
      float v[12], o12[12], w[12];
      {
        const float4 a0 = ld4(pn), a1 = ld4(pn + 4), a2 = ld4(pn + 8), b0 = ld4(pq), b1 = ld4(pq + 4), b2 = ld4(pq + 8);
        v[0] = a0.x; v[1] = a0.y; v[2] = a0.z; v[3] = a0.w; v[4] = a1.x; v[5] = a1.y; v[6] = a1.z; v[7] = a1.w;
        v[8] = a2.x; v[9] = a2.y; v[10] = a2.z; v[11] = a2.w;
        w[0] = b0.x; w[1] = b0.y; w[2] = b0.z; w[3] = b0.w; w[4] = b1.x; w[5] = b1.y; w[6] = b1.z; w[7] = b1.w;
        w[8] = b2.x; w[9] = b2.y; w[10] = b2.z; w[11] = b2.w;
      }
      // dY1[t] = now[t] + past[t + d2]
      // (a switch over the uniform dilation keeps every index static: a select chain over a register
      // array is turned into a local-memory array by the compiler)
#define TD_SHIFT(D) case D: _Pragma("unroll") for (int t = 0; t + D < 12; ++t) v[t] += w[t + D]; break;
      switch (a.d2) {
        TD_SHIFT(1) TD_SHIFT(2) TD_SHIFT(3) TD_SHIFT(4) TD_SHIFT(5) TD_SHIFT(6) TD_SHIFT(7) TD_SHIFT(8) TD_SHIFT(9) TD_SHIFT(10)
        TD_SHIFT(11)
        default: break;
      }
#undef TD_SHIFT
      const float* P = half ? Ptd : Pt;
#pragma unroll
      for (int j = 0; j < 12; ++j) o12[j] = 0.f;
#pragma unroll
      for (int t = 0; t < 12; ++t) {
        const float4 p0 = ld4(P + t * 12), p1 = ld4(P + t * 12 + 4), p2 = ld4(P + t * 12 + 8);
        o12[0] = fmaf(p0.x, v[t], o12[0]); o12[1] = fmaf(p0.y, v[t], o12[1]); o12[2] = fmaf(p0.z, v[t], o12[2]); o12[3] = fmaf(p0.w, v[t], o12[3]);
        o12[4] = fmaf(p1.x, v[t], o12[4]); o12[5] = fmaf(p1.y, v[t], o12[5]); o12[6] = fmaf(p1.z, v[t], o12[6]); o12[7] = fmaf(p1.w, v[t], o12[7]);
        o12[8] = fmaf(p2.x, v[t], o12[8]); o12[9] = fmaf(p2.y, v[t], o12[9]); o12[10] = fmaf(p2.z, v[t], o12[10]); o12[11] = fmaf(p2.w, v[t], o12[11]);
      }
      if (!half) {
        float* dy = DY + k * TD_PITCH + pp;
        float s = 0.f;
#pragma unroll
        for (int t = 0; t < 12; ++t) s += v[t];
        st4(dy, make_float4(v[0], v[1], v[2], v[3])); st4(dy + 4, make_float4(v[4], v[5], v[6], v[7]));
        st4(dy + 8, make_float4(v[8], v[9], v[10], v[11]));
        bsY += s;
      }
      const int row = half * b + k;
      const float4 u0 = make_float4(o12[0], o12[1], o12[2], o12[3]), u1 = make_float4(o12[4], o12[5], o12[6], o12[7]),
                   u2 = make_float4(o12[8], o12[9], o12[10], o12[11]);
      td_split4(GH, GL, sw128b32_off(row, pp, cbO), u0);
      td_split4(GH, GL, sw128b32_off(row, pp + 4, cbO), u1);
      td_split4(GH, GL, sw128b32_off(row, pp + 8, cbO), u2);
      if (n < a.N) {
        float* d = tcAb + ((int64_t)row * a.N + n) * 12;
        st4(d, u0); st4(d + 4, u1); st4(d + 8, u2);
      }
    }
    if (gz_on) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        td_split4(GH, GL, sw128b32_off(2 * b + gz_o0 + j, gz_nl * 12 + gz_t0, cbO), gz[j]);
      }
    }
    // the saved first-conv taps of the dPt contraction below: the first tap's rows are requested in front of the barrier, the
    // second tap's while the first is contracted (their L2 latency was exposed inside the short second MMA window)
    const int p_blk9 = tid % 9, p_grp = tid / 9, p_ib = (p_blk9 % 3) * 4, p_n = n0 + (p_grp & 7), p_kb = p_grp >> 3;
    const bool p_on = tid < TD_NGR * 9 && p_n < a.N;
    float4 hq[2][6];
#pragma unroll
    for (int m = 0; m < 6; ++m)
      hq[0][m] = p_on && p_kb + 4 * m < b ? ld4(A2b + ((int64_t)(p_kb + 4 * m) * a.N + p_n) * 12 + p_ib) : f4zero();
    fence_smem_to_async();
    __syncthreads();
    if (tid == TD_ISSUER) {
      tc_fence_after_sync();
      for (int ks = 0; ks < O / 8; ++ks) td_mma3(Dh, dGH, dGL, dW3H, dW3L, ks, ks, idescC, 1u);
      umma_commit(smem_u32(&bar));
    }
    // ---- while the tensor core runs: dPt partials from the dY1 rows and the saved first-conv taps
    if (tid < TD_NGR * 9) {
      // group g owns node g & 7 and the rows k = (g >> 3) + 4 m of both taps; 12 rows per group and
      // tile are summed in fp32, the per-tile sums in fp64
      const int blk9 = tid % 9, grp = tid / 9, tb = (blk9 / 3) * 4, ib = (blk9 % 3) * 4, nl = grp & 7, n = n0 + nl;
      float o16[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) o16[i] = 0.f;
      if (n < a.N) {
#pragma unroll
        for (int m = 0; m < 6; ++m)
          hq[1][m] = p_kb + 4 * m < b ? ld4(A2b + ((int64_t)(b + p_kb + 4 * m) * a.N + n) * 12 + ib) : f4zero();
#pragma unroll
        for (int tap = 0; tap < 2; ++tap) {
          const int sh = tap ? a.d1 : 0, kb = grp >> 3, sr = sh & 3, q0 = tb + (sh & ~3);
          const float4 (&h)[6] = hq[tap];
#pragma unroll
          for (int m = 0; m < 6; ++m) {
            if (kb + 4 * m < b) {
              const float* dy = DY + (kb + 4 * m) * TD_PITCH + nl * 12;
              // g[i] = dY1[tb + i + sh] (0 past the end): two aligned float4 and a uniform select
              const float4 ga = q0 < 12 ? ld4(dy + q0) : f4zero(), gb = (sr && q0 + 4 < 12) ? ld4(dy + q0 + 4) : f4zero();
              float g[4];
              if (sr == 0) { g[0] = ga.x; g[1] = ga.y; g[2] = ga.z; g[3] = ga.w; }
              else if (sr == 1) { g[0] = ga.y; g[1] = ga.z; g[2] = ga.w; g[3] = gb.x; }
              else if (sr == 2) { g[0] = ga.z; g[1] = ga.w; g[2] = gb.x; g[3] = gb.y; }
              else { g[0] = ga.w; g[1] = gb.x; g[2] = gb.y; g[3] = gb.z; }
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                o16[i * 4 + 0] = fmaf(g[i], h[m].x, o16[i * 4 + 0]); o16[i * 4 + 1] = fmaf(g[i], h[m].y, o16[i * 4 + 1]);
                o16[i * 4 + 2] = fmaf(g[i], h[m].z, o16[i * 4 + 2]); o16[i * 4 + 3] = fmaf(g[i], h[m].w, o16[i * 4 + 3]);
              }
            }
          }
        }
      }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) PRE[grp * 144 + (tb + i) * 12 + ib + j] = o16[i * 4 + j];
    }
    mbar_wait(smem_u32(&bar), phase); phase ^= 1;
    tc_fence_after_sync();
    __syncthreads();
    if (tid < 144) {
      double s = 0.0;
      for (int g2 = 0; g2 < TD_NGR; ++g2) s += (double)PRE[g2 * 144 + tid];
      pt_acc += s;
    }
    // the x rows of the LayerNorm backward (phase 5) are requested here: their L2 latency runs under phase 4
    float4 xq[2][3];
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const int r = tid + q * TD_THREADS, c = r / TD_NT, nl = r - c * TD_NT, n = n0 + nl;
      xq[q][0] = xq[q][1] = xq[q][2] = f4zero();
      if (r < C * TD_NT && n < a.N) {
        const float* src = xb + c * a.x_sc + (int64_t)n * 12;
        xq[q][0] = ld4(src); xq[q][1] = ld4(src + 4); xq[q][2] = ld4(src + 8);
      }
    }
    // ---- 4. D_h -> staging rows STG[C][.] (over the A tile, which the tensor core has finished reading)
    {
      int g = 0;
      for (int c0 = 0; c0 < C; c0 += 8, ++g) {
        if ((g & 3) != quarter) continue;
        float v[8];
        td_ld8_sum(Dh + lane_base + c0, v);
        if (l < TD_P) {
#pragma unroll
          for (int j = 0; j < 8; ++j) STG[(c0 + j) * TD_PITCH + l] = v[j];
        }
      }
    }
    __syncthreads();
    // ---- 5. LayerNorm backward on rows of 12 (msgat.py:122) + affine gradient partials
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const int r = tid + q * TD_THREADS;
      if (r < C * TD_NT) {
        const int c = r / TD_NT, nl = r - c * TD_NT, n = n0 + nl;
        if (n < a.N) {
          const float4 u = xq[q][0], v = xq[q][1], w = xq[q][2];
          float xh[12] = {u.x, u.y, u.z, u.w, v.x, v.y, v.z, v.w, w.x, w.y, w.z, w.w};
          uint32_t mb = 0;
#pragma unroll
          for (int t = 0; t < 12; ++t) mb |= (xh[t] > 0.f ? 1u : 0u) << t;
          MSK[c * TD_NT + nl] = a.x_is_relu ? mb : 0xfffu;
          float mu = 0.f;
#pragma unroll
          for (int t = 0; t < 12; ++t) mu += xh[t];
          mu *= (1.0f / 12.0f);
          float var = 0.f;
#pragma unroll
          for (int t = 0; t < 12; ++t) { xh[t] -= mu; var = fmaf(xh[t], xh[t], var); }
          const float rs = rsqrt_ln(var * (1.0f / 12.0f) + LN_EPS);
          float* qrow = STG + c * TD_PITCH + nl * 12;
          float gg[12], m1 = 0.f, m2 = 0.f;
          const float4 d0 = ld4(qrow), d1 = ld4(qrow + 4), d2 = ld4(qrow + 8);
          const float dhv[12] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w, d2.x, d2.y, d2.z, d2.w};
#pragma unroll
          for (int t = 0; t < 12; ++t) {
            xh[t] *= rs;
            const float dh = dhv[t];
            lnw_acc[t] = fmaf(dh, xh[t], lnw_acc[t]);
            lnb_acc[t] += dh;
            gg[t] = dh * LW[t];
            m1 += gg[t]; m2 = fmaf(gg[t], xh[t], m2);
          }
          m1 *= (1.0f / 12.0f); m2 *= (1.0f / 12.0f);
#pragma unroll
          for (int t = 0; t < 12; ++t) gg[t] = rs * (gg[t] - m1 - xh[t] * m2);
          st4(qrow, make_float4(gg[0], gg[1], gg[2], gg[3])); st4(qrow + 4, make_float4(gg[4], gg[5], gg[6], gg[7]));
          st4(qrow + 8, make_float4(gg[8], gg[9], gg[10], gg[11]));
        }
      }
    }
    __syncthreads();
    // ---- 6. dx = LN^T(dh) + D_x, one thread per (position, column-group quarter), coalesced stores
    {
      const int nl = l / 12, n = n0 + nl, t = l - nl * 12;
      const bool live = l < TD_P && n < a.N;
      const int64_t cs = (int64_t)a.N * 12;
      float* ob = dxb + (int64_t)bi * C * cs + (int64_t)n * 12 + t;
      int g = 0;
      for (int c0 = 0; c0 < C; c0 += 8, ++g) {
        if ((g & 3) != quarter) continue;
        float r[8];
        td_ld8_sum(Dx + lane_base + c0, r);
        if (live) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            float v = r[j] + STG[(c0 + j) * TD_PITCH + l];
            v = (MSK[(c0 + j) * TD_NT + nl] >> t) & 1u ? v : 0.f;   // x is a ReLU output: its gradient stops at x <= 0
            ob[(c0 + j) * cs] = v;
          }
        }
      }
    }
    tc_fence_before_sync();
    __syncthreads();
  }
  // ---- flush the per-block partials (layout of k_bwd_main: G row sums, then dY1 row sums)
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    float s = bsG[q];
    s += __shfl_xor_sync(0xffffffffu, s, 1); s += __shfl_xor_sync(0xffffffffu, s, 2); s += __shfl_xor_sync(0xffffffffu, s, 4);
    const int r = tid + q * TD_THREADS;
    if (r < O * TD_NT && (r & 7) == 0) a.bias_part[sco + (int64_t)blk * a.nbias + r / TD_NT] = s;
  }
  {
    float s = bsY;   // rows k on the even threads tid = 2 (k * 8 + nl)
    s += __shfl_xor_sync(0xffffffffu, s, 2); s += __shfl_xor_sync(0xffffffffu, s, 4); s += __shfl_xor_sync(0xffffffffu, s, 8);
    if (tid < 2 * b * TD_NT && (tid & 15) == 0) a.bias_part[sco + (int64_t)blk * a.nbias + O + (tid >> 4)] = s;
  }
  if (tid < 144) reinterpret_cast<double*>(reinterpret_cast<float*>(a.dPt_part) + sco)[(int64_t)blk * 144 + tid] = pt_acc;
#pragma unroll
  for (int t = 0; t < 12; ++t) {
    const float w = warp_sum(lnw_acc[t]), bb = warp_sum(lnb_acc[t]);
    if (lane == 0) { RED[warp * 24 + t] = w; RED[warp * 24 + 12 + t] = bb; }
  }
  __syncthreads();
  if (tid < 24) {
    float s = 0.f;
    for (int w = 0; w < TD_THREADS / 32; ++w) s += RED[w * 24 + tid];
    a.ln_part[sco + (int64_t)blk * 24 + tid] = s;
  }
  __syncthreads();   // RED, the per-sample operands and the accumulators are reused by the next logical block
  }
  if (warp == 0) tmem_dealloc(tm, 512);
}

bool tc_dgrad_supported(const Plan& p) {
  const char* off = getenv("MSGAT_TC_DISABLE");  // test switch: bit 2 = data-gradient kernel
  if (off && (atoi(off) & 4)) return false;
  // needs the first-conv taps (A2) kept by the tcgen05 forward kernels and the weight-gradient kernel for dA / GZ
  return (tc_apply3_supported(p) || tc_apply_supported(p)) && tc_wgrad_supported(p) && p.O == 3 * p.Oc;
}

int tc_dgrad_blocks_per_sample(const Plan& p) {
  const int ntile = (p.N + TD_NT - 1) / TD_NT;
  return tc_ctas_per_sample(p.B, ntile, 2, "MSGAT_TD_WAVES");
}

int launch_tc_dgrad(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, const float* x, const float* G, int64_t g_sm,
                    const float* saved, float* tcA, float* dx, int64_t dx_sm, float* bias_part, int nbias, float* ln_part,
                    double* dPt_part, void* stream) {
  TcDgradArgs a;
  a.x = x; a.x_sb = p.x_sb; a.x_sc = p.x_sc; a.x_sm = p.x_sm; a.g_sm = g_sm; a.sv_sm = p.saved_total; a.sc_sm = p.scratch_sm; a.dx_sm = dx_sm;
  a.G = G;
  a.C = p.C; a.N = p.N; a.O = p.O; a.b = p.Oc; a.ntile = (p.N + TD_NT - 1) / TD_NT; a.J = tc_dgrad_blocks_per_sample(p); a.B = p.B;
  a.d1 = p.dil[0]; a.d2 = p.dil[1]; a.nnz = p.nnz; a.x_is_relu = (p.flags & MSGAT_F_X_IS_RELU) ? 1 : 0; a.nbias = nbias;
  for (int m = 0; m < MSGAT_MAX_GROUP; ++m) {
    const msgat_meam_params* q = prm + (m < p.M ? m : 0);
    a.ln_w[m] = q->ln_w; a.ln_b[m] = q->ln_b; a.res_w[m] = q->res_w; a.c1_w[m] = q->conv_w[0]; a.c2_w[m] = q->conv_w[1]; a.sa_W[m] = q->sa_W;
  }
  a.Mc = saved + p.o_Mc; a.Pt = saved + p.o_Pt; a.Z = saved + p.o_Z; a.pw = saved + p.o_pw; a.A2 = saved + p.o_A2;
  a.row_ptr = adj->row_ptr; a.col_idx = adj->col_idx; a.csc_ptr = adj->csc_ptr; a.csc_row = adj->csc_row; a.csc_eid = adj->csc_eid;
  a.dx = dx; a.tcA = tcA; a.delta = nullptr; a.dSnz = nullptr; a.bias_part = bias_part; a.ln_part = ln_part; a.dPt_part = dPt_part;
  // one CTA per SM, every CTA on ONE component: sm_count / M CTAs per component walk its J * B logical blocks
  int G1 = msgat_sm_count() / p.M;
  if (G1 < 1) G1 = 1;
  if (G1 > a.J * p.B) G1 = a.J * p.B;
  MSGAT_LAUNCH(k_tc_dgrad, dim3(G1, p.M), dim3(TD_THREADS), tc_dgrad_smem(p.O, p.Oc), stream, a);
  return 0;
}
