// head.cu -- TPC head (msgat.py:158-160) fused with the time gate of MSGAT.forward
// (msgat.py:204): final LayerNorm_T, the Conv2d(T -> Tout, kernel [1, C]) that is a per-node
// linear map over (c, t), the multiplication by the TimeEmbedding gate and the sum over the
// M components (accumulate flag).  One read of the last MEAM's activation, HBM-bound.
#include "kernels.h"
#include "reduce.cuh"
#include "tile_ops.cuh"

#define HNT 16
#define HPS (HNT * 12 + 4)

struct HeadArgs {
  const float* x; int C, N, Tout, ntile, J, accumulate, x_is_relu, Bn;   // Bn: samples (the streaming kernels walk (sample, node tile) blocks)
  const float *ln_w, *ln_b, *fc_w, *fc_b, *gate; int64_t gate_sb;
  float *head_out, *out;
  // backward
  const float* dout; float *dx, *d_gate, *part; int pstride;
  int64_t dgate_sb;   // batch stride of d_gate ((n, o) contiguous)
};

// Grouped call: blockIdx.z = component m of M same-shaped heads (msgat.py:203-204).  The per-component buffers of
// HeadArgs (component 0) lie a fixed stride further for component m; parameters come from the tables.
struct HeadGrp {
  int M;
  int64_t x_sm, ho_sm, gate_sm, dx_sm, dgate_sm, part_sm;
  const float* ln_w[MSGAT_MAX_GROUP];
  const float* ln_b[MSGAT_MAX_GROUP];
  const float* fc_w[MSGAT_MAX_GROUP];
  const float* fc_b[MSGAT_MAX_GROUP];
};
MSGAT_DEV HeadArgs head_component(const HeadArgs& a0, const HeadGrp& g) {
  HeadArgs a = a0;
  if (g.M > 0) {
    const int m = blockIdx.z;
    a.x += m * g.x_sm; a.head_out += m * g.ho_sm;
    if (a.gate) a.gate += m * g.gate_sm;
    if (a.dx) a.dx += m * g.dx_sm;
    if (a.d_gate) a.d_gate += m * g.dgate_sm;
    if (a.part) a.part += m * g.part_sm;
    a.ln_w = g.ln_w[m]; a.ln_b = g.ln_b[m]; a.fc_w = g.fc_w[m]; a.fc_b = g.fc_b[m];
  }
  return a;
}

// weights staged as Wh[c][o][t] (t contiguous): fc.weight is [Tout, 12, 1, C]
MSGAT_DEV void stage_fc(float* Wh, const float* fc_w, int C, int Tout) {
  for (int i = threadIdx.x; i < C * Tout * 12; i += blockDim.x) {
    const int c = i / (Tout * 12), r = i - c * Tout * 12, o = r / 12, t = r - o * 12;
    Wh[i] = fc_w[((int64_t)o * 12 + t) * C + c];
  }
}

MSGAT_KERNEL MSGAT_LB(256, 1) k_head_fwd(HeadArgs a0, HeadGrp grp) {
  const HeadArgs a = head_component(a0, grp);
  MSGAT_DYN_SMEM(smem_raw);
  const int C = a.C, To = a.Tout, b = blockIdx.y, T = blockDim.x, tid = threadIdx.x;
  float* H = reinterpret_cast<float*>(smem_raw);
  float* Wh = H + C * HPS;
  float* LW = Wh + C * To * 12;
  stage_fc(Wh, a.fc_w, C, To);
  if (tid < 12) { LW[tid] = a.ln_w[tid]; LW[12 + tid] = a.ln_b[tid]; }
  const float* xb = a.x + (int64_t)b * C * a.N * 12;
  for (int tile = blockIdx.x; tile < a.ntile; tile += a.J) {
    const int n0 = tile * HNT;
    __syncthreads();
    tile_load_ln(xb, (int64_t)a.N * 12, C, a.N, n0, HNT, HPS, nullptr, H, LW, LW + 12, true, nullptr);
    __syncthreads();
    for (int it = tid; it < HNT * To; it += T) {
      const int nl = it / To, o = it - nl * To, n = n0 + nl;
      if (n >= a.N) continue;
      float s = __ldg(a.fc_b + o);
      for (int c = 0; c < C; ++c) {
        const float* h = H + c * HPS + nl * 12;
        const float* w = Wh + (c * To + o) * 12;
        s += dot4(ld4(h), ld4(w)) + dot4(ld4(h + 4), ld4(w + 4)) + dot4(ld4(h + 8), ld4(w + 8));
      }
      const int64_t g = ((int64_t)b * a.N + n) * To + o;
      a.head_out[g] = s;
      const float gt = a.gate ? a.gate[b * a.gate_sb + (int64_t)n * To + o] : 1.f;
      if (a.out) a.out[g] = a.accumulate ? a.out[g] + s * gt : s * gt;
    }
  }
}

#define HPAIRS 5  // (c, o) pairs of the fc weight gradient per thread

MSGAT_KERNEL MSGAT_LB(256, 1) k_head_bwd(HeadArgs a0, HeadGrp grp) {
  const HeadArgs a = head_component(a0, grp);
  MSGAT_DYN_SMEM(smem_raw);
  const int C = a.C, To = a.Tout, b = blockIdx.y, T = blockDim.x, tid = threadIdx.x;
  float* H = reinterpret_cast<float*>(smem_raw);
  float* Wh = H + C * HPS;
  float* LW = Wh + C * To * 12;
  float* ST = LW + 24;              // [C*HNT*2]
  float* DY = ST + C * HNT * 2;     // [HNT][To]
  float* RED = DY + HNT * To;       // [T/32][24]
  stage_fc(Wh, a.fc_w, C, To);
  if (tid < 12) { LW[tid] = a.ln_w[tid]; LW[12 + tid] = a.ln_b[tid]; }
  const float* xb = a.x + (int64_t)b * C * a.N * 12;
  float wacc[HPAIRS][12], bacc = 0.f, lnw_acc[12], lnb_acc[12];
  for (int k = 0; k < HPAIRS; ++k) for (int t = 0; t < 12; ++t) wacc[k][t] = 0.f;
  for (int t = 0; t < 12; ++t) { lnw_acc[t] = 0.f; lnb_acc[t] = 0.f; }
  for (int tile = blockIdx.x; tile < a.ntile; tile += a.J) {
    const int n0 = tile * HNT;
    __syncthreads();
    tile_load_ln(xb, (int64_t)a.N * 12, C, a.N, n0, HNT, HPS, nullptr, H, LW, LW + 12, true, ST);
    for (int it = tid; it < HNT * To; it += T) {
      const int nl = it / To, o = it - nl * To, n = n0 + nl;
      float dy = 0.f;
      if (n < a.N) {
        const int64_t g = ((int64_t)b * a.N + n) * To + o;
        const float d = a.dout[g];
        const float gt = a.gate ? a.gate[b * a.gate_sb + (int64_t)n * To + o] : 1.f;
        dy = d * gt;
        if (a.d_gate) a.d_gate[g + (int64_t)b * (a.dgate_sb - (int64_t)a.N * a.Tout)] = d * a.head_out[g];
      }
      DY[it] = dy;
    }
    __syncthreads();
    if (tid < To) {  // fc bias gradient
      float s = 0.f;
      for (int nl = 0; nl < HNT; ++nl) s += DY[nl * To + tid];
      bacc += s;
    }
    // fc weight gradient: d fc_w[o][t][c] += dy[n][o] * h[c][n][t]
    for (int k = 0; k < HPAIRS; ++k) {
      const int pr = tid + k * T;
      if (pr >= C * To) break;
      const int c = pr / To, o = pr - c * To;
      for (int nl = 0; nl < HNT; ++nl) {
        const float dy = DY[nl * To + o];
        const float* h = H + c * HPS + nl * 12;
        const float4 h0 = ld4(h), h1 = ld4(h + 4), h2 = ld4(h + 8);
        wacc[k][0] = fmaf(dy, h0.x, wacc[k][0]); wacc[k][1] = fmaf(dy, h0.y, wacc[k][1]);
        wacc[k][2] = fmaf(dy, h0.z, wacc[k][2]); wacc[k][3] = fmaf(dy, h0.w, wacc[k][3]);
        wacc[k][4] = fmaf(dy, h1.x, wacc[k][4]); wacc[k][5] = fmaf(dy, h1.y, wacc[k][5]);
        wacc[k][6] = fmaf(dy, h1.z, wacc[k][6]); wacc[k][7] = fmaf(dy, h1.w, wacc[k][7]);
        wacc[k][8] = fmaf(dy, h2.x, wacc[k][8]); wacc[k][9] = fmaf(dy, h2.y, wacc[k][9]);
        wacc[k][10] = fmaf(dy, h2.z, wacc[k][10]); wacc[k][11] = fmaf(dy, h2.w, wacc[k][11]);
      }
    }
    // dh[c][n][t] = sum_o fc_w[o][t][c] dy[n][o], then LayerNorm backward, one row per thread
    for (int r = tid; r < C * HNT; r += T) {
      const int c = r / HNT, nl = r - c * HNT, n = n0 + nl;
      if (n >= a.N) continue;
      float dh[12];
      for (int t = 0; t < 12; ++t) dh[t] = 0.f;
      for (int o = 0; o < To; ++o) {
        const float dy = DY[nl * To + o];
        const float* w = Wh + (c * To + o) * 12;
        const float4 w0 = ld4(w), w1 = ld4(w + 4), w2 = ld4(w + 8);
        dh[0] = fmaf(dy, w0.x, dh[0]); dh[1] = fmaf(dy, w0.y, dh[1]); dh[2] = fmaf(dy, w0.z, dh[2]); dh[3] = fmaf(dy, w0.w, dh[3]);
        dh[4] = fmaf(dy, w1.x, dh[4]); dh[5] = fmaf(dy, w1.y, dh[5]); dh[6] = fmaf(dy, w1.z, dh[6]); dh[7] = fmaf(dy, w1.w, dh[7]);
        dh[8] = fmaf(dy, w2.x, dh[8]); dh[9] = fmaf(dy, w2.y, dh[9]); dh[10] = fmaf(dy, w2.z, dh[10]); dh[11] = fmaf(dy, w2.w, dh[11]);
      }
      const float* src = xb + ((int64_t)c * a.N + n) * 12;
      const float mu = ST[2 * r], rs = ST[2 * r + 1];
      float xv[12], xh[12], m1 = 0.f, m2 = 0.f;
      for (int t = 0; t < 12; ++t) {
        xv[t] = src[t]; xh[t] = (xv[t] - mu) * rs;
        lnw_acc[t] += dh[t] * xh[t]; lnb_acc[t] += dh[t];
        const float gg = dh[t] * LW[t];
        m1 += gg; m2 += gg * xh[t];
      }
      m1 *= (1.0f / 12.0f); m2 *= (1.0f / 12.0f);
      float* d = a.dx + (((int64_t)b * C + c) * a.N + n) * 12;
      float o12[12];
      for (int t = 0; t < 12; ++t) {
        float v = rs * (dh[t] * LW[t] - m1 - xh[t] * m2);
        if (a.x_is_relu && !(xv[t] > 0.f)) v = 0.f;
        o12[t] = v;
      }
      st4(d, make_float4(o12[0], o12[1], o12[2], o12[3]));
      st4(d + 4, make_float4(o12[4], o12[5], o12[6], o12[7]));
      st4(d + 8, make_float4(o12[8], o12[9], o12[10], o12[11]));
    }
  }
  // partial layout per block: [To*12*C fc_w][To fc_b][24 ln]
  float* part = a.part + (int64_t)(b * a.J + blockIdx.x) * a.pstride;
  for (int k = 0; k < HPAIRS; ++k) {
    const int pr = tid + k * T;
    if (pr >= C * To) break;
    const int c = pr / To, o = pr - c * To;
    for (int t = 0; t < 12; ++t) part[((int64_t)o * 12 + t) * C + c] = wacc[k][t];
  }
  if (tid < To) part[To * 12 * C + tid] = bacc;
  __syncthreads();
  const int warp = tid >> 5, lane = tid & 31, nwarp = T >> 5;
  for (int t = 0; t < 12; ++t) {
    const float w = warp_sum(lnw_acc[t]), bb = warp_sum(lnb_acc[t]);
    if (lane == 0) { RED[warp * 24 + t] = w; RED[warp * 24 + 12 + t] = bb; }
  }
  __syncthreads();
  if (tid < 24) {
    float s = 0.f;
    for (int w = 0; w < nwarp; ++w) s += RED[w * 24 + tid];
    part[To * 12 * C + To + tid] = s;
  }
}

// ---------------------------------------------------------------------------------------
// Tout == 12 (every shipped model: main.py:34): the streaming pair of kernels.  A block owns 64
// consecutive nodes of one sample; warp w owns the channels w, w+8, ...; lane l owns the nodes
// l and l+32, whose 12-step rows it reads straight from HBM (a warp reads two runs of 1536
// contiguous bytes per channel) and keeps in registers -- no activation tile in shared memory,
// no block barrier inside the channel loop.  The fc weights of a channel are warp-uniform
// shared-memory reads, each reused for both nodes.
#define HS_NODES 64
#define HS_WARPS 8

#define HS_WP 148  // floats between the channels of the staged fc weight (16-byte aligned rows)
// fc.weight [12 o][12 t][1][C] (channel fastest) -> Wh[c][o][t].  A warp moves 8 channels x 4 consecutive (o,t)
// per step: four 32-byte sectors on the global side, 32 distinct banks on the shared side (pitch 148 = 20 mod 32);
// the loads of a batch of channel groups are issued before their stores, so a thread waits for one round trip
// per batch instead of one per element.
MSGAT_DEV void stage_fc12(float* Wh, const float* MSGAT_RESTRICT fc_w, int C) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, cl = lane >> 2, tl = lane & 3, nwarp = blockDim.x >> 5;
  for (int og = warp; og < 36; og += nwarp) {
    const int ot = og * 4 + tl;
    for (int c0 = 0; c0 < C; c0 += 72) {
      float v[9];
      MSGAT_UNROLL
      for (int k = 0; k < 9; ++k) {
        const int c = c0 + k * 8 + cl;
        v[k] = c < C ? __ldg(fc_w + (int64_t)ot * C + c) : 0.f;
      }
      MSGAT_UNROLL
      for (int k = 0; k < 9; ++k) {
        const int c = c0 + k * 8 + cl;
        if (c < C) Wh[c * HS_WP + ot] = v[k];
      }
    }
  }
}

struct Row12 { float4 a, b, c; };
MSGAT_DEV Row12 row_load(const float* p, bool live) {
  Row12 r;
  if (live) { r.a = ld4(p); r.b = ld4(p + 4); r.c = ld4(p + 8); }
  else { r.a = f4zero(); r.b = f4zero(); r.c = f4zero(); }
  return r;
}
// the thread's two rows (nodes lane and lane + 32 of the block) of one channel, global -> the warp's stage
MSGAT_DEV void rows_fetch(float* stage, const float* xc, int64_t nA, bool liveA, bool liveB, int lane) {
  const float* pa = liveA ? xc + nA * 12 : xc;          // dead rows are zero-filled; the address stays valid
  const float* pb = liveB ? xc + (nA + 32) * 12 : xc;
  float* sa = stage + lane * 12;
  float* sb = sa + 32 * 12;
  cp_async16(sa, pa, liveA); cp_async16(sa + 4, pa + 4, liveA); cp_async16(sa + 8, pa + 8, liveA);
  cp_async16(sb, pb, liveB); cp_async16(sb + 4, pb + 4, liveB); cp_async16(sb + 8, pb + 8, liveB);
}
MSGAT_DEV Row12 row_smem(const float* p) { Row12 r; r.a = ld4(p); r.b = ld4(p + 4); r.c = ld4(p + 8); return r; }
// (x - mean) * rstd of nn.LayerNorm([12]) (msgat.py:152): biased variance, eps 1e-5, same order as tile_load_ln
MSGAT_DEV void row_normalise(const Row12& x, float (&xh)[12], float& rs_out) {
  float4 a = x.a, b = x.b, d = x.c;
  const float mu = (a.x + a.y + a.z + a.w + b.x + b.y + b.z + b.w + d.x + d.y + d.z + d.w) * (1.0f / 12.0f);
  a.x -= mu; a.y -= mu; a.z -= mu; a.w -= mu; b.x -= mu; b.y -= mu; b.z -= mu; b.w -= mu;
  d.x -= mu; d.y -= mu; d.z -= mu; d.w -= mu;
  const float var = (dot4(a, a) + dot4(b, b) + dot4(d, d)) * (1.0f / 12.0f);
  const float rs = rsqrt_ln(var + LN_EPS);
  xh[0] = a.x * rs; xh[1] = a.y * rs; xh[2] = a.z * rs; xh[3] = a.w * rs;
  xh[4] = b.x * rs; xh[5] = b.y * rs; xh[6] = b.z * rs; xh[7] = b.w * rs;
  xh[8] = d.x * rs; xh[9] = d.y * rs; xh[10] = d.z * rs; xh[11] = d.w * rs;
  rs_out = rs;
}

MSGAT_KERNEL MSGAT_LB(256, 2) k_head_fwd12(HeadArgs a0, HeadGrp grp) {
  const HeadArgs a = head_component(a0, grp);
  MSGAT_DYN_SMEM(smem_raw);
  const int C = a.C, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  float* Wh = reinterpret_cast<float*>(smem_raw);   // [C][12 o][12 t]
  float* RED = Wh + C * HS_WP;                       // [HS_WARPS][HS_NODES][12] (after the channel loop)
  float* LW = RED + 2 * HS_WARPS * HS_NODES * 12;   // [24]
  float* XS = RED;                                  // [HS_WARPS][2 stages][HS_NODES][12] rows in flight (RED reuses it after the loop)
  float* ring = XS + warp * 2 * HS_NODES * 12;
  // The 41 KB of fc weights are staged ONCE per block; the block then walks the logical blocks (sample b, node tile)
  // lb = blockIdx.x, + gridDim.x, ... of its component (staging them per 64-node tile was 11 % of the kernel's samples).
  stage_fc12(Wh, a.fc_w, C);
  if (tid < 12) { LW[tid] = a.ln_w[tid]; LW[12 + tid] = a.ln_b[tid]; }
  __syncthreads();
  float lw[12], lb[12];
  MSGAT_UNROLL
  for (int t = 0; t < 12; ++t) { lw[t] = LW[t]; lb[t] = LW[12 + t]; }
  const int ntn = (a.N + HS_NODES - 1) / HS_NODES;
  for (int lbk = blockIdx.x; lbk < ntn * a.Bn; lbk += gridDim.x) {
  const int b = lbk / ntn;
  const int n0 = (lbk - b * ntn) * HS_NODES, nA = n0 + lane, nB = nA + 32;
  const bool liveA = nA < a.N, liveB = nB < a.N;
  const float* xb = a.x + (int64_t)b * C * a.N * 12;
  // two channels ahead of the one being consumed
  if (warp < C) rows_fetch(ring, xb + (int64_t)warp * a.N * 12, nA, liveA, liveB, lane);
  cp_async_commit();
  if (warp + HS_WARPS < C) rows_fetch(ring + HS_NODES * 12, xb + (int64_t)(warp + HS_WARPS) * a.N * 12, nA, liveA, liveB, lane);
  cp_async_commit();
  float accA[12], accB[12];
  MSGAT_UNROLL
  for (int o = 0; o < 12; ++o) { accA[o] = 0.f; accB[o] = 0.f; }
  int st = 0;
  for (int c = warp; c < C; c += HS_WARPS) {
    cp_async_wait<1>();   // this channel's rows have landed (each thread reads back only what it fetched itself)
    float* stage = ring + st * HS_NODES * 12;
    const Row12 xa = row_smem(stage + lane * 12), xc = row_smem(stage + (lane + 32) * 12);
    float hA[12], hB[12], rs;
    row_normalise(xa, hA, rs);
    row_normalise(xc, hB, rs);
    const int c2 = c + 2 * HS_WARPS;
    if (c2 < C) rows_fetch(stage, xb + (int64_t)c2 * a.N * 12, nA, liveA, liveB, lane);
    cp_async_commit();
    st ^= 1;
    MSGAT_UNROLL
    for (int t = 0; t < 12; ++t) { hA[t] = hA[t] * lw[t] + lb[t]; hB[t] = hB[t] * lw[t] + lb[t]; }
    const float* w = Wh + c * HS_WP;
    MSGAT_UNROLL
    for (int o = 0; o < 12; ++o) {
      const float4 w0 = ld4(w + o * 12), w1 = ld4(w + o * 12 + 4), w2 = ld4(w + o * 12 + 8);
      float sA = accA[o], sB = accB[o];
      sA = fmaf(hA[0], w0.x, sA); sA = fmaf(hA[1], w0.y, sA); sA = fmaf(hA[2], w0.z, sA); sA = fmaf(hA[3], w0.w, sA);
      sA = fmaf(hA[4], w1.x, sA); sA = fmaf(hA[5], w1.y, sA); sA = fmaf(hA[6], w1.z, sA); sA = fmaf(hA[7], w1.w, sA);
      sA = fmaf(hA[8], w2.x, sA); sA = fmaf(hA[9], w2.y, sA); sA = fmaf(hA[10], w2.z, sA); sA = fmaf(hA[11], w2.w, sA);
      sB = fmaf(hB[0], w0.x, sB); sB = fmaf(hB[1], w0.y, sB); sB = fmaf(hB[2], w0.z, sB); sB = fmaf(hB[3], w0.w, sB);
      sB = fmaf(hB[4], w1.x, sB); sB = fmaf(hB[5], w1.y, sB); sB = fmaf(hB[6], w1.z, sB); sB = fmaf(hB[7], w1.w, sB);
      sB = fmaf(hB[8], w2.x, sB); sB = fmaf(hB[9], w2.y, sB); sB = fmaf(hB[10], w2.z, sB); sB = fmaf(hB[11], w2.w, sB);
      accA[o] = sA; accB[o] = sB;
    }
  }
  cp_async_wait<0>();
  __syncthreads();   // every warp is done with its ring before RED overwrites it
  // fixed-order sum of the eight channel groups
  float* ra = RED + (warp * HS_NODES + lane) * 12;
  float* rb = ra + 32 * 12;
  st4(ra, make_float4(accA[0], accA[1], accA[2], accA[3])); st4(ra + 4, make_float4(accA[4], accA[5], accA[6], accA[7]));
  st4(ra + 8, make_float4(accA[8], accA[9], accA[10], accA[11]));
  st4(rb, make_float4(accB[0], accB[1], accB[2], accB[3])); st4(rb + 4, make_float4(accB[4], accB[5], accB[6], accB[7]));
  st4(rb + 8, make_float4(accB[8], accB[9], accB[10], accB[11]));
  __syncthreads();
  for (int it = tid; it < HS_NODES * 12; it += 256) {
    const int nl = it / 12, o = it - nl * 12, n = n0 + nl;
    if (n >= a.N) continue;
    float s = __ldg(a.fc_b + o);
    MSGAT_UNROLL
    for (int w = 0; w < HS_WARPS; ++w) s += RED[w * HS_NODES * 12 + it];
    const int64_t g = ((int64_t)b * a.N + n) * 12 + o;
    a.head_out[g] = s;
    const float gt = a.gate ? a.gate[b * a.gate_sb + (int64_t)n * 12 + o] : 1.f;
    if (a.out) a.out[g] = a.accumulate ? a.out[g] + s * gt : s * gt;
  }
  __syncthreads();   // RED is the next logical block's ring
  }
}

// LayerNorm([12]) backward of one row: dh -> dx (masked by x > 0 when x is a ReLU output), affine gradient
// partials.  The normalised row comes back from the warp buffer the thread wrote it to.
MSGAT_DEV void head_ln_bwd_row(const float (&dh)[12], const Row12& xr, const float* hrow, float rs, const float* LW,
                               float (&lnw_acc)[12], int x_is_relu, float* dst) {
  const float4 h0 = ld4(hrow), h1 = ld4(hrow + 4), h2 = ld4(hrow + 8);
  const float xh[12] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w, h2.x, h2.y, h2.z, h2.w};
  const float xv[12] = {xr.a.x, xr.a.y, xr.a.z, xr.a.w, xr.b.x, xr.b.y, xr.b.z, xr.b.w, xr.c.x, xr.c.y, xr.c.z, xr.c.w};
  float m1 = 0.f, m2 = 0.f;
  MSGAT_UNROLL
  for (int t = 0; t < 12; ++t) {
    lnw_acc[t] += dh[t] * xh[t];
    const float gg = dh[t] * LW[t];
    m1 += gg; m2 += gg * xh[t];
  }
  m1 *= (1.0f / 12.0f); m2 *= (1.0f / 12.0f);
  float o12[12];
  MSGAT_UNROLL
  for (int t = 0; t < 12; ++t) {
    float v = rs * (dh[t] * LW[t] - m1 - xh[t] * m2);
    if (x_is_relu && !(xv[t] > 0.f)) v = 0.f;
    o12[t] = v;
  }
  if (dst) {
    st4(dst, make_float4(o12[0], o12[1], o12[2], o12[3]));
    st4(dst + 4, make_float4(o12[4], o12[5], o12[6], o12[7]));
    st4(dst + 8, make_float4(o12[8], o12[9], o12[10], o12[11]));
  }
}

// Backward of the same block shape.  dh = W^T dy per row and the LayerNorm backward stay in the
// registers of the row's thread; the fc weight gradient of a channel, dW[c][o][t] =
// ln_w[t] * sum_n dy[n][o] xh[c][n][t] + ln_b[t] * sum_n dy[n][o], is contracted over the block's 64
// nodes by the warp that owns the channel: the normalised rows go through a per-warp buffer,
// 27 lanes own (node third, o quad, t quad) and combine the thirds with two shuffles.
MSGAT_KERNEL MSGAT_LB(256, 2) k_head_bwd12(HeadArgs a0, HeadGrp grp) {
  const HeadArgs a = head_component(a0, grp);
  MSGAT_DYN_SMEM(smem_raw);
  const int C = a.C, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  float* Wh = reinterpret_cast<float*>(smem_raw);   // [C][12 o][12 t]
  float* HB = Wh + C * HS_WP;                        // [HS_WARPS][HS_NODES][12]  normalised rows of the warp's channel
  float* DY = HB + HS_WARPS * HS_NODES * 12;        // [HS_NODES][12]   dy = dout * gate
  float* DYT = DY + HS_NODES * 12;                  // [12][HS_NODES]
  float* LW = DYT + HS_NODES * 12;                  // [24]
  float* BS = LW + 24;                              // [12] column sums of DY (fc bias gradient of the block)
  float* RED = BS + 12;                             // [HS_WARPS][24]
  float* XS = RED + HS_WARPS * 24;                  // [HS_WARPS][HS_NODES][12]  next channel's rows in flight
  float* WS = XS + HS_WARPS * HS_NODES * 12;        // [12 o][12 t] fc weights summed over the channels
  float* stage = XS + warp * HS_NODES * 12;
  // fc weights staged once per block; the block walks the logical blocks (sample b, node tile) of its component, each with its
  // own slice of partial sums (same layout as when every logical block was a CTA)
  stage_fc12(Wh, a.fc_w, C);
  if (tid < 12) { LW[tid] = a.ln_w[tid]; LW[12 + tid] = a.ln_b[tid]; }
  __syncthreads();
  // The LayerNorm bias gradient of a logical block, sum over its rows of dh[t] = sum_o fc_w[o][t][c] dy[n][o], is
  // sum_o (sum_n dy[n][o]) (sum_c fc_w[o][t][c]) = sum_o BS[o] WS[o][t]: no per-row accumulation (12 registers of every thread
  // across the channel loop, 24 additions per channel)
  if (tid < 144) {
    float sw = 0.f;
    for (int c = 0; c < C; ++c) sw += Wh[c * HS_WP + tid];
    WS[tid] = sw;
  }
  const int ntn = (a.N + HS_NODES - 1) / HS_NODES;
  for (int lbk = blockIdx.x; lbk < ntn * a.Bn; lbk += gridDim.x) {
  const int b = lbk / ntn;
  const int n0 = (lbk - b * ntn) * HS_NODES, nA = n0 + lane, nB = nA + 32;
  const bool liveA = nA < a.N, liveB = nB < a.N;
  const float* xb = a.x + (int64_t)b * C * a.N * 12;
  // the first channel's rows are in flight while the block stages dy
  if (warp < C) rows_fetch(stage, xb + (int64_t)warp * a.N * 12, nA, liveA, liveB, lane);
  cp_async_commit();
  {
    // the tile's 64 x 12 gradients are contiguous in dout / gate / head_out: three items per thread, all nine loads requested
    // before the first use (one exposed round trip per logical block instead of three)
    float dv[3], gv3[3], hv[3];
    MSGAT_UNROLL
    for (int u = 0; u < 3; ++u) {
      const int it = tid + u * 256;
      const bool ok = n0 + it / 12 < a.N;
      const int64_t g = ((int64_t)b * a.N + n0) * 12 + it;
      dv[u] = ok ? a.dout[g] : 0.f;
      gv3[u] = (ok && a.gate) ? a.gate[b * a.gate_sb + (int64_t)n0 * 12 + it] : 1.f;
      hv[u] = (ok && a.d_gate) ? a.head_out[g] : 0.f;
    }
    MSGAT_UNROLL
    for (int u = 0; u < 3; ++u) {
      const int it = tid + u * 256, nl = it / 12, o = it - nl * 12;
      const float dy = dv[u] * gv3[u];
      if (a.d_gate && n0 + nl < a.N)
        a.d_gate[((int64_t)b * a.N + n0) * 12 + it + (int64_t)b * (a.dgate_sb - (int64_t)a.N * a.Tout)] = dv[u] * hv[u];
      DY[it] = dy;
      DYT[o * HS_NODES + nl] = dy;
    }
  }
  __syncthreads();
  if (tid < 12) {
    float s = 0.f;
    for (int nl = 0; nl < HS_NODES; ++nl) s += DY[nl * 12 + tid];
    BS[tid] = s;
  }
  __syncthreads();
  float* part = a.part + (int64_t)lbk * a.pstride;
  float* hb = HB + warp * HS_NODES * 12;
  float lnw_acc[12];
  MSGAT_UNROLL
  for (int t = 0; t < 12; ++t) lnw_acc[t] = 0.f;
  const int third = lane / 9, item = lane - third * 9, oq = item / 3, tq = item - oq * 3;
  for (int c = warp; c < C; c += HS_WARPS) {
    cp_async_wait<0>();   // each thread reads back only the rows it fetched itself
    const Row12 xa = row_smem(stage + lane * 12), xc = row_smem(stage + (lane + 32) * 12);
    float hA[12], hB[12], rsA, rsB;
    row_normalise(xa, hA, rsA);
    row_normalise(xc, hB, rsB);
    if (c + HS_WARPS < C) rows_fetch(stage, xb + (int64_t)(c + HS_WARPS) * a.N * 12, nA, liveA, liveB, lane);
    cp_async_commit();
    {
      float* ra = hb + lane * 12;
      float* rb = ra + 32 * 12;
      st4(ra, make_float4(hA[0], hA[1], hA[2], hA[3])); st4(ra + 4, make_float4(hA[4], hA[5], hA[6], hA[7]));
      st4(ra + 8, make_float4(hA[8], hA[9], hA[10], hA[11]));
      st4(rb, make_float4(hB[0], hB[1], hB[2], hB[3])); st4(rb + 4, make_float4(hB[4], hB[5], hB[6], hB[7]));
      st4(rb + 8, make_float4(hB[8], hB[9], hB[10], hB[11]));
    }
    // dh[t] = sum_o fc_w[o][t][c] dy[n][o]
    float dA[12], dB[12];
    MSGAT_UNROLL
    for (int t = 0; t < 12; ++t) { dA[t] = 0.f; dB[t] = 0.f; }
    const float* w = Wh + c * HS_WP;
    MSGAT_UNROLL
    for (int o = 0; o < 12; ++o) {
      const float4 w0 = ld4(w + o * 12), w1 = ld4(w + o * 12 + 4), w2 = ld4(w + o * 12 + 8);
      const float ya = DYT[o * HS_NODES + lane], yb = DYT[o * HS_NODES + 32 + lane];
      dA[0] = fmaf(ya, w0.x, dA[0]); dA[1] = fmaf(ya, w0.y, dA[1]); dA[2] = fmaf(ya, w0.z, dA[2]); dA[3] = fmaf(ya, w0.w, dA[3]);
      dA[4] = fmaf(ya, w1.x, dA[4]); dA[5] = fmaf(ya, w1.y, dA[5]); dA[6] = fmaf(ya, w1.z, dA[6]); dA[7] = fmaf(ya, w1.w, dA[7]);
      dA[8] = fmaf(ya, w2.x, dA[8]); dA[9] = fmaf(ya, w2.y, dA[9]); dA[10] = fmaf(ya, w2.z, dA[10]); dA[11] = fmaf(ya, w2.w, dA[11]);
      dB[0] = fmaf(yb, w0.x, dB[0]); dB[1] = fmaf(yb, w0.y, dB[1]); dB[2] = fmaf(yb, w0.z, dB[2]); dB[3] = fmaf(yb, w0.w, dB[3]);
      dB[4] = fmaf(yb, w1.x, dB[4]); dB[5] = fmaf(yb, w1.y, dB[5]); dB[6] = fmaf(yb, w1.z, dB[6]); dB[7] = fmaf(yb, w1.w, dB[7]);
      dB[8] = fmaf(yb, w2.x, dB[8]); dB[9] = fmaf(yb, w2.y, dB[9]); dB[10] = fmaf(yb, w2.z, dB[10]); dB[11] = fmaf(yb, w2.w, dB[11]);
    }
    // LayerNorm backward, masked by the ReLU that produced x, one row per (thread, node)
    head_ln_bwd_row(dA, xa, hb + lane * 12, rsA, LW, lnw_acc, a.x_is_relu,
                    liveA ? a.dx + (((int64_t)b * C + c) * a.N + nA) * 12 : nullptr);
    head_ln_bwd_row(dB, xc, hb + (lane + 32) * 12, rsB, LW, lnw_acc, a.x_is_relu,
                    liveB ? a.dx + (((int64_t)b * C + c) * a.N + nB) * 12 : nullptr);
    __syncwarp();
    // fc weight gradient of channel c over the block's nodes
    float4 g0 = f4zero(), g1 = f4zero(), g2 = f4zero(), g3 = f4zero();
    if (lane < 27) {
      // a uniform trip count (node 64, which only the first third would reach, does not exist: 22, 21, 21 nodes) lets the compiler
      // keep the loads of the next step in flight under the 16 FMAs of this one
      const float* dyp = DY + third * 12 + oq * 4;
      const float* hp = hb + third * 12 + tq * 4;
      // software-pipelined by hand: the operands of step k + 1 are requested before the 16 FMAs of step k (with the loads at the
      // top of each unrolled body 60 % of this loop's samples waited for shared memory)
      float4 d4 = ld4(dyp), h4 = ld4(hp);
      MSGAT_UNROLL_N(4)
      for (int k = 1; k < 21; ++k) {
        const float4 dn = ld4(dyp + k * 36), hn = ld4(hp + k * 36);
        fma4(g0, d4.x, h4); fma4(g1, d4.y, h4); fma4(g2, d4.z, h4); fma4(g3, d4.w, h4);
        d4 = dn; h4 = hn;
      }
      fma4(g0, d4.x, h4); fma4(g1, d4.y, h4); fma4(g2, d4.z, h4); fma4(g3, d4.w, h4);
      if (third == 0) {
        const float4 d4 = ld4(DY + 63 * 12 + oq * 4), h4 = ld4(hb + 63 * 12 + tq * 4);
        fma4(g0, d4.x, h4); fma4(g1, d4.y, h4); fma4(g2, d4.z, h4); fma4(g3, d4.w, h4);
      }
    }
    float gv[16] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w, g2.x, g2.y, g2.z, g2.w, g3.x, g3.y, g3.z, g3.w};
    MSGAT_UNROLL
    for (int i = 0; i < 16; ++i) {
      const float v1 = __shfl_sync(0xffffffffu, gv[i], (lane + 9) & 31), v2 = __shfl_sync(0xffffffffu, gv[i], (lane + 18) & 31);
      gv[i] = (gv[i] + v1) + v2;
    }
    if (lane < 9) {
      // the partials of a logical block are CHANNEL-major, [c][o][t]: a lane's 4 x 4 block is four 16-byte stores into the 576
      // contiguous bytes of its channel (as [o][t][c], the gradient's own layout, it was 16 four-byte stores C floats apart: 12 %
      // of the kernel's samples); k_reduce transposes while it sums (RJob::tr)
      float* pc = part + c * 144 + oq * 48 + tq * 4;
      const float4 lw = ld4(LW + tq * 4), lb = ld4(LW + 12 + tq * 4);
      MSGAT_UNROLL
      for (int i = 0; i < 4; ++i) {
        const float bs = BS[oq * 4 + i];
        st4(pc + i * 12, make_float4(fmaf(lw.x, gv[i * 4], lb.x * bs), fmaf(lw.y, gv[i * 4 + 1], lb.y * bs),
                                     fmaf(lw.z, gv[i * 4 + 2], lb.z * bs), fmaf(lw.w, gv[i * 4 + 3], lb.w * bs)));
      }
    }
    __syncwarp();
  }
  if (tid < 12) part[144 * C + tid] = BS[tid];
  MSGAT_UNROLL
  for (int t = 0; t < 12; ++t) {
    const float w = warp_sum(lnw_acc[t]);
    if (lane == 0) RED[warp * 24 + t] = w;
  }
  __syncthreads();
  if (tid < 12) {
    float s = 0.f;
    for (int w = 0; w < HS_WARPS; ++w) s += RED[w * 24 + tid];
    part[144 * C + 12 + tid] = s;
  } else if (tid < 24) {
    float s = 0.f;
    for (int o = 0; o < 12; ++o) s = fmaf(BS[o], WS[o * 12 + tid - 12], s);
    part[144 * C + 12 + tid] = s;
  }
  __syncthreads();   // DY, BS, RED and the row buffers are the next logical block's
  }
}

// blocks of a streaming launch per component: two per SM over all components, at most one per (sample, node tile)
static int head_stream_grid(int N, int B, int M) {
  const int lb = ((N + HS_NODES - 1) / HS_NODES) * B;
  int g = (2 * msgat_sm_count()) / (M < 1 ? 1 : M);
  if (g < 1) g = 1;
  return g > lb ? lb : g;
}
static bool head_streaming(const msgat_head_desc* d) {
  return d->Tout == 12 && ((size_t)d->C * HS_WP + HS_WARPS * HS_NODES * 12 + 2 * HS_NODES * 12 + 24 + 12 + HS_WARPS * 24 + HS_WARPS * HS_NODES * 12 + 144) * 4 <= 110 * 1024;
}

static int head_J(const msgat_head_desc* d, int ntile) {
  int J = (2 * msgat_sm_count()) / d->B;
  return J > ntile ? ntile : (J < 1 ? 1 : J);
}
static int head_check(const msgat_head_desc* d) {
  if (d->B <= 0 || d->C <= 0 || d->N <= 0 || d->Tout <= 0 || d->Tout > 64 || d->C * d->Tout > 256 * HPAIRS) {
    msgat_set_error("head: unsupported shape C=%d Tout=%d", d->C, d->Tout);
    return -40;
  }
  size_t smem = ((size_t)d->C * HPS + (size_t)d->C * d->Tout * 12 + 24 + (size_t)d->C * HNT * 2 + HNT * d->Tout + 8 * 24) * 4;
  if (smem > 220 * 1024) { msgat_set_error("head: C=%d Tout=%d needs %zu B of shared memory", d->C, d->Tout, smem); return -41; }
  return 0;
}

// out[b,n,o] = sum_m head_out[m][b][n][o] * gate[b][m][n][o]   (msgat.py:204), components in index order
MSGAT_KERNEL k_head_combine(const float* ho, const float* gate, float* out, int M, int B, int NT, int64_t gate_sb, int64_t gate_sm) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)B * NT) return;
  const int b = (int)(i / NT), r = (int)(i - (int64_t)b * NT);
  float s = 0.f;
  for (int m = 0; m < M; ++m) {
    const float y = ho[((int64_t)m * B + b) * NT + r];
    s += gate ? y * gate[b * gate_sb + m * gate_sm + r] : y;
  }
  out[i] = s;
}

extern "C" {

int msgat_head_forward(const msgat_head_desc* d, const float* x, const float* ln_w, const float* ln_b, const float* fc_w,
                       const float* fc_b, const float* gate, float* head_out, float* out, void* stream) {
  if (int rc = head_check(d)) return rc;
  msgat_prof_tag("@C%dO%d", d->C, d->Tout);
  HeadArgs a = {};
  a.x = x; a.C = d->C; a.N = d->N; a.Tout = d->Tout; a.ntile = (d->N + HNT - 1) / HNT; a.J = head_J(d, a.ntile);
  a.accumulate = d->accumulate; a.ln_w = ln_w; a.ln_b = ln_b; a.fc_w = fc_w; a.fc_b = fc_b; a.gate = gate;
  a.gate_sb = d->gate_sb; a.head_out = head_out; a.out = out;
  if (head_streaming(d)) {
    size_t smem = ((size_t)d->C * HS_WP + 2 * HS_WARPS * HS_NODES * 12 + 24) * 4;
    a.Bn = d->B;
    MSGAT_LAUNCH(k_head_fwd12, dim3(head_stream_grid(d->N, d->B, 1)), dim3(256), smem, stream, a, HeadGrp{});
  } else {
    size_t smem = ((size_t)d->C * HPS + (size_t)d->C * d->Tout * 12 + 24) * 4;
    MSGAT_LAUNCH(k_head_fwd, dim3(a.J, d->B), dim3(256), smem, stream, a, HeadGrp{});
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { msgat_set_error("head_forward: %s", cudaGetErrorString(e)); return -100; }
  return 0;
}

int64_t msgat_head_scratch_floats(const msgat_head_desc* d) {
  const int ntile = (d->N + HNT - 1) / HNT;
  const int64_t blocks = head_streaming(d) ? (d->N + HS_NODES - 1) / HS_NODES : head_J(d, ntile);
  return blocks * d->B * (d->Tout * 12 * d->C + d->Tout + 24);
}

int msgat_head_backward(const msgat_head_desc* d, const float* x, const float* ln_w, const float* ln_b, const float* fc_w,
                        const float* gate, const float* head_out, const float* dout, float* dx, float* d_ln_w,
                        float* d_ln_b, float* d_fc_w, float* d_fc_b, float* d_gate, float* scratch, void* stream) {
  if (int rc = head_check(d)) return rc;
  msgat_prof_tag("@C%dO%d", d->C, d->Tout);
  HeadArgs a = {};
  a.x = x; a.C = d->C; a.N = d->N; a.Tout = d->Tout; a.ntile = (d->N + HNT - 1) / HNT; a.J = head_J(d, a.ntile);
  a.x_is_relu = d->x_is_relu; a.ln_w = ln_w; a.ln_b = ln_b; a.fc_w = fc_w; a.gate = gate; a.gate_sb = d->gate_sb;
  a.head_out = const_cast<float*>(head_out); a.dout = dout; a.dx = dx; a.d_gate = d_gate; a.part = scratch;
  a.dgate_sb = (int64_t)d->N * d->Tout;
  const int nW = d->Tout * 12 * d->C;
  a.pstride = nW + d->Tout + 24;
  int nblk = a.J;
  if (head_streaming(d)) {
    nblk = (d->N + HS_NODES - 1) / HS_NODES;
    size_t smem = ((size_t)d->C * HS_WP + HS_WARPS * HS_NODES * 12 + 2 * HS_NODES * 12 + 24 + 12 + HS_WARPS * 24 + HS_WARPS * HS_NODES * 12 + 144) * 4;
    a.Bn = d->B;
    MSGAT_LAUNCH(k_head_bwd12, dim3(head_stream_grid(d->N, d->B, 1)), dim3(256), smem, stream, a, HeadGrp{});
  } else {
    size_t smem = ((size_t)d->C * HPS + (size_t)d->C * d->Tout * 12 + 24 + (size_t)d->C * HNT * 2 + HNT * d->Tout + 8 * 24) * 4;
    MSGAT_LAUNCH(k_head_bwd, dim3(a.J, d->B), dim3(256), smem, stream, a, HeadGrp{});
  }
  ReduceArgs r;
  int nj = 0, tot = 0;
  const int nparts = nblk * d->B;
  auto add = [&](float* dst, const float* src, int n, int tr = 0) {
    if (!dst) return;
    RJob jb = {};
    jb.dst[0] = dst; jb.src = src; jb.n = n; jb.nparts = nparts; jb.dmul = 1; jb.doff = 0; jb.pstride = (int64_t)a.pstride; jb.f64 = 0;
    jb.tr = tr;
    r.j[nj] = jb;
    r.starts[nj] = tot; tot += n; ++nj;
  };
  add(d_fc_w, scratch, nW, head_streaming(d) ? 144 : 0);   // k_head_bwd12 writes channel-major partials
  add(d_fc_b, scratch + nW, d->Tout);
  add(d_ln_w, scratch + nW + d->Tout, 12);
  add(d_ln_b, scratch + nW + d->Tout + 12, 12);
  r.starts[nj] = tot; r.njobs = nj; r.sc_sm = 0;
  if (tot > 0) LAUNCH_REDUCE(tot, 1, stream, r);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { msgat_set_error("head_backward: %s", cudaGetErrorString(e)); return -100; }
  return 0;
}


// ---- grouped heads: all M components of MSGAT.forward (msgat.py:203-204) in one launch per kernel ----------------
int msgat_heads_forward(const msgat_heads_desc* d, const float* x, const float* const* ln_w, const float* const* ln_b,
                        const float* const* fc_w, const float* const* fc_b, const float* gate, float* head_out, float* out,
                        void* stream) {
  msgat_head_desc d1 = {d->B, d->C, d->N, d->Tout, 0, d->x_is_relu, d->gate_sb};
  if (int rc = head_check(&d1)) return rc;
  if (d->M < 1 || d->M > MSGAT_MAX_GROUP) { msgat_set_error("heads: 1..%d components", MSGAT_MAX_GROUP); return -42; }
  msgat_prof_tag("@C%dO%d", d->C, d->Tout);
  HeadArgs a = {};
  a.x = x; a.C = d->C; a.N = d->N; a.Tout = d->Tout; a.ntile = (d->N + HNT - 1) / HNT; a.J = head_J(&d1, a.ntile);
  a.gate = nullptr; a.head_out = head_out; a.out = nullptr;   // the gated sum over the components is k_head_combine's
  HeadGrp g = {};
  g.M = d->M; g.x_sm = d->x_sm; g.ho_sm = (int64_t)d->B * d->N * d->Tout;
  for (int m = 0; m < d->M; ++m) { g.ln_w[m] = ln_w[m]; g.ln_b[m] = ln_b[m]; g.fc_w[m] = fc_w[m]; g.fc_b[m] = fc_b[m]; }
  if (head_streaming(&d1)) {
    size_t smem = ((size_t)d->C * HS_WP + 2 * HS_WARPS * HS_NODES * 12 + 24) * 4;
    a.Bn = d->B;
    MSGAT_LAUNCH(k_head_fwd12, dim3(head_stream_grid(d->N, d->B, d->M), 1, d->M), dim3(256), smem, stream, a, g);
  } else {
    size_t smem = ((size_t)d->C * HPS + (size_t)d->C * d->Tout * 12 + 24) * 4;
    MSGAT_LAUNCH(k_head_fwd, dim3(a.J, d->B, d->M), dim3(256), smem, stream, a, g);
  }
  const int64_t n = (int64_t)d->B * d->N * d->Tout;
  int blocks = (int)((n + 255) / 256);
  MSGAT_LAUNCH(k_head_combine, dim3(blocks), dim3(256), 0, stream, (const float*)head_out, gate, out, d->M, d->B, d->N * d->Tout,
               d->gate_sb, d->gate_sm);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { msgat_set_error("heads_forward: %s", cudaGetErrorString(e)); return -100; }
  return 0;
}

int64_t msgat_heads_scratch_floats(const msgat_heads_desc* d) {
  msgat_head_desc d1 = {d->B, d->C, d->N, d->Tout, 0, d->x_is_relu, d->gate_sb};
  return msgat_head_scratch_floats(&d1);   // per component
}

int msgat_heads_backward(const msgat_heads_desc* d, const float* x, const float* const* ln_w, const float* const* ln_b,
                         const float* const* fc_w, const float* gate, const float* head_out, const float* dout, float* dx,
                         float* const* d_ln_w, float* const* d_ln_b, float* const* d_fc_w, float* const* d_fc_b, float* d_gate,
                         float* scratch, void* stream) {
  msgat_head_desc d1 = {d->B, d->C, d->N, d->Tout, 0, d->x_is_relu, d->gate_sb};
  if (int rc = head_check(&d1)) return rc;
  if (d->M < 1 || d->M > MSGAT_MAX_GROUP) { msgat_set_error("heads: 1..%d components", MSGAT_MAX_GROUP); return -42; }
  msgat_prof_tag("@C%dO%d", d->C, d->Tout);
  HeadArgs a = {};
  a.x = x; a.C = d->C; a.N = d->N; a.Tout = d->Tout; a.ntile = (d->N + HNT - 1) / HNT; a.J = head_J(&d1, a.ntile);
  a.x_is_relu = d->x_is_relu; a.gate = gate; a.gate_sb = d->gate_sb;
  a.head_out = const_cast<float*>(head_out); a.dout = dout; a.dx = dx; a.d_gate = d_gate; a.part = scratch;
  a.dgate_sb = d->gate_sb;   // d_gate has the gate's own [B, M, N, Tout] strides
  const int nW = d->Tout * 12 * d->C;
  a.pstride = nW + d->Tout + 24;
  const int64_t part_sm = msgat_head_scratch_floats(&d1);
  HeadGrp g = {};
  g.M = d->M; g.x_sm = d->x_sm; g.ho_sm = (int64_t)d->B * d->N * d->Tout; g.gate_sm = d->gate_sm;
  g.dx_sm = (int64_t)d->B * d->C * d->N * 12; g.dgate_sm = d->gate_sm; g.part_sm = part_sm;
  for (int m = 0; m < d->M; ++m) { g.ln_w[m] = ln_w[m]; g.ln_b[m] = ln_b[m]; g.fc_w[m] = fc_w[m]; g.fc_b[m] = nullptr; }
  int nblk = a.J;
  if (head_streaming(&d1)) {
    nblk = (d->N + HS_NODES - 1) / HS_NODES;
    size_t smem = ((size_t)d->C * HS_WP + HS_WARPS * HS_NODES * 12 + 2 * HS_NODES * 12 + 24 + 12 + HS_WARPS * 24 + HS_WARPS * HS_NODES * 12 + 144) * 4;
    a.Bn = d->B;
    MSGAT_LAUNCH(k_head_bwd12, dim3(head_stream_grid(d->N, d->B, d->M), 1, d->M), dim3(256), smem, stream, a, g);
  } else {
    size_t smem = ((size_t)d->C * HPS + (size_t)d->C * d->Tout * 12 + 24 + (size_t)d->C * HNT * 2 + HNT * d->Tout + 8 * 24) * 4;
    MSGAT_LAUNCH(k_head_bwd, dim3(a.J, d->B, d->M), dim3(256), smem, stream, a, g);
  }
  ReduceArgs r;
  int nj = 0, tot = 0;
  const int nparts = nblk * d->B;
  auto add = [&](float* const* dst, const float* src, int n, int tr = 0) {
    if (!dst || !dst[0]) return;
    RJob jb = {};
    for (int m = 0; m < d->M; ++m) jb.dst[m] = dst[m];
    jb.src = src; jb.n = n; jb.nparts = nparts; jb.dmul = 1; jb.doff = 0; jb.pstride = (int64_t)a.pstride; jb.f64 = 0;
    jb.tr = tr;
    r.j[nj] = jb;
    r.starts[nj] = tot; tot += n; ++nj;
  };
  add(d_fc_w, scratch, nW, head_streaming(&d1) ? 144 : 0);   // k_head_bwd12 writes channel-major partials
  add(d_fc_b, scratch + nW, d->Tout);
  add(d_ln_w, scratch + nW + d->Tout, 12);
  add(d_ln_b, scratch + nW + d->Tout + 12, 12);
  r.starts[nj] = tot; r.njobs = nj; r.sc_sm = part_sm;
  if (tot > 0) LAUNCH_REDUCE(tot, d->M, stream, r);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { msgat_set_error("heads_backward: %s", cudaGetErrorString(e)); return -100; }
  return 0;
}

}  // extern "C"
