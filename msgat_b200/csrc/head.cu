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
  const float* x; int C, N, Tout, ntile, J, accumulate, x_is_relu;
  const float *ln_w, *ln_b, *fc_w, *fc_b, *gate; int64_t gate_sb;
  float *head_out, *out;
  // backward
  const float* dout; float *dx, *d_gate, *part; int pstride;
};

// weights staged as Wh[c][o][t] (t contiguous): fc.weight is [Tout, 12, 1, C]
MSGAT_DEV void stage_fc(float* Wh, const float* fc_w, int C, int Tout) {
  for (int i = threadIdx.x; i < C * Tout * 12; i += blockDim.x) {
    const int c = i / (Tout * 12), r = i - c * Tout * 12, o = r / 12, t = r - o * 12;
    Wh[i] = fc_w[((int64_t)o * 12 + t) * C + c];
  }
}

MSGAT_KERNEL MSGAT_LB(256, 1) k_head_fwd(HeadArgs a) {
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
      a.out[g] = a.accumulate ? a.out[g] + s * gt : s * gt;
    }
  }
}

#define HPAIRS 5  // (c, o) pairs of the fc weight gradient per thread

MSGAT_KERNEL MSGAT_LB(256, 1) k_head_bwd(HeadArgs a) {
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
        if (a.d_gate) a.d_gate[g] = d * a.head_out[g];
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

static int head_J(const msgat_head_desc* d, int ntile) {
  int J = (2 * 148) / d->B;
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

extern "C" {

int msgat_head_forward(const msgat_head_desc* d, const float* x, const float* ln_w, const float* ln_b, const float* fc_w,
                       const float* fc_b, const float* gate, float* head_out, float* out, void* stream) {
  if (int rc = head_check(d)) return rc;
  msgat_prof_tag("@C%dO%d", d->C, d->Tout);
  HeadArgs a = {};
  a.x = x; a.C = d->C; a.N = d->N; a.Tout = d->Tout; a.ntile = (d->N + HNT - 1) / HNT; a.J = head_J(d, a.ntile);
  a.accumulate = d->accumulate; a.ln_w = ln_w; a.ln_b = ln_b; a.fc_w = fc_w; a.fc_b = fc_b; a.gate = gate;
  a.gate_sb = d->gate_sb; a.head_out = head_out; a.out = out;
  size_t smem = ((size_t)d->C * HPS + (size_t)d->C * d->Tout * 12 + 24) * 4;
  MSGAT_LAUNCH(k_head_fwd, dim3(a.J, d->B), dim3(256), smem, stream, a);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { msgat_set_error("head_forward: %s", cudaGetErrorString(e)); return -100; }
  return 0;
}

int64_t msgat_head_scratch_floats(const msgat_head_desc* d) {
  const int ntile = (d->N + HNT - 1) / HNT;
  return (int64_t)head_J(d, ntile) * d->B * (d->Tout * 12 * d->C + d->Tout + 24);
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
  const int nW = d->Tout * 12 * d->C;
  a.pstride = nW + d->Tout + 24;
  size_t smem = ((size_t)d->C * HPS + (size_t)d->C * d->Tout * 12 + 24 + (size_t)d->C * HNT * 2 + HNT * d->Tout + 8 * 24) * 4;
  MSGAT_LAUNCH(k_head_bwd, dim3(a.J, d->B), dim3(256), smem, stream, a);
  ReduceArgs r;
  int nj = 0, tot = 0;
  const int nparts = a.J * d->B;
  auto add = [&](float* dst, const float* src, int n) {
    if (!dst) return;
    r.j[nj] = RJob{dst, src, n, nparts, 1, 0, (int64_t)a.pstride, 0};
    r.starts[nj] = tot; tot += n; ++nj;
  };
  add(d_fc_w, scratch, nW);
  add(d_fc_b, scratch + nW, d->Tout);
  add(d_ln_w, scratch + nW + d->Tout, 12);
  add(d_ln_b, scratch + nW + d->Tout + 12, 12);
  r.starts[nj] = tot; r.njobs = nj;
  if (tot > 0) LAUNCH_REDUCE(tot, stream, r);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { msgat_set_error("head_backward: %s", cudaGetErrorString(e)); return -100; }
  return 0;
}

}  // extern "C"
