// reduce.cuh -- one launch that sums every per-block partial into the caller's gradient tensors.
//
// A thread block owns 32 consecutive output elements; its 8 warps split the partials
// (warp w takes parts w, w+8, ...), so every warp-wide load is one coalesced 128-byte row
// and eight independent chains are in flight per element.  The eight warp sums are combined
// in a fixed order: the result does not depend on scheduling.
#pragma once
#include "compat.cuh"

#define MAX_RJOBS 28
#define RED_WARPS 8
// Grouped: blockIdx.y = component m writes dst[m] from src + m * sc_sm (floats).
// f64: src holds doubles (pstride in doubles).  tr > 0: the partials are [n / tr][tr] (e.g. channel-major fc weight gradients, written
// by their producer as contiguous 16-byte quads) and the destination is the transposed [tr][n / tr] tensor.
struct RJob { float* dst[MSGAT_MAX_GROUP]; const float* src; int n, nparts, dmul, doff; int64_t pstride; int f64; int tr; };
struct ReduceArgs { RJob j[MAX_RJOBS]; int starts[MAX_RJOBS + 1]; int njobs; int64_t sc_sm; };
#define LAUNCH_REDUCE(tot, M, stream, r) \
  MSGAT_LAUNCH(k_reduce, dim3(((tot) + 31) / 32, (M)), dim3(32 * RED_WARPS), RED_WARPS * 32 * sizeof(double), stream, r)

MSGAT_KERNEL MSGAT_LB(32 * RED_WARPS, 4) k_reduce(ReduceArgs a) {
  MSGAT_DYN_SMEM(red_raw);
  double (*part)[32] = reinterpret_cast<double (*)[32]>(red_raw);  // [RED_WARPS][32]
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int idx = blockIdx.x * 32 + lane;
  const bool live = idx < a.starts[a.njobs];
  int q = 0, e = 0;
  double s = 0.0;  // partials of cancellation-heavy scalars (alpha, Wc ...) are summed in fp64
  if (live) {
    while (idx >= a.starts[q + 1]) ++q;
    e = idx - a.starts[q];
    const float* src = a.j[q].src + (int64_t)blockIdx.y * a.sc_sm + e;
    const int64_t ps = a.j[q].pstride;
    const int np = a.j[q].nparts;
    int p = a.j[q].f64 ? np : w;
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    for (; p + 3 * RED_WARPS < np; p += 4 * RED_WARPS) {
      const float v0 = src[(int64_t)p * ps], v1 = src[(int64_t)(p + RED_WARPS) * ps];
      const float v2 = src[(int64_t)(p + 2 * RED_WARPS) * ps], v3 = src[(int64_t)(p + 3 * RED_WARPS) * ps];
      s0 += (double)v0; s1 += (double)v1; s2 += (double)v2; s3 += (double)v3;
    }
    for (; p < np; p += RED_WARPS) s0 += (double)src[(int64_t)p * ps];
    if (a.j[q].f64) {
      const double* sd = reinterpret_cast<const double*>(a.j[q].src + (int64_t)blockIdx.y * a.sc_sm) + e;
      s0 = s1 = s2 = s3 = 0.0;
      for (p = w; p < np; p += RED_WARPS) s0 += sd[(int64_t)p * ps];
    }
    s = (s0 + s1) + (s2 + s3);
  }
  part[w][lane] = s;
  __syncthreads();
  if (w == 0 && live) {
    double t = 0.0;
    for (int k = 0; k < RED_WARPS; ++k) t += part[k][lane];
    const RJob& jb = a.j[q];
    const int de = jb.tr > 0 ? (e % jb.tr) * (jb.n / jb.tr) + e / jb.tr : e;
    jb.dst[blockIdx.y][(int64_t)de * jb.dmul + jb.doff] = (float)t;
  }
}
