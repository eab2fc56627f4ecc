// reduce.cuh -- one launch that sums every per-block partial into the caller's gradient tensors.
#pragma once
#include "compat.cuh"

#define MAX_RJOBS 28
struct RJob { float* dst; const float* src; int n, nparts, dmul, doff; int64_t pstride; };
struct ReduceArgs { RJob j[MAX_RJOBS]; int starts[MAX_RJOBS + 1]; int njobs; };

MSGAT_KERNEL k_reduce(ReduceArgs a) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= a.starts[a.njobs]) return;
  int q = 0;
  while (idx >= a.starts[q + 1]) ++q;
  const RJob& jb = a.j[q];
  const int e = idx - a.starts[q];
  double s = 0.0;  // partials of cancellation-heavy scalars (alpha, Wc ...) are summed in fp64
  for (int p = 0; p < jb.nparts; ++p) s += (double)jb.src[(int64_t)p * jb.pstride + e];
  jb.dst[(int64_t)e * jb.dmul + jb.doff] = (float)s;
}

