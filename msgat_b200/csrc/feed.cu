// feed.cu -- the data feed of the training loop on the device: the sliding-window sampler of
// data_loader.py:92-115 (TimeSeriesSlice.__getitem__ + the DataLoader's collate) as one gather kernel over a
// time series that stays resident in HBM.
//
// The reference keeps the normalised series [C, N, L] on the host, cuts M windows of tau steps per sample with
// Python slicing, stacks them (data_loader.py:110), collates B samples and copies X[B,M,C,N,tau] to the GPU every
// step (engine.py:50): 8.1 MB of H2D per PEMSD7 batch.  Here a batch is B time indices; for sample b at step t:
//   X[b,m,c,n,k] = inputs[c,n, t - hours[m] * tau + k]        k < tau          data_loader.py:110
//   Y[b,n,k]     = target[n, t + k]                           k < q            data_loader.py:111
//   H[b]         = (t / tau) % 24,  D[b] = (t / tau / 24) % 7                  data_loader.py:108-109,112
// Pure index arithmetic, bit-exact.  A window that leaves [0, L) is a caller error (the reference would return a
// short slice and fail in the collate); it is written as NaN so that it cannot pass unnoticed.
#include <math.h>
#include "kernels.h"
#include "compat.cuh"

#define FEED_MAX_HOURS 16

struct FeedArgs {
  const float *inputs, *target;
  const int64_t* t;
  int B, C, N, M, tau, q;
  int64_t L;
  int hours[FEED_MAX_HOURS];
  float *X, *Y;
  int64_t *H, *D;
  int64_t nx, ny;
};

MSGAT_KERNEL MSGAT_LB(256, 4) k_window_gather(FeedArgs a) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const float nan = nanf("");
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.nx + a.ny; i += stride) {
    if (i < a.nx) {
      const int k = (int)(i % a.tau);
      int64_t r = i / a.tau;
      const int n = (int)(r % a.N); r /= a.N;
      const int c = (int)(r % a.C); r /= a.C;
      const int m = (int)(r % a.M);
      const int b = (int)(r / a.M);
      const int64_t s = a.t[b] - (int64_t)a.hours[m] * a.tau + k;
      a.X[i] = (s >= 0 && s < a.L) ? a.inputs[((int64_t)c * a.N + n) * a.L + s] : nan;
    } else {
      const int64_t j = i - a.nx;
      const int k = (int)(j % a.q);
      const int64_t r = j / a.q;
      const int n = (int)(r % a.N);
      const int b = (int)(r / a.N);
      const int64_t s = a.t[b] + k;
      a.Y[j] = (s >= 0 && s < a.L) ? a.target[(int64_t)n * a.L + s] : nan;
    }
  }
  if (blockIdx.x == 0 && (int)threadIdx.x < a.B) {
    const int64_t h = a.t[threadIdx.x] / a.tau;   // t >= 0: truncation = floor (data_loader.py:108-109 rounding_mode="trunc")
    if (a.H) a.H[threadIdx.x] = h % 24;
    if (a.D) a.D[threadIdx.x] = (h / 24) % 7;
  }
}

extern "C" int msgat_window_gather(const float* inputs, const float* target, const int64_t* t, int B, int C, int N, int64_t L,
                                   const int32_t* hours, int M, int tau, int q, float* X, int64_t* H, int64_t* D, float* Y,
                                   void* stream) {
  if (!inputs || !target || !t || !hours || !X || !Y || B <= 0 || C <= 0 || N <= 0 || L <= 0 || M <= 0 || tau <= 0 || q <= 0) {
    msgat_set_error("msgat_window_gather: bad arguments");
    return -70;
  }
  if (M > FEED_MAX_HOURS) { msgat_set_error("msgat_window_gather: at most %d windows per sample (got %d)", FEED_MAX_HOURS, M); return -71; }
  if (B > 256) { msgat_set_error("msgat_window_gather: at most 256 samples per call (got %d)", B); return -72; }
  FeedArgs a;
  a.inputs = inputs; a.target = target; a.t = t; a.B = B; a.C = C; a.N = N; a.M = M; a.tau = tau; a.q = q; a.L = L;
  for (int m = 0; m < FEED_MAX_HOURS; ++m) a.hours[m] = m < M ? hours[m] : 0;
  a.X = X; a.Y = Y; a.H = H; a.D = D;
  a.nx = (int64_t)B * M * C * N * tau; a.ny = (int64_t)B * N * q;
  int64_t g = (a.nx + a.ny + 256 * 8 - 1) / (256 * 8);
  const int64_t cap = (int64_t)msgat_sm_count() * 8;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  MSGAT_LAUNCH(k_window_gather, dim3((unsigned)g), dim3(256), 0, stream, a);
  return 0;
}
