// loss.cu -- the step plumbing around the hot path, on the device:
//   k_loss   Huber loss of engine.py:56 (loss.py:51-52), its gradient, and the three streaming metric sums of
//            metrics.py:20-35 (MAE / MAPE with the y > mask filter / RMSE) in ONE pass over pred and target,
//            no host synchronisation (the reference does four .item() per step, engine.py:66 + metrics.py:24,30,34);
//   k_scale  dpred * (upstream scalar gradient), the backward of the mean;
//   k_adam   torch.optim.Adam(lr, betas, eps, weight_decay) of engine.py:106 over ONE flat parameter buffer
//            (coupled L2 weight decay, bias correction from a device-side step counter so that the step is
//            CUDA-graph capturable), with the 1 / world_size of the gradient all-reduce folded into the read.
// Reductions are deterministic: per-block fp64 partials, summed in block order by the last block to finish.
#include "kernels.h"
#include "compat.cuh"

#define LOSS_THREADS 256
#define LOSS_MAX_BLOCKS 1024

struct LossArgs {
  const float *pred, *target;
  int64_t n;
  float delta, mask_value, inv_n;
  float* loss;
  float* dpred;
  double* metrics;   // [4] += sum |r|, 100 sum_{y > mask} |r / y|, sum r^2, n
  double* part;      // [gridDim.x][4]
  unsigned int* counter;
};

MSGAT_DEV void loss_elem(float p, float y, const LossArgs& a, float& d, double& hub, double& ae, double& ape, double& se) {
  const float r = p - y, ab = fabsf(r);
  const bool quad = ab <= a.delta;                       // loss.py:52: torch.where(|r| <= delta, l2, l1)
  hub += quad ? 0.5 * (double)r * (double)r : (double)a.delta * (double)ab - 0.5 * (double)a.delta * (double)a.delta;
  d = (quad ? r : (r > 0.f ? a.delta : (r < 0.f ? -a.delta : 0.f))) * a.inv_n;
  ae += (double)ab;                                      // metrics.py:24
  if (y > a.mask_value) ape += (double)fabsf(r / y);     // metrics.py:28-30
  se += (double)r * (double)r;                           // metrics.py:34
}

MSGAT_KERNEL MSGAT_LB(LOSS_THREADS, 4) k_loss(LossArgs a) {
  MSGAT_DYN_SMEM(smem_raw);
  double* red = reinterpret_cast<double*>(smem_raw);   // [LOSS_THREADS / 32][4]
  double hub = 0.0, ae = 0.0, ape = 0.0, se = 0.0;
  const int64_t n4 = a.n >> 2;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 p = ld4(a.pred + 4 * i), y = ld4(a.target + 4 * i);
    float4 d;
    loss_elem(p.x, y.x, a, d.x, hub, ae, ape, se);
    loss_elem(p.y, y.y, a, d.y, hub, ae, ape, se);
    loss_elem(p.z, y.z, a, d.z, hub, ae, ape, se);
    loss_elem(p.w, y.w, a, d.w, hub, ae, ape, se);
    if (a.dpred) st4(a.dpred + 4 * i, d);
  }
  if (blockIdx.x == 0 && (int64_t)threadIdx.x < (a.n & 3)) {   // tail of a length that is not a multiple of 4
    const int64_t i = (n4 << 2) + threadIdx.x;
    float d;
    loss_elem(a.pred[i], a.target[i], a, d, hub, ae, ape, se);
    if (a.dpred) a.dpred[i] = d;
  }
  // block sum in a fixed order: lanes by shuffle tree, warps in index order
  double v[4] = {hub, ae, ape, se};
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int k = 0; k < 4; ++k) {
    for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
    if (lane == 0) red[warp * 4 + k] = v[k];
  }
  __syncthreads();
  int& last = *reinterpret_cast<int*>(red + (LOSS_THREADS / 32) * 4);
  if (threadIdx.x == 0) {
    double s[4] = {0.0, 0.0, 0.0, 0.0};
    for (int w = 0; w < (int)blockDim.x / 32; ++w)
      for (int k = 0; k < 4; ++k) s[k] += red[w * 4 + k];
    for (int k = 0; k < 4; ++k) a.part[(int64_t)blockIdx.x * 4 + k] = s[k];
    __threadfence();
    last = atomicAdd(a.counter, 1u) == gridDim.x - 1 ? 1 : 0;
  }
  __syncthreads();
  if (!last || threadIdx.x != 0) return;
  __threadfence();
  double s[4] = {0.0, 0.0, 0.0, 0.0};
  for (unsigned int b = 0; b < gridDim.x; ++b)   // block order: independent of which block finished last
    for (int k = 0; k < 4; ++k) s[k] += a.part[(int64_t)b * 4 + k];
  a.loss[0] = (float)(s[0] / (double)a.n);       // .mean(), loss.py:52
  if (a.metrics) {
    a.metrics[0] += s[1];
    a.metrics[1] += 100.0 * s[2];
    a.metrics[2] += s[3];
    a.metrics[3] += (double)a.n;
  }
  *a.counter = 0u;   // ready for the next launch (and for every replay of a captured graph)
}

MSGAT_KERNEL k_scale(const float* src, const float* scalar, float* dst, int64_t n) {
  const float s = scalar[0];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) dst[i] = src[i] * s;
}

struct AdamArgs {
  float *p, *m, *v;
  const float* g;
  int64_t n;
  float lr, beta1, beta2, eps, weight_decay, grad_scale;
  int* step;                // device-side step counter (number of steps taken so far)
  unsigned int* counter;
};

// torch.optim.Adam's single-tensor update (engine.py:106: lr 1e-3, weight_decay 5e-4, amsgrad off):
//   g += wd * p;  m = b1 m + (1 - b1) g;  v = b2 v + (1 - b2) g^2;
//   p -= lr / (1 - b1^t) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)
MSGAT_KERNEL MSGAT_LB(256, 4) k_adam(AdamArgs a) {
  const int t = *a.step + 1;
  const float c1 = 1.f - powf(a.beta1, (float)t), c2s = sqrtf(1.f - powf(a.beta2, (float)t));
  const float step_size = a.lr / c1;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += (int64_t)gridDim.x * blockDim.x) {
    const float p = a.p[i];
    const float g = fmaf(a.weight_decay, p, a.g[i] * a.grad_scale);
    const float m = a.beta1 * a.m[i] + (1.f - a.beta1) * g;
    const float v = a.beta2 * a.v[i] + (1.f - a.beta2) * g * g;
    a.m[i] = m; a.v[i] = v;
    a.p[i] = p - step_size * (m / (sqrtf(v) / c2s + a.eps));
  }
  // every block has read the counter before it arrives here: the last one to arrive advances it
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(a.counter, 1u) == gridDim.x - 1) { *a.step = t; *a.counter = 0u; }
  }
}

static int grid_for(int64_t n, int per_block) {
  int64_t g = (n + per_block - 1) / per_block;
  const int64_t cap = (int64_t)msgat_sm_count() * 4;
  if (g > cap) g = cap;
  if (g > LOSS_MAX_BLOCKS) g = LOSS_MAX_BLOCKS;
  return g < 1 ? 1 : (int)g;
}

extern "C" {

int64_t msgat_loss_scratch_bytes(void) { return (int64_t)LOSS_MAX_BLOCKS * 4 * sizeof(double) + 16; }

int msgat_loss_forward(const float* pred, const float* target, int64_t n, float delta, float mask_value, float* loss,
                       float* dpred, double* metrics, void* scratch, void* stream) {
  if (n <= 0 || !pred || !target || !loss || !scratch) { msgat_set_error("msgat_loss_forward: bad arguments"); return -60; }
  LossArgs a;
  a.pred = pred; a.target = target; a.n = n; a.delta = delta; a.mask_value = mask_value; a.inv_n = (float)(1.0 / (double)n);
  a.loss = loss; a.dpred = dpred; a.metrics = metrics;
  a.counter = reinterpret_cast<unsigned int*>(scratch);   // zero before the first use (the kernel re-arms it)
  a.part = reinterpret_cast<double*>(reinterpret_cast<char*>(scratch) + 16);
  MSGAT_LAUNCH(k_loss, dim3(grid_for(n, LOSS_THREADS * 4 * 4)), dim3(LOSS_THREADS), ((LOSS_THREADS / 32) * 4 + 1) * sizeof(double), stream, a);
  return 0;
}

int msgat_loss_backward(const float* dpred, const float* gout, int64_t n, float* out, void* stream) {
  if (n <= 0 || !dpred || !gout || !out) { msgat_set_error("msgat_loss_backward: bad arguments"); return -61; }
  MSGAT_LAUNCH(k_scale, dim3(grid_for(n, 1024)), dim3(256), 0, stream, dpred, gout, out, n);
  return 0;
}

int msgat_adam_step(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2, float eps,
                    float weight_decay, float grad_scale, int32_t* step, void* scratch, void* stream) {
  if (n <= 0 || !p || !g || !m || !v || !step || !scratch) { msgat_set_error("msgat_adam_step: bad arguments"); return -62; }
  AdamArgs a;
  a.p = p; a.g = g; a.m = m; a.v = v; a.n = n;
  a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.weight_decay = weight_decay; a.grad_scale = grad_scale;
  a.step = step; a.counter = reinterpret_cast<unsigned int*>(scratch);
  MSGAT_LAUNCH(k_adam, dim3(grid_for(n, 1024)), dim3(256), 0, stream, a);
  return 0;
}

}  // extern "C"

// ---- the allocation-free variant (requested by the run's operators; no counterpart in the reference) ------------
// Same kernel and grid as msgat_loss_forward, with the scratch (arrival counter + per-block partial sums) in a
// file-scope device array instead of caller memory: nothing on this path calls cudaMalloc / cudaFree or an
// allocating library routine, so it can run under CUDA graph capture or in harnesses that forbid allocation.
// Every device that loads the library holds the MSGAT_LOSS_STATIC_SCRATCH_BYTES array until exit.  All calls on a
// device share it: two calls that overlap on the GPU (different streams, no ordering between them) race.
#define MSGAT_LOSS_STATIC_MAX_N (1ll << 28)   // elements per call; the grid is capped, the limit keeps the contract explicit
#define MSGAT_LOSS_STATIC_SCRATCH_BYTES (LOSS_MAX_BLOCKS * 4 * 8 + 16)
#ifdef MSGAT_EMU
static unsigned char g_loss_static_scratch[MSGAT_LOSS_STATIC_SCRATCH_BYTES];
static int loss_static_scratch(void** out) { *out = g_loss_static_scratch; return 0; }
#else
static __device__ __align__(256) unsigned char g_loss_static_scratch[MSGAT_LOSS_STATIC_SCRATCH_BYTES];
#define LOSS_STATIC_MAX_DEVICES 64
// In host code g_loss_static_scratch is only a handle of the symbol, not a device address: the address is asked of
// the runtime once per device, with that device current, and cached.
static int loss_static_scratch(void** out) {
  static std::atomic<void*> cache[LOSS_STATIC_MAX_DEVICES];
  const int dev = msgat_current_device();
  if (dev < 0 || dev >= LOSS_STATIC_MAX_DEVICES) { msgat_set_error("msgat_loss_forward_static: device ordinal out of range"); return -64; }
  void* p = cache[dev].load(std::memory_order_acquire);
  if (!p) {
    const cudaError_t e = cudaGetSymbolAddress(&p, g_loss_static_scratch);
    if (e != cudaSuccess || !p) {
      msgat_set_error("msgat_loss_forward_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
      return -65;
    }
    cache[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return 0;
}
#endif

extern "C" {

int64_t msgat_loss_static_max_n(void) { return MSGAT_LOSS_STATIC_MAX_N; }

int msgat_loss_forward_static(const float* pred, const float* target, int64_t n, float delta, float mask_value, float* loss,
                              float* dpred, double* metrics, void* stream) {
  if (n <= 0 || !pred || !target || !loss) { msgat_set_error("msgat_loss_forward_static: bad arguments"); return -60; }
  if (n > MSGAT_LOSS_STATIC_MAX_N) {
    msgat_set_error("msgat_loss_forward_static: n exceeds the limit of 2^28 elements of the static-scratch variant; "
                    "use msgat_loss_forward with caller-provided scratch");
    return -63;
  }
  void* scratch = nullptr;
  const int rc = loss_static_scratch(&scratch);
  if (rc != 0) return rc;
  // the array keeps what the last call left: the arrival counter has to start at zero (the kernel re-arms it, but a
  // call that was aborted half-way would not have); the partial sums are fully overwritten before they are read
  if (cudaMemsetAsync(scratch, 0, 16, (cudaStream_t)stream) != 0) { msgat_set_error("msgat_loss_forward_static: cudaMemsetAsync failed"); return -66; }
  LossArgs a;
  a.pred = pred; a.target = target; a.n = n; a.delta = delta; a.mask_value = mask_value; a.inv_n = (float)(1.0 / (double)n);
  a.loss = loss; a.dpred = dpred; a.metrics = metrics;
  a.counter = reinterpret_cast<unsigned int*>(scratch);
  a.part = reinterpret_cast<double*>(reinterpret_cast<char*>(scratch) + 16);
  MSGAT_LAUNCH(k_loss, dim3(grid_for(n, LOSS_THREADS * 4 * 4)), dim3(LOSS_THREADS), ((LOSS_THREADS / 32) * 4 + 1) * sizeof(double), stream, a);
  return 0;
}

}  // extern "C"
