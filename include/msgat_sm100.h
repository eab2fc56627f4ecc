/* msgat_sm100.h -- C ABI of libmsgat_sm100.so, the B200 (sm_100a) implementation of the
 * MS-GAT multirelational attention hot path.
 *
 * The reference (luokn/ms-gat) is pure Python/PyTorch and has no FFI: its seam for this
 * path is the nn.Module API of src/models.  The entry points below are
 * what the torch.autograd.Functions behind those modules bind through ctypes
 * (msgat_b200/_lib.py); INTEGRATION.md shows the reference-side stub.  Each entry point
 * names the reference code it replaces.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer to fp32 (or int32 where typed so) unless noted;
 *   - signals are [B, C, N, T] with T == 12 innermost (reference layout, msgat.py:121);
 *     the (n, t) plane must be contiguous, batch and channel strides are explicit;
 *   - the library never allocates caller-visible memory: outputs, saved statistics and
 *     scratch are caller-owned buffers whose sizes msgat_meam_sizes() reports;
 *   - all work is ordered on `stream` (a cudaStream_t passed as void*), nothing
 *     synchronises the device; msgat_meam_forward / _backward run their per-sample
 *     kernels on a side stream of the library that is forked from and joined back
 *     into `stream` with events inside the call (legal under stream capture;
 *     MSGAT_NO_FORK=1 in the environment disables it);
 *   - return value 0 on success, negative on error; msgat_last_error() gives the text.
 *     No CPU fallback exists: without an sm_100 device msgat_device_check() fails.
 */
#ifndef MSGAT_SM100_H
#define MSGAT_SM100_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MSGAT_T 12        /* in_timesteps is always timesteps_per_hour = 12 (main.py:45) */
#define MSGAT_RANK 10     /* hard-coded rank of Wt1/Wt2 (attention.py:58-59) */
#define MSGAT_MAX_CONV 4  /* longest TACN chain: msgat96 [1,1,2,2] (msgat.py:229) */
#define MSGAT_MAX_GROUP 8 /* components (TPCs) one grouped call may carry: `-i 1,2,3,24,168` has 5 (main.py:28) */

/* msgat_meam_desc.flags */
#define MSGAT_F_LN 1          /* LayerNorm([T]) on the input first          (msgat.py:122) */
#define MSGAT_F_RES 2         /* + Conv2d 1x1 residual of the raw input      (msgat.py:131) */
#define MSGAT_F_RELU 4        /* ReLU on the sum                             (msgat.py:131) */
#define MSGAT_F_CMIX 8        /* CACN: Conv2d 1x1 (+bias) after the attention (msgat.py:94) */
#define MSGAT_F_SMIX 16       /* GACN: x W^T after the attention             (msgat.py:27)  */
#define MSGAT_F_DOUT_MASKED 32 /* backward: dout already carries the ReLU mask */
#define MSGAT_F_X_IS_RELU 64  /* backward: x is the ReLU output of the previous MEAM; dx is
                                 returned already multiplied by (x > 0) so that the previous
                                 MEAM can set MSGAT_F_DOUT_MASKED */

#define MSGAT_PREC_FP32 0 /* fp32-grade results (1e-5 bar): CUDA-core fp32; the layer-2 channel mixes (forward, data and
                             weight gradients) on tcgen05 kind::tf32 with error-compensated operands (hi/lo split,
                             "3xTF32"); the N x N attention scores on tcgen05 with a three-way operand split, their
                             second contraction in fp32 FFMA2 with fp64 sums (spatial_tc32.cu) */
#define MSGAT_PREC_TF32 1 /* 2e-3 bar on outputs and parameter gradients: single-pass tcgen05 kind::tf32 channel mixes in
                             the forward, both contractions of the N x N score passes on tcgen05 (spatial_tc.cu) */

/* One MEAM-shaped operator.  With all three widths set, LN|RES|RELU|CMIX|SMIX it is
 * MEAM.forward (msgat.py:121-131); with a single width and no flags it is the bare
 * ChannelAttention / TemporalAttention / GraphAttention (attention.py:91,62,32); with
 * CMIX / n_conv / SMIX it is CACN / TACN / GACN (msgat.py:96,76,25). */
typedef struct msgat_meam_desc {
  int32_t B, C, N;      /* batch, input channels, nodes */
  int32_t Oc, Ot, Os;   /* output widths of the channel / temporal / spatial branch, 0 = absent.
                           Without CMIX Oc == C, without convs Ot == C, without SMIX Os == C. */
  int32_t flags;
  int32_t n_conv;       /* TACN: number of dilated causal convs, 0..MSGAT_MAX_CONV */
  int32_t dil[MSGAT_MAX_CONV];
  int32_t precision;    /* MSGAT_PREC_* */
  int32_t M;            /* grouped call: number of same-shaped components (0 or 1 = one operator).  MSGAT.forward
                           loops over M independent TPCs (msgat.py:203-204); a grouped call runs the same layer
                           of all of them in one launch per kernel, component m with its own parameters. */
  int64_t x_sb, x_sc;   /* element strides of x over batch and channel */
  int64_t x_sm;         /* element stride of x over components (grouped call; X[:, m] of [B,M,C,N,12] is x_sm = C*N*12) */
} msgat_meam_desc;

/* Parameters, named after the reference state_dict. */
typedef struct msgat_meam_params {
  const float *ln_w, *ln_b;               /* ln.weight, ln.bias            [12]            */
  const float *res_w, *res_b;             /* res.weight [O,C,1,1], res.bias [O]            */
  const float *ca_Wc, *ca_alpha;          /* cacn.seq.0.Wc [12,12], .alpha [N]             */
  const float *ca_w, *ca_b;               /* cacn.seq.1.weight [Oc,C,1,1], .bias [Oc]      */
  const float *ta_Wt1, *ta_Wt2, *ta_alpha; /* tacn.seq.0.Wt1/Wt2 [10,N], .alpha [C]        */
  const float *conv_w[MSGAT_MAX_CONV];    /* tacn.seq.{1,3,..}.weight [Ot,Cin,1,2]         */
  const float *conv_b[MSGAT_MAX_CONV];    /* tacn.seq.{1,3,..}.bias   [Ot]                 */
  const float *sa_W;                      /* gacn.W [Os,C]                                 */
  const float *sa_Wg, *sa_alpha;          /* gacn.gatt.Wg [12,12], .alpha [C]              */
} msgat_meam_params;

/* Gradients of the same parameters (written, not accumulated). Null = not wanted. */
typedef struct msgat_meam_grads {
  float *ln_w, *ln_b, *res_w, *res_b, *ca_Wc, *ca_alpha, *ca_w, *ca_b;
  float *ta_Wt1, *ta_Wt2, *ta_alpha;
  float *conv_w[MSGAT_MAX_CONV], *conv_b[MSGAT_MAX_CONV];
  float *sa_W, *sa_Wg, *sa_alpha;
} msgat_meam_grads;

/* Sparse view of the (constant) normalised adjacency, built once by msgat_adj_build
 * (replaces the dense [N,N] fp32 parameter of msgat.py:190 on the data path). */
typedef struct msgat_adj {
  int32_t N, nnz;
  const int32_t *row_ptr;  /* [N+1]   CSR                                           */
  const int32_t *col_idx;  /* [nnz]                                                 */
  const float *val;        /* [nnz]                                                 */
  const int32_t *csc_ptr;  /* [N+1]   CSC of the same pattern                       */
  const int32_t *csc_row;  /* [nnz]   row of each CSC entry                         */
  const int32_t *csc_eid;  /* [nnz]   index of that entry in the CSR arrays         */
} msgat_adj;

const char *msgat_version(void);
const char *msgat_last_error(void);
/* 0 iff the current device is compute capability 10.x. */
int msgat_device_check(void);

/* Launch accounting (bench.py): number of kernels this library has launched since load;
 * optional per-kernel CUDA-event timing on the launching stream. */
long long msgat_launch_count(void);
int msgat_profile(int enable);
int msgat_profile_report(char *buf, long long cap);

/* Adjacency preparation.  count: per-row non-zeros of dense adj [N,N] into row_nnz[N] and
 * col_nnz[N].  fill: given exclusive scans row_ptr/csc_ptr [N+1] (caller computes them),
 * write col_idx/val and csc_row/csc_eid. */
int msgat_adj_count(const float *adj, int32_t N, int32_t *row_nnz, int32_t *col_nnz, void *stream);
int msgat_adj_fill(const float *adj, int32_t N, const int32_t *row_ptr, const int32_t *csc_ptr,
                   int32_t *col_idx, float *val, int32_t *csc_row, int32_t *csc_eid, void *stream);

/* Sizes, in floats, of the `saved` buffer (forward -> backward) and of `scratch`
 * (max of forward and backward needs) for this descriptor and adjacency nnz, PER COMPONENT:
 * a grouped call (desc.M > 1) takes M times as much of each, component-major. */
int msgat_meam_sizes(const msgat_meam_desc *d, int32_t nnz, int64_t *saved_floats, int64_t *scratch_floats);

/* Forward of the MEAM-shaped operator: x [B,C,N,12] -> out [B,Oc+Ot+Os,N,12] (contiguous).
 * Grouped (desc.M > 1): `p` points to M parameter structs, component m reads x + m * x_sm and writes
 * out[m] of [M,B,O,N,12]; saved / scratch are M consecutive per-component buffers. */
int msgat_meam_forward(const msgat_meam_desc *d, const msgat_meam_params *p, const msgat_adj *adj,
                       const float *x, float *out, float *saved, float *scratch, void *stream);

/* Backward: dout [B,O,N,12] contiguous -> dx [B,C,N,12] contiguous and parameter grads.
 * `out` is the forward output (ReLU mask) and may be null when DOUT_MASKED or no RELU.
 * Grouped: dout / out are [M,B,O,N,12], dx is [M,B,C,N,12], `p` and `g` point to M structs. */
int msgat_meam_backward(const msgat_meam_desc *d, const msgat_meam_params *p, const msgat_adj *adj,
                        const float *x, const float *out, const float *dout, const float *saved,
                        float *scratch, float *dx, const msgat_meam_grads *g, void *stream);

/* TPC head (msgat.py:158-160) fused with the time-embedding gate of MSGAT.forward
 * (msgat.py:204): y[b,n,o] = (sum_{c,t} fc_w[o,t,0,c] LN_T(x)[b,c,n,t] + fc_b[o]);
 * out[b,n,o] (+)= y * gate[b,n,o] (gate null = 1).  head_out keeps y for the backward. */
typedef struct msgat_head_desc {
  int32_t B, C, N, Tout;
  int32_t accumulate;   /* forward: out += instead of out = */
  int32_t x_is_relu;    /* backward: return dx * (x > 0) */
  int64_t gate_sb;      /* batch stride of gate (elements); (n,o) contiguous */
} msgat_head_desc;
int msgat_head_forward(const msgat_head_desc *d, const float *x, const float *ln_w, const float *ln_b,
                       const float *fc_w, const float *fc_b, const float *gate, float *head_out,
                       float *out, void *stream);
int64_t msgat_head_scratch_floats(const msgat_head_desc *d);
int msgat_head_backward(const msgat_head_desc *d, const float *x, const float *ln_w, const float *ln_b,
                        const float *fc_w, const float *gate, const float *head_out, const float *dout,
                        float *dx, float *d_ln_w, float *d_ln_b, float *d_fc_w, float *d_fc_b,
                        float *d_gate, float *scratch, void *stream);

/* Grouped heads: the heads of all M components of MSGAT.forward (msgat.py:203-204) and the gated sum over them.
 * x is [M,B,C,N,12] with component stride x_sm; gate [B,M,N,Tout] with batch stride gate_sb and component stride
 * gate_sm (null = 1); head_out [M,B,N,Tout] keeps the un-gated heads for the backward; out [B,N,Tout].
 * The parameter arguments are host arrays of M device pointers.  Backward: dx [M,B,C,N,12], d_gate with the
 * gate's strides (may be null), scratch = M * msgat_heads_scratch_floats() floats. */
typedef struct msgat_heads_desc {
  int32_t B, C, N, Tout, M;
  int32_t x_is_relu;
  int64_t x_sm, gate_sb, gate_sm;
} msgat_heads_desc;
int msgat_heads_forward(const msgat_heads_desc *d, const float *x, const float *const *ln_w, const float *const *ln_b,
                        const float *const *fc_w, const float *const *fc_b, const float *gate, float *head_out, float *out,
                        void *stream);
int64_t msgat_heads_scratch_floats(const msgat_heads_desc *d);
int msgat_heads_backward(const msgat_heads_desc *d, const float *x, const float *const *ln_w, const float *const *ln_b,
                         const float *const *fc_w, const float *gate, const float *head_out, const float *dout, float *dx,
                         float *const *d_ln_w, float *const *d_ln_b, float *const *d_fc_w, float *const *d_fc_b, float *d_gate,
                         float *scratch, void *stream);

/* Step plumbing on the device.
 * msgat_loss_forward: Huber loss of engine.py:56 (loss.py:51-52): loss[0] = mean over n elements of
 *   0.5 r^2 if |r| <= delta else delta |r| - 0.5 delta^2, r = pred - target; dpred (may be null) = d loss / d pred;
 *   metrics (may be null) is the running state of metrics.py:20-35, double[4] += {sum |r|,
 *   100 * sum_{target > mask_value} |r / target|, sum r^2, n}: MAE = m[0]/m[3], MAPE = m[1]/m[3],
 *   RMSE = sqrt(m[2]/m[3]).  No host synchronisation (the reference reads four scalars back per step).
 *   scratch: msgat_loss_scratch_bytes() bytes, zeroed once by the caller before its first use.
 * msgat_loss_backward: out = dpred * gout[0] (gout: the device scalar autograd hands to the mean).
 * msgat_adam_step: torch.optim.Adam(lr, betas, eps, weight_decay) of engine.py:106 over one flat fp32 buffer;
 *   g is read as g * grad_scale (1 / world size of the gradient all-reduce); step[0] is the device-side count of
 *   steps taken so far and is advanced by the kernel; scratch: 16 bytes, zeroed once. */
int64_t msgat_loss_scratch_bytes(void);
int msgat_loss_forward(const float *pred, const float *target, int64_t n, float delta, float mask_value, float *loss,
                       float *dpred, double *metrics, void *scratch, void *stream);
int msgat_loss_backward(const float *dpred, const float *gout, int64_t n, float *out, void *stream);
int msgat_adam_step(float *p, const float *g, float *m, float *v, int64_t n, float lr, float beta1, float beta2,
                    float eps, float weight_decay, float grad_scale, int32_t *step, void *scratch, void *stream);

/* Data feed on the device: the sliding-window sampler of data_loader.py:92-115
 * (TimeSeriesSlice.__getitem__ + collate) over a time series resident in HBM.  inputs [C,N,L] (normalised series),
 * target [N,L] (raw channel 0), t [B] device int64 time indices; hours [M] HOST array (windows start at
 * t - hours[m] * tau); X [B,M,C,N,tau], Y [B,N,q], H/D [B] int64 (hour of day, day of week; may be null).
 * M <= 16, B <= 256 per call; a window outside [0,L) is written as NaN. */
int msgat_window_gather(const float *inputs, const float *target, const int64_t *t, int B, int C, int N, int64_t L,
                        const int32_t *hours, int M, int tau, int q, float *X, int64_t *H, int64_t *D, float *Y,
                        void *stream);

/* msgat_loss_forward_static: msgat_loss_forward without the scratch argument, for callers that cannot allocate
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  It has NO counterpart in
 * the reference.  Same kernel, grid and result as
 * msgat_loss_forward; its scratch is a 32 784-byte file-scope device array that every GPU loading the library holds
 * until exit, and the call allocates nothing (no cudaMalloc / cudaFree / allocating library routine; one 16-byte
 * cudaMemsetAsync on `stream` clears the arrival counter).  Limit: n <= msgat_loss_static_max_n() = 2^28 elements,
 * beyond which it returns -63 with a message.  All calls on a device share the array: calls that overlap on the GPU
 * (different streams without an ordering between them) race; calls on one stream are safe. */
int64_t msgat_loss_static_max_n(void);
int msgat_loss_forward_static(const float *pred, const float *target, int64_t n, float delta, float mask_value,
                              float *loss, float *dpred, double *metrics, void *stream);

/* Self-test of the tcgen05 (kind::tf32) building block of the tensor-core path; see
 * csrc/tc_probe.cu.  mode 0: D[256,N] = A[K,256]^T B[K,N]; mode 1: D[128,N] = A[Ra,P] B[Rb,P]^T. */
int msgat_tc_probe(int mode, const float *A, const float *B, float *D, int K, int N, int Ra, int Rb, int P,
                   void *stream);

#ifdef __cplusplus
}
#endif
#endif
