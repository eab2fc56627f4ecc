// kernels.h -- internal host-side entry points shared by the translation units.
#pragma once
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include "plan.h"
#include "compat.cuh"

int launch_sstats(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, float* saved, void* stream);
int launch_scol(const Plan& p, const float* Wg, const float* saved, const float* delta, float* dq, void* stream);
// l2_sparse_bwd.cu: transposed sparse aggregation GZ, dS at the non-zeros and its row sums, all components in one launch
int launch_l2_gz(const Plan& p, const msgat_adj* adj, const float* G, int64_t g_sm, const float* saved, float* tcA, float* delta,
                 float* dSnz, void* stream);
// spatial_tc.cu: the same two passes on tcgen05 (MSGAT_PREC_TF32)
bool score_tc_supported(const Plan& p);
int launch_sstats_tc(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, float* saved, void* stream);
int launch_scol_tc(const Plan& p, const float* saved, const float* delta, float* dq, void* stream);
// spatial_tc32.cu: the scores on tcgen05 at fp32 grade (three-way operand split), the rest as in spatial.cu (MSGAT_PREC_FP32)
bool score_tc32_supported(const Plan& p);
int launch_sstats_tc32(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, float* saved, void* stream);
int launch_scol_tc32(const Plan& p, const float* saved, const float* delta, float* dq, void* stream);
int meam_forward_impl(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, const float* x, float* out,
                      float* saved, float* scratch, void* stream);
int meam_backward_impl(Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, const float* x, const float* out,
                       const float* dout, const float* saved, float* scratch, float* dx, const msgat_meam_grads* g,
                       void* stream);
int64_t meam_backward_scratch(Plan& p);
void msgat_set_error(const char* fmt, ...);
// api.cu: side = msgat_fork(stream) makes the device's side stream wait for everything queued on `stream` and returns it (or
// `stream` itself when forking is off); msgat_join(stream, side) makes `stream` wait for the side work.  Capture-safe.
#ifdef MSGAT_EMU
static inline void* msgat_fork(void* stream) { return stream; }
static inline void msgat_join(void*, void*) {}
#else
void* msgat_fork(void* stream);
void msgat_join(void* stream, void* side);
#endif
// CTAs per sample of a one-CTA-per-SM tcgen05 kernel whose CTAs stay inside a sample (per-sample weights):
// `waves` rounds of 148 CTAs; more waves waste fewer SMs (148 is not a multiple of B) but repeat the set-up.
static inline int tc_ctas_per_sample(int B, int ntile, int waves, const char* env) {
  const char* e = getenv(env);
  if (e && atoi(e) > 0) waves = atoi(e);
  int J = (msgat_sm_count() * waves) / B;
  if (J > ntile) J = ntile;
  return J < 1 ? 1 : J;
}
// first-layer streaming kernels, meam_l1.cu (C = 1..4, full MEAM, two convs), natively grouped
bool l1_fwd_supported(const Plan& p);
int launch_l1_fwd(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, const float* x, float* out,
                  float* saved, void* stream);
bool l1_bwd_supported(const Plan& p);
int l1_bwd_blocks_per_sample(const Plan& p);
int l1_bwd_qsz(const Plan& p);
int launch_l1_bwd(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, const float* x, const float* G, int64_t g_sm,
                  const float* saved, float* DH, float* DXR, float* DA, float* delta, float* dSnz, float* dMc_part,
                  float* dPt_part, float* Qpart, void* stream);
int launch_l1_wfin(const Plan& p, const msgat_meam_params* prm, const msgat_meam_grads* g, const float* Qpart, int nparts, void* stream);
// layer-2 pipelined tcgen05 kernels, l2_*.cu (C % 8 == 0, full MEAM, two convs), natively grouped
bool l2_fwd_supported(const Plan& p);
int launch_l2_mix(const Plan& p, const msgat_meam_params* prm, const float* x, float* saved, float* scratch, void* stream);
int launch_l2_out(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, const float* x, float* out, float* saved,
                  void* stream);
// tensor-core (tcgen05 kind::tf32) kernels, tc_*.cu: MSGAT_PREC_TF32 on the layer-2 MEAM shape
bool tc_apply_supported(const Plan& p);
bool tc_apply3_supported(const Plan& p);
int launch_tc_apply3(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, const float* x, float* out,
                     float* saved, void* stream);
bool tc_wgrad_supported(const Plan& p);
int tc_wgrad_blocks_per_sample(const Plan& p);
// grouped: all p.M components in one launch (inputs x_sm / g_sm apart, saved / scratch slices)
int launch_tc_wgrad(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, const float* x, const float* G, int64_t g_sm,
                    const float* saved, const float* tcA, float* p_res, float* p_ws, float* p_c1past, float* p_c1now, float* p_c2past,
                    float* p_c2now, int64_t wg_stride, float* p_mc, int64_t mc_stride, void* stream);
bool tc_dgrad_supported(const Plan& p);
int tc_dgrad_blocks_per_sample(const Plan& p);
// grouped: all p.M components in one launch (inputs x_sm / g_sm apart, saved / scratch slices, dx dx_sm apart)
int launch_tc_dgrad(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, const float* x, const float* G, int64_t g_sm,
                    const float* saved, float* tcA, float* dx, int64_t dx_sm, float* bias_part, int nbias, float* ln_part,
                    double* dPt_part, void* stream);
int launch_tc_apply(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, const float* x, float* out,
                    float* saved, void* stream);
