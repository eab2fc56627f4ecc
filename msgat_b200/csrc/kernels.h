// kernels.h -- internal host-side entry points shared by the translation units.
#pragma once
#include <math.h>
#include <stdint.h>
#include "plan.h"

int launch_sstats(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, float* saved, void* stream);
int launch_scol(const Plan& p, const float* Wg, const float* saved, const float* delta, float* dq, void* stream);
int meam_forward_impl(const Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, const float* x, float* out,
                      float* saved, float* scratch, void* stream);
int meam_backward_impl(Plan& p, const msgat_meam_params* prm, const msgat_adj* adj, const float* x, const float* out,
                       const float* dout, const float* saved, float* scratch, float* dx, const msgat_meam_grads* g,
                       void* stream);
int64_t meam_backward_scratch(Plan& p);
void msgat_set_error(const char* fmt, ...);
