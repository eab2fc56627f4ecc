// placeholder, replaced below
#include "kernels.h"
int64_t meam_backward_scratch(Plan& p) { return 0; }
int meam_backward_impl(Plan& p, const msgat_meam_params*, const msgat_adj*, const float*, const float*, const float*,
                       const float*, float*, float*, const msgat_meam_grads*, void*) { msgat_set_error("nyi"); return -99; }
