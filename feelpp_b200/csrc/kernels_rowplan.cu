// kernels_rowplan.cu -- owner-computes assembly (placeholder until the plan builder lands).
#include "feelgpu_internal.h"
int feelgpu_rowplan_build(feelgpu_pattern_t, const int64_t*, const int64_t*) { return FEELGPU_OK; }
void feelgpu_rowplan_destroy(feelgpu_pattern_t) {}
bool feelgpu_rowplan_supports(feelgpu_pattern_t, const feelgpu_form2_desc*) { return false; }
int feelgpu_launch_form2_rowplan(feelgpu_pattern_t pat, const FormParams&, const feelgpu_form2_desc*, int, double*,
                                 cudaStream_t)
{
    pat->ctx->last_error = "row plan not built";
    return FEELGPU_ERR_UNSUPPORTED;
}
