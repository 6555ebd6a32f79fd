// xlb_host.cuh -- host-side helpers shared by the translation units of libxlb_b200.so
#pragma once
#include <cuda_runtime.h>

#include "xlb_common.cuh"

namespace xlb {
// defined in xlb_capi.cu
int set_error(int code, const char* msg);
int count_launch(cudaError_t err);
}  // namespace xlb

// per-(lattice, policy) step entry points, defined by xlb_step_inst.cu instances
#define XLB_DECL_STEP(tag)                                                                           \
    int xlb_step_##tag(int collision, const xlb::StepParams& p, int variant, cudaStream_t st);       \
    int xlb_step_aa_##tag(int collision, const xlb::StepParams& p, int local_step, cudaStream_t st);
#define XLB_FOR_ALL_POLICIES(Q)      \
    XLB_DECL_STEP(q##Q##_f32f32)     \
    XLB_DECL_STEP(q##Q##_f32f16)     \
    XLB_DECL_STEP(q##Q##_f64f64)     \
    XLB_DECL_STEP(q##Q##_f64f32)     \
    XLB_DECL_STEP(q##Q##_f64f16)
XLB_FOR_ALL_POLICIES(9)
XLB_FOR_ALL_POLICIES(19)
XLB_FOR_ALL_POLICIES(27)
