// One instantiation family of the tensor-core step kernel (csrc/tc_step_kernel.cuh); compiled once per WX_TC_FT:
// 0 = feature count read at run time, 5 / 20 = feature count fixed at compile time.
#include "tc_step_kernel.cuh"

#ifndef WX_TC_FT
#error "compile with -DWX_TC_FT=<0|5|20>"
#endif

#define WX_TC_CAT2(a, b) a##b
#define WX_TC_CAT(a, b) WX_TC_CAT2(a, b)

int WX_TC_CAT(wx_tc_launch_ft, WX_TC_FT)(wx_ctx* ctx, wxk::GridArgs& a, cudaStream_t st, const char* who) {
    return wxtcstep::tc_launch<WX_TC_FT>(ctx, a, st, who);
}
