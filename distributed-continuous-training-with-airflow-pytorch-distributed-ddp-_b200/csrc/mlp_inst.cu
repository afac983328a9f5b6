// Instantiations of the step kernels for one group of feature counts (built four times, -DWX_F_GROUP=0..3).
#include "mlp_kernels.cuh"

#ifndef WX_F_GROUP
#error "build with -DWX_F_GROUP=0..3"
#endif

namespace wxk {

template <int F>
static grid_kernel_t pick_mode(int mode, int loader) {
    switch (mode) {
        case MODE_TRAIN:
            if (loader == LOAD_RAW_F32) return mlp_grid_kernel<F, MODE_TRAIN, LOAD_RAW_F32>;
            if (loader == LOAD_RAW_F64) return mlp_grid_kernel<F, MODE_TRAIN, LOAD_RAW_F64>;
            return mlp_grid_kernel<F, MODE_TRAIN, LOAD_ROWMAJOR>;
        case MODE_EVAL: return mlp_grid_kernel<F, MODE_EVAL, LOAD_ROWMAJOR>;
        default: return mlp_grid_kernel<F, MODE_PREDICT, LOAD_ROWMAJOR>;
    }
}

#define WX_GROUP_F(X) X(WX_F_GROUP * 8 + 1) X(WX_F_GROUP * 8 + 2) X(WX_F_GROUP * 8 + 3) X(WX_F_GROUP * 8 + 4) \
    X(WX_F_GROUP * 8 + 5) X(WX_F_GROUP * 8 + 6) X(WX_F_GROUP * 8 + 7) X(WX_F_GROUP * 8 + 8)

#define WX_CAT2(a, b) a##b
#define WX_CAT(a, b) WX_CAT2(a, b)

grid_kernel_t WX_CAT(pick_grid_kernel_g, WX_F_GROUP)(int f, int mode, int loader, int* smem) {
    switch (f) {
#define X(FV)                                  \
    case FV:                                   \
        *smem = smem_bytes<FV>(loader, mode);  \
        return pick_mode<FV>(mode, loader);
        WX_GROUP_F(X)
#undef X
        default: return nullptr;
    }
}

small_kernel_t WX_CAT(pick_small_kernel_g, WX_F_GROUP)(int f, int* smem) {
    switch (f) {
#define X(FV)                                                                       \
    case FV:                                                                        \
        *smem = (4 * Dim<FV>::SLAB + WARPS * 2 * R * Dim<FV>::FS) * 4;              \
        return mlp_small_steps_kernel<FV>;
        WX_GROUP_F(X)
#undef X
        default: return nullptr;
    }
}

small_kernel_t WX_CAT(pick_resident_kernel_g, WX_F_GROUP)(int f, int* smem) {
    switch (f) {
#define X(FV)                                                                       \
    case FV:                                                                        \
        *smem = (4 * Dim<FV>::SLAB + WARPS * 2 * R * Dim<FV>::FS) * 4;              \
        return mlp_resident_steps_kernel<FV>;
        WX_GROUP_F(X)
#undef X
        default: return nullptr;
    }
}

}  // namespace wxk
