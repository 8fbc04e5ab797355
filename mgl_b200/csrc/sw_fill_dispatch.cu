/* sw_fill_dispatch.cu -- (kind, G, K) -> the matching explicit instantiation. */
#include "sw_fill_kernel.cuh"
#include "sw_fill_pipe.cuh"
#include "sw_kernels.h"

namespace mglsw {

#define MGLSW_DECL(T, G, K) extern template cudaError_t launch_fill<T, G, K>(const FillArgs &, u32, cudaStream_t);
MGLSW_FOR_K_1_24(MGLSW_DECL, S16x2, 8)
MGLSW_FOR_K_1_24(MGLSW_DECL, S32, 8)
MGLSW_FOR_K_7_24(MGLSW_DECL, S32, 32)
#undef MGLSW_DECL
#define MGLSW_DECL_PIPE(T, G, K) extern template cudaError_t launch_fill_pipe<T, K>(const FillArgs &, u32, bool, cudaStream_t);
MGLSW_FOR_K_7_24(MGLSW_DECL_PIPE, S32, 32)
#undef MGLSW_DECL_PIPE

#define MGLSW_CASE(T, GG, KK) if (G == GG && K == KK) return launch_fill<T, GG, KK>(a, first_unit, stream);

#define MGLSW_CASE_PIPE(T, GG, KK) if (K == KK) return launch_fill_pipe<T, KK>(a, first_unit, pipe > 1, stream);

cudaError_t launch_fill_any(int kind, int G, int K, int pipe, const FillArgs &a, u32 first_unit, cudaStream_t stream)
{
    if (pipe) {
        if (kind == KIND_S32 && G == MGLSW_G_LONG) { MGLSW_FOR_K_7_24(MGLSW_CASE_PIPE, S32, 32) }
        return cudaErrorInvalidValue;
    }
    if (kind == KIND_S16X2) { MGLSW_FOR_K_1_24(MGLSW_CASE, S16x2, 8) }
    else if (kind == KIND_S32) {
        MGLSW_FOR_K_1_24(MGLSW_CASE, S32, 8)
        MGLSW_FOR_K_7_24(MGLSW_CASE, S32, 32)
    }
    return cudaErrorInvalidValue;
}

}  // namespace mglsw
