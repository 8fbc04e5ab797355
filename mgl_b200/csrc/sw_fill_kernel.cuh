/*
 * sw_fill_kernel.cuh -- the CUDA fill kernel: one warp drives 32 FillLanes
 * (sw_fill.h), 32 / G lane groups per warp, neighbours exchange (H, F) with
 * __shfl_up_sync.  sm_100a only.
 */
#ifndef MGL_B200_SW_FILL_KERNEL_CUH
#define MGL_B200_SW_FILL_KERNEL_CUH
#include <cuda_runtime.h>
#include "sw_config.h"
#include "sw_fill.h"

namespace mglsw {

template <class T, int G, int K>
#ifndef MGLSW_FILL_MINB
#define MGLSW_FILL_MINB 3
#endif
__global__ void __launch_bounds__(MGLSW_FILL_THREADS, MGLSW_FILL_MINB)
sw_fill_kernel(const __grid_constant__ FillArgs a, const u32 first_unit)
{
    const u32 warp = blockIdx.x * (MGLSW_FILL_THREADS / 32) + (threadIdx.x >> 5);
    if (warp >= a.n_warps) return;
    const u32 wl = threadIdx.x & 31u;
    const u32 ui = first_unit + warp * (32 / G) + wl / G;
    const UnitDesc &u = a.units[ui];

    /* one checkpoint region per warp, assembled here and copied out every MGLSW_RB steps */
    __shared__ uint4 s_stage[MGLSW_FILL_THREADS / 32][FillLane<T, G, K>::LS16 * 32];

    FillLane<T, G, K> L;
    L.setup(a, u, wl, ui);
    L.stage = reinterpret_cast<u32 *>(s_stage[threadIdx.x >> 5]);
    const u32 steps = u.rb_count * MGLSW_RB;     /* same for every lane of the warp */
    const u32 nstrips = u.nstrips;
    for (u32 s = 0; s < nstrips; s++) {
        L.strip_begin(a.c, s);
        for (u32 g = 0; g < steps / 4; g++) {           /* steps is a multiple of MGLSW_RB = 8 */
            const u32 t0 = g * 4u;
            if (!__all_sync(0xffffffffu, L.quiet4(t0))) {
                /* some lane has not reached its row 1 yet, is past its last row, or captures a high half in these four steps */
#pragma unroll 1
                for (u32 s4 = 0; s4 < 4; s4++) {
                    const u32 h = __shfl_up_sync(0xffffffffu, L.h_out, 1, G);
                    const u32 f = __shfl_up_sync(0xffffffffu, L.f_out, 1, G);
                    L.template step<true>(a.c, t0 + s4, h, f);
                }
            } else {
#pragma unroll
                for (u32 s4 = 0; s4 < 4; s4++) {
                    const u32 h = __shfl_up_sync(0xffffffffu, L.h_out, 1, G);
                    const u32 f = __shfl_up_sync(0xffffffffu, L.f_out, 1, G);
                    L.template step<false>(a.c, t0 + s4, h, f);
                }
            }
        }
        L.strip_end();
        __syncwarp();                            /* carry column of strip s is read by lane 0 in strip s + 1 */
    }
    /* the last region's bulk store still reads this block's shared memory */
    if (wl == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    __syncwarp();
}

template <class T, int G, int K>
cudaError_t launch_fill(const FillArgs &a, u32 first_unit, cudaStream_t stream)
{
    const u32 wpb = MGLSW_FILL_THREADS / 32;
    const u32 blocks = (a.n_warps + wpb - 1) / wpb;
    if (blocks == 0) return cudaSuccess;
    sw_fill_kernel<T, G, K><<<blocks, MGLSW_FILL_THREADS, 0, stream>>>(a, first_unit);
    return cudaGetLastError();
}

}  // namespace mglsw
#endif
