/*
 * sw_fill_pipe.cuh -- the fill of LONG pairs (>= MGLSW_PIPE_MIN_STRIPS column strips): one pair per
 * block, its strips pipelined over the block's four warps.
 *
 * sw_fill_kernel gives a pair to one warp, which sweeps the strips one after the other: a 10 kbp x
 * 10 kbp pair is 14 strips x 10 000 steps on ONE warp (45 ms), and a chunk of a few hundred such pairs
 * (their checkpoints fill the HBM budget) leaves most warp slots of the GPU empty.  Here warp w takes
 * strips w, w + 4, w + 8 ...: strip s + 1 runs ~36 steps behind strip s, consuming the carry column
 * (H, F of strip s's last column, one entry per row, in HBM/L2 as before) as it is produced.  The same
 * FillLane code runs each strip; only the schedule differs, so results, checkpoints and the emulator
 * are unchanged.  It is the GPU analogue of the reference's band loop (sw_avx.cpp:71-80) run as a
 * software pipeline.
 *
 * Synchronisation is inside one block (all four warps are resident together): per warp a 64-bit
 * progress word in shared memory, (strip + 1) << 32 | rows whose carry entries are written; the
 * consumer of strip s spins on the producer's word before each group of four steps.  The dependency
 * chain strip s -> s - 1 -> ... -> 0 is acyclic and every warp walks its strips in increasing order,
 * so some warp can always run.  A spin that exceeds ~1 s traps (an error, never a hang).
 */
#ifndef MGL_B200_SW_FILL_PIPE_CUH
#define MGL_B200_SW_FILL_PIPE_CUH
#include <cuda_runtime.h>
#include "sw_config.h"
#include "sw_fill.h"

namespace mglsw {

template <class T, int K>
__global__ void __launch_bounds__(MGLSW_FILL_THREADS, 3)
sw_fill_pipe_kernel(const __grid_constant__ FillArgs a, const u32 first_unit)
{
    constexpr int G = 32, W = MGLSW_FILL_THREADS / 32;
    __shared__ unsigned long long s_prog[W];
    __shared__ uint4 s_stage[W][FillLane<T, G, K>::LS16 * 32];
    if (threadIdx.x < W) s_prog[threadIdx.x] = 0ull;
    __syncthreads();
    if (blockIdx.x >= a.n_warps) return;                 /* n_warps = units of this launch (one per block) */
    const u32 w = threadIdx.x >> 5, wl = threadIdx.x & 31u;
    const u32 ui = first_unit + blockIdx.x;
    const UnitDesc &u = a.units[ui];

    FillLane<T, G, K> L;
    L.setup(a, u, wl, ui);
    L.stage = reinterpret_cast<u32 *>(s_stage[w]);
    const u32 steps = u.rb_count * MGLSW_RB;
    const u32 nstrips = u.nstrips;
    const u32 rows = u.tl[0];
    volatile unsigned long long *prog = s_prog;
    for (u32 s = w; s < nstrips; s += W) {
        L.strip_begin(a.c, s);
        const u32 pw = (s + W - 1) % W;                  /* the warp that runs strip s - 1 */
        for (u32 g = 0; g < steps / 4; g++) {
            const u32 t0 = g * 4u;
            if (s > 0) {
                /* lane 0 uses the carry entries of rows t0 .. t0 + 3 in this group and fetches row t0 + 4 ahead */
                const u32 need_rows = t0 + 4u < rows ? t0 + 4u : rows;
                const unsigned long long need = ((unsigned long long)s << 32) | need_rows;      /* strip s - 1 is word value s */
                u32 spins = 0;
                while (prog[pw] < need) {
                    __nanosleep(100);
                    if (++spins > (1u << 23)) __trap();
                }
                __threadfence_block();
            }
            if (!__all_sync(0xffffffffu, L.quiet4(t0))) {
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
            if (s + 1 < nstrips) {
                /* lane G - 1 has finished row t - (G - 1) at step t: rows <= t0 + 3 - 31 have their carry entry */
                const u32 t_end = t0 + 3u;
                u32 done = t_end >= (u32)(G - 1) ? t_end - (u32)(G - 1) : 0u;
                if (done > rows) done = rows;
                __threadfence_block();                   /* the carry stores of lane G - 1 before the progress word */
                __syncwarp();
                if (wl == 0) prog[w] = ((unsigned long long)(s + 1) << 32) | done;
            }
        }
        L.strip_end();
        __syncwarp();
    }
    /* the last region's bulk store still reads this block's shared memory */
    if (wl == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    __syncwarp();
}

template <class T, int K>
cudaError_t launch_fill_pipe(const FillArgs &a, u32 first_unit, cudaStream_t stream)
{
    if (a.n_warps == 0) return cudaSuccess;
    sw_fill_pipe_kernel<T, K><<<a.n_warps, MGLSW_FILL_THREADS, 0, stream>>>(a, first_unit);
    return cudaGetLastError();
}

}  // namespace mglsw
#endif
