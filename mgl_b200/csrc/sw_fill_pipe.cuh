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

#include <atomic>

#include "sw_config.h"
#include "sw_fill.h"

namespace mglsw {

/*
 * W warps share one pair.  W = 4 (three blocks per SM) is the throughput shape: a launch with many
 * long pairs fills the GPU with blocks anyway.  W = 8 is the latency shape: a launch with only a few
 * long pairs (a mixed batch has ~60 per chunk) would otherwise decide how long its chunk holds a
 * work slot; eight warps finish a 13-strip pair in two rounds instead of four.
 */
template <class T, int K, int W>
__device__ __forceinline__ void fill_pipe_body(const FillArgs &a, const u32 first_unit)
{
    constexpr int G = 32;
    /* dynamic shared memory (W = 8 needs more than the 48 KB a static array may have): progress words, then one staging region per warp */
    extern __shared__ __align__(128) unsigned char s_dyn[];
    unsigned long long *s_prog = reinterpret_cast<unsigned long long *>(s_dyn);
    uint4 (*s_stage)[FillLane<T, G, K>::LS16 * 32] = reinterpret_cast<uint4 (*)[FillLane<T, G, K>::LS16 * 32]>(s_dyn + 128);
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
__global__ void __launch_bounds__(128, 3)
sw_fill_pipe_kernel(const __grid_constant__ FillArgs a, const u32 first_unit)
{
    fill_pipe_body<T, K, 4>(a, first_unit);
}

/* 168 registers as in the other fill kernels: one of these blocks and one four-warp block share an SM */
template <class T, int K>
__global__ void __maxnreg__(168)
sw_fill_pipe_wide_kernel(const __grid_constant__ FillArgs a, const u32 first_unit)
{
    fill_pipe_body<T, K, MGLSW_PIPE_WIDE_WARPS>(a, first_unit);
}

template <class T, int K>
cudaError_t launch_fill_pipe(const FillArgs &a, u32 first_unit, bool wide, cudaStream_t stream)
{
    if (a.n_warps == 0) return cudaSuccess;
    constexpr size_t per_warp = (size_t)FillLane<T, 32, K>::LS16 * 512u;
    const size_t smem = 128u + per_warp * (wide ? MGLSW_PIPE_WIDE_WARPS : 4);
    /* function attributes are per device: a handle over several GPUs launches this from one process on each of them */
    static std::atomic<unsigned long long> done{0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (wide && !(done.load(std::memory_order_relaxed) >> (dev & 63) & 1ull)) {
        const cudaError_t e = cudaFuncSetAttribute(sw_fill_pipe_wide_kernel<T, K>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                   (int)(128u + per_warp * MGLSW_PIPE_WIDE_WARPS));
        if (e != cudaSuccess) return e;
        done.fetch_or(1ull << (dev & 63), std::memory_order_relaxed);
    }
    if (wide) sw_fill_pipe_wide_kernel<T, K><<<a.n_warps, 32 * MGLSW_PIPE_WIDE_WARPS, smem, stream>>>(a, first_unit);
    else      sw_fill_pipe_kernel<T, K><<<a.n_warps, 128, smem, stream>>>(a, first_unit);
    return cudaGetLastError();
}

}  // namespace mglsw
#endif
