/*
 * sw_trace_kernel.cu -- endpoint scan + checkpoint replay + traceback (one thread per pair), and
 * the two small kernels that turn per-pair CIGAR scratch into one dense array.
 * Replaces calculateCigar / calculateCigar_avx (sw.cpp:149-255, sw_avx.cpp:324-430).
 */
#include "sw_kernels.h"
#include "sw_trace.h"
#include "sw_config.h"

namespace mglsw {

/*
 * Thread 2 v + h handles half h of unit v: the threads of a warp walk pairs of one shape bucket
 * whose checkpoints sit next to each other, instead of 32 unrelated pairs in input order.
 */
#ifndef MGLSW_TRACE_MINB
#define MGLSW_TRACE_MINB 4
#endif
__global__ void __launch_bounds__(128, MGLSW_TRACE_MINB) sw_traceback_kernel(const __grid_constant__ TraceArgs a)
{
    __shared__ u64 s_nib[MGLSW_TRACE_BUF * 128];   /* the replayed block: records and boundary H values, one column per thread */
    const u32 tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= 2u * a.n_units) return;
    const UnitDesc &u = a.units[tid >> 1];
    const u32 half = tid & 1u;
    const u32 p = u.pair[half];
    if (!(u.flags & UNIT_ACTIVE) || p == NO_PAIR) return;
    const u32 G = (u.flags & UNIT_G32) ? MGLSW_G_LONG : MGLSW_G_SHORT;
    u32 *end = a.scratch + a.ops_end[p];
    const u32 *cb = a.colbest + (size_t)(tid >> 1) * 4u;
    PairResult r;
    if (u.flags & UNIT_S32) r = traceback_pair<S32>(u, half, G, u.regs, a.ckpt, a.arena, a.c32, cb, a.lastrow, a.strategy[p], end, s_nib + threadIdx.x, 128u);
    else                    r = traceback_pair<S16x2>(u, half, G, u.regs, a.ckpt, a.arena, a.c16, cb, a.lastrow, a.strategy[p], end, s_nib + threadIdx.x, 128u);
    a.offsets[p] = r.offset;
    a.n_ops[p] = r.n_ops;
}

/*
 * Score-only runs: no traceback, only the cell it would start from (sw.cpp:155-170: the corner, the
 * last-column maximum, or the better of last column and last row, by strategy) and the score there.
 */
__global__ void __launch_bounds__(128) sw_endpoint_kernel(const __grid_constant__ TraceArgs a)
{
    const u32 tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= 2u * a.n_units) return;
    const UnitDesc &u = a.units[tid >> 1];
    const u32 half = tid & 1u;
    const u32 p = u.pair[half];
    if (!(u.flags & UNIT_ACTIVE) || p == NO_PAIR) return;
    const u32 G = (u.flags & UNIT_G32) ? MGLSW_G_LONG : MGLSW_G_SHORT;
    const u32 *cb = a.colbest + (size_t)(tid >> 1) * 4u;
    StartCell st;
    int h;
    if (u.flags & UNIT_S32) { st = select_start<S32>(u, half, G, u.regs, cb, a.lastrow, a.strategy[p]); h = st.score - a.c32.bias - a.c32.s_match; }
    else                    { st = select_start<S16x2>(u, half, G, u.regs, cb, a.lastrow, a.strategy[p]); h = st.score - a.c16.bias - a.c16.s_match; }
    a.score3[3 * (size_t)p] = h;
    a.score3[3 * (size_t)p + 1] = st.i;
    a.score3[3 * (size_t)p + 2] = st.j;
}

/*
 * The exact-substring shortcut, opt-in.  The reference skips it on purpose -- "we skip substring
 * check since it performed by SWNativeAlignerWrapper" (...MicrosoftSmithWaterman.cpp:60) -- because GATK's wrapper
 * around the native binding (external to the reference: org.broadinstitute.hellbender...SWNativeAlignerWrapper.align)
 * answers SOFTCLIP and IGNORE calls itself when the alt occurs verbatim in the ref: offset = the LAST index at which
 * it occurs, CIGAR = <alt length>M.  A batched entry that bypasses that wrapper has no such pre-check.  (On every case
 * tried the DP's own tie rules -- the last-column maximum with ">=", sw.cpp:100-104 -- pick the last occurrence as
 * well, so the two agree; the option exists so that the wrapper's rule, not an observation, decides.)  With the option on,
 * this kernel overrides the traceback's result for exactly those pairs: one warp per pair, 32 candidate starts at a
 * time from the back of the ref, so the first group with a hit holds the last occurrence.
 */
__global__ void __launch_bounds__(128) sw_exact_match_kernel(const __grid_constant__ TraceArgs a)
{
    const u32 wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (wid >= 2u * a.n_units) return;
    const UnitDesc &u = a.units[wid >> 1];
    const u32 half = wid & 1u;
    const u32 p = u.pair[half];
    if (!(u.flags & UNIT_ACTIVE) || p == NO_PAIR || (half == 1u && (u.flags & UNIT_S32))) return;
    const int strategy = a.strategy[p];
    if (strategy != OS_SOFTCLIP && strategy != OS_IGNORE) return;
    const int tl = (int)u.tl[half], ql = (int)u.ql;
    if (ql > tl) return;
    const uint8_t *ref = a.arena + u.ref_off[half], *alt = a.arena + u.alt_off[half];
    for (int base = tl - ql; base >= 0; base -= 32) {
        const int i = base - (int)lane;
        bool hit = i >= 0;
        if (hit) {
            const uint8_t *r = ref + i;
            for (int k = 0; k < ql; k++) if (__ldg(r + k) != __ldg(alt + k)) { hit = false; break; }
        }
        const u32 m = __ballot_sync(0xffffffffu, hit);
        if (m) {
            if (lane == 0) {
                const int idx = base - (__ffs(m) - 1);                   /* the lowest lane with a hit = the largest start */
                a.offsets[p] = idx;
                a.n_ops[p] = 1u;
                a.scratch[a.ops_end[p] - 1u] = ((u32)ql << 4) | OP_M;
            }
            return;
        }
    }
}

cudaError_t launch_exact_match(const TraceArgs &a, cudaStream_t stream)
{
    if (a.n_units == 0) return cudaSuccess;
    sw_exact_match_kernel<<<(2u * a.n_units * 32u + 127) / 128, 128, 0, stream>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_endpoints(const TraceArgs &a, cudaStream_t stream)
{
    if (a.n_units == 0) return cudaSuccess;
    sw_endpoint_kernel<<<(2u * a.n_units + 127) / 128, 128, 0, stream>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_traceback(const TraceArgs &a, cudaStream_t stream)
{
    if (a.n_units == 0) return cudaSuccess;
    sw_traceback_kernel<<<(2u * a.n_units + 127) / 128, 128, 0, stream>>>(a);
    return cudaGetLastError();
}

/*
 * Exclusive scan of the per-pair element counts, two levels.  Level 1: each block scans 1024
 * counts (coalesced 16-byte loads, warp shuffles) into block-local offsets and one block total.
 * Level 2: one block scans the block totals in place and writes the grand total to starts[n].
 * The compaction adds the two.
 */
#define SCAN_TILE 1024u
__global__ void __launch_bounds__(256) sw_scan_tiles_kernel(const u32 *__restrict__ n_ops, u32 *__restrict__ starts,
                                                            u32 *__restrict__ tile_sum, u32 n)
{
    __shared__ u32 wsum[8];
    const u32 base = blockIdx.x * SCAN_TILE + threadIdx.x * 4u;
    u32 v[4];
#pragma unroll
    for (int k = 0; k < 4; k++) v[k] = base + k < n ? n_ops[base + k] : 0u;
    const u32 mine = v[0] + v[1] + v[2] + v[3];
    u32 incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const u32 x = __shfl_up_sync(0xffffffffu, incl, d); if ((threadIdx.x & 31u) >= (u32)d) incl += x; }
    if ((threadIdx.x & 31u) == 31u) wsum[threadIdx.x >> 5] = incl;
    __syncthreads();
    u32 woff = 0, total = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) { if ((u32)w < (threadIdx.x >> 5)) woff += wsum[w]; total += wsum[w]; }
    u32 run = woff + incl - mine;
#pragma unroll
    for (int k = 0; k < 4; k++) { if (base + k < n) starts[base + k] = run; run += v[k]; }
    if (threadIdx.x == 0) tile_sum[blockIdx.x] = total;
}

__global__ void __launch_bounds__(1024) sw_scan_totals_kernel(u32 *__restrict__ tile_sum, u32 *__restrict__ starts, u32 n_tiles, u32 n)
{
    __shared__ u32 part[1024];
    const u32 tid = threadIdx.x;
    const u32 per = (n_tiles + 1023u) / 1024u;
    const u32 lo = tid * per < n_tiles ? tid * per : n_tiles;
    const u32 hi = lo + per < n_tiles ? lo + per : n_tiles;
    u32 sum = 0;
    for (u32 k = lo; k < hi; k++) sum += tile_sum[k];
    part[tid] = sum;
    __syncthreads();
    for (u32 d = 1; d < 1024; d <<= 1) {
        const u32 v = tid >= d ? part[tid - d] : 0u;
        __syncthreads();
        part[tid] += v;
        __syncthreads();
    }
    u32 run = part[tid] - sum;
    for (u32 k = lo; k < hi; k++) { const u32 c = tile_sum[k]; tile_sum[k] = run; run += c; }
    if (tid == 1023) starts[n] = part[1023];
}

cudaError_t launch_scan_counts(const u32 *n_ops, u32 *starts, u32 *tile_sum, u32 n, cudaStream_t stream)
{
    const u32 n_tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    if (n_tiles) sw_scan_tiles_kernel<<<n_tiles, 256, 0, stream>>>(n_ops, starts, tile_sum, n);
    sw_scan_totals_kernel<<<1, 1024, 0, stream>>>(tile_sum, starts, n_tiles, n);
    return cudaGetLastError();
}

__global__ void __launch_bounds__(256) sw_compact_kernel(const u32 *__restrict__ scratch, const u64 *__restrict__ ops_end,
                                                         const u32 *__restrict__ n_ops, const u32 *__restrict__ starts,
                                                         const u32 *__restrict__ tile_off, u32 *__restrict__ ops_out, u32 n)
{
    const u32 p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const u32 cnt = n_ops[p];
    const u32 *src = scratch + (ops_end[p] - cnt);
    u32 *dst = ops_out + tile_off[p / SCAN_TILE] + starts[p];
    for (u32 k = 0; k < cnt; k++) dst[k] = src[k];
}

cudaError_t launch_compact(const u32 *scratch, const u64 *ops_end, const u32 *n_ops, const u32 *starts, const u32 *tile_off,
                           u32 *ops_out, u32 n, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    sw_compact_kernel<<<(n + 255) / 256, 256, 0, stream>>>(scratch, ops_end, n_ops, starts, tile_off, ops_out, n);
    return cudaGetLastError();
}

/*
 * Text CIGARs on the device: what the reference builds with to_string(length) + state per
 * element, elements of length <= 0 skipped (sw.cpp:251-252; the traceback never emits those).  Two passes over a
 * pair's packed elements: count the characters, then (after the same scan that places packed elements) write them.
 */
__device__ __forceinline__ u32 dec_digits(u32 v)
{
    u32 d = 1;
    while (v >= 10u) { v /= 10u; d++; }
    return d;
}

__global__ void __launch_bounds__(256) sw_text_len_kernel(const u32 *__restrict__ scratch, const u64 *__restrict__ ops_end,
                                                          const u32 *__restrict__ n_ops, u32 *__restrict__ n_chars, u32 n)
{
    const u32 p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const u32 cnt = n_ops[p];
    const u32 *src = scratch + (ops_end[p] - cnt);
    u32 chars = 0;
    for (u32 k = 0; k < cnt; k++) chars += dec_digits(src[k] >> 4) + 1u;
    n_chars[p] = chars;
}

__global__ void __launch_bounds__(256) sw_text_format_kernel(const u32 *__restrict__ scratch, const u64 *__restrict__ ops_end,
                                                             const u32 *__restrict__ n_ops, const u32 *__restrict__ starts,
                                                             const u32 *__restrict__ tile_off, char *__restrict__ text_out, u32 n)
{
    const u32 p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const u32 cnt = n_ops[p];
    const u32 *src = scratch + (ops_end[p] - cnt);
    char *dst = text_out + tile_off[p / SCAN_TILE] + starts[p];
    for (u32 k = 0; k < cnt; k++) {
        const u32 e = src[k];
        u32 len = e >> 4;
        const u32 d = dec_digits(len);
        for (u32 q = d; q-- > 0;) { dst[q] = (char)('0' + len % 10u); len /= 10u; }
        dst[d] = "MIDNSHP=X"[e & 15u];                        /* BAM numbering, as mgl_sw_format_cigar */
        dst += d + 1u;
    }
}

cudaError_t launch_text_len(const u32 *scratch, const u64 *ops_end, const u32 *n_ops, u32 *n_chars, u32 n, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    sw_text_len_kernel<<<(n + 255) / 256, 256, 0, stream>>>(scratch, ops_end, n_ops, n_chars, n);
    return cudaGetLastError();
}

cudaError_t launch_text_format(const u32 *scratch, const u64 *ops_end, const u32 *n_ops, const u32 *starts, const u32 *tile_off,
                               char *text_out, u32 n, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    sw_text_format_kernel<<<(n + 255) / 256, 256, 0, stream>>>(scratch, ops_end, n_ops, starts, tile_off, text_out, n);
    return cudaGetLastError();
}

}  // namespace mglsw
