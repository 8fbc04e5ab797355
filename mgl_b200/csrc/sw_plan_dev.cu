/*
 * sw_plan_dev.cu -- the chunk planner on the device: every stage of sw_plan_dev.h inside ONE kernel of 16 small
 * thread blocks per chunk that meet at grid barriers, with a grid-wide LSD radix sort of the 30-bit bucket keys and
 * two grid-wide scans between the stages.  Hand-written, no library calls.
 *
 * Why this shape.  The planner runs while the fill kernels of earlier chunks own every SM: their three blocks per SM
 * take the whole register file, so a planner block starts only when a fill block retires, and then shares its SM's
 * issue slots with two fill blocks.  Measured on the way here (65 536-pair plans under load, profiles/r05*):
 *   - each stage its own grid-wide kernel, library sort and scans between them (~30 dependent launches): every launch
 *     paid the launch latency plus the wait for SM space, 1.1-1.2 ms per plan against 0.2 ms on an idle GPU -- as
 *     long as the chunk's own fill, which made the planner the bottleneck of the end-to-end pipeline;
 *   - one block doing everything between __syncthreads(): one retirement, once, but ~800 instructions per pair on
 *     one shared SM: 2.7-4 ms per plan.
 * Sixteen blocks of 256 threads x <= 64 registers and 9 KiB of shared memory each fit beside two resident fill blocks
 * on sixteen SMs, need sixteen retirements once, and split the work sixteen ways; the ~30 barriers of a plan cost
 * microseconds each.  Every loop over pairs is coalesced (the sort ranks with MATCH.ANY inside 32-item groups).
 */
#include "sw_plan_dev.h"
#include "sw_plan_dev_api.h"

namespace mglsw {

namespace {

constexpr int PT = 256;                  /* threads per planner block */
constexpr int NW = PT / 32;              /* warps per block */
constexpr int GB = 16;                   /* planner blocks, all resident at once (they meet at grid barriers) */
constexpr int GW = GB * NW;              /* warps of the grid */
constexpr int RADIX_BITS = 8;            /* four passes over the 30-bit key */
constexpr int RADIX = 1 << RADIX_BITS;
constexpr int SEG = RADIX * GW / GB;     /* counters each block scans per pass */

#define PD_FOR(i, n) for (u32 i = blockIdx.x * PT + threadIdx.x; i < (n); i += GB * PT)

/*
 * Barrier over the GB blocks of the grid: the pattern of cooperative groups' grid.sync() -- block barrier, one
 * thread publishes with a fence and an atomic arrival, spins until everyone of this round has arrived, fences again,
 * block barrier.  The blocks need not have been launched cooperatively: they depend on nobody but each other, and
 * a block that is not resident yet becomes so as soon as any fill block retires.  *bar is zeroed by the host before
 * the launch; `round` counts this block's barriers.
 */
__device__ __forceinline__ void grid_sync(unsigned *bar, unsigned &round)
{
    __syncthreads();
    round++;
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(bar, 1u);
        while (*(volatile unsigned *)bar < round * GB) __nanosleep(40);
        __threadfence();
    }
    __syncthreads();
}

/* exclusive scan of one value per thread over the block; *total = the sum.  ws: NW shared words of type V */
template <class V>
__device__ __forceinline__ V block_excl_scan(V v, V *ws, V *total)
{
    const u32 lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    V incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const V x = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= (u32)d) incl += x; }
    if (lane == 31) ws[warp] = incl;
    __syncthreads();
    V woff = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < NW; w++) { const V s = ws[w]; if ((u32)w < warp) woff += s; tot += s; }
    __syncthreads();                     /* ws is reused by the next call */
    *total = tot;
    return woff + incl - v;
}

/* the slice of `groups` 32-item groups that this warp of the grid walks (contiguous, in order) */
__device__ __forceinline__ void warp_slice(u32 groups, u32 &g0, u32 &g1)
{
    const u32 gw = blockIdx.x * NW + (threadIdx.x >> 5), per = (groups + GW - 1) / GW;
    g0 = gw * per < groups ? gw * per : groups;
    g1 = g0 + per < groups ? g0 + per : groups;
}

/* device buffers of the sort and the scans that blocks share */
struct Shared {
    u32 *cnt;                            /* RADIX * GW digit counters, index digit * GW + warp of the grid */
    u32 *btot;                           /* GB: each block's total of its segment of cnt */
    u32 *wtot;                           /* GW: per-warp totals of a u32 scan */
    unsigned long long *wtot3;           /* GW * 3: per-warp totals of the three 64-bit scans */
    unsigned *bar;
};

/*
 * One stable LSD pass over the whole grid.  Every WARP owns a contiguous slice of the input and walks it 32 items at
 * a time with coalesced loads; inside a group, lanes with the same digit find each other with MATCH.ANY: the lowest
 * such lane bumps the warp's counter of that digit by the size of the peer group, a lane's rank among its peers is
 * the number of peers below it.  Counters are per (digit, warp of the grid) and are scanned in that order -- each
 * block scans one segment, block totals are added after a barrier -- so equal digits keep their input order: warp
 * slices are in input order, groups inside a slice too, ranks inside a group too.
 */
__device__ void radix_pass(const u32 *kin, const u32 *vin, u32 *kout, u32 *vout, u32 n, int shift,
                           u32 *hist /* NW * RADIX */, u32 *ws, const Shared &sh, unsigned &round)
{
    const u32 t = threadIdx.x, lane = t & 31u, warp = t >> 5, gw = blockIdx.x * NW + warp;
    u32 *mine = hist + warp * RADIX;
    u32 g0, g1;
    warp_slice((n + 31u) / 32u, g0, g1);
#pragma unroll
    for (int j = 0; j < RADIX / 32; j++) mine[lane + 32 * j] = 0u;
    __syncwarp();
    for (u32 g = g0; g < g1; g++) {
        const u32 i = g * 32u + lane;
        const bool valid = i < n;
        const u32 mask = __ballot_sync(0xffffffffu, valid);
        if (valid) {
            const u32 d = (kin[i] >> shift) & (RADIX - 1);
            const u32 peers = __match_any_sync(mask, d);
            if (lane == (u32)__ffs(peers) - 1u) mine[d] += (u32)__popc(peers);
        }
        __syncwarp();
    }
#pragma unroll
    for (int j = 0; j < RADIX / 32; j++) { const u32 d = lane + 32 * j; sh.cnt[d * GW + gw] = mine[d]; }
    grid_sync(sh.bar, round);
    /* this block's segment of the counters: exclusive scan in place, total to btot */
    {
        constexpr int PER = SEG / PT;
        u32 *seg = sh.cnt + blockIdx.x * SEG + t * PER;
        u32 c[PER], sum = 0;
#pragma unroll
        for (int j = 0; j < PER; j++) { c[j] = seg[j]; sum += c[j]; }
        u32 total;
        u32 run = block_excl_scan<u32>(sum, ws, &total);
#pragma unroll
        for (int j = 0; j < PER; j++) { seg[j] = run; run += c[j]; }
        if (t == 0) sh.btot[blockIdx.x] = total;
    }
    grid_sync(sh.bar, round);
    /* this warp's starting places: its counters plus the totals of the blocks whose segments come before */
#pragma unroll
    for (int j = 0; j < RADIX / 32; j++) {
        const u32 d = lane + 32 * j, k = d * GW + gw, owner = k / SEG;
        u32 pre = 0;
        for (u32 bq = 0; bq < owner; bq++) pre += sh.btot[bq];
        mine[d] = sh.cnt[k] + pre;
    }
    __syncwarp();
    for (u32 g = g0; g < g1; g++) {
        const u32 i = g * 32u + lane;
        const bool valid = i < n;
        const u32 mask = __ballot_sync(0xffffffffu, valid);
        if (valid) {
            const u32 key = kin[i], val = vin[i];
            const u32 d = (key >> shift) & (RADIX - 1);
            const u32 peers = __match_any_sync(mask, d);
            const u32 base = mine[d];
            const u32 at = base + (u32)__popc(peers & ((1u << lane) - 1u));
            kout[at] = key;
            vout[at] = val;
            __syncwarp(mask);                                /* every peer has read the counter */
            if (lane == (u32)__ffs(peers) - 1u) mine[d] = base + (u32)__popc(peers);
        }
        __syncwarp();
    }
    grid_sync(sh.bar, round);
}

/* exclusive sum of x[0, n) into out over the whole grid: warp slices of 32-item groups, coalesced */
__device__ void grid_excl_sum_u32(const u32 *x, u32 *out, u32 n, const Shared &sh, unsigned &round)
{
    const u32 lane = threadIdx.x & 31u, gw = blockIdx.x * NW + (threadIdx.x >> 5);
    u32 g0, g1;
    warp_slice((n + 31u) / 32u, g0, g1);
    u32 sum = 0;
    for (u32 g = g0; g < g1; g++) { const u32 i = g * 32u + lane; sum += i < n ? x[i] : 0u; }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, d);
    if (lane == 0) sh.wtot[gw] = sum;
    grid_sync(sh.bar, round);
    u32 run = 0;
    for (u32 w = lane; w < gw; w += 32) run += sh.wtot[w];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) run += __shfl_xor_sync(0xffffffffu, run, d);
    for (u32 g = g0; g < g1; g++) {
        const u32 i = g * 32u + lane;
        const u32 v = i < n ? x[i] : 0u;
        u32 incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const u32 y = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= (u32)d) incl += y; }
        if (i < n) out[i] = run + incl - v;
        run += __shfl_sync(0xffffffffu, incl, 31);
    }
    grid_sync(sh.bar, round);
}

/* the same for the three 64-bit reports of stage G */
__device__ void grid_excl_sum_u3(const U3 *x, U3 *out, u32 n, const Shared &sh, unsigned &round)
{
    typedef unsigned long long ull;
    const u32 lane = threadIdx.x & 31u, gw = blockIdx.x * NW + (threadIdx.x >> 5);
    u32 g0, g1;
    warp_slice((n + 31u) / 32u, g0, g1);
    ull sa = 0, sb = 0, sc = 0;
    for (u32 g = g0; g < g1; g++) { const u32 i = g * 32u + lane; if (i < n) { const U3 v = x[i]; sa += v.a; sb += v.b; sc += v.c; } }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) { sa += __shfl_xor_sync(0xffffffffu, sa, d); sb += __shfl_xor_sync(0xffffffffu, sb, d); sc += __shfl_xor_sync(0xffffffffu, sc, d); }
    if (lane == 0) { sh.wtot3[gw * 3] = sa; sh.wtot3[gw * 3 + 1] = sb; sh.wtot3[gw * 3 + 2] = sc; }
    grid_sync(sh.bar, round);
    ull ra = 0, rb = 0, rc = 0;
    for (u32 w = lane; w < gw; w += 32) { ra += sh.wtot3[w * 3]; rb += sh.wtot3[w * 3 + 1]; rc += sh.wtot3[w * 3 + 2]; }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) { ra += __shfl_xor_sync(0xffffffffu, ra, d); rb += __shfl_xor_sync(0xffffffffu, rb, d); rc += __shfl_xor_sync(0xffffffffu, rc, d); }
    for (u32 g = g0; g < g1; g++) {
        const u32 i = g * 32u + lane;
        U3 v; v.a = v.b = v.c = 0;
        if (i < n) v = x[i];
        ull ia = v.a, ib = v.b, ic = v.c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const ull ya = __shfl_up_sync(0xffffffffu, ia, d), yb = __shfl_up_sync(0xffffffffu, ib, d), yc = __shfl_up_sync(0xffffffffu, ic, d);
            if (lane >= (u32)d) { ia += ya; ib += yb; ic += yc; }
        }
        if (i < n) { U3 o; o.a = ra + ia - v.a; o.b = rb + ib - v.b; o.c = rc + ic - v.c; out[i] = o; }
        ra += __shfl_sync(0xffffffffu, ia, 31); rb += __shfl_sync(0xffffffffu, ib, 31); rc += __shfl_sync(0xffffffffu, ic, 31);
    }
    grid_sync(sh.bar, round);
}

__global__ void __launch_bounds__(PT, 4) sw_plan_kernel(const PlanDevIn in, const PlanDevLayout lay, char *work, UnitDesc *units)
{
    __shared__ u32 hist[RADIX * NW];                     /* the sort's counters of this block's warps */
    __shared__ u32 ws32[NW];
    __shared__ u32 s_n16, s_n32, s_need;
    __shared__ unsigned long long s_cells;
    u32 *key0 = (u32 *)(work + lay.key0), *key1 = (u32 *)(work + lay.key1), *idx0 = (u32 *)(work + lay.idx0), *idx1 = (u32 *)(work + lay.idx1);
    u32 *uhead = (u32 *)(work + lay.uhead), *urank = (u32 *)(work + lay.urank);
    U3 *u3 = (U3 *)(work + lay.u3), *u3x = (U3 *)(work + lay.u3x);
    PlanState *st = (PlanState *)(work + lay.state);
    Shared sh;
    sh.cnt = (u32 *)(work + lay.cnt); sh.btot = (u32 *)(work + lay.btot); sh.wtot = (u32 *)(work + lay.wtot);
    sh.wtot3 = (unsigned long long *)(work + lay.wtot3); sh.bar = (unsigned *)(work + lay.bar);
    const u32 n = in.n, t = threadIdx.x;
    unsigned round = 0;

    /* the state starts from zero */
    if (blockIdx.x == 0) for (u32 w = t; w < sizeof(PlanState) / 4; w += PT) reinterpret_cast<u32 *>(st)[w] = 0u;
    if (t == 0) { s_n16 = 0; s_n32 = 0; s_need = 0; s_cells = 0; }
    grid_sync(sh.bar, round);

    /* A: the chunk's 16-bit bias is the largest any candidate needs (one atomic per warp, one per block); the pair's
       bounds wait for stage B in the arrays the sort will overwrite later (key1 / idx1) */
    {
        u32 need = 0;
        PD_FOR(p, n) { const u32 v = pd_stage_a(in, p, st, key1, idx1); need = v > need ? v : need; }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) { const u32 x = __shfl_down_sync(0xffffffffu, need, d); need = x > need ? x : need; }
        if ((t & 31u) == 0) atomicMax(&s_need, need);
        __syncthreads();
        if (t == 0 && s_need) atomicMax(&st->need_bias, s_need);
    }
    grid_sync(sh.bar, round);

    /* B: keys, and the counts the summary carries */
    {
        u32 n16 = 0, n32 = 0;
        unsigned long long cells = 0;
        PD_FOR(p, n) {
            const int kind = pd_stage_b(in, p, st, key1, idx1, key0, idx0);
            n16 += kind == KIND_S16X2; n32 += kind != KIND_S16X2;
            cells += (unsigned long long)in.tl[p] * in.ql[p];
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            n16 += __shfl_down_sync(0xffffffffu, n16, d);
            n32 += __shfl_down_sync(0xffffffffu, n32, d);
            cells += __shfl_down_sync(0xffffffffu, cells, d);
        }
        if ((t & 31u) == 0) { atomicAdd(&s_n16, n16); atomicAdd(&s_n32, n32); atomicAdd(&s_cells, cells); }
        __syncthreads();
        if (t == 0) { atomicAdd(&st->pairs_s16, s_n16); atomicAdd(&st->pairs_s32, s_n32); atomicAdd(&st->cells, s_cells); }
    }
    grid_sync(sh.bar, round);

    /* sort (key, pair) ascending: four stable passes of 8 bits, ping-pong; the result is back in key0 / idx0 */
    static_assert(MGLSW_KEY_BITS <= 32, "four 8-bit passes");
    radix_pass(key0, idx0, key1, idx1, n, 0, hist, ws32, sh, round);
    radix_pass(key1, idx1, key0, idx0, n, 8, hist, ws32, sh, round);
    radix_pass(key0, idx0, key1, idx1, n, 16, hist, ws32, sh, round);
    radix_pass(key1, idx1, key0, idx0, n, 24, hist, ws32, sh, round);
    const u32 *skey = key0, *order = idx0;

    /* C */
    PD_FOR(q, n) pd_stage_c(q, skey, st, uhead);
    grid_sync(sh.bar, round);

    /* unit ranks */
    grid_excl_sum_u32(uhead, urank, n, sh, round);

    /* E */
    if (blockIdx.x == 0 && t == 0) pd_stage_e(n, skey, urank, uhead, st);
    grid_sync(sh.bar, round);

    /* F */
    PD_FOR(q, n) pd_stage_f(in, q, order, uhead, urank, st, units);
    grid_sync(sh.bar, round);

    /* G */
    const u32 nu = st->sum.n_units;
    PD_FOR(u, nu) pd_stage_g(u, st, units, u3);
    grid_sync(sh.bar, round);

    /* placements */
    grid_excl_sum_u3(u3, u3x, nu, sh, round);

    /* H */
    PD_FOR(u, nu) pd_stage_h(in, u, st, units, u3, u3x);
    grid_sync(sh.bar, round);

    /* I */
    if (blockIdx.x == 0 && t == 0) pd_stage_i(in, st);
}

size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

}  // namespace

u32 plan_dev_unit_capacity(u32 n) { return n + 3u * MGLSW_PLAN_MAX_LAUNCHES + 4u; }

PlanDevLayout plan_dev_layout(u32 n)
{
    if (n == 0) n = 1;
    PlanDevLayout w;
    w.n_cap = n;
    w.ucap = plan_dev_unit_capacity(n);
    size_t at = 0;
    auto take = [&](size_t bytes) { const size_t o = at; at += align256(bytes); return o; };
    w.key0 = take((size_t)n * 4); w.key1 = take((size_t)n * 4);
    w.idx0 = take((size_t)n * 4); w.idx1 = take((size_t)n * 4);
    w.uhead = take((size_t)n * 4); w.urank = take((size_t)n * 4);
    w.u3 = take((size_t)w.ucap * sizeof(U3)); w.u3x = take((size_t)w.ucap * sizeof(U3));
    w.state = take(sizeof(PlanState));
    w.cnt = take((size_t)RADIX * GW * 4); w.btot = take((size_t)GB * 4); w.wtot = take((size_t)GW * 4); w.wtot3 = take((size_t)GW * 24);
    w.bar = take(256);
    w.total = at;
    return w;
}

cudaError_t plan_dev_run(const PlanDevIn &in, const PlanDevLayout &lay, UnitDesc *units, void *work, PlanSummary *h_summary, cudaStream_t st)
{
    if (in.n == 0 || in.n > lay.n_cap) return cudaErrorInvalidValue;
    cudaError_t e = cudaMemsetAsync((char *)work + lay.bar, 0, 4, st);      /* the grid barrier's arrival count */
    if (e != cudaSuccess) return e;
    sw_plan_kernel<<<GB, PT, 0, st>>>(in, lay, (char *)work, units);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    const PlanState *state = (const PlanState *)((const char *)work + lay.state);
    return cudaMemcpyAsync(h_summary, &state->sum, sizeof(PlanSummary), cudaMemcpyDeviceToHost, st);
}

}  // namespace mglsw
