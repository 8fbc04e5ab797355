/*
 * sw_plan_dev.cu -- the chunk planner on the device: the stage bodies of sw_plan_dev.h as kernels, a radix sort of
 * the bucket keys and four scans between them (CUB device primitives from the CUDA toolkit: plumbing, like the
 * engine's cudaMemcpyAsync -- the hot kernels are the fill and the traceback).  Everything runs on one stream,
 * ahead of the chunk's fill; the host reads back PlanSummary (a few kilobytes) and never sees a unit descriptor.
 */
#include <algorithm>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include "sw_plan_dev.h"
#include "sw_plan_dev_api.h"

namespace mglsw {

namespace {

constexpr int PT = 256;       /* threads per block of every planner kernel */
/*
 * Small grids, grid-stride loops.  The planner runs while the fill kernels of earlier chunks own every SM (their
 * three blocks per SM take the whole register file), so a planner block starts only when a fill block retires;
 * measured with one thread per item (256 blocks per kernel): 2.6 ms for the ~30 dependent kernels of one plan,
 * each waiting for ~128 fill blocks to leave.  A few dozen blocks per kernel need a few dozen retirements.
 */
constexpr unsigned MAX_BLOCKS = 32;
#define PD_FOR(i, n) for (u32 i = blockIdx.x * PT + threadIdx.x; i < (n); i += gridDim.x * PT)

struct MaxU32 { __host__ __device__ __forceinline__ u32 operator()(u32 a, u32 b) const { return a > b ? a : b; } };
struct AddU3 { __host__ __device__ __forceinline__ U3 operator()(const U3 &a, const U3 &b) const { return a + b; } };

__global__ void __launch_bounds__(PT) pd_a_kernel(const PlanDevIn in, PlanState *st, u64 *oplen)
{
    PD_FOR(p, in.n) pd_stage_a(in, p, st, oplen);
}

__global__ void __launch_bounds__(PT) pd_b_kernel(const PlanDevIn in, PlanState *st, u64 *key, u32 *idx)
{
    u32 n16 = 0, n32 = 0;
    unsigned long long cells = 0;
    PD_FOR(p, in.n) {
        const int kind = pd_stage_b(in, p, st, key, idx);
        n16 += kind == KIND_S16X2; n32 += kind != KIND_S16X2;
        cells += (unsigned long long)in.tl[p] * in.ql[p];
    }
    /* counts and cells: one atomic per warp */
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        n16 += __shfl_down_sync(0xffffffffu, n16, d);
        n32 += __shfl_down_sync(0xffffffffu, n32, d);
        cells += __shfl_down_sync(0xffffffffu, cells, d);
    }
    if ((threadIdx.x & 31) == 0 && (n16 | n32)) {
        if (n16) atomicAdd(&st->pairs_s16, n16);
        if (n32) atomicAdd(&st->pairs_s32, n32);
        atomicAdd(&st->cells, cells);
    }
}

__global__ void __launch_bounds__(PT) pd_c_kernel(u32 n, const u64 *skey, PlanState *st, u32 *runmark)
{
    PD_FOR(q, n) pd_stage_c(n, q, skey, st, runmark);
}

__global__ void __launch_bounds__(PT) pd_d_kernel(u32 n, const u64 *skey, const u32 *run_start, u32 *uhead)
{
    PD_FOR(q, n) pd_stage_d(q, skey, run_start, uhead);
}

__global__ void pd_e_kernel(u32 n, const u64 *skey, const u32 *urank, const u32 *uhead, PlanState *st)
{
    if (blockIdx.x == 0 && threadIdx.x == 0) pd_stage_e(n, skey, urank, uhead, st);
}

__global__ void __launch_bounds__(PT) pd_f_kernel(const PlanDevIn in, const u32 *order, const u32 *run_start, const u32 *uhead,
                                                  const u32 *urank, const PlanState *st, UnitDesc *units)
{
    PD_FOR(q, in.n) pd_stage_f(in, q, order, run_start, uhead, urank, st, units);
}

/* launched over the unit capacity (the host does not know the unit count yet): units beyond the plan report zero */
__global__ void __launch_bounds__(PT) pd_g_kernel(u32 ucap, PlanState *st, UnitDesc *units, U3 *u3)
{
    const u32 nu = st->sum.n_units;
    PD_FOR(u, ucap) {
        if (u < nu) pd_stage_g(u, st, units, u3);
        else { U3 z; z.a = z.b = z.c = 0; u3[u] = z; }
    }
}

__global__ void __launch_bounds__(PT) pd_h_kernel(const PlanDevIn in, u32 ucap, PlanState *st, UnitDesc *units, const U3 *u3, const U3 *u3x)
{
    const u32 nu = st->sum.n_units < ucap ? st->sum.n_units : ucap;
    PD_FOR(u, nu) pd_stage_h(in, u, st, units, u3, u3x);
}

__global__ void pd_i_kernel(const PlanDevIn in, PlanState *st, const u64 *ops_end)
{
    if (blockIdx.x == 0 && threadIdx.x == 0) pd_stage_i(in, st, ops_end);
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
    w.key = take((size_t)n * 8); w.skey = take((size_t)n * 8);
    w.idx = take((size_t)n * 4); w.order = take((size_t)n * 4);
    w.runmark = take((size_t)n * 4); w.run_start = take((size_t)n * 4);
    w.uhead = take((size_t)n * 4); w.urank = take((size_t)n * 4);
    w.oplen = take((size_t)n * 8);
    w.u3 = take((size_t)w.ucap * sizeof(U3)); w.u3x = take((size_t)w.ucap * sizeof(U3));
    w.state = take(sizeof(PlanState));
    /* the largest temporary any of the five CUB calls asks for at this capacity */
    size_t need = 0, b = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, b, (const u64 *)nullptr, (u64 *)nullptr, (const u32 *)nullptr, (u32 *)nullptr, (int)n, 0, 64);
    need = b > need ? b : need;
    cub::DeviceScan::InclusiveScan(nullptr, b, (const u32 *)nullptr, (u32 *)nullptr, MaxU32(), (int)n);
    need = b > need ? b : need;
    cub::DeviceScan::ExclusiveSum(nullptr, b, (const u32 *)nullptr, (u32 *)nullptr, (int)n);
    need = b > need ? b : need;
    U3 zero; zero.a = zero.b = zero.c = 0;
    cub::DeviceScan::ExclusiveScan(nullptr, b, (const U3 *)nullptr, (U3 *)nullptr, AddU3(), zero, (int)w.ucap);
    need = b > need ? b : need;
    cub::DeviceScan::InclusiveSum(nullptr, b, (const u64 *)nullptr, (u64 *)nullptr, (int)n);
    need = b > need ? b : need;
    w.cub_bytes = need + 256;
    w.cub = take(w.cub_bytes);
    w.total = at;
    return w;
}

cudaError_t plan_dev_run(const PlanDevIn &in, const PlanDevLayout &w, UnitDesc *units, u64 *ops_end, void *work,
                         PlanSummary *h_summary, cudaStream_t st)
{
    const u32 n = in.n;
    if (n == 0 || n > w.n_cap) return cudaErrorInvalidValue;
    char *base = (char *)work;
    u64 *key = (u64 *)(base + w.key), *skey = (u64 *)(base + w.skey), *oplen = (u64 *)(base + w.oplen);
    u32 *idx = (u32 *)(base + w.idx), *order = (u32 *)(base + w.order), *runmark = (u32 *)(base + w.runmark);
    u32 *run_start = (u32 *)(base + w.run_start), *uhead = (u32 *)(base + w.uhead), *urank = (u32 *)(base + w.urank);
    U3 *u3 = (U3 *)(base + w.u3), *u3x = (U3 *)(base + w.u3x);
    PlanState *state = (PlanState *)(base + w.state);
    void *tmp = base + w.cub;
    size_t tb = w.cub_bytes;
    const u32 ucap = plan_dev_unit_capacity(n);           /* of THIS chunk (<= the capacity the buffers were sized for) */
    const unsigned gp = std::min<unsigned>((n + PT - 1) / PT, MAX_BLOCKS), gu = std::min<unsigned>((ucap + PT - 1) / PT, MAX_BLOCKS);
    cudaError_t e;
    if ((e = cudaMemsetAsync(state, 0, sizeof(PlanState), st)) != cudaSuccess) return e;
    pd_a_kernel<<<gp, PT, 0, st>>>(in, state, oplen);
    pd_b_kernel<<<gp, PT, 0, st>>>(in, state, key, idx);
    if ((e = cub::DeviceRadixSort::SortPairs(tmp, tb, (const u64 *)key, skey, (const u32 *)idx, order, (int)n, 0, 64, st)) != cudaSuccess) return e;
    pd_c_kernel<<<gp, PT, 0, st>>>(n, skey, state, runmark);
    tb = w.cub_bytes;
    if ((e = cub::DeviceScan::InclusiveScan(tmp, tb, (const u32 *)runmark, run_start, MaxU32(), (int)n, st)) != cudaSuccess) return e;
    pd_d_kernel<<<gp, PT, 0, st>>>(n, skey, run_start, uhead);
    tb = w.cub_bytes;
    if ((e = cub::DeviceScan::ExclusiveSum(tmp, tb, (const u32 *)uhead, urank, (int)n, st)) != cudaSuccess) return e;
    pd_e_kernel<<<1, 32, 0, st>>>(n, skey, urank, uhead, state);
    pd_f_kernel<<<gp, PT, 0, st>>>(in, order, run_start, uhead, urank, state, units);
    pd_g_kernel<<<gu, PT, 0, st>>>(ucap, state, units, u3);
    U3 zero; zero.a = zero.b = zero.c = 0;
    tb = w.cub_bytes;
    if ((e = cub::DeviceScan::ExclusiveScan(tmp, tb, (const U3 *)u3, u3x, AddU3(), zero, (int)ucap, st)) != cudaSuccess) return e;
    pd_h_kernel<<<gu, PT, 0, st>>>(in, ucap, state, units, u3, u3x);
    tb = w.cub_bytes;
    if ((e = cub::DeviceScan::InclusiveSum(tmp, tb, (const u64 *)oplen, ops_end, (int)n, st)) != cudaSuccess) return e;
    pd_i_kernel<<<1, 32, 0, st>>>(in, state, ops_end);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    return cudaMemcpyAsync(h_summary, &state->sum, sizeof(PlanSummary), cudaMemcpyDeviceToHost, st);
}

}  // namespace mglsw
