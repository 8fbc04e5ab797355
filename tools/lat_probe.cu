/*
 * tools/lat_probe.cu -- developer microbenchmark (not on the product path): what bounds the fill's inner loop?
 *   1. dependent-issue latency of the DPX / IMAD instructions the cell uses (one warp, clock64)
 *   2. throughput of the bare cell recurrence (sw_cell.h, no memory traffic, no shuffles) as a function of
 *      resident warps per SM sub-partition and of the number of independent F chains per thread
 * Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I mgl_b200/csrc -o tools/lat_probe tools/lat_probe.cu
 */
#include <cstdio>
#include <cuda_runtime.h>
#include "sw_cell.h"

using namespace mglsw;

/* helpers of the byte-permute ("ACGT") cell variant that was measured and dropped */
__host__ __device__ inline bool is_acgt(unsigned b) { return b == 0x41u || b == 0x43u || b == 0x47u || b == 0x54u; }
__host__ __device__ inline unsigned base_table(unsigned b) { return is_acgt(b) ? (0x01010101u ^ (1u << (8u * ((b >> 1) & 3u)))) : 0x01010101u; }
__device__ inline unsigned byte_perm_sw(unsigned x, unsigned y, unsigned s)
{
    unsigned r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(x), "r"(y), "r"(s));
    return r;
}

template <int W>
__device__ __forceinline__ unsigned chain_op(unsigned x, unsigned a, unsigned b)
{
    if (W == 0) return __vimax3_u16x2(x, a, b);
    if (W == 1) return __viaddmax_u16x2(x, a, b);
    if (W == 2) { unsigned r; asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(x), "r"(a), "r"(b)); return r; }
    if (W == 3) return (unsigned)__vimax3_s32((int)x, (int)a, (int)b);
    if (W == 4) return (unsigned)__viaddmax_s32((int)x, (int)a, (int)b);
    if (W == 5) { unsigned r; asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(x), "r"(a), "r"(b)); return r; }
    if (W == 6) return __vminu2(x, a);
    /* W == 7: the cell's F chain: max3 -> imad -> addmax */
    unsigned h = __vimax3_u16x2(x, a, b);
    unsigned g; asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(g) : "r"(h), "r"(1u + (a & 0u)), "r"(b));
    return __viaddmax_u16x2(x, a, g);
}

template <int W>
__global__ void lat_kernel(unsigned *out, long long *cyc, unsigned a, unsigned b, int n)
{
    unsigned x = threadIdx.x + a;
    long long t0 = clock64();
    for (int i = 0; i < n; i++) {
#pragma unroll
        for (int k = 0; k < 16; k++) x = chain_op<W>(x, a, b);
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
    if (x == 0x12345u) out[0] = x;
}

/* bare recurrence: K cells per step in CH independent chains (K / CH cells each), like FillLane::step without anything else */
template <class T, int K, int CH>
__global__ void __launch_bounds__(128) cell_kernel(unsigned *out, CellConsts c, int steps)
{
    unsigned e[K], hm[K], bq[K];
#pragma unroll
    for (int k = 0; k < K; k++) { e[k] = threadIdx.x + k; hm[k] = threadIdx.x * 3 + k; bq[k] = (threadIdx.x + k) & 3; }
    unsigned ra = threadIdx.x & 3, f0[CH], d0[CH];
#pragma unroll
    for (int q = 0; q < CH; q++) { f0[q] = threadIdx.x + q; d0[q] = q; }
    for (int s = 0; s < steps; s++) {
        ra = (ra * 5u + 1u) & 3u;
#pragma unroll
        for (int q = 0; q < CH; q++) {
            unsigned f = f0[q], diag = d0[q];
#pragma unroll
            for (int k = q * (K / CH); k < (q + 1) * (K / CH); k++) {
                const unsigned old = hm[k];
                sw_cell<T>(c, ra, bq[k], diag, e[k], f, hm[k]);
                diag = old;
            }
            f0[q] = f; d0[q] = diag;
        }
    }
    unsigned s = 0;
#pragma unroll
    for (int k = 0; k < K; k++) s ^= e[k] ^ hm[k];
    if (s == 0x12345678u) out[0] = s;
}

/* the same with the byte-permute mismatch indicator (ACGT fast path): tables from a shared-memory array, one PRMT per cell */
template <class T, int V>
__device__ __forceinline__ void cell_v(const CellConsts &c, unsigned m, unsigned hdm, unsigned &e, unsigned &f, unsigned &hm)
{
    const unsigned d = m * c.neg_delta + hdm;
    const unsigned h = T::max3(d, e, f);
    const unsigned hgo = V >= 1 ? h + c.neg_go : h * c.one + c.neg_go;
    e = T::addmax(e, c.neg_ge, hgo);
    f = T::addmax(f, c.neg_ge, hgo);
    hm = V >= 2 ? h + c.match : h * c.one + c.match;
}

template <class T, int K, int CH, int V = 0>
__global__ void __launch_bounds__(128) cellp_kernel(unsigned *out, CellConsts c, int steps)
{
    __shared__ unsigned lut[256];
    for (unsigned b = threadIdx.x; b < 256u; b += 128u) lut[b] = base_table(b);
    __syncthreads();
    unsigned e[K], hm[K], bq[K];
#pragma unroll
    for (int k = 0; k < K; k++) { e[k] = threadIdx.x + k; hm[k] = threadIdx.x * 3 + k; bq[k] = ((threadIdx.x + k) & 3) | 0x8080u | ((4u + ((threadIdx.x + 2 * k) & 3)) << 8); }
    unsigned ra = threadIdx.x & 3, f0[CH], d0[CH];
    const unsigned char acgt[4] = {0x41, 0x43, 0x54, 0x47};
#pragma unroll
    for (int q = 0; q < CH; q++) { f0[q] = threadIdx.x + q; d0[q] = q; }
    unsigned ta = lut[0x41], tb = lut[0x43];
    for (int s = 0; s < steps; s++) {
        ra = (ra * 5u + 1u) & 3u;
        const unsigned ta_ = ta, tb_ = tb;
        ta = lut[acgt[ra]]; tb = lut[acgt[(ra + 1) & 3]];
#pragma unroll
        for (int q = 0; q < CH; q++) {
            unsigned f = f0[q], diag = d0[q];
#pragma unroll
            for (int k = q * (K / CH); k < (q + 1) * (K / CH); k++) {
                const unsigned old = hm[k];
                cell_v<T, V>(c, byte_perm_sw(ta_, tb_, bq[k]), diag, e[k], f, hm[k]);
                diag = old;
            }
            f0[q] = f; d0[q] = diag;
        }
    }
    unsigned s = 0;
#pragma unroll
    for (int k = 0; k < K; k++) s ^= e[k] ^ hm[k];
    if (s == 0x12345678u) out[0] = s;
}

template <class T, int K, int CH, int V = 0>
static void run_cellp(const char *name, unsigned *out, int sms)
{
    Scores sc; sc.match = 10; sc.mismatch = -15; sc.go = 30; sc.ge = 5;
    CellConsts c = make_consts<T>(sc, 20000);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int bps = 1; bps <= 4; bps++) {
        const size_t smem = (200 * 1024) / bps;
        cudaFuncSetAttribute(cellp_kernel<T, K, CH, V>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        const int steps = 20000;
        cellp_kernel<T, K, CH, V><<<sms * bps, 128, smem>>>(out, c, 100);
        cudaEventRecord(e0);
        cellp_kernel<T, K, CH, V><<<sms * bps, 128, smem>>>(out, c, steps);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
        const double cells = (double)sms * bps * 128 * (double)steps * K * T::PAIRS;
        std::printf("%-14s K=%d chains=%d warps/subpartition=%d: %.2f TCUPS (byte-permute cell, variant %d)\n", name, K, CH, bps, cells / (ms * 1e-3) / 1e12, V);
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
}

template <class T, int K, int CH>
static void run_cell(const char *name, unsigned *out, int sms)
{
    Scores sc; sc.match = 10; sc.mismatch = -15; sc.go = 30; sc.ge = 5;
    CellConsts c = make_consts<T>(sc, 20000);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    int clk = 0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    for (int bps = 1; bps <= 4; bps++) {        /* blocks of 4 warps per SM = warps per sub-partition */
        const size_t smem = (200 * 1024) / bps;
        cudaFuncSetAttribute(cell_kernel<T, K, CH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        const int steps = 20000;
        cell_kernel<T, K, CH><<<sms * bps, 128, smem>>>(out, c, 100);
        cudaEventRecord(e0);
        cell_kernel<T, K, CH><<<sms * bps, 128, smem>>>(out, c, steps);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
        const double cells = (double)sms * bps * 128 * (double)steps * K * T::PAIRS;
        int nb = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, cell_kernel<T, K, CH>, 128, smem);
        std::printf("%-14s K=%d chains=%d warps/subpartition=%d (occupancy allows %d blocks): %.2f TCUPS, %.1f warp-instr-equivalents (8/cell) issue %.0f%%\n",
                    name, K, CH, bps, nb, cells / (ms * 1e-3) / 1e12, 8.0,
                    100.0 * (cells / T::PAIRS / 32.0 * 8.0) / (ms * 1e-3) / ((double)sms * 4 * clk * 1e3));
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
}

int main()
{
    unsigned *out; long long *cyc;
    cudaMalloc(&out, 64); cudaMalloc(&cyc, 64);
    cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
    const int sms = prop.multiProcessorCount;
    const char *names[8] = {"vimnmx3.u16x2", "viaddmnmx.u16x2", "imad", "vimnmx3.s32", "viaddmnmx.s32", "lop3", "vimnmx.u16x2", "max3->imad->addmax"};
    long long h = 0;
    const int n = 4096;
#define LAT(W) lat_kernel<W><<<1, 32>>>(out, cyc, 3u, 5u, 64); lat_kernel<W><<<1, 32>>>(out, cyc, 3u, 5u, n); \
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); std::printf("latency %-22s %.2f cycles per dependent op%s\n", names[W], (double)h / (n * 16.0), W == 7 ? " (3 instructions)" : "");
    LAT(0) LAT(1) LAT(2) LAT(3) LAT(4) LAT(5) LAT(6) LAT(7)
    run_cell<S16x2, 20, 1>("cell u16x2", out, sms);
    run_cell<S16x2, 20, 2>("cell u16x2", out, sms);
    run_cell<S16x2, 20, 4>("cell u16x2", out, sms);
    run_cellp<S16x2, 20, 2>("cellp u16x2", out, sms);
    run_cellp<S16x2, 20, 4>("cellp u16x2", out, sms);
    run_cellp<S16x2, 20, 1, 1>("cellp u16x2", out, sms);
    run_cellp<S16x2, 20, 2, 1>("cellp u16x2", out, sms);
    run_cellp<S16x2, 20, 1, 2>("cellp u16x2", out, sms);
    run_cellp<S16x2, 20, 2, 2>("cellp u16x2", out, sms);
    run_cellp<S32, 24, 1>("cellp s32", out, sms);
    run_cellp<S32, 24, 2>("cellp s32", out, sms);
    run_cellp<S32, 24, 2, 1>("cellp s32", out, sms);
    run_cell<S32, 24, 1>("cell s32", out, sms);
    run_cell<S32, 24, 2>("cell s32", out, sms);
    std::printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
