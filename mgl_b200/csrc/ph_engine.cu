/*
 * ph_engine.cu -- PairHMM on the B200: kernels, device buffers and the C-ABI of include/mgl_pairhmm.h.
 *
 * One warp per (read, haplotype) pair, running PhLane (ph_lane.h).  The job of the reference's JNI entry
 * (...MicrosoftPairHmm.cc:77-222): float for every pair, the pairs whose float result is below MIN_ACCEPTED again in
 * double, log10(result) - log10(INITIAL_CONSTANT) out.  Here that is two launches per precision shape with no host
 * round trip between them: the float kernel writes the final value or appends the pair to a redo list (atomic
 * counter), the double kernel is launched over the worst case and its warps beyond the counter leave at once.
 *
 * Lanes per pair (32, or 16 with two pairs per warp) and rows per lane are picked per launch from the longest read, so
 * a 150-base read is 10 rows on each of 15 lanes in one band; reads beyond 256 bases go band by band through a carry
 * row in shared memory.
 * Roofline: FP32 (FP64 for the redo pass) FMA pipe -- 8 flops-instructions and 2 selects per cell, no O(cells)
 * memory traffic at all (a pair reads readLen * 5 + hapLen bytes and writes 8).
 */
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/mgl_pairhmm.h"
#include "ph_lane.h"
#include "ph_tables.h"

/* column steps per loop trip (the loop-carried registers rotate once per trip).  Measured [B200, r02q]: 10 rows per lane
   1 814 / 1 646 / 1 697 GCUPS with 1 / 2 / 4 steps per trip, 7 rows per lane 1 463 / 1 482 / 1 533. */
#ifndef MGLPH_UNROLL
#define MGLPH_UNROLL (RPL >= 10 ? 1 : 4)
#endif

namespace mglph {

struct JobDev {
    const uint8_t *reads, *haps;
    const int64_t *read_off;      /* R + 1: byte offset of each read's five tracks */
    const int64_t *hap_off;       /* H + 1 */
    const int *read_list;         /* the reads of this launch (one shape bucket); its length x H work items */
    int R, H;
};

template <class N>
struct KArgs {
    JobDev job;
    Tables<N> tb;
    N log10_initial;
    double *out;                  /* R * H */
    unsigned int *redo;           /* float pass: pairs to recompute in double; double pass: the list to run (or NULL: all pairs) */
    unsigned int *redo_count;
    unsigned int items;           /* work items of this launch (not FROM_LIST) */
    int carry_cols;               /* columns the shared-memory carry row of one warp holds */
};

template <class N, int G> struct ShflT;
template <int G> struct ShflT<float, G> { static __device__ __forceinline__ float up(float v) { return __shfl_up_sync(0xffffffffu, v, 1, G); } };
template <int G> struct ShflT<double, G> { static __device__ __forceinline__ double up(double v) { return __shfl_up_sync(0xffffffffu, v, 1, G); } };

/*
 * G lanes per pair: 32 (one pair per warp) or 16 (two pairs per warp, each on RPL up to 10 rows per lane: the
 * per-step overhead -- shuffles, skew test, register rotation -- is spread over twice the cells and the skew is 15
 * columns instead of 31).  Work item i = warp * (32 / G) + group; FROM_LIST: item i is pair redo[i] for
 * i < *redo_count (the double pass behind a float pass).  The groups of a warp run in lock step over the longer of
 * their haplotypes and the larger of their band counts; a group that has finished sits the rest out.
 */
template <class N, int RPL, int G, bool FROM_LIST, bool ESCALATE>
__global__ void __launch_bounds__(128) ph_forward_kernel(const KArgs<N> a)
{
    constexpr int PPW = 32 / G;
    constexpr int kUnroll = MGLPH_UNROLL;
    extern __shared__ __align__(16) unsigned char s_raw[];
    const int warp_in_block = threadIdx.x >> 5, wl = threadIdx.x & 31, grp = wl / G, lane = wl % G;
    const unsigned int count = FROM_LIST ? *a.redo_count : a.items;
    /* the pass over the redo list does not know its length on the host: a bounded grid walks it (a float pass that
       escalates nothing costs a few hundred blocks that leave at once, not one block per four pairs of the whole job) */
    const unsigned int w_stride = FROM_LIST ? gridDim.x * (blockDim.x >> 5) : 0u;
    for (unsigned int w = blockIdx.x * (blockDim.x >> 5) + warp_in_block; w * PPW < count; w += w_stride) {
    unsigned int item = w * PPW + (unsigned int)grp;
    const bool valid = item < count;
    if (!valid) item = w * PPW;                          /* an odd tail: shadow the warp's first pair, write nothing */
    unsigned int p;
    int r, h;
    if (FROM_LIST) {
        p = a.redo[item];
        r = (int)(p / (unsigned int)a.job.H); h = (int)(p % (unsigned int)a.job.H);
    } else {
        r = a.job.read_list[item / (unsigned int)a.job.H]; h = (int)(item % (unsigned int)a.job.H);
        p = (unsigned int)r * (unsigned int)a.job.H + (unsigned int)h;
    }
    PairView pv;
    pv.rl = (int)((a.job.read_off[r + 1] - a.job.read_off[r]) / 5);
    pv.hl = (int)(a.job.hap_off[h + 1] - a.job.hap_off[h]);
    pv.rs = a.job.reads + a.job.read_off[r];
    pv.q = pv.rs + pv.rl; pv.ins = pv.q + pv.rl; pv.del = pv.ins + pv.rl; pv.gcp = pv.del + pv.rl;
    pv.hap = a.job.haps + MGLPH_HAP_PAD + a.job.hap_off[h];

    Top<N> *carry = reinterpret_cast<Top<N> *>(s_raw) + (size_t)(warp_in_block * PPW + grp) * (a.carry_cols + 1);
    const N y_initial = a.tb.initial / (N)pv.hl;                       /* compute_prob_scalar.cc:104 */
    const int nb = ph_bands(pv.rl, RPL, G);
    int nb_max = nb, hl_max = pv.hl;
    if (PPW > 1) {
        nb_max = max(nb_max, __shfl_xor_sync(0xffffffffu, nb_max, G));
        hl_max = max(hl_max, __shfl_xor_sync(0xffffffffu, hl_max, G));
    }
    const int hl_own = pv.hl;
    PhLane<N, RPL, G> L;
    for (int b = 0; b < nb_max; b++) {
        const bool live = b < nb;
        if (live) L.band_begin(a.tb, pv, b, lane, y_initial);
        pv.hl = live ? hl_own : 0;                                     /* a finished group: every step is outside 1..hl */
        const bool park = live && b + 1 < nb;
        const bool any_park = b + 1 < nb_max;
        unsigned hb_next = ph_hap_byte(pv, 0, lane);
#pragma unroll kUnroll
        for (int t = 0; t < hl_max + G - 1; t++) {
            const unsigned hb = hb_next;
            hb_next = ph_hap_byte(pv, t + 1, lane);
            Top<N> top;
            top.m = ShflT<N, G>::up(L.bottom.m);
            top.x = ShflT<N, G>::up(L.bottom.x);
            top.y = ShflT<N, G>::up(L.bottom.y);
            if (lane == 0) {
                if (b == 0) { top.m = (N)0; top.x = (N)0; top.y = y_initial; }
                else if (t + 1 <= pv.hl) top = carry[t + 1];
            }
            const bool active = L.step(pv, t, top, hb);
            if (any_park) {
                if (park && active && lane == G - 1) carry[t - (G - 1) + 1] = L.bottom;
                __syncwarp();
            }
        }
    }
    const int ll = ph_last_lane(pv.rl, RPL, G);
    if (valid && lane == ll) {
        const N res = L.result;
        if (ESCALATE && (double)res < (double)MGLPH_MIN_ACCEPTED) {    /* ...MicrosoftPairHmm.cc:182 */
            a.redo[atomicAdd(a.redo_count, 1u)] = p;
        } else {
            a.out[p] = log10((double)res) - (double)a.log10_initial;    /* .cc:188,206 */
        }
    }
    if (!FROM_LIST) break;
    __syncwarp();                                        /* the carry row is reused by the warp's next pair */
    }
}

/* shape = G * 100 + RPL; `items` = pairs (or the worst-case length of the redo list) */
template <class N, bool FROM_LIST, bool ESCALATE>
static cudaError_t launch_forward(int shape, unsigned int items, int carry_cols, KArgs<N> a, cudaStream_t st)
{
    if (items == 0) return cudaSuccess;
    a.items = items;
    a.carry_cols = carry_cols;
    const int g = shape / 100;
    const unsigned int ppw = 32u / (unsigned int)g;
    const unsigned int warps = (items + ppw - 1) / ppw;
    const unsigned int blocks = FROM_LIST ? std::min<unsigned int>((warps + 3) / 4, 148u * 8u) : (warps + 3) / 4;
    const size_t smem = (size_t)4 * ppw * (size_t)(carry_cols + 1) * sizeof(Top<N>);
#define PH_LAUNCH(R_, G_)                                                                                               \
    do {                                                                                                                \
        if (smem > 48 * 1024) {                                                                                         \
            cudaError_t e = cudaFuncSetAttribute(ph_forward_kernel<N, R_, G_, FROM_LIST, ESCALATE>,                     \
                                                 cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);               \
            if (e != cudaSuccess) return e;                                                                             \
        }                                                                                                               \
        ph_forward_kernel<N, R_, G_, FROM_LIST, ESCALATE><<<blocks, 128, smem, st>>>(a);                                \
    } while (0)
    switch (shape) {
    case 3201: PH_LAUNCH(1, 32); break;
    case 3202: PH_LAUNCH(2, 32); break;
    case 3203: PH_LAUNCH(3, 32); break;
    case 3204: PH_LAUNCH(4, 32); break;
    case 3205: PH_LAUNCH(5, 32); break;
    case 3206: PH_LAUNCH(6, 32); break;
    case 3208: PH_LAUNCH(8, 32); break;
    case 1605: if constexpr (sizeof(N) == 4) { PH_LAUNCH(5, 16); break; } else return cudaErrorInvalidValue;
    case 1606: if constexpr (sizeof(N) == 4) { PH_LAUNCH(6, 16); break; } else return cudaErrorInvalidValue;
    case 1607: if constexpr (sizeof(N) == 4) { PH_LAUNCH(7, 16); break; } else return cudaErrorInvalidValue;
    case 1608: if constexpr (sizeof(N) == 4) { PH_LAUNCH(8, 16); break; } else return cudaErrorInvalidValue;
    case 1610: if constexpr (sizeof(N) == 4) { PH_LAUNCH(10, 16); break; } else return cudaErrorInvalidValue;
    default: return cudaErrorInvalidValue;
    }
#undef PH_LAUNCH
    return cudaGetLastError();
}

/* whole-warp shape for reads of up to max_rl bases (one band up to 256 rows) */
static int shape_g32(int max_rl)
{
    int rpl = (max_rl + 31) / 32;
    if (rpl == 7 || rpl > 8) rpl = 8;
    return 3200 + (rpl < 1 ? 1 : rpl);
}
/* float pass: reads of 65..160 bases run two pairs per warp */
static int shape_float(int max_rl)
{
    if (max_rl > 64 && max_rl <= 160) {
        const int rpl = (max_rl + 15) / 16;
        return 1600 + (rpl == 9 ? 10 : rpl);
    }
    return shape_g32(max_rl);
}

struct Buf {
    void *p = nullptr; size_t cap = 0;
    cudaError_t reserve(size_t bytes)
    {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        const size_t want = bytes + bytes / 4 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

class Engine {
public:
    explicit Engine(int device) : device_(device) {}
    ~Engine()
    {
        cudaSetDevice(device_);
        for (Buf *b : {&reads_, &haps_, &roff_, &hoff_, &out_, &redo_, &tab_, &order_}) b->release();
        if (stream_) cudaStreamDestroy(stream_);
        if (ev0_) cudaEventDestroy(ev0_);
        if (ev1_) cudaEventDestroy(ev1_);
    }
    std::string err;
    std::mutex mu;
    bool force_g32 = false;       /* developer knob (MGL_PAIRHMM_G32=1): float pass on whole-warp shapes only */

    int init()
    {
        cudaError_t e;
        int n = 0;
        if (cudaGetDeviceCount(&n) != cudaSuccess || device_ < 0 || device_ >= n) { err = "no such CUDA device"; return MGL_PH_E_NO_DEVICE; }
        cudaDeviceProp prop;
        if ((e = cudaGetDeviceProperties(&prop, device_)) != cudaSuccess) return cuda_fail(e, "device properties");
        if (prop.major != 10) { err = "built for sm_100a (B200) only"; return MGL_PH_E_NO_DEVICE; }
        if ((e = cudaSetDevice(device_)) != cudaSuccess) return cuda_fail(e, "cudaSetDevice");
        { const char *v = std::getenv("MGL_PAIRHMM_G32"); force_g32 = v && v[0] == '1'; }
        if ((e = cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking)) != cudaSuccess) return cuda_fail(e, "stream");
        if ((e = cudaEventCreate(&ev0_)) != cudaSuccess || (e = cudaEventCreate(&ev1_)) != cudaSuccess) return cuda_fail(e, "event");
        /* tables: float ph2pr | float mm | double ph2pr | double mm */
        const HostTables<float> &tf = tables_f();
        const HostTables<double> &td = tables_d();
        const size_t fb = (128 + MGLPH_MM_ENTRIES) * sizeof(float), db = (128 + MGLPH_MM_ENTRIES) * sizeof(double);
        const size_t foff = 0, doff = (fb + 255) / 256 * 256;
        if ((e = tab_.reserve(doff + db)) != cudaSuccess) return cuda_fail(e, "cudaMalloc tables");
        std::vector<unsigned char> host(doff + db, 0);
        std::memcpy(host.data() + foff, tf.ph2pr.data(), 128 * sizeof(float));
        std::memcpy(host.data() + foff + 128 * sizeof(float), tf.mm.data(), MGLPH_MM_ENTRIES * sizeof(float));
        std::memcpy(host.data() + doff, td.ph2pr.data(), 128 * sizeof(double));
        std::memcpy(host.data() + doff + 128 * sizeof(double), td.mm.data(), MGLPH_MM_ENTRIES * sizeof(double));
        if ((e = cudaMemcpy(tab_.p, host.data(), host.size(), cudaMemcpyHostToDevice)) != cudaSuccess) return cuda_fail(e, "H2D tables");
        tf_.ph2pr = (const float *)((unsigned char *)tab_.p + foff); tf_.mm = tf_.ph2pr + 128; tf_.initial = tf.initial;
        td_.ph2pr = (const double *)((unsigned char *)tab_.p + doff); td_.mm = td_.ph2pr + 128; td_.initial = td.initial;
        return MGL_PH_OK;
    }

    /* validate the three buffers, build the offset tables, upload */
    int upload(const int32_t *lengths, const uint8_t *reads, const uint8_t *haps, mgl_pairhmm_stats &st)
    {
        if (!lengths) { err = "NULL lengths"; return MGL_PH_E_INVALID; }
        const int32_t R = lengths[0];
        if (R < 0) { err = "negative read count"; return MGL_PH_E_INVALID; }
        const int32_t H = lengths[1 + R];
        if (H < 0) { err = "negative haplotype count"; return MGL_PH_E_INVALID; }
        if ((R > 0 && !reads) || (H > 0 && !haps)) { err = "NULL buffer"; return MGL_PH_E_INVALID; }
        if ((uint64_t)R * (uint64_t)H > 0xfffffff0ull) { err = "more than 2^32 pairs in one call"; return MGL_PH_E_INVALID; }
        std::vector<int64_t> ro((size_t)R + 1, 0), ho((size_t)H + 1, 0);
        max_rl_ = 0; max_hl_ = 0;
        uint64_t sum_rl = 0, sum_hl = 0;
        for (int32_t r = 0; r < R; r++) {
            const int32_t n = lengths[1 + r];
            if (n <= 0) { err = "empty read"; return MGL_PH_E_INVALID; }
            ro[(size_t)r + 1] = ro[(size_t)r] + 5ll * n;
            max_rl_ = n > max_rl_ ? n : max_rl_;
            sum_rl += (uint64_t)n;
        }
        for (int32_t h = 0; h < H; h++) {
            const int32_t n = lengths[2 + R + h];
            if (n <= 0) { err = "empty haplotype"; return MGL_PH_E_INVALID; }
            ho[(size_t)h + 1] = ho[(size_t)h] + n;
            max_hl_ = n > max_hl_ ? n : max_hl_;
            sum_hl += (uint64_t)n;
        }
        R_ = R; H_ = H;
        /*
         * Shape buckets: a read's shape (lanes per pair x rows per lane) follows from its own length, so the reads are
         * grouped by shape and each group is one launch -- a few 250-base reads do not push a region of 101-base
         * reads onto the widest shape.  order = read indices, stable within a bucket.
         */
        buckets_.clear();
        std::vector<int> order((size_t)R);
        {
            std::vector<std::pair<int, int>> key((size_t)R);
            for (int32_t r = 0; r < R; r++) key[(size_t)r] = {shape_g32(lengths[1 + r]) * 10000 + (force_g32 ? 0 : shape_float(lengths[1 + r])), r};
            std::stable_sort(key.begin(), key.end(), [](const std::pair<int, int> &x, const std::pair<int, int> &y) { return x.first < y.first; });
            for (int32_t r = 0; r < R; r++) {
                order[(size_t)r] = key[(size_t)r].second;
                const int n = lengths[1 + key[(size_t)r].second];
                if (buckets_.empty() || buckets_.back().key != key[(size_t)r].first)
                    buckets_.push_back(Bucket{key[(size_t)r].first, force_g32 ? shape_g32(n) : shape_float(n), shape_g32(n), r, 0, 0});
                buckets_.back().count++;
                buckets_.back().max_rl = std::max(buckets_.back().max_rl, n);
            }
        }
        cells_ = sum_rl * sum_hl;
        cudaError_t e;
        if ((e = cudaSetDevice(device_)) != cudaSuccess) return cuda_fail(e, "cudaSetDevice");
        const size_t np = (size_t)R * (size_t)H;
        if ((e = reads_.reserve((size_t)ro[(size_t)R] + 16)) != cudaSuccess || (e = haps_.reserve((size_t)ho[(size_t)H] + 2 * MGLPH_HAP_PAD + (size_t)max_hl_)) != cudaSuccess ||
            (e = roff_.reserve(ro.size() * 8)) != cudaSuccess || (e = hoff_.reserve(ho.size() * 8)) != cudaSuccess ||
            (e = order_.reserve(order.size() * 4 + 4)) != cudaSuccess ||
            (e = out_.reserve(np * 8 + 8)) != cudaSuccess || (e = redo_.reserve(np * 4 + 16)) != cudaSuccess)
            return cuda_fail(e, "cudaMalloc");
        if (np == 0) return MGL_PH_OK;
        if ((e = cudaMemcpyAsync(reads_.p, reads, (size_t)ro[(size_t)R], cudaMemcpyHostToDevice, stream_)) != cudaSuccess ||
            (e = cudaMemcpyAsync((uint8_t *)haps_.p + MGLPH_HAP_PAD, haps, (size_t)ho[(size_t)H], cudaMemcpyHostToDevice, stream_)) != cudaSuccess ||
            (e = cudaMemcpyAsync(roff_.p, ro.data(), ro.size() * 8, cudaMemcpyHostToDevice, stream_)) != cudaSuccess ||
            (e = cudaMemcpyAsync(order_.p, order.data(), order.size() * 4, cudaMemcpyHostToDevice, stream_)) != cudaSuccess ||
            (e = cudaMemcpyAsync(hoff_.p, ho.data(), ho.size() * 8, cudaMemcpyHostToDevice, stream_)) != cudaSuccess)
            return cuda_fail(e, "H2D");
        if ((e = cudaStreamSynchronize(stream_)) != cudaSuccess) return cuda_fail(e, "H2D");   /* ro / ho are locals */
        st.h2d_bytes += (uint64_t)ro[(size_t)R] + (uint64_t)ho[(size_t)H] + (ro.size() + ho.size()) * 8;
        return MGL_PH_OK;
    }

    /* the kernels of one job on stream_, bracketed by events */
    int run(int use_double, mgl_pairhmm_stats &st)
    {
        const size_t np = (size_t)R_ * (size_t)H_;
        st.pairs += np; st.cells += cells_;
        if (np == 0) return MGL_PH_OK;
        cudaError_t e;
        if ((e = cudaSetDevice(device_)) != cudaSuccess) return cuda_fail(e, "cudaSetDevice");
        /* the redo pass runs one shape for whatever the float pass hands over: that of the longest read */
        const int shape_redo = shape_g32(max_rl_);
        const bool multi = max_rl_ > 256;
        JobDev job;
        job.reads = (const uint8_t *)reads_.p; job.haps = (const uint8_t *)haps_.p;
        job.read_off = (const int64_t *)roff_.p; job.hap_off = (const int64_t *)hoff_.p;
        job.R = R_; job.H = H_; job.read_list = nullptr;
        unsigned int *redo_count = (unsigned int *)redo_.p, *redo = redo_count + 4;
        const int carry_cols = multi ? max_hl_ + 1 : 0;
        if (4 * (size_t)(carry_cols + 1) * sizeof(Top<double>) > 220 * 1024) {
            err = "haplotype of " + std::to_string(max_hl_) + " bases with reads longer than 256: carry row exceeds shared memory";
            return MGL_PH_E_TOO_LONG;
        }
        KArgs<float> af;
        af.job = job; af.tb = tf_; af.log10_initial = tables_f().log10_initial; af.out = (double *)out_.p;
        af.redo = redo; af.redo_count = redo_count; af.carry_cols = carry_cols;
        KArgs<double> ad;
        ad.job = job; ad.tb = td_; ad.log10_initial = tables_d().log10_initial; ad.out = (double *)out_.p;
        ad.redo = redo; ad.redo_count = redo_count; ad.carry_cols = carry_cols;
        if ((e = cudaEventRecord(ev0_, stream_)) != cudaSuccess) return cuda_fail(e, "event record");
        /* one launch per shape bucket (upload()): 150-base reads two pairs per warp on 10 rows per lane, a 250-base read
           one pair per warp on 8; the double kernels stay one pair per warp (two registers per value) */
        if (!use_double && (e = cudaMemsetAsync(redo_count, 0, 16, stream_)) != cudaSuccess) return cuda_fail(e, "memset");
        for (const Bucket &bk : buckets_) {
            const unsigned int items = (unsigned int)bk.count * (unsigned int)H_;
            const int cc = bk.max_rl > 256 ? carry_cols : 0;
            af.job.read_list = ad.job.read_list = (const int *)order_.p + bk.first;
            if (use_double) {
                if ((e = launch_forward<double, false, false>(bk.shape_d, items, cc, ad, stream_)) != cudaSuccess) return cuda_fail(e, "double launch");
            } else {
                if ((e = launch_forward<float, false, true>(bk.shape_f, items, cc, af, stream_)) != cudaSuccess) return cuda_fail(e, "float launch");
            }
            st.launches += 1;
        }
        if (use_double) st.pairs_double += np;
        else {
            if ((e = launch_forward<double, true, false>(shape_redo, (unsigned int)np, carry_cols, ad, stream_)) != cudaSuccess) return cuda_fail(e, "double launch");
            st.launches += 1;
        }
        if ((e = cudaEventRecord(ev1_, stream_)) != cudaSuccess) return cuda_fail(e, "event record");
        return MGL_PH_OK;
    }

    int fetch(double *out, int use_double, mgl_pairhmm_stats &st)
    {
        const size_t np = (size_t)R_ * (size_t)H_;
        if (np == 0) return MGL_PH_OK;
        cudaError_t e;
        unsigned int count = 0;
        if (out && (e = cudaMemcpyAsync(out, out_.p, np * 8, cudaMemcpyDeviceToHost, stream_)) != cudaSuccess) return cuda_fail(e, "D2H");
        if (!use_double && (e = cudaMemcpyAsync(&count, redo_.p, 4, cudaMemcpyDeviceToHost, stream_)) != cudaSuccess) return cuda_fail(e, "D2H");
        if ((e = cudaStreamSynchronize(stream_)) != cudaSuccess) return cuda_fail(e, "kernel execution");
        float ms = 0;
        if (cudaEventElapsedTime(&ms, ev0_, ev1_) == cudaSuccess) st.device_ms += ms;
        if (!use_double) st.pairs_double += count;
        if (out) st.d2h_bytes += np * 8;
        return MGL_PH_OK;
    }

    int cuda_fail(cudaError_t e, const char *what)
    {
        err = std::string(what) + ": " + cudaGetErrorString(e);
        cudaGetLastError();
        return MGL_PH_E_CUDA;
    }

private:
    int device_;
    cudaStream_t stream_ = nullptr;
    cudaEvent_t ev0_ = nullptr, ev1_ = nullptr;
    Buf reads_, haps_, roff_, hoff_, out_, redo_, tab_, order_;
    struct Bucket { int key, shape_f, shape_d, first, count, max_rl; };
    std::vector<Bucket> buckets_;
    Tables<float> tf_{};
    Tables<double> td_{};
    int R_ = 0, H_ = 0, max_rl_ = 0, max_hl_ = 0;
    uint64_t cells_ = 0;
};

/*
 * ---- mgl_pairhmm_compute_noalloc: the same job out of a static device array (include/mgl_pairhmm.h) ----
 *
 * No counterpart in the reference.  For callers that cannot call cudaMalloc: everything Engine keeps in cudaMalloc'ed
 * buffers is one file-scope __device__ array here.  g_noalloc_scratch, as host code sees it, is only a host-side
 * stand-in for the array: its device address comes from cudaGetSymbolAddress, once per device with that device
 * current (noalloc_setup), and only that address is ever handed to a kernel or a copy.
 * The array is shared by every call and keeps the last call's contents; what must start at zero (the redo counter)
 * is cleared per call, everything else is overwritten before it is read.
 */
struct NoAllocScratch {
    unsigned int redo_count[64];                                   /* [0]: the float pass's escalation counter */
    float tab_f[128 + MGLPH_MM_ENTRIES];
    double tab_d[128 + MGLPH_MM_ENTRIES];
    int64_t read_off[MGL_PH_NOALLOC_MAX_READS + 1];
    int64_t hap_off[MGL_PH_NOALLOC_MAX_HAPS + 1];
    int read_list[MGL_PH_NOALLOC_MAX_READS];                       /* identity: one shape for every read */
    int32_t lengths[MGL_PH_NOALLOC_MAX_READS + MGL_PH_NOALLOC_MAX_HAPS + 2];
    unsigned int redo[MGL_PH_NOALLOC_MAX_PAIRS];
    double out[MGL_PH_NOALLOC_MAX_PAIRS];
    uint8_t reads[5 * MGL_PH_NOALLOC_MAX_READ_BASES + 16];
    uint8_t haps[2 * MGL_PH_NOALLOC_MAX_HAP_BASES + 2 * MGLPH_HAP_PAD];   /* pad | bases | pad + the longest haplotype (ph_hap_byte) */
};
static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(NoAllocScratch)];

/* exclusive scan of mul * len[0, n) into off[0, n]; one block of 128 threads, a contiguous slice per thread */
static __device__ void ph_noalloc_scan(const int32_t *len, int n, int mul, int64_t *off, int64_t *part)
{
    const int per = (n + 127) / 128, b = min(n, (int)threadIdx.x * per), e = min(n, b + per);
    int64_t sum = 0;
    for (int i = b; i < e; i++) sum += (int64_t)mul * len[i];
    part[threadIdx.x] = sum;
    __syncthreads();
    int64_t base = 0;
    for (int k = 0; k < (int)threadIdx.x; k++) base += part[k];
    if (threadIdx.x == 0) off[0] = 0;
    for (int i = b; i < e; i++) { base += (int64_t)mul * len[i]; off[i + 1] = base; }
    __syncthreads();
}

/* lengths (as uploaded) -> byte offsets of the reads' five tracks and of the haplotypes, and the identity read list */
__global__ void __launch_bounds__(128) ph_noalloc_offsets_kernel(NoAllocScratch *s, int R, int H)
{
    __shared__ int64_t part[128];
    ph_noalloc_scan(s->lengths + 1, R, 5, s->read_off, part);
    ph_noalloc_scan(s->lengths + 2 + R, H, 1, s->hap_off, part);
    for (int r = (int)threadIdx.x; r < R; r += 128) s->read_list[r] = r;
}

static constexpr int kNoAllocMaxDevices = 64;
static NoAllocScratch *g_noalloc_dev[kNoAllocMaxDevices];          /* device address of the array, per device, once set up */
static std::mutex g_noalloc_mu;
static thread_local char g_noalloc_err[256];

static int noalloc_fail(int rc, const char *what, cudaError_t e = cudaSuccess)
{
    if (e != cudaSuccess) { std::snprintf(g_noalloc_err, sizeof g_noalloc_err, "%s: %s", what, cudaGetErrorString(e)); cudaGetLastError(); }
    else std::snprintf(g_noalloc_err, sizeof g_noalloc_err, "%s", what);
    return rc;
}

/* once per device, with that device current: the array's device address and the quality tables */
static int noalloc_setup(int device)
{
    if (g_noalloc_dev[device]) return MGL_PH_OK;
    cudaError_t e;
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return noalloc_fail(MGL_PH_E_CUDA, "device properties", e);
    if (prop.major != 10) return noalloc_fail(MGL_PH_E_NO_DEVICE, "built for sm_100a (B200) only");
    NoAllocScratch *s = nullptr;
    if ((e = cudaGetSymbolAddress((void **)&s, g_noalloc_scratch)) != cudaSuccess || !s)
        return noalloc_fail(MGL_PH_E_CUDA, "cudaGetSymbolAddress(static scratch)", e);
    const HostTables<float> &tf = tables_f();
    const HostTables<double> &td = tables_d();
    if ((e = cudaMemcpy(s->tab_f, tf.ph2pr.data(), 128 * sizeof(float), cudaMemcpyHostToDevice)) != cudaSuccess ||
        (e = cudaMemcpy(s->tab_f + 128, tf.mm.data(), MGLPH_MM_ENTRIES * sizeof(float), cudaMemcpyHostToDevice)) != cudaSuccess ||
        (e = cudaMemcpy(s->tab_d, td.ph2pr.data(), 128 * sizeof(double), cudaMemcpyHostToDevice)) != cudaSuccess ||
        (e = cudaMemcpy(s->tab_d + 128, td.mm.data(), MGLPH_MM_ENTRIES * sizeof(double), cudaMemcpyHostToDevice)) != cudaSuccess)
        return noalloc_fail(MGL_PH_E_CUDA, "H2D tables", e);
    g_noalloc_dev[device] = s;
    return MGL_PH_OK;
}

static int noalloc_compute(int device, const int32_t *lengths, const uint8_t *reads, const uint8_t *haps, int use_double,
                           double *out, cudaStream_t st, int flags)
{
    g_noalloc_err[0] = 0;
    if (!lengths || !out) return noalloc_fail(MGL_PH_E_INVALID, "NULL lengths / out");
    const int32_t R = lengths[0];
    if (R < 0) return noalloc_fail(MGL_PH_E_INVALID, "negative read count");
    if (R > MGL_PH_NOALLOC_MAX_READS) return noalloc_fail(MGL_PH_E_TOO_LARGE, "more reads than MGL_PH_NOALLOC_MAX_READS");
    const int32_t H = lengths[1 + R];
    if (H < 0) return noalloc_fail(MGL_PH_E_INVALID, "negative haplotype count");
    if (H > MGL_PH_NOALLOC_MAX_HAPS) return noalloc_fail(MGL_PH_E_TOO_LARGE, "more haplotypes than MGL_PH_NOALLOC_MAX_HAPS");
    if ((R > 0 && !reads) || (H > 0 && !haps)) return noalloc_fail(MGL_PH_E_INVALID, "NULL buffer");
    const uint64_t np = (uint64_t)R * (uint64_t)H;
    if (np > MGL_PH_NOALLOC_MAX_PAIRS) return noalloc_fail(MGL_PH_E_TOO_LARGE, "more pairs than MGL_PH_NOALLOC_MAX_PAIRS");
    int max_rl = 0, max_hl = 0;
    uint64_t sum_rl = 0, sum_hl = 0;
    for (int32_t r = 0; r < R; r++) {
        const int32_t n = lengths[1 + r];
        if (n <= 0) return noalloc_fail(MGL_PH_E_INVALID, "empty read");
        max_rl = n > max_rl ? n : max_rl; sum_rl += (uint64_t)n;
        if (sum_rl > MGL_PH_NOALLOC_MAX_READ_BASES) return noalloc_fail(MGL_PH_E_TOO_LARGE, "more read bases than MGL_PH_NOALLOC_MAX_READ_BASES");
    }
    for (int32_t h = 0; h < H; h++) {
        const int32_t n = lengths[2 + R + h];
        if (n <= 0) return noalloc_fail(MGL_PH_E_INVALID, "empty haplotype");
        max_hl = n > max_hl ? n : max_hl; sum_hl += (uint64_t)n;
        if (sum_hl > MGL_PH_NOALLOC_MAX_HAP_BASES) return noalloc_fail(MGL_PH_E_TOO_LARGE, "more haplotype bases than MGL_PH_NOALLOC_MAX_HAP_BASES");
    }
    if (np == 0) return MGL_PH_OK;
    const int carry_cols = max_rl > 256 ? max_hl + 1 : 0;
    if (4 * (size_t)(carry_cols + 1) * sizeof(Top<double>) > 220 * 1024)
        return noalloc_fail(MGL_PH_E_TOO_LONG, "haplotype too long for the carry row of reads longer than 256 bases");

    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    cudaError_t e;
    int n_dev = 0, prev = -1;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || device < 0 || device >= n_dev || device >= kNoAllocMaxDevices) {
        cudaGetLastError();
        return noalloc_fail(MGL_PH_E_NO_DEVICE, "no such CUDA device");
    }
    if ((e = cudaGetDevice(&prev)) != cudaSuccess || (e = cudaSetDevice(device)) != cudaSuccess) return noalloc_fail(MGL_PH_E_CUDA, "cudaSetDevice", e);
    struct Restore { int d; ~Restore() { if (d >= 0) cudaSetDevice(d); } } restore{prev};
    int rc = noalloc_setup(device);
    if (rc != MGL_PH_OK) return rc;
    NoAllocScratch *s = g_noalloc_dev[device];

    if ((e = cudaMemcpyAsync(s->lengths, lengths, ((size_t)R + (size_t)H + 2) * sizeof(int32_t), cudaMemcpyDefault, st)) != cudaSuccess ||
        (e = cudaMemcpyAsync(s->reads, reads, 5 * (size_t)sum_rl, cudaMemcpyDefault, st)) != cudaSuccess ||
        (e = cudaMemcpyAsync(s->haps + MGLPH_HAP_PAD, haps, (size_t)sum_hl, cudaMemcpyDefault, st)) != cudaSuccess)
        return noalloc_fail(MGL_PH_E_CUDA, "input copy", e);
    ph_noalloc_offsets_kernel<<<1, 128, 0, st>>>(s, R, H);
    if ((e = cudaGetLastError()) != cudaSuccess) return noalloc_fail(MGL_PH_E_CUDA, "offsets launch", e);

    JobDev job;
    job.reads = s->reads; job.haps = s->haps;
    job.read_off = s->read_off; job.hap_off = s->hap_off;
    job.read_list = s->read_list; job.R = R; job.H = H;
    KArgs<double> ad;
    ad.job = job; ad.tb.ph2pr = s->tab_d; ad.tb.mm = s->tab_d + 128; ad.tb.initial = tables_d().initial;
    ad.log10_initial = tables_d().log10_initial; ad.out = s->out;
    ad.redo = s->redo; ad.redo_count = s->redo_count; ad.carry_cols = carry_cols;
    if (use_double) {
        if ((e = launch_forward<double, false, false>(shape_g32(max_rl), (unsigned int)np, carry_cols, ad, st)) != cudaSuccess)
            return noalloc_fail(MGL_PH_E_CUDA, "double launch", e);
    } else {
        KArgs<float> af;
        af.job = job; af.tb.ph2pr = s->tab_f; af.tb.mm = s->tab_f + 128; af.tb.initial = tables_f().initial;
        af.log10_initial = tables_f().log10_initial; af.out = s->out;
        af.redo = s->redo; af.redo_count = s->redo_count; af.carry_cols = carry_cols;
        /* the counter is the one thing that must start at zero: the array still holds the previous call's */
        if ((e = cudaMemsetAsync(s->redo_count, 0, sizeof s->redo_count, st)) != cudaSuccess) return noalloc_fail(MGL_PH_E_CUDA, "memset", e);
        if ((e = launch_forward<float, false, true>(shape_float(max_rl), (unsigned int)np, carry_cols, af, st)) != cudaSuccess)
            return noalloc_fail(MGL_PH_E_CUDA, "float launch", e);
        if ((e = launch_forward<double, true, false>(shape_g32(max_rl), (unsigned int)np, carry_cols, ad, st)) != cudaSuccess)
            return noalloc_fail(MGL_PH_E_CUDA, "double launch", e);
    }
    if ((e = cudaMemcpyAsync(out, s->out, (size_t)np * sizeof(double), cudaMemcpyDefault, st)) != cudaSuccess)
        return noalloc_fail(MGL_PH_E_CUDA, "output copy", e);
    if (!(flags & MGL_PH_NOALLOC_NO_SYNC) && (e = cudaStreamSynchronize(st)) != cudaSuccess)
        return noalloc_fail(MGL_PH_E_CUDA, "kernel execution", e);
    return MGL_PH_OK;
}

}  // namespace mglph

struct mgl_pairhmm_handle { mglph::Engine *engine; };

extern "C" {

int mgl_pairhmm_create(int device, mgl_pairhmm_handle **out)
{
    if (!out) return MGL_PH_E_INVALID;
    *out = nullptr;
    mglph::Engine *e = new mglph::Engine(device);
    const int rc = e->init();
    if (rc != MGL_PH_OK) { std::fprintf(stderr, "mgl_pairhmm: %s\n", e->err.c_str()); delete e; return rc; }
    *out = new mgl_pairhmm_handle{e};
    return MGL_PH_OK;
}

void mgl_pairhmm_destroy(mgl_pairhmm_handle *h)
{
    if (!h) return;
    delete h->engine;
    delete h;
}

const char *mgl_pairhmm_last_error(mgl_pairhmm_handle *h) { return h ? h->engine->err.c_str() : "NULL handle"; }

int mgl_pairhmm_compute(mgl_pairhmm_handle *h, const int32_t *lengths, const uint8_t *reads, const uint8_t *haps,
                        int use_double, double *out, mgl_pairhmm_stats *stats)
{
    if (!h || !out) return MGL_PH_E_INVALID;
    mgl_pairhmm_stats st;
    std::memset(&st, 0, sizeof st);
    std::lock_guard<std::mutex> lk(h->engine->mu);
    int rc = h->engine->upload(lengths, reads, haps, st);
    if (rc == MGL_PH_OK) rc = h->engine->run(use_double, st);
    if (rc == MGL_PH_OK) rc = h->engine->fetch(out, use_double, st);
    if (stats) *stats = st;
    return rc;
}

int mgl_pairhmm_resident_prepare(mgl_pairhmm_handle *h, const int32_t *lengths, const uint8_t *reads, const uint8_t *haps)
{
    if (!h) return MGL_PH_E_INVALID;
    mgl_pairhmm_stats st;
    std::memset(&st, 0, sizeof st);
    std::lock_guard<std::mutex> lk(h->engine->mu);
    return h->engine->upload(lengths, reads, haps, st);
}

int mgl_pairhmm_resident_run(mgl_pairhmm_handle *h, int use_double, mgl_pairhmm_stats *stats)
{
    if (!h) return MGL_PH_E_INVALID;
    mgl_pairhmm_stats st;
    std::memset(&st, 0, sizeof st);
    std::lock_guard<std::mutex> lk(h->engine->mu);
    int rc = h->engine->run(use_double, st);
    if (rc == MGL_PH_OK) rc = h->engine->fetch(nullptr, use_double, st);
    if (stats) *stats = st;
    return rc;
}

int mgl_pairhmm_resident_fetch(mgl_pairhmm_handle *h, double *out)
{
    if (!h || !out) return MGL_PH_E_INVALID;
    mgl_pairhmm_stats st;
    std::memset(&st, 0, sizeof st);
    std::lock_guard<std::mutex> lk(h->engine->mu);
    return h->engine->fetch(out, 1, st);
}

int mgl_pairhmm_compute_noalloc(int device, const int32_t *lengths, const uint8_t *reads, const uint8_t *haps,
                                int use_double, double *out, void *stream, int flags)
{
    return mglph::noalloc_compute(device, lengths, reads, haps, use_double, out, (cudaStream_t)stream, flags);
}

uint64_t mgl_pairhmm_noalloc_scratch_bytes(void) { return sizeof(mglph::NoAllocScratch); }

const char *mgl_pairhmm_noalloc_last_error(void) { return mglph::g_noalloc_err; }

}  // extern "C"
