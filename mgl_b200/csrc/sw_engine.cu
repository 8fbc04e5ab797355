/* sw_engine.cu -- see sw_engine.h. */
#include "sw_engine.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <condition_variable>
#include <cstring>
#include <mutex>
#include <thread>
#include <sched.h>

#include "../../include/mgl_sw.h"
#include "sw_cell.h"
#include "sw_config.h"
#include "sw_kernels.h"
#include "sw_layout.h"
#include "sw_pack.h"

namespace mglsw {

/* ---------------------------------------------------------------- 2-bit transport form -> arena bytes */

/*
 * One thread per 16 bases: a u32 of the packed form (sw_pack.h) becomes the 16 original bytes, one 16-byte
 * store.  `dst` is 16-byte aligned (the range starts at arena_lo & ~15); the last thread stores byte by byte
 * so that nothing beyond arena_hi is written.  Pure HBM streaming: 1.25 B moved per base.
 */
__global__ void __launch_bounds__(256) sw_unpack2_kernel(const u32 *__restrict__ packed, uint8_t *__restrict__ dst, u64 nbases)
{
    const u64 t = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    const u64 base = t * 16;
    if (base >= nbases) return;
    const u32 w = packed[t];
    u32 out[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const u32 b = (w >> (8 * k)) & 0xffu;
        const u32 sel = (b & 3u) | (b & 0xcu) << 2 | (b & 0x30u) << 4 | (b & 0xc0u) << 6;   /* one code per selector nibble */
        out[k] = __byte_perm(PACK2_BASES, 0u, sel);
    }
    if (base + 16 <= nbases) {
        *reinterpret_cast<uint4 *>(dst + base) = make_uint4(out[0], out[1], out[2], out[3]);
    } else {
        for (u64 k = 0; base + k < nbases; k++) dst[base + k] = (uint8_t)(out[k >> 2] >> (8 * (k & 3)));
    }
}

/* ---------------------------------------------------------------- helper threads */

HelperPool::~HelperPool()
{
    { std::lock_guard<std::mutex> lk(mu_); stop_ = true; }
    cv_.notify_all();
    for (std::thread &t : th_) t.join();
}

void HelperPool::worker()
{
    std::unique_lock<std::mutex> lk(mu_);
    for (;;) {
        cv_.wait(lk, [&] { return stop_ || next_ < total_; });
        if (stop_) return;
        const int i = next_++;
        const std::function<void(int)> *fn = task_;
        lk.unlock();
        (*fn)(i);
        lk.lock();
        if (--pending_ == 0) done_cv_.notify_all();
    }
}

void HelperPool::run(int T, const std::function<void(int)> &fn)
{
    if (T <= 1) { fn(0); return; }
    std::lock_guard<std::mutex> one(run_mu_);           /* one run at a time: the calling thread and the staging thread share the pool */
    {
        std::lock_guard<std::mutex> lk(mu_);
        while ((int)th_.size() < T - 1) th_.emplace_back(&HelperPool::worker, this);
        task_ = &fn; next_ = 1; total_ = T; pending_ = T - 1;
    }
    cv_.notify_all();
    fn(0);
    std::unique_lock<std::mutex> lk(mu_);
    done_cv_.wait(lk, [&] { return pending_ == 0; });
    task_ = nullptr; total_ = 0; next_ = 0;
}

/* ---------------------------------------------------------------- buffers */

cudaError_t DevBuf::reserve(size_t bytes)
{
    if (bytes <= cap) return cudaSuccess;
    release();
    size_t want = bytes + std::min<size_t>(bytes / 8, (size_t)256 << 20) + 256;     /* headroom against regrowth, bounded for the multi-GB checkpoint buffers */
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) { p = nullptr; cap = 0; return e; }
    cap = want;
    return cudaSuccess;
}
void DevBuf::release() { if (p) cudaFree(p); p = nullptr; cap = 0; }

cudaError_t PinBuf::reserve(size_t bytes)
{
    if (bytes <= cap) return cudaSuccess;
    release();
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaHostAlloc(&p, want, cudaHostAllocDefault);
    if (e != cudaSuccess) { p = nullptr; cap = 0; return e; }
    cap = want;
    return cudaSuccess;
}
void PinBuf::release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }

void ChunkDev::release()
{
    units.release(); pair_unit.release(); ops_end.release(); strategy.release();
    h_units.release(); h_pair_unit.release(); h_ops_end.release(); h_strategy.release();
    packed.release(); h_packed.release();
    table.release(); work.release(); h_table.release(); h_sum.release(); h_arena.release();
    if (ev_plan) { cudaEventDestroy(ev_plan); ev_plan = nullptr; }
    for (auto &x : ev_dbg) if (x) { cudaEventDestroy(x); x = nullptr; }
    if (ev_arena) { cudaEventDestroy(ev_arena); ev_arena = nullptr; }
    if (ev_table) { cudaEventDestroy(ev_table); ev_table = nullptr; }
}

void WorkSlot::release()
{
    ckpt.release(); carry.release(); lastrow.release(); colbest.release(); scratch.release();
    startcell.release(); worklist.release(); wstate.release();
    offsets.release(); n_ops.release(); starts.release(); tiles.release(); ops.release(); score3.release(); n_chars.release();
    h_n_ops.release(); h_total.release(); h_res.release(); h_ops.release();
    if (ev_counts) cudaEventDestroy(ev_counts);
    if (ev_done) cudaEventDestroy(ev_done);
    if (ev_fill) cudaEventDestroy(ev_fill);
    if (ev_post) cudaEventDestroy(ev_post);
    if (ev_fork) cudaEventDestroy(ev_fork);
    for (int k = 0; k < MGLSW_SIDE; k++) {
        if (ev_side[k]) cudaEventDestroy(ev_side[k]);
        if (side[k]) cudaStreamDestroy(side[k]);
        ev_side[k] = nullptr; side[k] = nullptr;
    }
    if (stream) cudaStreamDestroy(stream);
    if (post) cudaStreamDestroy(post);
    ev_counts = ev_done = ev_fill = ev_post = ev_fork = nullptr; stream = post = nullptr;
}

/* ---------------------------------------------------------------- engine */

Engine::Engine(int device) : device_(device)
{
    if (cudaSetDevice(device_) != cudaSuccess) { err_ = "cudaSetDevice failed"; return; }
    int prio_lo = 0, prio_hi = 0;
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
    /*
     * The two events align_batch's collecting thread waits on.  A plain event makes that thread spin on a CPU for the
     * whole batch; with one process per GPU on a box of few CPUs per GPU that CPU is one the planner threads lack.
     * MGL_SW_BLOCKING_SYNC=1 / 0 forces sleeping / spinning waits; unset: sleep when the box has fewer than
     * MGLSW_SPIN_MIN_CPUS_PER_GPU hardware threads per visible GPU.
     */
    bool blocking = false;
    {
        const char *v = std::getenv("MGL_SW_BLOCKING_SYNC");
        int ndev = 1;
        if (v && (v[0] == '0' || v[0] == '1')) blocking = v[0] == '1';
        else if (cudaGetDeviceCount(&ndev) == cudaSuccess && ndev > 0)
            blocking = std::thread::hardware_concurrency() < (unsigned)ndev * MGLSW_SPIN_MIN_CPUS_PER_GPU;
    }
    const unsigned wait_flags = cudaEventDisableTiming | (blocking ? cudaEventBlockingSync : 0u);
    wait_flags_ = wait_flags;
    cudaStreamCreateWithPriority(&plan_stream_, cudaStreamNonBlocking, prio_hi);
    cudaStreamCreateWithPriority(&upload_stream_, cudaStreamNonBlocking, prio_lo);
    for (auto &s : slots_) {
        cudaStreamCreateWithPriority(&s.stream, cudaStreamNonBlocking, prio_lo);
        cudaStreamCreateWithPriority(&s.post, cudaStreamNonBlocking, prio_hi);
        cudaEventCreateWithFlags(&s.ev_counts, wait_flags);
        cudaEventCreateWithFlags(&s.ev_done, wait_flags);
        cudaEventCreateWithFlags(&s.ev_fill, cudaEventDisableTiming);
        cudaEventCreateWithFlags(&s.ev_post, cudaEventDisableTiming);
        cudaEventCreateWithFlags(&s.ev_fork, cudaEventDisableTiming);
        for (int k = 0; k < MGLSW_SIDE; k++) {
            cudaStreamCreateWithPriority(&s.side[k], cudaStreamNonBlocking, prio_lo);
            cudaEventCreateWithFlags(&s.ev_side[k], cudaEventDisableTiming);
        }
        cudaEventRecord(s.ev_post, s.post);
    }
    for (auto &e : tev_) cudaEventCreate(&e);
}

Engine::~Engine()
{
    cudaSetDevice(device_);
    drop_resident();
    for (ChunkDev *cd : ring_) { cd->release(); delete cd; }
    for (auto &s : slots_) s.release();
    if (plan_stream_) cudaStreamDestroy(plan_stream_);
    if (upload_stream_) cudaStreamDestroy(upload_stream_);
    for (auto &e : tev_) if (e) cudaEventDestroy(e);
    arena_.release();
}

int Engine::fail(int code, const std::string &msg) { err_ = msg; return code; }

int Engine::cuda_fail(cudaError_t e, const char *what)
{
    err_ = std::string(what) + ": " + cudaGetErrorString(e);
    cudaGetLastError();
    return MGL_SW_E_CUDA;
}

/*
 * device arena = front pad | the caller's bytes [base, base + span) | tail pad long enough for the rows a lane reads
 * past its ref.  Only the byte range a batch's pairs touch is kept on the device (a shard of a multi-GPU call touches
 * 1 / n of the caller's arena); arena_dev() is biased by -base so that device code keeps using the caller's offsets.
 */
int Engine::reserve_arena(const BatchIn &in)
{
    u32 max_tl = 0;
    for (int64_t p = 0; p < in.n; p++) max_tl = std::max(max_tl, in.ref_len[p]);
    return reserve_arena(0, in.arena_bytes, max_tl);
}

int Engine::reserve_arena(u64 base, u64 span, u32 max_tl)
{
    arena_base_ = base;
    const size_t need = MGLSW_ARENA_FRONT_PAD + span + arena_tail_pad(max_tl);
    if (need > arena_.cap) {
        cudaError_t e = arena_.reserve(need);
        if (e != cudaSuccess) return cuda_fail(e, "cudaMalloc arena");
        if ((e = cudaMemset(arena_.p, 0, arena_.cap)) != cudaSuccess) return cuda_fail(e, "cudaMemset arena");
    }
    return MGL_SW_OK;
}

u64 Engine::budget()
{
    if (trace_budget_bytes) return trace_budget_bytes;
    /* taken once, from the memory free before this engine owns any work buffer: chunk boundaries (and so the sizes the
       slots grow to) must not move from one batch to the next as the engine's own buffers eat into the free figure */
    if (budget_auto_) return std::max<u64>(budget_auto_ / (u64)nslots_, 256ull << 20);
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) return 4ull << 30;
    budget_auto_ = std::min<u64>((u64)(free_b * 0.72), (u64)MGLSW_SLOTS * (48ull << 30));   /* all slots + descriptors stay well inside HBM */
    return std::max<u64>(budget_auto_ / (u64)nslots_, 256ull << 20);
}

/*
 * Work slots of a batch.  Four let the traceback of a chunk hide behind the fills of the next three: right for short
 * pairs, whose chunks are capped by pair count.  Longer pairs are capped by the HBM their checkpoints take, and there
 * fewer slots mean larger chunks inside the same memory: launches of whole waves instead of slivers, and, for the
 * pipelined long-pair kernel (one pair per block), enough pairs to fill every block slot -- a chunk of 320 pairs left
 * a third of them empty.  Two slots for long and medium pairs, three for mixed batches.
 * Slots whose buffers no longer fit the per-slot share are released first, so that a handle may alternate batch kinds.
 */
void Engine::choose_slots(u64 total_est, int64_t n)
{
    /* measured [B200, profiles/r10a]: config 4 (140 MB of checkpoints per pair) 1 706 GCUPS on four slots, 1 845 on two;
       config 3 (1 MB per pair) 1 748 / 1 868; the mixed config 5 (0.28 MB per pair on average) 1 607 / 1 715 on three / 1 646 on two */
    const u64 per_pair = n > 0 ? total_est / (u64)n : 0;
    const int want = slots_opt >= 2 && slots_opt <= MGLSW_SLOTS ? slots_opt : (per_pair >= (512ull << 10) ? 2 : (per_pair >= 196608u ? 3 : MGLSW_SLOTS));
    if (want == nslots_) return;
    nslots_ = want;
    const u64 share = budget() + budget() / 4 + (512ull << 20);
    for (int k = 0; k < MGLSW_SLOTS; k++) {
        WorkSlot &ws = slots_[k];
        if (k >= nslots_ || ws.ckpt.cap > share) { cudaStreamSynchronize(ws.post); cudaStreamSynchronize(ws.stream); ws.ckpt.release(); ws.scratch.release(); ws.ops.release(); }
    }
}

/*
 * How many pairs the chunk starting at `first` takes: at most chunk_pairs, and as many as keep
 * the estimated checkpoints inside the budget.
 */
/* CPUs this process may run on (a rank of an 8-GPU job is often confined to a few): sizes the short-lived scan threads */
static unsigned allowed_cpus()
{
    cpu_set_t set;
    CPU_ZERO(&set);
    if (sched_getaffinity(0, sizeof set, &set) == 0) { const int c = CPU_COUNT(&set); if (c > 0) return (unsigned)c; }
    return std::max(1u, std::thread::hardware_concurrency());
}

/* checkpoint bytes of one pair with the planner's geometry (sw_plan.cpp), as if it ran alone in 32-bit arithmetic */
static inline u64 pair_ckpt_bytes(u32 tl, u32 ql)
{
    if (ql <= (u32)MGLSW_G_SHORT * MGLSW_KMAX)       /* one strip of 8 lanes: no division on the path every short pair of a 1 M-pair call takes */
        return (u64)row_blocks(tl, MGLSW_G_SHORT) * lane_stride16((ql + MGLSW_G_SHORT - 1u) / MGLSW_G_SHORT) * 16u * MGLSW_G_SHORT;
    const u32 G = MGLSW_G_LONG;
    const u32 ns = ceil_div_u32(ql, (u32)MGLSW_G_LONG * MGLSW_KMAX);
    const u32 K = regs_per_lane(ql, ns * G);
    return (u64)ns * row_blocks(tl, G) * lane_stride16(K) * 16u * G;
}

/*
 * Pairs per chunk for a batch.  Short pairs: chunk_pairs (a chunk is a millisecond of work, more buys nothing).
 * Longer or mixed pairs: a chunk holds its work slot at least as long as its slowest pair takes (a 10 kbp pair
 * is milliseconds on its own), so the chunk must carry enough other work to fill that time: four times as many
 * pairs, which in practice means "as many as the checkpoint budget holds" (measured on config 5: 16 Ki pairs per
 * chunk 313 GCUPS, 32 Ki 509, 64 Ki 827, 128 Ki 1052).
 */
u32 Engine::chunk_cap(u64 total_est, int64_t n) const
{
    const bool short_pairs = n > 0 && total_est / (u64)n < 196608u;
    return short_pairs ? chunk_pairs : (u32)std::min<u64>((u64)chunk_pairs * 4u, 1u << 22);
}

int64_t Engine::chunk_take(const BatchIn &in, int64_t first, u64 limit, u32 max_pairs, const u32 *est256) const
{
    const int64_t n = std::min<int64_t>(in.n - first, max_pairs ? max_pairs : chunk_pairs);
    u64 est = 0;
    int64_t take = 0;
    for (; take < n; take++) {
        const u64 bytes = est256 ? ((u64)est256[first + take] << 8) : pair_ckpt_bytes(in.ref_len[first + take], in.alt_len[first + take]);
        if (take > 0 && est + bytes > limit) break;
        est += bytes;
    }
    return take;
}

/*
 * Plan pairs [first, first + take) into cd and stage the descriptors in pinned memory.  Touches
 * nothing of the engine but read-only options, so planner threads run it concurrently.
 */
int Engine::plan_range(const BatchIn &in, int64_t first, int64_t take, ChunkDev &cd, std::string &err) const
{
    std::vector<PairIn> &pairs = cd.pairs_scratch;
    pairs.resize((size_t)take);
    u64 lo = ~0ull, hi = 0;                /* the byte range of the arena this chunk touches */
    for (int64_t k = 0; k < take; k++) {
        PairIn &p = pairs[(size_t)k];
        p.ref_off = in.ref_off[first + k]; p.alt_off = in.alt_off[first + k];
        p.tl = in.ref_len[first + k]; p.ql = in.alt_len[first + k];
        p.strategy = in.strategies ? in.strategies[first + k] : in.strategy;
        if (p.ref_off > in.arena_bytes || p.tl > in.arena_bytes - p.ref_off || p.alt_off > in.arena_bytes || p.ql > in.arena_bytes - p.alt_off) {
            err = "pair " + std::to_string(first + k) + " lies outside the arena";
            return MGL_SW_E_INVALID;
        }
        lo = std::min<u64>(lo, std::min(p.ref_off, p.alt_off));
        hi = std::max<u64>(hi, std::max(p.ref_off + p.tl, p.alt_off + p.ql));
    }
    cd.arena_lo = lo; cd.arena_hi = hi; cd.packed_ok = false;
    const PlanStatus ps = plan_chunk(pairs.data(), (u32)take, in.scores, allow_s16, cd.plan);
    if (ps == PLAN_BAD_SCORES) { err = "score set out of range"; return MGL_SW_E_SCORES; }
    if (ps == PLAN_BAD_PAIR) { err = "empty or over-long sequence in batch"; return MGL_SW_E_PAIR; }
    cd.first = first;
    cd.n = (u32)take;
    if (in.score_only) {
        /*
         * Nothing reads the checkpoints of a score-only run.  The fill kernels are the same binaries (a "do not
         * store" flag inside the step loop cost the general path 7 % on config 3 when it was tried), so their
         * regions are folded onto a small sink instead: every warp writes region w mod S of S regions that together
         * fit the L2, and the stores never reach HBM as O(cells) traffic.
         */
        u64 maxw = 0;
        for (const UnitDesc &u : cd.plan.units) maxw = std::max<u64>(maxw, warp_ckpt_words(u.nstrips, u.rb_count, u.regs));
        const u64 S = std::max<u64>(1, std::min<u64>(64, (64ull << 20) / (maxw * 4 + 1)));
        for (const Launch &L : cd.plan.launches) {
            const u32 upw = 32u / (u32)L.G;
            for (u32 w = 0; w < L.n_warps; w++)
                for (u32 g = 0; g < upw; g++) cd.plan.units[L.first_unit + w * upw + g].ckpt_off = (w % S) * maxw;
        }
        cd.plan.ckpt_words = S * maxw;
    }

    /* stage descriptors in pinned memory so the uploads are asynchronous */
    cudaError_t e;
    const size_t nu = cd.plan.units.size();
    if ((e = cd.h_units.reserve(nu * sizeof(UnitDesc))) != cudaSuccess || (e = cd.h_pair_unit.reserve(take * sizeof(u32))) != cudaSuccess ||
        (e = cd.h_ops_end.reserve(take * sizeof(u64))) != cudaSuccess || (e = cd.h_strategy.reserve(take * sizeof(int))) != cudaSuccess) {
        err = std::string("pinned alloc: ") + cudaGetErrorString(e);
        return MGL_SW_E_CUDA;
    }
    std::memcpy(cd.h_units.p, cd.plan.units.data(), nu * sizeof(UnitDesc));
    std::memcpy(cd.h_pair_unit.p, cd.plan.pair_unit.data(), take * sizeof(u32));
    std::memcpy(cd.h_ops_end.p, cd.plan.ops_end.data(), take * sizeof(u64));
    int *st = (int *)cd.h_strategy.p;
    for (int64_t k = 0; k < take; k++) st[k] = pairs[(size_t)k].strategy;
    return MGL_SW_OK;
}

/*
 * Transport form of the chunk's byte range (sw_pack.h), packed on a planner thread from a 16-byte aligned
 * start so that the unpack kernel's stores are aligned.  The few bytes in front of arena_lo are the caller's
 * too (they belong to an earlier chunk, which writes the same values).  A range with any byte outside ACGT,
 * or one too small to be worth a second launch, goes up raw.
 */
int Engine::pack_chunk(const BatchIn &in, ChunkDev &cd, std::string &err) const
{
    cd.packed_ok = false;
    if (cd.arena_hi <= cd.arena_lo || cd.arena_hi - cd.arena_lo < MGLSW_PACK_MIN_BYTES) return MGL_SW_OK;
    const u64 lo16 = cd.arena_lo & ~15ull, nb = cd.arena_hi - lo16;
    const cudaError_t e = cd.h_packed.reserve(pack2_bytes(nb));
    if (e != cudaSuccess) { err = std::string("pinned alloc: ") + cudaGetErrorString(e); return MGL_SW_E_CUDA; }
    cd.packed_ok = pack2(in.arena + lo16, nb, (uint8_t *)cd.h_packed.p);
    return MGL_SW_OK;
}

/*
 * The same on T short-lived threads (the staging thread of the device planner calls this for the chunks it decides to
 * pack): the range is cut at multiples of 128 bytes, so every piece starts on a whole packed word.
 */
int Engine::pack_chunk_mt(const BatchIn &in, ChunkDev &cd, int T, std::string &err)
{
    cd.packed_ok = false;
    if (cd.arena_hi <= cd.arena_lo || cd.arena_hi - cd.arena_lo < MGLSW_PACK_MIN_BYTES) return MGL_SW_OK;
    const u64 lo16 = cd.arena_lo & ~15ull, nb = cd.arena_hi - lo16;
    const cudaError_t e = cd.h_packed.reserve(pack2_bytes(nb));
    if (e != cudaSuccess) { err = std::string("pinned alloc: ") + cudaGetErrorString(e); return MGL_SW_E_CUDA; }
    T = (int)std::min<u64>((u64)std::max(1, T), std::max<u64>(1, nb >> 20));
    std::vector<char> ok((size_t)T, 0);
    auto part = [&](int t) {
        const u64 a = (nb * (u64)t / (u64)T) & ~127ull, b = t + 1 == T ? nb : (nb * (u64)(t + 1) / (u64)T) & ~127ull;
        ok[(size_t)t] = b <= a || pack2(in.arena + lo16 + a, b - a, (uint8_t *)cd.h_packed.p + a / 4) ? 1 : 0;
    };
    pool_.run(T, part);
    bool all = true;
    for (char c : ok) all = all && c;
    cd.packed_ok = all;
    return MGL_SW_OK;
}

int Engine::plan_next(const BatchIn &in, int64_t first, ChunkDev &cd)
{
    std::string err;
    const int rc = plan_range(in, first, chunk_take(in, first, budget(), chunk_cap_, nullptr), cd, err);
    return rc == MGL_SW_OK ? rc : fail(rc, err);
}

/*
 * Device planner, host half.  The chunk's slice of the caller's pair table, plus the running end of each pair's CIGAR
 * scratch (36 bytes per pair), is staged in pinned memory -- the caller's arrays are ordinary memory, and a copy from
 * them would be staged by the driver, synchronously -- and goes up on the planner's stream, followed by the planning
 * kernel and the copy home of the summary.  The host touches each pair twice here (the memcpy, and one loop for the
 * min / max of its offsets and the scratch sum) and never again.
 */
int Engine::issue_plan_dev(const BatchIn &in, int64_t first, int64_t take, ChunkDev &cd, BatchStats &bs, std::string &err, bool upload_arena, bool want_pack)
{
    auto cfail = [&](cudaError_t ce, const char *what) { err = std::string(what) + ": " + cudaGetErrorString(ce); cudaGetLastError(); return (int)MGL_SW_E_CUDA; };
    cudaError_t e;
    const u32 n = (u32)take;
    cd.first = first; cd.n = n; cd.packed_ok = false;
    cd.plan.reset();
    if (!cd.ev_plan && (e = cudaEventCreateWithFlags(&cd.ev_plan, wait_flags_)) != cudaSuccess) return cfail(e, "event create");
    if (n > cd.lay.n_cap || !cd.work.p) {
        cd.lay = plan_dev_layout(std::max<u32>(n + n / 8, 4096));
        if ((e = cd.work.reserve(cd.lay.total)) != cudaSuccess) return cfail(e, "cudaMalloc planner work buffer");
    }
    const size_t tbytes = (size_t)n * 36;
    if ((e = cd.h_table.reserve(tbytes)) != cudaSuccess) return cfail(e, "pinned alloc");
    if ((e = cd.h_sum.reserve(sizeof(PlanSummary))) != cudaSuccess) return cfail(e, "pinned alloc");
    if ((e = cd.table.reserve(tbytes)) != cudaSuccess) return cfail(e, "cudaMalloc pair table");
    if ((e = cd.units.reserve((size_t)plan_dev_unit_capacity(n) * sizeof(UnitDesc))) != cudaSuccess) return cfail(e, "cudaMalloc units");
    /* table = ref_off | alt_off | ops_end (u64 each) | ref_len | alt_len | strategy (u32 each) */
    char *h = (char *)cd.h_table.p;
    u64 *h_ro = (u64 *)h, *h_ao = (u64 *)(h + (size_t)n * 8), *h_oe = (u64 *)(h + (size_t)n * 16);
    u32 *h_tl = (u32 *)(h + (size_t)n * 24), *h_ql = (u32 *)(h + (size_t)n * 28);
    int *h_st = (int *)(h + (size_t)n * 32);
    std::memcpy(h_ro, in.ref_off + first, (size_t)n * 8);
    std::memcpy(h_ao, in.alt_off + first, (size_t)n * 8);
    std::memcpy(h_tl, in.ref_len + first, (size_t)n * 4);
    std::memcpy(h_ql, in.alt_len + first, (size_t)n * 4);
    if (in.strategies) std::memcpy(h_st, in.strategies + first, (size_t)n * 4);
    else std::fill(h_st, h_st + n, in.strategy);
    u64 lo = ~0ull, hi = 0, acc = 0;           /* the byte range of the arena this chunk touches; CIGAR scratch: tl + ql + 4 elements per pair, in chunk order */
    for (u32 k = 0; k < n; k++) {
        const u64 ro = h_ro[k], ao = h_ao[k];
        const u32 tl = h_tl[k], ql = h_ql[k];
        lo = std::min(lo, std::min(ro, ao));
        hi = std::max(hi, std::max(ro + tl, ao + ql));
        acc += (u64)tl + ql + 4;
        h_oe[k] = acc;
    }
    cd.arena_lo = std::min(lo, in.arena_bytes); cd.arena_hi = std::min(hi, in.arena_bytes);   /* offsets outside the arena are refused by the planner */
    if (cd.arena_lo > cd.arena_hi) cd.arena_lo = cd.arena_hi;
    const bool tl_dbg = std::getenv("MGL_SW_TIMELINE") != nullptr;
    /*
     * The table rides the SAME stream as the arena bytes (table 0, arena 0, table 1, arena 1, ...): the copy engine
     * then serves them in exactly that order.  On the planner's own stream a table waited behind every arena copy this
     * thread had queued ahead -- 5 ms for the second chunk of a batch, with the GPU idle (profiles/r06c).
     */
    if (!cd.ev_table && (e = cudaEventCreateWithFlags(&cd.ev_table, cudaEventDisableTiming)) != cudaSuccess) return cfail(e, "event create");
    if (tl_dbg) {
        if (!cd.ev_dbg[0]) for (auto &x : cd.ev_dbg) cudaEventCreate(&x);
        cudaEventRecord(cd.ev_dbg[0], upload_stream_);
    }
    if ((e = cudaMemcpyAsync(cd.table.p, cd.h_table.p, tbytes, cudaMemcpyHostToDevice, upload_stream_)) != cudaSuccess) return cfail(e, "H2D pair table");
    if ((e = cudaEventRecord(cd.ev_table, upload_stream_)) != cudaSuccess) return cfail(e, "event record");
    if ((e = cudaStreamWaitEvent(plan_stream_, cd.ev_table, 0)) != cudaSuccess) return cfail(e, "stream wait");
    if (tl_dbg) cudaEventRecord(cd.ev_dbg[1], plan_stream_);
    bs.h2d_bytes += tbytes;
    const char *d = (const char *)cd.table.p;
    PlanDevIn pin;
    pin.ref_off = (const u64 *)d; pin.alt_off = (const u64 *)(d + (size_t)n * 8); pin.ops_end = (const u64 *)(d + (size_t)n * 16);
    pin.tl = (const u32 *)(d + (size_t)n * 24); pin.ql = (const u32 *)(d + (size_t)n * 28);
    pin.strategy = (const int *)(d + (size_t)n * 32);
    pin.n = n; pin.arena_bytes = in.arena_bytes; pin.s = in.scores; pin.allow_s16 = allow_s16 ? 1 : 0; pin.score_only = in.score_only ? 1 : 0;
    if ((e = plan_dev_run(pin, cd.lay, (UnitDesc *)cd.units.p, cd.work.p, (PlanSummary *)cd.h_sum.p, plan_stream_)) != cudaSuccess)
        return cfail(e, "device planner");
    if (tl_dbg) cudaEventRecord(cd.ev_dbg[2], plan_stream_);
    if ((e = cudaEventRecord(cd.ev_plan, plan_stream_)) != cudaSuccess) return cfail(e, "event record");
    /*
     * The chunk's bytes of the arena go up NOW, on their own stream, as far ahead of the chunk's kernels as the ring
     * lets this thread run: the upload depends on nothing the planner computes, and on a box whose GPUs share the
     * host's memory path (eight GPUs: 23 GB/s each, tools/h2d_ceiling.py) the link is what a step waits for -- it
     * must never idle behind a plan or a slot.  The chunk's fill waits for ev_arena.
     */
    cd.arena_ahead = false;
    if (upload_arena) {
        if (!cd.ev_arena && (e = cudaEventCreateWithFlags(&cd.ev_arena, cudaEventDisableTiming)) != cudaSuccess) return cfail(e, "event create");
        if (want_pack) {
            /* the 2-bit transport from this thread: worth its CPU time when the link is what the step waits for (the caller decides) */
            const int prc = pack_chunk_mt(in, cd, pack_threads > 0 ? pack_threads : 2, err);
            if (prc != MGL_SW_OK) return prc;
        }
        if (cd.packed_ok) {
            const u64 lo16 = cd.arena_lo & ~15ull, nb = cd.arena_hi - lo16;
            const size_t pb = pack2_bytes(nb);
            if ((e = cd.packed.reserve(pb)) != cudaSuccess) return cfail(e, "cudaMalloc packed arena");
            if ((e = cudaMemcpyAsync(cd.packed.p, cd.h_packed.p, pb, cudaMemcpyHostToDevice, upload_stream_)) != cudaSuccess) return cfail(e, "H2D packed arena");
            sw_unpack2_kernel<<<(unsigned)((nb / 16 + 256) / 256), 256, 0, upload_stream_>>>((const u32 *)cd.packed.p, (uint8_t *)arena_dev() + lo16, nb);
            if ((e = cudaGetLastError()) != cudaSuccess) return cfail(e, "sw_unpack2_kernel");
            bs.h2d_bytes += pb;
            bs.other_launches++;
        } else if (cd.arena_hi > cd.arena_lo) {
            const u64 nb = cd.arena_hi - cd.arena_lo;
            const uint8_t *src = in.arena + cd.arena_lo;
            /*
             * A caller's arena in ordinary pageable memory (a JVM direct ByteBuffer): cudaMemcpyAsync would stage it through
             * the driver's own pinned buffer on THIS thread, synchronously, at ~11 GB/s (631 MB: 58 ms against 17 ms of GPU
             * work).  A few short-lived threads copy the chunk's range into a pinned buffer of the ring entry instead
             * (memory bandwidth scales with threads; the driver's path does not), and the upload is asynchronous again.
             */
            if (arena_pageable_ && nb <= (96ull << 20) && copy_threads > 0) {
                if ((e = cd.h_arena.reserve(nb)) != cudaSuccess) return cfail(e, "pinned alloc");
                const int T = (int)std::min<u64>((u64)std::min<int>(copy_threads, std::max(1, (int)std::thread::hardware_concurrency() / 2)), std::max<u64>(1, nb >> 20));
                auto part = [&](int t) { const u64 lo = nb * t / T, hi = nb * (t + 1) / T; std::memcpy((uint8_t *)cd.h_arena.p + lo, src + lo, hi - lo); };
                pool_.run(T, part);
                src = (const uint8_t *)cd.h_arena.p;
            }
            if ((e = cudaMemcpyAsync((uint8_t *)arena_dev() + cd.arena_lo, src, nb, cudaMemcpyHostToDevice, upload_stream_)) != cudaSuccess)
                return cfail(e, "H2D arena");
            bs.h2d_bytes += nb;
        }
        if ((e = cudaEventRecord(cd.ev_arena, upload_stream_)) != cudaSuccess) return cfail(e, "event record");
        cd.arena_ahead = true;
    }
    cd.d_units = (const UnitDesc *)cd.units.p; cd.d_ops_end = pin.ops_end; cd.d_strategy = pin.strategy;
    bs.other_launches += 1;
    return MGL_SW_OK;
}

int Engine::collect_plan_dev(ChunkDev &cd)
{
    cudaError_t e;
    if ((e = cudaEventSynchronize(cd.ev_plan)) != cudaSuccess) return cuda_fail(e, "device planner");
    if (cd.ev_dbg[0] && std::getenv("MGL_SW_TIMELINE")) {
        float a = 0, b = 0;
        cudaEventSynchronize(cd.ev_dbg[2]);
        cudaEventElapsedTime(&a, cd.ev_dbg[0], cd.ev_dbg[1]);
        cudaEventElapsedTime(&b, cd.ev_dbg[1], cd.ev_dbg[2]);
        std::fprintf(stderr, "[mgl_sw] plan of chunk at pair %lld (n=%u): table H2D %.3f ms, kernels + summary D2H %.3f ms on the planner stream\n", (long long)cd.first, cd.n, a, b);
    }
    const PlanSummary &S = *(const PlanSummary *)cd.h_sum.p;
    if (S.err_invalid) return fail(MGL_SW_E_INVALID, "pair " + std::to_string(cd.first + (int64_t)(0xffffffffu - S.err_invalid)) + " lies outside the arena");
    if (S.err_pair) return fail(MGL_SW_E_PAIR, "empty or over-long sequence in batch (pair " + std::to_string(cd.first + (int64_t)(0xffffffffu - S.err_pair)) + ")");
    if (S.overflow || S.n_launches > MGLSW_PLAN_MAX_LAUNCHES) return fail(MGL_SW_E_INVALID, "device planner: launch table overflow");
    ChunkPlan &P = cd.plan;
    P.reset();
    for (u32 l = 0; l < S.n_launches; l++) {
        const PlanLaunchDev &D = S.launch[l];
        Launch L;
        L.kind = (int)D.kind; L.G = (int)D.G; L.K = (int)D.K; L.pipe = (int)D.pipe; L.nstrips = 0;
        L.first_unit = D.first_unit; L.n_units = D.n_units; L.n_warps = D.n_warps; L.cells = 0;
        P.launches.push_back(L);
    }
    P.n_units = S.n_units;
    P.ckpt_words = S.ckpt_words; P.carry_entries = S.carry_entries; P.lastrow_words = S.lastrow_words; P.scratch_words = S.scratch_words;
    P.cells = S.cells; P.pairs_s16 = S.pairs_s16; P.pairs_s32 = S.pairs_s32; P.s16_bias = (int)S.s16_bias;
    return MGL_SW_OK;
}

int Engine::upload_chunk(const BatchIn &in, ChunkDev &cd, cudaStream_t st, bool upload_arena, BatchStats &bs)
{
    cudaError_t e;
    const size_t nu = cd.plan.units.size(), n = cd.n;
    if ((e = cd.units.reserve(nu * sizeof(UnitDesc))) != cudaSuccess) return cuda_fail(e, "cudaMalloc units");
    if ((e = cd.pair_unit.reserve(n * sizeof(u32))) != cudaSuccess) return cuda_fail(e, "cudaMalloc");
    if ((e = cd.ops_end.reserve(n * sizeof(u64))) != cudaSuccess) return cuda_fail(e, "cudaMalloc");
    if ((e = cd.strategy.reserve(n * sizeof(int))) != cudaSuccess) return cuda_fail(e, "cudaMalloc");
    if ((e = cudaMemcpyAsync(cd.units.p, cd.h_units.p, nu * sizeof(UnitDesc), cudaMemcpyHostToDevice, st)) != cudaSuccess) return cuda_fail(e, "H2D units");
    if ((e = cudaMemcpyAsync(cd.pair_unit.p, cd.h_pair_unit.p, n * sizeof(u32), cudaMemcpyHostToDevice, st)) != cudaSuccess) return cuda_fail(e, "H2D");
    if ((e = cudaMemcpyAsync(cd.ops_end.p, cd.h_ops_end.p, n * sizeof(u64), cudaMemcpyHostToDevice, st)) != cudaSuccess) return cuda_fail(e, "H2D");
    if ((e = cudaMemcpyAsync(cd.strategy.p, cd.h_strategy.p, n * sizeof(int), cudaMemcpyHostToDevice, st)) != cudaSuccess) return cuda_fail(e, "H2D");
    bs.h2d_bytes += nu * sizeof(UnitDesc) + n * (sizeof(u32) + sizeof(u64) + sizeof(int));
    cd.d_units = (const UnitDesc *)cd.units.p; cd.d_ops_end = (const u64 *)cd.ops_end.p; cd.d_strategy = (const int *)cd.strategy.p;
    return upload_arena_range(in, cd, st, upload_arena, bs);
}

/* the bytes of the caller's arena this chunk's pairs touch, onto the chunk's stream (packed form when the planner made one) */
int Engine::upload_arena_range(const BatchIn &in, ChunkDev &cd, cudaStream_t st, bool upload_arena, BatchStats &bs)
{
    cudaError_t e;
    if (upload_arena && cd.packed_ok) {
        const u64 lo16 = cd.arena_lo & ~15ull, nb = cd.arena_hi - lo16;
        const size_t pb = pack2_bytes(nb);
        if ((e = cd.packed.reserve(pb)) != cudaSuccess) return cuda_fail(e, "cudaMalloc packed arena");
        if ((e = cudaMemcpyAsync(cd.packed.p, cd.h_packed.p, pb, cudaMemcpyHostToDevice, st)) != cudaSuccess) return cuda_fail(e, "H2D packed arena");
        sw_unpack2_kernel<<<(unsigned)((nb / 16 + 256) / 256), 256, 0, st>>>((const u32 *)cd.packed.p, (uint8_t *)arena_dev() + lo16, nb);
        if ((e = cudaGetLastError()) != cudaSuccess) return cuda_fail(e, "sw_unpack2_kernel");
        bs.h2d_bytes += pb;
        bs.other_launches++;
    } else if (upload_arena) {
        const u64 lo = cd.arena_lo, hi = cd.arena_hi;
        if (hi > lo) {
            if ((e = cudaMemcpyAsync((uint8_t *)arena_dev() + lo, in.arena + lo, hi - lo, cudaMemcpyHostToDevice, st)) != cudaSuccess)
                return cuda_fail(e, "H2D arena");
            bs.h2d_bytes += hi - lo;
        }
    }
    return MGL_SW_OK;
}

int Engine::ensure_slot(WorkSlot &ws, const ChunkPlan &plan, u32 n, bool score_only, bool text)
{
    cudaError_t e;
    if (text && (e = ws.n_chars.reserve((size_t)n * sizeof(u32))) != cudaSuccess) return cuda_fail(e, "cudaMalloc");
    if (score_only) {                       /* a small shared checkpoint sink (plan_range), no CIGAR scratch: the fill's side buffers and three words per pair */
        if ((e = ws.ckpt.reserve((plan.ckpt_words + 4) * sizeof(u32))) != cudaSuccess) return cuda_fail(e, "cudaMalloc checkpoints");
        if ((e = ws.carry.reserve((plan.carry_entries + 1) * sizeof(Carry))) != cudaSuccess) return cuda_fail(e, "cudaMalloc carry");
        if ((e = ws.lastrow.reserve((plan.lastrow_words + 1) * sizeof(u32))) != cudaSuccess) return cuda_fail(e, "cudaMalloc lastrow");
        if ((e = ws.colbest.reserve(((size_t)plan.n_units + 1) * 4 * sizeof(u32))) != cudaSuccess) return cuda_fail(e, "cudaMalloc colbest");
        if ((e = ws.score3.reserve((size_t)n * 3 * sizeof(int))) != cudaSuccess) return cuda_fail(e, "cudaMalloc");
        return MGL_SW_OK;
    }
    if ((e = ws.ckpt.reserve((plan.ckpt_words + 4) * sizeof(u32))) != cudaSuccess) return cuda_fail(e, "cudaMalloc checkpoints");
    if ((e = ws.carry.reserve((plan.carry_entries + 1) * sizeof(Carry))) != cudaSuccess) return cuda_fail(e, "cudaMalloc carry");
    if ((e = ws.lastrow.reserve((plan.lastrow_words + 1) * sizeof(u32))) != cudaSuccess) return cuda_fail(e, "cudaMalloc lastrow");
    if ((e = ws.colbest.reserve(((size_t)plan.n_units + 1) * 4 * sizeof(u32))) != cudaSuccess) return cuda_fail(e, "cudaMalloc colbest");
    if ((e = ws.scratch.reserve((plan.scratch_words + 1) * sizeof(u32))) != cudaSuccess) return cuda_fail(e, "cudaMalloc scratch");
    if ((e = ws.offsets.reserve((size_t)n * sizeof(int))) != cudaSuccess) return cuda_fail(e, "cudaMalloc");
    if ((e = ws.n_ops.reserve((size_t)n * sizeof(u32))) != cudaSuccess) return cuda_fail(e, "cudaMalloc");
    if ((e = ws.starts.reserve(((size_t)n + 1) * sizeof(u32))) != cudaSuccess) return cuda_fail(e, "cudaMalloc");
    if ((e = ws.tiles.reserve(((size_t)n / 1024 + 2) * sizeof(u32))) != cudaSuccess) return cuda_fail(e, "cudaMalloc");
    if ((e = ws.ops.reserve((plan.scratch_words + 1) * sizeof(u32))) != cudaSuccess) return cuda_fail(e, "cudaMalloc ops");
    if ((e = ws.h_n_ops.reserve(((size_t)n + 1) * sizeof(u32))) != cudaSuccess) return cuda_fail(e, "pinned alloc");
    if ((e = ws.startcell.reserve(((size_t)plan.n_units * 2 + 1) * sizeof(int4))) != cudaSuccess) return cuda_fail(e, "cudaMalloc");
    if ((e = ws.worklist.reserve(((size_t)plan.n_units * 2 + MGLSW_WC_WORDS) * sizeof(u32))) != cudaSuccess) return cuda_fail(e, "cudaMalloc");
    if ((e = ws.wstate.reserve(((size_t)plan.n_units * 2 + 1) * 2 * 3 * sizeof(uint4))) != cudaSuccess) return cuda_fail(e, "cudaMalloc");
    return MGL_SW_OK;
}

/*
 * fill launches on the slot's fill stream, then traceback -> scan -> compact on its (higher
 * priority) post stream.  The slot's previous chunk must have left the post stream before the
 * fill overwrites the trace.
 */
int Engine::run_chunk(ChunkDev &cd, WorkSlot &ws, BatchStats &bs, bool timed, bool score_only, bool text)
{
    cudaError_t e;
    cudaStream_t st = ws.stream, pt = ws.post;
    const Scores &sc = resident_scores_;
    if ((e = cudaStreamWaitEvent(st, ws.ev_post, 0)) != cudaSuccess) return cuda_fail(e, "stream wait");
    if (timed) cudaEventRecord(tev_[0], st);
    /* the buckets are independent: launch i goes to stream i mod (1 + MGLSW_SIDE), side streams fork from and join `st` */
    const int nside = (int)std::min<size_t>(MGLSW_SIDE, cd.plan.launches.size() > 0 ? cd.plan.launches.size() - 1 : 0);
    if (nside > 0) {
        if ((e = cudaEventRecord(ws.ev_fork, st)) != cudaSuccess) return cuda_fail(e, "event record");
        for (int k = 0; k < nside; k++)
            if ((e = cudaStreamWaitEvent(ws.side[k], ws.ev_fork, 0)) != cudaSuccess) return cuda_fail(e, "stream wait");
    }
    /* long pairs of the whole chunk (they spread over several (K) launches): few -> the latency shape of the pipelined kernel */
    u32 pipe_units = 0;
    for (const Launch &L : cd.plan.launches) if (L.pipe) pipe_units += L.n_warps;
    int li = 0;
    for (const Launch &L : cd.plan.launches) {
        const int lane = li++ % (nside + 1);
        cudaStream_t ls = lane == 0 ? st : ws.side[lane - 1];
        FillArgs a;
        a.arena = arena_dev();
        a.units = cd.d_units;
        a.ckpt = (u32 *)ws.ckpt.p;
        a.carry = (Carry *)ws.carry.p;
        a.lastrow = (u32 *)ws.lastrow.p;
        a.colbest = (u32 *)ws.colbest.p;
        a.n_warps = L.n_warps;
        a.c = L.kind == KIND_S16X2 ? make_consts<S16x2>(sc, cd.plan.s16_bias) : make_consts<S32>(sc, 0);
        const int pipe = !L.pipe ? 0 : (pipe_wide == 2 || (pipe_wide == 1 && pipe_units <= MGLSW_PIPE_WIDE_MAX_UNITS)) ? 2 : 1;
        if ((e = launch_fill_any(L.kind, L.G, L.K, pipe, a, L.first_unit, ls)) != cudaSuccess)
            return cuda_fail(e, "fill launch (is a kernel built for this kind/G/K?)");
        bs.fill_launches++;
    }
    for (int k = 0; k < nside; k++) {
        if ((e = cudaEventRecord(ws.ev_side[k], ws.side[k])) != cudaSuccess) return cuda_fail(e, "event record");
        if ((e = cudaStreamWaitEvent(st, ws.ev_side[k], 0)) != cudaSuccess) return cuda_fail(e, "stream wait");
    }
    if (timed) cudaEventRecord(tev_[1], st);
    if ((e = cudaEventRecord(ws.ev_fill, st)) != cudaSuccess) return cuda_fail(e, "event record");
    if ((e = cudaStreamWaitEvent(pt, ws.ev_fill, 0)) != cudaSuccess) return cuda_fail(e, "stream wait");
    if (timed) cudaEventRecord(tev_[2], pt);
    TraceArgs t;
    t.units = cd.d_units;
    t.ops_end = cd.d_ops_end;
    t.strategy = cd.d_strategy;
    t.ckpt = (const u32 *)ws.ckpt.p;
    t.arena = arena_dev();
    t.c16 = make_consts<S16x2>(sc, cd.plan.s16_bias);
    t.c32 = make_consts<S32>(sc, 0);
    t.colbest = (const u32 *)ws.colbest.p;
    t.lastrow = (const u32 *)ws.lastrow.p;
    t.scratch = (u32 *)ws.scratch.p;
    t.offsets = (int *)ws.offsets.p;
    t.n_ops = (u32 *)ws.n_ops.p;
    t.n_units = cd.plan.n_units;
    t.score3 = (int *)ws.score3.p;
    t.work_count = (u32 *)ws.worklist.p;
    t.worklist = (u32 *)ws.worklist.p + MGLSW_WC_WORDS;
    t.startcell = (int4 *)ws.startcell.p;
    t.no_rounds = cd.plan.n_units <= MGLSW_TRACE_ROUNDS_MIN_UNITS ? 1u : 0u;
    t.wstate[0] = (uint4 *)ws.wstate.p;
    t.wstate[1] = (uint4 *)ws.wstate.p + ((size_t)cd.plan.n_units * 2 + 1) * 3;
    if (score_only) {
        if ((e = launch_endpoints(t, pt)) != cudaSuccess) return cuda_fail(e, "endpoint launch");
        if ((e = cudaEventRecord(ws.ev_post, pt)) != cudaSuccess) return cuda_fail(e, "event record");
        bs.other_launches += 1;
        bs.chunks++;
        bs.cells += cd.plan.cells;
        bs.pairs += cd.n;
        bs.pairs_s16 += cd.plan.pairs_s16;
        bs.pairs_s32 += cd.plan.pairs_s32;
        return MGL_SW_OK;
    }
    if ((e = launch_start(t, pt)) != cudaSuccess) return cuda_fail(e, "start-cell launch");
    bool has_short = false, has_long = false;
    for (const Launch &L : cd.plan.launches) { if (L.G == MGLSW_G_LONG) has_long = true; else has_short = true; }
    if (t.no_rounds) { has_long = has_long || has_short; has_short = false; }
    if ((e = launch_traceback(t, has_short, has_long, pipe_units > 0, pt)) != cudaSuccess) return cuda_fail(e, "traceback launch");
    if (exact_shortcut) {
        if ((e = launch_exact_match(t, pt)) != cudaSuccess) return cuda_fail(e, "exact-match launch");
        bs.other_launches += 1;
    }
    if (timed) cudaEventRecord(tev_[3], pt);
    if (text) {
        /* characters per pair -> the same scan -> characters; the packed elements stay in the scratch buffer */
        if ((e = launch_text_len((const u32 *)ws.scratch.p, cd.d_ops_end, (const u32 *)ws.n_ops.p, (u32 *)ws.n_chars.p, cd.n, pt)) != cudaSuccess)
            return cuda_fail(e, "text length launch");
        if ((e = launch_scan_counts((const u32 *)ws.n_chars.p, (u32 *)ws.starts.p, (u32 *)ws.tiles.p, cd.n, pt)) != cudaSuccess) return cuda_fail(e, "scan launch");
        if ((e = launch_text_format((const u32 *)ws.scratch.p, cd.d_ops_end, (const u32 *)ws.n_ops.p, (const u32 *)ws.starts.p,
                                    (const u32 *)ws.tiles.p, (char *)ws.ops.p, cd.n, pt)) != cudaSuccess) return cuda_fail(e, "text format launch");
        bs.other_launches += 1;
    } else {
        if ((e = launch_scan_counts((const u32 *)ws.n_ops.p, (u32 *)ws.starts.p, (u32 *)ws.tiles.p, cd.n, pt)) != cudaSuccess) return cuda_fail(e, "scan launch");
        if ((e = launch_compact((const u32 *)ws.scratch.p, cd.d_ops_end, (const u32 *)ws.n_ops.p,
                                (const u32 *)ws.starts.p, (const u32 *)ws.tiles.p, (u32 *)ws.ops.p, cd.n, pt)) != cudaSuccess) return cuda_fail(e, "compact launch");
    }
    if (timed) cudaEventRecord(tev_[4], pt);
    if ((e = cudaEventRecord(ws.ev_post, pt)) != cudaSuccess) return cuda_fail(e, "event record");
    bs.other_launches += 5 + (has_short ? MGLSW_TRACE_ROUNDS : 0) + (has_long ? 1 : 0) - (has_short ? 0 : 1);
    bs.chunks++;
    bs.cells += cd.plan.cells;
    bs.pairs += cd.n;
    bs.trace_bytes += cd.plan.ckpt_words * 4;
    bs.pairs_s16 += cd.plan.pairs_s16;
    bs.pairs_s32 += cd.plan.pairs_s32;
    if (timed) {
        if ((e = cudaEventSynchronize(tev_[4])) != cudaSuccess) return cuda_fail(e, "chunk sync");
        float a = 0, b = 0, c = 0;
        cudaEventElapsedTime(&a, tev_[0], tev_[1]);
        cudaEventElapsedTime(&b, tev_[2], tev_[3]);
        cudaEventElapsedTime(&c, tev_[3], tev_[4]);
        bs.fill_ms += a; bs.trace_ms += b; bs.post_ms += c; bs.device_ms += a + b + c;
    }
    return MGL_SW_OK;
}

/* ---------------------------------------------------------------- streaming mode */

namespace {
struct InFlight { bool busy = false; int64_t first = 0; u32 n = 0; };
}

int Engine::align_batch(const BatchIn &in, BatchOut &out, BatchStats *stats)
{
    const auto fn_t0 = std::chrono::steady_clock::now();
    const bool tl_setup = std::getenv("MGL_SW_TIMELINE") != nullptr;
    auto setup_mark = [&](const char *what) {
        if (tl_setup) std::fprintf(stderr, "[mgl_sw] setup: %s at %.2f ms\n", what,
                                   std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - fn_t0).count());
    };
    BatchStats bs;
    std::memset(&bs, 0, sizeof bs);
    if (in.n < 0 || (in.n > 0 && (!in.arena || !in.ref_off || !in.alt_off || !in.ref_len || !in.alt_len)))
        return fail(MGL_SW_E_INVALID, "NULL batch array");
    const bool so = in.score_only, tx = in.text && !so;
    if (so ? !out.score3 : (!out.offsets || !out.cigar_start || (tx ? (!out.text && out.text_cap) : (!out.ops && out.ops_cap))))
        return fail(MGL_SW_E_INVALID, "NULL result array");
    const u64 out_cap = tx ? out.text_cap : out.ops_cap;       /* elements: characters of a text run, packed words otherwise */
    const size_t esz = tx ? 1 : sizeof(u32);
    if (check_scores(in.scores) != PLAN_OK) return fail(MGL_SW_E_SCORES, "score set out of range");
    cudaError_t e;
    if ((e = cudaSetDevice(device_)) != cudaSuccess) return cuda_fail(e, "cudaSetDevice");
    resident_scores_ = in.scores;
    out.ops_total = 0;
    if (!so && !out.keep_first_start) out.cigar_start[0] = 0;
    if (in.n == 0) { if (stats) *stats = bs; return MGL_SW_OK; }
    /* one parallel pass over the pair table: longest ref, total length, per-pair checkpoint estimate (units of 256 B).
       The engine's own sleeping helper threads (HelperPool), not OpenMP: an OpenMP team keeps spinning after the loop and
       takes the cores the planner threads need in the next few milliseconds (measured: first chunks planned 3x slower). */
    u32 max_tl = 0;
    u64 total_len = 0, total_est = 0;
    u64 span_lo = ~0ull, span_hi = 0;                 /* the byte range of the caller's arena this batch touches (wrapped offsets are refused by plan_range) */
    estpre_.resize((size_t)in.n);                     /* per slice: inclusive prefix sums of the per-pair checkpoint estimate (bytes) */
    const int T = in.n > 65536 ? (scan_threads > 0 ? scan_threads : (int)std::min<unsigned>(16u, std::max(1u, allowed_cpus() / 2))) : 1;
    std::vector<u64> slice_off((size_t)T + 1, 0);
    {
        std::vector<u32> tmax((size_t)T, 0);
        std::vector<u64> tsum((size_t)T, 0), tes((size_t)T, 0), tlo((size_t)T, ~0ull), thi((size_t)T, 0);
        auto scan = [&](int t) {
            const int64_t lo = in.n * t / T, hi = in.n * (t + 1) / T;
            u32 m = 0; u64 sum = 0, es = 0, alo = ~0ull, ahi = 0;
            u64 *pre = estpre_.data();
            for (int64_t p = lo; p < hi; p++) {
                const u32 tl = in.ref_len[p], ql = in.alt_len[p];
                const u64 ro = in.ref_off[p], ao = in.alt_off[p];
                m = std::max(m, tl);
                sum += (u64)tl + ql;
                es += pair_ckpt_bytes(tl, ql);
                pre[p] = es;
                alo = std::min(alo, std::min(ro, ao));
                ahi = std::max(ahi, std::max(ro + tl, ao + ql));
            }
            tmax[(size_t)t] = m; tsum[(size_t)t] = sum; tes[(size_t)t] = es; tlo[(size_t)t] = alo; thi[(size_t)t] = ahi;
        };
        pool_.run(T, scan);
        for (int t = 0; t < T; t++) {
            max_tl = std::max(max_tl, tmax[(size_t)t]); total_len += tsum[(size_t)t];
            slice_off[(size_t)t + 1] = slice_off[(size_t)t] + tes[(size_t)t];
            span_lo = std::min(span_lo, tlo[(size_t)t]); span_hi = std::max(span_hi, thi[(size_t)t]);
        }
        total_est = slice_off[(size_t)T];
        span_hi = std::min<u64>(span_hi, in.arena_bytes);
        span_lo = std::min(span_lo, span_hi) & ~63ull;
    }
    const u64 span = span_hi - span_lo;
    choose_slots(total_est, in.n);
    {
        cudaPointerAttributes pa;
        const cudaError_t pe = cudaPointerGetAttributes(&pa, in.arena);
        arena_pageable_ = pe != cudaSuccess || pa.type == cudaMemoryTypeUnregistered;
        if (pe != cudaSuccess) cudaGetLastError();
        /*
         * Result arrays in pageable memory (a JVM's buffers): a device-to-host copy into them is staged by the driver and
         * only returns when the chunk's kernels have finished -- the thread that launches the chunks would wait for every one
         * of them and nothing would overlap (config 2: 33.7 ms per 1 M pairs).  Such results come home into pinned buffers
         * of the work slot and are copied out by this thread when the chunk is collected.
         */
        auto pageable = [](const void *p) {
            if (!p) return false;
            cudaPointerAttributes a;
            const cudaError_t x = cudaPointerGetAttributes(&a, p);
            if (x != cudaSuccess) { cudaGetLastError(); return true; }
            return a.type == cudaMemoryTypeUnregistered;
        };
        res_pageable_ = pageable(so ? (const void *)out.score3 : (const void *)out.offsets);
        ops_pageable_ = !so && pageable(tx ? (const void *)out.text : (const void *)out.ops);
    }
    /* checkpoint estimate of pairs [0, x) */
    auto est_before = [&](int64_t x) -> u64 {
        if (x <= 0) return 0;
        int t = (int)(((x - 1) * T) / in.n);
        while (t + 1 < T && in.n * (t + 1) / T <= x - 1) t++;
        while (t > 0 && in.n * t / T > x - 1) t--;
        return estpre_[(size_t)(x - 1)] + slice_off[(size_t)t];
    };
    /* pairs the chunk starting at f0 takes: at most cap, and as many as keep the estimated checkpoints inside limit */
    auto take_at = [&](int64_t f0, u32 cap, u64 limit) -> int64_t {
        int64_t t = std::min<int64_t>(in.n - f0, cap);
        const u64 base = est_before(f0);
        if (est_before(f0 + t) - base <= limit) return t;
        int64_t lo = 1, hi = t;                           /* largest t with the sum inside the limit, at least one pair */
        while (lo < hi) { const int64_t mid = (lo + hi + 1) / 2; if (est_before(f0 + mid) - base <= limit) lo = mid; else hi = mid - 1; }
        return lo;
    };
    setup_mark("pair table scanned");
    { const int rc0 = reserve_arena(span_lo, span, max_tl); if (rc0 != MGL_SW_OK) return rc0; }
    setup_mark("arena reserved");

    /* arena upload policy: per-chunk byte ranges when the arena is laid out pair after pair,
       one up-front copy when pairs share or scatter their sequences */
    const bool whole = span > total_len + total_len / 2 + 4096 || total_len > span + span / 2;
    /*
     * 2-bit transport (sw_pack.h) only pays while it overlaps: a planner packs its chunk when it is at least
     * pack_lead chunks ahead of the one being uploaded -- the GPU or the host link is the slower side and the
     * thread would otherwise wait for a ring entry -- and sends it raw when the uploader is close behind
     * (planning is the slower side; packing would only add to it).  pack_arena 2 packs every chunk.
     */
    const bool pack_batch = pack_arena != 0 && !whole;
    std::atomic<int> upload_cursor{0};
    std::atomic<int> link_behind{0};                      /* 0 .. 4: how often, lately, a chunk's bytes were not on the device when its slot was free */
    if (whole) {
        if ((e = cudaMemcpyAsync((void *)(arena_dev() + span_lo), in.arena + span_lo, span, cudaMemcpyHostToDevice, slots_[0].stream)) != cudaSuccess)
            return cuda_fail(e, "H2D arena");
        if ((e = cudaStreamSynchronize(slots_[0].stream)) != cudaSuccess) return cuda_fail(e, "H2D arena");
        bs.h2d_bytes += span;
    }

    InFlight fl[MGLSW_SLOTS];
    const int NS = nslots_;
    u64 ops_running = 0;
    bool overflow = false;
    int rc = MGL_SW_OK;

    auto finish = [&](int si) -> int {
        WorkSlot &ws = slots_[si];
        InFlight &f = fl[si];
        if (!f.busy) return MGL_SW_OK;
        f.busy = false;
        cudaError_t ee;
        if ((ee = cudaEventSynchronize(ws.ev_counts)) != cudaSuccess) return cuda_fail(ee, "chunk execution");
        if (so) {                                         /* the three words per pair came home with ev_counts */
            if (res_pageable_) std::memcpy(out.score3 + 3 * f.first, ws.h_res.p, (size_t)f.n * 3 * sizeof(int));
            return MGL_SW_OK;
        }
        const u32 *cnt = (const u32 *)ws.h_n_ops.p;
        const u64 total = cnt[f.n];
        if (res_pageable_) std::memcpy(out.offsets + f.first, ws.h_res.p, (size_t)f.n * sizeof(int));   /* home since ev_counts */
        void *dst = nullptr;
        if (!overflow && ops_running + total <= out_cap) {
            if (total) {
                dst = tx ? (void *)(out.text + ops_running) : (void *)(out.ops + ops_running);
                if (ops_pageable_ && (ee = ws.h_ops.reserve(total * esz)) != cudaSuccess) return cuda_fail(ee, "pinned alloc");
                if ((ee = cudaMemcpyAsync(ops_pageable_ ? ws.h_ops.p : dst, ws.ops.p, total * esz, cudaMemcpyDeviceToHost, ws.post)) != cudaSuccess)
                    return cuda_fail(ee, "D2H ops");
                bs.d2h_bytes += total * esz;
            }
        } else overflow = true;
        cudaEventRecord(ws.ev_done, ws.post);
        cudaEventRecord(ws.ev_post, ws.post);
        u64 run = ops_running;
        for (u32 k = 0; k < f.n; k++) { run += cnt[k]; out.cigar_start[f.first + k + 1] = run; }
        ops_running = run;
        if ((ee = cudaEventSynchronize(ws.ev_done)) != cudaSuccess) return cuda_fail(ee, "D2H ops");
        if (ops_pageable_ && dst) std::memcpy(dst, ws.h_ops.p, total * esz);
        return MGL_SW_OK;
    };

    /*
     * Chunk boundaries first (a cheap scan), then the chunks are planned by a few host threads
     * running ahead of the GPU into a ring of staging entries, while this thread uploads,
     * launches and collects them in order.  Entry c % R is reused once chunk c - R has been
     * collected; R = planners + slots + 1 keeps every planner busy without blocking the collector.
     */
    /*
     * Who plans: device kernels (sw_plan_dev.cu) unless the caller asked for the host planner (whose threads then also do
     * the packing of the 2-bit transport; with device planning the staging thread packs).  With device planning ONE helper thread stages
     * each chunk's slice of the pair table and queues its planning kernels, as far ahead as the ring allows
     * (DEV_AHEAD chunks beyond the slots: under load a plan's ~30 small dependent kernels get SM space mostly when a
     * fill kernel drains, about one plan per chunk time, so the look-ahead must cover several chunks), and this
     * thread reads back a summary per chunk.
     */
    /* a batch of a few pairs (alignNative is a batch of one) is planned here, on the calling thread, in microseconds: the
       planning kernel's grid barriers and the summary's way back cost 0.17 ms however small the chunk */
    const bool dev_plan = host_plan == 0 && in.n > MGLSW_DEV_PLAN_MIN_PAIRS;
    const int DEV_AHEAD = 7;
    std::vector<std::pair<int64_t, int64_t>> bounds;
    int nthreads = plan_threads > 0 ? plan_threads : (int)std::min<unsigned>(8u, std::max(1u, std::thread::hardware_concurrency() / 2));
    if (dev_plan) nthreads = 2;                           /* the ramp below counts chunks in planners: two rounds of two */
    {
        const u64 limit = so ? ~0ull : budget();          /* a score-only run parks no checkpoints */
        /*
         * Ramp.  A batch of short pairs (a chunk is two or three kernels): the planners all start at once, so the
         * first round of chunks (one per planner) is a quarter of chunk_pairs, the second round half: the GPU gets
         * work after a quarter of a full plan and the full-size chunks are ready before it runs dry.  Longer or
         * mixed pairs spread a chunk over up to 66 kernels, and small chunks would cut those into slivers
         * (measured: config 3 e2e 1296 -> 1112 GCUPS): only the first three chunks are reduced (1/8, 1/4, 1/2).
         */
        const bool short_pairs = in.n >= 4 * (int64_t)chunk_pairs && total_est / (u64)in.n < 196608u;
        const u32 cap_full = chunk_cap(total_est, in.n);         /* the cap for THIS batch; the ramp stays in units of the option */
        for (int64_t f0 = 0; f0 < in.n;) {
            const size_t k = bounds.size();
            const u32 floor_ = std::min<u32>(chunk_pairs, 4096u);
            u32 cap = cap_full;
            if (short_pairs) {
                /* two small starters (1/16, 1/8) so the GPU has work within a fraction of a millisecond, then the ramp */
                if (k < 2) cap = std::max<u32>(chunk_pairs >> (4 - k), floor_);
                else cap = k < 2 + (size_t)nthreads ? std::max<u32>(chunk_pairs >> 2, floor_) : (k < 2 + 2 * (size_t)nthreads ? std::max<u32>(chunk_pairs >> 1, floor_) : chunk_pairs);
            } else if (k < 3) cap = std::max<u32>(chunk_pairs >> (3 - k), floor_);
            const int64_t t = take_at(f0, cap, limit);
            bounds.emplace_back(f0, t);
            f0 += t;
        }
    }
    const int nchunks = (int)bounds.size();
    nthreads = dev_plan ? 1 : std::min(nthreads, nchunks);
    setup_mark("chunk bounds");
    const int R = (dev_plan ? DEV_AHEAD : nthreads) + NS + 1;
    while ((int)ring_.size() < R) ring_.push_back(new ChunkDev());
    std::vector<int> ready((size_t)nchunks, 0);          /* 0 pending, 1 planned, 2 failed */
    std::vector<int> prc((size_t)nchunks, MGL_SW_OK);
    std::vector<std::string> perr((size_t)nchunks);
    std::mutex mu;
    std::condition_variable cv;
    int collected = 0;                                   /* chunks [0, collected) have released their ring entry */
    bool stop = false;
    std::atomic<int> next_chunk{0};
    BatchStats bs_stage;                                 /* what the staging thread of the device planner counts; merged at the end */
    std::memset(&bs_stage, 0, sizeof bs_stage);
    auto planner = [&]() {
        cudaSetDevice(device_);
        for (;;) {
            const int k = next_chunk.fetch_add(1);
            if (k >= nchunks) return;
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return stop || k - collected < R; });
                if (stop) return;
            }
            /*
             * Device planning, pack_arena 1: this thread packs chunk k when the link is what the step waits for AND the
             * packing hides behind it: the launching thread has lately found chunks' bytes still on their way when their
             * slot was free (link_behind; eight GPUs sharing the host's memory path get 23 GB/s each), and at least
             * MGLSW_PACK_BACKLOG of the uploads queued before chunk k have not finished, so the link has work while this
             * thread packs.  Never one of the first chunks, which the GPU is waiting for.  On a link that keeps up (one GPU:
             * 55 GB/s) the bytes are there long before their chunk's turn and nothing is packed.
             */
            bool want_pack = false;
            if (dev_plan && pack_batch) {
                if (pack_arena == 2) want_pack = true;
                else if (k >= MGLSW_PACK_FIRST_CHUNK && link_behind.load(std::memory_order_relaxed) >= 2) {
                    int backlog = 0;
                    for (int d = 1; d <= 2 * MGLSW_PACK_BACKLOG && k - d >= 0; d++) {
                        const ChunkDev &q = *ring_[(size_t)((k - d) % R)];
                        if (q.arena_ahead && q.ev_arena && cudaEventQuery(q.ev_arena) == cudaErrorNotReady) backlog++;
                    }
                    want_pack = backlog >= MGLSW_PACK_BACKLOG;
                }
            }
            int prc_k = dev_plan ? issue_plan_dev(in, bounds[(size_t)k].first, bounds[(size_t)k].second, *ring_[(size_t)(k % R)], bs_stage, perr[(size_t)k], !whole, want_pack)
                                 : plan_range(in, bounds[(size_t)k].first, bounds[(size_t)k].second, *ring_[(size_t)(k % R)], perr[(size_t)k]);
            if (!dev_plan && prc_k == MGL_SW_OK && pack_batch && (pack_arena == 2 || k - upload_cursor.load(std::memory_order_relaxed) >= pack_lead))
                prc_k = pack_chunk(in, *ring_[(size_t)(k % R)], perr[(size_t)k]);
            {
                std::lock_guard<std::mutex> lk(mu);
                prc[(size_t)k] = prc_k;
                ready[(size_t)k] = prc_k == MGL_SW_OK ? 1 : 2;
            }
            cv.notify_all();
        }
    };
    std::vector<std::thread> workers;
    if (nchunks > 1) for (int t = 0; t < nthreads; t++) workers.emplace_back(planner);
    auto release_upto = [&](int k) {                     /* chunks below k are collected */
        { std::lock_guard<std::mutex> lk(mu); if (k > collected) collected = k; }
        cv.notify_all();
    };
    const bool tl_on = std::getenv("MGL_SW_TIMELINE") != nullptr;     /* host timeline of the chunk pipeline on stderr */
    const auto tl_t0 = std::chrono::steady_clock::now();
    auto tl_ms = [&]() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tl_t0).count(); };
    setup_mark("planners started");
    if (tl_on) std::fprintf(stderr, "[mgl_sw] %d chunks, %d planner threads, setup took %.2f ms\n", nchunks, nthreads,
                            std::chrono::duration<double, std::milli>(tl_t0 - fn_t0).count());
    int c = 0;
    for (; c < nchunks && rc == MGL_SW_OK; c++) {
        const int si = c % NS;
        WorkSlot &ws = slots_[si];
        ChunkDev &cd = *ring_[(size_t)(c % R)];
        upload_cursor.store(c, std::memory_order_relaxed);
        if (nchunks == 1) {
            std::string err;
            rc = dev_plan ? issue_plan_dev(in, bounds[0].first, bounds[0].second, cd, bs_stage, err, !whole, pack_batch && pack_arena == 2) : plan_range(in, bounds[0].first, bounds[0].second, cd, err);
            if (rc != MGL_SW_OK) { fail(rc, err); break; }
        } else {
            std::unique_lock<std::mutex> lk(mu);
            cv.wait(lk, [&] { return ready[(size_t)c] != 0; });
            if (ready[(size_t)c] == 2) { rc = fail(prc[(size_t)c], perr[(size_t)c]); break; }
        }
        if (dev_plan && (rc = collect_plan_dev(cd)) != MGL_SW_OK) break;     /* the summary of this chunk's plan */
        const double t_ready = tl_ms();
        if ((rc = finish(si)) != MGL_SW_OK) break;              /* chunk c - NS: already home */
        release_upto(c - NS + 1);
        const double t_slot = tl_ms();
        const int64_t first = cd.first;
        if ((rc = ensure_slot(ws, cd.plan, cd.n, so, tx)) != MGL_SW_OK) break;
        if (dev_plan) {
            /* the staging thread sent the chunk's arena bytes ahead; its fill launches wait for them */
            if (cd.arena_ahead) {
                const int lb = link_behind.load(std::memory_order_relaxed);
                link_behind.store(cudaEventQuery(cd.ev_arena) == cudaErrorNotReady ? std::min(lb + 1, 4) : std::max(lb - 1, 0), std::memory_order_relaxed);
                if ((e = cudaStreamWaitEvent(ws.stream, cd.ev_arena, 0)) != cudaSuccess) { rc = cuda_fail(e, "stream wait"); break; }
            }
        } else if ((rc = upload_chunk(in, cd, ws.stream, !whole, bs)) != MGL_SW_OK) break;
        if ((rc = run_chunk(cd, ws, bs, false, so, tx)) != MGL_SW_OK) break;
        const u32 n = cd.n;
        if (so) {
            if (res_pageable_ && (e = ws.h_res.reserve((size_t)n * 3 * sizeof(int))) != cudaSuccess) { rc = cuda_fail(e, "pinned alloc"); break; }
            if ((e = cudaMemcpyAsync(res_pageable_ ? ws.h_res.p : (void *)(out.score3 + 3 * first), ws.score3.p, (size_t)n * 3 * sizeof(int), cudaMemcpyDeviceToHost, ws.post)) != cudaSuccess) { rc = cuda_fail(e, "D2H scores"); break; }
            bs.d2h_bytes += (size_t)n * 12;
            cudaEventRecord(ws.ev_counts, ws.post);
            cudaEventRecord(ws.ev_post, ws.post);
            fl[si].busy = true; fl[si].first = first; fl[si].n = n;
            if ((rc = finish((c + 1) % NS)) != MGL_SW_OK) break;
            release_upto(c - NS + 2);
            continue;
        }
        if (res_pageable_ && (e = ws.h_res.reserve((size_t)n * sizeof(int))) != cudaSuccess) { rc = cuda_fail(e, "pinned alloc"); break; }
        if ((e = cudaMemcpyAsync(res_pageable_ ? ws.h_res.p : (void *)(out.offsets + first), ws.offsets.p, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, ws.post)) != cudaSuccess) { rc = cuda_fail(e, "D2H offsets"); break; }
        if ((e = cudaMemcpyAsync(ws.h_n_ops.p, tx ? ws.n_chars.p : ws.n_ops.p, (size_t)n * sizeof(u32), cudaMemcpyDeviceToHost, ws.post)) != cudaSuccess) { rc = cuda_fail(e, "D2H counts"); break; }
        if ((e = cudaMemcpyAsync((u32 *)ws.h_n_ops.p + n, (u32 *)ws.starts.p + n, sizeof(u32), cudaMemcpyDeviceToHost, ws.post)) != cudaSuccess) { rc = cuda_fail(e, "D2H total"); break; }
        bs.d2h_bytes += (size_t)n * 8 + 4;
        cudaEventRecord(ws.ev_counts, ws.post);
        cudaEventRecord(ws.ev_post, ws.post);
        fl[si].busy = true; fl[si].first = first; fl[si].n = n;
        const double t_launched = tl_ms();
        /* while the newest chunks run, bring the oldest one home: its slot is the next chunk's */
        if ((rc = finish((c + 1) % NS)) != MGL_SW_OK) break;
        release_upto(c - NS + 2);
        if (tl_on) std::fprintf(stderr, "[mgl_sw] chunk %d n=%u: planned by %.2f ms, slot free %.2f, launched %.2f, previous chunk home %.2f\n",
                                c, n, t_ready, t_slot, t_launched, tl_ms());
    }
    { std::lock_guard<std::mutex> lk(mu); stop = true; }
    cv.notify_all();
    for (std::thread &t : workers) t.join();
    for (int k = 0; k < NS && rc == MGL_SW_OK; k++) rc = finish((c + k) % NS);   /* oldest first */
    if (rc != MGL_SW_OK) {
        cudaDeviceSynchronize();
        cudaGetLastError();
        return rc;
    }
    out.ops_total = ops_running;
    bs.h2d_bytes += bs_stage.h2d_bytes; bs.other_launches += bs_stage.other_launches;
    if (stats) *stats = bs;
    if (tl_on) std::fprintf(stderr, "[mgl_sw] batch done at %.2f ms\n", tl_ms());
    if (overflow) return fail(MGL_SW_E_OPS_CAPACITY, "ops buffer too small: need " + std::to_string(ops_running));
    return MGL_SW_OK;
}

/* ---------------------------------------------------------------- device-resident mode */

void Engine::drop_resident()
{
    for (ChunkDev *cd : resident_) { cd->release(); delete cd; }
    resident_.clear();
    for (DevBuf &b : resident_results_) b.release();
    resident_results_.clear();
}

int Engine::prepare_resident(const BatchIn &in)
{
    cudaError_t e;
    if ((e = cudaSetDevice(device_)) != cudaSuccess) return cuda_fail(e, "cudaSetDevice");
    if (in.n <= 0 || !in.arena) return fail(MGL_SW_E_INVALID, "empty batch");
    if (check_scores(in.scores) != PLAN_OK) return fail(MGL_SW_E_SCORES, "score set out of range");
    drop_resident();
    resident_scores_ = in.scores;
    { const int rc0 = reserve_arena(in); if (rc0 != MGL_SW_OK) return rc0; }
    if ((e = cudaMemcpy((void *)arena_dev(), in.arena, in.arena_bytes, cudaMemcpyHostToDevice)) != cudaSuccess) return cuda_fail(e, "H2D arena");
    BatchStats bs;
    std::memset(&bs, 0, sizeof bs);
    {
        u64 total_est = 0;
        for (int64_t k = 0; k < in.n; k++) total_est += pair_ckpt_bytes(in.ref_len[k], in.alt_len[k]);
        choose_slots(total_est, in.n);
        chunk_cap_ = chunk_cap(total_est, in.n);
    }
    int64_t first = 0;
    while (first < in.n) {
        ChunkDev *cd = new ChunkDev();
        resident_.push_back(cd);
        int rc = plan_next(in, first, *cd);
        if (rc != MGL_SW_OK) { drop_resident(); return rc; }
        rc = upload_chunk(in, *cd, slots_[0].stream, false, bs);
        if (rc != MGL_SW_OK) { drop_resident(); return rc; }
        first += cd->n;
    }
    if ((e = cudaStreamSynchronize(slots_[0].stream)) != cudaSuccess) return cuda_fail(e, "H2D descriptors");
    for (ChunkDev *cd : resident_) {        /* pinned staging no longer needed */
        cd->h_units.release(); cd->h_pair_unit.release(); cd->h_ops_end.release(); cd->h_strategy.release();
    }
    return MGL_SW_OK;
}

int Engine::run_resident(BatchStats *stats)
{
    cudaError_t e;
    if ((e = cudaSetDevice(device_)) != cudaSuccess) return cuda_fail(e, "cudaSetDevice");
    if (resident_.empty()) return fail(MGL_SW_E_INVALID, "no resident batch");
    BatchStats bs;
    std::memset(&bs, 0, sizeof bs);
    if (!overlap) {
        /* one chunk at a time on slot 0, every phase bracketed by events: per-kernel times */
        WorkSlot &ws = slots_[0];
        for (ChunkDev *cd : resident_) {
            int rc = ensure_slot(ws, cd->plan, cd->n);
            if (rc != MGL_SW_OK) return rc;
            rc = run_chunk(*cd, ws, bs, true);
            if (rc != MGL_SW_OK) return rc;
        }
        if (stats) *stats = bs;
        return MGL_SW_OK;
    }
    /*
     * Whole-batch throughput: chunks rotate through the work slots, so the traceback of chunk c
     * (post stream) runs while the next chunks fill.  One start event gates every fill stream; the
     * time is start -> the latest of the post streams' last events.
     */
    for (size_t c = 0; c < resident_.size(); c++) {        /* grow each slot to its largest chunk before timing */
        int rc = ensure_slot(slots_[c % nslots_], resident_[c]->plan, resident_[c]->n);
        if (rc != MGL_SW_OK) return rc;
    }
    if ((e = cudaEventRecord(tev_[0], slots_[0].stream)) != cudaSuccess) return cuda_fail(e, "event record");
    for (int k = 1; k < MGLSW_SLOTS; k++)
        if ((e = cudaStreamWaitEvent(slots_[k].stream, tev_[0], 0)) != cudaSuccess) return cuda_fail(e, "stream wait");
    for (size_t c = 0; c < resident_.size(); c++) {
        int rc = run_chunk(*resident_[c], slots_[c % nslots_], bs, false);
        if (rc != MGL_SW_OK) return rc;
    }
    float a = 0;
    for (int k = 0; k < MGLSW_SLOTS; k++) cudaEventRecord(tev_[5 + k], slots_[k].post);
    for (int k = 0; k < MGLSW_SLOTS; k++) {
        if ((e = cudaEventSynchronize(tev_[5 + k])) != cudaSuccess) return cuda_fail(e, "batch execution");
        float b = 0;
        cudaEventElapsedTime(&b, tev_[0], tev_[5 + k]);
        a = b > a ? b : a;
    }
    const float b = a;
    bs.device_ms = a > b ? a : b;
    if (stats) *stats = bs;
    return MGL_SW_OK;
}

int Engine::fetch_resident(BatchOut &out)
{
    cudaError_t e;
    if ((e = cudaSetDevice(device_)) != cudaSuccess) return cuda_fail(e, "cudaSetDevice");
    if (resident_.empty()) return fail(MGL_SW_E_INVALID, "no resident batch");
    BatchStats bs;
    std::memset(&bs, 0, sizeof bs);
    WorkSlot &ws = slots_[0];
    u64 run = 0;
    bool overflow = false;
    out.cigar_start[0] = 0;
    std::vector<u32> cnt;
    for (ChunkDev *cd : resident_) {
        int rc = ensure_slot(ws, cd->plan, cd->n);
        if (rc != MGL_SW_OK) return rc;
        if ((rc = run_chunk(*cd, ws, bs, false)) != MGL_SW_OK) return rc;
        cnt.resize((size_t)cd->n + 1);
        if ((e = cudaMemcpyAsync(out.offsets + cd->first, ws.offsets.p, (size_t)cd->n * sizeof(int), cudaMemcpyDeviceToHost, ws.post)) != cudaSuccess) return cuda_fail(e, "D2H");
        if ((e = cudaMemcpyAsync(cnt.data(), ws.n_ops.p, (size_t)cd->n * sizeof(u32), cudaMemcpyDeviceToHost, ws.post)) != cudaSuccess) return cuda_fail(e, "D2H");
        if ((e = cudaMemcpyAsync(cnt.data() + cd->n, (u32 *)ws.starts.p + cd->n, sizeof(u32), cudaMemcpyDeviceToHost, ws.post)) != cudaSuccess) return cuda_fail(e, "D2H");
        if ((e = cudaStreamSynchronize(ws.post)) != cudaSuccess) return cuda_fail(e, "chunk execution");
        const u64 total = cnt[cd->n];
        if (!overflow && run + total <= out.ops_cap) {
            if (total && (e = cudaMemcpy(out.ops + run, ws.ops.p, total * sizeof(u32), cudaMemcpyDeviceToHost)) != cudaSuccess) return cuda_fail(e, "D2H ops");
        } else overflow = true;
        for (u32 k = 0; k < cd->n; k++) { run += cnt[k]; out.cigar_start[cd->first + k + 1] = run; }
    }
    out.ops_total = run;
    if (overflow) return fail(MGL_SW_E_OPS_CAPACITY, "ops buffer too small: need " + std::to_string(run));
    return MGL_SW_OK;
}

}  // namespace mglsw
