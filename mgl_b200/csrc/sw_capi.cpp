/*
 * sw_capi.cpp -- the extern "C" surface declared in include/mgl_sw.h.
 * Thin: argument checks, locking, and translation to the engine.
 */
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <condition_variable>
#include <mutex>
#include <new>
#include <algorithm>
#include <string>
#include <thread>
#include <vector>

#include "../../include/mgl_sw.h"
#include "sw_engine.h"
#include "sw_kernels.h"
#include "sw_pack.h"

using namespace mglsw;

/*
 * One handle = one engine per GPU of its device list (SURVEY.md 8b "handle create/destroy (device list, streams)").
 * `engine` is engines[0]: the resident_* entries and a one-device handle use only it.
 */
struct mgl_sw_handle {
    Engine *engine = nullptr;
    std::vector<Engine *> engines;
    std::vector<PinBuf> shard_out;     /* multi-device calls: per shard, pinned staging for its CIGAR elements / characters */
    std::mutex mu;
    std::string err;
    bool err_is_own = false;           /* last error came from the sharding layer, not from an engine */
};

namespace {

int to_batch(const mgl_sw_batch *b, BatchIn &in)
{
    if (!b || b->n < 0) return MGL_SW_E_INVALID;
    in.n = b->n; in.arena = b->arena; in.arena_bytes = b->arena_bytes;
    in.ref_off = b->ref_off; in.alt_off = b->alt_off; in.ref_len = b->ref_len; in.alt_len = b->alt_len;
    in.strategies = b->strategies; in.strategy = b->strategy;
    in.scores = normalise_scores(b->match, b->mismatch, b->gap_open, b->gap_extend);
    return MGL_SW_OK;
}

void to_stats(const BatchStats &s, mgl_sw_stats *o)
{
    if (!o) return;
    o->cells = s.cells; o->pairs = s.pairs; o->pairs_s16 = s.pairs_s16; o->pairs_s32 = s.pairs_s32;
    o->fill_launches = s.fill_launches; o->other_launches = s.other_launches; o->chunks = s.chunks;
    o->trace_bytes = s.trace_bytes; o->h2d_bytes = s.h2d_bytes; o->d2h_bytes = s.d2h_bytes;
    o->fill_ms = s.fill_ms; o->trace_ms = s.trace_ms; o->post_ms = s.post_ms; o->device_ms = s.device_ms;
}

/*
 * The process-wide handle behind the entry points that have no handle argument: the reference's one-pair JNI call
 * (mgl_sw_align_one) and the batched JNI symbol.  Created by the first caller over every visible GPU -- or the
 * comma-separated list in MGL_SW_DEVICES, or the single device MGL_SW_DEVICE names -- and destroyed by
 * mgl_sw_shutdown (doneNative), which waits for the calls in flight; the next caller creates it again.
 */
std::mutex g_def_mu;
std::condition_variable g_def_cv;
mgl_sw_handle *g_default = nullptr;
int g_def_users = 0;
bool g_def_closing = false;

int make_default(mgl_sw_handle **out)
{
    std::vector<int> devs;
    if (const char *list = std::getenv("MGL_SW_DEVICES")) {
        for (const char *c = list; *c;) {
            char *end = nullptr;
            const long v = std::strtol(c, &end, 10);
            if (end == c) break;
            devs.push_back((int)v);
            c = *end == ',' ? end + 1 : end;
        }
    } else if (const char *one = std::getenv("MGL_SW_DEVICE")) {
        devs.push_back(std::atoi(one));
    } else {
        const int n = mgl_sw_device_count();
        for (int d = 0; d < n; d++) devs.push_back(d);
    }
    if (devs.empty()) return MGL_SW_E_NO_DEVICE;
    return mgl_sw_create_multi(devs.data(), (int)devs.size(), out);
}

/* sub-batch [lo, hi) of a batch: the same arena, the pair arrays offset */
BatchIn shard_of(const BatchIn &in, int64_t lo, int64_t hi)
{
    BatchIn s = in;
    s.n = hi - lo;
    s.ref_off = in.ref_off + lo; s.alt_off = in.alt_off + lo; s.ref_len = in.ref_len + lo; s.alt_len = in.alt_len + lo;
    if (in.strategies) s.strategies = in.strategies + lo;
    return s;
}

/*
 * Contiguous cut of [0, n) into `parts` ranges of near-equal DP cells (SURVEY.md 8e "partition on the host by
 * cumulative cell count"): cells are summed per slice on a few short-lived threads, then each boundary is found
 * inside the slice that holds it.  bounds has parts + 1 entries.
 */
void cut_by_cells(const BatchIn &in, int parts, std::vector<int64_t> &bounds)
{
    bounds.assign((size_t)parts + 1, in.n);
    bounds[0] = 0;
    if (parts <= 1 || in.n == 0) return;
    const int T = in.n > 65536 ? (int)std::min<unsigned>(8u, std::max(1u, std::thread::hardware_concurrency() / 2)) : 1;
    std::vector<u64> part((size_t)T, 0);
    auto sum = [&](int t) {
        u64 c = 0;
        for (int64_t p = in.n * t / T, e = in.n * (t + 1) / T; p < e; p++) c += (u64)in.ref_len[p] * in.alt_len[p];
        part[(size_t)t] = c;
    };
    {
        std::vector<std::thread> th;
        for (int t = 1; t < T; t++) th.emplace_back(sum, t);
        sum(0);
        for (std::thread &x : th) x.join();
    }
    u64 total = 0;
    for (u64 c : part) total += c;
    int t = 0;
    u64 before = 0;                                   /* cells of slices [0, t) */
    for (int k = 1; k < parts; k++) {
        const u64 target = (u64)((long double)total * k / parts);
        while (t + 1 < T && before + part[(size_t)t] < target) { before += part[(size_t)t]; t++; }
        u64 run = before;
        int64_t p = in.n * t / T;
        const int64_t e = in.n * (t + 1) / T;
        for (; p < e && run < target; p++) run += (u64)in.ref_len[p] * in.alt_len[p];
        bounds[(size_t)k] = std::max(p, bounds[(size_t)k - 1]);
    }
}

/*
 * One batch over every engine of the handle: shard s = pairs [bounds[s], bounds[s + 1]) runs on engine s on its own
 * host thread; nothing is exchanged between GPUs (pairs are independent: no collective).  Offsets and scores land
 * in the caller's arrays directly; CIGAR elements (or characters) of a shard land in the shard's pinned staging
 * buffer -- where shard s's elements start in the caller's array is known only when shards 0 .. s-1 are complete --
 * and are gathered in input order at the end, cigar_start rebased shard by shard.
 */
int run_multi(mgl_sw_handle *h, const BatchIn &in, BatchOut &out, BatchStats &total)
{
    const int S = (int)h->engines.size();
    const bool so = in.score_only, tx = in.text && !so;
    const size_t esz = tx ? 1 : sizeof(u32);
    const u64 cap = tx ? out.text_cap : out.ops_cap;
    h->err_is_own = false;
    if (in.n < 0 || (in.n > 0 && (!in.arena || !in.ref_off || !in.alt_off || !in.ref_len || !in.alt_len))) { h->err = "NULL batch array"; h->err_is_own = true; return MGL_SW_E_INVALID; }
    if (so ? !out.score3 : (!out.offsets || !out.cigar_start)) { h->err = "NULL result array"; h->err_is_own = true; return MGL_SW_E_INVALID; }
    std::vector<int64_t> bounds;
    cut_by_cells(in, S, bounds);
    std::vector<BatchStats> st((size_t)S);
    std::vector<int> rc((size_t)S, MGL_SW_OK);
    std::vector<u64> produced((size_t)S, 0);
    if ((int)h->shard_out.size() < S) h->shard_out.resize((size_t)S);
    u64 total_len = 0;
    if (!so) { out.cigar_start[0] = 0; }
    auto work = [&](int s) {
        std::memset(&st[(size_t)s], 0, sizeof(BatchStats));
        const int64_t lo = bounds[(size_t)s], hi = bounds[(size_t)s + 1];
        if (hi <= lo) return;
        const BatchIn sub = shard_of(in, lo, hi);
        BatchOut o;
        if (so) { o.offsets = nullptr; o.cigar_start = nullptr; o.ops = nullptr; o.ops_cap = 0; o.score3 = out.score3 + 3 * lo; }
        else {
            /* staging capacity: twice the shard's share of the caller's capacity; a shard that needs more (and less
               than the caller's whole capacity) is run again with exactly what it asked for */
            u64 want = std::min<u64>(cap, std::max<u64>(4096, 2 * (u64)((long double)cap * (hi - lo) / in.n) + 1024));
            const bool direct = lo == 0;               /* the first shard's elements start at 0: straight into the caller's array */
            if (direct) want = cap;
            for (int attempt = 0; attempt < 2; attempt++) {
                PinBuf &pb = h->shard_out[(size_t)s];
                cudaSetDevice(h->engines[(size_t)s]->device());
                if (!direct && pb.reserve((size_t)want * esz + 16) != cudaSuccess) { rc[(size_t)s] = MGL_SW_E_CUDA; return; }
                void *dst = direct ? (tx ? (void *)out.text : (void *)out.ops) : pb.p;
                o.offsets = out.offsets + lo;
                o.cigar_start = out.cigar_start + lo;
                o.keep_first_start = lo > 0;           /* element lo belongs to the previous shard's last pair */
                o.ops = tx ? nullptr : (u32 *)dst; o.ops_cap = tx ? 0 : want;
                o.text = tx ? (char *)dst : nullptr; o.text_cap = tx ? want : 0;
                rc[(size_t)s] = h->engines[(size_t)s]->align_batch(sub, o, &st[(size_t)s]);
                produced[(size_t)s] = o.ops_total;
                if (rc[(size_t)s] != MGL_SW_E_OPS_CAPACITY || o.ops_total > cap || attempt == 1) break;
                want = o.ops_total;
            }
            return;
        }
        rc[(size_t)s] = h->engines[(size_t)s]->align_batch(sub, o, &st[(size_t)s]);
    };
    {
        std::vector<std::thread> th;
        for (int s = 1; s < S; s++) th.emplace_back(work, s);
        work(0);
        for (std::thread &x : th) x.join();
    }
    std::memset(&total, 0, sizeof total);
    int first_rc = MGL_SW_OK;
    for (int s = 0; s < S; s++) {
        const BatchStats &b = st[(size_t)s];
        total.cells += b.cells; total.pairs += b.pairs; total.pairs_s16 += b.pairs_s16; total.pairs_s32 += b.pairs_s32;
        total.fill_launches += b.fill_launches; total.other_launches += b.other_launches; total.chunks += b.chunks;
        total.trace_bytes += b.trace_bytes; total.h2d_bytes += b.h2d_bytes; total.d2h_bytes += b.d2h_bytes;
        if (rc[(size_t)s] != MGL_SW_OK && rc[(size_t)s] != MGL_SW_E_OPS_CAPACITY && first_rc == MGL_SW_OK) {
            first_rc = rc[(size_t)s];
            h->err = "shard " + std::to_string(s) + " (device " + std::to_string(h->engines[(size_t)s]->device()) + "): " + h->engines[(size_t)s]->last_error();
            h->err_is_own = true;
        }
        total_len += produced[(size_t)s];
    }
    (void)total_len;
    if (first_rc != MGL_SW_OK) return first_rc;
    if (so) return MGL_SW_OK;
    /* gather in input order: shard s's elements start where shards 0 .. s-1 end */
    std::vector<u64> base((size_t)S + 1, 0);
    bool overflow = false;
    for (int s = 0; s < S; s++) { base[(size_t)s + 1] = base[(size_t)s] + produced[(size_t)s]; overflow = overflow || rc[(size_t)s] == MGL_SW_E_OPS_CAPACITY; }
    out.ops_total = base[(size_t)S];
    if (overflow || out.ops_total > cap) { h->err = "ops buffer too small: need " + std::to_string(out.ops_total); h->err_is_own = true; return MGL_SW_E_OPS_CAPACITY; }
    /* every shard's copy and rebase in P pieces, all pieces at once: this is the one serial tail of a multi-device call
       (18 MB of elements per 1 M config-2 pairs and shard; one thread per shard took 3 ms of a 20 ms call) */
    const int P = std::max(1, std::min(8, (int)std::thread::hardware_concurrency() / (2 * S)));
    auto gather = [&](int item) {
        const int s = item / P, q = item % P;
        const int64_t lo = bounds[(size_t)s], hi = bounds[(size_t)s + 1];
        if (hi <= lo) return;
        char *dst = tx ? out.text : (char *)out.ops;
        if (lo > 0) {
            const u64 nbytes = produced[(size_t)s] * esz, a = nbytes * (u64)q / (u64)P, b = nbytes * (u64)(q + 1) / (u64)P;
            if (b > a) std::memcpy(dst + base[(size_t)s] * esz + a, (const char *)h->shard_out[(size_t)s].p + a, (size_t)(b - a));
        }
        if (s > 0) {
            const u64 add = base[(size_t)s];
            const int64_t n = hi - lo, pa = lo + 1 + n * q / P, pb = lo + 1 + n * (q + 1) / P;
            for (int64_t p = pa; p < pb; p++) out.cigar_start[p] += add;
        }
    };
    {
        std::vector<std::thread> th;
        for (int item = 1; item < S * P; item++) th.emplace_back(gather, item);
        gather(0);
        for (std::thread &x : th) x.join();
    }
    return MGL_SW_OK;
}

/* the call on a handle: one engine directly, several through run_multi */
int run_batch(mgl_sw_handle *h, const BatchIn &in, BatchOut &out, BatchStats &bs)
{
    if (h->engines.size() <= 1) { h->err_is_own = false; return h->engine->align_batch(in, out, &bs); }
    return run_multi(h, in, out, bs);
}

}  // namespace

extern "C" {

int mgl_sw_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

const char *mgl_sw_version(void) { return "mgl_sw_b200 0.1 (sm_100a)"; }

int mgl_sw_create(int device, mgl_sw_handle **out) { return mgl_sw_create_multi(&device, 1, out); }

int mgl_sw_create_multi(const int *devices, int n_devices, mgl_sw_handle **out)
{
    if (!out) return MGL_SW_E_INVALID;
    *out = nullptr;
    if (!devices || n_devices < 1 || n_devices > 64) return MGL_SW_E_INVALID;
    const int n = mgl_sw_device_count();
    if (n <= 0) return MGL_SW_E_NO_DEVICE;
    for (int k = 0; k < n_devices; k++) {
        const int device = devices[k];
        if (device < 0 || device >= n) return MGL_SW_E_NO_DEVICE;
        cudaDeviceProp prop;
        if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return MGL_SW_E_NO_DEVICE;
        if (prop.major != 10) {       /* the kernels are sm_100a only: refuse rather than fall back */
            std::fprintf(stderr, "mgl_sw: device %d is sm_%d%d, this library is built for sm_100a only\n", device, prop.major, prop.minor);
            return MGL_SW_E_NO_DEVICE;
        }
    }
    mgl_sw_handle *h = new (std::nothrow) mgl_sw_handle();
    if (!h) return MGL_SW_E_INVALID;
    for (int k = 0; k < n_devices; k++) h->engines.push_back(new Engine(devices[k]));
    h->engine = h->engines[0];
    if (n_devices > 1) {
        /* the engines share the host's cores: each plans with its share of them */
        const int per = (int)std::max(1u, std::min(8u, std::thread::hardware_concurrency() / 2u / (unsigned)n_devices));
        for (Engine *e : h->engines) { e->plan_threads = per; e->copy_threads = std::min(4, std::max(1, per)); e->scan_threads = std::max(per, std::min(4, (int)(std::thread::hardware_concurrency() / (unsigned)n_devices))); }
    }
    *out = h;
    return MGL_SW_OK;
}

int mgl_sw_handle_devices(mgl_sw_handle *h, int *devices, int cap)
{
    if (!h) return MGL_SW_E_INVALID;
    const int n = (int)h->engines.size();
    for (int k = 0; k < n && k < cap && devices; k++) devices[k] = h->engines[(size_t)k]->device();
    return n;
}

void mgl_sw_destroy(mgl_sw_handle *h)
{
    if (!h) return;
    for (size_t k = 0; k < h->shard_out.size(); k++) {
        if (k < h->engines.size()) cudaSetDevice(h->engines[k]->device());
        h->shard_out[k].release();
    }
    for (Engine *e : h->engines) delete e;
    delete h;
}

const char *mgl_sw_last_error(mgl_sw_handle *h)
{
    if (!h) return "no handle";
    return h->err_is_own ? h->err.c_str() : h->engine->last_error().c_str();
}

int mgl_sw_default_handle(mgl_sw_handle **out)
{
    if (!out) return MGL_SW_E_INVALID;
    std::unique_lock<std::mutex> lk(g_def_mu);
    g_def_cv.wait(lk, [] { return !g_def_closing; });
    if (!g_default) {
        const int rc = make_default(&g_default);
        if (rc != MGL_SW_OK) { g_default = nullptr; *out = nullptr; return rc; }
    }
    g_def_users++;
    *out = g_default;
    return MGL_SW_OK;
}

void mgl_sw_default_release(void)
{
    std::lock_guard<std::mutex> lk(g_def_mu);
    if (g_def_users > 0) g_def_users--;
    g_def_cv.notify_all();
}

void mgl_sw_shutdown(void)
{
    std::unique_lock<std::mutex> lk(g_def_mu);
    if (!g_default) return;
    g_def_closing = true;
    g_def_cv.wait(lk, [] { return g_def_users == 0; });
    mgl_sw_handle *h = g_default;
    g_default = nullptr;
    g_def_closing = false;
    lk.unlock();
    g_def_cv.notify_all();
    mgl_sw_destroy(h);
}

int mgl_sw_set_option(mgl_sw_handle *h, const char *name, int64_t value)
{
    if (!h || !name) return MGL_SW_E_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    for (Engine *e : h->engines) {
        if (!std::strcmp(name, "allow_s16")) e->allow_s16 = value != 0;
        else if (!std::strcmp(name, "overlap")) e->overlap = value != 0;
        else if (!std::strcmp(name, "plan_threads")) { if (value < 0 || value > 64) return MGL_SW_E_INVALID; e->plan_threads = (int)value; }
        else if (!std::strcmp(name, "chunk_pairs")) { if (value < 1 || value > (1 << 22)) return MGL_SW_E_INVALID; e->chunk_pairs = (u32)value; }
        else if (!std::strcmp(name, "pipe_wide")) { if (value < 0 || value > 2) return MGL_SW_E_INVALID; e->pipe_wide = (int)value; }
        else if (!std::strcmp(name, "pack_arena")) { if (value < 0 || value > 2) return MGL_SW_E_INVALID; e->pack_arena = (int)value; }
        else if (!std::strcmp(name, "pack_threads")) { if (value < 0 || value > 32) return MGL_SW_E_INVALID; e->pack_threads = (int)value; }
        else if (!std::strcmp(name, "pack_lead")) { if (value < 0 || value > 1024) return MGL_SW_E_INVALID; e->pack_lead = (int)value; }
        else if (!std::strcmp(name, "copy_threads")) { if (value < 0 || value > 32) return MGL_SW_E_INVALID; e->copy_threads = (int)value; }
        else if (!std::strcmp(name, "exact_match_shortcut")) e->exact_shortcut = value != 0;
        else if (!std::strcmp(name, "slots")) { if (value != 0 && (value < 2 || value > MGLSW_SLOTS)) return MGL_SW_E_INVALID; e->slots_opt = (int)value; }
        else if (!std::strcmp(name, "host_plan")) { if (value < 0 || value > 1) return MGL_SW_E_INVALID; e->host_plan = (int)value; }
        else if (!std::strcmp(name, "trace_budget_mb")) { if (value < 0) return MGL_SW_E_INVALID; e->trace_budget_bytes = (u64)value << 20; }
        else return MGL_SW_E_INVALID;
    }
    return MGL_SW_OK;
}

int mgl_sw_align_batch(mgl_sw_handle *h, const mgl_sw_batch *batch, mgl_sw_result *result, mgl_sw_stats *stats)
{
    if (!h || !result) return MGL_SW_E_INVALID;
    BatchIn in;
    int rc = to_batch(batch, in);
    if (rc != MGL_SW_OK) return rc;
    BatchOut out;
    out.offsets = result->offsets; out.cigar_start = result->cigar_start; out.ops = result->ops;
    out.ops_cap = result->ops_cap; out.ops_total = 0;
    BatchStats bs;
    std::memset(&bs, 0, sizeof bs);
    std::lock_guard<std::mutex> lk(h->mu);
    rc = run_batch(h, in, out, bs);
    result->ops_total = out.ops_total;
    to_stats(bs, stats);
    return rc;
}

int mgl_sw_align_batch_text(mgl_sw_handle *h, const mgl_sw_batch *batch, mgl_sw_text_result *result, mgl_sw_stats *stats)
{
    if (!h || !result) return MGL_SW_E_INVALID;
    BatchIn in;
    int rc = to_batch(batch, in);
    if (rc != MGL_SW_OK) return rc;
    in.text = true;
    BatchOut out;
    out.offsets = result->offsets; out.cigar_start = result->text_start; out.ops = nullptr; out.ops_cap = 0; out.ops_total = 0;
    out.text = result->text; out.text_cap = result->text_cap;
    BatchStats bs;
    std::memset(&bs, 0, sizeof bs);
    std::lock_guard<std::mutex> lk(h->mu);
    rc = run_batch(h, in, out, bs);
    result->text_total = out.ops_total;
    to_stats(bs, stats);
    return rc;
}

int mgl_sw_score_batch(mgl_sw_handle *h, const mgl_sw_batch *batch, int32_t *score3, mgl_sw_stats *stats)
{
    if (!h || !score3) return MGL_SW_E_INVALID;
    BatchIn in;
    int rc = to_batch(batch, in);
    if (rc != MGL_SW_OK) return rc;
    in.score_only = true;
    BatchOut out;
    out.offsets = nullptr; out.cigar_start = nullptr; out.ops = nullptr; out.ops_cap = 0; out.ops_total = 0;
    out.score3 = score3;
    BatchStats bs;
    std::memset(&bs, 0, sizeof bs);
    std::lock_guard<std::mutex> lk(h->mu);
    rc = run_batch(h, in, out, bs);
    to_stats(bs, stats);
    return rc;
}

int mgl_sw_resident_prepare(mgl_sw_handle *h, const mgl_sw_batch *batch)
{
    if (!h) return MGL_SW_E_INVALID;
    BatchIn in;
    int rc = to_batch(batch, in);
    if (rc != MGL_SW_OK) return rc;
    std::lock_guard<std::mutex> lk(h->mu);
    return h->engine->prepare_resident(in);
}

int mgl_sw_resident_run(mgl_sw_handle *h, mgl_sw_stats *stats)
{
    if (!h) return MGL_SW_E_INVALID;
    BatchStats bs;
    std::memset(&bs, 0, sizeof bs);
    std::lock_guard<std::mutex> lk(h->mu);
    int rc = h->engine->run_resident(&bs);
    to_stats(bs, stats);
    return rc;
}

int mgl_sw_resident_fetch(mgl_sw_handle *h, mgl_sw_result *result)
{
    if (!h || !result || !result->offsets || !result->cigar_start) return MGL_SW_E_INVALID;
    BatchOut out;
    out.offsets = result->offsets; out.cigar_start = result->cigar_start; out.ops = result->ops;
    out.ops_cap = result->ops_cap; out.ops_total = 0;
    std::lock_guard<std::mutex> lk(h->mu);
    int rc = h->engine->fetch_resident(out);
    result->ops_total = out.ops_total;
    return rc;
}

void mgl_sw_resident_drop(mgl_sw_handle *h)
{
    if (!h) return;
    std::lock_guard<std::mutex> lk(h->mu);
    h->engine->drop_resident();
}

int mgl_sw_format_cigar(const uint32_t *ops, int64_t n_ops, char *out, int64_t out_cap)
{
    static const char kOp[] = "MIDNSHP=X";
    int64_t pos = 0;
    for (int64_t k = 0; k < n_ops; k++) {
        char tmp[16];
        const uint32_t op = ops[k] & 15u;
        const int len = std::snprintf(tmp, sizeof tmp, "%u%c", ops[k] >> 4, op < 9 ? kOp[op] : '?');
        if (pos + len > out_cap) return -1;
        std::memcpy(out + pos, tmp, (size_t)len);
        pos += len;
    }
    if (pos < out_cap) out[pos] = 0;
    return (int)pos;
}

/*
 * One pair per call is what the reference's JNI offers (...SmithWaterman.cpp:44-71), and a JVM calls it from many
 * threads at once.  Calls that arrive while a batch is running are therefore COALESCED: the first caller becomes
 * the leader, takes every request that is pending (its own included), runs them as one mgl_sw_align_batch per
 * score set, hands the results back and wakes the others.  A lone caller pays no waiting time (no linger window):
 * batching comes only from the requests that queue up behind the batch in flight.
 */
namespace {
struct OneReq {
    const uint8_t *ref, *alt;
    int32_t ref_len, alt_len, match, mismatch, gap_open, gap_extend, strategy;
    std::vector<uint32_t> ops;
    int32_t offset = 0;
    int rc = MGL_SW_OK;
    bool done = false;
};
std::mutex g_one_mu;
std::condition_variable g_one_cv;
std::vector<OneReq *> g_one_queue;
bool g_one_leader = false;

int run_one_group(mgl_sw_handle *h, std::vector<OneReq *> &grp)
{
    const size_t n = grp.size();
    size_t bytes = 0;
    for (OneReq *r : grp) bytes += (size_t)r->ref_len + (size_t)r->alt_len;
    std::vector<uint8_t> arena(bytes);
    std::vector<uint64_t> ro(n), ao(n), cs(n + 1);
    std::vector<uint32_t> rl(n), al(n);
    std::vector<int32_t> st(n), off(n);
    size_t at = 0, cap = 0;
    for (size_t k = 0; k < n; k++) {
        OneReq *r = grp[k];
        ro[k] = at; std::memcpy(arena.data() + at, r->ref, (size_t)r->ref_len); at += (size_t)r->ref_len;
        ao[k] = at; std::memcpy(arena.data() + at, r->alt, (size_t)r->alt_len); at += (size_t)r->alt_len;
        rl[k] = (uint32_t)r->ref_len; al[k] = (uint32_t)r->alt_len; st[k] = r->strategy;
        cap += (size_t)r->ref_len + (size_t)r->alt_len + 4;
    }
    std::vector<uint32_t> ops(cap);
    mgl_sw_batch b;
    std::memset(&b, 0, sizeof b);
    b.n = (int64_t)n; b.arena = arena.data(); b.arena_bytes = arena.size();
    b.ref_off = ro.data(); b.ref_len = rl.data(); b.alt_off = ao.data(); b.alt_len = al.data();
    b.strategies = st.data(); b.strategy = st[0];
    b.match = grp[0]->match; b.mismatch = grp[0]->mismatch; b.gap_open = grp[0]->gap_open; b.gap_extend = grp[0]->gap_extend;
    mgl_sw_result res;
    res.offsets = off.data(); res.cigar_start = cs.data(); res.ops = ops.data(); res.ops_cap = ops.size(); res.ops_total = 0;
    const int rc = mgl_sw_align_batch(h, &b, &res, nullptr);
    for (size_t k = 0; k < n; k++) {
        OneReq *r = grp[k];
        r->rc = rc;
        if (rc == MGL_SW_OK) {
            r->offset = off[k];
            r->ops.assign(ops.begin() + (ptrdiff_t)cs[k], ops.begin() + (ptrdiff_t)cs[k + 1]);
        }
    }
    return rc;
}

void run_one_batch(mgl_sw_handle *h, std::vector<OneReq *> &batch)
{
    /* a batch carries one score set: group the requests by theirs (in practice all calls of a run share it) */
    std::vector<char> taken(batch.size(), 0);
    for (size_t i = 0; i < batch.size(); i++) {
        if (taken[i]) continue;
        std::vector<OneReq *> grp;
        for (size_t j = i; j < batch.size(); j++) {
            OneReq *a = batch[i], *c = batch[j];
            if (!taken[j] && a->match == c->match && a->mismatch == c->mismatch && a->gap_open == c->gap_open && a->gap_extend == c->gap_extend) {
                taken[j] = 1;
                grp.push_back(c);
            }
        }
        /* one caller's bad pair (an over-long sequence, say) must not fail the unrelated callers coalesced with it:
           a group that fails is run again one request at a time, so that only the offender gets the error */
        if (run_one_group(h, grp) != MGL_SW_OK && grp.size() > 1)
            for (OneReq *r : grp) { std::vector<OneReq *> one(1, r); run_one_group(h, one); }
    }
}
}  // namespace

int mgl_sw_align_one(const uint8_t *ref, int32_t ref_len, const uint8_t *alt, int32_t alt_len,
                     int32_t match, int32_t mismatch, int32_t gap_open, int32_t gap_extend,
                     int32_t strategy, char *cigar_out, int32_t cigar_cap, int32_t *cigar_len, int32_t *offset)
{
    if (!ref || !alt || ref_len <= 0 || alt_len <= 0 || !cigar_out || !offset) return MGL_SW_E_INVALID;
    mgl_sw_handle *h = nullptr;
    { const int rc = mgl_sw_default_handle(&h); if (rc != MGL_SW_OK) return rc; }
    OneReq req;
    req.ref = ref; req.alt = alt; req.ref_len = ref_len; req.alt_len = alt_len;
    req.match = match; req.mismatch = mismatch; req.gap_open = gap_open; req.gap_extend = gap_extend; req.strategy = strategy;
    {
        std::unique_lock<std::mutex> lk(g_one_mu);
        g_one_queue.push_back(&req);
        while (!req.done) {
            if (!g_one_leader) {
                /* lead: take everything that is pending, run it outside the lock */
                g_one_leader = true;
                std::vector<OneReq *> batch;
                batch.swap(g_one_queue);
                lk.unlock();
                try {
                    run_one_batch(h, batch);
                } catch (...) {
                    /* out of host memory while assembling the batch: every request of it fails, nobody waits forever */
                    for (OneReq *r : batch) r->rc = MGL_SW_E_INVALID;
                }
                lk.lock();
                for (OneReq *r : batch) r->done = true;
                g_one_leader = false;
                g_one_cv.notify_all();
            } else {
                g_one_cv.wait(lk);
            }
        }
    }
    mgl_sw_default_release();
    if (req.rc != MGL_SW_OK) return req.rc;
    const int len = mgl_sw_format_cigar(req.ops.data(), (int64_t)req.ops.size(), cigar_out, cigar_cap);
    if (len < 0) return MGL_SW_E_OPS_CAPACITY;
    if (cigar_len) *cigar_len = len;
    *offset = req.offset;
    return MGL_SW_OK;
}

int mgl_sw_pack2bit_host(const uint8_t *src, uint64_t n, uint8_t *dst, uint64_t dst_cap, int scalar)
{
    if (!dst || (n && !src) || dst_cap < mglsw::pack2_bytes((size_t)n)) return -1;
    return (scalar ? mglsw::pack2_scalar(src, (size_t)n, dst) : mglsw::pack2(src, (size_t)n, dst)) ? 1 : 0;
}

void *mgl_sw_host_alloc(uint64_t bytes)
{
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}

void mgl_sw_host_free(void *p) { if (p) cudaFreeHost(p); }

int mgl_sw_plan_summary(const mgl_sw_batch *batch, int allow_s16, uint64_t *out8)
{
    BatchIn in;
    int rc = to_batch(batch, in);
    if (rc != MGL_SW_OK || !out8) return MGL_SW_E_INVALID;
    std::vector<PairIn> pairs((size_t)in.n);
    for (int64_t k = 0; k < in.n; k++) {
        pairs[(size_t)k].ref_off = in.ref_off[k]; pairs[(size_t)k].alt_off = in.alt_off[k];
        pairs[(size_t)k].tl = in.ref_len[k]; pairs[(size_t)k].ql = in.alt_len[k];
        pairs[(size_t)k].strategy = in.strategies ? in.strategies[k] : in.strategy;
    }
    ChunkPlan plan;
    const PlanStatus ps = plan_chunk(pairs.data(), (u32)in.n, in.scores, allow_s16 != 0, plan);
    if (ps == PLAN_BAD_SCORES) return MGL_SW_E_SCORES;
    if (ps == PLAN_BAD_PAIR) return MGL_SW_E_PAIR;
    uint64_t n16 = 0, n32 = 0, padded = 0;
    for (const UnitDesc &u : plan.units) {
        if (!(u.flags & UNIT_ACTIVE)) continue;
        const uint64_t k = 1 + (u.pair[1] != NO_PAIR ? 1 : 0);
        if (u.flags & UNIT_S32) n32 += k; else n16 += k;
    }
    for (const Launch &L : plan.launches) padded += (uint64_t)L.n_units;
    out8[0] = plan.launches.size(); out8[1] = plan.units.size(); out8[2] = n16; out8[3] = n32;
    out8[4] = plan.ckpt_words; out8[5] = plan.cells; out8[6] = plan.scratch_words; out8[7] = padded;
    return MGL_SW_OK;
}

int mgl_sw_int_peak_count(void) { return int_peak_count(); }
const char *mgl_sw_int_peak_name(int which) { return int_peak_name(which); }
int mgl_sw_int_peak(int device, int which, double *ops_per_s, double *sm_mhz)
{
    if (!ops_per_s) return MGL_SW_E_INVALID;
    if (mgl_sw_device_count() <= 0) return MGL_SW_E_NO_DEVICE;
    return run_int_peak(which, device, ops_per_s, sm_mhz) == cudaSuccess ? MGL_SW_OK : MGL_SW_E_CUDA;
}

}  // extern "C"
