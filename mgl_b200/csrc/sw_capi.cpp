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
#include <string>
#include <vector>

#include "../../include/mgl_sw.h"
#include "sw_engine.h"
#include "sw_kernels.h"
#include "sw_pack.h"

using namespace mglsw;

struct mgl_sw_handle {
    Engine *engine;
    std::mutex mu;
    std::string err;
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

mgl_sw_handle *g_default = nullptr;
std::once_flag g_default_once;
int g_default_rc = MGL_SW_E_NO_DEVICE;

void make_default()
{
    const char *env = std::getenv("MGL_SW_DEVICE");
    g_default_rc = mgl_sw_create(env ? std::atoi(env) : 0, &g_default);
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

int mgl_sw_create(int device, mgl_sw_handle **out)
{
    if (!out) return MGL_SW_E_INVALID;
    *out = nullptr;
    int n = mgl_sw_device_count();
    if (n <= 0 || device < 0 || device >= n) return MGL_SW_E_NO_DEVICE;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return MGL_SW_E_NO_DEVICE;
    if (prop.major != 10) {       /* the kernels are sm_100a only: refuse rather than fall back */
        std::fprintf(stderr, "mgl_sw: device %d is sm_%d%d, this library is built for sm_100a only\n", device, prop.major, prop.minor);
        return MGL_SW_E_NO_DEVICE;
    }
    mgl_sw_handle *h = new (std::nothrow) mgl_sw_handle();
    if (!h) return MGL_SW_E_INVALID;
    h->engine = new Engine(device);
    *out = h;
    return MGL_SW_OK;
}

void mgl_sw_destroy(mgl_sw_handle *h)
{
    if (!h) return;
    delete h->engine;
    delete h;
}

const char *mgl_sw_last_error(mgl_sw_handle *h) { return h ? h->engine->last_error().c_str() : "no handle"; }

int mgl_sw_set_option(mgl_sw_handle *h, const char *name, int64_t value)
{
    if (!h || !name) return MGL_SW_E_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    if (!std::strcmp(name, "allow_s16")) h->engine->allow_s16 = value != 0;
    else if (!std::strcmp(name, "overlap")) h->engine->overlap = value != 0;
    else if (!std::strcmp(name, "plan_threads")) { if (value < 0 || value > 64) return MGL_SW_E_INVALID; h->engine->plan_threads = (int)value; }
    else if (!std::strcmp(name, "chunk_pairs")) { if (value < 1 || value > (1 << 22)) return MGL_SW_E_INVALID; h->engine->chunk_pairs = (u32)value; }
    else if (!std::strcmp(name, "pipe_wide")) { if (value < 0 || value > 2) return MGL_SW_E_INVALID; h->engine->pipe_wide = (int)value; }
    else if (!std::strcmp(name, "pack_arena")) { if (value < 0 || value > 2) return MGL_SW_E_INVALID; h->engine->pack_arena = (int)value; }
    else if (!std::strcmp(name, "pack_lead")) { if (value < 0 || value > 1024) return MGL_SW_E_INVALID; h->engine->pack_lead = (int)value; }
    else if (!std::strcmp(name, "trace_budget_mb")) { if (value < 0) return MGL_SW_E_INVALID; h->engine->trace_budget_bytes = (u64)value << 20; }
    else return MGL_SW_E_INVALID;
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
    rc = h->engine->align_batch(in, out, &bs);
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
    rc = h->engine->align_batch(in, out, &bs);
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
    rc = h->engine->align_batch(in, out, &bs);
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

void run_one_group(std::vector<OneReq *> &grp)
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
    const int rc = mgl_sw_align_batch(g_default, &b, &res, nullptr);
    for (size_t k = 0; k < n; k++) {
        OneReq *r = grp[k];
        r->rc = rc;
        if (rc == MGL_SW_OK) {
            r->offset = off[k];
            r->ops.assign(ops.begin() + (ptrdiff_t)cs[k], ops.begin() + (ptrdiff_t)cs[k + 1]);
        }
    }
}

void run_one_batch(std::vector<OneReq *> &batch)
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
        run_one_group(grp);
    }
}
}  // namespace

int mgl_sw_align_one(const uint8_t *ref, int32_t ref_len, const uint8_t *alt, int32_t alt_len,
                     int32_t match, int32_t mismatch, int32_t gap_open, int32_t gap_extend,
                     int32_t strategy, char *cigar_out, int32_t cigar_cap, int32_t *cigar_len, int32_t *offset)
{
    if (!ref || !alt || ref_len <= 0 || alt_len <= 0 || !cigar_out || !offset) return MGL_SW_E_INVALID;
    std::call_once(g_default_once, make_default);
    if (g_default_rc != MGL_SW_OK || !g_default) return MGL_SW_E_NO_DEVICE;
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
                run_one_batch(batch);
                lk.lock();
                for (OneReq *r : batch) r->done = true;
                g_one_leader = false;
                g_one_cv.notify_all();
            } else {
                g_one_cv.wait(lk);
            }
        }
    }
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
