// pgm_api.cu — C ABI of the gmove hot path (include/pgm_gmove.h): context, device memory,
// kernel sequencing. No CPU fallback: every compute entry point needs a live CUDA device.
//
// Pipeline of one batch (all on one stream, no host synchronisation until pgm_download):
//   count: k_read_prep -> k_order -> [k_oor_check] -> k_decode | k_decode_moves -> k_scan_records -> k_tile_hist
//   emit : k_quota || k_scan_hist -> k_rank (tables of one digit) | k_msd_scatter -> k_msd_rank (two digits)
//          -> k_scan_sel -> [k_stats_wide] -> [k_stats] -> k_gather -> [k_text_len -> k_scan_text -> k_text_write]
// Element counts that only the device knows (passing / selected events) stay on the device: grids
// are sized by capacities derived from the batch and the kernels read the counts themselves.
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>   // header-only: resolves to no-ops unless a tool (nsys, ncu --nvtx) injects itself
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <string>
#include <vector>

#include "pgm_device.cuh"
#include "pgm_scan.cuh"
#include "pgm_decode.cuh"
#include "pgm_emit.cuh"
#include "pgm_signal.cuh"
#include "pgm_sort.cuh"
#include "pgm_text.cuh"
#include "pgm_model.cuh"
#include "pgm_svb.cuh"

namespace {

thread_local std::string g_create_error;

struct DBuf {
    void *p = nullptr;
    size_t cap = 0;
    bool guarded = false;        // PGM_GUARD=1: PGM_GUARD_BYTES of 0xA5 in front of p and behind p + cap (cap = the bytes asked for)
    const char *name = nullptr;
};
#define PGM_GUARD_BYTES 4096


typedef BatchScalars Scalars;  // device -> host summary of one batch (pgm_emit.cuh)
#define MISC_WORDS 160          // word 0 = k_stats_wide queue, 1 = k_decode record ticket, 2 = k_stats ticket, 3 = k_read_prep CTAs done, 4 = k_msd_rank bucket ticket,
                                // 8..63 = look-back scan tickets, 96..159 = record length classes (k_read_prep)

enum State { ST_IDLE = 0, ST_LOADED, ST_COUNTED, ST_EMITTED };

}  // namespace

struct pgm_ctx {
    pgm_opts opts;
    std::string kmer_store;
    DevOpts dopts;
    int device = 0;
    int n_sm = 148;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    std::string err;
    State state = ST_IDLE;
    uint64_t launches = 0;

    // per-context persistent device state
    DBuf lut, consumed, counts, keep, bucket_base, sel_base, gscratch;
    DBuf totals_snap;                 // kmer_frequency_map right after this context's newest batch (consumed is shared by siblings)
    uint32_t lb_ticket = 0;          // look-back scans launched since the last zeroing (each gets its own ticket + status words)
    size_t lb_used = 0, lb_words = 0;   // status words handed out / available in zbatch
    const uint32_t *user_prior = nullptr, *user_consumed = nullptr;
    const uint32_t *gathered = nullptr;       // pgm_set_gathered
    // model aggregation log (pgm_model_begin / pgm_model_finalize): one entry per batch, the values in two growing arenas
    // (nothing is allocated per batch once the arenas have their size)
    struct ModelBatch { size_t vals_off, n_vals, lens_off, n_ev; std::vector<unsigned long long> h_slot_off; std::vector<uint32_t> h_ev_off; };
    DBuf model_vals, model_lens;              // int64 values (N of "%.8f") / uint32 event lengths of every logged batch, back to back
    size_t model_vals_used = 0, model_lens_used = 0;
    DBuf model_tab;                           // device copy of one batch's slot table while it is grouped
    bool model_on = false, model_logged = false;   // model_logged: the emitted batch is in the log already
    std::vector<ModelBatch> model_log;
    DBuf model_bad, model_grouped, model_off, model_dst, model_rows;
    std::vector<pgm_model_row> h_model_rows;
    std::vector<int64_t> h_dwell;
    cudaStream_t aux = nullptr;               // short scans that do not depend on each other run beside the main stream
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    size_t text_written_cap = 0;              // capacity of `text` when pgm_emit formatted into it (0: it did not)
    uint32_t n_ranks = 1, rank = 0;

    // context family: siblings share the per-k-mer totals (`consumed` of the root) and take turns in record order
    pgm_ctx *root = nullptr;                  // null: this context is the root of its family
    bool guard = false;                       // PGM_GUARD=1 at pgm_create: every device buffer exact-sized between guard zones, checked by pgm_download
    std::vector<DBuf *> guard_list;
    bool guard_selftest = false;              // PGM_GUARD=2: guard_check first writes one byte behind `cal`, to show that the check sees it
    cudaEvent_t ev_quota = nullptr;           // root: recorded after the newest k_quota / reset of the family
    bool quota_recorded = false;
    std::vector<pgm_ctx *> pipe;              // pgm_submit / pgm_wait: the internal siblings, one per batch in flight
    uint64_t submitted = 0, waited = 0;

    // batch (device view) + owned copies
    DevBatch db{};
    bool attached = false;
    DBuf raw, align, bases, reads;
    DBuf svb_blob, svb_off, svb_words;        // pgm_upload_svb: streams, their offsets, {ticket, error} words
    bool svb_used = false;
    uint32_t gather_ctas = 0;                 // resident k_gather CTAs per SM (occupancy query, first pgm_emit)
    DBuf h_svb_err;
    uint32_t ev_cap = 0;
    uint64_t samples_cap = 0;

    // per-record
    DBuf cal, rflags, n_pass, ev_base;
    // events / sort / selection (rows: pgm_sort.cuh)
    DBuf rows_tmp;                    // per-record temporaries of the decoders
    DBuf rows_a;                      // two-digit tables: the batch split by the high digit
    DBuf sel;                         // selected events, grouped by slot
    DBuf rec_src0;                    // [R] index in raw[] of every record's sample 0
    DBuf sc_start, sc_end, sc_code, sc_indel;   // ss decoder scratch
    uint32_t tmp_shift = 0;           // rows_tmp index = align_off >> tmp_shift
    DBuf zbatch;                       // zeroed every batch: scalars | misc words | scan status | read_has_sel[R] | tile histogram
    size_t has_sel_off = 0, hist_off = 0;   // byte offsets inside zbatch
    uint32_t sort_ntiles = 0;
    DBuf sel_start, sel_len, sel_read;   // structure-of-arrays form of `sel`, made by pgm_download for the host (k_unpack_sel)
    DBuf sel_off, samples, rec_packed, rec_order;
    DBuf ev_tlen, ev_toff, text, slot_toff;   // PGM_OUT_TEXT
    DBuf hp_prior, hp_consumed;               // pgm_set_prior_host
    bool need_wide = true;  // some record may have an in-range code span above SIG_SMEM_BINS

    // pinned host results
    Scalars *h_scalars = nullptr;
    DBuf h_slot_sel, h_slot_total, h_ev_len, h_ev_read, h_ev_start, h_samples, h_rflags, h_text, h_slot_toff, h_ev_toff;
    // pinned staging for pageable inputs is left to the caller (pass pinned buffers for speed)

    // optional per-stage event timing
    bool prof = false;
    int prof_only = -1;      // >= 0: only this stage is bracketed (pgm_profile_enable(ctx, PGM_PROFILE_ONLY(stage)))
    struct Span { int stage; cudaEvent_t a, b; uint32_t launches; };
    std::vector<Span> spans;
    std::vector<cudaEvent_t> ev_pool;
};

namespace {

int fail(pgm_ctx *c, int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (c) c->err = buf; else g_create_error = buf;
    return code;
}

// NVTX range around a C-ABI phase (the reference has no tracing; an nsys timeline of the e2e leg shows
// upload / count / emit / download of two batches overlapping). Domain "pgm_gmove".
struct PhaseRange {
    static nvtxDomainHandle_t domain() { static nvtxDomainHandle_t d = nvtxDomainCreateA("pgm_gmove"); return d; }
    explicit PhaseRange(const char *name) {
        nvtxEventAttributes_t a = {};
        a.version = NVTX_VERSION; a.size = NVTX_EVENT_ATTRIB_STRUCT_SIZE;
        a.messageType = NVTX_MESSAGE_TYPE_ASCII; a.message.ascii = name;
        nvtxDomainRangePushEx(domain(), &a);
    }
    ~PhaseRange() { nvtxDomainRangePop(domain()); }
};
#define PHASE(name) PhaseRange phase_range_(name)

#define CU(c, call)                                                                                        \
    do {                                                                                                   \
        cudaError_t e_ = (call);                                                                           \
        if (e_ != cudaSuccess) return fail((c), PGM_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e_)); \
    } while (0)

void free_dev(DBuf &b) {
    if (b.p) cudaFree(b.guarded ? (void *)((char *)b.p - PGM_GUARD_BYTES) : b.p);
    b.p = nullptr; b.cap = 0; b.guarded = false;
}

// Guard mode (environment PGM_GUARD=1, read by pgm_create; a debugging aid where compute-sanitizer is not available):
// a buffer is exactly the bytes asked for, filled with 0xA5 (so stale or never-written data is not zero), with
// PGM_GUARD_BYTES of 0xA5 on both sides; guard_check() after every pgm_download / pgm_summary / pgm_model_finalize
// fails the call when a kernel wrote into a guard zone. Slow (reallocates whenever a size changes): tests only.
int ensure_(pgm_ctx *c, DBuf &b, size_t bytes, const char *name) {
    if (c->guard) {
        if (b.p && b.guarded && b.cap == bytes) return PGM_OK;
        free_dev(b);
        char *raw = nullptr;
        cudaError_t e = cudaMalloc(&raw, bytes + 2 * PGM_GUARD_BYTES);
        if (e == cudaSuccess) e = cudaMemset(raw, 0xA5, bytes + 2 * PGM_GUARD_BYTES);
        if (e == cudaSuccess) e = cudaDeviceSynchronize();      // the fill runs on the legacy stream: the context's non-blocking streams do not wait for it
        if (e != cudaSuccess) return fail(c, PGM_ERR_NOMEM, "guarded cudaMalloc(%zu) for %s failed: %s", bytes, name, cudaGetErrorString(e));
        b.p = raw + PGM_GUARD_BYTES; b.cap = bytes; b.guarded = true; b.name = name;
        if (std::find(c->guard_list.begin(), c->guard_list.end(), &b) == c->guard_list.end()) c->guard_list.push_back(&b);
        return PGM_OK;
    }
    if (bytes <= b.cap && b.p) return PGM_OK;
    free_dev(b);
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&b.p, want);
    if (e != cudaSuccess) { b.p = nullptr; return fail(c, PGM_ERR_NOMEM, "cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(e)); }
    b.cap = want;
    return PGM_OK;
}
#define ensure(c, b, bytes) ensure_((c), (b), (bytes), #b)

// All guard zones of the context's buffers still hold 0xA5? (synchronises; guard mode only)
int guard_check(pgm_ctx *c, const char *where) {
    if (!c->guard) return PGM_OK;
    if (cudaStreamSynchronize(c->stream) != cudaSuccess) return PGM_OK;      // the caller reports CUDA errors itself
    std::vector<unsigned char> h(2 * PGM_GUARD_BYTES);
    if (c->guard_selftest && c->cal.p && c->cal.guarded) CU(c, cudaMemset((char *)c->cal.p + c->cal.cap + 3, 0x00, 1));
    for (DBuf *b : c->guard_list) {
        if (!b->p || !b->guarded) continue;
        CU(c, cudaMemcpy(h.data(), (char *)b->p - PGM_GUARD_BYTES, PGM_GUARD_BYTES, cudaMemcpyDeviceToHost));
        CU(c, cudaMemcpy(h.data() + PGM_GUARD_BYTES, (char *)b->p + b->cap, PGM_GUARD_BYTES, cudaMemcpyDeviceToHost));
        for (size_t i = 0; i < h.size(); i++)
            if (h[i] != 0xA5) {
                const bool front = i < PGM_GUARD_BYTES;
                return fail(c, PGM_ERR_STATE, "PGM_GUARD: %s: buffer %s (%zu bytes): byte %zu %s the buffer was overwritten (0x%02x)", where, b->name ? b->name : "?",
                            b->cap, front ? PGM_GUARD_BYTES - i : i - PGM_GUARD_BYTES, front ? "in front of" : "behind", h[i]);
            }
    }
    return PGM_OK;
}

int ensure_host(pgm_ctx *c, DBuf &b, size_t bytes) {
    if (bytes <= b.cap && b.p) return PGM_OK;
    if (b.p) { cudaFreeHost(b.p); b.p = nullptr; b.cap = 0; }
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaHostAlloc(&b.p, want, cudaHostAllocDefault);
    if (e != cudaSuccess) { b.p = nullptr; return fail(c, PGM_ERR_NOMEM, "cudaHostAlloc(%zu) failed: %s", want, cudaGetErrorString(e)); }
    b.cap = want;
    return PGM_OK;
}

// Room for `more` elements of `elem` bytes behind the `used` ones of a growing device array: the contents move to a buffer
// of at least twice the size when it is full (amortised: a long run settles on its final size after a few batches).
int arena_reserve(pgm_ctx *c, DBuf &b, size_t used, size_t more, size_t elem) {
    const size_t need = (used + more) * elem;
    if (need <= b.cap && b.p) return PGM_OK;
    size_t want = std::max(need + need / 2, (size_t)1 << 20);
    void *np = nullptr;
    cudaError_t e = cudaMalloc(&np, want);
    if (e != cudaSuccess) return fail(c, PGM_ERR_NOMEM, "cudaMalloc(%zu) for the model log failed: %s", want, cudaGetErrorString(e));
    if (b.p && used) {
        e = cudaMemcpyAsync(np, b.p, used * elem, cudaMemcpyDeviceToDevice, c->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) { cudaFree(np); return fail(c, PGM_ERR_CUDA, "moving the model log failed: %s", cudaGetErrorString(e)); }
    }
    if (b.p) cudaFree(b.p);
    b.p = np; b.cap = want;
    return PGM_OK;
}

void release(DBuf &b) { free_dev(b); }
void release_host(DBuf &b) { if (b.p) cudaFreeHost(b.p); b.p = nullptr; b.cap = 0; }

template <typename T> T *P(DBuf &b) { return reinterpret_cast<T *>(b.p); }


#define ZB_MISC_OFF 64
#define ZB_STATUS_OFF (ZB_MISC_OFF + MISC_WORDS * 4)
inline Scalars *SCALARS(pgm_ctx *c) { return reinterpret_cast<Scalars *>(c->zbatch.p); }
inline uint32_t *MISC(pgm_ctx *c) { return reinterpret_cast<uint32_t *>(reinterpret_cast<uint8_t *>(c->zbatch.p) + ZB_MISC_OFF); }
inline uint8_t *HAS_SEL(pgm_ctx *c) { return reinterpret_cast<uint8_t *>(c->zbatch.p) + c->has_sel_off; }
inline uint32_t *HIST(pgm_ctx *c) { return reinterpret_cast<uint32_t *>(reinterpret_cast<uint8_t *>(c->zbatch.p) + c->hist_off); }

#define RET(x) do { int rc_ = (x); if (rc_ != PGM_OK) return rc_; } while (0)

inline uint32_t cdiv(uint64_t a, uint64_t b) { return (uint32_t)((a + b - 1) / b); }

uint32_t bit_length(uint32_t x) { uint32_t b = 0; while (x) { b++; x >>= 1; } return b; }

// Brackets a pipeline stage with a CUDA event pair on the context's stream when profiling is on.
struct StageTimer {
    pgm_ctx *c;
    int stage;
    cudaEvent_t a = nullptr, b = nullptr;
    uint64_t l0;
    static cudaEvent_t get(pgm_ctx *c) {
        if (!c->ev_pool.empty()) { cudaEvent_t e = c->ev_pool.back(); c->ev_pool.pop_back(); return e; }
        cudaEvent_t e = nullptr;
        cudaEventCreate(&e);
        return e;
    }
    cudaStream_t s;
    StageTimer(pgm_ctx *c_, int stage_, cudaStream_t s_ = nullptr) : c(c_), stage(stage_), l0(c_->launches), s(s_ ? s_ : c_->stream) {
        if (!c->prof || (c->prof_only >= 0 && c->prof_only != stage)) return;
        a = get(c); b = get(c);
        cudaEventRecord(a, s);
    }
    void stop() {
        if (!a) return;
        cudaEventRecord(b, s);
        c->spans.push_back({stage, a, b, (uint32_t)(c->launches - l0)});
        a = nullptr;
    }
    ~StageTimer() { stop(); }
};

}  // namespace

// ================================================================================================
extern "C" {

int pgm_abi_version(void) { return PGM_ABI_VERSION; }

const char *pgm_last_error(const pgm_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

void pgm_destroy(pgm_ctx *c) {
    if (!c) return;
    cudaSetDevice(c->device);
    for (pgm_ctx *k : c->pipe) pgm_destroy(k);
    c->pipe.clear();
    if (c->root) { c->consumed.p = nullptr; c->consumed.cap = 0; }   // the totals belong to the root
    if (c->ev_quota) cudaEventDestroy(c->ev_quota);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->aux) { cudaStreamSynchronize(c->aux); cudaStreamDestroy(c->aux); }
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    if (c->ev_join) cudaEventDestroy(c->ev_join);
    c->model_log.clear();
    for (DBuf *mbuf : {&c->model_bad, &c->model_grouped, &c->model_off, &c->model_dst, &c->model_rows, &c->model_vals, &c->model_lens, &c->model_tab}) release(*mbuf);
    DBuf *dev[] = {&c->lut, &c->consumed, &c->counts, &c->keep, &c->bucket_base, &c->sel_base, &c->totals_snap,
                   &c->gscratch, &c->raw, &c->align, &c->bases, &c->reads, &c->cal, &c->rflags, &c->n_pass,
                   &c->ev_base, &c->rows_tmp, &c->rows_a, &c->sel, &c->rec_src0, &c->sc_start, &c->sc_end, &c->sc_code,
                   &c->sc_indel, &c->zbatch, &c->sel_start, &c->sel_len, &c->sel_read, &c->sel_off, &c->samples, &c->rec_packed, &c->rec_order, &c->ev_tlen, &c->ev_toff, &c->text, &c->slot_toff, &c->hp_prior, &c->hp_consumed, &c->svb_blob, &c->svb_off, &c->svb_words};
    for (DBuf *b : dev) release(*b);
    DBuf *host[] = {&c->h_slot_sel, &c->h_slot_total, &c->h_ev_len, &c->h_ev_read, &c->h_ev_start, &c->h_samples, &c->h_rflags, &c->h_text, &c->h_slot_toff, &c->h_ev_toff, &c->h_svb_err};
    for (DBuf *b : host) release_host(*b);
    if (c->h_scalars) cudaFreeHost(c->h_scalars);
    for (auto &sp : c->spans) { cudaEventDestroy(sp.a); cudaEventDestroy(sp.b); }
    for (cudaEvent_t e : c->ev_pool) cudaEventDestroy(e);
    if (c->own_stream && c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

int pgm_create(const pgm_opts *opts, int device, void *stream, pgm_ctx **out) {
    if (!opts || !out) return fail(nullptr, PGM_ERR_ARG, "pgm_create: null argument");
    *out = nullptr;
    if (opts->kmer_size < 1 || opts->kmer_size > 15) return fail(nullptr, PGM_ERR_ARG, "kmer_size %u outside 1..15", opts->kmer_size);
    if (opts->n_slots < 1) return fail(nullptr, PGM_ERR_ARG, "n_slots must be >= 1");
    if (opts->scaling > 1) return fail(nullptr, PGM_ERR_ARG, "scaling must be 0 or 1");
    if (opts->output > PGM_OUT_TEXT_ONLY) return fail(nullptr, PGM_ERR_ARG, "unknown output kind %u", opts->output);
    const uint64_t n_ids = 1ull << (2 * opts->kmer_size);
    if (!opts->kmer_list && (uint64_t)opts->index_lo + opts->n_slots > n_ids)
        return fail(nullptr, PGM_ERR_ARG, "k-mer window [%u,+%u) exceeds 4^k", opts->index_lo, opts->n_slots);
    if (opts->kmer_list && opts->kmer_size > 12) return fail(nullptr, PGM_ERR_ARG, "--kmer_file lists need kmer_size <= 12");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(nullptr, PGM_ERR_CUDA, "no CUDA device (%s); the gmove hot path has no CPU fallback", cudaGetErrorString(e));
    if (device < 0 || device >= ndev) return fail(nullptr, PGM_ERR_ARG, "device %d out of range (have %d)", device, ndev);
    cudaDeviceProp prop;
    if ((e = cudaSetDevice(device)) != cudaSuccess || (e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess)
        return fail(nullptr, PGM_ERR_CUDA, "cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
    if (prop.major != 10) return fail(nullptr, PGM_ERR_CUDA, "device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);

    if (bit_length(opts->n_slots - 1) > 2 * MSD_MAX_LO_BITS)
        return fail(nullptr, PGM_ERR_ARG, "n_slots %u needs more than two digits of %d bits", opts->n_slots, MSD_MAX_LO_BITS);
    if ((e = cudaFuncSetAttribute(k_rank, cudaFuncAttributeMaxDynamicSharedMemorySize, RANK_SMEM)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(k_msd_scatter, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MSD_SCATTER_SMEM(MSD_MAX_LO_BITS))) != cudaSuccess ||
        (e = cudaFuncSetAttribute(k_msd_rank, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MSD_RANK_SMEM(MSD_MAX_LO_BITS))) != cudaSuccess ||
        (e = cudaFuncSetAttribute(k_stats, cudaFuncAttributeMaxDynamicSharedMemorySize, STATS_DYN_SMEM)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(k_gather<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GATHER_SMEM)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(k_gather<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GATHER_SMEM)) != cudaSuccess)
        return fail(nullptr, PGM_ERR_CUDA, "cudaFuncSetAttribute(k_stats / k_gather): %s", cudaGetErrorString(e));

    pgm_ctx *c = new pgm_ctx();
    c->opts = *opts;
    c->device = device;
    c->n_sm = prop.multiProcessorCount;
    if (stream) { c->stream = (cudaStream_t)stream; c->own_stream = false; }
    else {
        if ((e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess) {
            delete c;
            return fail(nullptr, PGM_ERR_CUDA, "cudaStreamCreate: %s", cudaGetErrorString(e));
        }
        c->own_stream = true;
    }
    if ((e = cudaEventCreateWithFlags(&c->ev_quota, cudaEventDisableTiming)) != cudaSuccess ||
        (e = cudaStreamCreateWithFlags(&c->aux, cudaStreamNonBlocking)) != cudaSuccess ||
        (e = cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming)) != cudaSuccess ||
        (e = cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming)) != cudaSuccess) {
        delete c;
        return fail(nullptr, PGM_ERR_CUDA, "cudaStreamCreate (aux): %s", cudaGetErrorString(e));
    }
    DevOpts &d = c->dopts;
    memset(&d, 0, sizeof d);
    d.k = opts->kmer_size; d.m = opts->sig_move_offset; d.s_off = opts->kmer_start_offset; d.margin = opts->signal_print_margin;
    d.limit = opts->sample_limit; d.max_dur = opts->max_dur; d.min_dur = opts->min_dur; d.pick_margin = opts->kmer_pick_margin;
    d.scaling = opts->scaling; d.n_slots = opts->n_slots; d.index_lo = opts->index_lo;
    d.pa_min = opts->pa_min; d.pa_max = opts->pa_max;
    d.fourth = opts->flag_rna ? 'U' : 'T';
    if (const char *dbg = getenv("PGM_DEBUG_SLOW_STATS")) d.debug |= (dbg[0] && dbg[0] != '0') ? 1u : 0u;
    if (const char *gd = getenv("PGM_GUARD")) { c->guard = gd[0] && gd[0] != '0'; c->guard_selftest = gd[0] == '2'; }
    int rc = PGM_OK;
    auto bail = [&](int code) { std::string m = c->err; pgm_destroy(c); g_create_error = m; return code; };
    if (opts->kmer_list) {
        // --kmer_file: the list is verbatim text (gmove.cpp:406-414). Lines made of A,C,G and ONE of
        // T/U get a LUT entry; any other line can never equal a substring of an ACGT/ACGU sequence
        // and simply keeps count 0 (documented limitation: 'N' or mixed T/U lines do not match).
        c->kmer_store.assign(opts->kmer_list, (size_t)opts->n_slots * opts->kmer_size);
        c->opts.kmer_list = c->kmer_store.data();
        bool has_u = c->kmer_store.find('U') != std::string::npos, has_t = c->kmer_store.find('T') != std::string::npos;
        d.fourth = (has_u && !has_t) ? 'U' : (has_t ? 'T' : (opts->flag_rna ? 'U' : 'T'));
        std::vector<uint32_t> lut(n_ids, PGM_INVALID);
        for (uint32_t s = 0; s < opts->n_slots; s++) {
            uint32_t id = 0; bool ok = true;
            for (uint32_t j = 0; j < opts->kmer_size; j++) {
                const char ch = c->kmer_store[(size_t)s * opts->kmer_size + j];
                uint32_t code = ch == 'A' ? 0 : ch == 'C' ? 1 : ch == 'G' ? 2 : ((uint32_t)ch == d.fourth) ? 3 : 0xFF;
                if (code == 0xFF) { ok = false; break; }
                id = (id << 2) | code;
            }
            if (!ok) continue;
            if (lut[id] != PGM_INVALID) continue;   // a repeated line: the reference's std::map keeps ONE file and ONE count per k-mer string
            lut[id] = s;                            // (gmove.cpp:460-473): the first slot of the string takes the events, the others stay empty
        }
        if ((rc = ensure(c, c->lut, n_ids * 4)) != PGM_OK) return bail(rc);
        if (cudaMemcpy(c->lut.p, lut.data(), n_ids * 4, cudaMemcpyHostToDevice) != cudaSuccess) { c->err = "LUT upload failed"; return bail(PGM_ERR_CUDA); }
        d.use_lut = 1; d.lut = P<uint32_t>(c->lut);
    }
    const size_t sb = (size_t)opts->n_slots * 4;
    DBuf *slotbufs[] = {&c->consumed, &c->counts, &c->keep, &c->totals_snap};
    for (DBuf *b : slotbufs) if ((rc = ensure(c, *b, sb)) != PGM_OK) return bail(rc);
    if ((rc = ensure(c, c->bucket_base, sb + 4)) != PGM_OK || (rc = ensure(c, c->sel_base, sb + 4)) != PGM_OK) return bail(rc);
    if (cudaMemset(c->consumed.p, 0, sb) != cudaSuccess) { c->err = "memset failed"; return bail(PGM_ERR_CUDA); }
    if (cudaHostAlloc((void **)&c->h_scalars, sizeof(Scalars), cudaHostAllocDefault) != cudaSuccess) { c->err = "cudaHostAlloc failed"; return bail(PGM_ERR_NOMEM); }
    *out = c;
    return PGM_OK;
}

int pgm_reset(pgm_ctx *c) {
    if (!c) return PGM_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaMemsetAsync(c->consumed.p, 0, (size_t)c->opts.n_slots * 4, c->stream));
    {
        pgm_ctx *root = c->root ? c->root : c;
        if (root->ev_quota) { CU(c, cudaEventRecord(root->ev_quota, c->stream)); root->quota_recorded = true; }
    }
    c->user_prior = c->user_consumed = nullptr; c->gathered = nullptr;
    if (c->state != ST_IDLE) c->state = ST_LOADED;  // the loaded batch stays usable: count/emit can be re-run
    return PGM_OK;
}

void *pgm_stream(pgm_ctx *c) { return c ? (void *)c->stream : nullptr; }
uint64_t pgm_launch_count(const pgm_ctx *c) { return c ? c->launches : 0; }

int pgm_sync(pgm_ctx *c) {
    if (!c) return PGM_ERR_ARG;
    CU(c, cudaStreamSynchronize(c->stream));
    return PGM_OK;
}

// digits of the slot: tables up to 2^SORT_MAX_BITS slots are ranked in one pass (hi_bits = 0), larger ones are split
// by the high digit first (pgm_sort.cuh). `first` = bits of the digit the tile histogram counts.
struct SortPlan { uint32_t bits, hi_bits, lo_bits, first_bits, first_shift; };
static SortPlan sort_plan(const pgm_ctx *c) {
    SortPlan p;
    p.bits = bit_length(c->opts.n_slots - 1);
    if (p.bits == 0) p.bits = 1;
    if (p.bits <= SORT_MAX_BITS) { p.hi_bits = 0; p.lo_bits = p.bits; p.first_bits = p.bits; p.first_shift = 0; }
    else {
        p.hi_bits = (p.bits + 1) / 2;
        if (const char *e = getenv("PGM_MSD_HI_BITS")) { const int v = atoi(e); if (v >= (int)p.bits - MSD_MAX_LO_BITS && v >= 1 && v <= SORT_MAX_BITS && v < (int)p.bits) p.hi_bits = (uint32_t)v; }
        p.lo_bits = p.bits - p.hi_bits; p.first_bits = p.hi_bits; p.first_shift = p.lo_bits;
    }
    return p;
}

// ------------------------------------------------------------------------------------------------
static int prepare_batch(pgm_ctx *c, const pgm_batch *b) {
    if (b->mode > PGM_MODE_MOVES_I8) return fail(c, PGM_ERR_ARG, "unknown batch mode %u", b->mode);
    if (b->n_align >= (1ull << 32)) return fail(c, PGM_ERR_ARG, "n_align must be < 2^32 per batch");
    if (b->n_raw >= (1ull << 40)) return fail(c, PGM_ERR_ARG, "n_raw must be < 2^40 per batch");
    if (b->n_reads && (!b->reads || (b->n_raw && !b->raw) || (b->n_align && !b->align) || (b->n_bases && !b->bases)))
        return fail(c, PGM_ERR_ARG, "null array in batch");
    const uint32_t R = b->n_reads;
    // capacity of the passing-event list: every event consumes >= 1 alignment byte (ss: >= 2) and
    // needs a k-mer that starts at its own base of the record's sequence.
    uint64_t cap = (b->mode == PGM_MODE_SS) ? b->n_align / 2 + R : b->n_align;
    if (b->n_bases < cap) cap = b->n_bases;
    if (cap >= 0x7FFFFFF0ull) return fail(c, PGM_ERR_ARG, "batch too large: split it");
    c->ev_cap = (uint32_t)cap;
    // every selected sample is one raw sample of one event, widened by --margin on both sides
    c->samples_cap = b->n_raw + 2ull * c->opts.signal_print_margin * cap;

    RET(ensure(c, c->cal, (size_t)R * sizeof(ReadCal)));
    RET(ensure(c, c->rflags, R));
    RET(ensure(c, c->n_pass, (size_t)R * 4));
    RET(ensure(c, c->rec_packed, (size_t)R * 4 + 4));
    RET(ensure(c, c->rec_order, (size_t)R * 4 + 4));
    RET(ensure(c, c->ev_base, (size_t)(R + 1) * 4));
    const size_t e4 = (size_t)cap * 4;
    const SortPlan sp = sort_plan(c);
    RET(ensure(c, c->rec_src0, (size_t)R * 8 + 8));
    // rows_tmp: one 16-byte row per possible passing event of a record, at the record's alignment offset (ss: half of it)
    c->tmp_shift = (b->mode == PGM_MODE_SS) ? 1u : 0u;
    RET(ensure(c, c->rows_tmp, (((size_t)b->n_align >> c->tmp_shift) + R + 1) * 16));
    RET(ensure(c, c->sel, (size_t)cap * 16 + 16));
    if (sp.hi_bits) RET(ensure(c, c->rows_a, (size_t)cap * 16 + 16));
    RET(ensure(c, c->sel_off, (size_t)(cap + 1) * 8));
    const uint32_t ntiles = cdiv(cap ? cap : 1, SORT_TILE);
    c->sort_ntiles = ntiles;
    {   // per-batch zero region: scalars | misc words | look-back status words | read_has_sel[R]; then (not zeroed) the tile histogram
        c->lb_words = (size_t)R / LB_TILE + c->opts.n_slots / LB_TILE + 2 * ((size_t)cap / LB_TILE) + 16 +
                      (((size_t)ntiles << sp.first_bits) / LB_TILE + 2);
        c->has_sel_off = (ZB_STATUS_OFF + c->lb_words * 8 + 255) & ~(size_t)255;
        c->hist_off = (c->has_sel_off + (size_t)R + 255) & ~(size_t)255;
        RET(ensure(c, c->zbatch, c->hist_off + ((size_t)ntiles << sp.first_bits) * 4));
    }
    RET(ensure(c, c->samples, (size_t)c->samples_cap * 8));
    if (c->opts.output != PGM_OUT_F64) {
        RET(ensure(c, c->ev_tlen, e4 + 4));
        RET(ensure(c, c->ev_toff, (size_t)(cap + 1) * 8));
        RET(ensure(c, c->slot_toff, (size_t)(c->opts.n_slots + 1) * 8));
    }
    if (b->mode == PGM_MODE_SS) {
        RET(ensure(c, c->sc_start, (size_t)b->n_align * 4 + 4));
        RET(ensure(c, c->sc_end, (size_t)b->n_align * 4 + 4));
        RET(ensure(c, c->sc_code, (size_t)b->n_align + 4));
        RET(ensure(c, c->sc_indel, (size_t)b->n_align * 4 + 4));
    }
    if (c->opts.scaling) RET(ensure(c, c->gscratch, (size_t)c->n_sm * 4 * 65536 * 4));
    return PGM_OK;
}

// H2D copies of a host batch; `with_raw` false leaves raw[] to be produced on the device (pgm_upload_svb)
static int upload_common(pgm_ctx *c, const pgm_batch *b, bool with_raw) {
    RET(prepare_batch(c, b));
    RET(ensure(c, c->raw, (size_t)b->n_raw * 2 + 32));
    RET(ensure(c, c->align, (size_t)b->n_align + 32));
    RET(ensure(c, c->bases, (size_t)b->n_bases + 32));
    RET(ensure(c, c->reads, (size_t)b->n_reads * sizeof(pgm_read)));
    if (with_raw && b->n_raw) CU(c, cudaMemcpyAsync(c->raw.p, b->raw, (size_t)b->n_raw * 2, cudaMemcpyHostToDevice, c->stream));
    if (b->n_align) CU(c, cudaMemcpyAsync(c->align.p, b->align, (size_t)b->n_align, cudaMemcpyHostToDevice, c->stream));
    if (b->n_bases) CU(c, cudaMemcpyAsync(c->bases.p, b->bases, (size_t)b->n_bases, cudaMemcpyHostToDevice, c->stream));
    if (b->n_reads) CU(c, cudaMemcpyAsync(c->reads.p, b->reads, (size_t)b->n_reads * sizeof(pgm_read), cudaMemcpyHostToDevice, c->stream));
    c->db.n_reads = b->n_reads; c->db.mode = b->mode; c->db.n_raw = b->n_raw; c->db.n_align = b->n_align; c->db.n_bases = b->n_bases;
    c->db.raw = P<int16_t>(c->raw); c->db.align = P<uint8_t>(c->align); c->db.bases = P<uint8_t>(c->bases); c->db.reads = P<pgm_read>(c->reads);
    c->attached = false;
    c->need_wide = false;
    c->svb_used = false;
    if (c->opts.scaling) {
        // in-range code span ~ (pa_max - pa_min) / |scale|; anything close to the shared-memory budget goes the wide way
        const double span = c->opts.pa_max - c->opts.pa_min;
        for (uint32_t i = 0; i < b->n_reads && !c->need_wide; i++) {
            const double sc = fabs(b->reads[i].range / b->reads[i].digitisation);
            if (!(span / sc < (double)(SIG_SMEM_BINS - 8))) c->need_wide = true;
        }
    }
    c->state = ST_LOADED;
    return PGM_OK;
}

int pgm_upload(pgm_ctx *c, const pgm_batch *b) {
    PHASE("pgm_upload");
    if (!c || !b) return PGM_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    return upload_common(c, b, true);
}

int pgm_upload_svb(pgm_ctx *c, const pgm_batch *b, const pgm_svb *svb) {
    PHASE("pgm_upload_svb");
    if (!c || !b || !svb) return PGM_ERR_ARG;
    if (b->raw) return fail(c, PGM_ERR_ARG, "pgm_upload_svb: batch.raw must be NULL (the samples are decoded on the device)");
    if (b->n_reads && (!svb->off || (svb->n_bytes && !svb->bytes))) return fail(c, PGM_ERR_ARG, "null array in pgm_svb");
    CU(c, cudaSetDevice(c->device));
    pgm_batch hb = *b;
    static const int16_t dummy = 0;
    if (hb.n_raw) hb.raw = &dummy;     // prepare_batch only checks for presence
    RET(upload_common(c, &hb, false));
    c->state = ST_IDLE;
    const uint32_t R = b->n_reads;
    RET(ensure(c, c->svb_blob, (size_t)svb->n_bytes + 64));
    RET(ensure(c, c->svb_off, (size_t)(R + 1) * 8));
    RET(ensure(c, c->svb_words, 16));
    RET(ensure_host(c, c->h_svb_err, 16));
    cudaStream_t st = c->stream;
    if (svb->n_bytes) CU(c, cudaMemcpyAsync(c->svb_blob.p, svb->bytes, (size_t)svb->n_bytes, cudaMemcpyHostToDevice, st));
    CU(c, cudaMemsetAsync(c->svb_words.p, 0, 16, st));
    if (R) {
        if (svb->off[R] > svb->n_bytes) return fail(c, PGM_ERR_ARG, "pgm_svb.off[n_reads] exceeds n_bytes");
        CU(c, cudaMemcpyAsync(c->svb_off.p, svb->off, (size_t)(R + 1) * 8, cudaMemcpyHostToDevice, st));
        StageTimer t(c, PGM_STAGE_SVB);
        SvbIO io;
        io.blob = P<uint8_t>(c->svb_blob); io.blob_off = P<unsigned long long>(c->svb_off); io.reads = P<pgm_read>(c->reads);
        io.raw = P<int16_t>(c->raw); io.n_raw = b->n_raw; io.n_reads = R;
        io.ticket = P<uint32_t>(c->svb_words); io.err = P<uint32_t>(c->svb_words) + 1;
        const uint32_t grid = (uint32_t)std::min<uint64_t>(cdiv(R, SVB_WARPS), (uint64_t)c->n_sm * 8);
        k_svb_decode<<<grid, SVB_WARPS * 32, 0, st>>>(io);
        c->launches++;
        CU(c, cudaGetLastError());
    }
    c->svb_used = true;
    c->state = ST_LOADED;
    return PGM_OK;
}

int pgm_attach_device(pgm_ctx *c, const pgm_batch *b) {
    if (!c || !b) return PGM_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    if (((uintptr_t)b->raw & 15) != 0) return fail(c, PGM_ERR_ARG, "device raw[] must be 16-byte aligned");
    RET(prepare_batch(c, b));
    c->db.n_reads = b->n_reads; c->db.mode = b->mode; c->db.n_raw = b->n_raw; c->db.n_align = b->n_align; c->db.n_bases = b->n_bases;
    c->db.raw = b->raw; c->db.align = b->align; c->db.bases = b->bases; c->db.reads = b->reads;
    c->attached = true;
    c->need_wide = true;
    c->state = ST_LOADED;
    return PGM_OK;
}

// status words + ticket of the next look-back scan of this batch (`tiles` = its grid size)
static LbBuf next_lb(pgm_ctx *c, size_t tiles) {
    LbBuf st;
    st.status = reinterpret_cast<unsigned long long *>(reinterpret_cast<uint8_t *>(c->zbatch.p) + ZB_STATUS_OFF) + c->lb_used;
    st.ticket = MISC(c) + 8 + c->lb_ticket++;
    c->lb_used += tiles + 1;
    return st;
}

static EvSource ev_source(pgm_ctx *c) {
    EvSource s;
    s.rows_tmp = P<uint4>(c->rows_tmp); s.ev_base = P<uint32_t>(c->ev_base); s.reads = c->db.reads;
    s.n_reads = c->db.n_reads; s.shift = c->tmp_shift;
    return s;
}

int pgm_count(pgm_ctx *c) {
    PHASE("pgm_count");
    if (!c) return PGM_ERR_ARG;
    if (c->state != ST_LOADED) return fail(c, PGM_ERR_STATE, "pgm_count needs pgm_upload/pgm_attach_device first");
    CU(c, cudaSetDevice(c->device));
    const uint32_t R = c->db.n_reads;
    cudaStream_t st = c->stream;
    const SortPlan sp = sort_plan(c);
    {   // the decoder reads the totals the newest k_quota of the family wrote (a sibling's batch in front of this one)
        pgm_ctx *root = c->root ? c->root : c;
        if (root->quota_recorded) CU(c, cudaStreamWaitEvent(st, root->ev_quota, 0));
    }
    {
        StageTimer t(c, PGM_STAGE_MISC);
        CU(c, cudaMemsetAsync(c->counts.p, 0, (size_t)c->opts.n_slots * 4, st));
        // scalars, tickets, scan status, read_has_sel[R] (the tile histogram behind them is written whole by k_tile_hist)
        CU(c, cudaMemsetAsync(c->zbatch.p, 0, c->hist_off, st));
        c->lb_ticket = 0;
        c->lb_used = 0;
        if (!R) CU(c, cudaMemsetAsync(P<uint32_t>(c->ev_base), 0, 4, st));
    }
    c->user_prior = c->user_consumed = nullptr; c->gathered = nullptr;
    DevOpts d = c->dopts;
    d.mode = c->db.mode;
    if (R) {
        {
            StageTimer t(c, PGM_STAGE_PREP);
            PrepOrder po;
            po.cls_count = MISC(c) + 96; po.packed = P<uint32_t>(c->rec_packed); po.order = P<uint32_t>(c->rec_order);
            k_read_prep<<<cdiv(R, PGM_BLOCK), PGM_BLOCK, 0, st>>>(d, c->db, P<ReadCal>(c->cal), P<uint8_t>(c->rflags), &SCALARS(c)->err_flag, po,
                                                                  P<unsigned long long>(c->rec_src0));
            k_order<<<cdiv(R, PGM_BLOCK), PGM_BLOCK, 0, st>>>(R, po);
            c->launches += 2;
            if (c->db.mode == PGM_MODE_MOVES_I8) { k_oor_check<<<R, PGM_BLOCK, 0, st>>>(c->db, P<ReadCal>(c->cal), P<uint8_t>(c->rflags)); c->launches++; }
        }
        {
            StageTimer t(c, PGM_STAGE_DECODE_COUNT);
            DecodeIO io;
            memset(&io, 0, sizeof io);
            io.consumed = P<uint32_t>(c->consumed);
            io.counts = P<uint32_t>(c->counts);
            io.ev_base = P<uint32_t>(c->ev_base);
            io.ev_cap = c->ev_cap;
            io.rflags = P<uint8_t>(c->rflags);
            io.examined = &SCALARS(c)->examined;
            io.err_flag = &SCALARS(c)->err_flag;
            io.rows_tmp = P<uint4>(c->rows_tmp); io.tmp_shift = c->tmp_shift;
            io.sc_start = P<uint32_t>(c->sc_start); io.sc_end = P<uint32_t>(c->sc_end);
            io.sc_code = P<uint8_t>(c->sc_code); io.sc_indel = P<int32_t>(c->sc_indel);
            io.n_pass = P<uint32_t>(c->n_pass);
            io.ticket = MISC(c) + 1; io.order = P<uint32_t>(c->rec_order);
            if (c->db.mode == PGM_MODE_SS) {
                const uint32_t want = cdiv(R, DEC_WARPS), full = (uint32_t)c->n_sm * (2048 / PGM_BLOCK);
                k_decode<<<want < full ? want : full, PGM_BLOCK, 0, st>>>(d, c->db, io);
            } else {   // move tables: one CTA per record at a time
                const uint32_t full = (uint32_t)c->n_sm * DM_MIN_CTAS;
                k_decode_moves<<<R < full ? R : full, DM_THREADS, 0, st>>>(d, c->db, io);
            }
            c->launches++;
            t.stop();
            StageTimer t2(c, PGM_STAGE_PLACE);
            k_scan_records<<<cdiv(R, LB_TILE), PGM_BLOCK, 0, st>>>(io.n_pass, io.ev_base, R, next_lb(c, cdiv(R, LB_TILE)));
            c->launches++;
            if (c->ev_cap) {   // tile histogram of the first digit, over the dense (record, position) order
                const uint32_t grid = std::min<uint32_t>(c->sort_ntiles, (uint32_t)c->n_sm * 8);
                k_tile_hist<<<grid, PGM_BLOCK, 0, st>>>(ev_source(c), P<uint32_t>(c->ev_base) + R, sp.first_shift, sp.first_bits, HIST(c));
                c->launches++;
            }
        }
        CU(c, cudaGetLastError());
    }
    c->state = ST_COUNTED;
    return PGM_OK;
}

int pgm_counts_device(pgm_ctx *c, uint32_t **counts_dev) {
    if (!c || !counts_dev) return PGM_ERR_ARG;
    *counts_dev = P<uint32_t>(c->counts);
    return PGM_OK;
}
int pgm_totals_device(pgm_ctx *c, uint32_t **consumed_dev) {
    if (!c || !consumed_dev) return PGM_ERR_ARG;
    *consumed_dev = P<uint32_t>(c->consumed);
    return PGM_OK;
}
int pgm_counts_host(pgm_ctx *c, uint32_t *out) {
    if (!c || !out) return PGM_ERR_ARG;
    if (c->state != ST_COUNTED) return fail(c, PGM_ERR_STATE, "pgm_counts_host needs pgm_count first");
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaMemcpyAsync(out, c->counts.p, (size_t)c->opts.n_slots * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    return PGM_OK;
}
int pgm_totals_host(pgm_ctx *c, uint32_t *out) {
    if (!c || !out) return PGM_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaMemcpyAsync(out, c->consumed.p, (size_t)c->opts.n_slots * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    return PGM_OK;
}
int pgm_set_prior_host(pgm_ctx *c, const uint32_t *prior, const uint32_t *consumed) {
    if (!c || !prior || !consumed) return PGM_ERR_ARG;
    if (c->state != ST_COUNTED) return fail(c, PGM_ERR_STATE, "pgm_set_prior_host belongs between pgm_count and pgm_emit");
    CU(c, cudaSetDevice(c->device));
    const size_t sb = (size_t)c->opts.n_slots * 4;
    RET(ensure(c, c->hp_prior, sb));
    RET(ensure(c, c->hp_consumed, sb));
    CU(c, cudaMemcpyAsync(c->hp_prior.p, prior, sb, cudaMemcpyHostToDevice, c->stream));
    CU(c, cudaMemcpyAsync(c->hp_consumed.p, consumed, sb, cudaMemcpyHostToDevice, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));   // the host arrays may go away
    c->user_prior = P<uint32_t>(c->hp_prior); c->user_consumed = P<uint32_t>(c->hp_consumed); c->gathered = nullptr;
    return PGM_OK;
}
void *pgm_alloc_pinned(size_t bytes) {
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) return nullptr;
    return p;
}
void pgm_free_pinned(void *p) { if (p) cudaFreeHost(p); }
int pgm_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

int pgm_warmup(int device) {
    if (cudaSetDevice(device) != cudaSuccess || cudaFree(nullptr) != cudaSuccess) { cudaGetLastError(); return PGM_ERR_CUDA; }
    return PGM_OK;
}
int pgm_set_gathered(pgm_ctx *c, const uint32_t *all_counts_dev, uint32_t n_ranks, uint32_t rank) {
    if (!c || !all_counts_dev || !n_ranks || rank >= n_ranks) return PGM_ERR_ARG;
    if (c->state != ST_COUNTED) return fail(c, PGM_ERR_STATE, "pgm_set_gathered belongs between pgm_count and pgm_emit");
    c->gathered = all_counts_dev; c->n_ranks = n_ranks; c->rank = rank;
    c->user_prior = nullptr; c->user_consumed = nullptr;
    return PGM_OK;
}

int pgm_set_prior(pgm_ctx *c, const uint32_t *prior_dev, const uint32_t *consumed_dev) {
    if (!c || !prior_dev || !consumed_dev) return PGM_ERR_ARG;
    if (c->state != ST_COUNTED) return fail(c, PGM_ERR_STATE, "pgm_set_prior belongs between pgm_count and pgm_emit");
    c->user_prior = prior_dev; c->user_consumed = consumed_dev; c->gathered = nullptr;
    return PGM_OK;
}


static TextIO make_text_io(pgm_ctx *c) {
    TextIO t;
    t.samples = P<double>(c->samples); t.sel = P<uint4>(c->sel); t.sel_off = P<unsigned long long>(c->sel_off);
    t.n_sel = P<uint32_t>(c->sel_base) + c->opts.n_slots;
    t.ev_tlen = P<uint32_t>(c->ev_tlen); t.ev_toff = P<unsigned long long>(c->ev_toff);
    t.text = P<char>(c->text); t.text_cap = c->text.cap;
    return t;
}

int pgm_emit(pgm_ctx *c) {
    PHASE("pgm_emit");
    if (!c) return PGM_ERR_ARG;
    if (c->state != ST_COUNTED) return fail(c, PGM_ERR_STATE, "pgm_emit needs pgm_count first");
    CU(c, cudaSetDevice(c->device));
    const uint32_t R = c->db.n_reads, NS = c->opts.n_slots;
    cudaStream_t st = c->stream;
    DevOpts d = c->dopts;
    d.mode = c->db.mode;
    const uint32_t *n_ev = P<uint32_t>(c->ev_base) + R;  // device scalar: passing events
    // The scan of the tile histogram (k_tile_hist) does not depend on the quota: it runs on the auxiliary stream while
    // k_quota runs here; both are a few microseconds of latency, not of bandwidth.
    const bool sorting = R && c->ev_cap;
    const SortPlan sp = sort_plan(c);
    const uint32_t first_bins = 1u << sp.first_bits;
    if (sorting) {
        CU(c, cudaEventRecord(c->ev_fork, st));
        CU(c, cudaStreamWaitEvent(c->aux, c->ev_fork, 0));
        StageTimer t(c, PGM_STAGE_SORT, c->aux);
        k_scan_hist<<<cdiv((uint64_t)first_bins * c->sort_ntiles, LB_TILE), PGM_BLOCK, 0, c->aux>>>(HIST(c), n_ev, first_bins,
                                                                                               next_lb(c, cdiv((uint64_t)first_bins * c->sort_ntiles, LB_TILE)));
        c->launches++;
        t.stop();
        CU(c, cudaEventRecord(c->ev_join, c->aux));
    }
    // quota: keep / bucket_base / sel_base / totals after the batch, one launch
    {
        StageTimer t(c, PGM_STAGE_SELECT);
        QuotaIO q;
        q.n_slots = NS; q.limit = c->opts.sample_limit;
        q.prior = c->user_prior ? c->user_prior : P<uint32_t>(c->consumed);
        q.user_consumed = c->user_consumed;
        q.gathered = c->gathered; q.n_ranks = c->n_ranks; q.rank = c->rank;
        q.counts = P<uint32_t>(c->counts); q.keep = P<uint32_t>(c->keep); q.consumed = P<uint32_t>(c->consumed);
        q.bucket_base = P<uint32_t>(c->bucket_base); q.sel_base = P<uint32_t>(c->sel_base); q.snapshot = P<uint32_t>(c->totals_snap);
        k_quota<<<cdiv(NS, LB_TILE), PGM_BLOCK, 0, st>>>(q, next_lb(c, cdiv(NS, LB_TILE)));
        c->launches++;
        pgm_ctx *root = c->root ? c->root : c;
        if (root->ev_quota) { CU(c, cudaEventRecord(root->ev_quota, st)); root->quota_recorded = true; }
    }
    if (sorting) {
        // stable bucketing by slot fused with the selection (pgm_sort.cuh)
        CU(c, cudaStreamWaitEvent(st, c->ev_join, 0));
        StageTimer tsort(c, PGM_STAGE_SORT);
        const uint32_t ntiles = c->sort_ntiles;
        uint8_t *has_sel = c->opts.scaling ? HAS_SEL(c) : nullptr;
        if (!sp.hi_bits) {
            RankIO r;
            r.n_events = n_ev; r.src = ev_source(c); r.hist_scanned = HIST(c); r.bits = sp.bits;
            r.bucket_base = P<uint32_t>(c->bucket_base); r.keep = P<uint32_t>(c->keep); r.sel_base = P<uint32_t>(c->sel_base);
            r.sel = P<uint4>(c->sel); r.read_has_sel = has_sel;
            k_rank<<<std::min<uint32_t>(ntiles, (uint32_t)c->n_sm * 2), RANK_BLOCK, (RANK_WARPS + 1) * (1u << sp.bits) * 4, st>>>(r);
            c->launches++;
        } else {
            MsdScatterIO ms;
            ms.n_events = n_ev; ms.src = ev_source(c); ms.hist_scanned = HIST(c);
            ms.shift = sp.lo_bits; ms.bits = sp.hi_bits; ms.rows_out = P<uint4>(c->rows_a);
            k_msd_scatter<<<std::min<uint32_t>(ntiles, (uint32_t)c->n_sm * 2), RANK_BLOCK, MSD_SCATTER_SMEM(sp.hi_bits), st>>>(ms);
            MsdRankIO mr;
            mr.n_events = n_ev; mr.rows_in = P<uint4>(c->rows_a); mr.hist_scanned = HIST(c);
            mr.hi_bits = sp.hi_bits; mr.lo_bits = sp.lo_bits; mr.n_slots = NS;
            mr.keep = P<uint32_t>(c->keep); mr.sel_base = P<uint32_t>(c->sel_base);
            mr.sel = P<uint4>(c->sel); mr.read_has_sel = has_sel;
            mr.ticket = MISC(c) + 4;
            const uint32_t n_buckets = cdiv(NS, 1u << sp.lo_bits);
            k_msd_rank<<<std::min<uint32_t>(n_buckets, (uint32_t)c->n_sm * MSD_MIN_CTAS), MSD_BLOCK, MSD_RANK_SMEM(sp.lo_bits), st>>>(mr);
            c->launches += 2;
        }
        CU(c, cudaGetLastError());
    }
    // Sample offsets of the selected events: needed by k_gather, not by the statistics, so with scaling the scan runs
    // beside k_stats.
    const bool beside = sorting && c->opts.scaling;
    {
        cudaStream_t ss = beside ? c->aux : st;
        if (beside) {
            CU(c, cudaEventRecord(c->ev_fork, st));
            CU(c, cudaStreamWaitEvent(c->aux, c->ev_fork, 0));
        }
        StageTimer t(c, PGM_STAGE_SCAN, ss);
        k_scan_sel<<<cdiv((uint64_t)c->ev_cap + 1, LB_TILE), PGM_BLOCK, 0, ss>>>(P<uint4>(c->sel), P<unsigned long long>(c->sel_off),
                                                                               P<uint32_t>(c->sel_base) + NS, n_ev, SCALARS(c),
                                                                               next_lb(c, cdiv((uint64_t)c->ev_cap + 1, LB_TILE)));
        c->launches++;
        t.stop();
        if (beside) CU(c, cudaEventRecord(c->ev_join, c->aux));
    }
    if (R && c->ev_cap) {
        if (c->opts.scaling) {
            StageTimer t(c, PGM_STAGE_STATS);
            if (c->need_wide) {
                k_stats_wide<<<c->n_sm * 4, PGM_BLOCK, 0, st>>>(d, c->db, P<uint8_t>(c->rflags), HAS_SEL(c), P<ReadCal>(c->cal), MISC(c),
                                                                P<uint32_t>(c->gscratch));
                c->launches++;
            }
            StatsIO si;
            si.rflags = P<uint8_t>(c->rflags); si.read_has_sel = HAS_SEL(c); si.cal = P<ReadCal>(c->cal); si.ticket = MISC(c) + 2; si.order = P<uint32_t>(c->rec_order);
            const uint32_t sgrid = (uint32_t)std::min<uint64_t>(cdiv(R, STATS_WARPS), (uint64_t)c->n_sm * 3);
            k_stats<<<sgrid, STATS_WARPS * 32, STATS_DYN_SMEM, st>>>(d, c->db, si);
            c->launches++;
        }
        if (beside) CU(c, cudaStreamWaitEvent(st, c->ev_join, 0));
        StageTimer t(c, PGM_STAGE_GATHER);
        GatherIO g;
        g.sel = P<uint4>(c->sel);
        g.sel_off = P<uint64_t>(c->sel_off); g.n_sel = P<uint32_t>(c->sel_base) + NS; g.cal = P<ReadCal>(c->cal);
        g.samples = P<double>(c->samples); g.samples_cap = c->samples_cap;
        if (!c->gather_ctas) {   // resident CTAs per SM of the instantiation in use (registers and shared memory decide)
            int nb = 0;
            const size_t sm = GATHER_SMEM;
            if (c->opts.scaling) CU(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_gather<true>, GATHER_WARPS * 32, sm));
            else CU(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_gather<false>, GATHER_WARPS * 32, sm));
            c->gather_ctas = nb > 0 ? (uint32_t)nb : 1u;
        }
        const uint32_t ggrid = (uint32_t)c->n_sm * c->gather_ctas;
        if (c->opts.scaling) k_gather<true><<<ggrid, GATHER_WARPS * 32, GATHER_SMEM, st>>>(d, c->db, g);
        else k_gather<false><<<ggrid, GATHER_WARPS * 32, GATHER_SMEM, st>>>(d, c->db, g);
        c->launches++;
    }
    if (c->opts.output != PGM_OUT_F64) {
        StageTimer t(c, PGM_STAGE_TEXT);
        TextIO tio = make_text_io(c);
        k_text_len<<<c->n_sm * 8, PGM_BLOCK, 0, st>>>(tio);
        k_scan_text<<<cdiv((uint64_t)c->ev_cap + 1, LB_TILE), PGM_BLOCK, 0, st>>>(tio.ev_tlen, tio.ev_toff, tio.n_sel, &SCALARS(c)->n_text,
                                                                                next_lb(c, cdiv((uint64_t)c->ev_cap + 1, LB_TILE)));
        c->launches += 2;
        // The exact text size is a device-side number. When the buffer of an earlier batch is still around, write
        // into it right away (the kernel drops what does not fit): pgm_download then only has to check that the
        // text fitted, and the formatting overlaps whatever the host does before it calls pgm_download.
        c->text_written_cap = 0;
        if (c->text.cap) {
            k_text_write<<<c->n_sm * 8, PGM_BLOCK, 0, st>>>(tio);
            k_slot_text_off<<<cdiv((uint64_t)NS + 1, PGM_BLOCK), PGM_BLOCK, 0, st>>>(P<uint32_t>(c->sel_base), tio.ev_toff, NS, P<unsigned long long>(c->slot_toff));
            c->launches += 2;
            c->text_written_cap = c->text.cap;
        }
    }
    CU(c, cudaGetLastError());
    c->model_logged = false;
    c->state = ST_EMITTED;
    return PGM_OK;
}

// K7: appends the emitted batch's selected samples (as the integers their "%.8f" text spells) and event lengths to the
// model log. Everything is enqueued on the context's stream; the host tables are valid after the next synchronisation.
static int model_log_batch(pgm_ctx *c, size_t ne, size_t nsamp) {
    cudaStream_t st = c->stream;
    const uint32_t NS = c->opts.n_slots;
    if (ne) {   // the rows as arrays (the dwell log wants the lengths)
        RET(ensure(c, c->sel_start, ne * 4)); RET(ensure(c, c->sel_len, ne * 4)); RET(ensure(c, c->sel_read, ne * 4));
        k_unpack_sel<<<c->n_sm * 4, PGM_BLOCK, 0, st>>>(P<uint4>(c->sel), P<uint32_t>(c->sel_base) + NS, P<unsigned long long>(c->rec_src0),
                                                       P<uint32_t>(c->sel_start), P<uint32_t>(c->sel_len), P<uint32_t>(c->sel_read));
        c->launches++;
    }
    pgm_ctx::ModelBatch mb;
    RET(arena_reserve(c, c->model_vals, c->model_vals_used, nsamp ? nsamp : 1, 8));
    RET(arena_reserve(c, c->model_lens, c->model_lens_used, ne ? ne : 1, 4));
    RET(ensure(c, c->model_tab, (size_t)(NS + 1) * 8));
    mb.vals_off = c->model_vals_used; mb.n_vals = nsamp; mb.lens_off = c->model_lens_used; mb.n_ev = ne;
    mb.h_slot_off.resize(NS + 1);
    mb.h_ev_off.resize(NS + 1);
    k_model_scale<<<c->n_sm * 8, PGM_BLOCK, 0, st>>>(P<double>(c->samples), &SCALARS(c)->n_sel_samples, P<long long>(c->model_vals) + mb.vals_off, P<uint32_t>(c->model_bad));
    k_slot_text_off<<<cdiv((uint64_t)NS + 1, PGM_BLOCK), PGM_BLOCK, 0, st>>>(P<uint32_t>(c->sel_base), P<unsigned long long>(c->sel_off), NS, P<unsigned long long>(c->model_tab));
    c->launches += 2;
    CU(c, cudaGetLastError());
    CU(c, cudaMemcpyAsync(mb.h_slot_off.data(), c->model_tab.p, (size_t)(NS + 1) * 8, cudaMemcpyDeviceToHost, st));
    if (ne) CU(c, cudaMemcpyAsync(P<uint32_t>(c->model_lens) + mb.lens_off, c->sel_len.p, ne * 4, cudaMemcpyDeviceToDevice, st));
    CU(c, cudaMemcpyAsync(mb.h_ev_off.data(), c->sel_base.p, (size_t)(NS + 1) * 4, cudaMemcpyDeviceToHost, st));
    c->model_vals_used += nsamp; c->model_lens_used += ne;
    c->model_log.push_back(std::move(mb));
    c->model_logged = true;
    return PGM_OK;
}

int pgm_download(pgm_ctx *c, pgm_result *out) {
    PHASE("pgm_download");
    if (!c || !out) return PGM_ERR_ARG;
    if (c->state != ST_EMITTED) return fail(c, PGM_ERR_STATE, "pgm_download needs pgm_emit first");
    CU(c, cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    const uint32_t R = c->db.n_reads, NS = c->opts.n_slots;
    CU(c, cudaMemcpyAsync(c->h_scalars, SCALARS(c), sizeof(Scalars), cudaMemcpyDeviceToHost, st));
    if (c->svb_used) CU(c, cudaMemcpyAsync(c->h_svb_err.p, P<uint32_t>(c->svb_words) + 1, 4, cudaMemcpyDeviceToHost, st));
    RET(ensure_host(c, c->h_slot_sel, (size_t)NS * 4));
    RET(ensure_host(c, c->h_slot_total, (size_t)NS * 4));
    RET(ensure_host(c, c->h_rflags, R ? R : 1));
    CU(c, cudaMemcpyAsync(c->h_slot_sel.p, c->keep.p, (size_t)NS * 4, cudaMemcpyDeviceToHost, st));
    CU(c, cudaMemcpyAsync(c->h_slot_total.p, c->totals_snap.p, (size_t)NS * 4, cudaMemcpyDeviceToHost, st));
    if (R) CU(c, cudaMemcpyAsync(c->h_rflags.p, c->rflags.p, R, cudaMemcpyDeviceToHost, st));
    CU(c, cudaStreamSynchronize(st));
    const Scalars s = *c->h_scalars;
    if (s.err_flag & 2u) return fail(c, PGM_ERR_INPUT, "an event is longer than %u samples (max_dur / --margin too large for the event rows)", EV_MAX_LEN);
    if (s.err_flag) return fail(c, PGM_ERR_INPUT, "batch has records whose offsets fall outside the arrays (or ss records with a trim offset)");
    if (c->svb_used && *P<uint32_t>(c->h_svb_err))
        return fail(c, PGM_ERR_INPUT, "svb-zd stream of a record does not decode to its raw_len samples");
    if (s.n_pass > c->ev_cap || s.n_sel_samples > c->samples_cap)
        return fail(c, PGM_ERR_INPUT, "internal capacity exceeded (events %u/%u, samples %llu/%llu)", s.n_pass, c->ev_cap,
                    s.n_sel_samples, (unsigned long long)c->samples_cap);
    const size_t ne = s.n_sel, nsamp = (size_t)s.n_sel_samples;
    RET(ensure_host(c, c->h_ev_len, ne * 4 + 4));
    RET(ensure_host(c, c->h_ev_read, ne * 4 + 4));
    RET(ensure_host(c, c->h_ev_start, ne * 4 + 4));
    const bool as_text = c->opts.output != PGM_OUT_F64;
    const bool with_events = c->opts.output != PGM_OUT_TEXT_ONLY;   // per-event arrays travel to the host
    if (!as_text) RET(ensure_host(c, c->h_samples, nsamp * 8 + 8));
    if (ne && with_events) {   // the rows as the three arrays of pgm_result
        RET(ensure(c, c->sel_start, ne * 4)); RET(ensure(c, c->sel_len, ne * 4)); RET(ensure(c, c->sel_read, ne * 4));
        k_unpack_sel<<<c->n_sm * 4, PGM_BLOCK, 0, st>>>(P<uint4>(c->sel), P<uint32_t>(c->sel_base) + NS, P<unsigned long long>(c->rec_src0),
                                                       P<uint32_t>(c->sel_start), P<uint32_t>(c->sel_len), P<uint32_t>(c->sel_read));
        c->launches++;
        CU(c, cudaGetLastError());
    }
    if (ne && with_events) {
        CU(c, cudaMemcpyAsync(c->h_ev_len.p, c->sel_len.p, ne * 4, cudaMemcpyDeviceToHost, st));
        CU(c, cudaMemcpyAsync(c->h_ev_read.p, c->sel_read.p, ne * 4, cudaMemcpyDeviceToHost, st));
        CU(c, cudaMemcpyAsync(c->h_ev_start.p, c->sel_start.p, ne * 4, cudaMemcpyDeviceToHost, st));
    }
    if (nsamp && !as_text) CU(c, cudaMemcpyAsync(c->h_samples.p, c->samples.p, nsamp * 8, cudaMemcpyDeviceToHost, st));
    if (as_text) {   // the length is known now: format into an exactly sized buffer, then bring it over
        const size_t nt = (size_t)s.n_text;
        RET(ensure_host(c, c->h_text, nt + 16));
        RET(ensure_host(c, c->h_slot_toff, (size_t)(NS + 1) * 8));
        RET(ensure_host(c, c->h_ev_toff, (ne + 1) * 8));
        if (nt + 16 > c->text_written_cap) {   // first batch, or the text outgrew the buffer pgm_emit wrote into
            RET(ensure(c, c->text, nt + nt / 8 + 16));
            TextIO tio = make_text_io(c);
            if (ne) { k_text_write<<<c->n_sm * 8, PGM_BLOCK, 0, st>>>(tio); c->launches++; }
            k_slot_text_off<<<cdiv((uint64_t)NS + 1, PGM_BLOCK), PGM_BLOCK, 0, st>>>(P<uint32_t>(c->sel_base), tio.ev_toff, NS, P<unsigned long long>(c->slot_toff));
            c->launches++;
            CU(c, cudaGetLastError());
        }
        if (nt) CU(c, cudaMemcpyAsync(c->h_text.p, c->text.p, nt, cudaMemcpyDeviceToHost, st));
        CU(c, cudaMemcpyAsync(c->h_slot_toff.p, c->slot_toff.p, (size_t)(NS + 1) * 8, cudaMemcpyDeviceToHost, st));
        if (with_events) CU(c, cudaMemcpyAsync(c->h_ev_toff.p, c->ev_toff.p, (ne + 1) * 8, cudaMemcpyDeviceToHost, st));
    }
    if (c->model_on && !c->model_logged) RET(model_log_batch(c, ne, nsamp));
    CU(c, cudaStreamSynchronize(st));
    memset(out, 0, sizeof *out);
    out->n_slots = NS; out->n_reads = R;
    out->n_events_examined = s.examined; out->n_events_pass = s.n_pass;
    out->n_sel_events = ne; out->n_sel_samples = nsamp;
    out->slot_sel = P<uint32_t>(c->h_slot_sel); out->slot_total = P<uint32_t>(c->h_slot_total);
    if (with_events) { out->ev_len = P<uint32_t>(c->h_ev_len); out->ev_read = P<uint32_t>(c->h_ev_read); out->ev_start = P<uint32_t>(c->h_ev_start); }
    out->samples = as_text ? nullptr : P<double>(c->h_samples); out->read_flags = P<uint8_t>(c->h_rflags);
    out->text = as_text ? P<char>(c->h_text) : nullptr; out->slot_text_off = as_text ? P<uint64_t>(c->h_slot_toff) : nullptr;
    out->ev_text_off = (as_text && with_events) ? P<uint64_t>(c->h_ev_toff) : nullptr;
    c->state = ST_IDLE;
    return guard_check(c, "pgm_download");
}

int pgm_summary(pgm_ctx *c, pgm_result *out) {
    if (!c || !out) return PGM_ERR_ARG;
    if (c->state != ST_EMITTED) return fail(c, PGM_ERR_STATE, "pgm_summary needs pgm_emit first");
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaMemcpyAsync(c->h_scalars, SCALARS(c), sizeof(Scalars), cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    const Scalars s = *c->h_scalars;
    memset(out, 0, sizeof *out);
    out->n_slots = c->opts.n_slots; out->n_reads = c->db.n_reads;
    out->n_events_examined = s.examined; out->n_events_pass = s.n_pass;
    out->n_sel_events = s.n_sel; out->n_sel_samples = s.n_sel_samples;
    if (s.err_flag) return fail(c, PGM_ERR_INPUT, "batch has records whose offsets fall outside the arrays (or ss records with a trim offset)");
    return guard_check(c, "pgm_summary");
}

static const char *const kStageNames[PGM_N_STAGES] = {"prep", "decode", "scan", "place", "sort", "select", "stats", "gather", "misc", "text", "svb"};

const char *pgm_stage_name(int stage) { return (stage >= 0 && stage < PGM_N_STAGES) ? kStageNames[stage] : "?"; }

int pgm_profile_enable(pgm_ctx *c, int on) {
    if (!c) return PGM_ERR_ARG;
    c->prof = on != 0;
    c->prof_only = (on >= 2 && on - 2 < PGM_N_STAGES) ? on - 2 : -1;
    return PGM_OK;
}

int pgm_profile_read(pgm_ctx *c, float *ms, uint32_t *launches) {
    if (!c || !ms || !launches) return PGM_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaStreamSynchronize(c->stream));
    for (int i = 0; i < PGM_N_STAGES; i++) { ms[i] = 0.f; launches[i] = 0; }
    for (auto &sp : c->spans) {
        float t = 0.f;
        if (cudaEventElapsedTime(&t, sp.a, sp.b) == cudaSuccess) { ms[sp.stage] += t; launches[sp.stage] += sp.launches; }
        c->ev_pool.push_back(sp.a);
        c->ev_pool.push_back(sp.b);
    }
    c->spans.clear();
    return PGM_OK;
}

static void model_clear(pgm_ctx *c) {
    c->model_log.clear();
    c->model_vals_used = c->model_lens_used = 0;
}

int pgm_model_begin(pgm_ctx *c) {
    if (!c) return PGM_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaStreamSynchronize(c->stream));
    model_clear(c);
    RET(ensure(c, c->model_bad, 4));
    CU(c, cudaMemsetAsync(c->model_bad.p, 0, 4, c->stream));
    c->model_on = true;
    return PGM_OK;
}

int pgm_model_finalize(pgm_ctx *c, double stddev_limit, const pgm_model_row **rows_out) {
    PHASE("pgm_model_finalize");
    if (!c || !rows_out) return PGM_ERR_ARG;
    if (!c->model_on) return fail(c, PGM_ERR_STATE, "pgm_model_finalize needs pgm_model_begin first");
    static_assert(sizeof(ModelRowDev) == sizeof(pgm_model_row), "device and ABI rows must agree");
    CU(c, cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    const uint32_t NS = c->opts.n_slots;
    // where every batch's slot range goes in the grouped array: file order = batch order inside a slot
    std::vector<unsigned long long> off(NS + 1, 0);
    for (auto &mb : c->model_log)
        for (uint32_t s = 0; s < NS; s++) off[s + 1] += mb.h_slot_off[s + 1] - mb.h_slot_off[s];
    for (uint32_t s = 0; s < NS; s++) off[s + 1] += off[s];
    const unsigned long long total = off[NS];
    RET(ensure(c, c->model_off, (size_t)(NS + 1) * 8));
    RET(ensure(c, c->model_rows, (size_t)NS * sizeof(ModelRowDev)));
    CU(c, cudaMemcpyAsync(c->model_off.p, off.data(), (size_t)(NS + 1) * 8, cudaMemcpyHostToDevice, st));
    const long long *grouped = nullptr;
    std::vector<unsigned long long> dst;   // must outlive the copies below
    if (c->model_log.size() == 1) grouped = P<long long>(c->model_vals) + c->model_log[0].vals_off;   // one batch: its order is the grouped order
    else if (total) {
        RET(ensure(c, c->model_grouped, (size_t)total * 8));
        RET(ensure(c, c->model_dst, (size_t)(NS + 1) * 8 * 2 * c->model_log.size()));
        // per batch: [its slot table | where each of its slots goes]
        dst.assign((size_t)(2 * NS + 2) * c->model_log.size(), 0);
        std::vector<unsigned long long> cur(off.begin(), off.end() - 1);
        for (size_t bi = 0; bi < c->model_log.size(); bi++) {
            unsigned long long *tab = dst.data() + bi * (2 * NS + 2), *to = tab + NS + 1;
            memcpy(tab, c->model_log[bi].h_slot_off.data(), (size_t)(NS + 1) * 8);
            for (uint32_t s = 0; s < NS; s++) {
                to[s] = cur[s];
                cur[s] += tab[s + 1] - tab[s];
            }
        }
        CU(c, cudaMemcpyAsync(c->model_dst.p, dst.data(), dst.size() * 8, cudaMemcpyHostToDevice, st));
        for (size_t bi = 0; bi < c->model_log.size(); bi++) {
            const unsigned long long *tab = P<unsigned long long>(c->model_dst) + bi * (2 * NS + 2);
            k_model_group<<<c->n_sm * 4, PGM_BLOCK, 0, st>>>(P<long long>(c->model_vals) + c->model_log[bi].vals_off, tab, tab + NS + 1, NS, P<long long>(c->model_grouped));
            c->launches++;
        }
        grouped = P<long long>(c->model_grouped);
    }
    k_model_rows<<<NS < (uint32_t)c->n_sm * 8 ? NS : (uint32_t)c->n_sm * 8, PGM_BLOCK, 0, st>>>(grouped, P<unsigned long long>(c->model_off), NS, stddev_limit,
                                                                                              P<ModelRowDev>(c->model_rows));
    c->launches++;
    CU(c, cudaGetLastError());
    c->h_model_rows.resize(NS);
    CU(c, cudaMemcpyAsync(c->h_model_rows.data(), c->model_rows.p, (size_t)NS * sizeof(pgm_model_row), cudaMemcpyDeviceToHost, st));
    CU(c, cudaStreamSynchronize(st));
    *rows_out = c->h_model_rows.data();
    return guard_check(c, "pgm_model_finalize");
}

int pgm_model_log(pgm_ctx *c) {
    PHASE("pgm_model_log");
    if (!c) return PGM_ERR_ARG;
    if (!c->model_on) return fail(c, PGM_ERR_STATE, "pgm_model_log needs pgm_model_begin first");
    if (c->state != ST_EMITTED) return fail(c, PGM_ERR_STATE, "pgm_model_log needs pgm_emit first");
    if (c->model_logged) return PGM_OK;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaMemcpyAsync(c->h_scalars, SCALARS(c), sizeof(Scalars), cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    const Scalars s = *c->h_scalars;
    if (s.err_flag) return fail(c, PGM_ERR_INPUT, "the batch is flawed (err_flag %u): see pgm_download", s.err_flag);
    RET(model_log_batch(c, s.n_sel, (size_t)s.n_sel_samples));
    CU(c, cudaStreamSynchronize(c->stream));
    return PGM_OK;
}

int pgm_model_log_size(pgm_ctx *c, uint32_t *n_batches) {
    if (!c || !n_batches) return PGM_ERR_ARG;
    *n_batches = (uint32_t)c->model_log.size();
    return PGM_OK;
}

int pgm_model_log_entry(pgm_ctx *c, uint32_t i, const int64_t **vals_dev, const uint64_t **slot_off_host) {
    if (!c || !vals_dev || !slot_off_host || i >= c->model_log.size()) return PGM_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaStreamSynchronize(c->stream));
    static_assert(sizeof(unsigned long long) == sizeof(uint64_t) && sizeof(long long) == sizeof(int64_t), "64-bit types");
    *vals_dev = reinterpret_cast<const int64_t *>(P<long long>(c->model_vals) + c->model_log[i].vals_off);
    *slot_off_host = reinterpret_cast<const uint64_t *>(c->model_log[i].h_slot_off.data());
    return PGM_OK;
}

int pgm_model_log_append(pgm_ctx *c, const int64_t *vals_dev, uint64_t n_vals, const uint64_t *slot_off_host) {
    if (!c || !slot_off_host || (n_vals && !vals_dev)) return PGM_ERR_ARG;
    if (!c->model_on) return fail(c, PGM_ERR_STATE, "pgm_model_log_append needs pgm_model_begin first");
    const uint32_t NS = c->opts.n_slots;
    if (slot_off_host[NS] - slot_off_host[0] != n_vals) return fail(c, PGM_ERR_ARG, "pgm_model_log_append: the slot table does not add up to n_vals");
    CU(c, cudaSetDevice(c->device));
    RET(arena_reserve(c, c->model_vals, c->model_vals_used, n_vals ? n_vals : 1, 8));
    pgm_ctx::ModelBatch mb;
    mb.vals_off = c->model_vals_used; mb.n_vals = n_vals; mb.lens_off = c->model_lens_used; mb.n_ev = 0;
    mb.h_slot_off.resize(NS + 1);
    for (uint32_t s = 0; s <= NS; s++) mb.h_slot_off[s] = slot_off_host[s] - slot_off_host[0];
    mb.h_ev_off.assign(NS + 1, 0);
    if (n_vals) CU(c, cudaMemcpyAsync(P<long long>(c->model_vals) + mb.vals_off, vals_dev, n_vals * 8, cudaMemcpyDeviceToDevice, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    c->model_vals_used += n_vals;
    c->model_log.push_back(std::move(mb));
    return PGM_OK;
}

int pgm_model_log_reserve(pgm_ctx *c, uint64_t n_vals_total) {
    if (!c) return PGM_ERR_ARG;
    if (!c->model_on) return fail(c, PGM_ERR_STATE, "pgm_model_log_reserve needs pgm_model_begin first");
    CU(c, cudaSetDevice(c->device));
    if (n_vals_total <= c->model_vals_used) return PGM_OK;
    return arena_reserve(c, c->model_vals, c->model_vals_used, n_vals_total - c->model_vals_used, 8);
}

int pgm_model_dwell(pgm_ctx *c, const int64_t **twice_median_out) {
    if (!c || !twice_median_out) return PGM_ERR_ARG;
    if (!c->model_on) return fail(c, PGM_ERR_STATE, "pgm_model_dwell needs pgm_model_begin first");
    CU(c, cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    const uint32_t NS = c->opts.n_slots;
    // per slot: two leading zeros (see k_dwell_group), then every batch's events in file order
    std::vector<unsigned long long> off(NS + 1, 0);
    for (uint32_t s = 0; s < NS; s++) {
        unsigned long long n = 2;
        for (auto &mb : c->model_log) n += mb.h_ev_off[s + 1] - mb.h_ev_off[s];
        off[s + 1] = off[s] + n;
    }
    const unsigned long long total = off[NS];
    RET(ensure(c, c->model_off, (size_t)(NS + 1) * 8));
    RET(ensure(c, c->model_rows, (size_t)NS * sizeof(ModelRowDev)));
    RET(ensure(c, c->model_grouped, (size_t)total * 8));
    const size_t nb = c->model_log.size();
    RET(ensure(c, c->model_dst, ((size_t)NS * 8 + (size_t)(NS + 1) * 4) * (nb + 1) + 64));
    CU(c, cudaMemcpyAsync(c->model_off.p, off.data(), (size_t)(NS + 1) * 8, cudaMemcpyHostToDevice, st));
    // device layout: nb destination tables (u64[NS]) followed by nb event tables (u32[NS+1])
    std::vector<unsigned long long> dst((size_t)NS * nb, 0), cur(NS);
    std::vector<uint32_t> evt((size_t)(NS + 1) * nb, 0);
    for (uint32_t s = 0; s < NS; s++) cur[s] = off[s] + 2;
    for (size_t bi = 0; bi < nb; bi++) {
        memcpy(evt.data() + bi * (NS + 1), c->model_log[bi].h_ev_off.data(), (size_t)(NS + 1) * 4);
        for (uint32_t s = 0; s < NS; s++) {
            dst[bi * NS + s] = cur[s];
            cur[s] += c->model_log[bi].h_ev_off[s + 1] - c->model_log[bi].h_ev_off[s];
        }
    }
    unsigned long long *d_dst = P<unsigned long long>(c->model_dst);
    uint32_t *d_evt = reinterpret_cast<uint32_t *>(d_dst + (size_t)NS * nb);
    if (nb) {
        CU(c, cudaMemcpyAsync(d_dst, dst.data(), dst.size() * 8, cudaMemcpyHostToDevice, st));
        CU(c, cudaMemcpyAsync(d_evt, evt.data(), evt.size() * 4, cudaMemcpyHostToDevice, st));
    }
    k_dwell_heads<<<cdiv(NS, PGM_BLOCK), PGM_BLOCK, 0, st>>>(P<unsigned long long>(c->model_off), NS, P<long long>(c->model_grouped));
    c->launches++;
    for (size_t bi = 0; bi < nb; bi++) {
        k_dwell_group<<<c->n_sm * 4, PGM_BLOCK, 0, st>>>(P<uint32_t>(c->model_lens) + c->model_log[bi].lens_off, d_evt + bi * (NS + 1), d_dst + bi * NS, NS,
                                                         P<long long>(c->model_grouped));
        c->launches++;
    }
    k_model_rows<<<NS < (uint32_t)c->n_sm * 8 ? NS : (uint32_t)c->n_sm * 8, PGM_BLOCK, 0, st>>>(P<long long>(c->model_grouped), P<unsigned long long>(c->model_off), NS,
                                                                                              (double)INFINITY, P<ModelRowDev>(c->model_rows));
    c->launches++;
    CU(c, cudaGetLastError());
    std::vector<pgm_model_row> rows(NS);
    CU(c, cudaMemcpyAsync(rows.data(), c->model_rows.p, (size_t)NS * sizeof(pgm_model_row), cudaMemcpyDeviceToHost, st));
    CU(c, cudaStreamSynchronize(st));
    c->h_dwell.resize(NS);
    for (uint32_t s = 0; s < NS; s++) c->h_dwell[s] = (off[s + 1] - off[s] > 2) ? rows[s].median_num : -1;   // no event: the file is empty
    *twice_median_out = c->h_dwell.data();
    return PGM_OK;
}

int pgm_debug_raw(pgm_ctx *c, int16_t *out, uint64_t n) {
    if (!c || (!out && n)) return PGM_ERR_ARG;
    if (c->state == ST_IDLE) return fail(c, PGM_ERR_STATE, "pgm_debug_raw needs a loaded batch");
    if (n > c->db.n_raw) return fail(c, PGM_ERR_ARG, "pgm_debug_raw: n exceeds the batch's n_raw");
    CU(c, cudaSetDevice(c->device));
    if (n) CU(c, cudaMemcpyAsync(out, c->db.raw, (size_t)n * 2, cudaMemcpyDeviceToHost, c->stream));
    if (c->svb_used) CU(c, cudaMemcpyAsync(c->h_svb_err.p, P<uint32_t>(c->svb_words) + 1, 4, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    if (c->svb_used && *P<uint32_t>(c->h_svb_err))
        return fail(c, PGM_ERR_INPUT, "svb-zd stream of a record does not decode to its raw_len samples");
    return PGM_OK;
}

int pgm_debug_format(pgm_ctx *c, const double *vals, uint32_t n, char *text_out, uint64_t cap, uint64_t *len_out) {
    if (!c || !len_out || (n && !vals)) return PGM_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    double *dx = nullptr; char *dt = nullptr; unsigned long long *dl = nullptr;
    CU(c, cudaMalloc(&dx, (size_t)n * 8 + 8));
    CU(c, cudaMalloc(&dl, 8));
    CU(c, cudaMemcpyAsync(dx, vals, (size_t)n * 8, cudaMemcpyHostToDevice, c->stream));
    k_text_debug<<<1, 32, 0, c->stream>>>(dx, n, nullptr, dl);
    unsigned long long len = 0;
    CU(c, cudaMemcpyAsync(&len, dl, 8, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    *len_out = len;
    if (text_out && len <= cap && len) {
        CU(c, cudaMalloc(&dt, (size_t)len + 8));
        k_text_debug<<<1, 32, 0, c->stream>>>(dx, n, dt, dl);
        CU(c, cudaMemcpyAsync(text_out, dt, (size_t)len, cudaMemcpyDeviceToHost, c->stream));
        CU(c, cudaStreamSynchronize(c->stream));
        cudaFree(dt);
    }
    cudaFree(dx); cudaFree(dl);
    return PGM_OK;
}

int pgm_create_sibling(pgm_ctx *parent, void *stream, pgm_ctx **out) {
    if (!parent || !out) return PGM_ERR_ARG;
    *out = nullptr;
    pgm_ctx *root = parent->root ? parent->root : parent;
    pgm_ctx *k = nullptr;
    const int rc = pgm_create(&root->opts, root->device, stream, &k);
    if (rc != PGM_OK) return fail(parent, rc, "pgm_create_sibling: %s", g_create_error.c_str());
    release(k->consumed);
    k->consumed = root->consumed;            // one kmer_frequency_map for the family
    k->root = root;
    *out = k;
    return PGM_OK;
}

#define PGM_PIPE_DEPTH 2
int pgm_submit(pgm_ctx *c, const pgm_batch *host_batch, const pgm_svb *svb, uint64_t *batch_id) {
    PHASE("pgm_submit");
    if (!c || !host_batch) return PGM_ERR_ARG;
    if (c->root) return fail(c, PGM_ERR_STATE, "pgm_submit works on the root context of a family");
    if (c->model_on) return fail(c, PGM_ERR_STATE, "pgm_submit does not feed the model log: use the split-phase calls after pgm_model_begin");
    if (c->submitted - c->waited >= PGM_PIPE_DEPTH) return fail(c, PGM_ERR_STATE, "%d batches in flight: pgm_wait for batch %llu first", PGM_PIPE_DEPTH, (unsigned long long)c->waited);
    CU(c, cudaSetDevice(c->device));
    while (c->pipe.size() < PGM_PIPE_DEPTH) {
        pgm_ctx *k = nullptr;
        RET(pgm_create_sibling(c, nullptr, &k));
        c->pipe.push_back(k);
    }
    pgm_ctx *k = c->pipe[c->submitted % PGM_PIPE_DEPTH];
    int rc = svb ? pgm_upload_svb(k, host_batch, svb) : pgm_upload(k, host_batch);
    if (rc == PGM_OK) rc = pgm_count(k);
    if (rc == PGM_OK) rc = pgm_emit(k);
    if (rc != PGM_OK) return fail(c, rc, "%s", k->err.c_str());
    c->launches = 0;
    for (pgm_ctx *q : c->pipe) c->launches += q->launches;
    if (batch_id) *batch_id = c->submitted;
    c->submitted++;
    return PGM_OK;
}

int pgm_wait(pgm_ctx *c, uint64_t batch_id, pgm_result *out) {
    PHASE("pgm_wait");
    if (!c || !out) return PGM_ERR_ARG;
    if (batch_id != c->waited || c->waited >= c->submitted) return fail(c, PGM_ERR_STATE, "pgm_wait takes the batches in the order they were submitted (next: %llu)", (unsigned long long)c->waited);
    pgm_ctx *k = c->pipe[batch_id % PGM_PIPE_DEPTH];
    const int rc = pgm_download(k, out);
    c->waited++;
    if (rc != PGM_OK) return fail(c, rc, "%s", k->err.c_str());
    return PGM_OK;
}

int pgm_process(pgm_ctx *c, const pgm_batch *host_batch, pgm_result *out) {
    if (!c || !host_batch || !out) return PGM_ERR_ARG;
    RET(pgm_upload(c, host_batch));
    RET(pgm_count(c));
    RET(pgm_emit(c));
    return pgm_download(c, out);
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// pgm_format_noalloc: the dump text of n doubles without any allocation (include/pgm_gmove.h). No counterpart in the
// reference; added for callers that cannot call cudaMalloc (CUDA graph capture,
// tight serving loops). A variant of k_scan_text (pgm_text.cuh): the same single-pass look-back scan, with the length
// of a value computed in the scan's load step and its characters written in the store step, and with the scan's
// scratch (one ticket + one status word per tile of LB_TILE values) in a file-scope __device__ array instead of the
// context's per-batch zero region. Every GPU that loads this library holds the array until the process exits.
#define NOALLOC_TILES ((PGM_NOALLOC_MAX_VALUES + LB_TILE - 1) / LB_TILE)
#define NOALLOC_STATUS_OFF 256                                     // bytes: the ticket has a 256-byte line of its own
#define NOALLOC_SCRATCH_BYTES (NOALLOC_STATUS_OFF + NOALLOC_TILES * 8)
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

__global__ void __launch_bounds__(PGM_BLOCK) k_format_noalloc(const double *x, uint32_t n, char *text, unsigned long long cap, unsigned long long *len_out, LbBuf st) {
    lb_scan<unsigned long long>(
        n, st, [&](uint64_t i) { return (unsigned long long)fmt8_len(x[i]); },
        [&](uint64_t i, unsigned long long at, unsigned long long len) {
            if (text && at + len <= cap) {
                const double v = x[i];
                fmt8_write(fmt8_prepare(v), v, (i + 1 == n) ? ';' : ',', text + at);      // gmove.cpp:674-678
            }
        },
        [&](unsigned long long tot) { *len_out = tot; });
}

namespace {
#define NOALLOC_MAX_DEVICES 64
// In host code g_noalloc_scratch is a host-side stand-in, not a device address: the address a kernel can use comes
// from cudaGetSymbolAddress, asked once per device with that device current.
std::atomic<void *> g_noalloc_addr[NOALLOC_MAX_DEVICES];
}  // namespace

extern "C" int pgm_format_noalloc(int device, void *stream, const double *vals_dev, uint32_t n, char *text_dev, uint64_t cap, unsigned long long *len_dev) {
    if (device < 0 || device >= NOALLOC_MAX_DEVICES) return fail(nullptr, PGM_ERR_ARG, "pgm_format_noalloc: device %d outside [0, %d)", device, NOALLOC_MAX_DEVICES);
    if (!len_dev || (n && !vals_dev)) return fail(nullptr, PGM_ERR_ARG, "pgm_format_noalloc: vals_dev / len_dev is NULL");
    if (n > PGM_NOALLOC_MAX_VALUES)
        return fail(nullptr, PGM_ERR_ARG, "pgm_format_noalloc: n = %u exceeds PGM_NOALLOC_MAX_VALUES = %u (the static scratch holds one status word per %u values)",
                    n, (unsigned)PGM_NOALLOC_MAX_VALUES, (unsigned)LB_TILE);
    CU(nullptr, cudaSetDevice(device));
    void *base = g_noalloc_addr[device].load(std::memory_order_acquire);
    if (!base) {                                                     // set-up, once per device (a second thread asking at the same time gets the same answer)
        CU(nullptr, cudaGetSymbolAddress(&base, g_noalloc_scratch));
        g_noalloc_addr[device].store(base, std::memory_order_release);
    }
    cudaStream_t s = (cudaStream_t)stream;
    const uint32_t tiles = (n + LB_TILE - 1) / LB_TILE;
    // the array keeps the last call's contents: the ticket and the status words this call will use start at zero
    CU(nullptr, cudaMemsetAsync(base, 0, NOALLOC_STATUS_OFF + (size_t)(tiles ? tiles : 1) * 8, s));
    LbBuf st;
    st.ticket = (uint32_t *)base;
    st.status = (unsigned long long *)((unsigned char *)base + NOALLOC_STATUS_OFF);
    k_format_noalloc<<<tiles ? tiles : 1, PGM_BLOCK, 0, s>>>(vals_dev, n, text_dev, (unsigned long long)cap, len_dev, st);
    CU(nullptr, cudaGetLastError());
    return PGM_OK;
}
