/* pgm_gmove.h — C ABI of the B200-native `poregen gmove` hot path.
 *
 * The reference (hiruna72/poregen) exposes NO library/FFI interface for this path: its only
 * boundaries are the command line `poregen gmove [opts] reads.blow5 alignment output_dir`
 * (reference src/gmove.cpp:80-104, 330-360) and the C++ symbol `int gmove(int, char**)`
 * (src/main.c:43,76). This header is the one host<->device seam the new build introduces under
 * that CLI. Each entry point names the reference code it replaces.
 *
 * Plain C: pointers and sizes only, no C++/torch types. The library never calls exit(), never
 * throws across the ABI, and has NO CPU fallback: every compute entry point fails with
 * PGM_ERR_CUDA when no sm_100 device is usable.
 *
 * Data model (one "batch" = consecutive alignment records, in alignment-file order):
 *   raw[]    int16 signal of all reads, concatenated               (slow5_rec_t::raw_signal)
 *   align[]  per read: move string '0'/'1' (TSV col 5, gmove.cpp:574), or the mv:B:c moves after
 *            the stride element as int8 (gmove.cpp:1100-1123), or the ss:Z: string (gmove.cpp:778)
 *   bases[]  per read: basecalled sequence (TSV col 3 / SAM SEQ) or, for ss records, the target
 *            slice faidx_fetch_seq(tid, st_k, end_k-1) (gmove.cpp:805). NOT T->U converted: the
 *            device applies gmove.cpp:815-817 itself for records flagged PGM_READ_RNA.
 *   reads[]  one pgm_read per record (offsets into the three arrays + per-record scalars)
 */
#ifndef PGM_GMOVE_H
#define PGM_GMOVE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PGM_ABI_VERSION 4   /* v4 (round 2): pgm_submit / pgm_wait, pgm_create_sibling, pgm_summary, pgm_model_log*, pgm_format_noalloc */

/* ---- status codes ---- */
#define PGM_OK 0
#define PGM_ERR_ARG 1       /* bad argument / unsupported option combination */
#define PGM_ERR_CUDA 2      /* CUDA runtime failure or no usable device */
#define PGM_ERR_NOMEM 3     /* host or device allocation failed */
#define PGM_ERR_INPUT 4     /* malformed batch (bad ss string, offsets out of range, ...) */
#define PGM_ERR_STATE 5     /* calls made in the wrong order */

/* ---- alignment encodings (pgm_batch.mode) ---- */
#define PGM_MODE_MOVES_CHAR 0 /* TSV move table: bytes '0'/'1'              gmove.cpp:539-706   */
#define PGM_MODE_SS 1         /* PAF ss:Z: string "<n>,|<n>I|<n>D"          gmove.cpp:707-975   */
#define PGM_MODE_MOVES_I8 2   /* SAM/BAM mv:B:c moves (int8 0/1); a read with ANY out-of-range pA
                                 sample is dropped whole                    gmove.cpp:1061-1266 */

/* ---- pgm_read.flags ---- */
#define PGM_READ_RNA 1u  /* ss record with target_start > target_end (gmove.cpp:793)            */
#define PGM_READ_SKIP 2u /* host decided to skip the record (e.g. ss target slice shorter than
                            k, gmove.cpp:806-808); it produces no events and no ':' delimiter   */

/* ---- result read_flags bits (device decided) ---- */
#define PGM_RF_SKIPPED 1u   /* record produced nothing and gets no ':' (fastq_len<10
                               gmove.cpp:616-619,1178-1181; MOVES_I8 out-of-range 1158-1160;
                               host PGM_READ_SKIP)                                               */
#define PGM_RF_BAD_ALIGN 2u /* malformed alignment (bad ss char, fewer than m+1 moves): the
                               reference exits / has undefined behaviour; we emit nothing        */
#define PGM_RF_BAD_CALIB 4u /* non-finite calibration; reference behaviour undefined             */

/* ---- output kinds (pgm_opts.output) ---- */
#define PGM_OUT_F64 0  /* selected samples as IEEE doubles (bit-exact with the reference vector) */
#define PGM_OUT_TEXT 1 /* the exact "%.8f," / "%.8f;" text of gmove.cpp:674-678, formatted on the
                          device, INSTEAD of the doubles (pgm_result.samples is NULL)             */
#define PGM_OUT_TEXT_ONLY 2 /* PGM_OUT_TEXT without the per-event arrays on the host (ev_len, ev_read, ev_start,
                               ev_text_off stay NULL): all a writer needs unless -d asks for record delimiters */

/* Mirrors the fields of the reference's opt_t that the gmove path reads
 * (src/poregen.h:48-79, defaults src/poregen.cpp:209-237) after gmove()'s own option handling
 * (gmove.cpp:240-327, 428-440, 479-491). */
typedef struct pgm_opts {
    uint32_t kmer_size;           /* -k            [9]   1..15                                    */
    uint32_t sig_move_offset;     /* -m            [0]                                            */
    uint32_t kmer_start_offset;   /* -s            [0]   moves modes only (gmove.cpp:631,1194)    */
    uint32_t signal_print_margin; /* --margin      [0]                                            */
    uint32_t sample_limit;        /* --sample_limit[100]                                          */
    uint32_t max_dur;             /* --max_dur     [70]                                           */
    uint32_t min_dur;             /* --min_dur     [5]                                            */
    uint32_t kmer_pick_margin;    /* --kmer_pick_margin [2], ss mode only (gmove.cpp:204-211)     */
    uint32_t flag_rna;            /* --rna: generated alphabet is ACGU (gmove.cpp:421-425)        */
    uint32_t scaling;             /* --scaling: 0 none (the real default, gmove.cpp:229), 1 medmad*/
    uint32_t output;              /* PGM_OUT_*                                                    */
    uint32_t n_slots;             /* number of dumped k-mers = index_end - index_start + 1; at
                                     most 2^20 (all 10-mers): two sort digits of 10 bits           */
    uint32_t index_lo;            /* generated list: id of the first dumped k-mer = index_start-1 */
    uint32_t reserved0;
    double pa_min;                /* --pa_min [40.0]                                              */
    double pa_max;                /* --pa_max [180.0]                                             */
    const char *kmer_list;        /* NULL: k-mers are the lexicographic ACGT/ACGU enumeration
                                     (poregen.cpp:248-267) and slot s is id index_lo+s.
                                     else: n_slots*kmer_size chars, the --kmer_file lines of the
                                     dumped range in order (gmove.cpp:395-419). A line that repeats
                                     an earlier one gets a slot that stays empty: the FIRST slot of
                                     a k-mer string takes its events (the reference keys files and
                                     counts by the string, so repeats share one of each).           */
} pgm_opts;

/* One alignment record. 88 bytes, no padding. */
typedef struct pgm_read {
    uint64_t raw_off;      /* index of the read's first sample in raw[]                           */
    uint64_t align_off;    /* byte offset in align[]                                              */
    uint64_t seq_off;      /* byte offset in bases[]                                              */
    uint32_t raw_len;      /* slow5_rec_t::len_raw_signal                                         */
    uint32_t align_len;    /* bytes: moves (without the stride element) or strlen(ss)             */
    uint32_t seq_len;      /* bytes in bases[] (fastq_seq.size(); ss: fetched length)             */
    uint32_t trim;         /* moves: trim_offset / ts:i (gmove.cpp:576,1098); ss: must be 0       */
    uint32_t stride;       /* moves: stride (TSV col 4 / mv[0]); ss: ignored                      */
    int32_t fastq_len;     /* moves: TSV col 2 / l_qseq, only used for the "<10 => skip" rule     */
    uint32_t query_start;  /* ss: PAF col 3 = first raw index (gmove.cpp:779)                     */
    int32_t st_k;          /* ss: min(target_start,target_end) (gmove.cpp:792-794)                */
    int32_t end_k;         /* ss: max(target_start,target_end)                                    */
    uint32_t flags;        /* PGM_READ_*                                                          */
    double digitisation;   /* slow5_rec_t calibration, used as TO_PICOAMPS (poregen.h:30)         */
    double offset;
    double range;
} pgm_read;

/* Host (pgm_process / pgm_upload) or device (pgm_attach_device) pointers. */
typedef struct pgm_batch {
    uint32_t n_reads;
    uint32_t mode;         /* PGM_MODE_*                                                          */
    uint64_t n_raw;        /* samples in raw[]                                                    */
    uint64_t n_align;      /* bytes in align[]  (must be < 2^32)                                  */
    uint64_t n_bases;      /* bytes in bases[]                                                    */
    const int16_t *raw;
    const uint8_t *align;
    const uint8_t *bases;
    const pgm_read *reads;
} pgm_batch;

/* Result of one batch. All pointers are pinned host memory owned by the context and stay valid
 * until the next pgm_process/pgm_download/pgm_destroy on it.
 * Selected events are grouped by slot (k-mer), slots ascending; inside a slot they are in the
 * reference's emission order = (record order, position in record) (gmove.cpp:650-652,680-685). */
typedef struct pgm_result {
    uint32_t n_slots;
    uint32_t n_reads;
    uint64_t n_events_examined; /* alignment events looked at (before any filter)                 */
    uint64_t n_events_pass;     /* passed dur/k-mer/indel filters and not already over quota      */
    uint64_t n_sel_events;
    uint64_t n_sel_samples;
    const uint32_t *slot_sel;   /* [n_slots]  events selected per slot in THIS batch              */
    const uint32_t *slot_total; /* [n_slots]  kmer_frequency_map after this batch (gmove.cpp:680) */
    const uint32_t *ev_len;     /* [n_sel_events] samples per event, grouped order                */
    const uint32_t *ev_read;    /* [n_sel_events] record index inside the batch                   */
    const uint32_t *ev_start;   /* [n_sel_events] first sample, relative to raw_off+trim          */
    const double *samples;      /* [n_sel_samples] the values gmove.cpp:675-678 prints            */
    const uint8_t *read_flags;  /* [n_reads] PGM_RF_*                                             */
    const char *text;           /* PGM_OUT_TEXT: concatenated dump text, grouped by slot          */
    const uint64_t *slot_text_off; /* [n_slots+1] byte range of slot s in text                    */
    const uint64_t *ev_text_off;   /* PGM_OUT_TEXT: [n_sel_events+1] byte offset of every event's
                                      text (the writer of `-d` delimiters needs it, gmove.cpp:692-694) */
} pgm_result;

typedef struct pgm_ctx pgm_ctx;

/* Replaces the option block handed to process_move_table_* (gmove.cpp:515-520) and the
 * kmer_frequency_map / kmer_file_pointer_array initialisation (gmove.cpp:460-477).
 * `device` is a CUDA ordinal. `stream` is a cudaStream_t to enqueue on, or NULL for a private one. */
int pgm_create(const pgm_opts *opts, int device, void *stream, pgm_ctx **out);
void pgm_destroy(pgm_ctx *ctx);
const char *pgm_last_error(const pgm_ctx *ctx); /* ctx may be NULL: error of a failed pgm_create */
int pgm_abi_version(void);

/* One call = one pass of process_move_table_file / _paf / _bam over the batch's records
 * (gmove.cpp:539-706, 707-975, 1061-1266): host buffers in, host result out. Synchronous. */
int pgm_process(pgm_ctx *ctx, const pgm_batch *host_batch, pgm_result *out);

/* Split-phase form of the same pass (device-resident timing, multi-GPU quota exchange):
 *   pgm_upload | pgm_attach_device  ->  pgm_count  ->  [pgm_set_prior]  ->  pgm_emit  ->  pgm_download
 * Everything except pgm_download is asynchronous on the context's stream. */
int pgm_upload(pgm_ctx *ctx, const pgm_batch *host_batch);          /* H2D copies only            */
int pgm_attach_device(pgm_ctx *ctx, const pgm_batch *device_batch); /* inputs already in HBM      */
int pgm_count(pgm_ctx *ctx);  /* decode + filters + per-slot counts of passing events             */
int pgm_emit(pgm_ctx *ctx);   /* stable bucketing, sample_limit, Med-MAD, sample gather/format    */
int pgm_download(pgm_ctx *ctx, pgm_result *out); /* D2H + stream sync                             */
int pgm_sync(pgm_ctx *ctx);
/* The four counters of a pgm_result: n_events_examined, n_events_pass, n_sel_events, n_sel_samples of the batch pgm_emit
 * just finished, without moving any array to the host (every pointer of *out is NULL). Synchronises; the batch stays
 * emitted, so pgm_download can still follow. For callers that keep the result in HBM (benchmarks, K7). */
int pgm_summary(pgm_ctx *ctx, pgm_result *out);

struct pgm_svb; /* defined below (pgm_upload_svb) */

/* Batches in flight (pgm_submit / pgm_wait). The reference's loop handles one record after the other; a host that
 * streams a file wants batch j+1 on its way to the device, and its kernels running, while batch j's dump text travels back.
 * pgm_submit queues one whole pass (upload or svb upload when `svb` is not NULL, count, emit) of the NEXT batch in alignment-
 * file order and returns at once; at most two batches are in flight. pgm_wait(batch_id) blocks until that batch is done and
 * hands out its result like pgm_download; batches are waited for in submission order. The per-k-mer totals are carried from
 * batch to batch on the device (batch j+1's decoder waits for batch j's quota kernel, nothing else is serialised), so the
 * output is the reference's single loop. The caller's (pinned) input buffers must stay untouched until pgm_wait returns for
 * that batch; a result stays valid until the second pgm_submit after it. Not combinable with pgm_model_begin or the
 * multi-GPU exchange (use the split-phase calls on sibling contexts for those). */
int pgm_submit(pgm_ctx *ctx, const pgm_batch *host_batch, const struct pgm_svb *svb, uint64_t *batch_id);
int pgm_wait(pgm_ctx *ctx, uint64_t batch_id, pgm_result *out);
/* A sibling of `parent`: same options and device, its own stream (NULL: a private one), batch buffers and results, and the
 * SAME per-k-mer totals. Split-phase calls on siblings may be interleaved freely as long as the batches are counted in
 * alignment-file order: a sibling's pgm_count waits (on the device) for the quota kernel of the family's previous pgm_emit.
 * Destroy siblings before the parent. */
int pgm_create_sibling(pgm_ctx *parent, void *stream, pgm_ctx **out);

/* Compressed signal. Replaces the host-side signal decompression slow5lib runs inside
 * slow5_get / slow5_rec_parse (slow5lib/src/slow5.c:2876-2905 -> slow5_press.c:1085-1130 ptr_depress_svb_zd ->
 * thirdparty/streamvbyte/src/streamvbyte_decode.c:57-104 and streamvbyte_zigzag.c:34-40): record i's stream
 * bytes[off[i] .. off[i+1]) is the raw_signal field of its BLOW5 record exactly as stored when the file's
 * signal_press is svb-zd (u32 count | (count+3)/4 key bytes | 1..4 data bytes per value). pgm_upload_svb copies
 * the streams instead of raw[] (host_batch->raw must be NULL; n_raw and reads[].raw_off/raw_len describe the
 * layout the samples are decoded into on the device - make raw_off a multiple of 8 for 16-byte stores) and
 * decodes them with one kernel; every later call is the same as after pgm_upload. A stream whose count differs
 * from raw_len, or that does not consume exactly its bytes (slow5lib's own check, slow5_press.c:1097-1103),
 * makes pgm_download fail with PGM_ERR_INPUT. */
typedef struct pgm_svb {
    uint64_t n_bytes;      /* bytes in bytes[]                                                     */
    const uint8_t *bytes;  /* concatenated streams (pinned host memory for speed)                  */
    const uint64_t *off;   /* [n_reads+1] byte offsets into bytes[], ascending                     */
} pgm_svb;
int pgm_upload_svb(pgm_ctx *ctx, const pgm_batch *host_batch, const pgm_svb *svb);

/* Multi-GPU quota exchange (north_star: allgather of the 4^k count vector + exclusive scan over
 * ranks). After pgm_count, counts[s] = passing events of slot s in this rank's batch (device
 * uint32[n_slots]). The caller allgathers them and hands back, as DEVICE pointers:
 *   prior[s]    = quota of slot s already consumed before this rank's batch
 *                 (= total before this round + sum over lower ranks of their counts)
 *   consumed[s] = min(sample_limit, total before this round + sum over ALL ranks) = the
 *                 kmer_frequency_map after the round.
 * Without these calls the context behaves as a single rank and carries the totals itself. */
int pgm_counts_device(pgm_ctx *ctx, uint32_t **counts_dev);
int pgm_totals_device(pgm_ctx *ctx, uint32_t **consumed_dev);
int pgm_set_prior(pgm_ctx *ctx, const uint32_t *prior_dev, const uint32_t *consumed_dev);
/* Fused form: hand over the all-gathered counts themselves (device uint32[n_ranks][n_slots], rank
 * order = record order) and this context's rank; the quota kernel of pgm_emit sums the lower ranks
 * (prior) and all ranks (totals after the round) per slot on the fly, on top of the totals the
 * context carried into the round. The exchange is then one all-gather and no extra launch. */
int pgm_set_gathered(pgm_ctx *ctx, const uint32_t *all_counts_dev, uint32_t n_ranks, uint32_t rank);

/* Host-side form of the same exchange for callers that do not link the CUDA runtime (the C++ CLI):
 * pgm_counts_host copies counts[n_slots] out after pgm_count (synchronises the stream);
 * pgm_set_prior_host uploads prior[] / consumed[] into buffers owned by the context. */
int pgm_counts_host(pgm_ctx *ctx, uint32_t *counts_out);
int pgm_totals_host(pgm_ctx *ctx, uint32_t *consumed_out);
int pgm_set_prior_host(pgm_ctx *ctx, const uint32_t *prior, const uint32_t *consumed);

/* Pinned host memory for batch arrays (uploads from pageable memory are staged by the driver and slow). */
void *pgm_alloc_pinned(size_t bytes);
void pgm_free_pinned(void *p);
int pgm_device_count(void);
/* Creates the device's CUDA context ahead of time (callable from any thread, before pgm_create): a command line
 * can hide the few hundred milliseconds behind its own start-up. */
int pgm_warmup(int device);

/* Forget all per-k-mer totals (start a new run with the same options). */
int pgm_reset(pgm_ctx *ctx);

/* Introspection for benchmarks: the stream (cudaStream_t) and the number of kernel launches
 * issued so far. */
void *pgm_stream(pgm_ctx *ctx);
uint64_t pgm_launch_count(const pgm_ctx *ctx);

/* Per-stage device timing with CUDA events on the context's stream (for roofline reporting).
 * pgm_profile_enable(ctx, 1) starts bracketing every stage with an event pair;
 * pgm_profile_read synchronises, adds up the elapsed time and launch count of every stage since
 * the last read into ms[PGM_N_STAGES] / launches[PGM_N_STAGES], and clears the record. */
#define PGM_STAGE_PREP 0         /* k_read_prep, k_order, k_oor_check                              */
#define PGM_STAGE_DECODE_COUNT 1 /* k_decode_moves / k_decode: alignment -> event rows, filters, per-k-mer counts */
#define PGM_STAGE_SCAN 2         /* k_scan_sel (sample offsets of the selected events)             */
#define PGM_STAGE_PLACE 3        /* k_scan_records, k_tile_hist: dense order of the passing events */
#define PGM_STAGE_SORT 4         /* k_scan_hist, k_rank | k_msd_scatter + k_msd_rank               */
#define PGM_STAGE_SELECT 5       /* k_quota (per-k-mer quota, bucket and selection bases)          */
#define PGM_STAGE_STATS 6        /* k_stats (+ k_stats_wide): exact Med-MAD of every record that owns a selected event */
#define PGM_STAGE_GATHER 7       /* k_gather: the selected samples as doubles, in output order     */
#define PGM_STAGE_MISC 8         /* memsets                                                        */
#define PGM_STAGE_TEXT 9         /* PGM_OUT_TEXT: k_text_len, k_scan_text (k_text_write runs in pgm_download) */
#define PGM_STAGE_SVB 10         /* k_svb_decode (pgm_upload_svb)                                   */
#define PGM_N_STAGES 11
#define PGM_PROFILE_ONLY(stage) (2 + (stage)) /* pgm_profile_enable(ctx, PGM_PROFILE_ONLY(s)): bracket stage s alone. Every
                                                  event pair costs a few microseconds of stream time; a timed region that
                                                  wants one kernel's duration should not pay for ten. */
int pgm_profile_enable(pgm_ctx *ctx, int on);
int pgm_profile_read(pgm_ctx *ctx, float *ms, uint32_t *launches);
const char *pgm_stage_name(int stage);

/* Model aggregation (K7; reference scripts/poregen.sh:54-85, calculate_mean_stddev_all): per k-mer, over
 * all values of dump/<KMER> except the first (`tail -n +2`), `datamash median` and `datamash sstdev`, the
 * latter clamped to stddev_limit. The script reads the dump TEXT, so the values are the "%.8f" decimals:
 * value = N * 1e-8 with N the printed integer. pgm_model_begin starts (or restarts) a log of every
 * following batch's selected samples (taken in pgm_download); pgm_model_finalize groups the log by k-mer
 * in file order and returns rows[n_slots] (host memory owned by the context, valid until the next call).
 * median = median_num / 2e8 exactly (median_num = sum of the two middle N, or twice the middle one). */
typedef struct pgm_model_row {
    double median, stddev;      /* NaN where datamash would print nothing / nan                    */
    int64_t median_num;
    uint64_t n;                 /* values used: the file's values minus the first                    */
    uint32_t flags, pad;        /* PGM_MODEL_*                                                       */
} pgm_model_row;
#define PGM_MODEL_EMPTY 1u      /* fewer than two values in the file                                 */
#define PGM_MODEL_NO_SD 2u      /* exactly two values: one left after the drop, sstdev undefined     */
#define PGM_MODEL_RANGE 4u      /* some |value| * 1e8 does not fit int64 (or inf/nan): row not computed */
#define PGM_MODEL_CLAMPED 8u    /* stddev was above stddev_limit                                     */
int pgm_model_begin(pgm_ctx *ctx);
int pgm_model_finalize(pgm_ctx *ctx, double stddev_limit, const pgm_model_row **rows_out);
/* Logs the batch pgm_emit just finished without downloading anything (pgm_download logs it too, once). Synchronises. */
int pgm_model_log(pgm_ctx *ctx);
/* K7 across GPUs (SURVEY.md 8e "Output merge"; BASELINE.json configs[4]). A rank's log holds one entry per logged batch: the
 * batch's values (int64 N of the "%.8f" text, device memory) grouped by k-mer in file order, and the table
 * slot_off[n_slots+1] (host memory) of where each k-mer's values start. The owner of a k-mer range collects, round by
 * round and rank by rank (= file order), every rank's values of that range - the transport is the caller's (NCCL
 * all-to-all in poregen_b200.gmove.model_exchange, device-to-device copies in the one-GPU tests) - appends them to a
 * context of its own with pgm_model_log_append (slot_off_host: full table, zero-length outside the owned range; vals_dev:
 * the values behind slot_off_host[0] ... slot_off_host[n_slots]) and calls pgm_model_finalize there: the rows of its range
 * are then what one context that saw every batch would have produced. Event lengths are not exchanged (pgm_model_dwell stays
 * a one-context call). Pointers handed out stay valid until the next call that logs or clears. */
int pgm_model_log_size(pgm_ctx *ctx, uint32_t *n_batches);
int pgm_model_log_entry(pgm_ctx *ctx, uint32_t i, const int64_t **vals_dev, const uint64_t **slot_off_host);
int pgm_model_log_append(pgm_ctx *ctx, const int64_t *vals_dev, uint64_t n_vals, const uint64_t *slot_off_host);
/* Room for n_vals_total logged values in all (the log grows by reallocation otherwise: an owner that is about to append the
 * pieces of every rank and round knows the total from the exchanged counts and asks for it once). */
int pgm_model_log_reserve(pgm_ctx *ctx, uint64_t n_vals_total);

/* Dwell medians (reference scripts/poregen.sh:33-52, calculate_dwell_times_medians): per dump file, `datamash
 * median` over the comma counts of its ';'-separated fields = (event length - 1) of every event, plus one 0 for the empty field
 * behind the last ';'. Uses the log pgm_model_begin started (event lengths are logged beside the values). Returns
 * twice_median[n_slots] (host memory owned by the context): the sum of the two middle counts (or twice the middle one), so the
 * median is twice_median / 2 exactly; -1 for a k-mer without events (empty file: datamash prints nothing). Can be called before
 * or after pgm_model_finalize. */
int pgm_model_dwell(pgm_ctx *ctx, const int64_t **twice_median_out);

/* Test hook for the device formatter: the text of ONE event made of the n host doubles vals[]
 * ("%.8f," ... "%.8f;", gmove.cpp:674-678) into text_out[cap]; *len_out = its length (also when it
 * does not fit, in which case nothing is written). */
int pgm_debug_format(pgm_ctx *ctx, const double *vals, uint32_t n, char *text_out, uint64_t cap, uint64_t *len_out);

/* Test hook: the first n samples of the loaded batch's raw[] as the device holds them (after pgm_upload_svb: as
 * decoded), and the decoder's verdict (PGM_ERR_INPUT for a malformed stream). Synchronises. */
int pgm_debug_raw(pgm_ctx *ctx, int16_t *out, uint64_t n);

/* The dump text ("%.8f," ... "%.8f;": one event of n values, as pgm_debug_format writes it) of n doubles that are
 * already in device memory, for callers that cannot allocate. This entry point has NO counterpart in the reference:
 * it was added for code under CUDA graph capture, tight serving loops and
 * harnesses that forbid cudaMalloc. It calls no cudaMalloc / cudaFree / allocating routine and takes no context:
 * its scan scratch is a static __device__ array of NOALLOC_SCRATCH_BYTES (33,024 bytes) that every GPU loading the
 * library holds until exit, which bounds n at PGM_NOALLOC_MAX_VALUES; a larger n is rejected with PGM_ERR_ARG and a
 * message in pgm_last_error(NULL).
 *   vals_dev[n]   device doubles;  text_dev[cap] device characters (NULL: lengths only);  *len_dev (device, 8 bytes)
 *   = the length of the whole text, always. A value whose characters do not all fit below cap is not written.
 * One memset node + one kernel on `stream` (NULL = the legacy default stream); returns without synchronising.
 * Makes `device` current. The first call on a device looks the array's address up (cudaGetSymbolAddress) and may load
 * the module: make it outside stream capture (n = 0 will do).
 * All calls on a device share the one scratch array, and each call clears what it uses: two calls whose work overlaps
 * on the GPU (different streams without an ordering between them) RACE. Keep calls on one stream or order them. */
#define PGM_NOALLOC_MAX_VALUES (1u << 24)
int pgm_format_noalloc(int device, void *stream, const double *vals_dev, uint32_t n, char *text_dev, uint64_t cap,
                       unsigned long long *len_dev);

#ifdef __cplusplus
}
#endif
#endif /* PGM_GMOVE_H */
