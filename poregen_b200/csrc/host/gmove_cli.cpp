// gmove_cli.cpp — `poregen gmove` on B200: the reference's command line, option semantics, stderr banner and
// on-disk output (dump/<KMER>, freq.txt) over the CUDA library (include/pgm_gmove.h).
//
// Host side only: option handling (reference src/gmove.cpp:213-537), alignment parsing (TSV :557-577,
// PAF :977-1052 + FASTQ slice :805, text SAM :1084-1134), slow5lib record decode on host threads
// (slow5_get is pread-based and thread-safe, slow5lib/src/slow5.c:2460), batching into pinned buffers, and
// the dump writer. All arithmetic on the signal happens on the device; there is no CPU fallback.
//
// Extensions that do not change existing semantics: --gpus INT (records of every batch are split into
// contiguous ranges over the GPUs; the per-k-mer counts are exchanged so sample_limit is applied in file
// order), --batch-samples INT, --threads INT.
#include <dirent.h>
#include <errno.h>
#include <getopt.h>
#include <inttypes.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/resource.h>
#include <sys/stat.h>
#include <sys/time.h>

#include <algorithm>
#include <atomic>
#include <functional>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include <slow5/slow5.h>

#include "../../../include/pgm_gmove.h"
#include "bam_reader.hpp"

#define GMOVE_VERSION "0.1.0"   // reference src/poregen.h POREGEN_VERSION

namespace {

int g_verbose = 1;
void msg(const char *lvl, const char *fmt, ...) {
    fprintf(stderr, "[%s] ", lvl);
    va_list ap;
    va_start(ap, fmt);
    vfprintf(stderr, fmt, ap);
    va_end(ap);
    fputc('\n', stderr);
}
#define DIE(...) do { msg("ERROR", __VA_ARGS__); exit(EXIT_FAILURE); } while (0)

double realtime() {
    struct timeval tp;
    gettimeofday(&tp, nullptr);
    return tp.tv_sec + tp.tv_usec * 1e-6;
}
[[maybe_unused]] double cputime() {
    struct rusage r;
    getrusage(RUSAGE_SELF, &r);
    return r.ru_utime.tv_sec + r.ru_stime.tv_sec + 1e-6 * (r.ru_utime.tv_usec + r.ru_stime.tv_usec);
}
[[maybe_unused]] long peakrss() {
    struct rusage r;
    getrusage(RUSAGE_SELF, &r);
    return r.ru_maxrss * 1024;
}

// Options after gmove()'s own handling. Defaults: reference src/poregen.cpp:209-237 and gmove.cpp:229 (scaling 0).
struct Opt {
    int kmer_size = 9, sig_move_offset = 0, kmer_start_offset = 0;
    uint32_t signal_print_margin = 0, sample_limit = 100, file_limit = 500, index_start = 1, index_end = 500;
    int max_dur = 70, min_dur = 5, kmer_pick_margin = 2;
    double pa_min = 40.0, pa_max = 180.0;
    int flag_rna = 0, delimit_files = 0, scaling = 0;
    // extensions
    int gpus = 1, threads = 0;
    uint64_t batch_samples = 64ull << 20;
    const char *dwell_out = nullptr;   // --dwell FILE: kmer<TAB>median dwell (scripts/poregen.sh:33-52)
    const char *model_out = nullptr;   // --model FILE: kmer<TAB>median<TAB>stddev (scripts/poregen.sh:54-85)
    double stddev_limit = 1e300;
    const char *ont_model_out = nullptr;   // --ont_model FILE: the --model rows through apply_transformation (scripts/poregen.sh:87-129)
    const char *transform = "17.569300789355,84.112089074928,2.5,4";   // A,B,C,D of the workflow (scripts/poregen.sh:175)
    int host_signal = 0;          // --host-signal-decode: let slow5lib decompress the signal (default: svb-zd streams go to the GPU)
};

void print_help(FILE *fp, const Opt &o) {
    fprintf(fp, "Usage: poregen gmove reads.blow5 event_alignment_file output_dir\n");
    fprintf(fp, "\nbasic options:\n");
    fprintf(fp, "   -k INT                     kmer_size [%d]\n", o.kmer_size);
    fprintf(fp, "   -m INT                     move start offset [%d]\n", o.sig_move_offset);
    fprintf(fp, "   -s INT                     kmer start offset [%d]\n", o.kmer_start_offset);
    fprintf(fp, "   --scaling INT              scaling [%d] (0-no scaling, 1-medmad)\n", 1);
    fprintf(fp, "   --margin INT               signal print margin on both sides of the sub signal[%u] \n", o.signal_print_margin);
    fprintf(fp, "   --sample_limit INT         maximum number of instances to output for a kmer [%u] \n", o.sample_limit);
    fprintf(fp, "   --file_limit INT           maximum number of kmer files to output [%u] \n", o.file_limit);
    fprintf(fp, "   --kmer_file FILE           kmer file (optional) \n");
    fprintf(fp, "   --index_start INT          1-based closed interval index of start kmer [%u] \n", o.file_limit);
    fprintf(fp, "   --index_end INT            1-based closed interval index of end kmer [%u] \n", o.file_limit);
    fprintf(fp, "   --fastq FILE               fastq file (optional - should be provided with .paf) \n");
    fprintf(fp, "   -d                         delimit output files per read\n");
    fprintf(fp, "   --max_dur                  maximum move duration allowed for samples [%d]\n", o.max_dur);
    fprintf(fp, "   --min_dur                  maximum move duration allowed for samples [%d]\n", o.min_dur);
    fprintf(fp, "   --pa_min                   minimum pA level a sampling signal should have [%.3f]\n", o.pa_min);
    fprintf(fp, "   --pa_max                  maximum pA level a sampling signal should have [%.3f]\n", o.pa_max);
    fprintf(fp, "   --kmer_pick_margin         distance in bases from an indel when picking a kmer as sample [%d]\n", o.kmer_pick_margin);
    fprintf(fp, "   --rna                      dataset is rna\n");
    fprintf(fp, "   -h                         help\n");
    fprintf(fp, "   --verbose INT              verbosity level [%d]\n", g_verbose);
    fprintf(fp, "   --version                  print version\n");
    fprintf(fp, "\nB200 options:\n");
    fprintf(fp, "   --gpus INT                 GPUs to spread every batch over [%d]\n", o.gpus);
    fprintf(fp, "   --threads INT              host threads decoding SLOW5 records [all cores]\n");
    fprintf(fp, "   --batch-samples INT        raw samples per device batch [%" PRIu64 "]\n", o.batch_samples);
    fprintf(fp, "   --model FILE               also write kmer<TAB>median<TAB>stddev per dumped k-mer (calculate_mean_stddev_all of scripts/poregen.sh)\n");
    fprintf(fp, "   --ont_model FILE           with --model: also write the ONT .model file (apply_transformation of scripts/poregen.sh)\n");
    fprintf(fp, "   --transform A,B,C,D        level_mean*A+B, stddev rescaled to [C,D] [17.569300789355,84.112089074928,2.5,4]\n");
    fprintf(fp, "   --stddev_limit FLOAT       upper clamp of the stddev column [none]\n");
    fprintf(fp, "   --dwell FILE               per-k-mer median of (event length - 1), as calculate_dwell_times_medians prints it\n");
    fprintf(fp, "   --host-signal-decode       decompress svb-zd signals with slow5lib on the host instead of on the GPU\n");
}

// 0 created / exists and empty, -1 exists and not empty, -2 cannot create (reference gmove.cpp:126-140)
int create_dir(const char *name) {
    struct stat st;
    if (stat(name, &st) == -1) return mkdir(name, 0700) == -1 ? -2 : 0;
    DIR *d = opendir(name);
    if (!d) return 0;
    size_t n = 0;
    while (readdir(d)) n++;
    closedir(d);
    return n > 2 ? -1 : 0;
}

// lexicographic enumeration over the alphabet (reference src/poregen.cpp:248-267)
void generate_kmers(const char *alphabet, int k, std::vector<std::string> &out) {
    const uint64_t n = 1ull << (2 * k);
    out.reserve(n);
    std::string s((size_t)k, 'A');
    for (uint64_t id = 0; id < n; id++) {
        for (int j = 0; j < k; j++) s[j] = alphabet[(id >> (2 * (k - 1 - j))) & 3];
        out.push_back(s);
    }
}

// ---------------------------------------------------------------------------------------------------
// FASTA/FASTQ random access (what the reference gets from htslib's faidx: fai_load, gmove.cpp:724, and
// faidx_fetch_seq(tid, st_k, end_k - 1), gmove.cpp:805). An existing <file>.fai is used; otherwise the index
// is built in memory (htslib would also write it next to the file; this does not).
struct Fai {
    struct Ent { long long len, offset, line_blen, line_len; };
    std::unordered_map<std::string, Ent> ent;
    FILE *fp = nullptr;

    bool load(const char *fn) {
        fp = fopen(fn, "r");
        if (!fp) return false;
        std::string idx = std::string(fn) + ".fai";
        if (FILE *fi = fopen(idx.c_str(), "r")) {
            char *line = nullptr;
            size_t cap = 0;
            while (getline(&line, &cap, fi) != -1) {
                char name[4096];
                Ent e;
                if (sscanf(line, "%4095s %lld %lld %lld %lld", name, &e.len, &e.offset, &e.line_blen, &e.line_len) == 5) ent.emplace(name, e);
            }
            free(line);
            fclose(fi);
            return true;
        }
        return build();
    }
    // one pass over a FASTA ('>') or FASTQ ('@' ... '+' ...) file with uniform line lengths per record
    bool build() {
        char *line = nullptr;
        size_t cap = 0;
        ssize_t n;
        long long pos = 0;
        std::string name;
        Ent e{0, 0, 0, 0};
        int state = 0;   // 0 expect header, 1 in sequence, 2 in qualities
        long long qleft = 0;
        auto flush = [&]() { if (!name.empty()) ent.emplace(name, e); name.clear(); };
        while ((n = getline(&line, &cap, fp)) != -1) {
            long long blen = n;
            while (blen > 0 && (line[blen - 1] == '\n' || line[blen - 1] == '\r')) blen--;
            if (state == 2) {
                qleft -= blen;
                if (qleft <= 0) state = 0;
            } else if (state == 0 || ((line[0] == '>' || line[0] == '@') && state == 1 && line[0] == '>')) {
                if (line[0] == '>' || line[0] == '@') {
                    flush();
                    size_t l = 1;
                    while (l < (size_t)blen && !isspace((unsigned char)line[l])) l++;
                    name.assign(line + 1, l - 1);
                    e = Ent{0, pos + n, 0, 0};
                    state = 1;
                }
            } else if (state == 1) {
                if (line[0] == '+') { state = 2; qleft = e.len; if (qleft == 0) state = 0; }
                else {
                    if (e.line_blen == 0) { e.line_blen = blen; e.line_len = n; }
                    e.len += blen;
                }
            }
            pos += n;
        }
        flush();
        free(line);
        return true;
    }
    // 0-based inclusive coordinates clamped to the sequence; *len = -2 when the name is absent
    std::string fetch(const std::string &name, long long beg, long long end, int *len) const {
        auto it = ent.find(name);
        if (it == ent.end()) {
            *len = -2;
            fprintf(stderr, "[E::fai_fetch] The sequence \"%s\" was not found\n", name.c_str());
            return std::string();
        }
        const Ent &e = it->second;
        if (end < beg) beg = end;
        if (beg < 0) beg = 0; else if (e.len <= beg) beg = e.len;
        if (end < 0) end = 0; else if (e.len <= end) end = e.len - 1;
        long long want = end + 1 - beg;
        if (want < 0) want = 0;
        std::string s;
        s.reserve((size_t)want);
        if (want > 0 && e.line_blen > 0) {
            fseeko(fp, (off_t)(e.offset + beg / e.line_blen * e.line_len + beg % e.line_blen), SEEK_SET);
            int c;
            while ((long long)s.size() < want && (c = fgetc(fp)) != EOF)
                if (isgraph(c)) s.push_back((char)c);
        }
        *len = (int)s.size();
        return s;
    }
};

// ---------------------------------------------------------------------------------------------------
// One alignment record as parsed from the file (before the signal is fetched).
struct Rec {
    std::string read_id, align, bases;
    uint32_t trim = 0, stride = 0, query_start = 0, flags = 0;
    int32_t fastq_len = 0, st_k = 0, end_k = 0;
    uint64_t signal_len = 0;
    bool check_len = false;
};

// TSV move table: read_id, fastq_len, seq, stride, moves, signal_len, trim_offset (gmove.cpp:570-576)
bool parse_tsv(char *line, Rec &r) {
    char *save = nullptr;
    char *f[7];
    for (int i = 0; i < 7; i++) {
        f[i] = strtok_r(i ? nullptr : line, "\t", &save);
        if (!f[i]) return false;
    }
    r.read_id = f[0];
    r.fastq_len = atoi(f[1]);
    r.bases = f[2];
    r.stride = (uint32_t)atoi(f[3]);
    r.align = f[4];
    r.signal_len = strtoull(f[5], nullptr, 10);
    r.trim = (uint32_t)atoi(f[6]);
    r.check_len = true;
    return true;
}

// PAF with ss:Z: (gmove.cpp:977-1052) + the target slice of the FASTQ (gmove.cpp:779-817)
void parse_paf(char *line, Rec &r, const Fai &fai, const Opt &o) {
    char *save = nullptr;
    std::vector<char *> f;
    for (char *p = strtok_r(line, "\t\r\n", &save); p; p = strtok_r(nullptr, "\t\r\n", &save)) f.push_back(p);
    if (f.size() < 12) DIE("PAF record with fewer than 12 columns");
    r.read_id = f[0];
    r.query_start = (uint32_t)atoi(f[2]);
    if (strcmp(f[4], "+") != 0 && strcmp(f[4], "-") != 0) DIE("PAF strand must be + or -");
    const std::string tid = f[5];
    const int tstart = atoi(f[7]), tend = atoi(f[8]);
    const char *ss = nullptr;
    for (size_t i = 12; i < f.size(); i++)
        if (strncmp("ss:Z:", f[i], 5) == 0) ss = f[i] + 5;
    if (!ss) DIE("ss:Z: tag not found in paf record for %s", r.read_id.c_str());
    r.align = ss;
    const bool rna = tstart > tend;                                   // gmove.cpp:793
    if (rna && !o.flag_rna) { msg("ERROR", "Data is detected as RNA. Please provide --rna"); exit(1); }
    r.st_k = rna ? tend : tstart;
    r.end_k = rna ? tstart : tend;
    int len = 0;
    r.bases = fai.fetch(tid, r.st_k, (long long)r.end_k - 1, &len);   // gmove.cpp:805
    r.flags = rna ? PGM_READ_RNA : 0u;
    if (len < o.kmer_size) r.flags |= PGM_READ_SKIP;                  // gmove.cpp:806-808 `continue`
}

// text SAM with the move table tags ns:i, ts:i, mv:B:c (gmove.cpp:1084-1134). Returns false for header lines.
bool parse_sam(char *line, Rec &r) {
    if (line[0] == '@') return false;
    char *save = nullptr;
    std::vector<char *> f;
    for (char *p = strtok_r(line, "\t\r\n", &save); p; p = strtok_r(nullptr, "\t\r\n", &save)) f.push_back(p);
    if (f.size() < 11) DIE("SAM record with fewer than 11 columns");
    r.read_id = f[0];
    const char *ns = nullptr, *ts = nullptr, *mv = nullptr;
    for (size_t i = 11; i < f.size(); i++) {
        if (strncmp(f[i], "ns:i:", 5) == 0) ns = f[i] + 5;
        else if (strncmp(f[i], "ts:i:", 5) == 0) ts = f[i] + 5;
        else if (strncmp(f[i], "mv:B:", 5) == 0) mv = f[i] + 5;
    }
    if (!ns) DIE("tag 'ns' is not found. Please check your SAM/BAM file: ");
    if (!ts) DIE("tag 'ts' is not found. Please check your SAM/BAM file: ");
    if (!mv) DIE("NULL returned for tag mv: ");
    if (mv[0] != 'c' || mv[1] != ',') DIE("tag 'mv' specification is incorrect");
    r.signal_len = strtoull(ns, nullptr, 10);
    r.trim = (uint32_t)strtoull(ts, nullptr, 10);
    r.check_len = true;
    const char *p = mv + 2;
    r.stride = (uint32_t)strtol(p, (char **)&p, 10);                 // element 0 is the stride
    r.align.clear();
    while (*p == ',') { p++; r.align.push_back((char)strtol(p, (char **)&p, 10)); }
    // SEQ goes through htslib's 4-bit code and back through "NACNGNNNT" (gmove.cpp:1128-1134): only A C G T survive
    const char *seq = f[9];
    r.bases.clear();
    if (strcmp(seq, "*") != 0)
        for (const char *q = seq; *q; q++) {
            const char c = (char)toupper((unsigned char)*q);
            r.bases.push_back((c == 'A' || c == 'C' || c == 'G' || c == 'T') ? c : 'N');
        }
    r.fastq_len = (int32_t)r.bases.size();
    return true;
}

// The same record from a BAM file (bam_reader.hpp), with the reference's checks (gmove.cpp:1092-1127).
void bam_to_rec(const BamRecord &b, Rec &r) {
    r.read_id = b.name;
    if (!b.has_ns) DIE("tag 'ns' is not found. Please check your SAM/BAM file: ");
    if (!b.has_ts) DIE("tag 'ts' is not found. Please check your SAM/BAM file: ");
    if (!b.has_mv) DIE("NULL returned for tag mv: ");
    if (b.mv_subtype != 'c') DIE("tag 'mv' specification is incorrect");
    if (b.mv.empty()) DIE("mv array length is 0: ");
    r.signal_len = (uint64_t)b.ns;
    r.trim = (uint32_t)b.ts;
    r.check_len = true;
    r.stride = (uint32_t)b.mv[0];
    r.align.assign(b.mv.begin() + 1, b.mv.end());
    r.bases.clear();
    for (char c : b.seq) r.bases.push_back((c == 'A' || c == 'C' || c == 'G' || c == 'T') ? c : 'N');   // "NACNGNNNT" of gmove.cpp:1128-1134
    r.fastq_len = (int32_t)r.bases.size();
}

#include "reform.inc"
#include "transform.inc"

// ---------------------------------------------------------------------------------------------------
// Pinned, growable array for batch data.
template <typename T>
struct Pinned {
    T *p = nullptr;
    size_t n = 0, cap = 0;
    ~Pinned() { pgm_free_pinned(p); }
    void reserve(size_t want) {
        if (want <= cap) return;
        size_t nc = std::max(want, cap + cap / 2 + 1024);
        T *q = (T *)pgm_alloc_pinned(nc * sizeof(T));
        if (!q) DIE("pinned host allocation of %zu bytes failed", nc * sizeof(T));
        if (n) memcpy(q, p, n * sizeof(T));
        pgm_free_pinned(p);
        p = q; cap = nc;
    }
    void resize(size_t want) { reserve(want); n = want; }
    void clear() { n = 0; }
};

struct Batch {
    std::vector<Rec> recs;
    Pinned<int16_t> raw;
    Pinned<uint8_t> align, bases;
    Pinned<pgm_read> reads;
    Pinned<uint8_t> svb;          // device signal decode: the records' svb-zd streams instead of raw[]
    Pinned<uint64_t> svb_off;
    bool use_svb = false;
    uint64_t samples = 0;
    void clear() { recs.clear(); raw.clear(); align.clear(); bases.clear(); reads.clear(); svb.clear(); svb_off.clear(); use_svb = false; samples = 0; }
};

struct Gpu {
    pgm_ctx *ctx = nullptr;
    pgm_result res;
    uint32_t r0 = 0, r1 = 0;   // records of the batch this GPU owns
    std::vector<uint32_t> live_prefix;   // -d: non-skipped records among the first i of this GPU's share (write_batch)
};

struct Run {
    Opt opt;
    std::vector<std::string> kmers;     // full list; dumped range = [index_start-1, index_end)
    std::string dump_dir;
    slow5_file_t *sp = nullptr;
    bool device_signal = false;         // BLOW5 with signal_press svb-zd: the streams are decoded by k_svb_decode
    std::vector<Gpu> gpus;
    std::vector<uint32_t> totals;       // kmer_frequency_map of the dumped range
    std::vector<uint32_t> first_slot;   // slot -> first slot with the same k-mer string (itself unless a --kmer_file repeats a line)
    uint32_t n_slots = 0;
    uint32_t mode = PGM_MODE_MOVES_CHAR;
    uint64_t n_records = 0, n_samples = 0, n_examined = 0, n_selected = 0;
    double t_read = 0, t_dev = 0, t_write = 0, t_create = 0, t_parse = 0, t_start = 0, t_close = 0;
};

// slow5lib entry points below the public header (declared in slow5lib/src/slow5_extra.h:60 and
// include/slow5/slow5_press.h:105; both are exported by libslow5.a): the record's bytes as stored, and the
// record-level decompressor. slow5_get is these two followed by slow5_rec_parse (slow5lib/src/slow5.c:2519-2530).
extern "C" void *slow5_get_mem(const char *read_id, size_t *n, const slow5_file_t *s5p);

// One BLOW5 record as slow5_rec_parse reads it (slow5lib/src/slow5.c:2801-2905), stopping at raw_signal: with
// signal_press svb-zd the field holds the compressed stream and len_raw_signal its BYTE length; the sample count
// is the stream's first word (slow5_press.c:1086-1087).
struct RawRec {
    char *mem = nullptr;          // malloc'd by slow5lib
    const uint8_t *stream = nullptr;
    uint64_t stream_len = 0;
    uint32_t n_samples = 0;
    double digitisation = 0, offset = 0, range = 0;
};

static bool fetch_raw_record(const Run &R, const char *read_id, RawRec *out) {
    size_t bytes = 0;
    char *mem = (char *)slow5_get_mem(read_id, &bytes, R.sp);
    if (!mem) return false;
    const enum slow5_press_method rp = R.sp->compress ? R.sp->compress->record_press->method : SLOW5_COMPRESS_NONE;
    if (rp != SLOW5_COMPRESS_NONE) {
        size_t nb = 0;
        char *plain = (char *)slow5_ptr_depress_solo(rp, mem, bytes, &nb);
        free(mem);
        if (!plain || !nb) { free(plain); return false; }
        mem = plain; bytes = nb;
    }
    size_t o = 0;
    uint16_t idlen = 0;
    if (bytes < 2) { free(mem); return false; }
    memcpy(&idlen, mem, 2); o = 2 + (size_t)idlen + 4;                 // read_id, read_group
    if (o + 4 * 8 + 8 > bytes) { free(mem); return false; }
    double samp = 0;
    memcpy(&out->digitisation, mem + o, 8); memcpy(&out->offset, mem + o + 8, 8); memcpy(&out->range, mem + o + 16, 8);
    memcpy(&samp, mem + o + 24, 8);
    uint64_t len = 0;
    memcpy(&len, mem + o + 32, 8);
    o += 40;
    if (len < 4 || o + len > bytes) { free(mem); return false; }
    out->mem = mem; out->stream = (const uint8_t *)mem + o; out->stream_len = len;
    memcpy(&out->n_samples, out->stream, 4);
    return true;
}

// Device signal decode: the batch carries every record's svb-zd stream (pgm_upload_svb); nothing is decompressed here.
void assemble_svb(Run &R, Batch &b) {
    const size_t n = b.recs.size();
    std::vector<RawRec> recs(n);
    int nt = R.opt.threads > 0 ? R.opt.threads : (int)std::thread::hardware_concurrency();
    nt = std::max(1, std::min<int>(nt, (int)n));
    std::atomic<size_t> next(0);
    std::atomic<int> failed(0);
    auto work = [&]() {
        for (;;) {
            const size_t i = next.fetch_add(1);
            if (i >= n) break;
            if (!fetch_raw_record(R, b.recs[i].read_id.c_str(), &recs[i])) failed = 1;         // gmove.cpp:582
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < nt; t++) th.emplace_back(work);
    work();
    for (auto &t : th) t.join();
    if (failed) { fprintf(stderr, "Error in when fetching the read\n"); exit(EXIT_FAILURE); }
    uint64_t nraw = 0, nal = 0, nb = 0, nsv = 0;
    b.reads.resize(n);
    b.svb_off.resize(n + 1);
    for (size_t i = 0; i < n; i++) {
        const Rec &r = b.recs[i];
        const RawRec &s = recs[i];
        if (r.check_len && s.n_samples != r.signal_len)                                         // assert gmove.cpp:589
            DIE("read %s: signal length %" PRIu64 " in the alignment, %" PRIu64 " in the SLOW5 record", r.read_id.c_str(), r.signal_len, (uint64_t)s.n_samples);
        pgm_read &d = b.reads.p[i];
        memset(&d, 0, sizeof d);
        d.raw_off = nraw; d.raw_len = s.n_samples;
        d.align_off = nal; d.align_len = (uint32_t)r.align.size();
        d.seq_off = nb; d.seq_len = (uint32_t)r.bases.size();
        d.trim = r.trim; d.stride = r.stride; d.fastq_len = r.fastq_len; d.query_start = r.query_start;
        d.st_k = r.st_k; d.end_k = r.end_k; d.flags = r.flags;
        d.digitisation = s.digitisation; d.offset = s.offset; d.range = s.range;
        b.svb_off.p[i] = nsv;
        nraw += ((uint64_t)s.n_samples + 7) & ~(uint64_t)7;
        nal += (r.align.size() + 15) & ~(size_t)15;
        nb += r.bases.size();
        nsv += s.stream_len;
    }
    b.svb_off.p[n] = nsv;
    b.raw.n = nraw;                                  // layout only: raw[] exists on the device
    b.svb.resize(nsv); b.align.resize(nal); b.bases.resize(nb);
    next = 0;
    auto copy = [&]() {
        for (;;) {
            const size_t i = next.fetch_add(1);
            if (i >= n) break;
            const pgm_read &d = b.reads.p[i];
            memcpy(b.svb.p + b.svb_off.p[i], recs[i].stream, (size_t)recs[i].stream_len);
            memcpy(b.align.p + d.align_off, b.recs[i].align.data(), d.align_len);
            memcpy(b.bases.p + d.seq_off, b.recs[i].bases.data(), d.seq_len);
            free(recs[i].mem);
        }
    };
    th.clear();
    for (int t = 1; t < nt; t++) th.emplace_back(copy);
    copy();
    for (auto &t : th) t.join();
    b.samples = nraw;
    b.use_svb = true;
}

// Fetches the signal of every record of the batch on host threads and lays the batch out for the device:
// records start on 16-byte boundaries of raw[] (the kernels stream aligned 16-byte words).
void assemble(Run &R, Batch &b) {
    if (R.device_signal) { assemble_svb(R, b); return; }
    const size_t n = b.recs.size();
    std::vector<slow5_rec_t *> recs(n, nullptr);
    int nt = R.opt.threads > 0 ? R.opt.threads : (int)std::thread::hardware_concurrency();
    nt = std::max(1, std::min<int>(nt, (int)n));
    std::atomic<size_t> next(0);
    std::atomic<int> failed(0);
    auto work = [&]() {
        for (;;) {
            const size_t i = next.fetch_add(1);
            if (i >= n) break;
            if (slow5_get(b.recs[i].read_id.c_str(), &recs[i], R.sp) < 0) failed = 1;   // gmove.cpp:582
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < nt; t++) th.emplace_back(work);
    work();
    for (auto &t : th) t.join();
    if (failed) { fprintf(stderr, "Error in when fetching the read\n"); exit(EXIT_FAILURE); }
    // offsets
    uint64_t nraw = 0, nal = 0, nb = 0;
    b.reads.resize(n);
    for (size_t i = 0; i < n; i++) {
        const Rec &r = b.recs[i];
        const slow5_rec_t *s = recs[i];
        if (r.check_len && s->len_raw_signal != r.signal_len)                              // assert gmove.cpp:589
            DIE("read %s: signal length %" PRIu64 " in the alignment, %" PRIu64 " in the SLOW5 record", r.read_id.c_str(), r.signal_len, (uint64_t)s->len_raw_signal);
        pgm_read &d = b.reads.p[i];
        memset(&d, 0, sizeof d);
        d.raw_off = nraw; d.raw_len = (uint32_t)s->len_raw_signal;
        d.align_off = nal; d.align_len = (uint32_t)r.align.size();
        d.seq_off = nb; d.seq_len = (uint32_t)r.bases.size();
        d.trim = r.trim; d.stride = r.stride; d.fastq_len = r.fastq_len; d.query_start = r.query_start;
        d.st_k = r.st_k; d.end_k = r.end_k; d.flags = r.flags;
        d.digitisation = s->digitisation; d.offset = s->offset; d.range = s->range;
        nraw += (s->len_raw_signal + 7) & ~(uint64_t)7;
        nal += (r.align.size() + 15) & ~(size_t)15;
        nb += r.bases.size();
    }
    b.raw.resize(nraw); b.align.resize(nal); b.bases.resize(nb);
    next = 0;
    auto copy = [&]() {
        for (;;) {
            const size_t i = next.fetch_add(1);
            if (i >= n) break;
            const pgm_read &d = b.reads.p[i];
            memcpy(b.raw.p + d.raw_off, recs[i]->raw_signal, (size_t)d.raw_len * 2);
            const uint64_t pad = ((d.raw_len + 7) & ~7u) - d.raw_len;
            if (pad) memset(b.raw.p + d.raw_off + d.raw_len, 0, pad * 2);
            memcpy(b.align.p + d.align_off, b.recs[i].align.data(), d.align_len);
            memcpy(b.bases.p + d.seq_off, b.recs[i].bases.data(), d.seq_len);
            slow5_rec_free(recs[i]);
        }
    };
    th.clear();
    for (int t = 1; t < nt; t++) th.emplace_back(copy);
    copy();
    for (auto &t : th) t.join();
    b.samples = nraw;
}

void check(pgm_ctx *c, int rc) {
    if (rc != PGM_OK) DIE("%s", pgm_last_error(c));
}

// Appends the text of one batch to dump/<KMER>. Without -d a slot's text is one contiguous range per GPU,
// concatenated in GPU (= record) order. With -d a ':' follows every non-skipped record in every file that is
// still open, i.e. whose count is below sample_limit after that record (gmove.cpp:680-694).
void write_batch(Run &R, const Batch &b) {
    const Opt &o = R.opt;
    std::string out;
    if (o.delimit_files)
        for (Gpu &g : R.gpus) {
            const pgm_result &r = g.res;
            g.live_prefix.assign((size_t)r.n_reads + 1, 0);
            if (g.r1 == g.r0) continue;
            for (uint32_t rec = 0; rec < r.n_reads; rec++) g.live_prefix[rec + 1] = g.live_prefix[rec] + ((r.read_flags[rec] & PGM_RF_SKIPPED) ? 0u : 1u);
        }
    for (uint32_t s = 0; s < R.n_slots; s++) {
        out.clear();
        uint32_t count = R.totals[s];            // before this batch
        for (Gpu &g : R.gpus) {
            const pgm_result &r = g.res;
            if (g.r1 == g.r0) continue;
            const uint64_t t0 = r.slot_text_off[s], t1 = r.slot_text_off[s + 1];
            if (!o.delimit_files) { out.append(r.text + t0, (size_t)(t1 - t0)); count += r.slot_sel[s]; continue; }
            // events of the slot are the index range [e0, e1) of the grouped arrays
            uint64_t e0 = 0;
            {   // first event whose text starts at t0: slots are contiguous in event order
                const uint64_t *lo = std::lower_bound(r.ev_text_off, r.ev_text_off + r.n_sel_events, t0);
                e0 = (uint64_t)(lo - r.ev_text_off);
            }
            // Walk the slot's events once; between two of them the records without an event of this slot only contribute their
            // ':' (live = not skipped), counted from the prefix sums made once per batch (not one pass over the records per slot).
            const uint64_t e1 = e0 + r.slot_sel[s];
            const std::vector<uint32_t> &live = g.live_prefix;          // live[rec] = non-skipped records among [0, rec)
            uint32_t done = 0;                                          // records whose ':' decision has been made
            for (uint64_t e = e0; e < e1;) {
                const uint32_t rec = r.ev_read[e];
                if (o.sample_limit == 0 || count < o.sample_limit) out.append((size_t)(live[rec] - live[done]), ':');
                while (e < e1 && r.ev_read[e] == rec) {
                    out.append(r.text + r.ev_text_off[e], (size_t)(r.ev_text_off[e + 1] - r.ev_text_off[e]));
                    e++; count++;
                }
                // a file is closed only by the emission that reaches the limit (gmove.cpp:680-684); with limit 0 never
                if (!(r.read_flags[rec] & PGM_RF_SKIPPED) && (o.sample_limit == 0 || count < o.sample_limit)) out.push_back(':');
                done = rec + 1;
            }
            if (o.sample_limit == 0 || count < o.sample_limit) out.append((size_t)(live[r.n_reads] - live[done]), ':');
        }
        if (out.empty()) continue;
        const std::string path = R.dump_dir + "/" + R.kmers[o.index_start - 1 + s];
        FILE *fp = fopen(path.c_str(), "a");
        if (!fp) DIE("cannot append to %s: %s", path.c_str(), strerror(errno));
        fwrite(out.data(), 1, out.size(), fp);
        fclose(fp);
    }
}

// One pass of the reference's record loop over the batch: split over the GPUs by contiguous record ranges
// (balanced by samples), count, exchange the per-k-mer counts so every GPU knows how much of sample_limit
// earlier records consume, emit, download, write.
void process_batch(Run &R, Batch &b) {
    if (b.recs.empty()) return;
    double t0 = realtime();
    assemble(R, b);
    R.t_read += realtime() - t0;
    t0 = realtime();
    const uint32_t n = (uint32_t)b.recs.size();
    const uint32_t G = (uint32_t)R.gpus.size();
    // contiguous ranges with about the same number of samples
    {
        uint32_t r = 0;
        uint64_t acc = 0;
        for (uint32_t g = 0; g < G; g++) {
            R.gpus[g].r0 = r;
            const uint64_t target = b.samples * (g + 1) / G;
            while (r < n && (g + 1 == G || acc < target)) { acc += (b.reads.p[r].raw_len + 7) & ~7u; r++; }
            R.gpus[g].r1 = r;
        }
    }
    std::vector<std::vector<uint32_t>> counts(G, std::vector<uint32_t>(R.n_slots, 0));
    auto upload_count = [&](uint32_t g) {
        Gpu &gp = R.gpus[g];
        if (gp.r1 == gp.r0) return;
        pgm_batch pb;
        pb.n_reads = gp.r1 - gp.r0; pb.mode = R.mode;
        pb.n_raw = b.raw.n; pb.n_align = b.align.n; pb.n_bases = b.bases.n;
        pb.raw = b.raw.p; pb.align = b.align.p; pb.bases = b.bases.p; pb.reads = b.reads.p + gp.r0;
        if (G > 1) {   // ship only this GPU's slice of the big arrays: offsets become relative to the slice
            const pgm_read &a = b.reads.p[gp.r0], &z = b.reads.p[gp.r1 - 1];
            const uint64_t raw0 = a.raw_off, raw1 = z.raw_off + ((z.raw_len + 7) & ~7u);
            const uint64_t al0 = a.align_off, al1 = z.align_off + ((z.align_len + 15) & ~15u);
            const uint64_t sq0 = a.seq_off, sq1 = z.seq_off + z.seq_len;
            for (uint32_t i = gp.r0; i < gp.r1; i++) { b.reads.p[i].raw_off -= raw0; b.reads.p[i].align_off -= al0; b.reads.p[i].seq_off -= sq0; }
            pb.raw += raw0; pb.n_raw = raw1 - raw0;
            pb.align += al0; pb.n_align = al1 - al0;
            pb.bases += sq0; pb.n_bases = sq1 - sq0;
        }
        if (b.use_svb) {   // this GPU's streams, offsets relative to the first of them
            std::vector<uint64_t> off(pb.n_reads + 1);
            const uint64_t s0 = b.svb_off.p[gp.r0];
            for (uint32_t i = 0; i <= pb.n_reads; i++) off[i] = b.svb_off.p[gp.r0 + i] - s0;
            pgm_svb sv;
            sv.n_bytes = off[pb.n_reads]; sv.bytes = b.svb.p + s0; sv.off = off.data();
            pb.raw = nullptr;
            check(gp.ctx, pgm_upload_svb(gp.ctx, &pb, &sv));   // the offsets are pageable: copied before the call returns
        } else check(gp.ctx, pgm_upload(gp.ctx, &pb));
        check(gp.ctx, pgm_count(gp.ctx));
        if (G > 1) check(gp.ctx, pgm_counts_host(gp.ctx, counts[g].data()));
    };
    auto emit_download = [&](uint32_t g) {
        Gpu &gp = R.gpus[g];
        if (gp.r1 == gp.r0) return;
        check(gp.ctx, pgm_emit(gp.ctx));
        check(gp.ctx, pgm_download(gp.ctx, &gp.res));
    };
    auto on_all = [&](const std::function<void(uint32_t)> &f) {
        std::vector<std::thread> th;
        for (uint32_t g = 1; g < G; g++) th.emplace_back(f, g);
        f(0);
        for (auto &t : th) t.join();
    };
    on_all(upload_count);
    if (G > 1) {   // prior_g = totals + counts of lower ranks; consumed = totals + all counts (both capped)
        std::vector<uint64_t> run(R.totals.begin(), R.totals.end());
        std::vector<std::vector<uint32_t>> prior(G, std::vector<uint32_t>(R.n_slots));
        for (uint32_t g = 0; g < G; g++)
            for (uint32_t s = 0; s < R.n_slots; s++) {
                prior[g][s] = (uint32_t)std::min<uint64_t>(run[s], R.opt.sample_limit);
                run[s] += counts[g][s];
            }
        std::vector<uint32_t> consumed(R.n_slots);
        for (uint32_t s = 0; s < R.n_slots; s++) consumed[s] = (uint32_t)std::min<uint64_t>(run[s], R.opt.sample_limit);
        for (uint32_t g = 0; g < G; g++)
            if (R.gpus[g].r1 > R.gpus[g].r0) check(R.gpus[g].ctx, pgm_set_prior_host(R.gpus[g].ctx, prior[g].data(), consumed.data()));
    }
    on_all(emit_download);
    R.t_dev += realtime() - t0;
    t0 = realtime();
    write_batch(R, b);
    for (Gpu &g : R.gpus) {
        if (g.r1 == g.r0) continue;
        R.n_examined += g.res.n_events_examined;
        R.n_selected += g.res.n_sel_events;
        for (uint32_t s = 0; s < R.n_slots; s++) R.totals[s] += g.res.slot_sel[s];
    }
    R.t_write += realtime() - t0;
    R.n_records += n;
    R.n_samples += b.samples;
    b.clear();
}

enum { LO_SCALING = 3, LO_MARGIN, LO_SAMPLE_LIMIT, LO_FILE_LIMIT, LO_KMER_FILE, LO_INDEX_START, LO_INDEX_END, LO_FASTQ, LO_D,
       LO_MAX_DUR, LO_MIN_DUR, LO_PA_MIN, LO_PA_MAX, LO_PICK_MARGIN, LO_RNA, LO_VERBOSE, LO_HELP, LO_VERSION, LO_DEBUG_BREAK,
       LO_GPUS, LO_THREADS, LO_BATCH, LO_MODEL, LO_SDLIMIT, LO_HOSTSIG, LO_DWELL, LO_ONT_MODEL, LO_TRANSFORM };

const struct option long_options[] = {
    {"kmer_size", required_argument, 0, 'k'},          // 0
    {"sig_move_offset", required_argument, 0, 'm'},    // 1
    {"kmer_start_offset", required_argument, 0, 's'},  // 2
    {"scaling", required_argument, 0, 0},              // 3
    {"margin", required_argument, 0, 0},               // 4
    {"sample_limit", required_argument, 0, 0},         // 5
    {"file_limit", required_argument, 0, 0},           // 6
    {"kmer_file", required_argument, 0, 0},            // 7
    {"index_start", required_argument, 0, 0},          // 8
    {"index_end", required_argument, 0, 0},            // 9
    {"fastq", required_argument, 0, 0},                // 10
    {"", no_argument, 0, 'd'},                         // 11
    {"max_dur", required_argument, 0, 0},              // 12
    {"min_dur", required_argument, 0, 0},              // 13
    {"pa_min", required_argument, 0, 0},               // 14
    {"pa_max", required_argument, 0, 0},               // 15
    {"kmer_pick_margin", required_argument, 0, 0},     // 16
    {"rna", no_argument, 0, 0},                        // 17
    {"verbose", required_argument, 0, 'v'},            // 18
    {"help", no_argument, 0, 'h'},                     // 19
    {"version", no_argument, 0, 'V'},                  // 20
    {"debug-break", required_argument, 0, 0},          // 21
    {"gpus", required_argument, 0, 0},                 // 22 (B200 build)
    {"threads", required_argument, 0, 0},              // 23
    {"batch-samples", required_argument, 0, 0},        // 24
    {"model", required_argument, 0, 0},                // 25
    {"stddev_limit", required_argument, 0, 0},         // 26
    {"host-signal-decode", no_argument, 0, 0},         // 27
    {"dwell", required_argument, 0, 0},                // 28
    {"ont_model", required_argument, 0, 0},            // 29
    {"transform", required_argument, 0, 0},            // 30
    {0, 0, 0, 0}};

}  // namespace

// The library under this command line must be the one its structures were compiled against (include/pgm_gmove.h):
// a stale binary next to a fresh libpgm_gmove.so would otherwise corrupt pgm_read / pgm_opts / pgm_result silently.
static void check_abi() {
    if (pgm_abi_version() != PGM_ABI_VERSION)
        DIE("libpgm_gmove.so speaks ABI v%d, this command line was built for v%d: rebuild it (poregen_b200/csrc/host/build_cli.sh)", pgm_abi_version(), PGM_ABI_VERSION);
}

#ifndef PGM_SRC_HASH
#define PGM_SRC_HASH "unknown"
#endif
// sha256 over the command line's sources at build time (build_cli.sh); tests compare it with the sources they see
extern "C" const char *pgm_cli_build_info(void) { return PGM_SRC_HASH; }

// poregen reform (reference src/main.c:47,80), same contract as gmove below.
int reform(int argc, char *argv[]) {
    check_abi();
    return reform_impl(argc, argv);
}

// Same contract as the reference's gmove(argc, argv) (src/main.c:43,76): argv[0] is the word "gmove".
int gmove(int argc, char *argv[]) {
    check_abi();
    const double t_gmove0 = realtime();
    Run R;
    Opt &opt = R.opt;
    const char *kmer_file = nullptr, *fastq_file = nullptr;
    FILE *fp_help = stderr;
    int c, longindex = 0;
    optind = 1;
    while ((c = getopt_long(argc, argv, "k:m:s:d", long_options, &longindex)) >= 0) {
        if (c == 'k') {
            if (atoi(optarg) < 1) DIE("Kmer length should be larger than 0. You entered %d", atoi(optarg));
            opt.kmer_size = atoi(optarg);
        } else if (c == 'm') {
            if (atoi(optarg) < 0) DIE("signal move offset value must not be less than zero. You entered %d", atoi(optarg));
            opt.sig_move_offset = atoi(optarg);
        } else if (c == 's') {
            if (atoi(optarg) < 1) DIE("Kmer offset should be larger than 0. You entered %d", atoi(optarg));
            opt.kmer_start_offset = atoi(optarg);
        } else if (c == 'd') opt.delimit_files = 1;
        else if (c == 'v') g_verbose = atoi(optarg);
        else if (c == 'V') { fprintf(stdout, "gmove %s\n", GMOVE_VERSION); exit(EXIT_SUCCESS); }
        else if (c == 'h') fp_help = stdout;
        else if (c == 0) {
            switch (longindex) {
                case LO_SCALING: opt.scaling = atoi(optarg); break;
                case LO_MARGIN:
                    if (atoi(optarg) < 0) DIE("Signal print margin should be non negative. You entered %d", atoi(optarg));
                    opt.signal_print_margin = atoi(optarg); break;
                case LO_SAMPLE_LIMIT:
                    if (atoi(optarg) < 0) DIE("Maximum number of instances to output for a kmer should be non negative. You entered %d", atoi(optarg));
                    opt.sample_limit = atoi(optarg); break;
                case LO_FILE_LIMIT:   // option ORDER matters, as in the reference
                    if (atoi(optarg) < 0) DIE("Maximum number of kmer files to output should be non negative. You entered %d", atoi(optarg));
                    opt.file_limit = atoi(optarg); opt.index_end = opt.index_start + opt.file_limit - 1; break;
                case LO_KMER_FILE: kmer_file = optarg; break;
                case LO_INDEX_START:
                    if (atoi(optarg) < 1) DIE("kmer index start should be a positive number. You entered %d", atoi(optarg));
                    opt.index_start = atoi(optarg); opt.file_limit = opt.index_end - opt.index_start + 1; break;
                case LO_INDEX_END:
                    if (atoi(optarg) < 1) DIE("kmer index end should be a positive number. You entered %d", atoi(optarg));
                    opt.index_end = atoi(optarg); opt.file_limit = opt.index_end - opt.index_start + 1; break;
                case LO_FASTQ: fastq_file = optarg; break;
                case LO_MAX_DUR: opt.max_dur = atoi(optarg); break;
                case LO_MIN_DUR: opt.min_dur = atoi(optarg); break;
                case LO_PA_MIN: opt.pa_min = atof(optarg); break;
                case LO_PA_MAX: opt.pa_max = atof(optarg); break;
                case LO_PICK_MARGIN: opt.kmer_pick_margin = atoi(optarg); break;
                case LO_RNA: opt.flag_rna = 1; break;
                case LO_GPUS: opt.gpus = atoi(optarg); break;
                case LO_THREADS: opt.threads = atoi(optarg); break;
                case LO_BATCH: opt.batch_samples = strtoull(optarg, nullptr, 10); break;
                case LO_MODEL: opt.model_out = optarg; break;
                case LO_SDLIMIT: opt.stddev_limit = atof(optarg); break;
                case LO_HOSTSIG: opt.host_signal = 1; break;
                case LO_DWELL: opt.dwell_out = optarg; break;
                case LO_ONT_MODEL: opt.ont_model_out = optarg; break;
                case LO_TRANSFORM: opt.transform = optarg; break;
                default: break;
            }
        }
    }
    if (argc - optind != 3 || fp_help == stdout) {
        print_help(fp_help, opt);
        exit(fp_help == stdout ? EXIT_SUCCESS : EXIT_FAILURE);
    }
    const char *slow5file = argv[optind++], *move_table = argv[optind++], *output_dir = argv[optind++];
    if (opt.kmer_size <= opt.sig_move_offset) msg("WARNING", "signal move offset value should be less than the kmer length");
    if (opt.kmer_size > 15) DIE("kmer_size above 15 is not supported by the B200 build");

    // CUDA context creation (a few hundred ms) runs beside the host-side start-up: directory, k-mer files, indexes
    std::thread warm([&opt]() { for (int g = 0; g < opt.gpus; g++) pgm_warmup(g); });
    struct Joiner { std::thread &t; ~Joiner() { if (t.joinable()) t.join(); } } warm_join{warm};

    int rc = create_dir(output_dir);
    if (rc == -1) { msg("ERROR", "Output directory %s is not empty. Please remove it or specify another directory.", output_dir); return EXIT_FAILURE; }
    if (rc == -2) { msg("ERROR", "Could not create the output dir %s.", output_dir); return EXIT_FAILURE; }
    R.dump_dir = std::string(output_dir) + "/dump";
    rc = create_dir(R.dump_dir.c_str());
    if (rc == -1) { fprintf(stderr, "Output directory %s is not empty. Please remove it or specify another directory.", R.dump_dir.c_str()); return EXIT_FAILURE; }
    if (rc == -2) { fprintf(stderr, "Could not create the output dir %s.", R.dump_dir.c_str()); return EXIT_FAILURE; }

    std::vector<std::string> &kmers = R.kmers;
    if (kmer_file) {   // gmove.cpp:395-419: the last character of every line is chopped, the rest must be k long
        FILE *fp = fopen(kmer_file, "r");
        if (!fp) { fprintf(stderr, "Error in opening file %s\n", kmer_file); exit(EXIT_FAILURE); }
        char *line = nullptr;
        size_t cap = 0;
        ssize_t n;
        while ((n = getline(&line, &cap, fp)) != -1) {
            line[n - 1] = '\0';
            if (n != opt.kmer_size + 1) DIE("The length of kmers in %s have a different value (%zu) than the kmer size %d.", kmer_file, (size_t)(n - 1), opt.kmer_size);
            kmers.push_back(line);
        }
        free(line);
        fclose(fp);
    } else generate_kmers(opt.flag_rna ? "ACGU" : "ACGT", opt.kmer_size, kmers);

    uint32_t num_kmers = (uint32_t)kmers.size();
    fprintf(stderr, "num_kmers: %d\n", num_kmers);
    if (opt.file_limit < num_kmers) {
        num_kmers = opt.file_limit;
        fprintf(stderr, "only dumping %d kmers in kmer interval [%d-%d]\n", num_kmers, opt.index_start, opt.index_end);
    } else if (opt.file_limit > kmers.size() - opt.index_start + 1) {
        if (opt.index_end > kmers.size()) { opt.file_limit = (uint32_t)kmers.size() - opt.index_start + 1; opt.index_end = opt.index_start + opt.file_limit - 1; }
        else opt.file_limit = opt.index_end - opt.index_start + 1;
    }
    fprintf(stderr, "slow5_file_path: %s\n", slow5file);
    fprintf(stderr, "guppy_sam_output_file: %s\n", move_table);
    fprintf(stderr, "kmer_output_dir: %s\n", output_dir);
    fprintf(stderr, "kmer_size: %d\n", opt.kmer_size);
    fprintf(stderr, "sig_move_offset: %d\n", opt.sig_move_offset);
    fprintf(stderr, "signal_print_margin: %d\n", opt.signal_print_margin);
    fprintf(stderr, "kmer index closed interval : [%d-%d]\n", opt.index_start, opt.index_end);
    fprintf(stderr, "no.of output files: %d\n", opt.file_limit);
    fprintf(stderr, "sample limit : %d\n", opt.sample_limit);
    fprintf(stderr, "sample max duration : %d\n", opt.max_dur);
    fprintf(stderr, "sample min duration : %d\n", opt.min_dur);
    fprintf(stderr, "sample kmer_pick_margin : %d\n", opt.kmer_pick_margin);
    fprintf(stderr, "dataset: %s\n", opt.flag_rna ? "RNA" : "DNA");
    if (opt.index_end > kmers.size() || opt.index_start > opt.index_end)
        DIE("kmer interval [%u-%u] is outside the %zu kmers of the list", opt.index_start, opt.index_end, kmers.size());

    // one (empty) file per dumped k-mer, as the reference's fopen(..., "w") loop leaves them (gmove.cpp:460-473);
    // opened and closed one at a time, so 4^9 files do not need 262,144 descriptors
    R.n_slots = opt.index_end - opt.index_start + 1;
    {   // file creation is latency of the file system, not CPU: a few threads take it off the start-up time
        const uint32_t nt = std::min<uint32_t>(16, R.n_slots);
        std::atomic<uint32_t> failed{0xffffffffu};
        std::vector<std::thread> pool;
        for (uint32_t t = 0; t < nt; t++)
            pool.emplace_back([&, t]() {
                for (uint32_t i = opt.index_start - 1 + t; i < opt.index_end; i += nt) {
                    const std::string path = R.dump_dir + "/" + kmers[i];
                    FILE *fp = fopen(path.c_str(), "w");
                    if (!fp) { uint32_t exp = 0xffffffffu; failed.compare_exchange_strong(exp, i); return; }
                    fclose(fp);
                }
            });
        for (auto &th : pool) th.join();
        const uint32_t i = failed.load();
        if (i != 0xffffffffu) {
            fprintf(stderr, "Error in opening %dth kmer-file %s\n", i - (opt.index_start - 1) + 1, (R.dump_dir + "/" + kmers[i]).c_str());
            exit(EXIT_FAILURE);
        }
    }
    R.totals.assign(R.n_slots, 0);
    R.first_slot.resize(R.n_slots);
    {
        std::unordered_map<std::string, uint32_t> seen;
        for (uint32_t s = 0; s < R.n_slots; s++) R.first_slot[s] = seen.emplace(kmers[opt.index_start - 1 + s], s).first->second;
    }

    if (opt.scaling == 0) fprintf(stderr, "scaling: %s\n", "no scale");
    else if (opt.scaling == 1) fprintf(stderr, "scaling: %s\n", "medmad scale");
    else { print_help(fp_help, opt); exit(EXIT_FAILURE); }

    R.sp = slow5_open(slow5file, "r");
    if (!R.sp) { fprintf(stderr, "Error in opening file %s\n", slow5file); exit(EXIT_FAILURE); }
    if (slow5_idx_load(R.sp) < 0) { fprintf(stderr, "Error in loading index\n"); exit(EXIT_FAILURE); }
    R.device_signal = !opt.host_signal && R.sp->format == SLOW5_FORMAT_BINARY && R.sp->compress &&
                      R.sp->compress->signal_press->method == SLOW5_COMPRESS_SVB_ZD;

    const std::string mt(move_table);
    const std::string ext = mt.size() >= 4 ? mt.substr(mt.size() - 4) : std::string();
    Fai fai;
    if (ext == ".paf") {
        if (!fastq_file) { fprintf(stderr, ".paf input requires an additional .fastq file\n"); exit(EXIT_FAILURE); }
        if (!fai.load(fastq_file)) { fprintf(stderr, "Error in opening file %s\n", fastq_file); exit(EXIT_FAILURE); }
        R.mode = PGM_MODE_SS;
    } else if (ext == ".bam" || ext == ".sam") R.mode = PGM_MODE_MOVES_I8;   // gmove.cpp:515-520: both go through the SAM/BAM loop
    else R.mode = PGM_MODE_MOVES_CHAR;

    // device contexts
    const int ndev = pgm_device_count();
    if (ndev < 1) DIE("no CUDA device: the gmove hot path has no CPU fallback");
    // PGM_CLI_SHARE_DEVICES=1 (tests on a one-GPU box): --gpus N contexts are spread over the devices that exist
    const bool share = getenv("PGM_CLI_SHARE_DEVICES") != nullptr;
    if (opt.gpus < 1 || (opt.gpus > ndev && !share)) DIE("--gpus %d but %d device(s) present", opt.gpus, ndev);
    pgm_opts po;
    memset(&po, 0, sizeof po);
    po.kmer_size = opt.kmer_size; po.sig_move_offset = opt.sig_move_offset; po.kmer_start_offset = opt.kmer_start_offset;
    po.signal_print_margin = opt.signal_print_margin; po.sample_limit = opt.sample_limit;
    po.max_dur = (uint32_t)opt.max_dur; po.min_dur = (uint32_t)opt.min_dur; po.kmer_pick_margin = (uint32_t)opt.kmer_pick_margin;
    po.flag_rna = opt.flag_rna; po.scaling = opt.scaling; po.output = opt.delimit_files ? PGM_OUT_TEXT : PGM_OUT_TEXT_ONLY;   // -d needs every event's record
    po.n_slots = R.n_slots; po.index_lo = opt.index_start - 1;
    po.pa_min = opt.pa_min; po.pa_max = opt.pa_max;
    std::string list;
    if (kmer_file) {
        for (uint32_t i = opt.index_start - 1; i < opt.index_end; i++) list += kmers[i];
        po.kmer_list = list.data();
    }
    R.gpus.resize(opt.gpus);
    if (warm.joinable()) warm.join();
    const double t_c0 = realtime();
    R.t_start = t_c0 - t_gmove0;
    for (int g = 0; g < opt.gpus; g++)
        if (pgm_create(&po, g % ndev, nullptr, &R.gpus[g].ctx) != PGM_OK) DIE("%s", pgm_last_error(nullptr));
    if (opt.ont_model_out && !opt.model_out) DIE("%s", "--ont_model needs --model (it transforms that file)");
    if (opt.model_out || opt.dwell_out) {
        // a k-mer's values would be spread over the devices: the segmented median needs them in one place
        if (opt.gpus != 1) DIE("%s", "--model / --dwell need --gpus 1");
        if (opt.delimit_files) DIE("%s", "--model / --dwell cannot be combined with -d (the scripts split on ';' and ',' only)");
        if (pgm_model_begin(R.gpus[0].ctx) != PGM_OK) DIE("%s", pgm_last_error(R.gpus[0].ctx));
    }

    R.t_create = realtime() - t_c0;
    // record loop
    const double t_loop0 = realtime();
    FILE *fp = fopen(move_table, "r");
    if (!fp) { fprintf(stderr, "Error in opening file %s\n", move_table); exit(EXIT_FAILURE); }
    BamReader bam;
    const bool is_bam = R.mode == PGM_MODE_MOVES_I8 && BamReader::looks_like_bam(move_table);
    if (is_bam && !bam.open(move_table)) DIE("%s: %s", move_table, bam.err().c_str());
    BamRecord brec;
    Batch batch;
    char *line = nullptr;
    size_t cap = 0;
    uint64_t est = 0, progress = 0;
    const uint64_t max_records = 1u << 20;
    for (;;) {
        Rec r;
        bool have = true;
        if (is_bam) {
            const int rc = bam.next(brec);
            if (rc < 0) DIE("%s: %s", move_table, bam.err().c_str());
            if (rc == 0) break;
            bam_to_rec(brec, r);
        } else {
            if (getline(&line, &cap, fp) == -1) break;
            if (R.mode == PGM_MODE_SS) parse_paf(line, r, fai, opt);
            else if (R.mode == PGM_MODE_MOVES_I8) have = parse_sam(line, r);
            else if (!parse_tsv(line, r)) DIE("move table line with fewer than 7 columns");
        }
        if (!have) continue;
        est += r.check_len ? r.signal_len : (uint64_t)r.align.size() * 4;   // PAF: the signal length is only known after slow5_get
        batch.recs.push_back(std::move(r));
        if (++progress % 10000 == 0) fprintf(stderr, "*");
        if (est >= opt.batch_samples || batch.recs.size() >= max_records) {
            process_batch(R, batch);
            est = 0;
            bool all_full = R.n_slots == kmers.size();                         // gmove.cpp:558: every k-mer complete
            for (uint32_t s = 0; all_full && s < R.n_slots; s++) all_full = opt.sample_limit != 0 && R.totals[R.first_slot[s]] >= opt.sample_limit;
            if (all_full) break;
        }
    }
    process_batch(R, batch);
    free(line);
    fclose(fp);
    R.t_parse = realtime() - t_loop0 - R.t_read - R.t_dev - R.t_write;
    const double t_close0 = realtime();

    const std::string freq = std::string(output_dir) + "/freq.txt";
    FILE *ff = fopen(freq.c_str(), "w");
    if (!ff) { fprintf(stderr, "Error in opening %s\n", freq.c_str()); exit(EXIT_FAILURE); }
    // one line per k-mer LINE (gmove.cpp:525-534 walks the list); repeated lines of a --kmer_file share the string's one count
    for (uint32_t s = 0; s < R.n_slots; s++) fprintf(ff, "%s\t%" PRIu64 "\n", kmers[opt.index_start - 1 + s].c_str(), (uint64_t)R.totals[R.first_slot[s]]);
    fclose(ff);
    if (opt.model_out) {
        const pgm_model_row *rows = nullptr;
        if (pgm_model_finalize(R.gpus[0].ctx, opt.stddev_limit, &rows) != PGM_OK) DIE("%s", pgm_last_error(R.gpus[0].ctx));
        FILE *fm = fopen(opt.model_out, "w");
        if (!fm) { fprintf(stderr, "Error in opening %s\n", opt.model_out); exit(EXIT_FAILURE); }
        for (uint32_t s = 0; s < R.n_slots; s++) {
            // echo -e "$(basename "$file")\t$mean\t$stddev" with datamash's "%.14Lg" numbers (poregen.sh:82)
            const char *name = kmers[opt.index_start - 1 + s].c_str();
            if (rows[s].flags & (PGM_MODEL_EMPTY | PGM_MODEL_RANGE)) fprintf(fm, "%s\t\t\n", name);
            else if (rows[s].flags & PGM_MODEL_NO_SD) fprintf(fm, "%s\t%.14g\tnan\n", name, (double)rows[s].median_num / 2e8);
            else fprintf(fm, "%s\t%.14g\t%.14g\n", name, (double)rows[s].median_num / 2e8, rows[s].stddev);
        }
        fclose(fm);
        if (opt.ont_model_out) {
            std::vector<std::string> abcd;
            for (const char *p = opt.transform;; ) {
                const char *c = strchr(p, ',');
                abcd.push_back(c ? std::string(p, c) : std::string(p));
                if (!c) break;
                p = c + 1;
            }
            if (abcd.size() != 4) DIE("%s", "--transform wants A,B,C,D");
            if (apply_transformation(opt.model_out, opt.ont_model_out, abcd[0].c_str(), abcd[1].c_str(), abcd[2].c_str(), abcd[3].c_str())) exit(EXIT_FAILURE);
        }
    }
    if (opt.dwell_out) {
        const int64_t *tm = nullptr;
        if (pgm_model_dwell(R.gpus[0].ctx, &tm) != PGM_OK) DIE("%s", pgm_last_error(R.gpus[0].ctx));
        FILE *fd = fopen(opt.dwell_out, "w");
        if (!fd) { fprintf(stderr, "Error in opening %s\n", opt.dwell_out); exit(EXIT_FAILURE); }
        for (uint32_t s = 0; s < R.n_slots; s++) {   // echo -e "$filename\t$median" (poregen.sh:50)
            const char *name = kmers[opt.index_start - 1 + s].c_str();
            if (tm[s] < 0) fprintf(fd, "%s\t\n", name);
            else fprintf(fd, "%s\t%.14g\n", name, (double)tm[s] / 2.0);
        }
        fclose(fd);
    }
    for (Gpu &g : R.gpus) pgm_destroy(g.ctx);
    slow5_idx_unload(R.sp);
    slow5_close(R.sp);
    R.t_close = realtime() - t_close0;
    if (g_verbose > 1 || getenv("PGM_CLI_STATS"))
        fprintf(stderr, "\n[gmove] %" PRIu64 " records, %" PRIu64 " samples, %" PRIu64 " events examined, %" PRIu64 " selected; start %.3fs create %.3fs parse %.3fs read %.3fs device %.3fs write %.3fs close %.3fs\n",
                R.n_records, R.n_samples, R.n_examined, R.n_selected, R.t_start, R.t_create, R.t_parse, R.t_read, R.t_dev, R.t_write, R.t_close);
    return 0;
}

extern "C" int pgm_gmove_main(int argc, char **argv) { return gmove(argc, argv); }
extern "C" int pgm_reform_main(int argc, char **argv) { return reform(argc, argv); }

// poregen transform <input_file> <output_file> <A> <B> <C> <D>: apply_transformation with the script's argument order
extern "C" int pgm_transform_main(int argc, char **argv) {
    if (argc != 7) { fprintf(stderr, "Usage: apply_transformation <input_file> <output_file> <A> <B> <C> <D>\n"); return 1; }
    return apply_transformation(argv[1], argv[2], argv[3], argv[4], argv[5], argv[6]);
}

#ifndef PGM_NO_MAIN
// poregen's dispatcher (reference src/main.c:64-103), gmove only.
int main(int argc, char *argv[]) {
    const double t0 = realtime();
    const bool is_gmove = argc >= 2 && strcmp(argv[1], "gmove") == 0, is_reform = argc >= 2 && strcmp(argv[1], "reform") == 0;
    if (argc >= 2 && strcmp(argv[1], "--build-info") == 0) { printf("%s\n", pgm_cli_build_info()); return 0; }
    if (argc >= 2 && strcmp(argv[1], "transform") == 0) return pgm_transform_main(argc - 1, argv + 1);
    if (!is_gmove && !is_reform) {
        fprintf(stderr, "Usage: poregen <command> [options]\n\ncommand:\n         gmove     B200 build of `poregen gmove`\n"
                        "         reform    move table (SAM/BAM) -> per-k-mer TSV or PAF with ss:Z\n"
                        "         transform raw model -> ONT .model (apply_transformation of scripts/poregen.sh)\n");
        return (argc >= 2 && (strcmp(argv[1], "--help") == 0 || strcmp(argv[1], "-h") == 0)) ? 0 : 1;
    }
    if (is_reform) {
        const int ret = reform(argc - 1, argv + 1);
        return ret;
    }
    const int ret = gmove(argc - 1, argv + 1);
    fprintf(stderr, "\n[%s] Version: %s (B200)\n", __func__, GMOVE_VERSION);
    fprintf(stderr, "[%s] CMD:", __func__);
    for (int i = 0; i < argc; i++) fprintf(stderr, " %s", argv[i]);
    fprintf(stderr, "\n[%s] Real time: %.3f sec; CPU time: %.3f sec; Peak RAM: %.3f GB\n\n", __func__, realtime() - t0, cputime(), peakrss() / 1024.0 / 1024.0 / 1024.0);
    return ret;
}
#endif
