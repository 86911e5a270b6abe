// bam_reader.hpp — a small BAM reader for the move-table path (reference gmove.cpp:1061-1134 and
// reform.cpp:190-240 use htslib's sam_open / sam_read1 / bam_aux_get for this; htslib is not vendored by
// the reference and is not available here). Only what those call sites read is decoded:
// read name, flag, SEQ, and the aux tags ns (int), ts (int), mv (B:c array).
//
// Format (SAM/BAM specification, sections 4.1-4.2): a BGZF file is a series of gzip members, each holding
// at most 64 KiB of the BAM byte stream; the stream is "BAM\1", the header text, the reference list and
// then length-prefixed alignment records. Inflation goes through zlib, one member at a time.
#pragma once
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <zlib.h>

#include <string>
#include <vector>

struct BamRecord {
    std::string name;
    uint16_t flag = 0;
    std::string seq;               // decoded with "=ACMGRSVTWYHKDBN"
    bool has_ns = false, has_ts = false, has_mv = false;
    int64_t ns = 0, ts = 0;
    char mv_subtype = 0;           // the B array's element type ('c' is what the basecallers write)
    std::vector<int8_t> mv;        // all elements, the stride first
};

class BamReader {
  public:
    // true when the file starts with a gzip member (a BAM file is always BGZF-compressed)
    static bool looks_like_bam(const char *path) {
        FILE *f = fopen(path, "rb");
        if (!f) return false;
        unsigned char m[4] = {0, 0, 0, 0};
        const size_t n = fread(m, 1, 4, f);
        fclose(f);
        return n == 4 && m[0] == 0x1f && m[1] == 0x8b && m[2] == 8 && (m[3] & 4);
    }

    ~BamReader() {
        if (fp_) fclose(fp_);
        if (zinit_) inflateEnd(&zs_);
    }

    // Opens the file and consumes the header. err() says what went wrong.
    bool open(const char *path) {
        fp_ = fopen(path, "rb");
        if (!fp_) return fail("cannot open file");
        memset(&zs_, 0, sizeof zs_);
        if (inflateInit2(&zs_, -15) != Z_OK) return fail("zlib inflateInit2 failed");
        zinit_ = true;
        char magic[4];
        if (!read(magic, 4) || memcmp(magic, "BAM\1", 4) != 0) return fail("not a BAM file (magic)");
        int32_t l_text, n_ref;
        if (!read(&l_text, 4) || l_text < 0) return fail("truncated BAM header");
        header_.resize((size_t)l_text);
        if (l_text && !read(&header_[0], (size_t)l_text)) return fail("truncated BAM header text");
        if (!read(&n_ref, 4) || n_ref < 0) return fail("truncated BAM reference list");
        for (int32_t i = 0; i < n_ref; i++) {
            int32_t l_name, l_ref;
            if (!read(&l_name, 4) || l_name < 0) return fail("truncated BAM reference list");
            std::string nm((size_t)l_name, '\0');
            if (l_name && !read(&nm[0], (size_t)l_name)) return fail("truncated BAM reference list");
            if (!read(&l_ref, 4)) return fail("truncated BAM reference list");
        }
        return true;
    }

    // 1 = record read, 0 = end of file, -1 = error
    int next(BamRecord &r) {
        int32_t block_size;
        const int got = read_or_eof(&block_size, 4);
        if (got == 0) return 0;
        if (got < 0 || block_size < 32) return fail("truncated BAM record") ? -1 : -1;
        rec_.resize((size_t)block_size);
        if (!read(rec_.data(), (size_t)block_size)) { fail("truncated BAM record"); return -1; }
        const uint8_t *p = rec_.data();
        const uint32_t l_read_name = p[8];
        const uint32_t n_cigar = rd16(p + 12);
        r.flag = rd16(p + 14);
        const uint32_t l_seq = rd32(p + 16);
        size_t o = 32;
        const size_t need = o + l_read_name + 4ull * n_cigar + (l_seq + 1) / 2 + l_seq;
        if (need > rec_.size() || l_read_name == 0) { fail("malformed BAM record"); return -1; }
        r.name.assign((const char *)p + o, l_read_name - 1);
        o += l_read_name + 4ull * n_cigar;
        static const char code[] = "=ACMGRSVTWYHKDBN";
        r.seq.resize(l_seq);
        for (uint32_t i = 0; i < l_seq; i++) r.seq[i] = code[(p[o + (i >> 1)] >> ((~i & 1) << 2)) & 15];
        o += (l_seq + 1) / 2 + l_seq;
        r.has_ns = r.has_ts = r.has_mv = false;
        r.mv.clear();
        // aux fields: tag[2] type[1] value
        while (o + 3 <= rec_.size()) {
            const char t0 = (char)p[o], t1 = (char)p[o + 1], ty = (char)p[o + 2];
            o += 3;
            int64_t iv = 0;
            bool is_int = false;
            switch (ty) {
                case 'A': o += 1; break;
                case 'c': if (o + 1 > rec_.size()) goto bad; iv = (int8_t)p[o]; o += 1; is_int = true; break;
                case 'C': if (o + 1 > rec_.size()) goto bad; iv = p[o]; o += 1; is_int = true; break;
                case 's': if (o + 2 > rec_.size()) goto bad; iv = (int16_t)rd16(p + o); o += 2; is_int = true; break;
                case 'S': if (o + 2 > rec_.size()) goto bad; iv = rd16(p + o); o += 2; is_int = true; break;
                case 'i': if (o + 4 > rec_.size()) goto bad; iv = (int32_t)rd32(p + o); o += 4; is_int = true; break;
                case 'I': if (o + 4 > rec_.size()) goto bad; iv = rd32(p + o); o += 4; is_int = true; break;
                case 'f': o += 4; break;
                case 'Z': case 'H': while (o < rec_.size() && p[o]) o++; o++; break;
                case 'B': {
                    if (o + 5 > rec_.size()) goto bad;
                    const char sub = (char)p[o];
                    const uint32_t cnt = rd32(p + o + 1);
                    o += 5;
                    const size_t es = (sub == 'c' || sub == 'C') ? 1 : (sub == 's' || sub == 'S') ? 2 : (sub == 'i' || sub == 'I' || sub == 'f') ? 4 : 0;
                    if (es == 0 || o + es * (size_t)cnt > rec_.size()) goto bad;
                    if (t0 == 'm' && t1 == 'v') {
                        r.has_mv = true;
                        r.mv_subtype = sub;
                        if (es == 1) r.mv.assign((const int8_t *)p + o, (const int8_t *)p + o + cnt);
                    }
                    o += es * (size_t)cnt;
                    break;
                }
                default: goto bad;
            }
            if (is_int) {
                if (t0 == 'n' && t1 == 's') { r.has_ns = true; r.ns = iv; }
                else if (t0 == 't' && t1 == 's') { r.has_ts = true; r.ts = iv; }
            }
        }
        if (o > rec_.size()) goto bad;
        return 1;
    bad:
        fail("malformed BAM auxiliary field");
        return -1;
    }

    const std::string &err() const { return err_; }
    const std::string &header_text() const { return header_; }

  private:
    static uint16_t rd16(const uint8_t *p) { return (uint16_t)(p[0] | (p[1] << 8)); }
    static uint32_t rd32(const uint8_t *p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }
    bool fail(const char *m) { if (err_.empty()) err_ = m; return false; }

    // next BGZF member -> out_ (may be empty: the end-of-file marker block). false at end of file or on error.
    bool next_block() {
        unsigned char h[18];
        const size_t n = fread(h, 1, 18, fp_);
        if (n == 0) { eof_ = true; return false; }
        if (n != 18 || h[0] != 0x1f || h[1] != 0x8b || h[2] != 8 || !(h[3] & 4)) return fail("not a BGZF block");
        const uint32_t xlen = rd16(h + 10);
        // the BC subfield is the first extra subfield in every BGZF writer; scan to be safe
        std::vector<unsigned char> extra(xlen);
        memcpy(extra.data(), h + 12, xlen < 6 ? xlen : 6);
        if (xlen > 6 && fread(extra.data() + 6, 1, xlen - 6, fp_) != xlen - 6) return fail("truncated BGZF block");
        uint32_t bsize = 0;
        for (size_t i = 0; i + 4 <= xlen;) {
            const uint32_t slen = rd16(extra.data() + i + 2);
            if (extra[i] == 'B' && extra[i + 1] == 'C' && slen == 2 && i + 6 <= xlen) bsize = rd16(extra.data() + i + 4) + 1u;
            i += 4 + slen;
        }
        if (bsize < 12 + xlen + 8) return fail("BGZF block without a BC subfield");
        const uint32_t clen = bsize - 12 - xlen - 8;
        in_.resize(clen + 8);
        if (fread(in_.data(), 1, clen + 8, fp_) != clen + 8) return fail("truncated BGZF block");
        const uint32_t isize = rd32(in_.data() + clen + 4);
        out_.resize(isize);
        out_pos_ = 0;
        if (isize == 0) return true;
        if (inflateReset(&zs_) != Z_OK) return fail("zlib inflateReset failed");
        zs_.next_in = in_.data(); zs_.avail_in = clen;
        zs_.next_out = out_.data(); zs_.avail_out = isize;
        const int rc = inflate(&zs_, Z_FINISH);
        if (rc != Z_STREAM_END || zs_.avail_out != 0) return fail("BGZF block does not inflate");
        if ((uint32_t)crc32(crc32(0L, Z_NULL, 0), out_.data(), isize) != rd32(in_.data() + clen)) return fail("BGZF block checksum mismatch");
        return true;
    }

    // n bytes, or: 0 when the stream ends exactly here, -1 when it ends inside
    int read_or_eof(void *dst, size_t n) {
        size_t done = 0;
        while (done < n) {
            if (out_pos_ == out_.size()) {
                if (!next_block()) return (eof_ && done == 0 && err_.empty()) ? 0 : -1;
                continue;
            }
            const size_t k = out_.size() - out_pos_ < n - done ? out_.size() - out_pos_ : n - done;
            memcpy((char *)dst + done, out_.data() + out_pos_, k);
            out_pos_ += k; done += k;
        }
        return 1;
    }
    bool read(void *dst, size_t n) { return read_or_eof(dst, n) == 1; }

    FILE *fp_ = nullptr;
    z_stream zs_;
    bool zinit_ = false, eof_ = false;
    std::vector<unsigned char> in_, out_, rec_;
    size_t out_pos_ = 0;
    std::string header_, err_;
};
