/* oracle/shim/htslib/sam.h — TEST INFRASTRUCTURE, not product code.
 *
 * Minimal stand-in for the htslib-1.17 SAM API subset the reference's gmove
 * SAM/BAM path calls (reference src/gmove.cpp:1067-1133, 1186, 1196,
 * 1262-1264), so that the UNMODIFIED reference sources compile here without
 * htslib (not vendored; Makefile:69-81 downloads it). Text SAM only: records
 * are parsed into htslib's in-memory BAM layout for qname / 4-bit seq / aux
 * (types i, f, Z, A, B). BGZF/BAM input is not supported by this stand-in
 * (sam_open returns NULL for it), which is recorded in DESIGN.md.
 * bam_auxB2i follows htslib's documented contract: an index past the array end
 * returns 0 and sets errno=ERANGE.
 */
#ifndef ORACLE_SHIM_SAM_H
#define ORACLE_SHIM_SAM_H
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
#include <errno.h>

typedef struct htsFile { FILE *fp; char *line; size_t cap; int have_line; } htsFile;
typedef htsFile samFile;
typedef struct bam_hdr_t { int dummy; } bam_hdr_t;
typedef struct { int32_t l_qseq; uint16_t l_qname; uint16_t flag; } bam1_core_t;
typedef struct bam1_t { bam1_core_t core; uint8_t *data; int l_data; uint32_t m_data; int aux_off; } bam1_t;

static inline htsFile *sam_open(const char *fn, const char *mode) {
    (void)mode;
    FILE *fp = fopen(fn, "r");
    if (!fp) return NULL;
    int c0 = fgetc(fp), c1 = fgetc(fp);
    if (c0 == 0x1f && c1 == 0x8b) { fclose(fp); fprintf(stderr, "[shim] BGZF/BAM input is not supported by the htslib stand-in\n"); return NULL; }
    rewind(fp);
    htsFile *h = (htsFile *)calloc(1, sizeof(htsFile));
    h->fp = fp;
    return h;
}
static inline int sam_close(htsFile *h) { if (!h) return 0; fclose(h->fp); free(h->line); free(h); return 0; }

static inline bam_hdr_t *sam_hdr_read(htsFile *h) {
    /* consume '@' lines; keep the first alignment line buffered */
    ssize_t r;
    while ((r = getline(&h->line, &h->cap, h->fp)) != -1) {
        if (h->line[0] != '@') { h->have_line = 1; break; }
    }
    return (bam_hdr_t *)calloc(1, sizeof(bam_hdr_t));
}
static inline void bam_hdr_destroy(bam_hdr_t *h) { free(h); }
static inline bam1_t *bam_init1(void) { return (bam1_t *)calloc(1, sizeof(bam1_t)); }
static inline void bam_destroy1(bam1_t *b) { if (b) { free(b->data); free(b); } }

static inline void shim_bam_reserve(bam1_t *b, size_t need) {
    if (need > b->m_data) { b->m_data = (uint32_t)(need * 2 + 64); b->data = (uint8_t *)realloc(b->data, b->m_data); }
}
static inline void shim_bam_put(bam1_t *b, const void *p, size_t n) {
    shim_bam_reserve(b, (size_t)b->l_data + n);
    memcpy(b->data + b->l_data, p, n);
    b->l_data += (int)n;
}

static inline int shim_nt16(char c) {
    switch (c) {
        case '=': return 0; case 'A': case 'a': return 1; case 'C': case 'c': return 2; case 'M': case 'm': return 3;
        case 'G': case 'g': return 4; case 'R': case 'r': return 5; case 'S': case 's': return 6; case 'V': case 'v': return 7;
        case 'T': case 't': return 8; case 'W': case 'w': return 9; case 'Y': case 'y': return 10; case 'H': case 'h': return 11;
        case 'K': case 'k': return 12; case 'D': case 'd': return 13; case 'B': case 'b': return 14; default: return 15;
    }
}

static inline int sam_read1(htsFile *h, bam_hdr_t *hdr, bam1_t *b) {
    (void)hdr;
    if (!h->have_line) {
        ssize_t r = getline(&h->line, &h->cap, h->fp);
        if (r == -1) return -1;
    }
    h->have_line = 0;
    char *line = h->line;
    size_t L = strlen(line);
    while (L && (line[L - 1] == '\n' || line[L - 1] == '\r')) line[--L] = 0;
    if (L == 0) return -1;
    /* split on tabs */
    char *f[11]; int nf = 0; char *p = line; char *rest = NULL;
    while (nf < 11) {
        f[nf++] = p;
        char *t = strchr(p, '\t');
        if (!t) { p = NULL; break; }
        *t = 0; p = t + 1;
    }
    if (nf < 11) return -2;
    rest = p;
    b->l_data = 0;
    size_t lq = strlen(f[0]) + 1;
    shim_bam_put(b, f[0], lq);
    b->core.l_qname = (uint16_t)lq;
    b->core.flag = (uint16_t)atoi(f[1]);
    const char *seq = f[9];
    int ls = (strcmp(seq, "*") == 0) ? 0 : (int)strlen(seq);
    b->core.l_qseq = ls;
    size_t nb = (size_t)(ls + 1) / 2;
    shim_bam_reserve(b, (size_t)b->l_data + nb + (size_t)ls);
    uint8_t *s4 = b->data + b->l_data;
    memset(s4, 0, nb);
    for (int i = 0; i < ls; i++) s4[i >> 1] |= (uint8_t)(shim_nt16(seq[i]) << ((~i & 1) << 2));
    b->l_data += (int)nb;
    memset(b->data + b->l_data, 0xff, (size_t)ls);
    b->l_data += ls;
    b->aux_off = b->l_data;
    /* aux fields */
    while (rest && *rest) {
        char *t = strchr(rest, '\t');
        if (t) *t = 0;
        size_t fl = strlen(rest);
        if (fl >= 5 && rest[2] == ':' && rest[4] == ':') {
            char ty = rest[3]; const char *v = rest + 5;
            shim_bam_put(b, rest, 2);
            if (ty == 'i') {
                long long x = strtoll(v, NULL, 10);
                if (x >= 0) {
                    if (x <= 0xff) { uint8_t c = 'C', y = (uint8_t)x; shim_bam_put(b, &c, 1); shim_bam_put(b, &y, 1); }
                    else if (x <= 0xffff) { uint8_t c = 'S'; uint16_t y = (uint16_t)x; shim_bam_put(b, &c, 1); shim_bam_put(b, &y, 2); }
                    else { uint8_t c = 'I'; uint32_t y = (uint32_t)x; shim_bam_put(b, &c, 1); shim_bam_put(b, &y, 4); }
                } else {
                    if (x >= -128) { uint8_t c = 'c'; int8_t y = (int8_t)x; shim_bam_put(b, &c, 1); shim_bam_put(b, &y, 1); }
                    else if (x >= -32768) { uint8_t c = 's'; int16_t y = (int16_t)x; shim_bam_put(b, &c, 1); shim_bam_put(b, &y, 2); }
                    else { uint8_t c = 'i'; int32_t y = (int32_t)x; shim_bam_put(b, &c, 1); shim_bam_put(b, &y, 4); }
                }
            } else if (ty == 'f') {
                uint8_t c = 'f'; float y = strtof(v, NULL); shim_bam_put(b, &c, 1); shim_bam_put(b, &y, 4);
            } else if (ty == 'A') {
                uint8_t c = 'A'; shim_bam_put(b, &c, 1); shim_bam_put(b, v, 1);
            } else if (ty == 'Z' || ty == 'H') {
                uint8_t c = (uint8_t)ty; shim_bam_put(b, &c, 1); shim_bam_put(b, v, strlen(v) + 1);
            } else if (ty == 'B') {
                char sub = v[0];
                uint32_t n = 0;
                for (const char *q = v + 1; *q; q++) if (*q == ',') n++;
                uint8_t c = 'B'; shim_bam_put(b, &c, 1); shim_bam_put(b, &sub, 1); shim_bam_put(b, &n, 4);
                const char *q = v + 1;
                while (*q == ',') {
                    q++;
                    char *e;
                    if (sub == 'f') { float y = strtof(q, &e); shim_bam_put(b, &y, 4); }
                    else {
                        long long x = strtoll(q, &e, 10);
                        switch (sub) {
                            case 'c': { int8_t y = (int8_t)x; shim_bam_put(b, &y, 1); break; }
                            case 'C': { uint8_t y = (uint8_t)x; shim_bam_put(b, &y, 1); break; }
                            case 's': { int16_t y = (int16_t)x; shim_bam_put(b, &y, 2); break; }
                            case 'S': { uint16_t y = (uint16_t)x; shim_bam_put(b, &y, 2); break; }
                            case 'i': { int32_t y = (int32_t)x; shim_bam_put(b, &y, 4); break; }
                            default:  { uint32_t y = (uint32_t)x; shim_bam_put(b, &y, 4); break; }
                        }
                    }
                    q = e;
                }
            }
        }
        rest = t ? t + 1 : NULL;
    }
    return b->l_data;
}

#define bam_get_qname(b) ((char *)(b)->data)
#define bam_get_seq(b) ((b)->data + (b)->core.l_qname)
#define bam_seqi(s, i) ((s)[(i) >> 1] >> ((~(i) & 1) << 2) & 0xf)

static inline int shim_aux_type_size(uint8_t t) {
    switch (t) { case 'A': case 'c': case 'C': return 1; case 's': case 'S': return 2; case 'i': case 'I': case 'f': return 4; case 'd': return 8; default: return 0; }
}
static inline const uint8_t *shim_aux_skip(const uint8_t *s) {
    uint8_t t = *s++;
    if (t == 'Z' || t == 'H') { while (*s) s++; return s + 1; }
    if (t == 'B') { uint8_t sub = *s++; uint32_t n; memcpy(&n, s, 4); s += 4; return s + (size_t)n * shim_aux_type_size(sub); }
    return s + shim_aux_type_size(t);
}
/* returns a pointer to the TYPE byte, like htslib */
static inline uint8_t *bam_aux_get(const bam1_t *b, const char tag[2]) {
    const uint8_t *s = b->data + b->aux_off, *end = b->data + b->l_data;
    while (s + 3 <= end) {
        if (s[0] == (uint8_t)tag[0] && s[1] == (uint8_t)tag[1]) return (uint8_t *)(s + 2);
        s = shim_aux_skip(s + 2);
    }
    errno = ENOENT;
    return NULL;
}
static inline int64_t bam_aux2i(const uint8_t *s) {
    uint8_t t = *s++;
    switch (t) {
        case 'c': return (int8_t)*s;
        case 'C': return *s;
        case 's': { int16_t v; memcpy(&v, s, 2); return v; }
        case 'S': { uint16_t v; memcpy(&v, s, 2); return v; }
        case 'i': { int32_t v; memcpy(&v, s, 4); return v; }
        case 'I': { uint32_t v; memcpy(&v, s, 4); return v; }
        default: errno = EINVAL; return 0;
    }
}
static inline uint32_t bam_auxB_len(const uint8_t *s) {
    if (s[0] != 'B') { errno = EINVAL; return 0; }
    uint32_t n; memcpy(&n, s + 2, 4); return n;
}
static inline int64_t bam_auxB2i(const uint8_t *s, uint32_t idx) {
    uint32_t len = bam_auxB_len(s);
    if (idx >= len) { errno = ERANGE; return 0; }
    uint8_t sub = s[1];
    s += 6;
    switch (sub) {
        case 'c': return (int8_t)s[idx];
        case 'C': return s[idx];
        case 's': { int16_t v; memcpy(&v, s + 2 * (size_t)idx, 2); return v; }
        case 'S': { uint16_t v; memcpy(&v, s + 2 * (size_t)idx, 2); return v; }
        case 'i': { int32_t v; memcpy(&v, s + 4 * (size_t)idx, 4); return v; }
        case 'I': { uint32_t v; memcpy(&v, s + 4 * (size_t)idx, 4); return v; }
        default: errno = EINVAL; return 0;
    }
}
#endif
