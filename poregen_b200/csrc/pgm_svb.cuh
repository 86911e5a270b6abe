// pgm_svb.cuh — slow5lib's default signal compression ("svb-zd": StreamVByte over zig-zag deltas) decoded in HBM,
// so that the host ships the BLOW5 record's raw_signal field as stored (≈ 1.2–1.4 bytes per sample instead of 2) and
// spends no cycles on it.
//
// Stream of one record (reference slow5lib/src/slow5_press.c:1034-1130, ptr_compress_svb / ptr_depress_svb_zd):
//     u32 count | keys[(count+3)/4] | data[...]
// value i has code (keys[i/4] >> 2(i%4)) & 3 and takes code+1 data bytes, little endian
// (thirdparty/streamvbyte/src/streamvbyte_decode.c:36-78, the scalar decoder); the values are zig-zag coded
// differences, sample[i] = (int16)(prev += (v >> 1) ^ -(v & 1)) with prev = 0 at the record's start
// (thirdparty/streamvbyte/src/streamvbyte_zigzag.c:19-40).
//
// Both recurrences (byte offset of a value, running sum of differences) are prefix sums. One warp per record walks it
// in steps of 256 values, 8 per lane. Three steps are in flight: step c is unpacked from shared memory, the data bytes
// of step c+1 (≤ 1 KiB, any alignment) are on their way there as aligned 16-byte cp.async pieces, and step c+2 is
// placed: its lanes' byte counts come from the 16 key bits with three masked adds and ride in the low half of the ONE
// warp scan per step whose high half carries step c's difference sums (16 bits of them reach an int16). A lane's
// 8 byte offsets come from one multiplication on the codes spread to bytes; values of 1-2 bytes (every real signal)
// unpack without a branch. A lane writes its 8 samples with one 16-byte store when raw_off is a multiple of 8.
#pragma once
#include "pgm_device.cuh"

#define SVB_WARPS 8
#define SVB_CHUNK 256                    // values per warp iteration
#define SVB_NBUF 3
#define SVB_STAGE_WORDS (66 * 4)         // 65 aligned 16-byte words cover 1 KiB at any alignment, + 1 of slack for the look-ahead byte

struct SvbIO {
    const uint8_t *blob;                 // concatenated streams (device)
    const unsigned long long *blob_off;  // [R+1] byte range of record r's stream
    const pgm_read *reads;               // raw_off / raw_len: where the samples go
    int16_t *raw;
    unsigned long long n_raw;
    uint32_t n_reads;
    uint32_t *ticket;                    // zeroed before the launch
    uint32_t *err;                       // |= 1: a stream does not describe raw_len samples
};

// sum of the eight 2-bit fields of x (16 bits)
__device__ __forceinline__ uint32_t svb_sum2(uint32_t x) {
    x = (x & 0x3333u) + ((x >> 2) & 0x3333u);
    x = (x + (x >> 4)) & 0x0f0fu;
    return (x + (x >> 8)) & 0xffu;
}

// One step's 8 values of a lane: byte offsets from the key bits, bytes from the staged copy, zig-zag differences and
// their running sums a[0..7]. FULL: all 256 values of the step exist. WIDE: some value of the step has 3 or 4 bytes
// (a difference of 32768 or more in magnitude: never in real signal, so the 2-byte form carries no branch).
template <bool FULL, bool WIDE>
__device__ __forceinline__ int32_t svb_unpack8(const uint8_t *Sb, uint32_t k, uint32_t nv, uint32_t bpos, int32_t (&a)[8]) {
    // the 2-bit codes spread to bytes, their running sums by one multiplication (no byte exceeds 12), plus 0,1,2,3
    const uint32_t ka1 = (k & 0xffu) * 0x1001u, kb1 = (k >> 8) * 0x1001u;
    const uint32_t tA = (ka1 | (ka1 << 6)) & 0x03030303u, tB = (kb1 | (kb1 << 6)) & 0x03030303u;
    const uint32_t iA = tA * 0x01010101u, iB = tB * 0x01010101u;
    const uint32_t eA = iA - tA + 0x03020100u;
    const uint32_t eB = iB - tB + 0x03020100u + ((iA >> 24) + 4u) * 0x01010101u;
    int32_t run = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const uint32_t code = ((j < 4 ? tA : tB) >> (8 * (j & 3))) & 0xffu;
        const uint32_t bp = bpos + (((j < 4 ? eA : eB) >> (8 * (j & 3))) & 0xffu);
        uint32_t x;
        if (!WIDE) x = (uint32_t)Sb[bp] | (((uint32_t)Sb[bp + 1] & (0u - code)) << 8);    // code is 0 or 1 here
        else {
            x = Sb[bp];
            if (code) x |= (uint32_t)Sb[bp + 1] << 8;
            if (code > 1) x |= (uint32_t)Sb[bp + 2] << 16;
            if (code > 2) x |= (uint32_t)Sb[bp + 3] << 24;
        }
        int32_t d = (int32_t)((x >> 1) ^ (0u - (x & 1u)));
        if (!FULL && (uint32_t)j >= nv) d = 0;
        run += d;
        a[j] = run;
    }
    return run;
}

__device__ __forceinline__ void svb_cp16(uint32_t dst_smem, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst_smem), "l"(src) : "memory");
}

__global__ void __launch_bounds__(SVB_WARPS * 32) k_svb_decode(SvbIO io) {
    __shared__ __align__(16) uint32_t sm_stage[SVB_WARPS][SVB_NBUF][SVB_STAGE_WORDS];
    const uint32_t lane = threadIdx.x & 31;
    uint32_t(*S)[SVB_STAGE_WORDS] = sm_stage[threadIdx.x >> 5];
    const uint32_t s_base = (uint32_t)__cvta_generic_to_shared(&S[0][0]);
    for (;;) {
        uint32_t r = 0;
        if (lane == 0) r = atomicAdd(io.ticket, 1u);
        r = __shfl_sync(0xffffffffu, r, 0);
        if (r >= io.n_reads) break;
        const unsigned long long b0 = io.blob_off[r], b1 = io.blob_off[r + 1];
        const uint32_t want = io.reads[r].raw_len;
        const unsigned long long raw_off = io.reads[r].raw_off;
        const uint8_t *p = io.blob + b0;
        bool bad = b1 < b0 + 4 || raw_off + want > io.n_raw;
        uint32_t count = 0;
        if (!bad) {
            count = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
            bad = count != want;
        }
        const uint32_t keylen = (count + 3) >> 2;
        if (!bad) bad = b0 + 4 + keylen > b1;
        if (bad) {
            if (lane == 0) atomicOr(io.err, 1u);
            continue;
        }
        const uint8_t *keys = p + 4, *data = keys + keylen;
        const unsigned long long data_len = b1 - b0 - 4 - keylen;
        int16_t *out = io.raw + raw_off;
        const bool aligned = (raw_off & 7) == 0;
        const uint32_t nchunks = (count + SVB_CHUNK - 1) / SVB_CHUNK;
        auto load_keys = [&](uint32_t c) -> uint32_t {      // this lane's 16 key bits of step c, zero past the end
            const uint32_t i = c * (SVB_CHUNK / 4) + 2 * lane;
            uint32_t k = 0;
            if (i + 1 < keylen) k = (uint32_t)keys[i] | ((uint32_t)keys[i + 1] << 8);
            else if (i < keylen) k = keys[i];
            return k;
        };
        auto live = [&](uint32_t c) -> uint32_t {           // values of step c this lane owns
            const uint32_t v0 = c * SVB_CHUNK + 8 * lane;
            return (v0 < count) ? min(8u, count - v0) : 0u;
        };
        auto masked = [&](uint32_t kraw, uint32_t nv) -> uint32_t { return kraw & ((nv == 8) ? 0xffffu : ((1u << (2 * nv)) - 1u)); };
        unsigned long long dpos = 0;      // data bytes of the steps placed so far
        bool overrun = false;
        // places step c in the data (its lanes' byte offsets come from the caller's scan) and starts the copy of its
        // bytes into stage buffer `buf`: aligned 16-byte pieces, straight to shared memory. Returns the lane's position.
        auto issue = [&](uint32_t buf, uint32_t excl, uint32_t tot) -> uint32_t {
            uint32_t bpos = 0;
            if (dpos + tot > data_len) overrun = true;
            else {
                const uint8_t *g = data + dpos;
                const uint32_t sh = (uint32_t)((uintptr_t)g & 15);
                const uint8_t *gb = g - sh + 16u * lane;
                const uint32_t n16 = (sh + tot + 15) >> 4;
                const uint32_t dst = s_base + buf * (SVB_STAGE_WORDS * 4) + 16u * lane;
                if (lane < n16) svb_cp16(dst, gb);
                if (lane + 32u < n16) svb_cp16(dst + 512u, gb + 512);
                if (lane + 64u < n16) svb_cp16(dst + 1024u, gb + 1024);
                dpos += tot;
                bpos = sh + excl;
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
            return bpos;
        };
        // steps c (being unpacked), c+1 (copy in flight) and c+2 (placed in this iteration) live in three stage buffers
        uint32_t kA = 0, kB = 0, posA = 0, posB = 0;
        {
            const uint32_t nv0 = live(0);
            kA = masked(load_keys(0), nv0);
            const uint32_t s0 = svb_sum2(kA) + nv0;
            const uint32_t inc0 = warp_incl_scan_add(s0);
            posA = issue(0, inc0 - s0, __shfl_sync(0xffffffffu, inc0, 31));
            const uint32_t nv1 = live(1);
            kB = masked(nchunks > 1 ? load_keys(1) : 0u, nv1);
            const uint32_t s1 = svb_sum2(kB) + nv1;
            const uint32_t inc1 = warp_incl_scan_add(s1);
            posB = issue(1, inc1 - s1, __shfl_sync(0xffffffffu, inc1, 31));
        }
        // key bits are requested three steps before they are needed (a step is shorter than a trip to HBM)
        uint32_t kq = (nchunks > 2) ? load_keys(2) : 0u, kq2 = (nchunks > 3) ? load_keys(3) : 0u, kq3 = (nchunks > 4) ? load_keys(4) : 0u;
        uint32_t bufA = 0, bufC = 2;
        uint32_t prev = 0;                 // running sum of the differences, mod 2^16 (the samples are int16)
        for (uint32_t c = 0; c < nchunks && !overrun; c++) {
            asm volatile("cp.async.wait_group 1;" ::: "memory");     // step c has landed; c+1 may still be on its way
            __syncwarp();
            const uint8_t *Sb = reinterpret_cast<const uint8_t *>(S[bufA]);
            const bool full = (c + 1) * SVB_CHUNK <= count;                      // warp-uniform
            const bool wide = __any_sync(0xffffffffu, (kA & 0xAAAAu) != 0);      // some value of the step takes 3 or 4 bytes
            const uint32_t nv = full ? 8u : live(c);
            int32_t a[8];
            int32_t run;
            if (full && !wide) run = svb_unpack8<true, false>(Sb, kA, nv, posA, a);
            else if (!wide) run = svb_unpack8<false, false>(Sb, kA, nv, posA, a);
            else run = svb_unpack8<false, true>(Sb, kA, nv, posA, a);
            // ONE warp scan serves both recurrences: the byte counts of step c+2 in the low half (a step has at most
            // 1024 bytes), this step's difference sums in the high half (only 16 bits of them reach the output)
            const uint32_t nvC = live(c + 2);
            const uint32_t kC = masked(kq, nvC);
            const uint32_t sC = svb_sum2(kC) + nvC;
            kq = kq2; kq2 = kq3;
            kq3 = (c + 5 < nchunks) ? load_keys(c + 5) : 0u;
            const uint32_t v = ((uint32_t)run << 16) | sC;
            const uint32_t inc = warp_incl_scan_add(v);
            const uint32_t tot = __shfl_sync(0xffffffffu, inc, 31);
            const uint32_t posC = issue(bufC, (inc & 0xffffu) - sC, tot & 0xffffu);
            const uint32_t base = prev + (inc >> 16) - (uint32_t)run;
            prev += tot >> 16;
            const uint32_t v0 = c * SVB_CHUNK + 8 * lane;
            if (nv == 8 && aligned) {
                uint4 o;
                o.x = ((base + (uint32_t)a[0]) & 0xffffu) | ((base + (uint32_t)a[1]) << 16);
                o.y = ((base + (uint32_t)a[2]) & 0xffffu) | ((base + (uint32_t)a[3]) << 16);
                o.z = ((base + (uint32_t)a[4]) & 0xffffu) | ((base + (uint32_t)a[5]) << 16);
                o.w = ((base + (uint32_t)a[6]) & 0xffffu) | ((base + (uint32_t)a[7]) << 16);
                *reinterpret_cast<uint4 *>(out + v0) = o;
            } else {
#pragma unroll
                for (int j = 0; j < 8; j++)
                    if ((uint32_t)j < nv) out[v0 + j] = (int16_t)(uint16_t)(base + (uint32_t)a[j]);
            }
            kA = kB; posA = posB; kB = kC; posB = posC;
            bufC = bufA;
            bufA = (bufA == SVB_NBUF - 1) ? 0u : bufA + 1;
        }
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncwarp();       // the next record reuses the stage buffers
        // slow5lib rejects a stream whose decoder did not consume exactly the stored bytes (slow5_press.c:1097-1103)
        if (overrun || dpos != data_len) {
            if (lane == 0) atomicOr(io.err, 1u);
        }
    }
}
