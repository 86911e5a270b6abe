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
// in chunks of 256 values, 8 per lane: the lane's 16 key bits give its byte count with three masked adds, one warp
// scan places the lanes in the chunk's data, the chunk's data bytes (≤ 1 KiB, any alignment) go through shared
// memory as aligned 16-byte words — the loads of chunk c+1 are issued before chunk c is unpacked — and a second warp
// scan carries the differences. A lane writes its 8 samples with one 16-byte store when raw_off is a multiple of 8.
#pragma once
#include "pgm_device.cuh"

#define SVB_WARPS 8
#define SVB_CHUNK 256                    // values per warp iteration
#define SVB_STAGE_WORDS (66 * 4)         // 65 aligned 16-byte words cover 1 KiB at any alignment, + 1 for the funnel shift

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

__device__ __forceinline__ uint4 svb_ldg16(const uint8_t *p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}

// sum of the eight 2-bit fields of x (16 bits)
__device__ __forceinline__ uint32_t svb_sum2(uint32_t x) {
    x = (x & 0x3333u) + ((x >> 2) & 0x3333u);
    x = (x + (x >> 4)) & 0x0f0fu;
    return (x + (x >> 8)) & 0xffu;
}

__global__ void __launch_bounds__(SVB_WARPS * 32) k_svb_decode(SvbIO io) {
    __shared__ __align__(16) uint32_t sm_stage[SVB_WARPS][SVB_STAGE_WORDS];
    const uint32_t lane = threadIdx.x & 31;
    uint32_t *S = sm_stage[threadIdx.x >> 5];
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
        auto load_keys = [&](uint32_t c) -> uint32_t {      // this lane's 16 key bits of chunk c, zero past the end
            const uint32_t i = c * (SVB_CHUNK / 4) + 2 * lane;
            uint32_t k = 0;
            if (i < keylen) k = keys[i];
            if (i + 1 < keylen) k |= (uint32_t)keys[i + 1] << 8;
            return k;
        };
        // state of the chunk whose data loads are in flight ("next") and of the one being unpacked ("cur")
        uint32_t kq = load_keys(0), kq2 = (nchunks > 1) ? load_keys(1) : 0u;
        unsigned long long dpos = 0;      // data bytes consumed by the chunks before "next"
        uint32_t n_k, n_nv, n_o, n_sh, n_n16;
        uint4 L[3];
        bool overrun = false;
        auto stage_next = [&](uint32_t c) {   // keys of chunk c -> byte placement; issue the loads of its data
            const uint32_t v0 = c * SVB_CHUNK + 8 * lane;
            n_nv = (v0 < count) ? min(8u, count - v0) : 0u;
            n_k = kq & ((n_nv == 8) ? 0xffffu : ((1u << (2 * n_nv)) - 1u));
            const uint32_t s = svb_sum2(n_k) + n_nv;
            const uint32_t inc = warp_incl_scan_add(s);
            const uint32_t tot = __shfl_sync(0xffffffffu, inc, 31);
            n_o = inc - s;
            if (dpos + tot > data_len) { overrun = true; n_n16 = 0; return; }
            const uint8_t *g = data + dpos;
            n_sh = (uint32_t)((uintptr_t)g & 15);
            const uint8_t *gb = g - n_sh;
            n_n16 = (n_sh + tot + 15) >> 4;
#pragma unroll
            for (int t = 0; t < 3; t++) {
                const uint32_t i = lane + 32u * t;
                if (i < n_n16) L[t] = svb_ldg16(gb + 16u * i);
            }
            dpos += tot;
            kq = kq2;
            kq2 = (c + 2 < nchunks) ? load_keys(c + 2) : 0u;
        };
        int32_t prev = 0;
        if (nchunks) stage_next(0);
        for (uint32_t c = 0; c < nchunks && !overrun; c++) {
            // the loaded words of chunk c go to shared memory; then chunk c+1 is placed and its loads issued
            __syncwarp();
#pragma unroll
            for (int t = 0; t < 3; t++) {
                const uint32_t i = lane + 32u * t;
                if (i < n_n16) reinterpret_cast<uint4 *>(S)[i] = L[t];
            }
            const uint32_t k = n_k, nv = n_nv;
            uint32_t bpos = n_sh + n_o;
            if (c + 1 < nchunks) stage_next(c + 1);
            __syncwarp();
            int32_t a[8];
            int32_t run = 0;
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const uint32_t code = (k >> (2 * j)) & 3u;
                const uint32_t w = bpos >> 2;
                uint32_t x = __funnelshift_r(S[w], S[w + 1], (bpos & 3u) * 8u);
                x &= 0xffffffffu >> (8u * (3u - code));
                const bool live = (uint32_t)j < nv;
                if (live) bpos += code + 1u;
                const int32_t d = live ? (int32_t)((x >> 1) ^ (0u - (x & 1u))) : 0;
                run += d;
                a[j] = run;
            }
            // carry of the differences across lanes and chunks (int32 wrap-around = the reference's arithmetic)
            uint32_t inc = warp_incl_scan_add((uint32_t)run);
            const int32_t base = prev + (int32_t)(inc - (uint32_t)run);
            prev += (int32_t)__shfl_sync(0xffffffffu, inc, 31);
            const uint32_t v0 = c * SVB_CHUNK + 8 * lane;
            if (nv == 8 && aligned) {
                uint4 o;
                o.x = ((uint32_t)(base + a[0]) & 0xffffu) | ((uint32_t)(base + a[1]) << 16);
                o.y = ((uint32_t)(base + a[2]) & 0xffffu) | ((uint32_t)(base + a[3]) << 16);
                o.z = ((uint32_t)(base + a[4]) & 0xffffu) | ((uint32_t)(base + a[5]) << 16);
                o.w = ((uint32_t)(base + a[6]) & 0xffffu) | ((uint32_t)(base + a[7]) << 16);
                *reinterpret_cast<uint4 *>(out + v0) = o;
            } else {
#pragma unroll
                for (int j = 0; j < 8; j++)
                    if ((uint32_t)j < nv) out[v0 + j] = (int16_t)(base + a[j]);
            }
        }
        // slow5lib rejects a stream whose decoder did not consume exactly the stored bytes (slow5_press.c:1097-1103)
        if (overrun || dpos != data_len) {
            if (lane == 0) atomicOr(io.err, 1u);
        }
    }
}
