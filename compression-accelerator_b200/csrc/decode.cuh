// decode.cuh -- Snappy raw-stream decompressor, one warp per stream.
//
// Behind uncompress(source, compressed_length, dest) of the reference's driver
// (tests/compression.c:32-39; the reference hardware left the command a stub,
// CompressionAccelerator.scala:385-388).  Follows tests/snappy/format_description.txt and the tag
// semantics of SnappyDecompressor::DecompressAllTags (tests/snappy/snappy.cc:811-911): varint32
// length, then literal / copy-1 / copy-2 / copy-4 elements.  Element boundaries can only be found
// by walking the stream, so a stream is decoded by ONE warp: the walk is warp-uniform, the bytes of
// every element are moved by all 32 lanes (overlapping copies replicate the offset-long pattern,
// so even offset < length copies move in parallel).  Parallelism comes from many streams at once
// (the page-shaped batch); a single large stream is decoded correctly but at one warp's pace.
// Hardened like the reference's decompressor (snappy_unittest.cc:536-): every length and offset is
// checked; a malformed stream yields kDecodeError, never an out-of-bounds access.
#pragma once
#include "kernels.h"
#include "snappy_device.cuh"

namespace sb {

constexpr uint32_t kDecodeError = 0xffffffffu;

struct DecodeParams {
    const uint8_t* in;            // stream j at in + j*in_stride
    uint64_t in_stride;
    const uint32_t* in_sizes;     // compressed length of stream j (nullptr: every stream is in_size bytes)
    uint32_t in_size;
    uint32_t n_jobs;
    uint8_t* out;                 // stream j decodes to out + j*out_stride
    uint64_t out_stride;
    uint64_t out_cap;             // capacity per stream
    uint32_t* out_lens;           // decoded length of stream j, kDecodeError if malformed (may be nullptr)
    unsigned long long* out_len64;// single-stream flavour: length or ~0ull (may be nullptr)
    // indexed flavour (framed streams, frame.cuh): when in_offs is set, stream j = in + in_offs[j]
    // (in_sizes[j] bytes) decodes to out + out_offs[j] with capacity caps[j]; kinds[j] == 1 marks
    // a stored (uncompressed) chunk that is copied as it is
    const unsigned long long* in_offs = nullptr;
    const unsigned long long* out_offs = nullptr;
    const uint32_t* caps = nullptr;
    const uint32_t* kinds = nullptr;
    // decode_par.cuh: blocks of one large stream have no preamble of their own (their decoded length
    // is caps[j]); the block decode is skipped when *skip_if != 0, the serial walk runs only when
    // *only_if != 0 -- both flags live on the device, the host never waits to choose
    uint32_t headerless = 0;
    const uint32_t* skip_if = nullptr;
    const uint32_t* only_if = nullptr;
};

// n bytes d[0..n) = s[0..n), s read-only input (any alignment), all lanes of the warp
__device__ __forceinline__ void decode_literal(uint8_t* __restrict__ d, const uint8_t* __restrict__ s,
                                               uint32_t n, uint32_t lane) {
    if (n <= 64) {
        for (uint32_t i = lane; i < n; i += 32) d[i] = __ldg(s + i);
        return;
    }
    uint32_t head = (uint32_t)((16u - ((uintptr_t)d & 15u)) & 15u);
    if (lane < head) d[lane] = __ldg(s + lane);
    const uint32_t body = (n - head) >> 4;
    uint4* d4 = reinterpret_cast<uint4*>(d + head);
    const uint8_t* sb_ = s + head;
    const uint32_t* w = reinterpret_cast<const uint32_t*>((uintptr_t)sb_ & ~(uintptr_t)3);
    const uint32_t sh = ((uint32_t)(uintptr_t)sb_ & 3u) << 3;
    for (uint32_t j = lane; j < body; j += 32) {
        const uint32_t* q = w + (j << 2);
        const uint32_t a0 = __ldg(q), a1 = __ldg(q + 1), a2 = __ldg(q + 2), a3 = __ldg(q + 3);
        const uint32_t a4 = sh ? __ldg(q + 4) : 0u;
        uint4 v;
        v.x = shf_r(a0, a1, sh);
        v.y = shf_r(a1, a2, sh);
        v.z = shf_r(a2, a3, sh);
        v.w = shf_r(a3, a4, sh);
        d4[j] = v;
    }
    const uint32_t done = head + (body << 4);
    if (done + lane < n) d[done + lane] = __ldg(s + done + lane);
}

// n bytes at d[0..n) = the bytes `off` back, pattern-replicated when off < n (snappy.cc:744-790
// IncrementalCopy).  Plain (L1-cached) loads: the bytes were written by lanes of this same warp, and the
// __syncwarp() after every element orders those writes before these reads (same CTA: no L1 staleness).
__device__ __forceinline__ void decode_copy(uint8_t* d, uint32_t off, uint32_t n, uint32_t lane) {
    const uint8_t* s = d - off;
    if (off >= n) {
        for (uint32_t i = lane; i < n; i += 32) d[i] = s[i];
    } else {
        for (uint32_t i = lane; i < n; i += 32) d[i] = s[i % off];
    }
}

// One stream, input and output in global memory.  Returns the decoded length or kDecodeError.
__device__ __forceinline__ uint32_t decode_stream_global(const uint8_t* __restrict__ src, uint32_t n,
                                                         uint8_t* dst, uint64_t out_cap, uint32_t lane,
                                                         bool headerless = false) {
    bool bad = false;
    // varint32 preamble (snappy.cc:1006-1009 on the way in, :864- on the way out)
    uint32_t ip = 0, ulen = 0, shift = 0;
    if (headerless) ulen = (uint32_t)out_cap;
    else for (;;) {
        if (ip >= n || shift > 28) { bad = true; break; }
        const uint32_t b = __ldg(src + ip++);
        if (shift == 28 && (b & 0x7fu) > 15u) { bad = true; break; }   // would not fit 32 bits
        ulen |= (b & 0x7fu) << shift;
        if (b < 128) break;
        shift += 7;
    }
    if (!bad && (uint64_t)ulen > out_cap) bad = true;
    uint32_t op = 0;
    while (!bad && ip < n) {
        const uint32_t tag = __ldg(src + ip++);
        const uint32_t kind = tag & 3u;
        if (kind == 0) {                                         // literal
            uint32_t len = (tag >> 2) + 1;
            if (len > 60) {
                const uint32_t extra = len - 60;
                if (ip + extra > n) { bad = true; break; }
                uint32_t v = 0;
                for (uint32_t i = 0; i < extra; ++i) v |= (uint32_t)__ldg(src + ip + i) << (8 * i);
                if (v == 0xffffffffu) { bad = true; break; }
                len = v + 1;
                ip += extra;
            }
            if (len > n - ip || len > ulen - op) { bad = true; break; }
            decode_literal(dst + op, src + ip, len, lane);
            ip += len;
            op += len;
        } else {                                                 // copy-1 / copy-2 / copy-4
            uint32_t len, off;
            if (kind == 1) {
                if (ip + 1 > n) { bad = true; break; }
                len = 4 + ((tag >> 2) & 7u);
                off = ((tag >> 5) << 8) | __ldg(src + ip);
                ip += 1;
            } else if (kind == 2) {
                if (ip + 2 > n) { bad = true; break; }
                len = (tag >> 2) + 1;
                off = __ldg(src + ip) | ((uint32_t)__ldg(src + ip + 1) << 8);
                ip += 2;
            } else {
                if (ip + 4 > n) { bad = true; break; }
                len = (tag >> 2) + 1;
                off = __ldg(src + ip) | ((uint32_t)__ldg(src + ip + 1) << 8) |
                      ((uint32_t)__ldg(src + ip + 2) << 16) | ((uint32_t)__ldg(src + ip + 3) << 24);
                ip += 4;
            }
            if (off == 0 || off > op || len > ulen - op) { bad = true; break; }
            decode_copy(dst + op, off, len, lane);
            op += len;
        }
        __syncwarp();                                            // the next element may read these bytes
    }
    if (!bad && op != ulen) bad = true;
    return bad ? kDecodeError : op;
}

#ifndef SB_DECODE_MINB
#define SB_DECODE_MINB 16                // 32 registers: 64 warps per SM (the walk is latency bound)
#endif
__global__ void __launch_bounds__(128, SB_DECODE_MINB) decode_kernel(DecodeParams p) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t job = blockIdx.x * 4 + (threadIdx.x >> 5);
    if (job >= p.n_jobs) return;
    if ((p.skip_if && *p.skip_if) || (p.only_if && !*p.only_if)) return;
    const uint8_t* __restrict__ src = p.in + (p.in_offs ? p.in_offs[job] : (uint64_t)job * p.in_stride);
    const uint32_t n = p.in_sizes ? p.in_sizes[job] : p.in_size;
    uint8_t* dst = p.out + (p.out_offs ? p.out_offs[job] : (uint64_t)job * p.out_stride);
    const uint64_t cap = p.caps ? (uint64_t)p.caps[job] : p.out_cap;
    uint32_t got;
    if (p.kinds && p.kinds[job] == 1) {                  // stored chunk
        got = n <= cap ? n : kDecodeError;
        if (got != kDecodeError) decode_literal(dst, src, n, lane);
    } else {
        got = decode_stream_global(src, n, dst, cap, lane, p.headerless != 0);
    }
    if (lane == 0) {
        if (p.out_lens) p.out_lens[job] = got;
        if (p.out_len64) *p.out_len64 = got == kDecodeError ? ~0ull : (unsigned long long)got;
    }
}

// ---- page-sized streams: input and output staged in shared memory ------------------------------
// A copy element reads bytes this warp wrote a moment ago; from global memory that is an L2 round
// trip per element on the walk's critical path.  For streams that fit (the page-shaped batch of
// benchmark/hw_results.csv) the warp keeps its compressed input AND its output in shared memory:
// every element costs shared-memory latency, and global memory sees one coalesced read of the
// stream and one coalesced write of the page.  Streams that do not fit the staging area take
// decode_stream_global.
constexpr uint32_t kStageWarps = 4;
constexpr uint32_t kStageOutMax = 8192;          // largest decoded size the staged kernel is launched for

// unaligned 4 bytes from shared memory (the staging areas are padded: reading 3 bytes past is safe)
__device__ __forceinline__ uint32_t stage_ld32(const uint8_t* s, uint32_t pos) {
    const uint32_t* w = reinterpret_cast<const uint32_t*>(s) + (pos >> 2);
    return shf_r(w[0], w[1], (pos & 3u) << 3);
}

__global__ void __launch_bounds__(32 * kStageWarps) decode_staged_kernel(DecodeParams p, uint32_t stage_in,
                                                                         uint32_t stage_out) {
    extern __shared__ __align__(16) uint8_t smem[];
    const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const uint32_t job = blockIdx.x * kStageWarps + wid;
    if (job >= p.n_jobs) return;
    uint8_t* s_in = smem + (size_t)wid * (stage_in + stage_out);
    uint8_t* s_out = s_in + stage_in;
    const uint8_t* __restrict__ src = p.in + (uint64_t)job * p.in_stride;
    const uint32_t n = p.in_sizes ? p.in_sizes[job] : p.in_size;
    uint8_t* dst = p.out + (uint64_t)job * p.out_stride;
    uint32_t got;
    const uint32_t skew = (uint32_t)((uintptr_t)src & 15u);
    if (n == 0 || (uint64_t)skew + n + 20 > stage_in) {
        got = n ? decode_stream_global(src, n, dst, p.out_cap, lane) : kDecodeError;
    } else {
        // stage the stream: whole aligned 16-byte words (a word holding one valid byte lies in a valid page)
        {
            const uint4* g = reinterpret_cast<const uint4*>(src - skew);
            uint4* d = reinterpret_cast<uint4*>(s_in);
            const uint32_t words = (skew + n + 15) >> 4;
            for (uint32_t i = lane; i < words; i += 32) d[i] = __ldg(g + i);
        }
        __syncwarp();
        const uint8_t* in = s_in + skew;                 // in[0..n) is the stream
        bool bad = false;
        uint32_t ip = 0, ulen = 0, shift = 0;
        for (;;) {                                       // varint32 preamble
            if (ip >= n || shift > 28) { bad = true; break; }
            const uint32_t b = in[ip++];
            if (shift == 28 && (b & 0x7fu) > 15u) { bad = true; break; }
            ulen |= (b & 0x7fu) << shift;
            if (b < 128) break;
            shift += 7;
        }
        if (!bad && ((uint64_t)ulen > p.out_cap || ulen > stage_out)) {
            // cannot be staged (or does not fit the caller's buffer: the global walk reports that)
            got = decode_stream_global(src, n, dst, p.out_cap, lane);
        } else {
            uint32_t op = 0;
            while (!bad && ip < n) {
                const uint32_t w = stage_ld32(s_in, skew + ip);      // tag + up to 3 following bytes
                const uint32_t tag = w & 0xffu, kind = tag & 3u;
                ++ip;
                if (kind == 0) {                                     // literal
                    uint32_t len = (tag >> 2) + 1;
                    if (len > 60) {
                        const uint32_t extra = len - 60;
                        if (ip + extra > n) { bad = true; break; }
                        const uint32_t v = extra == 4 ? stage_ld32(s_in, skew + ip)
                                                      : (w >> 8) & (0xffffffu >> (8 * (3 - extra)));
                        if (v == 0xffffffffu) { bad = true; break; }
                        len = v + 1;
                        ip += extra;
                    }
                    if (len > n - ip || len > ulen - op) { bad = true; break; }
                    for (uint32_t i = lane; i < len; i += 32) s_out[op + i] = in[ip + i];
                    ip += len;
                    op += len;
                } else {                                             // copy-1 / copy-2 / copy-4
                    uint32_t len, off;
                    if (kind == 1) {
                        if (ip + 1 > n) { bad = true; break; }
                        len = 4 + ((tag >> 2) & 7u);
                        off = ((tag >> 5) << 8) | ((w >> 8) & 0xffu);
                        ip += 1;
                    } else if (kind == 2) {
                        if (ip + 2 > n) { bad = true; break; }
                        len = (tag >> 2) + 1;
                        off = (w >> 8) & 0xffffu;
                        ip += 2;
                    } else {
                        if (ip + 4 > n) { bad = true; break; }
                        len = (tag >> 2) + 1;
                        off = stage_ld32(s_in, skew + ip);
                        ip += 4;
                    }
                    if (off == 0 || off > op || len > ulen - op) { bad = true; break; }
                    const uint8_t* from = s_out + op - off;          // all source bytes precede op
                    if (off >= len) {
                        for (uint32_t i = lane; i < len; i += 32) s_out[op + i] = from[i];
                    } else {
                        for (uint32_t i = lane; i < len; i += 32) s_out[op + i] = from[i % off];
                    }
                    op += len;
                }
                __syncwarp();                                        // the next element may read these bytes
            }
            if (!bad && op != ulen) bad = true;
            got = bad ? kDecodeError : op;
            if (!bad) {                                              // one coalesced write of the page
                __syncwarp();
                if (((uintptr_t)dst & 15u) == 0) {
                    const uint4* s4 = reinterpret_cast<const uint4*>(s_out);
                    uint4* d4 = reinterpret_cast<uint4*>(dst);
                    const uint32_t words = op >> 4;
                    for (uint32_t i = lane; i < words; i += 32) d4[i] = s4[i];
                    for (uint32_t i = (words << 4) + lane; i < op; i += 32) dst[i] = s_out[i];
                } else {
                    for (uint32_t i = lane; i < op; i += 32) dst[i] = s_out[i];
                }
            }
        }
    }
    if (lane == 0) {
        if (p.out_lens) p.out_lens[job] = got;
        if (p.out_len64) *p.out_len64 = got == kDecodeError ? ~0ull : (unsigned long long)got;
    }
}

}  // namespace sb
