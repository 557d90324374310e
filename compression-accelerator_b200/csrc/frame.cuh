// frame.cuh -- Snappy framing format (tests/snappy/framing_format.txt) on the device.
//
// A framed stream is the 10-byte stream identifier (framing_format.txt:68-75) followed by chunks
// of at most 65,536 uncompressed bytes (:88-91), each: type byte, 3-byte LE length, masked
// CRC-32C of the UNCOMPRESSED bytes (:38-57), then an ordinary Snappy stream of the chunk (:80-86)
// -- exactly a fragment of the raw path with its own varint preamble, so the writer reuses the
// parse / size / scan / emit kernels with per-job headers and only adds the CRC pass and the
// chunk headers.  Chunks are independent: unlike a raw stream, a framed one decodes in parallel.
//
// CRC-32C (Castagnoli, reflected polynomial 0x82F63B78; RFC 3720 12.1, vectors B.4): one CTA per
// chunk, 512 threads: the chunk is staged in shared memory with coalesced 16-byte loads (one padding
// word per slice keeps the slices in different banks), each thread takes the raw remainder (init 0)
// of its 128-byte slice byte by byte from a lane-private copy of the 256-entry table; slices are combined in a tree with the linear operator "append 2^k
// zero bytes" (32x32 bit matrices, precomputed by the host into constant memory).  The chunk is
// laid out right-aligned in a virtual 64 KiB so every combine step shifts by a fixed amount; the
// effect of the all-ones initial value, crc_update(0xFFFFFFFF, len zero bytes), is a product of
// those operators over the bits of len.
#pragma once
#include "kernels.h"
#include "snappy_device.cuh"
#include "parse_pipe.cuh"   // GlobalSrc: unaligned, bounds-aware global loads

namespace sb {

constexpr uint32_t kCrcPoly = 0x82F63B78u;
constexpr uint32_t kCrcMaskDelta = 0xa282ead8u;      // framing_format.txt:46-53
constexpr uint32_t kFrameChunk = 65536;              // :88-91
constexpr uint32_t kCrcThreads = 512;
constexpr uint32_t kFrameIdLen = 10;

struct CrcOps { uint32_t pow[17][32]; };             // pow[k][j] = image of bit j under "append 2^k zero bytes"
__constant__ CrcOps g_crc_ops;

__host__ __device__ inline uint32_t crc_mask(uint32_t x) { return ((x >> 15) | (x << 17)) + kCrcMaskDelta; }
__host__ __device__ inline uint32_t crc_unmask(uint32_t m) { const uint32_t r = m - kCrcMaskDelta; return (r >> 17) | (r << 15); }

inline void crc_ops_build(CrcOps& o) {               // host, once
    for (int j = 0; j < 32; ++j) {
        uint32_t v = 1u << j;
        for (int b = 0; b < 8; ++b) v = (v >> 1) ^ ((v & 1u) ? kCrcPoly : 0u);
        o.pow[0][j] = v;
    }
    for (int k = 1; k < 17; ++k)
        for (int j = 0; j < 32; ++j) {
            uint32_t v = o.pow[k - 1][j], r = 0;
            for (int b = 0; b < 32; ++b) if ((v >> b) & 1u) r ^= o.pow[k - 1][b];
            o.pow[k][j] = r;
        }
}

__device__ __forceinline__ uint32_t crc_append_zeros(uint32_t v, int k) {   // 2^k zero bytes
    // four independent XOR chains: the tree of the CRC kernel waits on this function's latency
    uint32_t r0 = 0, r1 = 0, r2 = 0, r3 = 0;
    const uint32_t* m = g_crc_ops.pow[k];
#pragma unroll
    for (int b = 0; b < 8; ++b) {
        r0 ^= (0u - ((v >> b) & 1u)) & m[b];
        r1 ^= (0u - ((v >> (b + 8)) & 1u)) & m[b + 8];
        r2 ^= (0u - ((v >> (b + 16)) & 1u)) & m[b + 16];
        r3 ^= (0u - ((v >> (b + 24)) & 1u)) & m[b + 24];
    }
    return (r0 ^ r1) ^ (r2 ^ r3);
}

struct CrcParams {
    const uint8_t* base;                 // chunk c = base + (offs ? offs[c] : c*stride)
    const unsigned long long* offs;
    const uint32_t* lens;                // nullptr: min(fixed_len, n_total - c*stride)
    uint64_t stride, n_total;
    uint32_t fixed_len, n_chunks;
    uint32_t* out_masked;                // masked CRC-32C of chunk c
};

// Shared-memory image of the chunk: the aligned 16-byte words covering it, one padding word after every
// slice so that the threads' slices start in different banks.  The byte table is replicated 32 times,
// entry i of lane l at word 32*i + l: every lane reads its own bank, so a table look-up of the warp is
// one conflict-free wavefront (a shared 256-entry table serialises ~3.5-fold on random indices).
constexpr uint32_t kCrcSliceLog = 7;                               // 128-byte slices, 512 threads
constexpr uint32_t kCrcSlice = 1u << kCrcSliceLog;
static_assert(kCrcThreads * kCrcSlice == kFrameChunk, "one slice per thread");
constexpr uint32_t kCrcStageWords = (kFrameChunk + 32) / 4 + (kFrameChunk + 32) / kCrcSlice + 8;
constexpr uint32_t kCrcSmemBytes = kCrcStageWords * 4 + 256 * 32 * 4 + 2 * kCrcThreads * 4;
__device__ __forceinline__ uint32_t crc_stage_word(uint32_t y) { return (y >> 2) + (y >> kCrcSliceLog); }   // y % 4 == 0

__global__ void __launch_bounds__(kCrcThreads) crc32c_chunks_kernel(CrcParams p) {
    extern __shared__ __align__(16) uint32_t crc_smem[];
    uint32_t* sw = crc_smem;                                       // staged chunk
    uint32_t* T = crc_smem + kCrcStageWords;                       // T[32*i + lane]
    uint32_t (*part)[kCrcThreads] = reinterpret_cast<uint32_t (*)[kCrcThreads]>(crc_smem + kCrcStageWords + 256 * 32);
    const uint32_t t = threadIdx.x, lane = t & 31;
    for (uint32_t i = t >> 5; i < 256; i += kCrcThreads >> 5) {    // warp w fills entries w, w+16, ...: all 32 copies
        uint32_t v = i;
        for (int b = 0; b < 8; ++b) v = (v >> 1) ^ ((v & 1u) ? kCrcPoly : 0u);
        T[(i << 5) + lane] = v;
    }
    // persistent: the table is built once per CTA, chunks are taken round-robin.  Every warp stages
    // and reads only its own 32 slices (4 KiB of the image), so the warps of a CTA synchronise once
    // per chunk -- when the 16 warp remainders are merged -- and not between load and compute.
    __syncthreads();
    const uint32_t* Tl = T + lane;
    const uint32_t wid = t >> 5;
    uint32_t parity = 0;
    for (uint32_t c = blockIdx.x; c < p.n_chunks; c += gridDim.x, parity ^= 1u) {
        const uint64_t off = p.offs ? p.offs[c] : (uint64_t)c * p.stride;
        uint32_t len;
        if (p.lens) len = p.lens[c];
        else { const uint64_t left = p.n_total - off; len = left < p.fixed_len ? (uint32_t)left : p.fixed_len; }
        if (len > kFrameChunk) len = kFrameChunk;        // callers never ask for more
        const uint8_t* g = p.base + off;
        const uint32_t skew = (uint32_t)((uintptr_t)g & 15u);
        // right-aligned: virtual position v holds data byte v - pad = image byte v - pad + skew
        const uint32_t pad = kFrameChunk - len;
        uint32_t v0 = t * kCrcSlice, v1 = v0 + kCrcSlice, crc = 0;
        if (v0 < pad) v0 = pad;
        {   // the warp's part of the image: whole aligned 16-byte words (a word holding one valid byte
            // lies in a valid page); neighbouring warps write the same bytes into a shared word
            uint32_t w0 = wid * 32u * kCrcSlice, w1 = w0 + 32u * kCrcSlice;
            if (w0 < pad) w0 = pad;
            if (w0 < w1) {
                const uint32_t lo16 = (w0 - pad + skew) >> 4, hi16 = (w1 - pad + skew + 15) >> 4;
                const uint4* g4 = reinterpret_cast<const uint4*>(g - skew);
#pragma unroll 3
                for (uint32_t j = lo16 + lane; j < hi16; j += 32) {
                    const uint4 v = __ldg(g4 + j);
                    const uint32_t w = crc_stage_word(j << 4);     // 16 bytes never straddle a slice row
                    sw[w] = v.x; sw[w + 1] = v.y; sw[w + 2] = v.z; sw[w + 3] = v.w;
                }
            }
        }
        __syncwarp();
        if (v0 < v1) {
            uint32_t x = v0 - pad + skew;                          // position in the staged image
            const uint32_t xe = v1 - pad + skew;
            const uint8_t* sb8 = reinterpret_cast<const uint8_t*>(sw);
            for (; ((xe - x) & 3u) != 0; ++x)                      // leading bytes: the rest is whole words
                crc = Tl[((crc ^ sb8[(crc_stage_word(x & ~3u) << 2) + (x & 3u)]) & 0xffu) << 5] ^ (crc >> 8);
            const uint32_t sh = (x & 3u) << 3;
            uint32_t y = x & ~3u;
            uint32_t lo = sw[crc_stage_word(y)];
            for (; x < xe; x += 4) {
                y += 4;
                const uint32_t hi = sw[crc_stage_word(y)];
                crc ^= shf_r(lo, hi, sh);
                lo = hi;
                crc = Tl[(crc & 0xffu) << 5] ^ (crc >> 8);
                crc = Tl[(crc & 0xffu) << 5] ^ (crc >> 8);
                crc = Tl[(crc & 0xffu) << 5] ^ (crc >> 8);
                crc = Tl[(crc & 0xffu) << 5] ^ (crc >> 8);
            }
        }
        __syncwarp();                                              // the image rows may be overwritten by the next chunk
        // merge: neighbouring runs of kCrcSlice * 2^k bytes, left ++ right = shift(left) ^ right; first
        // inside the warp (every lane ends with the warp's remainder), then the 16 warps in warp 0
#pragma unroll 1
        for (int k = 0; k < 5; ++k) {
            const uint32_t other = __shfl_xor_sync(kFull, crc, 1u << k);
            const bool right = (lane >> k) & 1u;
            crc = crc_append_zeros(right ? other : crc, kCrcSliceLog + k) ^ (right ? crc : other);
        }
        if (lane == 0) part[parity][wid] = crc;
        __syncthreads();
        if (wid == 0) {
            uint32_t r = part[parity][lane & 15];
#pragma unroll 1
            for (int k = 0; k < 4; ++k) {
                const uint32_t other = __shfl_xor_sync(kFull, r, 1u << k);
                const bool right = (lane >> k) & 1u;
                r = crc_append_zeros(right ? other : r, kCrcSliceLog + 5 + k) ^ (right ? r : other);
            }
            if (lane == 0) {
                uint32_t init = 0xffffffffu;             // crc_update(~0, len zero bytes)
#pragma unroll 1
                for (int k = 0; k < 17; ++k) if ((len >> k) & 1u) init = crc_append_zeros(init, k);
                p.out_masked[c] = crc_mask(~(r ^ init));
            }
        }
    }
}

// ---- writer ------------------------------------------------------------------------------------
// sizes[j] (chunk's Snappy stream, varint included) -> bytes of the whole chunk (header + CRC + stream)
__global__ void frame_bump_sizes_kernel(uint32_t* sizes, uint32_t n) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n) sizes[j] += 8u;
}

// stream identifier + every chunk's 4-byte header and masked CRC (the Snappy stream follows at +8)
__global__ void frame_headers_kernel(uint8_t* dst, const unsigned long long* offsets, const uint32_t* sizes,
                                     const uint32_t* crcs, uint32_t n) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j == 0) {
        const uint8_t id[kFrameIdLen] = {0xff, 0x06, 0x00, 0x00, 0x73, 0x4e, 0x61, 0x50, 0x70, 0x59};
        for (uint32_t i = 0; i < kFrameIdLen; ++i) dst[i] = id[i];
    }
    if (j >= n) return;
    uint8_t* d = dst + offsets[j];
    const uint32_t l = sizes[j] - 4u, crc = crcs[j];     // chunk length field: CRC + stream
    d[0] = 0x00; d[1] = (uint8_t)l; d[2] = (uint8_t)(l >> 8); d[3] = (uint8_t)(l >> 16);
    d[4] = (uint8_t)crc; d[5] = (uint8_t)(crc >> 8); d[6] = (uint8_t)(crc >> 16); d[7] = (uint8_t)(crc >> 24);
}

// ---- reader ------------------------------------------------------------------------------------
struct FrameIndex {                      // one record per data chunk, filled by the walk
    unsigned long long* in_off;          // payload (after header + CRC) in the framed stream
    unsigned long long* out_off;         // where the chunk's bytes go
    uint32_t* in_len;                    // payload bytes
    uint32_t* ulen;                      // uncompressed bytes (from the payload's varint / the chunk length)
    uint32_t* crc;                       // stored masked CRC
    uint32_t* kind;                      // 0 compressed, 1 uncompressed
};

// index_base holds max_records records as parallel arrays (kFrameRecordBytes each, kernels.h):
// in_off, out_off (u64), in_len, ulen, crc, kind, then the decode's got_len, got_crc (u32), then the walk's
// per-chunk scratch (every chunk of the stream, data or not): position (u64), payload length,
// uncompressed length, CRC, kind (u32).
struct FrameScratch {
    FrameIndex ix; uint32_t* got_len; uint32_t* got_crc;
    unsigned long long* t_pos; uint32_t* t_plen; uint32_t* t_ulen; uint32_t* t_crc; uint32_t* t_kind;
};
__host__ __device__ inline FrameScratch frame_scratch(void* base, uint64_t m) {
    FrameScratch f;
    f.ix.in_off = (unsigned long long*)base;
    f.ix.out_off = f.ix.in_off + m;
    f.t_pos = f.ix.out_off + m;
    f.ix.in_len = (uint32_t*)(f.t_pos + m);
    f.ix.ulen = f.ix.in_len + m;
    f.ix.crc = f.ix.ulen + m;
    f.ix.kind = f.ix.crc + m;
    f.got_len = f.ix.kind + m;
    f.got_crc = f.got_len + m;
    f.t_plen = f.got_crc + m;
    f.t_ulen = f.t_plen + m;
    f.t_crc = f.t_ulen + m;
    f.t_kind = f.t_crc + m;
    return f;
}
constexpr uint32_t kFrameNotData = 0xffffffffu;

// The chunk chain is a linked list through the stream (16 K hops per GiB).  Only the hop itself is
// sequential -- the 4 bytes at a chunk's start give the next chunk's position -- so the walk is split:
//   frame_hop_kernel    one warp, all lanes in step: position of every chunk.  A hop is ~25 instructions and
//                       one round trip (two aligned words, the same in every lane); between hops one TMA bulk
//                       prefetch per 32 KiB keeps the next 512 KiB of the stream on its way into L2, so that
//                       round trip is an L2 hit.  (A lone warp runs one dependent instruction every ~6 cycles: the full parse of
//                       a header, ~180 instructions, cost more per hop than the memory did.)
//   frame_parse_kernel  one thread per chunk: identifier / type / length checks, CRC, the varint preamble
//                       of the chunk's own stream (framing_format.txt:20-21, :68-78, :80-106, :122-128)
//   frame_index_kernel  one CTA: data chunks compacted in order, output offsets by an exclusive scan
__global__ void __launch_bounds__(32) frame_hop_kernel(const uint8_t* __restrict__ src, uint64_t n,
                                                       unsigned long long* t_pos, uint64_t max_records,
                                                       FrameWalkResult* res) {
    const uint32_t lane = threadIdx.x;
    uint64_t pos = 0, cnt = 0, pf = 0;
    uint32_t err = 0;
    const uintptr_t end = (uintptr_t)(src + n);
    while (pos < n) {
        while (pf < n && pf < pos + (512u << 10)) {                          // stream -> L2: one TMA bulk prefetch
            const uintptr_t b = ((uintptr_t)src + pf) & ~(uintptr_t)15;      // of 32 KiB per round (UBLKPF)
            const uintptr_t e = (end & ~(uintptr_t)15) < b + 32768u ? (end & ~(uintptr_t)15) : b + 32768u;
            if (lane == 0 && e > b)
                asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(b), "r"((uint32_t)(e - b)) : "memory");
            pf += 32768;
        }
        if (pos + 4 > n) { err = 1; break; }
        const uintptr_t a0 = (uintptr_t)(src + pos);
        const uint32_t* wp = (const uint32_t*)(a0 & ~(uintptr_t)3);
        const uint32_t w0 = __ldcg(wp), w1 = ((uintptr_t)(wp + 1) < end) ? __ldcg(wp + 1) : 0u;
        const uint32_t len = shf_r(w0, w1, (uint32_t)(a0 & 3u) << 3) >> 8;
        if (pos + 4 + len > n) { err = 1; break; }
        if (cnt < max_records) { if (lane == 0) t_pos[cnt] = pos; }
        else if (!err) err = 2;                                              // keep counting: the host retries
        ++cnt;
        pos += 4 + (uint64_t)len;
    }
    if (n == 0) err = 1;                                                     // no stream identifier
    if (lane == 0) { res->n_all = cnt; res->n_chunks = 0; res->out_bytes = 0; res->error = err; }
}

__global__ void __launch_bounds__(256) frame_parse_kernel(const uint8_t* __restrict__ src, uint64_t n, FrameScratch f,
                                                          uint64_t max_records, FrameWalkResult* res) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t cnt = res->n_all < max_records ? res->n_all : max_records;
    if (i >= cnt || res->error == 1) return;
    const uint64_t pos = f.t_pos[i];
    const uintptr_t end = (uintptr_t)(src + n);
    const uintptr_t a0 = (uintptr_t)(src + pos);
    const uint32_t* wp = (const uint32_t*)(a0 & ~(uintptr_t)3);
    const uint32_t sh = (uint32_t)(a0 & 3u) << 3;
    uint32_t w[5];
#pragma unroll
    for (int k = 0; k < 5; ++k) w[k] = ((uintptr_t)(wp + k) < end) ? __ldg(wp + k) : 0u;
    const uint32_t d0 = shf_r(w[0], w[1], sh), d1 = shf_r(w[1], w[2], sh),
                   d2 = shf_r(w[2], w[3], sh), d3 = shf_r(w[3], w[4], sh);
    const uint32_t type = d0 & 0xffu, len = d0 >> 8;
    bool bad = (i == 0 && type != 0xff);                                     // :20-21
    uint32_t kind = kFrameNotData, plen = 0, ulen = 0;
    if (type == 0xff) {                                                      // :68-78  "sNaPpY"
        bad = bad || len != 6 || d1 != 0x50614e73u || (d2 & 0xffffu) != 0x5970u;
    } else if (type == 0x00 || type == 0x01) {
        if (len < 4) bad = true;
        else {
            kind = type; plen = len - 4; ulen = plen;
            if (type == 0x00) {                                              // varint32 preamble of the chunk's stream
                const unsigned long long vb = ((unsigned long long)d3 << 32) | d2;   // payload bytes 0..7
                ulen = 0;
                uint32_t shift = 0;
                for (uint32_t k = 0;; ++k) {
                    if (k >= plen || shift > 28) { bad = true; break; }
                    const uint32_t bt = (uint32_t)(vb >> (8 * k)) & 0xffu;
                    if (shift == 28 && (bt & 0x7fu) > 15u) { bad = true; break; }
                    ulen |= (bt & 0x7fu) << shift;
                    if (bt < 128) break;
                    shift += 7;
                }
            }
            if (ulen > kFrameChunk) bad = true;                              // :88-91, :104-106
        }
    } else if (type >= 0x02 && type <= 0x7f) {                               // :122-128 unskippable
        bad = true;
    }                                                                        // 0x80..0xfe: skip (:111-120, :131-138)
    f.t_kind[i] = kind; f.t_plen[i] = plen; f.t_ulen[i] = bad ? 0u : ulen; f.t_crc[i] = d1;
    if (bad) atomicExch(&res->pad, 1u);                                      // malformed (raised to error by the index pass)
}

__global__ void __launch_bounds__(1024) frame_index_kernel(FrameScratch f, uint64_t max_records, FrameWalkResult* res) {
    __shared__ unsigned long long warp_sum[32];
    __shared__ unsigned long long carry_s;
    const uint32_t tid = threadIdx.x, lane = tid & 31u, wid = tid >> 5;
    const uint64_t cnt = res->n_all < max_records ? res->n_all : max_records;
    if (tid == 0) carry_s = 0;                       // hi 24 bits: data chunks so far, lo 40 bits: output bytes so far
    __syncthreads();
    const bool malformed = res->error == 1 || res->pad != 0;
    if (!malformed) {
        for (uint64_t base = 0; base < cnt; base += 1024) {
            const uint64_t i = base + tid;
            const bool data = i < cnt && f.t_kind[i] != kFrameNotData;
            const unsigned long long v = data ? ((1ull << 40) | f.t_ulen[i]) : 0ull;
            unsigned long long x = v;
            for (int d = 1; d < 32; d <<= 1) {
                const unsigned long long y = __shfl_up_sync(kFull, x, d);
                if (lane >= (uint32_t)d) x += y;
            }
            if (lane == 31) warp_sum[wid] = x;
            __syncthreads();
            if (wid == 0) {
                unsigned long long sv = warp_sum[lane], t = sv;
                for (int d = 1; d < 32; d <<= 1) {
                    const unsigned long long y = __shfl_up_sync(kFull, t, d);
                    if (lane >= (uint32_t)d) t += y;
                }
                warp_sum[lane] = t - sv;
            }
            __syncthreads();
            const unsigned long long excl = carry_s + warp_sum[wid] + (x - v);
            if (data) {
                const uint64_t j = excl >> 40;
                f.ix.in_off[j] = f.t_pos[i] + 8; f.ix.out_off[j] = excl & ((1ull << 40) - 1);
                f.ix.in_len[j] = f.t_plen[i]; f.ix.ulen[j] = f.t_ulen[i]; f.ix.crc[j] = f.t_crc[i]; f.ix.kind[j] = f.t_kind[i];
            }
            __syncthreads();
            if (tid == 1023) carry_s = excl + v;
            __syncthreads();
        }
    }
    if (tid == 0) {
        res->n_chunks = carry_s >> 40;
        res->out_bytes = carry_s & ((1ull << 40) - 1);
        if (malformed) res->error = 1;
        res->pad = 0;
    }
}

// ---- reader, fused: the chain walk and the decode in ONE kernel ---------------------------------
// Only the hop is sequential, and a chunk can be decoded the moment its record exists, so the walk need not
// finish before the decode starts (round 2: 8.4 ms of walk in front of 13.5 ms of decode per GiB of text).
//   CTA 0, warp 0  hops (as frame_hop_kernel) and hands positions to warp 1 through a ring in shared memory
//   CTA 0, warp 1  parses 32 hopped headers at a time (as frame_parse_kernel), scans their output offsets
//                  (as frame_index_kernel), writes the records and publishes their number with a release store
//   every other warp of the grid takes a ticket, waits until that record is published (sleeping in proportion
//                  to its distance from the published count: no warp spins on the counter), decodes the chunk
// The hopper never waits for a decoder, and CTA 0 is resident from the start, so nothing can deadlock; the
// grid is persistent (16 CTAs per SM) and CTAs that start late just take later tickets.
struct FrameFusedCtl { unsigned int ticket, ready, finished, pad; };
constexpr uint32_t kHopRing = 1024;

__device__ __forceinline__ uint32_t ld_acquire_u32(const unsigned int* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u32(unsigned int* p, uint32_t v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__global__ void __launch_bounds__(128, SB_DECODE_MINB) frame_fused_kernel(const uint8_t* __restrict__ src, uint64_t n,
                                                                          uint8_t* dst, uint64_t dst_cap, FrameScratch f,
                                                                          uint64_t max_records, FrameWalkResult* res,
                                                                          FrameFusedCtl* ctl) {
    __shared__ unsigned long long ring[kHopRing];
    __shared__ volatile uint32_t s_hopped, s_parsed, s_hop_done, s_hop_err, s_abort;
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    const uintptr_t end = (uintptr_t)(src + n);
    if (blockIdx.x == 0) {
        if (threadIdx.x == 0) { s_hopped = 0; s_parsed = 0; s_hop_done = 0; s_hop_err = 0; s_abort = 0; }
        __syncthreads();
        if (wid == 0) {                                                      // ---- the hopper
            uint64_t pos = 0, pf = 0;
            uint32_t cnt = 0, err = n ? 0u : 1u;                             // n == 0: no stream identifier
            while (pos < n && !s_abort) {                                    // (the parser aborts the walk on a bad header)
                while (pf < n && pf < pos + (512u << 10)) {                  // stream -> L2, 32 KiB per TMA bulk prefetch
                    const uintptr_t b = ((uintptr_t)src + pf) & ~(uintptr_t)15;
                    const uintptr_t e = (end & ~(uintptr_t)15) < b + 32768u ? (end & ~(uintptr_t)15) : b + 32768u;
                    if (lane == 0 && e > b)
                        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(b), "r"((uint32_t)(e - b)) : "memory");
                    pf += 32768;
                }
                if (pos + 4 > n) { err = 1; break; }
                const uintptr_t a0 = (uintptr_t)(src + pos);
                const uint32_t* wp = (const uint32_t*)(a0 & ~(uintptr_t)3);
                const uint32_t w0 = __ldcg(wp), w1 = ((uintptr_t)(wp + 1) < end) ? __ldcg(wp + 1) : 0u;
                const uint32_t len = shf_r(w0, w1, (uint32_t)(a0 & 3u) << 3) >> 8;
                if (pos + 4 + len > n) { err = 1; break; }
                while (cnt - s_parsed >= kHopRing && !s_abort) __nanosleep(64);   // never laps the parser (it is 32 x faster)
                if (s_abort) break;
                if (lane == 0) { ring[cnt % kHopRing] = pos; __threadfence_block(); s_hopped = cnt + 1; }
                ++cnt;
                pos += 4 + (uint64_t)len;
            }
            if (lane == 0) { s_hop_err = err; __threadfence_block(); s_hop_done = 1; }
            return;
        }
        if (wid == 1) {                                                      // ---- the parser / indexer
            uint32_t done = 0, err = 0;
            unsigned long long carry = 0;                                    // hi 24 bits: data chunks, lo 40: output bytes
            for (;;) {
                // every lane loads, lane 0's values decide for the warp: the code below is warp-synchronous
                const uint32_t hd = __shfl_sync(kFull, (uint32_t)s_hop_done, 0);
                __threadfence_block();
                const uint32_t avail = __shfl_sync(kFull, (uint32_t)s_hopped, 0);
                if (avail == done) { if (hd) break; __nanosleep(64); continue; }
                const uint32_t m = avail - done < 32u ? avail - done : 32u;
                bool bad = false;
                uint32_t kind = kFrameNotData, plen = 0, ulen = 0, crc = 0;
                unsigned long long pos = 0;
                if (lane < m) {
                    pos = ring[(done + lane) % kHopRing];
                    const uintptr_t a0 = (uintptr_t)(src + pos);
                    const uint32_t* wp = (const uint32_t*)(a0 & ~(uintptr_t)3);
                    const uint32_t sh = (uint32_t)(a0 & 3u) << 3;
                    uint32_t w[5];
#pragma unroll
                    for (int k = 0; k < 5; ++k) w[k] = ((uintptr_t)(wp + k) < end) ? __ldcg(wp + k) : 0u;
                    const uint32_t d0 = shf_r(w[0], w[1], sh), d1 = shf_r(w[1], w[2], sh),
                                   d2 = shf_r(w[2], w[3], sh), d3 = shf_r(w[3], w[4], sh);
                    const uint32_t type = d0 & 0xffu, len = d0 >> 8;
                    crc = d1;
                    bad = (done + lane == 0 && type != 0xff);                // :20-21
                    if (type == 0xff) {                                      // :68-78  "sNaPpY"
                        bad = bad || len != 6 || d1 != 0x50614e73u || (d2 & 0xffffu) != 0x5970u;
                    } else if (type == 0x00 || type == 0x01) {
                        if (len < 4) bad = true;
                        else {
                            kind = type; plen = len - 4; ulen = plen;
                            if (type == 0x00) {                              // varint32 preamble of the chunk's stream
                                const unsigned long long vb = ((unsigned long long)d3 << 32) | d2;
                                ulen = 0;
                                uint32_t shift = 0;
                                for (uint32_t k = 0;; ++k) {
                                    if (k >= plen || shift > 28) { bad = true; break; }
                                    const uint32_t bt = (uint32_t)(vb >> (8 * k)) & 0xffu;
                                    if (shift == 28 && (bt & 0x7fu) > 15u) { bad = true; break; }
                                    ulen |= (bt & 0x7fu) << shift;
                                    if (bt < 128) break;
                                    shift += 7;
                                }
                            }
                            if (ulen > kFrameChunk) bad = true;              // :88-91, :104-106
                        }
                    } else if (type >= 0x02 && type <= 0x7f) {               // :122-128 unskippable
                        bad = true;
                    }
                    if (bad) { kind = kFrameNotData; ulen = 0; }
                }
                __syncwarp();
                if (lane == 0) s_parsed = done + m;                          // the ring slots are free again
                if (__any_sync(kFull, bad)) { err = 1; if (lane == 0) s_abort = 1; break; }
                const bool data = kind != kFrameNotData;
                const unsigned long long v = data ? ((1ull << 40) | ulen) : 0ull;
                unsigned long long x = v;
                for (int d = 1; d < 32; d <<= 1) {
                    const unsigned long long y = __shfl_up_sync(kFull, x, d);
                    if (lane >= (uint32_t)d) x += y;
                }
                const unsigned long long excl = carry + x - v;
                const uint64_t j = excl >> 40;
                if (data) {
                    if (j < max_records) {
                        f.ix.in_off[j] = pos + 8; f.ix.out_off[j] = excl & ((1ull << 40) - 1);
                        f.ix.in_len[j] = plen; f.ix.ulen[j] = ulen; f.ix.crc[j] = crc; f.ix.kind[j] = kind;
                    } else err = 2;                                          // index too small: keep counting, publish no more
                }
                err = __reduce_max_sync(kFull, err);
                carry += __shfl_sync(kFull, x, 31);
                done += m;
                __syncwarp();
                if (lane == 0) {
                    const uint64_t c = carry >> 40;
                    __threadfence();
                    st_release_u32(&ctl->ready, (uint32_t)(c < max_records ? c : max_records));
                }
            }
            if (lane == 0) {
                if (s_hop_err) err = 1;
                res->n_all = done; res->n_chunks = carry >> 40; res->out_bytes = carry & ((1ull << 40) - 1);
                res->error = err; res->pad = 0;
                __threadfence();
                st_release_u32(&ctl->finished, 1u);
            }
            return;
        }
    }
    // ---- decoders: every other warp of the grid.  (Keeping decoders off the hopper's SM was measured: no
    // change, 17.9 ms per GiB either way -- the read is bounded by two rounds of per-chunk decode latency,
    // 16,384 chunks on 9,472 warp slots, not by the hop.)
    for (;;) {
        uint32_t idx = 0;
        if (lane == 0) idx = atomicAdd(&ctl->ticket, 1u);
        idx = __shfl_sync(kFull, idx, 0);
        for (;;) {                                   // every lane loads, lane 0's value decides for the warp (the
            uint32_t r = __shfl_sync(kFull, ld_acquire_u32(&ctl->ready), 0);    // decode below is warp-synchronous)
            if (idx < r) break;
            if (__shfl_sync(kFull, ld_acquire_u32(&ctl->finished), 0)) {
                r = __shfl_sync(kFull, ld_acquire_u32(&ctl->ready), 0);
                if (idx < r) break;
                return;                                                      // no such chunk: done
            }
            const uint32_t gap = idx - r;                                    // records appear every ~0.4 us
            __nanosleep(gap < 4u ? 100u : gap < 3000u ? gap * 256u : 800000u);
        }
        const unsigned long long in_off = __ldcg(f.ix.in_off + idx), out_off = __ldcg(f.ix.out_off + idx);
        const uint32_t in_len = __ldcg(f.ix.in_len + idx), ulen = __ldcg(f.ix.ulen + idx), kind = __ldcg(f.ix.kind + idx);
        uint32_t got = kDecodeError;
        if (out_off + ulen <= dst_cap) {                                     // the host reports BUFFER_TOO_SMALL from out_bytes
            if (kind == 1) {
                if (in_len <= ulen) { decode_literal(dst + out_off, src + in_off, in_len, lane); got = in_len; }
            } else {
                got = decode_stream_global(src + in_off, in_len, dst + out_off, ulen, lane);
            }
        }
        if (lane == 0) f.got_len[idx] = got;
    }
}

// stored vs recomputed CRC and decode status of every chunk -> one flag
__global__ void frame_verify_kernel(FrameIndex ix, const uint32_t* got_len, const uint32_t* got_crc,
                                    uint32_t n, uint32_t* bad) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    if (got_len[j] != ix.ulen[j] || got_crc[j] != ix.crc[j]) atomicOr(bad, 1u);
}

}  // namespace sb
