// parse_tiny.cuh -- small buffers of a batch (the reference's own benchmark shape: 10 .. 2,000-byte
// buffers, benchmark/hw_results.csv, DataGenerator.scala:27; this kernel takes those up to 256 bytes): ONE THREAD compresses one buffer from
// start to finish -- parse and emission fused, no sequence list, no size / emit pass.
//
// Why: a warp or a CTA per 100-byte buffer is launch-bound (2.6 ns per buffer whatever its size: 107 M
// CTAs for 1 GiB of 10-byte buffers).  A buffer of up to
// 512 bytes needs a table of at most 512 uint16 entries (tests/snappy/snappy.cc:445-456), so 200-256 of
// them fit an SM's shared memory at once and each is a short serial chain; the lanes of a warp run
// different buffers in lock step.
//
// The table of thread t is interleaved with its neighbours' (entry h at base[h * T + t]): clearing it is a
// coalesced sweep by the whole CTA, and look-ups of different threads fall into different banks whatever
// their hash.  Input is read in place through L1 (the CTA's buffers are contiguous), elements are written
// at their final position.
//
// The per-buffer routine is plain C++ over a memory policy: tests/host_lane/tiny_host.cc compiles the same
// source for the host and tests/test_tiny_host.py pins it against the reference Snappy without a GPU.
#pragma once
#include <stdint.h>
#include <stddef.h>

#ifdef __CUDACC__
#define SB_THD __host__ __device__ __forceinline__
#else
#define SB_THD inline
#endif

namespace sb {
namespace tiny {

constexpr uint32_t kMul = 0x1e35a7bdu;      // tests/snappy/snappy.cc:92
constexpr uint32_t kMargin = 15;            // snappy.cc:556 kInputMarginBytes
constexpr uint32_t kMaxLen = 256;           // buffers this kernel takes (SNAPPY_B200_TINY_MAX overrides, up to 512).
                                            // Measured against the warp-per-buffer parser, which also writes its
                                            // elements itself (GB/s on 512 MiB of text / random / constant buffers):
                                            // 100 B 72/91/431 vs 54/73/110, 200 B 62/94/430 vs 64/123/212, 300 B
                                            // 39/76/221 vs 72/186/285, 512 B 34/73/301 vs 78/292/561 -- beyond ~250
                                            // bytes a warp's 32-position windows beat one serial thread.
constexpr uint32_t kMaxLenLimit = 512;

SB_THD uint32_t entries_for(uint32_t len) {                     // snappy.cc:445-456
    uint32_t e = 256;
    while (e < len) e <<= 1;
    return e > 16384u ? 16384u : e;
}
SB_THD uint32_t fsr(uint32_t lo, uint32_t hi, uint32_t sh) {
#ifdef __CUDA_ARCH__
    uint32_t r_;
    asm("shf.r.wrap.b32 %0, %1, %2, %3;" : "=r"(r_) : "r"(lo), "r"(hi), "r"(sh));   // not volatile: see snappy_device.cuh
    return r_;
#else
    return sh ? (lo >> sh) | (hi << (32u - sh)) : lo;
#endif
}
SB_THD uint32_t ctz32(uint32_t x) {
#ifdef __CUDA_ARCH__
    return (uint32_t)(__ffs((int)x) - 1);
#else
    return (uint32_t)__builtin_ctz(x);
#endif
}

// 4 bytes at src[p..p+4) from aligned words; touches only words that hold one of those bytes
template <class M>
SB_THD uint32_t load4(const uint8_t* src, uint32_t p) {
    const uintptr_t a = (uintptr_t)(src + p);
    const uint32_t* w = (const uint32_t*)(a & ~(uintptr_t)3);
    const uint32_t sh = (uint32_t)(a & 3u) << 3;
    const uint32_t w0 = M::ld32(w);
    return sh ? fsr(w0, M::ld32(w + 1), sh) : w0;
}

template <class M>
SB_THD uint32_t put_literal(uint8_t* out, uint32_t op, const uint8_t* src, uint32_t from, uint32_t n) {   // snappy.cc:337-378
    const uint32_t m = n - 1;
    if (m < 60) { M::st8(out + op, (uint8_t)(m << 2)); op += 1; }
    else if (m < 256) { M::st8(out + op, 60u << 2); M::st8(out + op + 1, (uint8_t)m); op += 2; }
    else { M::st8(out + op, 61u << 2); M::st8(out + op + 1, (uint8_t)m); M::st8(out + op + 2, (uint8_t)(m >> 8)); op += 3; }
    for (uint32_t i = 0; i < n; ++i) M::st8(out + op + i, M::ld8(src + from + i));
    return op + n;
}

template <class M>
SB_THD uint32_t put_copy64(uint8_t* out, uint32_t op, uint32_t off, uint32_t len, bool lt12) {            // snappy.cc:380-400
    if (lt12 && off < 2048) {
        M::st8(out + op, (uint8_t)(1u | ((len - 4) << 2) | ((off >> 8) << 5)));
        M::st8(out + op + 1, (uint8_t)off);
        return op + 2;
    }
    M::st8(out + op, (uint8_t)(2u | ((len - 1) << 2)));
    M::st8(out + op + 1, (uint8_t)off);
    M::st8(out + op + 2, (uint8_t)(off >> 8));
    return op + 3;
}

template <class M>
SB_THD uint32_t put_copy(uint8_t* out, uint32_t op, uint32_t off, uint32_t len) {                          // snappy.cc:402-431
    if (len < 12) return put_copy64<M>(out, op, off, len, true);
    while (len >= 68) { op = put_copy64<M>(out, op, off, 64, false); len -= 64; }
    if (len > 64) { op = put_copy64<M>(out, op, off, 60, false); len -= 60; }
    return put_copy64<M>(out, op, off, len, len < 12);
}

// One whole stream: varint32(len) (when with_header) + the single fragment's elements (len <= 65536;
// the kernel is used for len <= kMaxLen).  `tab` is the thread's table accessor, already all zero.
template <class M, class Tab>
SB_THD uint32_t compress_one(const uint8_t* src, uint32_t len, Tab tab, uint32_t entries, uint8_t* out, bool with_header) {
    uint32_t op = 0;
    if (with_header) {                                                      // snappy.cc:1006-1009
        uint32_t v = len;
        while (v >= 128) { M::st8(out + op++, (uint8_t)(v | 128)); v >>= 7; }
        M::st8(out + op++, (uint8_t)v);
    }
    uint32_t next_emit = 0;
    if (len >= kMargin) {                                                   // snappy.cc:557
        uint32_t lg = 8;
        while ((1u << lg) < entries) ++lg;
        const uint32_t shift = 32u - lg, ip_limit = len - kMargin;
        uint32_t ip = 1, skip = 32;
        for (;;) {                                                          // one scan + the copies behind it
            uint32_t cand;
            for (;;) {                                                      // snappy.cc:589-608
                const uint32_t step = skip >> 5;
                skip += step;
                const uint32_t next = ip + step;
                if (next > ip_limit) goto remainder;
                const uint32_t cur = load4<M>(src, ip);
                const uint32_t h = (cur * kMul) >> shift;
                cand = tab.ld(h);
                tab.st(h, (uint16_t)ip);
                if (cur == load4<M>(src, cand)) break;
                ip = next;
            }
            op = put_literal<M>(out, op, src, next_emit, ip - next_emit);   // snappy.cc:614 (ip > next_emit always)
            for (;;) {                                                      // snappy.cc:627-656
                uint32_t matched = 4;
                while (ip + matched + 4 <= len) {                           // FindMatchLength, 4 bytes a round
                    const uint32_t x = load4<M>(src, cand + matched) ^ load4<M>(src, ip + matched);
                    if (x) { matched += ctz32(x) >> 3; goto extended; }
                    matched += 4;
                }
                while (ip + matched < len && M::ld8(src + cand + matched) == M::ld8(src + ip + matched)) ++matched;
            extended:
                op = put_copy<M>(out, op, ip - cand, matched);
                ip += matched;
                next_emit = ip;
                if (ip >= ip_limit) goto remainder;
                tab.st((load4<M>(src, ip - 1) * kMul) >> shift, (uint16_t)(ip - 1));
                const uint32_t cur = load4<M>(src, ip);
                const uint32_t h = (cur * kMul) >> shift;
                cand = tab.ld(h);
                tab.st(h, (uint16_t)ip);
                if (cur != load4<M>(src, cand)) break;
            }
            ++ip;                                                           // snappy.cc:658-659
            skip = 32;
        }
    }
remainder:
    if (next_emit < len) op = put_literal<M>(out, op, src, next_emit, len - next_emit);   // snappy.cc:665-668
    return op;
}

}  // namespace tiny
}  // namespace sb

#ifdef __CUDACC__
#include "kernels.h"
#include "emit.cuh"

namespace sb {

struct TinyDev {
    static __device__ __forceinline__ uint32_t ld32(const uint32_t* p) { return __ldg(p); }
    static __device__ __forceinline__ uint32_t ld8(const uint8_t* p) { return __ldg(p); }
    static __device__ __forceinline__ void st8(uint8_t* p, uint8_t v) { *p = v; }
};
struct TinyTab {                           // entry h of this thread: base[h * stride]
    uint16_t* base;
    uint32_t stride;
    __device__ __forceinline__ uint32_t ld(uint32_t h) const { return base[h * stride]; }
    __device__ __forceinline__ void st(uint32_t h, uint16_t v) const { base[h * stride] = v; }
};

constexpr uint32_t kTinySmemBytes = 200u << 10;
__host__ __device__ inline uint32_t tiny_threads(uint32_t job_len) {      // buffers (= threads) per CTA
    const uint32_t per = tiny::entries_for(job_len) * 2u;
    uint32_t t = kTinySmemBytes / per;
    if (t > 256u) t = 256u;
    return t & ~31u ? t & ~31u : t;        // whole warps when there is room for one
}

__global__ void __launch_bounds__(256) tiny_kernel(JobParams p) {
    extern __shared__ __align__(16) uint8_t smem[];
    const uint32_t T = blockDim.x, tid = threadIdx.x;
    const uint32_t entries = tiny::entries_for(p.job_len);                  // the launch's buffer length
    {
        uint4* z = reinterpret_cast<uint4*>(smem);
        const uint32_t n16 = T * entries * 2u / 16u;
        for (uint32_t i = tid; i < n16; i += T) z[i] = make_uint4(0, 0, 0, 0);
    }
    __syncthreads();
    const uint64_t job = (uint64_t)blockIdx.x * T + tid;
    if (job >= p.n_jobs) return;
    const uint32_t len = job_len_of(p, (uint32_t)job);
    TinyTab tab{reinterpret_cast<uint16_t*>(smem) + tid, T};
    // a shorter last buffer uses a smaller table (snappy.cc:445-456): its entries are a prefix of the zeroed ones
    const uint32_t n = tiny::compress_one<TinyDev>(p.in + job * p.in_stride, len, tab, tiny::entries_for(len),
                                                   p.out + job * p.out_stride, p.with_header != 0);
    p.sizes[job] = n;
}

}  // namespace sb
#endif  // __CUDACC__
