// parse_lane.cuh -- exact greedy parse, ONE THREAD per fragment, every fragment of the call in
// flight at once ("lane parser").
//
// Why: Snappy's parse (tests/snappy/snappy.cc:539-671) is one loop-carried step per ~4.6 input
// bytes on text, and each step is a chain of dependent memory round trips (own bytes -> hash ->
// table -> candidate bytes -> compare).  A warp or a CTA per fragment does not shorten that chain,
// it only spends 32..160 threads' issue slots on it, and the 32 KB shared-memory table caps an SM at
// six fragments (888 per GPU).  Here the chain is cut to ONE round trip per step and the number of
// chains in flight is the number of fragments of the call (16,384 per GiB):
//
//   * the hash table lives in global memory and every entry is "fat": 16 bytes = position, epoch and
//     the 12 input bytes at that position.  The 4-byte hit test (snappy.cc:607-608, :656) and the
//     first 8 bytes of FindMatchLength (snappy-internal.h:95-143) are answered by the entry itself,
//     so no second, dependent load of the candidate's bytes is needed unless a match is longer than
//     12 bytes.  A 16-byte entry costs the same 32-byte DRAM sector as a 2-byte one.
//   * the table is never cleared: an entry is valid only if its epoch equals the call's epoch
//     (snappy.cc:478 zeroes the table per fragment; an invalid entry reads as "position 0", which
//     is what a zero entry means to the reference, SURVEY.md 8a item 1, with the fragment's first 12
//     bytes as its data).
//   * the thread keeps ip / skip / next_emit in registers and pushes one sequence record per copy;
//     it also adds up the encoded size of its fragment, so no size pass re-reads the records.
//
// Throughput is (fragments in flight) x (bytes per step) / (DRAM latency), bounded by the random
// 32-byte sector traffic of the tables (one read + one write-back per probe).  Streams with few
// fragments keep the pipelined parser (parse_pipe.cuh), whose per-fragment latency is lower.
//
// The per-fragment routine below is plain C++ over a memory policy, so that the same source is
// compiled for the host by tests/host_lane (g++) and checked against the reference Snappy byte for
// byte without a GPU; the product only ever runs the device instantiation.
#pragma once
#include <stdint.h>
#include <stddef.h>

#ifdef __CUDACC__
#define SB_HD __host__ __device__ __forceinline__
#else
#define SB_HD inline
#endif

namespace sb {
namespace lane {

typedef unsigned long long seq64;

constexpr uint32_t kLaneMul = 0x1e35a7bdu;     // tests/snappy/snappy.cc:92
constexpr uint32_t kLaneMargin = 15;           // snappy.cc:556 kInputMarginBytes
constexpr uint32_t kLaneMaxTable = 16384;      // snappy.h:199-200

struct W3 { uint32_t x, y, z; };               // 12 consecutive input bytes, little endian
struct Entry { uint32_t tag, g0, g1, g2; };    // tag = position | epoch << 16; g* = bytes at position

SB_HD uint32_t fsr(uint32_t lo, uint32_t hi, uint32_t sh) {   // (hi:lo) >> sh, sh in {0,8,16,24}
#ifdef __CUDA_ARCH__
    uint32_t r_;
    asm("shf.r.wrap.b32 %0, %1, %2, %3;" : "=r"(r_) : "r"(lo), "r"(hi), "r"(sh));   // not volatile: see snappy_device.cuh
    return r_;
#else
    return sh ? (lo >> sh) | (hi << (32u - sh)) : lo;
#endif
}
SB_HD uint32_t ctz64(unsigned long long x) {
#ifdef __CUDA_ARCH__
    return (uint32_t)(__ffsll((long long)x) - 1);
#else
    return (uint32_t)__builtin_ctzll(x);
#endif
}

// sizes of the elements EmitLiteral / EmitCopy produce (snappy.cc:337-431), for the running total
SB_HD uint32_t enc_literal(uint32_t len) {                    // len >= 1, <= 65536
    const uint32_t n = len - 1;
    return len + (n < 60 ? 1u : n < 256 ? 2u : 3u);
}
SB_HD uint32_t enc_copy(uint32_t offset, uint32_t len) {
    if (len < 12) return offset < 2048 ? 2u : 3u;
    const uint32_t k = (len >= 68) ? ((len - 4) >> 6) : 0u;   // 64-byte copies
    uint32_t rest = len - (k << 6), n = 3 * k;
    if (rest > 64) { n += 3; rest -= 60; }
    return n + ((rest < 12 && offset < 2048) ? 2u : 3u);
}
SB_HD seq64 pack_seq(uint32_t lit, uint32_t len, uint32_t off) {   // kernels.h seq_t layout
    return (seq64)lit | ((seq64)len << 17) | ((seq64)off << 33);
}

struct Out {
    seq64* seqs;
    uint32_t n, bytes;
};

// 12 bytes at src[p..p+12) from four aligned words (the address may have any alignment).  Touches
// only words that hold at least one byte of [p, p+13).
template <class M>
SB_HD W3 load12(const uint8_t* src, uint32_t p) {
    const uintptr_t a = (uintptr_t)(src + p);
    const uint32_t* w = (const uint32_t*)(a & ~(uintptr_t)3);
    const uint32_t sh = (uint32_t)(a & 3u) << 3;
    const uint32_t w0 = M::ld_in(w), w1 = M::ld_in(w + 1), w2 = M::ld_in(w + 2), w3 = M::ld_in(w + 3);
    W3 r;
    r.x = fsr(w0, w1, sh); r.y = fsr(w1, w2, sh); r.z = fsr(w2, w3, sh);
    return r;
}
template <class M>
SB_HD uint32_t load4(const uint8_t* src, uint32_t p) {
    const uintptr_t a = (uintptr_t)(src + p);
    const uint32_t* w = (const uint32_t*)(a & ~(uintptr_t)3);
    const uint32_t sh = (uint32_t)(a & 3u) << 3;
    const uint32_t w0 = M::ld_in(w);
    return sh ? fsr(w0, M::ld_in(w + 1), sh) : w0;
}

// Common prefix of src[a..] and src[b..] (a < b), b bounded by len: the part of FindMatchLength
// (snappy-internal.h:95-143) the fat entry could not answer.  Rare on text (matches > 12 bytes).
template <class M>
SB_HD uint32_t extend_slow(const uint8_t* src, uint32_t a, uint32_t b, uint32_t len) {
    uint32_t n = 0;
    while (b + n + 4 <= len) {
        const uint32_t x = load4<M>(src, a + n) ^ load4<M>(src, b + n);
        if (x) return n + (ctz64((unsigned long long)x) >> 3);
        n += 4;
    }
    while (b + n < len && M::ld_byte(src + a + n) == M::ld_byte(src + b + n)) ++n;
    return n;
}

template <class M>
SB_HD void push(Out& o, uint32_t lit, uint32_t len, uint32_t off) {
    M::st_seq(o.seqs + o.n, pack_seq(lit, len, off));
    ++o.n;
    o.bytes += (lit ? enc_literal(lit) : 0u) + (len ? enc_copy(off, len) : 0u);
}

// kVar selects how the table is reached (all variants produce the same records):
//   bit 0  prefetch.global.L2 of the next scan position's entry (fetches a 128-byte line)
//   bit 1  speculative LOAD of the next scan position's entry into registers (one 32-byte sector); a
//          missed probe then continues with no memory round trip at all.  The load is issued before
//          this probe's own store, so if both hash to the same slot the entry just stored is used.
//   bit 5  the probe is ONE 128-bit atomic exchange (old entry out, this position's entry in) instead of a
//          load followed by a store.  Measured (scripts/probes/randlat2.cu): a store issued right behind
//          the load of the same slot collapses to 0.8 G probes/s on the whole GPU whatever the
//          instruction flavour; the exchange sustains 15-19 G/s, a store one iteration late 17 G/s.
//   bits 2-4 belong to the memory policy (store ordering, entry stride, cache path).
template <class M, int kVar>
SB_HD void parse_fragment(const uint8_t* src, uint32_t len, void* table, uint32_t shift, uint32_t epoch,
                          Out& out) {
    constexpr bool kPfL2 = (kVar & 1) != 0, kSpec = (kVar & 2) != 0, kExch = (kVar & 32) != 0;
    uint32_t next_emit = 0;
    if (len >= kLaneMargin) {                                   // snappy.cc:557
        const uint32_t ip_limit = len - kLaneMargin;            // snappy.cc:558
        const uint32_t etag = epoch << 16;
        const W3 f0 = load12<M>(src, 0);                         // what an empty entry stands for
        uint32_t p = 1, skip = 32;                               // snappy.cc:560, :587
        bool chain = false;                                      // probe right after a copy (snappy.cc:649-656)
        bool have_spec = false, have_own = false;
        uint32_t h_prev = 0;
        W3 own = f0;
        Entry spec = {0, 0, 0, 0}, mine_prev = {0, 0, 0, 0};
        for (;;) {
            uint32_t next;
            if (!chain) {                                        // scan probe (snappy.cc:589-606)
                const uint32_t step = skip >> 5;
                skip += step;
                next = p + step;
                if (next > ip_limit) break;
            } else {                                             // a failed chain probe restarts the scan at ip+1
                next = p + 1;
                skip = 32;
            }
            if (!have_own) own = load12<M>(src, p);
            const uint32_t h = (own.x * kLaneMul) >> shift;
            void* slot = M::slot(table, h);
            Entry mine;
            mine.tag = p | etag; mine.g0 = own.x; mine.g1 = own.y; mine.g2 = own.z;
            // The bytes of the next scan position are requested together with the table access (next <=
            // ip_limit, so the 12 bytes exist): a missed probe continues without a round trip for its input,
            // which matters once the skip is large and every probe lands on a cold line (literal-heavy
            // input: 0.42 -> 0.34 ms for 256 MiB of random bytes).
            const W3 own_next = load12<M>(src, next);
            Entry e;                                             // snappy.cc:602-606 / :652-655: look up, then insert
            if (kSpec && have_spec) {
                e = (h == h_prev) ? mine_prev : spec;
                M::st_tab(slot, mine, 0u);
            } else if (kExch) {
                e = M::xchg_tab(slot, mine);
            } else {
                e = M::ld_tab(slot);
                M::st_tab(slot, mine, e.tag);
            }
            M::prefetch_in(src, p + M::kAhead, len);
            if (kPfL2 || kSpec) {                                // the next probe if this one misses
                const uint32_t g1 = own_next.x;
                void* slot1 = M::slot(table, (g1 * kLaneMul) >> shift);
                if (kSpec) spec = M::ld_tab(slot1);
                else M::prefetch_tab(slot1);
            }
            uint32_t cand = e.tag & 0xffffu;
            if ((e.tag >> 16) != epoch) { cand = 0; e.g0 = f0.x; e.g1 = f0.y; e.g2 = f0.z; }
            if (own.x != e.g0) {                                 // snappy.cc:607-608 / :656
                p = next;
                chain = false;
                own = own_next;
                have_own = true;
                if (kSpec) { have_spec = true; h_prev = h; mine_prev = mine; }
                continue;
            }
            // a copy (snappy.cc:627-641)
            const unsigned long long x = ((unsigned long long)(own.z ^ e.g2) << 32) | (own.y ^ e.g1);
            uint32_t matched;
            if (x) matched = 4 + (ctz64(x) >> 3);
            else matched = 12 + extend_slow<M>(src, cand + 12, p + 12, len);
            push<M>(out, p - next_emit, matched, p - cand);
            p += matched;
            next_emit = p;
            if (p >= ip_limit) break;                            // snappy.cc:643-645
            const W3 prev = load12<M>(src, p - 1);               // snappy.cc:649-651: insert ip-1
            Entry ins;
            ins.tag = (p - 1) | etag; ins.g0 = prev.x; ins.g1 = prev.y; ins.g2 = prev.z;
            M::st_tab(M::slot(table, (prev.x * kLaneMul) >> shift), ins, 0u);
            chain = true;
            have_spec = false;
            have_own = false;
        }
    }
    if (next_emit < len) push<M>(out, len - next_emit, 0, 0);    // snappy.cc:665-668
}

// table entries of a fragment of `len` bytes (snappy.cc:445-456) and the matching hash shift
SB_HD uint32_t entries_for(uint32_t len) {
    if (len > kLaneMaxTable) return kLaneMaxTable;
    uint32_t e = 256;
    while (e < len) e <<= 1;
    return e;
}
SB_HD uint32_t shift_for(uint32_t entries) {
    uint32_t lg = 0;
    while ((1u << lg) < entries) ++lg;
    return 32u - lg;
}

}  // namespace lane
}  // namespace sb

#ifdef __CUDACC__
#include "kernels.h"
#include "emit.cuh"

namespace sb {

// Device memory policy.  kVar bits (continued from parse_fragment):
//   bit 2  the store waits for the load of the same slot (register dependency) instead of chasing it
//   bit 3  entries padded to 32 bytes and written with ONE 256-bit store (a full DRAM sector, so the
//          L2 never has to merge a partial write with a fill)
//   bit 4  table loads through L1 (ld.global.ca) instead of L2-only
template <int kVar>
struct LaneDev {
    static constexpr uint32_t kStride = (kVar & 8) ? 32u : 16u;
    static constexpr uint32_t kAhead = 192u;                         // input prefetch distance (bytes)
    static __device__ __forceinline__ uint32_t ld_in(const uint32_t* p) { return __ldg(p); }
    static __device__ __forceinline__ uint32_t ld_byte(const uint8_t* p) { return __ldg(p); }
    static __device__ __forceinline__ void* slot(void* table, uint32_t h) {
        return reinterpret_cast<uint8_t*>(table) + (size_t)h * kStride;
    }
    static __device__ __forceinline__ lane::Entry ld_tab(const void* p) {
        uint4 v;
        if (kVar & 16) asm volatile("ld.global.ca.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
        else v = __ldcg(reinterpret_cast<const uint4*>(p));     // L2 only: a random sector, no reuse in L1
        lane::Entry e; e.tag = v.x; e.g0 = v.y; e.g1 = v.z; e.g2 = v.w;
        return e;
    }
    static __device__ __forceinline__ void st_tab(void* p, const lane::Entry& e, uint32_t dep) {
        uint32_t tag = e.tag;
        if (kVar & 4) asm volatile("" : "+r"(tag) : "r"(dep));
        if (kVar & 8) {
            asm volatile("st.global.cg.v8.b32 [%0], {%1,%2,%3,%4,%5,%5,%5,%5};" ::"l"(p), "r"(tag), "r"(e.g0),
                         "r"(e.g1), "r"(e.g2), "r"(0u)
                         : "memory");
        } else {
            __stcg(reinterpret_cast<uint4*>(p), make_uint4(tag, e.g0, e.g1, e.g2));
        }
    }
    static __device__ __forceinline__ lane::Entry xchg_tab(void* p, const lane::Entry& n) {
        lane::Entry o;
        asm volatile("{\n .reg .b128 o, n;\n mov.b128 n, {%4,%5,%6,%7};\n"
                     " atom.global.relaxed.gpu.exch.b128 o, [%8], n;\n mov.b128 {%0,%1,%2,%3}, o;\n}"
                     : "=r"(o.tag), "=r"(o.g0), "=r"(o.g1), "=r"(o.g2)
                     : "r"(n.tag), "r"(n.g0), "r"(n.g1), "r"(n.g2), "l"(p)
                     : "memory");
        return o;
    }
    static __device__ __forceinline__ void st_seq(lane::seq64* p, lane::seq64 v) { __stcg(p, v); }
    static __device__ __forceinline__ void prefetch_tab(const void* p) {
        asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
    }
    static __device__ __forceinline__ void prefetch_in(const uint8_t* src, uint32_t pos, uint32_t len) {
        if (pos < len) asm volatile("prefetch.global.L1 [%0];" ::"l"(src + pos));
    }
};

struct LaneTable {
    uint8_t* base;             // job j uses base + j * stride * (bytes per entry)
    uint32_t stride;           // entries per job = table_entries(job_len)
    uint32_t epoch;            // 1..65535, never repeated between two clears of the pool
    uint32_t fpw;              // fragments (active lanes) per warp
};

constexpr int kLaneMaxThreads = 128;   // up to four warps per CTA (32 CTAs per SM would otherwise cap the warps)

template <int kVar>
__global__ void __launch_bounds__(kLaneMaxThreads) parse_lane_kernel(JobParams p, LaneTable t) {
    const uint32_t lane_id = threadIdx.x & 31u;
    if (lane_id >= t.fpw) return;
    const uint32_t warp = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const uint32_t job = warp * t.fpw + lane_id;
    if (job >= p.n_jobs) return;
    const uint32_t len = job_len_of(p, job);
    const uint32_t entries = lane::entries_for(len);             // from THIS fragment's length
    lane::Out out;
    out.seqs = p.seqs + (uint64_t)job * seq_stride(p.job_len);
    out.n = 0;
    out.bytes = p.with_header ? varint32_len(len) : 0u;
    lane::parse_fragment<LaneDev<kVar>, kVar>(p.in + (uint64_t)job * p.in_stride, len,
                                              t.base + (uint64_t)job * t.stride * LaneDev<kVar>::kStride,
                                              32u - (31u - __clz(entries)), t.epoch, out);
    p.nseq[job] = out.n;
    p.sizes[job] = out.bytes;
}

}  // namespace sb
#endif  // __CUDACC__
