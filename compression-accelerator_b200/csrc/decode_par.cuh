// decode_par.cuh -- one LARGE raw Snappy stream decoded in parallel.
//
// A raw stream carries no index, and element boundaries exist only along the walk, so decode.cuh
// gives a stream to one warp: fine for the page batch, hopeless for a 1 GiB stream (~8 MB/s).
// But compressors that work like the reference (tests/snappy/snappy.cc:1018-1061: independent
// fragments of 65,536 input bytes) produce streams in which (a) an element starts exactly at every
// multiple of 65,536 output bytes and (b) no copy reaches back across such a boundary.  For those
// streams the blocks are independent once their compressed offsets are known.  Finding the offsets:
//
//   1. raw_spec_walk_kernel   the stream is cut into segments of S compressed bytes; one thread per
//                             segment walks the TAGS only, starting at the segment's first byte as
//                             if it were an element boundary.  Any byte is a valid tag, so the walk
//                             is always defined, and it falls in step with the true parse after a
//                             few elements.  It records the output bytes produced before every
//                             boundary within its first kSyncW bytes, its total, and where it leaves.
//   2. raw_exits / raw_sync / raw_chain   the true entry of segment k+1 is where segment k's walk
//                             leaves (true and speculative walks leave together once they have
//                             met); every segment follows the chain from its entry until it meets
//                             a speculative walk (its own after a few elements, or a later
//                             segment's), and one short pass hops from run to run.
//   3. raw_blocks_kernel      one thread per run walks again from its true entry with the true
//                             output offset and notes the compressed offset of every 64 KiB block;
//                             an element that straddles a block boundary clears the fast path.
//   4. decode_kernel          (decode.cuh, indexed + headerless) one warp per block; a copy reaching
//                             before its block or any other defect shows up as a wrong block length.
//   5. the serial walk of decode.cuh runs ONLY IF a flag says the stream is not block structured or
//      is malformed, so every valid stream still decodes and every malformed one is still reported
//      by the hardened walk.  No host synchronisation anywhere: the flag lives on the device.
#pragma once
#include "decode.cuh"

namespace sb {

constexpr uint32_t kSyncW = 256;                 // boundaries recorded at the head of each segment
constexpr uint32_t kNoBoundary = 0xffffffffu;
constexpr uint32_t kParBlock = 65536;
constexpr uint32_t kWalkWarps = 2;
#ifndef SB_SPEC_MAXLIT
#define SB_SPEC_MAXLIT 4096
#endif
constexpr uint32_t kSpecMaxLiteral = SB_SPEC_MAXLIT;   // longest literal a speculative walk believes
constexpr uint64_t kMinSegment = 2 * kSpecMaxLiteral;

struct RawIndex {
    // per segment
    unsigned long long* exit_pos;                // speculative walk: first boundary >= segment end
    unsigned long long* spec_out;                // output bytes it produced
    unsigned long long* distrust;                // 1 + last position where it refused an element (0: never)
    uint32_t* cum;                               // [n_seg][kSyncW] output bytes before the boundary at r, or kNoBoundary
    unsigned long long* entry_pos;               // first boundary >= segment start, assuming every earlier walk met the true chain
    unsigned long long* entry_out;               // output offset there (run leaders only)
    unsigned long long* run_exit;                // where the true chain from entry_pos leaves the run's last segment
    unsigned long long* run_delta;               // output bytes from entry_pos to run_exit
    uint32_t* run_end;                           // last segment of the run (the one whose speculative walk is met)
    uint32_t* lead;                              // 1: this segment starts a run of the true chain
    uint32_t* mlist;                             // segments whose run spans more than themselves, in order
    // per block
    unsigned long long* blk_in;                  // compressed offset of the block's first element (~0: not found)
    unsigned long long* blk_out;                 // 65536 * m
    uint32_t* blk_len;                           // compressed bytes
    uint32_t* blk_cap;                           // decoded bytes expected
    uint32_t* blk_got;                           // decoded bytes produced
    // scalars
    unsigned long long* ulen;                    // from the preamble
    uint32_t* serial;                            // != 0: take the serial walk
};

__host__ __device__ inline size_t raw_index_bytes(uint64_t n_seg, uint64_t n_blk) {
    return n_seg * (7 * 8 + 3 * 4 + (size_t)kSyncW * 4) + n_blk * (2 * 8 + 3 * 4) + 64;
}
__host__ __device__ inline RawIndex raw_index_at(void* base, uint64_t n_seg, uint64_t n_blk) {
    RawIndex x;
    x.ulen = (unsigned long long*)base;
    x.serial = (uint32_t*)((uint8_t*)base + 8);
    x.exit_pos = (unsigned long long*)((uint8_t*)base + 64);
    x.spec_out = x.exit_pos + n_seg;
    x.distrust = x.spec_out + n_seg;
    x.entry_pos = x.distrust + n_seg;
    x.entry_out = x.entry_pos + n_seg;
    x.run_exit = x.entry_out + n_seg;
    x.run_delta = x.run_exit + n_seg;
    x.blk_in = x.run_delta + n_seg;
    x.blk_out = x.blk_in + n_blk;
    x.cum = (uint32_t*)(x.blk_out + n_blk);
    x.run_end = x.cum + n_seg * kSyncW;
    x.lead = x.run_end + n_seg;
    x.mlist = x.lead + n_seg;
    x.blk_len = x.mlist + n_seg;
    x.blk_cap = x.blk_len + n_blk;
    x.blk_got = x.blk_cap + n_blk;
    return x;
}

// Size of the element at pos: compressed bytes and output bytes.  false: its header needs bytes
// beyond n (the walk stops; on the true chain that is a malformed stream, which the decode reports).
__device__ __forceinline__ bool raw_element(const uint8_t* __restrict__ s, uint64_t pos, uint64_t n,
                                            uint64_t& esz, uint64_t& olen) {
    const uint32_t tag = __ldg(s + pos), kind = tag & 3u;
    if (kind == 0) {
        uint64_t len = (tag >> 2) + 1;
        uint32_t extra = 0;
        if (len > 60) {
            extra = (uint32_t)len - 60;
            if (pos + 1 + extra > n) return false;
            uint64_t v = 0;
            for (uint32_t i = 0; i < extra; ++i) v |= (uint64_t)__ldg(s + pos + 1 + i) << (8 * i);
            len = v + 1;
        }
        esz = 1 + extra + len; olen = len;
    } else if (kind == 1) { esz = 2; olen = 4 + ((tag >> 2) & 7u); }
    else if (kind == 2) { esz = 3; olen = (tag >> 2) + 1; }
    else { esz = 5; olen = (tag >> 2) + 1; }
    return true;
}

// The same, without branches: 32 walkers in one warp stay in lockstep (a divergent 4-way tag switch
// would serialise them).  Reads the two aligned words that hold the tag and up to four length bytes.
__device__ __forceinline__ bool raw_element_lockstep(const uint8_t* __restrict__ s, uint64_t pos, uint64_t n,
                                                     uint64_t& esz, uint64_t& olen) {
    const uintptr_t a = (uintptr_t)(s + pos);
    const uint32_t* w = reinterpret_cast<const uint32_t*>(a & ~(uintptr_t)3);
    const uint32_t a0 = __ldg(w);
    const uint32_t a1 = (reinterpret_cast<const uint8_t*>(w + 1) < s + n) ? __ldg(w + 1) : 0u;
    const uint64_t q = (((uint64_t)a1 << 32) | a0) >> (((uint32_t)a & 3u) << 3);     // >= 5 bytes from pos
    const uint32_t tag = (uint32_t)q & 0xffu, kind = tag & 3u, len6 = tag >> 2;
    const uint32_t extra = max(len6, 59u) - 59u;                                     // 0, or 1..4 length bytes
    const uint32_t v = (uint32_t)(q >> 8) & (extra ? 0xffffffffu >> (32u - 8u * extra) : 0u);
    const uint64_t lit = extra ? (uint64_t)v + 1 : (uint64_t)len6 + 1;
    const uint32_t copy_sz = (0x05030200u >> (8u * kind)) & 0xffu;
    const uint32_t copy_out = kind == 1 ? 4u + (len6 & 7u) : len6 + 1u;
    esz = kind ? (uint64_t)copy_sz : 1 + extra + lit;
    olen = kind ? (uint64_t)copy_out : lit;
    return kind != 0 || pos + 1 + extra <= n;
}

__device__ __forceinline__ uint32_t raw_preamble(const uint8_t* __restrict__ s, uint64_t n, uint64_t& ulen) {
    uint32_t ip = 0, shift = 0;
    ulen = 0;
    for (;;) {
        if (ip >= n || shift > 28) return 0;
        const uint32_t b = __ldg(s + ip++);
        if (shift == 28 && (b & 0x7fu) > 15u) return 0;
        ulen |= (uint64_t)(b & 0x7fu) << shift;
        if (b < 128) return ip;
        shift += 7;
    }
}

// 1. speculative tag walk, one thread per segment, 32 walkers per warp in lockstep
__global__ void __launch_bounds__(64) raw_spec_walk_kernel(const uint8_t* __restrict__ s, uint64_t n, uint64_t seg,
                                                           uint32_t n_seg, RawIndex x) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_seg) return;
    uint32_t* cum = x.cum + (size_t)k * kSyncW;
    for (uint32_t i = 0; i < kSyncW; ++i) cum[i] = kNoBoundary;
    const uint64_t s0 = (uint64_t)k * seg, s1 = s0 + seg < n ? s0 + seg : n;
    uint64_t pos = s0, out = 0, refused = 0;
    if (k == 0) {
        uint64_t ulen;
        pos = raw_preamble(s, n, ulen);
        if (pos == 0) { x.exit_pos[0] = n; x.spec_out[0] = 0; x.distrust[0] = 0; return; }   // malformed: 2a notices
    }
    while (pos < s1) {
        if (pos - s0 < kSyncW) cum[pos - s0] = out < kNoBoundary ? (uint32_t)out : kNoBoundary;
        uint64_t esz, olen;
        if (!raw_element_lockstep(s, pos, n, esz, olen)) { pos = n; break; }
        // A long literal is far more often the sign of a walk that is not in step yet (a stray
        // 0xF4/0xF8/0xFC byte read as a tag would send it kilobytes to megabytes away) than a real
        // one.  Speculation only: step one byte and keep looking; a real long literal makes the
        // neighbouring segments part of one run of the true chain, which costs nothing (few, huge
        // elements).  With segments of at least 2 * kSpecMaxLiteral no element a speculative walk
        // accepts can jump over a whole segment.
        const bool refuse = k != 0 && olen > kSpecMaxLiteral;
        refused = refuse ? pos + 1 : refused;
        pos += refuse ? 1 : esz;
        out += refuse ? 0 : olen;
    }
    x.exit_pos[k] = pos;
    x.spec_out[k] = out;
    x.distrust[k] = refused;
}

// 2. chain the segments.
// 2a (one CTA): entry of segment k = where the speculative walk of segment k-1 leaves, except that an
//     element spanning a whole segment carries its end through.  These entries are TRUE as long as
//     every earlier walk has met the true chain.
// 2b (one thread per segment): follow the chain from that entry until it meets a speculative walk:
//     usually the segment's own, after a few elements; if not within kSyncW bytes, keep walking --
//     through following segments if need be -- until some segment's walk is met within its window (a
//     "run"; an incompressible stream is one run of few, huge elements).  Bounded: kRunMaxElems.
// 2c (one CTA): from segment 0, jump from run to run; a run's exit is a speculative exit, hence the
//     next run's 2a entry, so all leaders' entries are true.  Output offsets accumulate on the way.
constexpr uint32_t kRunMaxElems = 1u << 15;      // two full text-like segments of the smallest size
constexpr uint32_t kRunOpen = 0xffffffffu;

__global__ void __launch_bounds__(1024) raw_exits_kernel(const uint8_t* __restrict__ s, uint64_t n, uint64_t seg,
                                                         uint32_t n_seg, uint64_t out_cap, RawIndex x) {
    __shared__ unsigned long long tile[1024];
    __shared__ unsigned long long carry;
    const uint32_t t = threadIdx.x;
    if (t == 0) {
        uint64_t ulen;
        const uint32_t hdr = raw_preamble(s, n, ulen);
        *x.ulen = ulen;
        if (hdr == 0 || ulen > out_cap) *x.serial = 1;                       // the serial walk reports it
        carry = hdr;
    }
    __syncthreads();
    for (uint32_t base = 0; base < n_seg; base += 1024) {
        const uint32_t k = base + t;
        if (k < n_seg) tile[t] = x.exit_pos[k];
        __syncthreads();
        if (t == 0) {
            uint64_t e = carry;
            const uint32_t cnt = n_seg - base < 1024 ? n_seg - base : 1024;
            for (uint32_t i = 0; i < cnt; ++i) {
                const uint64_t s0 = (uint64_t)(base + i) * seg, s1 = s0 + seg < n ? s0 + seg : n;
                const uint64_t entry = e;
                if (e < s1) e = tile[i];                                     // else: an element spans the segment
                tile[i] = entry;
            }
            carry = e;
        }
        __syncthreads();
        if (k < n_seg) x.entry_pos[k] = tile[t];
        __syncthreads();
    }
}

__global__ void __launch_bounds__(32 * kWalkWarps) raw_sync_kernel(const uint8_t* __restrict__ s, uint64_t n, uint64_t seg,
                                                      uint32_t n_seg, RawIndex x) {
    // one walker per warp (32 divergent walks in one warp cost ~8x), kWalkWarps warps per CTA (32 CTAs per SM at most)
    const uint32_t k = blockIdx.x * kWalkWarps + (threadIdx.x >> 5);
    if ((threadIdx.x & 31) || k >= n_seg) return;
    uint64_t pos = x.entry_pos[k], oo = 0;
    uint32_t j = k, end = kRunOpen;
    uint64_t s0 = (uint64_t)j * seg, s1 = s0 + seg < n ? s0 + seg : n;
    unsigned long long delta = 0, leave = 0;
    if (pos < s1) {
        const uint64_t from = pos;
        for (uint32_t steps = 0;; ++steps) {
            // bounded: a walk from a wrong entry (they exist wherever an earlier speculation never met
            // the true chain) hops through small pseudo-elements for ever; a real run longer than this
            // is an incompressible region, whose elements are huge
            if (steps >= kRunMaxElems && pos - from < (uint64_t)steps * 64) break;
            const uint64_t rel = pos - s0;
            if (rel < kSyncW) {
                const uint32_t c = x.cum[(size_t)j * kSyncW + rel];
                // met -- and the two stay together only if the speculative walk refused nothing from
                // here on (it refuses long literals: a real one ahead would have thrown it off)
                if (c != kNoBoundary && x.distrust[j] <= pos) {
                    end = j; delta = oo + x.spec_out[j] - c; leave = x.exit_pos[j]; break;
                }
            }
            uint64_t esz, olen;
            if (!raw_element(s, pos, n, esz, olen)) break;                   // truncated: leave the run open
            pos += esz; oo += olen;
            if (pos >= n) { if (pos == n) { end = n_seg - 1; delta = oo; leave = n; } break; }
            if (pos >= s1) { j = (uint32_t)(pos / seg); s0 = (uint64_t)j * seg; s1 = s0 + seg < n ? s0 + seg : n; }
        }
    }
    x.run_end[k] = end; x.run_delta[k] = delta; x.run_exit[k] = leave;
}

// 2c.  Which segments lead a run of the true chain?  Every segment the chain enters (entry inside the
// segment) leads one, unless it lies inside the run of an earlier leader.  Runs of more than one
// segment are rare (a speculation that met nothing within its window; an incompressible region), so:
// collect them in order (parallel compaction), decide among THEM serially which are real (not covered
// by an earlier real one), clear the segments they cover, check that every leader's run ends where
// the next entry begins, and turn the leaders' output counts into offsets with a scan.
__global__ void __launch_bounds__(1024) raw_chain_kernel(uint64_t n, uint64_t seg, uint32_t n_seg, RawIndex x) {
    __shared__ unsigned long long warp_sum[32];
    __shared__ unsigned long long carry_s;
    __shared__ uint32_t warp_cnt[32];
    __shared__ uint32_t m_count, bad, last_leader;
    const uint32_t t = threadIdx.x, lane = t & 31, wid = t >> 5;
    if (t == 0) { m_count = 0; bad = 0; carry_s = 0; last_leader = 0; }
    __syncthreads();
    // A: entered segments are tentative leaders; multi-segment runs are listed in order
    for (uint32_t base = 0; base < n_seg; base += 1024) {
        const uint32_t k = base + t;
        bool entered = false, multi = false;
        if (k < n_seg) {
            const uint64_t s0 = (uint64_t)k * seg, s1 = s0 + seg < n ? s0 + seg : n;
            entered = x.entry_pos[k] < s1;
            multi = entered && x.run_end[k] != k;
            x.lead[k] = entered ? 1u : 0u;
        }
        const uint32_t bal = __ballot_sync(kFull, multi);
        if (lane == 0) warp_cnt[wid] = __popc(bal);
        __syncthreads();
        uint32_t before = 0;
        for (uint32_t w = 0; w < wid; ++w) before += warp_cnt[w];
        if (multi) x.mlist[m_count + before + __popc(bal & ((1u << lane) - 1u))] = k;
        __syncthreads();
        if (t == 0) { uint32_t tot = 0; for (uint32_t w = 0; w < 32; ++w) tot += warp_cnt[w]; m_count += tot; }
        __syncthreads();
    }
    // B: among the multi-segment runs, in order, the real ones are those not inside an earlier real one
    if (t == 0) {
        uint32_t covered_until = 0;                      // segments below this index are inside a real run
        for (uint32_t i = 0; i < m_count; ++i) {
            const uint32_t k = x.mlist[i];
            if (k < covered_until) { x.mlist[i] = kRunOpen; continue; }       // a walk from a wrong entry
            const uint32_t e = x.run_end[k];
            if (e == kRunOpen) { bad = 1; break; }
            covered_until = e + 1;
        }
    }
    __syncthreads();
    // C: segments inside a real multi-segment run do not lead
    for (uint32_t i = wid; i < m_count && !bad; i += 32) {
        const uint32_t k = x.mlist[i];
        if (k == kRunOpen) continue;
        const uint32_t e = x.run_end[k];
        for (uint32_t j = k + 1 + lane; j <= e && j < n_seg; j += 32) x.lead[j] = 0;
    }
    __syncthreads();
    // D: every leader's run must end where the next entry begins; E: offsets
    for (uint32_t base = 0; base < n_seg && !bad; base += 1024) {
        const uint32_t k = base + t;
        unsigned long long v = 0;
        if (k < n_seg && x.lead[k]) {
            const uint32_t e = x.run_end[k];
            if (e == kRunOpen) bad = 1;
            else {
                const uint64_t leave = x.run_exit[k];
                if (e + 1 < n_seg ? x.entry_pos[e + 1] != leave : leave != n) bad = 1;
                v = x.run_delta[k];
                atomicMax(&last_leader, k);
            }
        }
        unsigned long long xs = v;
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned long long y = __shfl_up_sync(kFull, xs, d);
            if (lane >= (uint32_t)d) xs += y;
        }
        if (lane == 31) warp_sum[wid] = xs;
        __syncthreads();
        if (wid == 0) {
            unsigned long long sv = warp_sum[lane], tv = sv;
            for (int d = 1; d < 32; d <<= 1) {
                const unsigned long long y = __shfl_up_sync(kFull, tv, d);
                if (lane >= (uint32_t)d) tv += y;
            }
            warp_sum[lane] = tv - sv;
        }
        __syncthreads();
        const unsigned long long excl = carry_s + warp_sum[wid] + (xs - v);
        if (k < n_seg) x.entry_out[k] = excl;
        __syncthreads();
        if (t == 1023) carry_s = excl + v;
        __syncthreads();
    }
    // the last run must leave exactly at the end of the stream (segments after it are spanned by its last element)
    if (t == 0 && (bad || carry_s != *x.ulen || x.run_exit[last_leader] != n)) *x.serial = 1;
}

// 3. compressed offset of every 64 KiB block: the leader of a run walks the whole run (lockstep, as 1.)
__global__ void __launch_bounds__(64) raw_blocks_kernel(const uint8_t* __restrict__ s, uint64_t n, uint64_t seg,
                                                        uint32_t n_seg, uint64_t n_blk, RawIndex x) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_seg || *x.serial || !x.lead[k]) return;
    const uint64_t stop = x.run_exit[k];
    uint64_t pos = x.entry_pos[k], o = x.entry_out[k];
    while (pos < stop) {
        if ((o & (kParBlock - 1)) == 0) {
            const uint64_t m = o / kParBlock;
            if (m < n_blk) x.blk_in[m] = pos;
        }
        uint64_t esz, olen;
        if (!raw_element_lockstep(s, pos, n, esz, olen)) { *x.serial = 1; return; }
        if ((o & (kParBlock - 1)) + olen > kParBlock) { *x.serial = 1; return; }   // element straddles a block
        pos += esz; o += olen;
    }
}

// block table: lengths from neighbouring offsets; a missing block clears the fast path
__global__ void raw_block_table_kernel(uint64_t n, uint64_t n_blk, RawIndex x) {
    const uint64_t m = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= n_blk) return;
    x.blk_out[m] = m * kParBlock; x.blk_len[m] = 0; x.blk_cap[m] = 0; x.blk_got[m] = 0;
    if (*x.serial) return;
    const uint64_t ulen = *x.ulen, blocks = (ulen + kParBlock - 1) / kParBlock;
    if (m >= blocks) return;
    const uint64_t a = x.blk_in[m], b = m + 1 < blocks ? x.blk_in[m + 1] : n;
    if (a == ~0ull || b == ~0ull || b < a || b - a > 0xffffffffull) { *x.serial = 1; return; }
    x.blk_len[m] = (uint32_t)(b - a);
    const uint64_t left = ulen - m * kParBlock;
    x.blk_cap[m] = left < kParBlock ? (uint32_t)left : kParBlock;
}

// after the block decode: every block must have produced exactly its share; then the length is final
__global__ void raw_verify_kernel(uint64_t n_blk, RawIndex x) {
    const uint64_t m = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= n_blk || *x.serial) return;
    const uint64_t blocks = (*x.ulen + kParBlock - 1) / kParBlock;
    if (m < blocks && x.blk_got[m] != x.blk_cap[m]) *x.serial = 1;
}
__global__ void raw_finish_kernel(RawIndex x, unsigned long long* out_len64) {
    if (threadIdx.x || blockIdx.x) return;
    if (!*x.serial && out_len64) *out_len64 = *x.ulen;
}

}  // namespace sb
