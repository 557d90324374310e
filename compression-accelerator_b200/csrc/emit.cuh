// emit.cuh -- from sequence lists to Snappy elements.
//   size_kernel / size_block_kernel  one warp (page-sized jobs) or one CTA per job: encoded bytes of the job (EmitLiteral + EmitCopy sizes,
//                tests/snappy/snappy.cc:337-431) [+ varint header in batch mode].
//   emit_kernel  one CTA per job: block scan of per-sequence input/output lengths, then every
//                thread encodes its own sequence at its final output position; long literals
//                are copied cooperatively with 16-byte stores.  Output is written once.
#pragma once
#include "kernels.h"
#include "snappy_device.cuh"

namespace sb {

__device__ __forceinline__ uint32_t job_len_of(const JobParams& p, uint32_t job) {
    uint64_t left = p.n_total - (uint64_t)job * p.in_stride;
    return left < p.job_len ? (uint32_t)left : p.job_len;
}

__device__ __forceinline__ uint32_t seq_encoded_size(seq_t q) {
    uint32_t lit = seq_lit(q), len = seq_len(q);
    uint32_t n = lit ? literal_header_size(lit) + lit : 0;
    if (len) n += copy_encoded_size(seq_off(q), len);
    return n;
}

__global__ void __launch_bounds__(128) size_kernel(JobParams p) {            // page-sized jobs: one warp each
    const int lane = threadIdx.x & 31;
    const uint32_t job = blockIdx.x * 4 + (threadIdx.x >> 5);
    if (job >= p.n_jobs) return;
    const seq_t* q = p.seqs + (uint64_t)job * seq_stride(p.job_len);
    const uint32_t n = p.nseq[job];
    if (n == kSeqEmitted) return;                                            // sizes[job] is final already
    uint32_t sum = 0;
    for (uint32_t i = lane; i < n; i += 32) sum += seq_encoded_size(q[i]);
    for (int d = 16; d; d >>= 1) sum += __shfl_xor_sync(kFull, sum, d);
    if (lane == 0) p.sizes[job] = sum + (p.with_header ? varint32_len(job_len_of(p, job)) : 0);
}

constexpr int kSizeThreads = 256;
__global__ void __launch_bounds__(kSizeThreads) size_block_kernel(JobParams p) {
    // one CTA per job: a lone warp walking a fragment's ~7 K sequences is a chain of ~230 dependent
    // loads (130 us for any number of fragments up to a full wave)
    __shared__ uint32_t part[kSizeThreads / 32];
    const uint32_t tid = threadIdx.x, job = blockIdx.x;
    const seq_t* q = p.seqs + (uint64_t)job * seq_stride(p.job_len);
    const uint32_t n = p.nseq[job];
    if (n == kSeqEmitted) return;
    uint32_t sum = 0;
    for (uint32_t i = tid; i < n; i += kSizeThreads) sum += seq_encoded_size(q[i]);
    for (int d = 16; d; d >>= 1) sum += __shfl_xor_sync(kFull, sum, d);
    if ((tid & 31) == 0) part[tid >> 5] = sum;
    __syncthreads();
    if (tid == 0) {
        uint32_t total = 0;
        for (int w = 0; w < kSizeThreads / 32; ++w) total += part[w];
        p.sizes[job] = total + (p.with_header ? varint32_len(job_len_of(p, job)) : 0);
    }
}

// Cooperative copy of n bytes global -> global, both arbitrarily aligned: destination-aligned
// 16-byte stores assembled from aligned 4-byte source words (funnel shift).
__device__ __forceinline__ void block_copy(uint8_t* __restrict__ d, const uint8_t* __restrict__ s,
                                           uint32_t n, uint32_t tid, uint32_t nthreads) {
    uint32_t head = (uint32_t)((16u - ((uintptr_t)d & 15u)) & 15u);
    if (head > n) head = n;
    if (tid < head) d[tid] = s[tid];
    const uint32_t body = (n - head) >> 4;
    if (body) {
        uint4* d4 = reinterpret_cast<uint4*>(d + head);
        const uint8_t* sb_ = s + head;
        const uint32_t* w = reinterpret_cast<const uint32_t*>((uintptr_t)sb_ & ~(uintptr_t)3);
        const uint32_t sh = ((uint32_t)(uintptr_t)sb_ & 3u) << 3;
        for (uint32_t j = tid; j < body; j += nthreads) {
            const uint32_t* q = w + (j << 2);
            uint32_t a0 = __ldg(q), a1 = __ldg(q + 1), a2 = __ldg(q + 2), a3 = __ldg(q + 3);
            uint32_t a4 = sh ? __ldg(q + 4) : 0u;      // never read past the source when aligned
            uint4 v;
            v.x = shf_r(a0, a1, sh);
            v.y = shf_r(a1, a2, sh);
            v.z = shf_r(a2, a3, sh);
            v.w = shf_r(a3, a4, sh);
            d4[j] = v;
        }
    }
    const uint32_t done = head + (body << 4);
    if (done + tid < n) d[done + tid] = s[done + tid];
}

constexpr int kEmitThreads = 256;
constexpr uint32_t kShortLiteral = 32;   // literals up to this many bytes are copied by their thread
constexpr int kMaxLong = kEmitThreads;   // long-literal work items per tile

__global__ void __launch_bounds__(kEmitThreads) emit_kernel(JobParams p) {
    __shared__ unsigned long long warp_tot[kEmitThreads / 32];
    __shared__ unsigned long long carry_s;
    __shared__ uint32_t long_src[kMaxLong], long_dst[kMaxLong], long_len[kMaxLong];
    __shared__ uint32_t n_long;
    const uint32_t job = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t n = p.nseq[job];
    if (n == kSeqEmitted) return;                            // the parse kernel wrote this job's stream (whole CTA leaves)
    const uint32_t len = job_len_of(p, job);
    const uint8_t* __restrict__ in = p.in + (uint64_t)job * p.in_stride;
    uint8_t* __restrict__ out = p.out + (p.offsets ? p.offsets[job] : (uint64_t)job * p.out_stride);
    const seq_t* __restrict__ q = p.seqs + (uint64_t)job * seq_stride(p.job_len);
    uint32_t hdr = 0;
    if (p.with_header) {
        hdr = varint32_len(len);
        if (tid == 0) put_varint32(out, len);
    }
    if (tid == 0) { carry_s = (unsigned long long)hdr; n_long = 0; }   // hi32 = input pos, lo32 = output pos
    __syncthreads();
    for (uint32_t base = 0; base < n; base += kEmitThreads) {
        const uint32_t i = base + tid;
        seq_t sq = i < n ? q[i] : 0ull;
        const uint32_t lit = seq_lit(sq), clen = seq_len(sq), off = seq_off(sq);
        const uint32_t lh = lit ? literal_header_size(lit) : 0;
        const uint32_t osz = lh + lit + (clen ? copy_encoded_size(off, clen) : 0);
        unsigned long long v = ((unsigned long long)(lit + clen) << 32) | osz, x = v;
        for (int d = 1; d < 32; d <<= 1) {
            unsigned long long y = __shfl_up_sync(kFull, x, d);
            if (lane >= d) x += y;
        }
        if (lane == 31) warp_tot[wid] = x;
        __syncthreads();
        unsigned long long pre = carry_s;
        for (uint32_t k = 0; k < wid; ++k) pre += warp_tot[k];
        const unsigned long long excl = pre + x - v;
        const uint32_t ipos = (uint32_t)(excl >> 32);
        uint32_t op = (uint32_t)excl;
        if (lit) {                                           // snappy.cc:337-378 EmitLiteral
            const uint32_t m = lit - 1;
            if (m < 60) out[op] = (uint8_t)(m << 2);
            else if (m < 256) { out[op] = 60u << 2; out[op + 1] = (uint8_t)m; }
            else { out[op] = 61u << 2; out[op + 1] = (uint8_t)m; out[op + 2] = (uint8_t)(m >> 8); }
            op += lh;
            if (lit <= kShortLiteral) {
                for (uint32_t b = 0; b < lit; ++b) out[op + b] = __ldg(in + ipos + b);
            } else {
                uint32_t slot = atomicAdd(&n_long, 1u);
                long_src[slot] = ipos; long_dst[slot] = op; long_len[slot] = lit;
            }
            op += lit;
        }
        if (clen) {                                          // snappy.cc:402-431 EmitCopy
            uint32_t rem = clen;
            if (rem >= 12) {
                while (rem >= 68) { op += put_copy_le64(out + op, off, 64, false); rem -= 64; }
                if (rem > 64) { op += put_copy_le64(out + op, off, 60, false); rem -= 60; }
            }
            put_copy_le64(out + op, off, rem, rem < 12);
        }
        __syncthreads();
        const uint32_t nl = n_long;
        for (uint32_t k = 0; k < nl; ++k)
            block_copy(out + long_dst[k], in + long_src[k], long_len[k], tid, kEmitThreads);
        __syncthreads();
        if (tid == kEmitThreads - 1) { carry_s = excl + v; n_long = 0; }
        __syncthreads();
    }
}


// ---- sparse (position-indexed) copy slots of the fast mode (parse_fast.cuh) ----------------------
// Slot i of a job describes the copy that starts at position 4i + (slot & 3), if any:
// offset << 16 | (len - 4) << 2 | (position & 3); zero = none.  The literal before a copy is the gap
// since the end of the previous copy, so both kernels carry a running "end of the last copy"
// (a prefix maximum: copy ends increase along the stream).
__device__ __forceinline__ uint32_t sparse_pos(uint32_t slot_index, uint32_t v) { return (slot_index << 2) + (v & 3u); }
__device__ __forceinline__ uint32_t sparse_len(uint32_t v) { return ((v >> 2) & 0x3fffu) + 4u; }
__device__ __forceinline__ uint32_t sparse_off(uint32_t v) { return v >> 16; }
__device__ __forceinline__ uint32_t sparse_slots(uint32_t len) { return (len + 3u) >> 2; }

// encoded bytes of the literal [from, to) (nothing when empty)
__device__ __forceinline__ uint32_t literal_encoded_size(uint32_t from, uint32_t to) {
    return to > from ? literal_header_size(to - from) + (to - from) : 0u;
}

// Both kernels give one CTA to a job and one contiguous segment of the slots to each of its warps.
// A warp walks its segment 32 slots at a time; the end of the previous copy comes from the previous
// non-empty lane (one ballot + one shuffle: copy ends increase along the stream).  Only the literal
// in front of a segment's FIRST copy depends on earlier segments, so a segment is summarised as
// (position of its first copy, end of its last copy, encoded bytes without that first literal) and
// the eight summaries are chained by a few threads.
constexpr int kSparseWarps = kEmitThreads / 32;
constexpr uint32_t kNoPos = 0xffffffffu;

struct SparseWalk {                      // per-lane view of 32 slots
    uint32_t v, pos, clen, off, end, before, lit, osz;
};

// before = end of the last copy in front of my slot (carry for lanes with no non-empty lane below)
__device__ __forceinline__ uint32_t sparse_fetch(const uint32_t* __restrict__ q, uint32_t i, uint32_t n) {
    return i < n ? __ldg(q + i) : 0u;
}
// `v` = slot i, fetched one round ahead by the walks below (a round's slots are an L2 / DRAM round trip that
// nothing else in the round can hide)
__device__ __forceinline__ void sparse_load(uint32_t v, uint32_t i,
                                            uint32_t carry, uint32_t lane, SparseWalk& k, uint32_t& nonempty) {
    k.v = v;
    k.pos = sparse_pos(i, k.v); k.clen = sparse_len(k.v); k.off = sparse_off(k.v);
    k.end = k.v ? k.pos + k.clen : 0u;
    nonempty = __ballot_sync(kFull, k.v != 0u);
    const uint32_t below = nonempty & ((1u << lane) - 1u);
    const uint32_t pe = __shfl_sync(kFull, k.end, below ? 31 - __clz(below) : 0);
    k.before = below ? pe : carry;
}

__device__ __forceinline__ void sparse_segment(uint32_t n, uint32_t warp, uint32_t& lo, uint32_t& hi) {
    const uint32_t per = (((n + kSparseWarps - 1) / kSparseWarps) + 31u) & ~31u;
    lo = min(n, warp * per);
    hi = min(n, lo + per);
}

// Summary of one segment; every lane returns the same values.
__device__ __forceinline__ void sparse_summary(const uint32_t* __restrict__ q, uint32_t n, uint32_t lo,
                                               uint32_t hi, uint32_t lane, uint32_t& first_pos,
                                               uint32_t& last_end, uint32_t& bytes) {
    first_pos = kNoPos; last_end = kNoPos;
    uint32_t sum = 0, carry = 0;
    const uint32_t lim = min(n, hi);
    uint32_t v_next = sparse_fetch(q, lo + lane, lim);
    for (uint32_t base = lo; base < hi; base += 32) {
        SparseWalk k;
        uint32_t nonempty;
        const uint32_t v = v_next;
        v_next = sparse_fetch(q, base + 32 + lane, lim);
        sparse_load(v, base + lane, carry, lane, k, nonempty);
        if (!nonempty) continue;
        const uint32_t fl = __ffs(nonempty) - 1, ll = 31 - __clz(nonempty);
        const bool seg_first = first_pos == kNoPos && lane == fl;      // its literal is added by the chain
        if (k.v) sum += (seg_first ? 0u : literal_encoded_size(k.before, k.pos)) + copy_encoded_size(k.off, k.clen);
        if (first_pos == kNoPos) first_pos = __shfl_sync(kFull, k.pos, fl);
        carry = last_end = __shfl_sync(kFull, k.end, ll);
    }
    bytes = __reduce_add_sync(kFull, sum);
}

// The size pass leaves its eight segment summaries behind the job's slots (the job's region of
// `seqs` is twice as large as the slot list), so the emit pass does not walk the slots a third time.
__device__ __forceinline__ bool sparse_keeps_summaries(uint32_t job_len) { return job_len >= 64u; }
__device__ __forceinline__ uint32_t sparse_summary_at(uint32_t job_len) { return ((job_len + 31u) >> 5) << 3; }

struct SparseChain {                     // shared: summaries in, carry-ins out
    uint32_t first_pos[kSparseWarps], last_end[kSparseWarps], bytes[kSparseWarps];
    uint32_t end_in[kSparseWarps + 1], out_in[kSparseWarps + 1];
};

__device__ __forceinline__ void sparse_chain(SparseChain& c, uint32_t out0) {   // one thread
    uint32_t end = 0, out = out0;
    for (int w = 0; w < kSparseWarps; ++w) {
        c.end_in[w] = end; c.out_in[w] = out;
        if (c.first_pos[w] != kNoPos) out += literal_encoded_size(end, c.first_pos[w]);
        out += c.bytes[w];
        if (c.last_end[w] != kNoPos) end = c.last_end[w];
    }
    c.end_in[kSparseWarps] = end; c.out_in[kSparseWarps] = out;
}

__global__ void __launch_bounds__(kEmitThreads) size_sparse_kernel(JobParams p) {
    __shared__ SparseChain c;
    const uint32_t job = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t len = job_len_of(p, job);
    uint32_t* q = reinterpret_cast<uint32_t*>(p.seqs + (uint64_t)job * seq_stride(p.job_len));
    const uint32_t n = sparse_slots(len);
    uint32_t lo, hi, fp, le, by;
    sparse_segment(n, wid, lo, hi);
    sparse_summary(q, n, lo, hi, lane, fp, le, by);
    if (lane == 0) {
        c.first_pos[wid] = fp; c.last_end[wid] = le; c.bytes[wid] = by;
        if (sparse_keeps_summaries(p.job_len)) {
            uint32_t* keep = q + sparse_summary_at(p.job_len) + 3 * wid;
            keep[0] = fp; keep[1] = le; keep[2] = by;
        }
    }
    __syncthreads();
    if (tid == 0) {
        sparse_chain(c, p.with_header ? varint32_len(len) : 0u);
        p.sizes[job] = c.out_in[kSparseWarps] + literal_encoded_size(c.end_in[kSparseWarps], len);   // + trailing literal
    }
}

// warp-cooperative copy of n bytes global -> global (any alignment)
__device__ __forceinline__ void warp_copy(uint8_t* __restrict__ d, const uint8_t* __restrict__ s, uint32_t n,
                                          uint32_t lane) {
    block_copy(d, s, n, lane, 32);
}

__global__ void __launch_bounds__(kEmitThreads) emit_sparse_kernel(JobParams p) {
    __shared__ SparseChain c;
    const uint32_t job = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t len = job_len_of(p, job);
    const uint8_t* __restrict__ in = p.in + (uint64_t)job * p.in_stride;
    uint8_t* __restrict__ out = p.out + (p.offsets ? p.offsets[job] : (uint64_t)job * p.out_stride);
    const uint32_t* __restrict__ q = reinterpret_cast<const uint32_t*>(p.seqs + (uint64_t)job * seq_stride(p.job_len));
    const uint32_t n = sparse_slots(len);
    uint32_t hdr = 0;
    if (p.with_header) {
        hdr = varint32_len(len);
        if (tid == 0) put_varint32(out, len);
    }
    uint32_t lo, hi, fp, le, by;
    sparse_segment(n, wid, lo, hi);
    if (sparse_keeps_summaries(p.job_len)) {                 // left by size_sparse_kernel
        const uint32_t* keep = q + sparse_summary_at(p.job_len) + 3 * wid;
        fp = __ldg(keep); le = __ldg(keep + 1); by = __ldg(keep + 2);
    } else {
        sparse_summary(q, n, lo, hi, lane, fp, le, by);
    }
    if (lane == 0) { c.first_pos[wid] = fp; c.last_end[wid] = le; c.bytes[wid] = by; }
    __syncthreads();
    if (tid == 0) sparse_chain(c, hdr);
    __syncthreads();
    // second walk, now with the true carries: write the elements
    uint32_t carry = c.end_in[wid], op0 = c.out_in[wid];
    const uint32_t lim = min(n, hi);
    uint32_t v_next = sparse_fetch(q, lo + lane, lim);
    for (uint32_t base = lo; base < hi; base += 32) {
        SparseWalk k;
        uint32_t nonempty;
        const uint32_t v = v_next;
        v_next = sparse_fetch(q, base + 32 + lane, lim);
        sparse_load(v, base + lane, carry, lane, k, nonempty);
        if (!nonempty) continue;
        k.lit = (k.v && k.pos > k.before) ? k.pos - k.before : 0u;
        const uint32_t lh = k.lit ? literal_header_size(k.lit) : 0u;
        k.osz = k.v ? lh + k.lit + copy_encoded_size(k.off, k.clen) : 0u;
        uint32_t sx = k.osz;                                  // inclusive prefix sum of the encoded sizes
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t y = __shfl_up_sync(kFull, sx, d);
            if (lane >= (uint32_t)d) sx += y;
        }
        uint32_t op = op0 + sx - k.osz;
        if (k.lit) {                                         // snappy.cc:337-378 EmitLiteral
            const uint32_t m = k.lit - 1;
            if (m < 60) out[op] = (uint8_t)(m << 2);
            else if (m < 256) { out[op] = 60u << 2; out[op + 1] = (uint8_t)m; }
            else { out[op] = 61u << 2; out[op + 1] = (uint8_t)m; out[op + 2] = (uint8_t)(m >> 8); }
            op += lh;
            if (k.lit <= kShortLiteral)
                for (uint32_t b = 0; b < k.lit; ++b) out[op + b] = __ldg(in + k.before + b);
        }
        const uint32_t lit_dst = op;
        op += k.lit;
        if (k.v) {                                           // snappy.cc:402-431 EmitCopy
            uint32_t rem = k.clen;
            if (rem >= 12) {
                while (rem >= 68) { op += put_copy_le64(out + op, k.off, 64, false); rem -= 64; }
                if (rem > 64) { op += put_copy_le64(out + op, k.off, 60, false); rem -= 60; }
            }
            put_copy_le64(out + op, k.off, rem, rem < 12);
        }
        uint32_t longs = __ballot_sync(kFull, k.lit > kShortLiteral);   // long literals: the whole warp copies
        while (longs) {
            const int l = __ffs(longs) - 1;
            longs &= longs - 1;
            warp_copy(out + __shfl_sync(kFull, lit_dst, l), in + __shfl_sync(kFull, k.before, l),
                      __shfl_sync(kFull, k.lit, l), lane);
        }
        op0 += __shfl_sync(kFull, sx, 31);
        carry = __shfl_sync(kFull, k.end, 31 - __clz(nonempty));
    }
    // trailing literal (snappy.cc:665-668): the whole block copies it
    const uint32_t from = c.end_in[kSparseWarps];
    if (from < len) {
        const uint32_t lit = len - from, m = lit - 1;
        uint32_t op = c.out_in[kSparseWarps];
        if (tid == 0) {
            if (m < 60) out[op] = (uint8_t)(m << 2);
            else if (m < 256) { out[op] = 60u << 2; out[op + 1] = (uint8_t)m; }
            else { out[op] = 61u << 2; out[op + 1] = (uint8_t)m; out[op + 2] = (uint8_t)(m >> 8); }
        }
        op += literal_header_size(lit);
        block_copy(out + op, in + from, lit, tid, kEmitThreads);
    }
}

}  // namespace sb
