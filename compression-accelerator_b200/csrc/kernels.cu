// kernels.cu -- Snappy block-compression kernels for B200 (sm_100a) and their launchers.
//
//   parse_kernel   one CTA per fragment (stream path) or per small buffer (batch path): the
//                  fragment is staged into shared memory by a 1-D TMA bulk copy, the uint16 hash
//                  table lives in shared memory, and Snappy's greedy parse with its skip
//                  heuristic (tests/snappy/snappy.cc:539-671) produces a sequence list.
//   size_kernel / scan_sizes_kernel / emit_kernel (emit.cuh): sizes, offsets (+ varint header,
//                  snappy.cc:1006-1061) and the final elements, written once.
#include "kernels.h"
#include "snappy_device.cuh"
#ifdef SNAPPY_B200_STATS
namespace sb { __device__ unsigned long long g_stats[16]; }
#endif
#include "parse_serial.cuh"
#include "parse_window.cuh"
#include "parse_pipe.cuh"
#include "parse_fast.cuh"
#include "parse_lane.cuh"
#include "parse_tiny.cuh"
#include "emit.cuh"
#include "decode.cuh"
#include "decode_par.cuh"
#include "frame.cuh"
#include "peer.cuh"
#include <type_traits>
#include <cstdlib>

namespace sb {

// Shared-memory layout of one parse CTA: [0,16) mbarrier | [128, 128+table) table | input.
constexpr uint32_t kSmemTableOff = 128;
constexpr uint32_t kInputPad = 32;   // slack so unaligned word reads past the end stay in bounds

__host__ __device__ inline uint32_t parse_smem_bytes(uint32_t job_len) {
    return kSmemTableOff + table_entries(job_len) * 2 + ((job_len + 15u) & ~15u) + kInputPad;
}

// Stage src[0,len) into shared memory: TMA bulk copy for the 16-byte-aligned body, ordinary
// loads for the tail (or for everything when the source is not 16-byte aligned), while the
// warp zeroes the hash table (snappy.cc:475-481).
__device__ __forceinline__ void stage_fragment(const uint8_t* __restrict__ src, uint32_t len,
                                               uint8_t* s_in, uint16_t* table, uint32_t entries,
                                               uint64_t* bar, int lane) {
    const bool aligned = ((uintptr_t)src & 15u) == 0;
    const uint32_t bulk = aligned ? (len & ~15u) : 0u;
    if (bulk) {
        if (lane == 0) {
            mbar_init(bar, 1);
            mbar_expect_tx(bar, bulk);
            tma_load_1d(s_in, src, bulk, bar);
        }
    }
    uint4* t4 = reinterpret_cast<uint4*>(table);
    const uint4 z = make_uint4(0, 0, 0, 0);
    for (uint32_t i = lane; i < entries / 8; i += 32) t4[i] = z;
    if (bulk) {
        for (uint32_t i = bulk + lane; i < len; i += 32) s_in[i] = src[i];
    } else {
        // unaligned source: whole words of the staged copy from aligned source words (a byte loop was 2 % of the
        // kernel's instructions on 1,000-byte buffers); a source word is read only if it holds a byte of the buffer
        const uint32_t r = (uint32_t)(uintptr_t)src & 3u, sh = r << 3;
        const uint32_t* w = reinterpret_cast<const uint32_t*>(src - r);
        uint32_t* s32 = reinterpret_cast<uint32_t*>(s_in);
        for (uint32_t j = lane; (j << 2) < len; j += 32) {
            const uint32_t a0 = __ldg(w + j);
            const uint32_t a1 = (r && (j << 2) + 4u - r < len) ? __ldg(w + j + 1) : 0u;
            s32[j] = shf_r(a0, a1, sh);
        }
        __syncwarp();                                        // the last word may spill into the padding zeroed below
    }
    uint32_t padded = ((len + 15u) & ~15u) + kInputPad;
    for (uint32_t i = len + lane; i < padded; i += 32) s_in[i] = 0;
    __syncwarp();
    if (bulk) mbar_wait(bar, 0);
}

// The job is one literal: preamble (snappy.cc:1006-1009) + EmitLiteral (:337-378) from the staged copy.
__device__ __noinline__ void emit_one_literal(uint8_t* dst, uint32_t with_header, uint32_t* size_slot, uint32_t* nseq_slot,
                                              uint32_t len, const uint8_t* s_in, int lane) {
    uint32_t op = 0;
    if (with_header) {
        op = varint32_len(len);
        if (lane == 0) put_varint32(dst, len);
    }
    op = warp_emit_literal(dst, op, s_in, 0, len, lane);
    if (lane == 0) { *size_slot = op; *nseq_slot = kSeqEmitted; }
}

// Batch path, short jobs: the warp that parsed the job turns its sequences into elements right away (what
// size_kernel + emit_kernel do with a warp / a 256-thread CTA per job in two more launches: for 1,000-byte text
// buffers those took 5.8 ms per GiB next to 11.4 ms of parsing, most of it CTA launch rate).  32 sequences a round:
// a warp scan of input / output lengths gives every lane its element's place (snappy.cc:337-431 for the bytes);
// literals come from the copy staged in shared memory, the long ones written by the whole warp.
// kSmem: `s_in` is the copy staged in shared memory (4-byte aligned, padded); otherwise the buffer itself in global
// memory, any alignment (the global-table parser reads its input in place).
constexpr uint32_t kLaneLiteral = 16;
template <bool kSmem>
__device__ __noinline__ void warp_emit_job(uint8_t* out, uint32_t with_header, uint32_t* size_slot, uint32_t* nseq_slot,
                                           uint32_t len, const uint8_t* s_in, const seq_t* q, uint32_t n, int lane) {
    unsigned long long carry = 0;                            // hi32 = input position, lo32 = output position
    if (with_header) {
        carry = varint32_len(len);
        if (lane == 0) put_varint32(out, len);
    }
    __syncwarp();                                            // the sequences were written by other lanes
    seq_t sq_next = (uint32_t)lane < n ? q[lane] : 0ull;     // one round ahead: the list comes back from L2
    for (uint32_t base = 0; base < n; base += 32) {
        const seq_t sq = sq_next;
        sq_next = base + 32 + lane < n ? q[base + 32 + lane] : 0ull;
        const uint32_t lit = seq_lit(sq), clen = seq_len(sq), off = seq_off(sq);
        const uint32_t lh = lit ? literal_header_size(lit) : 0;
        const uint32_t osz = lh + lit + (clen ? copy_encoded_size(off, clen) : 0);
        const unsigned long long v = ((unsigned long long)(lit + clen) << 32) | osz;
        unsigned long long x = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned long long y = __shfl_up_sync(kFull, x, d);
            if (lane >= d) x += y;
        }
        const unsigned long long excl = carry + x - v;
        const uint32_t ipos = (uint32_t)(excl >> 32);
        uint32_t op = (uint32_t)excl;
        const bool by_warp = lit > kLaneLiteral;
        if (lit && !by_warp) {                               // EmitLiteral, tag only (lit <= 60)
            out[op] = (uint8_t)((lit - 1) << 2);
            for (uint32_t b = 0; b < lit; ++b) out[op + 1 + b] = s_in[ipos + b];
        }
        op += lh + lit;
        if (clen) {                                          // EmitCopy
            uint32_t rem = clen;
            if (rem >= 12) {
                while (rem >= 68) { op += put_copy_le64(out + op, off, 64, false); rem -= 64; }
                if (rem > 64) { op += put_copy_le64(out + op, off, 60, false); rem -= 60; }
            }
            put_copy_le64(out + op, off, rem, rem < 12);
        }
        uint32_t longs = __ballot_sync(kFull, by_warp);
        while (longs) {
            const int k = __ffs(longs) - 1;
            longs &= longs - 1;
            const uint32_t dst = __shfl_sync(kFull, (uint32_t)excl, k), src = __shfl_sync(kFull, ipos, k);
            const uint32_t L = __shfl_sync(kFull, lit, k);
            if constexpr (kSmem) {
                warp_emit_literal(out, dst, s_in, src, L, lane);
            } else {                                         // EmitLiteral tag (snappy.cc:337-378), then the bytes
                const uint32_t m = L - 1, lhk = literal_header_size(L);
                if (lane == 0) {
                    if (m < 60) out[dst] = (uint8_t)(m << 2);
                    else if (m < 256) { out[dst] = 60u << 2; out[dst + 1] = (uint8_t)m; }
                    else { out[dst] = 61u << 2; out[dst + 1] = (uint8_t)m; out[dst + 2] = (uint8_t)(m >> 8); }
                }
                block_copy(out + dst + lhk, s_in + src, L, (uint32_t)lane, 32u);
            }
        }
        carry = __shfl_sync(kFull, excl + v, 31);
    }
    if (lane == 0) { *size_slot = (uint32_t)carry; *nseq_slot = kSeqEmitted; }
}

// One warp (one CTA) per job.  Packing several page-sized jobs into one CTA to spread the per-CTA
// shared-memory reservation (18 instead of 17 4 KiB pages per SM) was measured and lost: a CTA's
// resources return only when its slowest job ends, and the kernel compiled for more than one warp per
// CTA ran 12 % slower per page (85 -> 75-77 GB/s on 4 KiB text pages).
// kCtas = 32 caps the kernel at 64 registers (71 otherwise): 32 instead of 28 jobs per SM where shared memory leaves
// room for them -- buffers up to 2 KiB (text, GB/s: 300 B 72.6 -> 75.8, 1,000 B 85.5 -> 88.7, 2,000 B 93.8 -> 97.1;
// 4 KiB pages fit 17 per SM either way and lose 3 % on incompressible data, so they keep the uncapped build).
template <bool kWindow, int kCtas = 1>
__global__ void __launch_bounds__(32, kCtas) parse_kernel(JobParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int lane = threadIdx.x;
    const uint32_t job = blockIdx.x;
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem);
    uint16_t* table = reinterpret_cast<uint16_t*>(smem + kSmemTableOff);
    const uint32_t len = job_len_of(p, job);
    const uint32_t entries = table_entries(len);            // from THIS fragment's length
    uint8_t* s_in = smem + kSmemTableOff + table_entries(p.job_len) * 2;
    stage_fragment(p.in + (uint64_t)job * p.in_stride, len, s_in, table, entries, bar, lane);
    SeqWriter out{p.seqs + (uint64_t)job * seq_stride(p.job_len), 0};
    const int shift = 32 - (31 - __clz(entries));           // snappy.cc:548
    bool one_literal = false;
    if (kWindow) one_literal = parse_fragment_window(SmemSrc{s_in}, len, table, shift, out, lane);
    else parse_fragment_serial(s_in, len, table, shift, out, lane);
    // Batch path (every job at its fixed slot, no scan of sizes needed): a job that parsed to ONE literal --
    // incompressible data -- is written here, from the copy staged in shared memory, so that the emit kernel
    // does not read it from DRAM a second time (snappy.cc:1006-1009 preamble, :337-378 EmitLiteral).
    if (kWindow && one_literal && !p.offsets && !p.sparse) {
        emit_one_literal(p.out + (uint64_t)job * p.out_stride, p.with_header, p.sizes + job, p.nseq + job, len, s_in, lane);
        return;
    }
    if (kWindow && p.fuse_emit) {
        warp_emit_job<true>(p.out + (uint64_t)job * p.out_stride, p.with_header, p.sizes + job, p.nseq + job, len, s_in,
                            out.seqs, out.n, lane);
        return;
    }
    if (lane == 0) p.nseq[job] = out.n;
}

// ---- window parser with the table in GLOBAL memory (kept L2-resident) and the input read in place ----------
// No shared memory at all, so residency is set by registers and warp slots (up to ~28 warps per SM) instead of
// by the 32 KB table (6 fragments per SM in parse_pipe_kernel).  A persistent grid of one-warp CTAs; CTA b owns
// table b of the pool for its whole life (148 x W tables of 32 KB: W = 16 is 76 MB of the 126 MB L2) and takes
// fragments from a ticket counter.  Only this warp ever touches its table, so plain (L1-cached) accesses are
// coherent; each window costs two L2 round trips (table entries, candidate bytes) on top of the parser's own
// instruction latency.
__global__ void __launch_bounds__(32) parse_gtab_kernel(JobParams p, uint16_t* pool, unsigned int* ticket) {
    const int lane = threadIdx.x;
    uint16_t* table = pool + (size_t)blockIdx.x * kMaxTable;
    for (;;) {
        uint32_t job = 0;
        if (lane == 0) job = atomicAdd(ticket, 1u);
        job = __shfl_sync(kFull, job, 0);
        if (job >= p.n_jobs) return;
        const uint32_t len = job_len_of(p, job);
        const uint32_t entries = table_entries(len);             // from THIS fragment's length
        {
            uint4* t4 = reinterpret_cast<uint4*>(table);
            const uint4 z = make_uint4(0, 0, 0, 0);
            for (uint32_t i = lane; i < entries / 8; i += 32) t4[i] = z;
        }
        __syncwarp();
        SeqWriter out{p.seqs + (uint64_t)job * seq_stride(p.job_len), 0};
        const int shift = 32 - (31 - __clz(entries));
        GlobalSrc src;
        src.g = p.in + (uint64_t)job * p.in_stride;
        parse_fragment_window(src, len, table, shift, out, lane);
        if (p.fuse_emit) {                                   // batch path: this warp writes the buffer's elements too
            warp_emit_job<false>(p.out + (uint64_t)job * p.out_stride, p.with_header, p.sizes + job, p.nseq + job, len,
                                 src.g, out.seqs, out.n, lane);
            __syncwarp();
            continue;
        }
        if (lane == 0) p.nseq[job] = out.n;
        __syncwarp();
    }
}

// ---- pipelined parser: one resolver warp + P producer warps per fragment ----------------------
// kGlobal = false: fragment staged in shared memory by TMA (2 fragments per SM).
// kGlobal = true : producers gather the input from global memory; shared memory holds only the
//                  table and the ring (up to 6 fragments per SM).
template <bool kGlobal> struct PipeCfg;
#ifndef SB_PIPE_MIN_LEN
#define SB_PIPE_MIN_LEN 8192             // jobs up to this length take the single-warp parser
#endif
#ifndef SB_PIPE_P
#define SB_PIPE_P 4
#endif
#ifndef SB_PIPE_LEAD
#define SB_PIPE_LEAD 2
#endif
template <> struct PipeCfg<false> { static constexpr uint32_t P = 2, R = 8, LEAD = 3; };
#ifndef SB_PIPE_R
#define SB_PIPE_R 8
#endif
template <> struct PipeCfg<true>  { static constexpr uint32_t P = SB_PIPE_P, R = SB_PIPE_R, LEAD = SB_PIPE_LEAD; };

template <bool kGlobal>
__host__ __device__ inline uint32_t pipe_smem_bytes(uint32_t job_len) {
    uint32_t n = kSmemTableOff + table_entries(job_len) * 2 + (uint32_t)sizeof(PipeShared<PipeCfg<kGlobal>::R>);
    if (!kGlobal) n += ((job_len + 15u) & ~15u) + kInputPad;
    return n;
}

template <bool kGlobal>
__global__ void __launch_bounds__(32 * (1 + PipeCfg<kGlobal>::P)) parse_pipe_kernel(JobParams p) {
    constexpr uint32_t P = PipeCfg<kGlobal>::P, R = PipeCfg<kGlobal>::R, NT = 32 * (1 + P);
    static_assert(R % P == 0, "one producer must own all windows of a ring slot");
    typedef PipeShared<R> Ring;
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t job = blockIdx.x;
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem);
    uint16_t* table = reinterpret_cast<uint16_t*>(smem + kSmemTableOff);
    Ring* ps = reinterpret_cast<Ring*>(smem + kSmemTableOff + table_entries(p.job_len) * 2);
    uint8_t* s_in = reinterpret_cast<uint8_t*>(ps + 1);
    const uint32_t len = job_len_of(p, job);
    const uint32_t entries = table_entries(len);
    const uint8_t* __restrict__ src = p.in + (uint64_t)job * p.in_stride;

    uint32_t bulk = 0;
    if (!kGlobal) {                                         // TMA bulk copy of the 16-byte aligned body
        bulk = (((uintptr_t)src & 15u) == 0) ? (len & ~15u) : 0u;
        if (tid == 0 && bulk) {
            mbar_init(bar, 1);
            mbar_expect_tx(bar, bulk);
            tma_load_1d(s_in, src, bulk, bar);
        }
    }
    {
        uint4* t4 = reinterpret_cast<uint4*>(table);
        const uint4 z = make_uint4(0, 0, 0, 0);
        for (uint32_t i = tid; i < entries / 8; i += NT) t4[i] = z;
        if (!kGlobal) {
            for (uint32_t i = bulk + tid; i < len; i += NT) s_in[i] = src[i];
            const uint32_t padded = ((len + 15u) & ~15u) + kInputPad;
            for (uint32_t i = len + tid; i < padded; i += NT) s_in[i] = 0;
        }
        if (tid < R) ps->ready[tid] = 0;
        if (tid == 0) { ps->cur_win = 0; ps->err = p.err; mbar_init(&ps->bar, 1); }
    }
    __syncthreads();
    if (bulk) mbar_wait(bar, 0);

    typedef typename std::conditional<kGlobal, GlobalSrc, SmemSrc>::type Src;
    Src in_src;
    if constexpr (kGlobal) in_src.g = src; else in_src.s = s_in;
#ifdef SNAPPY_B200_STATS
    const long long t_start = clock64();
#endif
    const int shift = 32 - (31 - __clz(entries));
    // spread the resolver warps of co-resident CTAs over the SM's four schedulers
    const uint32_t rwarp = (blockIdx.x / 148u) % (1 + P);
    if (wid == rwarp) {
        PipeResolver<Src, R> pr;
        pr.next_emit = 0;
        pr.seqs = p.seqs + (uint64_t)job * seq_stride(p.job_len);
        pr.nseq = 0;
        pr.lane = (int)lane;
        if (len >= kMargin) {
            pr.src = in_src; pr.len = len; pr.ip_limit = len - kMargin;
            pr.table = table; pr.ps = ps; pr.shift = shift;
            pr.table_s = (uint32_t)__cvta_generic_to_shared(table);
            pr.run();
        }
        __syncwarp();
        if (lane == 0) { ps->cur_win = kDone; mbar_arrive(&ps->bar); }
        if (pr.next_emit < len) pr.push(len - pr.next_emit, 0, 0);
        if (lane == 0) p.nseq[job] = pr.nseq;
#ifdef SNAPPY_B200_STATS
        if (lane == 0) {
            atomicAdd(&g_stats[0], 1ull); atomicAdd(&g_stats[1], (unsigned long long)(clock64() - t_start));
            atomicAdd(&g_stats[2], pr.st_wait); atomicAdd(&g_stats[3], (unsigned long long)pr.st_win);
            atomicAdd(&g_stats[4], (unsigned long long)pr.st_stale); atomicAdd(&g_stats[5], (unsigned long long)pr.st_nslow);
            atomicAdd(&g_stats[6], pr.st_slow); atomicAdd(&g_stats[7], (unsigned long long)pr.st_serial);
            atomicAdd(&g_stats[8], (unsigned long long)pr.st_slowhit); atomicAdd(&g_stats[9], (unsigned long long)pr.nseq);
            atomicAdd(&g_stats[12], pr.st_pa); atomicAdd(&g_stats[13], pr.st_pb); atomicAdd(&g_stats[14], pr.st_pc);
            atomicAdd(&g_stats[10], (unsigned long long)pr.st_cstale); atomicAdd(&g_stats[11], (unsigned long long)pr.st_cpeer);
        }
#endif
    } else if (len >= kMargin) {
        const uint32_t pidx = wid > rwarp ? wid - 1 : wid;
        pipe_producer<Src, R, PipeCfg<kGlobal>::LEAD>(in_src, table, shift, len, (len + 31) >> 5, ps, pidx, P, (int)lane);
    }
}


// ---- fast mode: table warp + chaser warp + K evaluator warps per fragment (parse_fast.cuh) --------
__global__ void __launch_bounds__(32 * (2 + kFastMaxK), 2) parse_fast_kernel(JobParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, NT = blockDim.x;
    // Roles: warp 0 = table, warp 1 = chaser, warps 2.. = evaluators.
    const uint32_t K = (NT >> 5) - 2;
    const uint32_t role_t = 0, role_c = 1;
    const uint32_t e_idx = wid - 2;
    const bool idle = false;
    const uint32_t job = blockIdx.x;
    const FastLayout lay = fast_layout(p.job_len);
    FastCtl* ctl = reinterpret_cast<FastCtl*>(smem);
    uint16_t* table = reinterpret_cast<uint16_t*>(smem + lay.table);
    uint16_t* cring = reinterpret_cast<uint16_t*>(smem + lay.cring);
    uint32_t* ering = reinterpret_cast<uint32_t*>(smem + lay.ering);
    uint8_t* entry = smem + lay.entry;
    uint8_t* s_in = smem + lay.input;
    const uint32_t len = job_len_of(p, job);
    const uint32_t entries = table_entries(len);
    const uint32_t nwin = (len + 31) >> 5;
    const uint8_t* __restrict__ src = p.in + (uint64_t)job * p.in_stride;

    const uint32_t bulk = (((uintptr_t)src & 15u) == 0) ? (len & ~15u) : 0u;
    if (tid == 0) {
        ctl->t_done = 0;
        ctl->abort = 0;
        ctl->err = p.err;
        if (bulk) {                                          // TMA bulk copy of the 16-byte aligned body
            mbar_init(&ctl->bar, 1);
            mbar_expect_tx(&ctl->bar, bulk);
            tma_load_1d(s_in, src, bulk, &ctl->bar);
        }
    }
    {
        uint4* t4 = reinterpret_cast<uint4*>(table);
        const uint4 z = make_uint4(0, 0, 0, 0);
        for (uint32_t i = tid; i < entries / 8; i += NT) t4[i] = z;
        for (uint32_t i = tid; i < kRecRing * 32; i += NT) ering[i] = 0;
        for (uint32_t i = tid; i < nwin; i += NT) entry[i] = (uint8_t)kEntryNone;
        for (uint32_t i = bulk + tid; i < len; i += NT) s_in[i] = src[i];
        const uint32_t padded = ((len + 15u) & ~15u) + kFastPad;
        for (uint32_t i = len + tid; i < padded; i += NT) s_in[i] = 0;
    }
    __syncthreads();
    if (bulk) mbar_wait(&ctl->bar, 0);
    if (idle) return;

    const int shift = 32 - (31 - __clz(entries));
    uint32_t* slots = reinterpret_cast<uint32_t*>(p.seqs + (uint64_t)job * seq_stride(p.job_len));
#ifdef SNAPPY_B200_STATS
    const long long t_start = clock64();
#endif
    FSTAT_DECL();
    if (wid == role_t) {
        fast_table_warp(s_in, len, nwin, table, shift, cring, entry, ctl, (int)lane FSTAT_ARG);
        FSTAT_FLUSH(1, 1, lane);
#ifdef SNAPPY_B200_STATS
        if (lane == 0) { atomicAdd(&g_stats[0], (unsigned long long)(clock64() - t_start)); atomicAdd(&g_stats[9], 1ull); }
#endif
    } else if (wid == role_c) {
        fast_chaser(ering, entry, nwin, 2 * K, (int)lane FSTAT_ARG);
        FSTAT_FLUSH(2, 6, lane); FSTAT_FLUSH(3, 8, lane);
#ifdef SNAPPY_B200_STATS
        if (lane == 0) atomicAdd(&g_stats[7], (unsigned long long)(clock64() - t_start));
#endif
    } else {
        FastWin prev;
        prev.live = false;
        for (uint32_t i = 0, w = e_idx;; ++i, w += K) {
            FastWin cur;
            cur.live = false;
            if (w < nwin) fast_evaluate(s_in, len, cring, ering, entry, ctl, w, e_idx + K * (i & 1u), cur, (int)lane FSTAT_ARG);
            if (prev.live) fast_emit(entry, slots, prev, (int)lane FSTAT_ARG);
            if (!cur.live) break;
            prev = cur;
        }
        FSTAT_FLUSH(2, 2, lane); FSTAT_FLUSH(3, 3, lane); FSTAT_FLUSH(0, 4, lane); FSTAT_FLUSH(1, 5, lane);
    }
    // A watchdog expired (FAST_WATCHDOG): whatever the roles left in the slot list is not to be trusted.
    // With every slot empty the fragment is emitted as one literal -- a valid stream; the call itself fails
    // through the context's error word.
    __syncthreads();
    if (ctl->abort)
        for (uint32_t i = tid; i < sparse_slots(len); i += NT) slots[i] = 0u;
}

// ---- scan of fragment sizes + header ---------------------------------------------------------
__global__ void __launch_bounds__(1024) scan_sizes_kernel(const uint32_t* __restrict__ sizes,
                                                          uint32_t n, unsigned long long* offsets,
                                                          uint32_t hdr_len, uint32_t n_input,
                                                          uint8_t* dst,
                                                          unsigned long long* d_total,
                                                          const unsigned long long* carry_in,
                                                          unsigned long long* h_mirror) {
    __shared__ unsigned long long warp_sum[32];
    __shared__ unsigned long long carry_s;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) {
        carry_s = hdr_len + (carry_in ? *carry_in : 0ull);   // later chunks continue the running offset
        if (hdr_len) put_varint32(dst, n_input);             // snappy.cc:1006-1009
    }
    __syncthreads();
    for (uint32_t base = 0; base < n; base += 1024) {
        uint32_t i = base + tid;
        unsigned long long v = i < n ? sizes[i] : 0ull, x = v;
        for (int d = 1; d < 32; d <<= 1) {
            unsigned long long y = __shfl_up_sync(kFull, x, d);
            if (lane >= d) x += y;
        }
        if (lane == 31) warp_sum[wid] = x;
        __syncthreads();
        if (wid == 0) {
            unsigned long long s = warp_sum[lane], t = s;
            for (int d = 1; d < 32; d <<= 1) {
                unsigned long long y = __shfl_up_sync(kFull, t, d);
                if (lane >= d) t += y;
            }
            warp_sum[lane] = t - s;                           // exclusive over warps
        }
        __syncthreads();
        unsigned long long excl = carry_s + warp_sum[wid] + (x - v);
        if (i < n) offsets[i] = excl;
        __syncthreads();
        if (tid == 1023) carry_s = excl + v;
        __syncthreads();
    }
    if (tid == 0) {
        *d_total = carry_s;
        if (h_mirror) { *h_mirror = carry_s; __threadfence_system(); }   // mapped pinned host word
    }
}

// ---- debug self-check of the sequence lists (SNAPPY_B200_DEBUG=1) -----------------------------
// report[0] = number of bad jobs, report[1..4] = first bad job: index, nseq, sum of lengths, len
__global__ void __launch_bounds__(128) check_seqs_kernel(JobParams p, unsigned long long* report) {
    const uint32_t job = blockIdx.x * 128 + threadIdx.x;
    if (job >= p.n_jobs) return;
    const uint32_t len = job_len_of(p, job), n = p.nseq[job], cap = seq_stride(p.job_len);
    if (n == kSeqEmitted) return;                             // written by the parse kernel itself
    const seq_t* q = p.seqs + (uint64_t)job * cap;
    unsigned long long sum = 0;
    bool bad = n > cap;
    for (uint32_t i = 0; i < n && !bad; ++i) {
        const uint32_t lit = seq_lit(q[i]), cl = seq_len(q[i]), off = seq_off(q[i]);
        if (cl && (cl < 4 || off == 0 || off > sum + lit)) bad = true;
        if (!cl && i + 1 != n) bad = true;
        sum += lit + cl;
        if (sum > len) bad = true;
    }
    if (sum != len) bad = true;
    if (bad && atomicAdd(report, 1ull) == 0) {
        report[1] = job; report[2] = n; report[3] = sum; report[4] = len;
    }
}

cudaError_t launch_check(const JobParams& p, unsigned long long* report, cudaStream_t st) {
    if (p.n_jobs == 0) return cudaSuccess;
    check_seqs_kernel<<<(p.n_jobs + 127) / 128, 128, 0, st>>>(p, report);
    return cudaGetLastError();
}

// ---- launchers ----------------------------------------------------------------------------------
cudaError_t read_stats(unsigned long long* out16, bool reset) {
#ifdef SNAPPY_B200_STATS
    cudaError_t e = cudaMemcpyFromSymbol(out16, g_stats, 16 * sizeof(unsigned long long));
    if (e == cudaSuccess && reset) { unsigned long long z[16] = {0}; e = cudaMemcpyToSymbol(g_stats, z, sizeof z); }
    return e;
#else
    for (int i = 0; i < 16; ++i) out16[i] = 0;
    (void)reset;
    return cudaSuccess;
#endif
}

// kVar values of parse_lane.cuh that are compiled in (40 is the product default; the others are kept for
// A/B runs through SNAPPY_B200_LANE_VAR)
#define SB_LANE_VARIANTS(X) X(0) X(2) X(32) X(34) X(40) X(42)

cudaError_t kernels_init() {
    cudaError_t e = cudaFuncSetAttribute(parse_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)parse_smem_bytes(kBlock));
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(parse_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)parse_smem_bytes(kBlock));
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(parse_pipe_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)pipe_smem_bytes<false>(kBlock));
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(parse_pipe_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)pipe_smem_bytes<true>(kBlock));
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(parse_fast_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)fast_layout(kBlock).total);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(decode_staged_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)(kStageWarps * (3 * kStageOutMax + 128)));
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(tiny_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTinySmemBytes);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(crc32c_chunks_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kCrcSmemBytes);
    if (e != cudaSuccess) return e;
    // The exchange kernels run beside the parser's CTAs: ask for the parser's shared-memory carve-out
    // (the maximum) so that placing them never calls for another split of an SM.  (Measured neutral on
    // 2 GPUs; kept because a re-partition needs a drained SM.)
    {
        const void* beside[] = {(const void*)peer_publish_kernel, (const void*)peer_copy_kernel,
                                (const void*)peer_finish_kernel, (const void*)peer_wait_kernel};
        for (const void* f : beside) {
            e = cudaFuncSetAttribute(f, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
            if (e != cudaSuccess) return e;
        }
    }
    static CrcOps ops;                                      // "append 2^k zero bytes" operators of CRC-32C
    crc_ops_build(ops);
    return cudaMemcpyToSymbol(g_crc_ops, &ops, sizeof(ops));
}

bool page_parser_takes(uint32_t job_len) { return job_len <= SB_PIPE_MIN_LEN; }

cudaError_t launch_parse(const JobParams& p, int mode, cudaStream_t st,
                         unsigned long long* launches) {
    if (p.n_jobs == 0) return cudaSuccess;
    if (mode == 1) {
        const uint32_t K = fast_evaluators(p.job_len);      // 14 -> the 16-warp CTA (12 of them evaluate)
        parse_fast_kernel<<<p.n_jobs, 32 * (2 + K), fast_layout(p.job_len).total, st>>>(p);
    } else if (mode == 2) parse_kernel<false><<<p.n_jobs, 32, parse_smem_bytes(p.job_len), st>>>(p);
    else if (mode == 0 && p.job_len > SB_PIPE_MIN_LEN)   // small jobs fit many per SM: the single-warp parser wins
        parse_pipe_kernel<true><<<p.n_jobs, 32 * (1 + PipeCfg<true>::P), pipe_smem_bytes<true>(p.job_len), st>>>(p);
    else if (mode == 4)
        parse_pipe_kernel<false><<<p.n_jobs, 32 * (1 + PipeCfg<false>::P), pipe_smem_bytes<false>(p.job_len), st>>>(p);
    else if (p.job_len <= 2048) parse_kernel<true, 32><<<p.n_jobs, 32, parse_smem_bytes(p.job_len), st>>>(p);
    else parse_kernel<true><<<p.n_jobs, 32, parse_smem_bytes(p.job_len), st>>>(p);
    if (launches) ++*launches;
    return cudaGetLastError();
}

// ---- lane parser -----------------------------------------------------------------------------------
static int env_int(const char* name, int dflt) {
    const char* e = getenv(name);
    return e ? atoi(e) : dflt;
}
// tuning knobs: SNAPPY_B200_LANE_VAR = kVar bits of parse_lane.cuh, SNAPPY_B200_LANE_FPW = fragments per warp
static int lane_var() { static const int v = env_int("SNAPPY_B200_LANE_VAR", 40); return v; }

size_t lane_pool_bytes(uint64_t n_jobs, uint32_t job_len) {
    return (size_t)n_jobs * table_entries(job_len) * ((lane_var() & 8) ? 32u : 16u);
}

cudaError_t launch_parse_lane(const JobParams& p, void* pool, uint32_t epoch, cudaStream_t st,
                              unsigned long long* launches, bool beside_pipe) {
    if (p.n_jobs == 0) return cudaSuccess;
    {   // the kernel's preferred carve-out follows the launch's company (alone: all L1; beside the pipelined
        // parser: its maximum shared-memory split).  The attribute is per function, so it is switched here.
        static int current = -2;
        const int want = beside_pipe ? (int)cudaSharedmemCarveoutMaxShared : (int)cudaSharedmemCarveoutDefault;
        if (want != current) {
            cudaError_t e = cudaSuccess;
#define SB_LANE_ATTR(V) if (lane_var() == V) e = cudaFuncSetAttribute(parse_lane_kernel<V>, cudaFuncAttributePreferredSharedMemoryCarveout, want);
            SB_LANE_VARIANTS(SB_LANE_ATTR)
#undef SB_LANE_ATTR
            if (e != cudaSuccess) return e;
            current = want;
        }
    }
    static_assert(sizeof(lane::Entry) == kLaneEntryBytes, "entry size");
    static const int fpw_env = env_int("SNAPPY_B200_LANE_FPW", 0), wpc_env = env_int("SNAPPY_B200_LANE_WPC", 0);
    LaneTable t;
    t.base = reinterpret_cast<uint8_t*>(pool);
    t.stride = table_entries(p.job_len);
    t.epoch = epoch;
    t.fpw = fpw_env >= 1 && fpw_env <= 32 ? (uint32_t)fpw_env : 3u;   // measured on 1 GiB text: 3 -> 21.9 ms, 4 -> 22.4, 8 -> 26.1, 2 -> 31 (registers)
    const uint32_t wpc = wpc_env >= 1 && wpc_env <= kLaneMaxThreads / 32 ? (uint32_t)wpc_env : 2u;
    const uint32_t per_cta = t.fpw * wpc, grid = (p.n_jobs + per_cta - 1) / per_cta;
#define SB_LANE_CASE(V) case V: parse_lane_kernel<V><<<grid, 32 * wpc, 0, st>>>(p, t); break;
    switch (lane_var()) {
        SB_LANE_VARIANTS(SB_LANE_CASE)
        default: return cudaErrorInvalidValue;
    }
#undef SB_LANE_CASE
    if (launches) ++*launches;
    return cudaGetLastError();
}

// ---- window parser, table in global memory ---------------------------------------------------------------
// Warps (= tables) per SM.  Measured on 1 GiB of text, ms per parse: 8 -> 42.1, 16 -> 28.6, 24 -> 25.0, 28 -> 23.8,
// 32 -> 26.1 (throughput per SM stops growing at 28: L1, L2 and DRAM all near 55 % busy, every window issues
// ~330 uncoalesced L1 requests).
uint32_t gtab_tables() {
    static const int w = [] { const char* e = getenv("SNAPPY_B200_GTAB_W"); int v = e ? atoi(e) : 28; return v < 1 ? 1 : v > 32 ? 32 : v; }();
    return 148u * (uint32_t)w;
}
size_t gtab_table_bytes() { return (size_t)kMaxTable * sizeof(uint16_t); }

cudaError_t launch_parse_gtab(const JobParams& p, void* tables, uint32_t n_tables, unsigned int* ticket, cudaStream_t st,
                              unsigned long long* launches) {
    if (p.n_jobs == 0) return cudaSuccess;
    cudaError_t e = cudaMemsetAsync(ticket, 0, sizeof(unsigned int), st);
    if (e != cudaSuccess) return e;
    // Leave the shared-memory / L1 split to the driver's default for this kernel.  Measured, ms per GiB of text, by
    // preferred carve-out: default 24.2; 0 % (MaxL1) 44.0 and 5 % 29.7 -- every CTA reserves 1 KB of shared memory
    // for the system, so one-warp CTAs run short of it and too few are resident; 15 % 24.2, 30 % 24.3, 50 % 24.5;
    // 75 % 27.7 and 100 % 46.9 -- the L1 gets too small (the same with 7-10 KB of unused dynamic shared memory per
    // CTA, tried as a residency cap for the host entry's concurrent launches: 45.9).  Two or four warps per CTA
    // (a smaller reservation, more L1): 24.2, no change.
    parse_gtab_kernel<<<p.n_jobs < n_tables ? p.n_jobs : n_tables, 32, 0, st>>>(p, reinterpret_cast<uint16_t*>(tables), ticket);
    if (launches) ++*launches;
    return cudaGetLastError();
}

// ---- small buffers: one thread per buffer, parse + emission fused (parse_tiny.cuh) -------------------
bool tiny_takes(uint32_t job_len) {
    static const uint32_t max_len = [] {
        const int v = env_int("SNAPPY_B200_TINY_MAX", (int)tiny::kMaxLen);
        return v < 0 ? 0u : (uint32_t)v > tiny::kMaxLenLimit ? tiny::kMaxLenLimit : (uint32_t)v;
    }();
    return job_len <= max_len;
}

cudaError_t launch_tiny(const JobParams& p, cudaStream_t st, unsigned long long* launches) {
    if (p.n_jobs == 0) return cudaSuccess;
    const uint32_t T = tiny_threads(p.job_len);
    const size_t smem = (size_t)T * tiny::entries_for(p.job_len) * 2u;
    tiny_kernel<<<(p.n_jobs + T - 1) / T, T, smem, st>>>(p);
    if (launches) ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_size(const JobParams& p, cudaStream_t st, unsigned long long* launches) {
    if (p.n_jobs == 0) return cudaSuccess;
    if (p.sparse) size_sparse_kernel<<<p.n_jobs, kEmitThreads, 0, st>>>(p);
    else if (p.job_len > 8192) size_block_kernel<<<p.n_jobs, kSizeThreads, 0, st>>>(p);
    else size_kernel<<<(p.n_jobs + 3) / 4, 128, 0, st>>>(p);
    if (launches) ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_emit(const JobParams& p, cudaStream_t st, unsigned long long* launches) {
    if (p.n_jobs == 0) return cudaSuccess;
    if (p.sparse) emit_sparse_kernel<<<p.n_jobs, kEmitThreads, 0, st>>>(p);
    else emit_kernel<<<p.n_jobs, kEmitThreads, 0, st>>>(p);
    if (launches) ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_decode(const uint8_t* in, uint64_t in_stride, const uint32_t* in_sizes, uint32_t in_size,
                          uint32_t count, uint8_t* out, uint64_t out_stride, uint64_t out_cap,
                          uint32_t* out_lens, unsigned long long* out_len64, cudaStream_t st,
                          unsigned long long* launches) {
    if (count == 0) return cudaSuccess;
    DecodeParams p;
    p.in = in; p.in_stride = in_stride; p.in_sizes = in_sizes; p.in_size = in_size; p.n_jobs = count;
    p.out = out; p.out_stride = out_stride; p.out_cap = out_cap; p.out_lens = out_lens; p.out_len64 = out_len64;
    // page-shaped batches: stage stream and page in shared memory (decode.cuh)
    const uint64_t in_bound = in_sizes ? in_stride : (uint64_t)in_size;
    if (count >= kStageWarps && out_cap <= kStageOutMax && in_bound <= 2 * kStageOutMax) {
        const uint32_t stage_out = (uint32_t)((out_cap + 15) & ~15ull) + 16;
        const uint32_t stage_in = (uint32_t)((in_bound + 15) & ~15ull) + 48;
        const size_t smem = (size_t)kStageWarps * (stage_in + stage_out);
        decode_staged_kernel<<<(count + kStageWarps - 1) / kStageWarps, 32 * kStageWarps, smem, st>>>(p, stage_in, stage_out);
    } else {
        decode_kernel<<<(count + 3) / 4, 128, 0, st>>>(p);
    }
    if (launches) ++*launches;
    return cudaGetLastError();
}

// ---- one large raw stream (decode_par.cuh) ------------------------------------------------------
static uint64_t raw_segment_bytes(uint64_t n) {      // at least kMinSegment (decode_par.cuh), at most 16 K segments
    uint64_t s = kMinSegment;
    while (n / s > 65536) s <<= 1;
    return s;
}
size_t raw_decode_scratch_bytes(uint64_t n, uint64_t out_cap) {
    const uint64_t seg = raw_segment_bytes(n);
    return raw_index_bytes((n + seg - 1) / seg, out_cap / kParBlock + 1);
}
cudaError_t launch_decode_large(const uint8_t* in, uint64_t n, uint8_t* out, uint64_t out_cap, void* scratch,
                                unsigned long long* out_len64, cudaStream_t st, unsigned long long* launches) {
    const uint64_t seg = raw_segment_bytes(n);
    const uint32_t n_seg = (uint32_t)((n + seg - 1) / seg);
    const uint64_t n_blk = out_cap / kParBlock + 1;
    const RawIndex x = raw_index_at(scratch, n_seg, n_blk);
    cudaError_t e = cudaMemsetAsync(scratch, 0, 64, st);                       // ulen, serial flag
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(x.blk_in, 0xff, n_blk * sizeof(unsigned long long), st);
    if (e != cudaSuccess) return e;
    raw_spec_walk_kernel<<<(n_seg + 63) / 64, 64, 0, st>>>(in, n, seg, n_seg, x);
    raw_exits_kernel<<<1, 1024, 0, st>>>(in, n, seg, n_seg, out_cap, x);
    raw_sync_kernel<<<(n_seg + kWalkWarps - 1) / kWalkWarps, 32 * kWalkWarps, 0, st>>>(in, n, seg, n_seg, x);
    raw_chain_kernel<<<1, 1024, 0, st>>>(n, seg, n_seg, x);
    raw_blocks_kernel<<<(n_seg + 63) / 64, 64, 0, st>>>(in, n, seg, n_seg, n_blk, x);
    raw_block_table_kernel<<<(uint32_t)((n_blk + 255) / 256), 256, 0, st>>>(n, n_blk, x);
    DecodeParams p;
    p.in = in; p.in_stride = 0; p.in_sizes = x.blk_len; p.in_size = 0; p.n_jobs = (uint32_t)n_blk;
    p.out = out; p.out_stride = 0; p.out_cap = 0; p.out_lens = x.blk_got; p.out_len64 = nullptr;
    p.in_offs = x.blk_in; p.out_offs = x.blk_out; p.caps = x.blk_cap; p.headerless = 1; p.skip_if = x.serial;
    decode_kernel<<<(uint32_t)((n_blk + 3) / 4), 128, 0, st>>>(p);
    raw_verify_kernel<<<(uint32_t)((n_blk + 255) / 256), 256, 0, st>>>(n_blk, x);
    raw_finish_kernel<<<1, 32, 0, st>>>(x, out_len64);
    DecodeParams q;                                                             // the hardened serial walk, only if flagged
    q.in = in; q.in_stride = 0; q.in_sizes = nullptr; q.in_size = (uint32_t)n; q.n_jobs = 1;
    q.out = out; q.out_stride = 0; q.out_cap = out_cap; q.out_lens = nullptr; q.out_len64 = out_len64;
    q.only_if = x.serial;
    decode_kernel<<<1, 128, 0, st>>>(q);
    if (launches) *launches += 10;
    return cudaGetLastError();
}

// ---- framing ------------------------------------------------------------------------------------
cudaError_t launch_crc32c_chunks(const uint8_t* base, const unsigned long long* offs, const uint32_t* lens,
                                 uint64_t stride, uint64_t n_total, uint32_t fixed_len, uint32_t n_chunks,
                                 uint32_t* out_masked, cudaStream_t st, unsigned long long* launches) {
    if (n_chunks == 0) return cudaSuccess;
    CrcParams p;
    p.base = base; p.offs = offs; p.lens = lens; p.stride = stride; p.n_total = n_total;
    p.fixed_len = fixed_len; p.n_chunks = n_chunks; p.out_masked = out_masked;
    const uint32_t grid = n_chunks < 2 * 148 ? n_chunks : 2 * 148;   // persistent: two CTAs per SM
    crc32c_chunks_kernel<<<grid, kCrcThreads, kCrcSmemBytes, st>>>(p);
    if (launches) ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_frame_bump(uint32_t* sizes, uint32_t n, cudaStream_t st, unsigned long long* launches) {
    if (n == 0) return cudaSuccess;
    frame_bump_sizes_kernel<<<(n + 255) / 256, 256, 0, st>>>(sizes, n);
    if (launches) ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_frame_headers(uint8_t* dst, const unsigned long long* offsets, const uint32_t* sizes,
                                 const uint32_t* crcs, uint32_t n, cudaStream_t st, unsigned long long* launches) {
    frame_headers_kernel<<<(n + 256) / 256, 256, 0, st>>>(dst, offsets, sizes, crcs, n);
    if (launches) ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_frame_walk(const uint8_t* src, uint64_t n, void* index_base, uint64_t max_records,
                              FrameWalkResult* d_result, cudaStream_t st, unsigned long long* launches) {
    const FrameScratch f = frame_scratch(index_base, max_records);
    cudaError_t e = cudaMemsetAsync(d_result, 0, sizeof(FrameWalkResult), st);
    if (e != cudaSuccess) return e;
    frame_hop_kernel<<<1, 32, 0, st>>>(src, n, f.t_pos, max_records, d_result);
    frame_parse_kernel<<<(uint32_t)((max_records + 255) / 256), 256, 0, st>>>(src, n, f, max_records, d_result);
    frame_index_kernel<<<1, 1024, 0, st>>>(f, max_records, d_result);
    if (launches) *launches += 3;
    return cudaGetLastError();
}

cudaError_t launch_frame_fused(const uint8_t* src, uint64_t n, uint8_t* dst, uint64_t dst_cap, void* index_base,
                               uint64_t max_records, FrameWalkResult* d_result, void* d_ctl, cudaStream_t st,
                               unsigned long long* launches) {
    const FrameScratch f = frame_scratch(index_base, max_records);
    cudaError_t e = cudaMemsetAsync(d_result, 0, sizeof(FrameWalkResult), st);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(d_ctl, 0, sizeof(FrameFusedCtl), st);
    if (e != cudaSuccess) return e;
    const uint64_t want = (max_records + 2 + 3) / 4 + 1;                // warps for every record + the two of CTA 0
    const uint32_t grid = (uint32_t)(want < 148ull * SB_DECODE_MINB ? want : 148ull * SB_DECODE_MINB);
    frame_fused_kernel<<<grid, 128, 0, st>>>(src, n, dst, dst_cap, f, max_records, d_result, (FrameFusedCtl*)d_ctl);
    if (launches) ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_frame_decode(const uint8_t* src, uint8_t* dst, void* index_base, uint64_t max_records,
                                uint32_t n_chunks, uint32_t* d_bad, cudaStream_t st,
                                unsigned long long* launches, bool decoded_already) {
    if (n_chunks == 0) return cudaSuccess;
    const FrameScratch f = frame_scratch(index_base, max_records);
    cudaError_t e = cudaSuccess;
    if (!decoded_already) {
        DecodeParams p;
        p.in = src; p.in_stride = 0; p.in_sizes = f.ix.in_len; p.in_size = 0; p.n_jobs = n_chunks;
        p.out = dst; p.out_stride = 0; p.out_cap = 0; p.out_lens = f.got_len; p.out_len64 = nullptr;
        p.in_offs = f.ix.in_off; p.out_offs = f.ix.out_off; p.caps = f.ix.ulen; p.kinds = f.ix.kind;
        decode_kernel<<<(n_chunks + 3) / 4, 128, 0, st>>>(p);
        if (launches) ++*launches;
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    e = launch_crc32c_chunks(dst, f.ix.out_off, f.ix.ulen, 0, 0, 0, n_chunks, f.got_crc, st, launches);
    if (e != cudaSuccess) return e;
    frame_verify_kernel<<<(n_chunks + 255) / 256, 256, 0, st>>>(f.ix, f.got_len, f.got_crc, n_chunks, d_bad);
    if (launches) ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_scan(const uint32_t* sizes, uint32_t n_frag, unsigned long long* offsets,
                        uint32_t hdr_len, uint32_t n_input, uint8_t* dst,
                        unsigned long long* d_total, cudaStream_t st,
                        unsigned long long* launches, const unsigned long long* carry_in,
                        unsigned long long* h_mirror) {
    scan_sizes_kernel<<<1, 1024, 0, st>>>(sizes, n_frag, offsets, hdr_len, n_input, dst, d_total,
                                          carry_in, h_mirror);
    if (launches) ++*launches;
    return cudaGetLastError();
}

}  // namespace sb
