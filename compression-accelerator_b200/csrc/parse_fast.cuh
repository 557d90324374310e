// parse_fast.cuh -- SNAPPY_B200_FAST: a parallel parse of one fragment by a whole CTA.
//
// The exact parse (parse_pipe.cuh) is sequential because Snappy inserts a position into its hash
// table only if the parse probes it (tests/snappy/snappy.cc:587-659).  The fast mode drops that
// one rule -- EVERY position is looked up and inserted, window by window -- and with it the
// dependence of the table on the parse.  What is left splits into three roles that run as a
// pipeline over the fragment's 32-position windows, the fragment staged in shared memory by TMA:
//
//   table warp (1)      walks the windows in order: hash (snappy.cc:91-97), table look-up, table
//                       insert; leaves cand(p) = most recent position with p's hash in an EARLIER
//                       window in a small ring (a window's own positions become candidates only
//                       for later windows; within a window the highest position of a slot stays).
//   evaluator warps (K) take windows round-robin: 4-byte compare and match extension against
//                       cand(p) (snappy-internal.h:95-143; 16-byte compares per lane up to 64 bytes,
//                       warp-cooperative beyond), then -- by pointer jumping over the lanes -- for
//                       every possible entry lane of the window where the greedy parse leaves it.
//   chaser warp (1)     the only sequential step, kept to a dozen instructions per window (a lone
//                       warp issues one dependent instruction every ~10 cycles): one shuffle maps
//                       the entry lane to the next window's entry lane; it records every window's
//                       entry lane, and the window's evaluator then writes the window's copies.
//
// Copies go to a position-indexed ("sparse") list: a copy starting at position p owns slot p / 4
// (copies are >= 4 bytes long, so slots never collide and stay in stream order); every window's
// evaluator writes the window's 8 slots, empty ones as zero.  Literal lengths are the gaps between
// consecutive copies and are recovered by the size / emit kernels (emit.cuh), so the sequential
// step carries neither a sequence count nor a literal start.
//
// The stream is a valid Snappy stream, deterministic, and defined by the sequential model
// port_fast_compress kept with the test infrastructure, against which tests pin it byte for byte; its size
// stays within 2 % of the reference's on the reference's corpora (it is smaller on text, since more
// positions are candidates).  Copies are capped at kFastCap bytes per match.
#pragma once
#include "kernels.h"
#include "snappy_device.cuh"
#include "parse_window.cuh"
#include <stdio.h>

namespace sb {

constexpr uint32_t kFastCap = 1024;        // longest match reported (model: min(prefix, 1024))
constexpr uint32_t kFastLane = 64;         // per-lane extension limit; longer matches go cooperative
constexpr uint32_t kCandRing = 64;         // windows the table warp may run ahead of the chaser
constexpr uint32_t kRecRing = 64;          // record slots allocated; 2K are used: evaluator e owns slots e and e + K,
                                           // so a late record of a window the chaser jumped over can never land on
                                           // another warp's record
#ifndef SB_FAST_K
#define SB_FAST_K 14
#endif
constexpr uint32_t kFastMaxK = SB_FAST_K;  // evaluator warps of the stream path (+ table warp + chaser warp)
constexpr uint32_t kNoCand = 0xffffu;
constexpr uint32_t kFastPad = 96;          // zeroed slack after the input: 16-byte reads up to pos+80
constexpr uint32_t kEntryNone = 0xffu;     // entry[] : chaser has not reached the window
constexpr uint32_t kEntrySkip = 0xfeu;     // entry[] : window lies wholly inside a copy

struct FastLayout {                        // byte offsets into dynamic shared memory
    uint32_t table, cring, ering, entry, input, total;
};
__host__ __device__ inline FastLayout fast_layout(uint32_t job_len) {
    FastLayout l;
    l.table = 128;
    l.cring = l.table + table_entries(job_len) * 2;
    l.ering = l.cring + kCandRing * 32 * 2;
    l.entry = l.ering + kRecRing * 32 * 4;
    l.input = (l.entry + ((job_len + 31u) >> 5) + 15u) & ~15u;
    l.total = l.input + ((job_len + 15u) & ~15u) + kFastPad;
    return l;
}
__host__ __device__ inline uint32_t fast_evaluators(uint32_t job_len) {
    return job_len > 16384 ? kFastMaxK : job_len > 2048 ? 6u : 2u;
}

struct FastCtl {                           // first 128 bytes of shared memory
    uint64_t bar;                          // mbarrier of the TMA load
    volatile uint32_t t_done;              // windows whose candidates are in the ring
    volatile uint32_t abort;               // a watchdog expired: every wait loop leaves, the fragment becomes one literal
    unsigned int* err;                     // the context's error word (mapped host memory) or null
};
// The control block sits at the start of the kernel's dynamic shared memory: reachable from any role.
__device__ __forceinline__ FastCtl* fast_ctl() {
    extern __shared__ __align__(128) uint8_t fast_smem_base[];
    return reinterpret_cast<FastCtl*>(fast_smem_base);
}

// sparse sequence slot: offset << 16 | (len - 4) << 2 | (position & 3); 0 = no copy starts here
__host__ __device__ inline uint32_t sparse_pack(uint32_t pos, uint32_t len, uint32_t off) {
    return (off << 16) | ((len - 4) << 2) | (pos & 3u);
}

#ifdef SNAPPY_B200_STATS
// cycle accumulators live in registers (fs_acc[]) and reach g_stats once per role and fragment
#define FSTAT_T0() const long long fs_t0 = clock64()
#define FSTAT_ADD(i, lane) do { fs_acc[(i) & 3] += (unsigned long long)(clock64() - fs_t0); } while (0)
#define FSTAT_DECL() unsigned long long fs_acc[4] = {0, 0, 0, 0}
#define FSTAT_PARAM , unsigned long long* fs_acc
#define FSTAT_ARG , fs_acc
#define FSTAT_FLUSH(slot, i, lane) do { if ((lane) == 0) atomicAdd(&g_stats[i], fs_acc[(slot) & 3]); } while (0)
#else
#define FSTAT_DECL() do {} while (0)
#define FSTAT_PARAM
#define FSTAT_ARG
#define FSTAT_FLUSH(slot, i, lane) do {} while (0)
#define FSTAT_T0() do {} while (0)
#define FSTAT_ADD(i, lane) do {} while (0)
#endif
#ifndef SB_FAST_TGROUP
#define SB_FAST_TGROUP 8
#endif
#ifndef SB_FAST_SLEEP
#define SB_FAST_SLEEP 32
#endif
__device__ __forceinline__ void fast_backoff() {
#if SB_FAST_SLEEP > 0
    __nanosleep(SB_FAST_SLEEP);          // waiting warps must not eat the issue slots of the sequential roles
#endif
}
// Inside a wait loop: leaves the loop (break) when a watchdog has expired anywhere in the CTA, or expires
// itself -- raising the context's error word (the call then fails with SNAPPY_B200_CUDA_ERROR) instead of
// trapping, which would poison the whole CUDA context.  The roles then run to their end on whatever they
// hold, and the kernel turns the fragment into one literal (parse_fast_kernel's last lines).
#define FAST_WATCHDOG(spins, what, a, b)                                                          \
    if ((spins) > (1u << 24) || fast_ctl()->abort) {                                              \
        FastCtl* wc_ = fast_ctl();                                                                \
        if (!wc_->abort) { wc_->abort = 1u; if (wc_->err) *(volatile unsigned int*)wc_->err = 1u; } \
        break;                                                                                    \
    }

// common prefix of s[a..] and s[b..] (b < end), at most `cap`, 128 bytes per round
__device__ __forceinline__ uint32_t coop_match_capped(const uint8_t* s, uint32_t a, uint32_t b,
                                                      uint32_t end, uint32_t cap, int lane) {
    uint32_t n = 0;
    while (n < cap) {
        const uint32_t pb = b + n + (lane << 2);
        uint32_t m = 0;
        if (pb < end) {
            const uint32_t x = lds32u(s, a + n + (lane << 2)) ^ lds32u(s, pb);
            m = x ? (uint32_t)(__ffs(x) - 1) >> 3 : 4u;
            m = min(m, end - pb);
        }
        const uint32_t bal = __ballot_sync(kFull, m < 4);
        if (bal) {
            const int first = __ffs(bal) - 1;
            n += (first << 2) + __shfl_sync(kFull, m, first);
            break;
        }
        n += 128;
    }
    return min(n, cap);
}

__device__ __forceinline__ uint32_t prefix16(uint32_t x0, uint32_t x1, uint32_t x2, uint32_t x3) {
    uint32_t m = 16;
    if (x3) m = 12 + ((uint32_t)(__ffs(x3) - 1) >> 3);
    if (x2) m = 8 + ((uint32_t)(__ffs(x2) - 1) >> 3);
    if (x1) m = 4 + ((uint32_t)(__ffs(x1) - 1) >> 3);
    if (x0) m = (uint32_t)(__ffs(x0) - 1) >> 3;
    return m;
}

// ---- table warp ---------------------------------------------------------------------------------
// One window of the table walk.  Lanes run DOWN the window: when several lanes of one store
// instruction hit the same shared-memory address, B200 keeps the LOWEST lane's value (measured:
// scripts/probes/sts_conflict.cu; CUDA leaves it unspecified), so with this mapping the highest
// position of a slot stays, as in the model.  Any other arbitration would still give a valid stream,
// only not the model's bytes -- the parity tests would say so.
__device__ __forceinline__ void fast_table_window(const uint8_t* s, uint32_t len, uint32_t w,
                                                  uint16_t* table, int shift, uint16_t* cring, uint32_t rl) {
    const uint32_t pos = (w << 5) + rl;
    const bool valid = pos + 4 <= len;
    const uint32_t h = hash32(lds32u(s, pos), shift);
    uint32_t T = kNoCand;
    if (valid) {
        T = table[h];                         // state after the previous window: always < this window
        table[h] = (uint16_t)pos;
    }
    cring[((w & (kCandRing - 1)) << 5) + rl] = (uint16_t)(pos ? T : kNoCand);
}

__device__ __forceinline__ void fast_table_warp(const uint8_t* s, uint32_t len, uint32_t nwin,
                                                uint16_t* table, int shift, uint16_t* cring,
                                                const volatile uint8_t* entry, FastCtl* ctl,
                                                int lane FSTAT_PARAM) {
    const uint32_t rl = 31u - lane;
    for (uint32_t w0 = 0; w0 < nwin; w0 += SB_FAST_TGROUP) {
        const uint32_t w1 = min(w0 + SB_FAST_TGROUP, nwin);
        if (w1 > kCandRing) {                 // the slot of window w1-1 still belongs to window w1-1-64
            FSTAT_T0();
            for (uint32_t spins = 0; entry[w1 - 1 - kCandRing] == kEntryNone; ++spins) {
                fast_backoff();
                FAST_WATCHDOG(spins, "table warp waits for the chaser", w1, ctl->t_done);
            }
            FSTAT_ADD(1, lane);
        }
        if (w1 - w0 == SB_FAST_TGROUP) {
            // A lone warp lives on instruction-level parallelism, and the compiler may not move one
            // window's loads above another window's stores (all shared memory), so the group is
            // staged by hand: every window's word and hash first, then the table accesses in window
            // order (a store does not wait for the look-up before it), then the candidates.
            uint32_t h[SB_FAST_TGROUP], T[SB_FAST_TGROUP];
#pragma unroll
            for (uint32_t k = 0; k < SB_FAST_TGROUP; ++k) h[k] = hash32(lds32u(s, ((w0 + k) << 5) + rl), shift);
#pragma unroll
            for (uint32_t k = 0; k < SB_FAST_TGROUP; ++k) {
                const uint32_t pos = ((w0 + k) << 5) + rl;
                T[k] = kNoCand;
                if (pos + 4 <= len) {
                    T[k] = table[h[k]];
                    table[h[k]] = (uint16_t)pos;
                }
            }
#pragma unroll
            for (uint32_t k = 0; k < SB_FAST_TGROUP; ++k) {
                const uint32_t w = w0 + k;
                cring[((w & (kCandRing - 1)) << 5) + rl] = (uint16_t)((w | rl) ? T[k] : kNoCand);
            }
        } else {
            for (uint32_t w = w0; w < w1; ++w) fast_table_window(s, len, w, table, shift, cring, rl);
        }
        __threadfence_block();
        __syncwarp();
        if (lane == 0) ctl->t_done = w1;
    }
}

// ---- evaluator ------------------------------------------------------------------------------------
struct FastWin {                           // what an evaluator keeps of a window until it emits it
    uint32_t w, L, imm, mo;                // imm = next taken lane | end lane << 6 ; mo = len | offset << 16
    bool live;
};

__device__ __forceinline__ void fast_evaluate(const uint8_t* s, uint32_t len, const uint16_t* cring,
                                              uint32_t* ering, const volatile uint8_t* entry,
                                              FastCtl* ctl, uint32_t w, uint32_t slot, FastWin& st,
                                              int lane FSTAT_PARAM) {
    // (the record slot is free: this warp has already emitted the window that used it, w - 2K)
    FSTAT_T0();
#ifndef SB_FAST_TSLEEP
#define SB_FAST_TSLEEP 40
#endif
    for (uint32_t spins = 0; ctl->t_done <= w; ++spins) {
#if SB_FAST_TSLEEP > 0
        __nanosleep(SB_FAST_TSLEEP);       // hot spinning here took 60 % of the SM's issue slots (ncu), mostly from the table warp
#endif
        FAST_WATCHDOG(spins, "evaluator waits for candidates", w, ctl->t_done);
    }
    __syncwarp();
    FSTAT_ADD(2, lane);
    const uint32_t pos = (w << 5) + lane;
    uint32_t c = cring[((w & (kCandRing - 1)) << 5) + lane];
    // A window that lies wholly inside a copy is marked by the chaser without its record, so the table warp may
    // already have reused this ring slot for window w + 64 when its evaluator gets here: the candidates are then
    // positions AHEAD of the window.  Nothing of such a window is emitted, but on periodic input a candidate from
    // the future matches for kilobytes and the extension would read past the staged fragment (found by
    // scripts/fuzz_decode.py: period 64 / 128 / 256, a full 64 KiB fragment).  A real candidate is always earlier.
    if (c >= pos) c = kNoCand;
    uint32_t m = 0;
    uint32_t d0, d1, d2, d3, c0, c1, c2, c3;
    lds16u(s, pos, d0, d1, d2, d3);
    // a run (the 4 bytes one position back are the same 4 bytes) beats the table candidate: it is
    // the one kind of match a window cannot find among earlier windows' inserts
    if (c != kNoCand && lds32u(s, pos - 1) == d0) c = pos - 1;
    if (c != kNoCand) {
        lds16u(s, c, c0, c1, c2, c3);
        m = prefix16(c0 ^ d0, c1 ^ d1, c2 ^ d2, c3 ^ d3);
        for (uint32_t o = 16; m == o && o < kFastLane; o += 16) {
            lds16u(s, pos + o, d0, d1, d2, d3);
            lds16u(s, c + o, c0, c1, c2, c3);
            m = o + prefix16(c0 ^ d0, c1 ^ d1, c2 ^ d2, c3 ^ d3);
        }
        m = min(m, len - pos);
    }
    const uint32_t off = pos - c;
    const bool is_long = m >= kFastLane;
    const uint32_t longm = __ballot_sync(kFull, is_long);
    if (longm) {
        // runs of consecutive long lanes with one offset share one match: extend the run's head
        // cooperatively, the members follow from it (prefix(head) - distance to the head)
        const uint32_t off_prev = __shfl_up_sync(kFull, off, 1);
        const bool head = is_long && !(lane > 0 && ((longm >> (lane - 1)) & 1u) && off_prev == off);
        const uint32_t heads = __ballot_sync(kFull, head);
        const uint32_t rid = __popc(heads & ((2u << lane) - 1u));
        uint32_t hm = heads, r = 0;
        while (hm) {
            const int hl = __ffs(hm) - 1;
            hm &= hm - 1;
            ++r;
            const uint32_t a = __shfl_sync(kFull, c, hl) + kFastLane;
            const uint32_t b = __shfl_sync(kFull, pos, hl) + kFastLane;
            const uint32_t n = coop_match_capped(s, a, b, len, kFastCap + 32, lane);
            if (is_long && rid == r) m = kFastLane + n - (lane - hl);
        }
        m = min(m, kFastCap);
    }
    const bool hit = m >= 4;
    const uint32_t L = __ballot_sync(kFull, hit);
    const uint32_t ek = lane + m;                                   // end lane of my copy
    uint32_t nxt = 32;
    if (hit && ek < 32) {
        const uint32_t mm = L & (kFull << ek);
        if (mm) nxt = __ffs(mm) - 1;
    }
    st.w = w; st.L = L; st.live = true;
    st.imm = nxt | (ek << 6);
    st.mo = m | (off << 16);
    // pointer jumping (a chain has at most 8 copies): pk = next lane (32 = end of chain) | end lane of
    // the chain's last copy << 6
    uint32_t pk = nxt | (ek << 6);
#pragma unroll
    for (int round = 0; round < 3; ++round) {
        const uint32_t q = __shfl_sync(kFull, pk, pk & 31u);
        if ((pk & 63u) < 32u) pk = q;
    }
    // record of entry lane `lane`: where the parse enters the following windows
    const uint32_t me = L & (kFull << lane);
    const uint32_t q = __shfl_sync(kFull, pk, me ? __ffs(me) - 1 : 0);
    const uint32_t E = me ? (q >> 6) : 0u;                          // end of the last copy taken (0: none)
    const uint32_t rec = E >= 32 ? (E & 31u) | (((E - 32) >> 5) << 5) : 0u;   // next entry lane | windows skipped << 5
    reinterpret_cast<volatile uint32_t*>(ering)[(slot << 5) + lane] = ((w + 1) << 16) | rec;
    FSTAT_ADD(3, lane);      // wait + compute
}

// The window's 8 sparse slots: the copies the parse takes when it enters the window at lane e.
__device__ __forceinline__ void fast_emit(const volatile uint8_t* entry, uint32_t* slots,
                                          const FastWin& st, int lane FSTAT_PARAM) {
    uint32_t e;
    FSTAT_T0();
    for (uint32_t spins = 0;; ++spins) {
        e = entry[st.w];
        if (e != kEntryNone) break;
        fast_backoff();
        FAST_WATCHDOG(spins, "evaluator waits for its entry lane", st.w, 0);
    }
    __syncwarp();
    FSTAT_ADD(0, lane);
    uint32_t chosen = 0;
    if (e != kEntrySkip) {
        const uint32_t me = st.L & (kFull << e);
        uint32_t f = me ? __ffs(me) - 1 : 32;
        while (f < 32) {                                          // one shuffle per copy
            chosen |= 1u << f;
            f = __shfl_sync(kFull, st.imm, f) & 63u;
        }
    }
    // slot j (lane j < 8) holds the copy that starts in lanes 4j .. 4j+3, if any
    const uint32_t mine = sparse_pack(lane, st.mo & 0xffffu, st.mo >> 16);
    const uint32_t grp = (chosen >> ((lane & 7u) << 2)) & 15u;
    const uint32_t v = __shfl_sync(kFull, mine, ((lane & 7u) << 2) + (grp ? __ffs(grp) - 1 : 0));
    if (lane < 8) slots[(st.w << 3) + lane] = grp ? v : 0u;
    FSTAT_ADD(1, lane);
}

// ---- chaser ---------------------------------------------------------------------------------------
// One window at a time; returns when `stop` is reached (or the fragment ends).
__device__ __forceinline__ void fast_chase_steps(const volatile uint32_t* er, volatile uint8_t* entry,
                                                 uint32_t nwin, uint32_t R, uint32_t stop, uint32_t& w,
                                                 uint32_t& e, uint32_t& slot, int lane FSTAT_PARAM) {
    while (w < nwin && w < stop) {
        uint32_t v;
        for (uint32_t spins = 0;; ++spins) {
            v = __shfl_sync(kFull, er[(slot << 5) + lane], e);
            if ((v >> 16) == w + 1) break;                         // else: record not published yet
            FAST_WATCHDOG(spins, "chaser waits for a record", w, e);
        }
        entry[w] = (uint8_t)e;                                     // all lanes store the same byte
        ++w;
        slot = slot + 1 == R ? 0 : slot + 1;
        e = v & 31u;
        const uint32_t adv = (v >> 5) & 63u;
        if (adv) {                                                 // windows wholly inside the copy
            const uint32_t w2 = min(w + adv, nwin);
            for (uint32_t x = w + lane; x < w2; x += 32) entry[x] = (uint8_t)kEntrySkip;
            slot = (slot + (w2 - w)) % R;
            w = w2;
        }
    }
}

// The sequential step, four windows per round: the rows of the next four windows are read at once,
// the entry lane is pushed through them with four dependent shuffles, and only then are the four
// records validated together (all published, no copy jumping over a whole window).  A round that
// does not validate is redone window by window.  Nothing but the shuffles sits on the chain.
__device__ __forceinline__ void fast_chaser(const uint32_t* ering, volatile uint8_t* entry,
                                            uint32_t nwin, uint32_t R, int lane FSTAT_PARAM) {
    const volatile uint32_t* er = reinterpret_cast<const volatile uint32_t*>(ering);
    uint32_t w = 0, e = 0, slot = 0;                               // slot = w mod R
    while (w + 4 <= nwin) {
        const uint32_t s1 = slot + 1 >= R ? slot + 1 - R : slot + 1;
        const uint32_t s2 = slot + 2 >= R ? slot + 2 - R : slot + 2;
        const uint32_t s3 = slot + 3 >= R ? slot + 3 - R : slot + 3;
        const uint32_t r0 = er[(slot << 5) + lane], r1 = er[(s1 << 5) + lane];
        const uint32_t r2 = er[(s2 << 5) + lane], r3 = er[(s3 << 5) + lane];
        const uint32_t v0 = __shfl_sync(kFull, r0, e);
        const uint32_t v1 = __shfl_sync(kFull, r1, v0);            // shuffles use the index modulo 32
        const uint32_t v2 = __shfl_sync(kFull, r2, v1);
        const uint32_t v3 = __shfl_sync(kFull, r3, v2);
        // tags are w+1 .. w+4 in bits 16.., skips in bits 5..10: one compare per record
        const uint32_t t = (w + 1) << 16;
        const bool ok = ((v0 ^ t) < 32u) & ((v1 ^ (t + 0x10000u)) < 32u) & ((v2 ^ (t + 0x20000u)) < 32u) &
                        ((v3 ^ (t + 0x30000u)) < 32u);
#ifdef SNAPPY_B200_STATS
        fs_acc[3] += 1;
#endif
        if (ok) {
            const uint32_t packed = e | ((v0 & 31u) << 8) | ((v1 & 31u) << 16) | ((v2 & 31u) << 24);
            *reinterpret_cast<volatile uint32_t*>(entry + w) = packed;     // w is a multiple of 4 here
            e = v3 & 31u;
            w += 4;
            slot = slot + 4 >= R ? slot + 4 - R : slot + 4;
        } else {
            FSTAT_T0();
            fast_chase_steps(er, entry, nwin, R, w + 4, w, e, slot, lane FSTAT_ARG);
            // realign to a multiple of four windows (a jump may land anywhere): the packed store needs it
            while ((w & 3u) && w < nwin) fast_chase_steps(er, entry, nwin, R, (w + 3) & ~3u, w, e, slot, lane FSTAT_ARG);
            FSTAT_ADD(2, lane);
        }
    }
    fast_chase_steps(er, entry, nwin, R, nwin, w, e, slot, lane FSTAT_ARG);
}

}  // namespace sb
