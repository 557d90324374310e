// parse_fast.cuh -- SNAPPY_B200_FAST: a parallel parse of one fragment by a whole CTA.
//
// The exact parse (parse_pipe.cuh) is sequential because Snappy inserts a position into its hash
// table only if the parse probes it (tests/snappy/snappy.cc:587-659).  The fast mode drops that
// one rule -- EVERY position is looked up and inserted, in position order -- and with it the
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
//                       every possible entry lane of the window the end of the last copy and the
//                       number of copies the greedy parse takes from there.
//   chaser warp (1)     the only sequential step: one shared-memory look-up per window maps the
//                       entry lane to the next window's entry lane; it hands every window its
//                       entry, sequence index and literal start, and the window's evaluator
//                       writes the sequences (one shuffle per copy).
//
// The stream is a valid Snappy stream, deterministic, and defined by the sequential model
// port_fast_compress (oracle/snappy_port.c), against which tests pin it byte for byte; its size
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
constexpr uint32_t kFastMaxK = 14;         // evaluator warps per CTA (stream path)
constexpr uint32_t kNoCand = 0xffffu;
// Answers of the chaser live in a ring of 64.  Chains of long copies let the chaser run arbitrarily
// far ahead of a slow evaluator (it needs only one evaluated window per copy), so every evaluator
// publishes how far it has emitted (FastCtl::done) and the chaser never writes the answer of window
// v before every window <= v - 64 has been emitted; it re-reads the progress words only when it
// reaches that bound.
constexpr uint32_t kAnsRing = 64;
constexpr uint32_t kFastPad = 96;          // zeroed slack after the input: 16-byte reads up to pos+80

struct FastLayout {                        // byte offsets into dynamic shared memory
    uint32_t table, cring, ering, ans, input, total;
};
__host__ __device__ inline FastLayout fast_layout(uint32_t job_len, uint32_t K) {
    FastLayout l;
    l.table = 128;
    l.cring = l.table + table_entries(job_len) * 2;
    l.ering = l.cring + kCandRing * 32 * 2;
    l.ans = l.ering + 2 * K * 32 * 4;
    l.input = (l.ans + kAnsRing * 8 + 15u) & ~15u;
    l.total = l.input + ((job_len + 15u) & ~15u) + kFastPad;
    return l;
}
__host__ __device__ inline uint32_t fast_evaluators(uint32_t job_len) {
    return job_len > 16384 ? kFastMaxK : job_len > 2048 ? 6u : 2u;
}

struct FastCtl {                           // first 128 bytes of shared memory
    uint64_t bar;                          // mbarrier of the TMA load
    volatile uint32_t t_done;              // windows whose candidates are in the ring
    volatile uint32_t c_win;               // chaser progress (flow control only, may lag)
    volatile uint32_t done[16];            // per evaluator: all its windows below this index are emitted
};

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

__device__ __forceinline__ void fast_table_warp(const uint8_t* s, uint32_t len, uint32_t nwin,
                                                uint16_t* table, int shift, uint16_t* cring,
                                                FastCtl* ctl, int lane FSTAT_PARAM) {
    for (uint32_t w0 = 0; w0 < nwin; w0 += SB_FAST_TGROUP) {
        const uint32_t w1 = min(w0 + SB_FAST_TGROUP, nwin);
        if (w1 > kCandRing) {
            FSTAT_T0();
            for (uint32_t spins = 0; w1 > ctl->c_win + kCandRing; ++spins) {
                fast_backoff();
                if (spins > (1u << 24)) { if (lane == 0) printf("fast: table warp stuck blk %d w1 %u c_win %u t_done %u\n", blockIdx.x, w1, ctl->c_win, ctl->t_done); __trap(); }
            }
            FSTAT_ADD(1, lane);
        }
        for (uint32_t w = w0; w < w1; ++w) {
            // Lanes run DOWN the window.  When several lanes of one store instruction hit the same
            // shared-memory address, B200 keeps the LOWEST lane's value (measured:
            // scripts/probes/sts_conflict.cu; CUDA leaves it unspecified), so with this mapping the
            // highest position of a slot stays, as in the model.  Any other arbitration would still
            // give a valid stream, only not the model's bytes -- the parity tests would say so.
            const uint32_t rl = 31u - lane, pos = (w << 5) + rl;
            const bool valid = pos + 4 <= len;
            const uint32_t h = hash32(lds32u(s, pos), shift);
            uint32_t T = kNoCand;
            if (valid) {
                T = table[h];                 // state after the previous window: always < this window
                table[h] = (uint16_t)pos;
            }
            cring[((w & (kCandRing - 1)) << 5) + rl] = (uint16_t)(pos ? T : kNoCand);
        }
        __threadfence_block();
        __syncwarp();
        if (lane == 0) ctl->t_done = w1;
    }
}

// ---- evaluator ------------------------------------------------------------------------------------
struct FastWin {                           // what an evaluator keeps of a window until it emits it
    uint32_t w, slot, L, imm, mo;          // imm = next taken lane | end lane << 6 ; mo = len | offset << 16
    bool live;
};

__device__ __forceinline__ void fast_evaluate(const uint8_t* s, uint32_t len, const uint16_t* cring,
                                              uint32_t* ering, FastCtl* ctl, uint32_t w, uint32_t slot,
                                              FastWin& st, int lane FSTAT_PARAM) {
    FSTAT_T0();
    for (uint32_t spins = 0; ctl->t_done <= w; ++spins) {
        if (spins > (1u << 24)) { if (lane == 0) printf("fast: evaluator stuck on t_done blk %d w %u c_win %u t_done %u\n", blockIdx.x, w, ctl->c_win, ctl->t_done); __trap(); }
    }
    __syncwarp();
    FSTAT_ADD(2, lane);
    const uint32_t pos = (w << 5) + lane;
    uint32_t c = cring[((w & (kCandRing - 1)) << 5) + lane];
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
    st.w = w; st.slot = slot; st.L = L; st.live = true;
    st.imm = nxt | (ek << 6);
    st.mo = m | (off << 16);
    // pointer jumping: pk = next lane (32 = end of chain) | end of the chain's last copy << 6 | copies << 17
    uint32_t pk = nxt | (ek << 6) | (1u << 17);
#pragma unroll
    for (int round = 0; round < 3; ++round) {
        const uint32_t q = __shfl_sync(kFull, pk, pk & 31u);
        if ((pk & 63u) < 32u) pk = (q & 0x1ffffu) + ((pk >> 17) + (q >> 17) << 17);
    }
    const uint32_t me = L & (kFull << lane);
    const uint32_t q = __shfl_sync(kFull, pk, me ? __ffs(me) - 1 : 0);
    const uint32_t rec = me ? ((q >> 6) & 0x7ffu) | ((q >> 17) << 11) : 0u;
    reinterpret_cast<volatile uint32_t*>(ering)[(slot << 5) + lane] = ((w + 1) << 16) | rec;
    FSTAT_ADD(3, lane);      // wait + compute
}

// answers of the chaser: tag (window + 1) : 12 | skipped : 1 | entry lane : 5 | sequence index : 15 | literal start : 17
__device__ __forceinline__ unsigned long long fast_ans_pack(uint32_t w, uint32_t skipped, uint32_t e,
                                                            uint32_t sb, uint32_t ne) {
    return (unsigned long long)((w + 1) & 0xfffu) | ((unsigned long long)skipped << 12) |
           ((unsigned long long)e << 13) | ((unsigned long long)sb << 18) | ((unsigned long long)ne << 33);
}

__device__ __forceinline__ void fast_emit(const volatile unsigned long long* ans, volatile uint32_t* done,
                                          seq_t* seqs, const FastWin& st, int lane FSTAT_PARAM) {
    unsigned long long a;
    FSTAT_T0();
    for (uint32_t spins = 0;; ++spins) {
        a = ans[st.w & (kAnsRing - 1)];
        if ((uint32_t)(a & 0xfffu) == ((st.w + 1) & 0xfffu)) break;
        fast_backoff();
        if (spins > (1u << 24)) { if (lane == 0) printf("fast: evaluator stuck on answer blk %d w %u ans %llx\n", blockIdx.x, st.w, a); __trap(); }
    }
    __syncwarp();
    FSTAT_ADD(4, lane);
    if (lane == 0) *done = st.w + 1;                              // the answer is read: its slot may be reused
    if ((a >> 12) & 1u) return;                                   // window lies inside a copy
    const uint32_t e = (uint32_t)(a >> 13) & 31u;
    uint32_t idx = (uint32_t)(a >> 18) & 0x7fffu, cur = (uint32_t)(a >> 33) & 0x1ffffu;
    const uint32_t wbase = st.w << 5, pos = wbase + lane;
    const uint32_t me = st.L & (kFull << e);
    uint32_t f = me ? __ffs(me) - 1 : 32;
    while (f < 32) {                                              // one shuffle per copy
        if ((uint32_t)lane == f) seqs[idx] = seq_pack(pos - cur, st.mo & 0xffffu, st.mo >> 16);
        const uint32_t t = __shfl_sync(kFull, st.imm, f);
        cur = wbase + (t >> 6);
        f = t & 63u;
        ++idx;
    }
    FSTAT_ADD(5, lane);
}

// ---- chaser ---------------------------------------------------------------------------------------
// Windows whose answers may be written: v < (lowest emitted-progress word) + kAnsRing.
__device__ __forceinline__ uint32_t fast_safe_bound(FastCtl* ctl, uint32_t K, uint32_t w, int lane) {
    for (uint32_t spins = 0;; ++spins) {
        const uint32_t d = (uint32_t)lane < K ? ctl->done[lane] : 0xffffffffu;
        const uint32_t m = __reduce_min_sync(kFull, d);
        if (m > 0xffffffffu - kAnsRing) return 0xffffffffu;
        if (w < m + kAnsRing) return m + kAnsRing;
        if (spins > (1u << 26)) { if (lane == 0) printf("fast: chaser waits for emitters blk %d w %u min %u\n", blockIdx.x, w, m); __trap(); }
    }
}

__device__ __forceinline__ uint32_t fast_chaser(const uint32_t* ering, volatile unsigned long long* ans,
                                                FastCtl* ctl, uint32_t nwin, uint32_t K, uint32_t len,
                                                seq_t* seqs, int lane FSTAT_PARAM) {
    const volatile uint32_t* er = reinterpret_cast<const volatile uint32_t*>(ering);
    const uint32_t R = 2 * K;
    uint32_t w = 0, e = 0, ne = 0, sb = 0, slot = 0, safe = kAnsRing;
    uint32_t row = nwin ? er[lane] : 0u;                           // lane l holds the record of entry lane l
    while (w < nwin) {
#ifdef SNAPPY_B200_STATS
        const long long tc0 = clock64();
#endif
        uint32_t v = __shfl_sync(kFull, row, e);
#ifdef SNAPPY_B200_STATS
        const long long tc1 = clock64(); fs_acc[0] += (unsigned long long)(tc1 - tc0);
#endif
        if ((v >> 16) != w + 1) {                                  // record not published yet
            FSTAT_T0();
            for (uint32_t spins = 0;; ++spins) {
                row = er[(slot << 5) + lane];
                v = __shfl_sync(kFull, row, e);
                if ((v >> 16) == w + 1) break;
                if (spins > (1u << 26)) { if (lane == 0) printf("fast: chaser stuck blk %d w %u e %u v %x t_done %u\n", blockIdx.x, w, e, v, ctl->t_done); __trap(); }
            }
            FSTAT_ADD(6, lane);
        }
#ifdef SNAPPY_B200_STATS
        fs_acc[3] += 1;
#endif
        const uint32_t nslot = slot + 1 == R ? 0 : slot + 1;
        const uint32_t row_next = er[(nslot << 5) + lane];         // prefetch: usually ready long before
        const uint32_t E = v & 0x7ffu, cnt = (v >> 11) & 15u;      // end of the last copy, copies taken
        if (w >= safe) safe = fast_safe_bound(ctl, K, w, lane);
#ifdef SNAPPY_B200_STATS
        const long long tc2 = clock64();
#endif
        if (lane == 0) ans[w & (kAnsRing - 1)] = fast_ans_pack(w, 0, e, sb, ne);
#ifdef SNAPPY_B200_STATS
        const long long tc3 = clock64(); fs_acc[1] += (unsigned long long)(tc3 - tc2);
#endif
        uint32_t adv = 0;
        e = 0;
        if (cnt) {
            ne = (w << 5) + E;
            sb += cnt;
            if (E >= 32) { adv = (E - 32) >> 5; e = E & 31u; }
        }
        ++w; slot = nslot; row = row_next;
        const bool adv_taken = adv != 0;
        if (adv) {
            for (; adv && w < nwin; --adv) {                       // windows wholly inside the copy
                if (w >= safe) safe = fast_safe_bound(ctl, K, w, lane);
                if (lane == 0) ans[w & (kAnsRing - 1)] = fast_ans_pack(w, 1, 0, 0, 0);
                ++w; slot = slot + 1 == R ? 0 : slot + 1;
            }
            if (w < nwin) row = er[(slot << 5) + lane];
        }
#ifndef SB_FAST_CWIN_MASK
#define SB_FAST_CWIN_MASK 7u
#endif
        if (lane == 0 && ((w & SB_FAST_CWIN_MASK) == 0 || adv_taken)) ctl->c_win = w;   // flow control only: may lag
    }
    if (lane == 0) ctl->c_win = w;
    if (ne < len) {                                                // trailing literal (snappy.cc:665-668)
        if (lane == 0) seqs[sb] = seq_pack(len - ne, 0, 0);
        ++sb;
    }
    return sb;
}

}  // namespace sb
