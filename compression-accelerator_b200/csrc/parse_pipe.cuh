// parse_pipe.cuh -- the window parser of parse_window.cuh split over cooperating warps.
//
// A single warp per fragment is latency-bound: every instruction of the per-window set-up sits on
// the parse's critical path.  Here P "producer" warps run ahead of one "resolver" warp.  A
// producer does, for window w, the whole parallel set-up of parse_window.cuh (word, hash, table
// look-up, 4-byte compare, 12-byte pre-extension, peer detection, per-lane next-copy values) and
// leaves the result in a small shared-memory ring.  Its table look-ups are SPECULATIVE: inserts of
// windows the resolver has not finished yet are missing.  The resolver therefore re-reads
// table[h] for every lane when it enters the window (all earlier inserts are committed by then);
// lanes whose entry moved are simply marked for exact evaluation, which happens only if the parse
// actually reaches them.  The resolver itself never reads input bytes on its fast path, so the
// input does not have to live in shared memory: with GlobalSrc the producers gather it from
// global memory (latency hidden by running ahead) and a fragment needs only its 32 KB table plus
// the ring -- six fragments per SM instead of two.  Bytes produced are identical to the serial
// parse (tests/snappy/snappy.cc:539-671) by the same argument as parse_window.cuh.
//
// Ring protocol (per fragment): slot = w % R, R % P == 0 so one producer owns all windows of a
// slot.  Producer p takes the smallest w >= cur_win with w % P == p, waits until w < cur_win + R,
// fills the slot, then publishes ready[slot] = w + 1.  The resolver publishes cur_win whenever it
// moves to another window (also when a long copy jumps over several) and kDone at the end.
#pragma once
#include "parse_window.cuh"

namespace sb {

constexpr uint32_t kDone = 0xffffffffu;
#ifndef SB_PIPE_EARLY
#define SB_PIPE_EARLY 2                  // the speculative look-up may run this many windows ahead of the final one
#endif
#ifndef SB_PIPE_SLEEP
#define SB_PIPE_SLEEP 100
#endif

// Slow-path probe of position p against candidate cand in ONE round trip: lane 0 compares the
// 4-byte words (snappy.cc:607), lanes 1..31 the following 124 bytes (FindMatchLength).  Returns 0
// when the probe misses, else the match length (>= 4).
template <class Src>
__device__ __forceinline__ uint32_t probe_and_extend(const Src& src, uint32_t cand, uint32_t p,
                                                     uint32_t end, int lane) {
    const uint32_t pb = p + (lane << 2);
    uint32_t m = 0;
    if (pb < end) {
        const uint32_t x = src.ld32_upto(cand + (lane << 2), end) ^ src.ld32_upto(pb, end);
        m = x ? (uint32_t)(__ffs(x) - 1) >> 3 : 4u;
        m = min(m, end - pb);
    }
    const uint32_t bal = __ballot_sync(kFull, m < 4);
    if (bal & 1u) return 0;
    if (bal) {
        const int first = __ffs(bal) - 1;
        return (first << 2) + __shfl_sync(kFull, m, first);
    }
    return 128 + src_match_len(src, cand + 128, p + 128, end, lane);
}

constexpr uint32_t kLongMatch = 127;     // pre-extension gave up: evaluate exactly when reached

// ---- ring ---------------------------------------------------------------------------------------
template <uint32_t R>
struct alignas(16) PipeShared {
    uint4 hdr[R];                        // x = L, y = SLOWM, z = inval, w = any_peer (see parse_window.cuh)
    uint4 rec[R][32];                    // x = h | Tspec<<16, y = nx | ek<<16, z = ph (lanes with my hash), w = chain summary (lanes taken when the fast path enters at this lane)
    volatile uint32_t ready[R];          // window index + 1 whose record is complete
    volatile uint32_t cur_win;           // resolver's current window, kDone when the parse ended
    unsigned int* err;                   // the context's error word (mapped host memory) or null: watchdogs raise it
    uint64_t bar;                        // mbarrier (count 1): the resolver arrives whenever cur_win moves
};

// per-lane "where does the parse go after my copy" (see parse_window.cuh)
__device__ __forceinline__ uint32_t lane_next(uint32_t pos, uint32_t ek, uint32_t ip_limit,
                                              uint32_t L, uint32_t SLOWM) {
    if (pos + ek - (pos & 31u) >= ip_limit) return 256u;    // pos + 4 + ml >= ip_limit
    if (ek >= 32) return 64u;                                // my copy leaves the window
    const uint32_t m = L & (kFull << ek);
    if (!m) return 32u;
    uint32_t v = __ffs(m) - 1;
    if ((SLOWM >> v) & 1u) v |= 128u;
    return v;
}

template <class Src, uint32_t R, uint32_t LEAD>
__device__ __forceinline__ void pipe_producer(const Src& src, const uint16_t* table, int shift,
                                              uint32_t len, uint32_t nwin, PipeShared<R>* ps,
                                              uint32_t p, uint32_t P, int lane) {
    static_assert(LEAD <= R, "producers may not run ahead of the ring");
    const uint32_t ip_limit = len - kMargin;
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t next = p, parity = 0;
    // A window is produced in three phases so that the table look-up the record is built on happens
    // as late as possible (fewer pending inserts missed) without the global-memory latency behind it:
    //   A   (any time)           own 16 bytes per lane, hash, same-hash lanes -- table-independent
    //   B1  (w < cur_win+LEAD_A)  speculative look-up, candidate bytes, 4-byte compare, pre-extension
    //   B2  (w < cur_win+LEAD)    look-up again; only lanes whose entry moved since B1 redo their
    //                             candidate (usually none: no memory latency), then publish.
    constexpr uint32_t LEAD_A = LEAD + SB_PIPE_EARLY;
    uint32_t have = kDone, have_b1 = kDone;
    uint32_t d0 = 0, d1 = 0, d2 = 0, d3 = 0, h = 0, ph = 0, inval = 0, T = 0, ml = 0;
    bool hit = false;
    auto evaluate = [&](uint32_t pos, uint32_t cand) {       // compare + pre-extend against one candidate
        uint32_t c0, c1, c2, c3;
        src.ld16(cand, c0, c1, c2, c3);
        hit = (c0 == d0);
        const uint32_t x1 = c1 ^ d1, x2 = c2 ^ d2, x3 = c3 ^ d3;
        ml = 12;
        if (x3) ml = 8 + ((uint32_t)(__ffs(x3) - 1) >> 3);
        if (x2) ml = 4 + ((uint32_t)(__ffs(x2) - 1) >> 3);
        if (x1) ml = (uint32_t)(__ffs(x1) - 1) >> 3;
        if (hit && ml == 12) {                               // keep extending, 16 bytes a step
            ml = kLongMatch;
            for (uint32_t o = 16; o < 80 && pos + o + 16 <= len; o += 16) {
                uint32_t e0, e1, e2, e3, g0, g1, g2, g3;
                src.ld16(pos + o, e0, e1, e2, e3);
                src.ld16(cand + o, g0, g1, g2, g3);
                const uint32_t y0 = e0 ^ g0, y1 = e1 ^ g1, y2 = e2 ^ g2, y3 = e3 ^ g3;
                if (y0 | y1 | y2 | y3) {
                    uint32_t q = 12 + ((uint32_t)(__ffs(y3) - 1) >> 3);
                    if (y2) q = 8 + ((uint32_t)(__ffs(y2) - 1) >> 3);
                    if (y1) q = 4 + ((uint32_t)(__ffs(y1) - 1) >> 3);
                    if (y0) q = (uint32_t)(__ffs(y0) - 1) >> 3;
                    ml = o - 4 + q;
                    break;
                }
            }
        }
    };
    for (uint32_t idle = 0;;) {
        const uint32_t cw = ps->cur_win;
        if (cw == kDone) return;
        if (next < cw) next = cw + (p + P - cw % P) % P;
        if (next < nwin && have != next) {                   // ---- phase A
            have = next;
            d0 = d1 = d2 = d3 = 0;
            const bool v = (next << 5) + lane < ip_limit;
            if (v) src.ld16((next << 5) + lane, d0, d1, d2, d3);
            h = hash32(d0, shift);
            ph = __match_any_sync(kFull, h);                 // slow instruction: keep it off the late phase
            inval = ~__ballot_sync(kFull, v);
        }
        if (next < nwin && have_b1 != next && next < cw + LEAD_A) {   // ---- phase B1
            have_b1 = next;
            T = table[h];                                    // speculative: may miss pending inserts
            hit = false; ml = 0;
            if ((next << 5) + lane < ip_limit) evaluate((next << 5) + lane, T);
        }
        if (next >= cw + LEAD || next >= nwin) {
            // far enough ahead (or nothing left): sleep in hardware until the resolver moves on.  The
            // parity is only a belief; a completed phase returns at once and flips it.
            if (mbar_try_wait(&ps->bar, parity)) parity ^= 1u;
#if SB_PIPE_SLEEP > 0
            else __nanosleep(SB_PIPE_SLEEP);                 // try_wait times out quickly: do not burn issue slots
#endif
            if (++idle > (1u << 24)) {                       // watchdog: never hang the GPU, never poison the context
                if (lane == 0 && ps->err) *(volatile unsigned int*)ps->err = 1u;
                return;                                      // the resolver's own watchdog ends the fragment
            }
            continue;
        }
        idle = 0;
        const uint32_t w = next, si = w % R;                 // ---- phase B2 (have_b1 == w: LEAD <= LEAD_A)
        const uint32_t pos = (w << 5) + lane;
        {
            const uint32_t T2 = table[h];
            if (pos < ip_limit && T2 != T) { T = T2; evaluate(pos, T); }
        }
        const bool peer = (ph & lt) != 0;
        const uint32_t vpre_end = inval ? (uint32_t)__ffs(inval) - 1u : 32u;
        const uint32_t L = __ballot_sync(kFull, hit || peer) & lanes_below(vpre_end);
        const uint32_t SLOWM = __ballot_sync(kFull, peer || ml == kLongMatch);
        const uint32_t anyp = __ballot_sync(kFull, peer);
        const uint32_t ek = lane + 4 + (ml == kLongMatch ? 12u : ml);
        const uint32_t nx = lane_next(pos, ek, ip_limit, L, SLOWM);
        // chain summary: the lanes whose copies are taken when the parse enters the window's fast
        // path at this lane (pointer jumping over nx; a chain has at most 8 copies)
        uint32_t chain = 1u << lane, ptr = nx < 32 ? nx : (uint32_t)lane;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            chain |= __shfl_sync(kFull, chain, ptr);
            ptr = __shfl_sync(kFull, ptr, ptr);
        }
        ps->rec[si][lane] = make_uint4(h | (T << 16), nx | (ek << 16), ph, chain);
        if (lane == 0) ps->hdr[si] = make_uint4(L, SLOWM, inval, anyp);
        __threadfence_block();
        __syncwarp();
        if (lane == 0) ps->ready[si] = w + 1;
        next = w + P;
    }
}

template <class Src, uint32_t R>
struct PipeResolver {
    Src src;
    uint32_t len, ip_limit;
    uint16_t* table;
    uint32_t table_s;                    // the table's address in the shared window
    PipeShared<R>* ps;
    int shift, lane;
    seq_t* seqs;
    uint32_t nseq;
    uint32_t next_emit;
    // diagnostics (SNAPPY_B200_STATS builds only): windows, wait cycles, stale windows, slow evaluations, ...
    unsigned long long st_wait = 0, st_slow = 0, st_pa = 0, st_pb = 0, st_pc = 0;
    uint32_t st_win = 0, st_stale = 0, st_nslow = 0, st_serial = 0, st_slowhit = 0, st_cstale = 0, st_cpeer = 0;

    // Write the window's pending inserts: per hash slot the highest inserted lane wins (serial order).
    __device__ __forceinline__ void commit(uint16_t*, uint32_t h, uint32_t pos, uint32_t I, uint32_t ph) {
        const uint32_t gt = ~((2u << lane) - 1u);
        if (((I >> lane) & 1u) && (ph & I & gt) == 0)
            asm volatile("st.shared.u16 [%0], %1;" :: "r"(table_s + (h << 1)), "h"((uint16_t)pos) : "memory");
        __syncwarp();
    }

    __device__ __forceinline__ void push(uint32_t lit, uint32_t mlen, uint32_t off) {
        if (lane == 0) seqs[nseq] = seq_pack(lit, mlen, off);
        ++nseq;
    }

    __device__ __forceinline__ void run() {
        const uint32_t lt = (1u << lane) - 1u;
        uint32_t win = 0, sl = 1, I = 0;
        int scan0 = 1;
        for (;;) {
            // ---- take the producer's record of this window, validate the speculation ----------
            const uint32_t wbase = win << 5, pos = wbase + lane, si = win % R;
#ifdef SNAPPY_B200_STATS
            const long long t0 = clock64();
#endif
            bool stuck = false;
            for (uint32_t spins = 0; ps->ready[si] != win + 1; ++spins)
                if (spins > (1u << 24)) { stuck = true; break; }
            if (__any_sync(kFull, stuck)) {                  // watchdog: raise the context's error word and end the
                if (lane == 0 && ps->err) *(volatile unsigned int*)ps->err = 1u;   // fragment (the caller emits the
                return;                                      // rest as one literal, so the lists stay well formed)
            }
#ifdef SNAPPY_B200_STATS
            st_wait += clock64() - t0; ++st_win;
#endif
            __syncwarp();
            const uint4 rec = ps->rec[si][lane];
            const uint4 hdr = ps->hdr[si];
            const uint32_t h = rec.x & 0xffffu;
            uint32_t T;                                      // exact: earlier windows are committed
            {
                uint16_t t16;
                asm volatile("ld.shared.u16 %0, [%1];" : "=h"(t16) : "r"(table_s + (h << 1)) : "memory");
                T = t16;
            }
            uint32_t L = hdr.x, SLOWM = hdr.y, nxek = rec.y;
            const uint32_t inval = hdr.z, ph = rec.z;
            const uint32_t vpre_end = inval ? (uint32_t)__ffs(inval) - 1u : 32u;
            const uint32_t stale = __ballot_sync(kFull, T != (rec.x >> 16)) & lanes_below(vpre_end);
#ifdef SNAPPY_B200_STATS
            if (stale) ++st_stale;
#endif
            if (stale) {       // those lanes' compare/extension are void: evaluate them if reached
                L |= stale;
                SLOWM |= stale;
                nxek = (nxek & 0xffff0000u) | lane_next(pos, nxek >> 16, ip_limit, L, SLOWM);
            }
#ifdef SNAPPY_B200_STATS
            const long long t_a = clock64(); st_pa += t_a - t0;
#endif
            // ---- sequential part -------------------------------------------------------------------
            uint32_t nwin_next = kDone;
            bool done_window = false;
            while (!done_window) {
                const uint32_t vend0 = scan0 < 0 ? (uint32_t)(scan0 + 32) : 32u;
                const uint32_t m0 = L & lanes_from(sl) & lanes_below(vend0);
                uint32_t f = 32, chosen = 0, cur_emit = next_emit, mylit = 0;
                if (m0) { f = __ffs(m0) - 1; if ((SLOWM >> f) & 1u) f |= 128u; }
#ifdef SNAPPY_B200_STATS
                const long long t_c0 = clock64();
#endif
                if (!stale) {
                    if (f < 32) {                             // the producer's chain summary holds
                        chosen = __shfl_sync(kFull, rec.w, f);
                        const uint32_t below = chosen & lt;
                        const uint32_t v = __shfl_sync(kFull, nxek, 31u - __clz(chosen));
                        const uint32_t pe = __shfl_sync(kFull, nxek, below ? 31u - __clz(below) : 0u) >> 16;
                        mylit = pos - (below ? wbase + pe : cur_emit);
                        cur_emit = wbase + (v >> 16);
                        f = v & 0x1ffu;
                    }
                } else {
                    while (f < 32) {                          // one shuffle per copy
                        chosen |= 1u << f;
                        if ((uint32_t)lane == f) mylit = pos - cur_emit;
                        const uint32_t v = __shfl_sync(kFull, nxek, f);
                        cur_emit = wbase + (v >> 16);
                        f = v & 0x1ffu;
                    }
                }
#ifdef SNAPPY_B200_STATS
                const long long t_b0 = clock64(); st_pc += t_b0 - t_c0;
#endif
                // lanes strictly inside a taken copy (lane+1 .. ek-2) are not inserted
                const uint32_t ek = nxek >> 16;
                const uint32_t Sk = lanes_below(ek - 1 < 32 ? ek - 1 : 32) & ~((2u << lane) - 1u);
                const uint32_t SK = __reduce_or_sync(kFull, ((chosen >> lane) & 1u) ? Sk : 0u);
                uint32_t e_last = 0;
                if (chosen) {
                    if ((chosen >> lane) & 1u) {
                        const uint2 q = make_uint2(((ek - lane) << 17) | mylit, (pos - T) << 1);
                        *reinterpret_cast<uint2*>(seqs + nseq + __popc(chosen & lt)) = q;
                    }
                    nseq += __popc(chosen);
                    next_emit = cur_emit;
                    e_last = cur_emit - wbase;
                    scan0 = (int)e_last + 1;
                }
                uint32_t ip = 0;
                bool jump = false, e32 = false;        // e32: the copy ended exactly at the window end
                if (f == 32) {
                    const uint32_t vend = chosen ? 32u : vend0;
                    const uint32_t upto = vend < vpre_end ? vend : vpre_end;
                    I |= lanes_from(sl) & lanes_below(upto) & ~SK;
                    if (vpre_end <= vend && inval) return;
                    commit(table, h, pos, I, ph);
                    if (vend < 32) {
                        uint32_t skip = 64, cand;
#ifdef SNAPPY_B200_STATS
                        ++st_serial;
#endif
                        ip = wbase + vend;
                        if (!src_serial_scan_call(src, table, shift, ip_limit, ip, skip, cand, lane)) return;
                        const uint32_t matched = 4 + src_match_len(src, cand + 4, ip + 4, len, lane);
                        push(ip - next_emit, matched, ip - cand);
                        ip += matched;
                        next_emit = ip;
                        if (ip >= ip_limit) return;
                        jump = true;
                    } else {
                        nwin_next = win + 1; sl = 0; scan0 -= 32; I = 0;
                    }
                    done_window = true;
                } else if (f & 256u) {
                    return;
                } else if (f & 64u) {
                    I |= lanes_from(sl) & ~SK;
                    commit(table, h, pos, I, ph);
                    ip = next_emit;
                    jump = true;
                    e32 = (e_last == 32);
                    done_window = true;
                } else {
                    const uint32_t fs = f & 31u;
#ifdef SNAPPY_B200_STATS
                    ++st_nslow; const long long t1 = clock64();
                    if ((stale >> fs) & 1u) ++st_cstale; else if (__shfl_sync(kFull, ph & lt, fs)) ++st_cpeer;
#endif
                    I |= lanes_from(sl) & lanes_below(fs) & ~SK;
                    const uint32_t pk = __shfl_sync(kFull, ph, fs) & I & lanes_below(fs);   // inserted peers
                    const uint32_t cand = pk ? wbase + (31u - __clz(pk)) : __shfl_sync(kFull, T, fs);
                    const uint32_t p = wbase + fs;
                    I |= 1u << fs;
                    const uint32_t matched = probe_and_extend(src, cand, p, len, lane);
#ifdef SNAPPY_B200_STATS
                    st_slow += clock64() - t1; if (matched) ++st_slowhit;
#endif
                    if (!matched) { sl = fs + 1; continue; }             // missed probe (:607)
                    push(p - next_emit, matched, p - cand);
                    ip = p + matched;
                    next_emit = ip;
                    if (ip >= ip_limit) return;
                    const uint32_t e = ip - wbase;
                    if (e <= 32) I |= 1u << (e - 1);
                    if (e < 32) { sl = e; scan0 = (int)e + 1; continue; }
                    commit(table, h, pos, I, ph);
                    jump = true;
                    e32 = (e == 32);
                    done_window = true;
                }
                if (jump) {                                   // continue at the window holding ip-1
                    if (e32) { nwin_next = win + 1; sl = 0; scan0 = 1; I = 0; }   // ip-1 was lane 31: committed above
                    else {
                        const uint32_t a = (ip - 1) & 31u;  // insert ip-1 (:649-651) from its own window
                        nwin_next = (ip - 1) >> 5; I = 1u << a; sl = a + 1; scan0 = (int)a + 2;
                    }
                }
            }
#ifdef SNAPPY_B200_STATS
            st_pb += clock64() - t_a;
#endif
            win = nwin_next;
            if (lane == 0) { ps->cur_win = win; mbar_arrive(&ps->bar); }
        }
    }
};

}  // namespace sb
