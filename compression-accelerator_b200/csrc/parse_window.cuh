// parse_window.cuh -- lane-parallel restatement of Snappy's greedy parse, byte-exact.
//
// The fragment is cut into aligned windows of 32 positions; lane k of the warp owns position
// 32*win + k.  Per window the lanes do, in parallel, everything that does not depend on the
// parse: load the 4-byte word, hash it (tests/snappy/snappy.cc:91-97), look the hash table up,
// compare with the candidate (:607), pre-extend the match by up to 12 bytes
// (snappy-internal.h:95-143) and -- assuming the copy at their position is taken -- work out
// where the parse goes next (:643-659: the copy's end, then the next position whose probe
// hits).  The sequential part of the parse is then a pointer chase over those per-lane "next"
// values: one warp shuffle per COPY, no shared-memory access on its critical path.
//
// Exactness.  A probe's candidate is the most recent insert with its hash.  Inserts of earlier
// windows are in the table when a window is entered.  Inserts made inside the window can only
// shadow the table entry of a lane that shares its hash with a lower lane ("peer" lanes, found
// with match.any -- a shared-memory min-scatter was no faster and its scratch cost two pages per SM); those lanes, and matches longer than the 12 pre-extended
// bytes, are evaluated exactly when the chase reaches them (slow path).  Table inserts (:606,
// :649-655) are kept as a lane mask and committed when the parse leaves the window, the highest
// lane per slot winning as in serial order.  Scans that miss 32 times in a row (skip >= 64,
// step > 1, :590-597) fall back to the serial scan.
#pragma once
#include "parse_serial.cuh"

namespace sb {


// five aligned words starting at the word containing `pos` -> the 16 bytes at pos..pos+15
__device__ __forceinline__ void lds16u(const uint8_t* s, uint32_t pos, uint32_t& b0, uint32_t& b1,
                                       uint32_t& b2, uint32_t& b3) {
    const uint32_t* w = reinterpret_cast<const uint32_t*>(s) + (pos >> 2);
    const uint32_t sh = (pos & 3u) << 3;
    uint32_t a0 = w[0], a1 = w[1], a2 = w[2], a3 = w[3], a4 = w[4];
    b0 = shf_r(a0, a1, sh);
    b1 = shf_r(a1, a2, sh);
    b2 = shf_r(a2, a3, sh);
    b3 = shf_r(a3, a4, sh);
}

// one SHF each: the funnel shift clamps the count at 32
__device__ __forceinline__ uint32_t lanes_from(uint32_t s) { return shf_lc(0u, kFull, s); }
__device__ __forceinline__ uint32_t lanes_below(uint32_t e) { return shf_lc(kFull, 0u, e); }

// ---- input accessors ------------------------------------------------------------------------
struct SmemSrc {                         // fragment staged in shared memory (4-byte aligned, padded)
    static constexpr bool kInPlace = false;          // table in global memory (speculative look-ups, commit in registers)
    static constexpr bool kInputAhead = false;       // input read in place: keep its lines coming into L1
    const uint8_t* s;
    __device__ __forceinline__ void prefetch(uint32_t) const {}
    __device__ __forceinline__ void touch(uint32_t) const {}
    __device__ __forceinline__ uint32_t ld32(uint32_t pos) const { return lds32u(s, pos); }
    __device__ __forceinline__ uint32_t ld32_upto(uint32_t pos, uint32_t) const { return lds32u(s, pos); }
    __device__ __forceinline__ void ld16(uint32_t pos, uint32_t& b0, uint32_t& b1, uint32_t& b2,
                                         uint32_t& b3) const { lds16u(s, pos, b0, b1, b2, b3); }
};

struct GlobalSrc {                       // fragment read in place from global memory
    static constexpr bool kInPlace = true;
    static constexpr bool kInputAhead = true;
    const uint8_t* g;                    // fragment start (any alignment)
    __device__ __forceinline__ void prefetch(uint32_t pos) const {       // the 128-byte line holding g[pos] -> L1
        asm volatile("prefetch.global.L1 [%0];" ::"l"(g + pos));
    }
    // One 32-byte SECTOR -> L1 (a load whose result nobody waits for).  prefetch.global.L1 moves whole lines:
    // for 32 random candidates per window that is four times the traffic (ncu: 47.7 GB of DRAM reads for
    // 512 MiB of input with line prefetches).
    __device__ __forceinline__ void touch(uint32_t pos) const {
        uint32_t unused;
        asm volatile("ld.global.ca.u32 %0, [%1];" : "=r"(unused) : "l"((uintptr_t)(g + pos) & ~(uintptr_t)3));
    }
    __device__ __forceinline__ uint32_t ld32(uint32_t pos) const {
        const uintptr_t a = (uintptr_t)(g + pos);
        const uint32_t* w = reinterpret_cast<const uint32_t*>(a & ~(uintptr_t)3);
        const uint32_t sh = ((uint32_t)a & 3u) << 3;
        const uint32_t a0 = __ldg(w), a1 = sh ? __ldg(w + 1) : 0u;
        return shf_r(a0, a1, sh);
    }
    // as ld32, but never touches a word that lies entirely at or beyond position `end` (the
    // last fragment may end on the last mapped byte of the caller's buffer)
    __device__ __forceinline__ uint32_t ld32_upto(uint32_t pos, uint32_t end) const {
        const uintptr_t a = (uintptr_t)(g + pos);
        const uint32_t* w = reinterpret_cast<const uint32_t*>(a & ~(uintptr_t)3);
        const uint32_t r = (uint32_t)a & 3u;
        const uint32_t a0 = __ldg(w), a1 = (r && pos + 4 - r < end) ? __ldg(w + 1) : 0u;
        return shf_r(a0, a1, r << 3);
    }
    __device__ __forceinline__ void ld16(uint32_t pos, uint32_t& b0, uint32_t& b1, uint32_t& b2,
                                         uint32_t& b3) const {
        const uintptr_t a = (uintptr_t)(g + pos);
        const uint32_t* w = reinterpret_cast<const uint32_t*>(a & ~(uintptr_t)3);
        const uint32_t sh = ((uint32_t)a & 3u) << 3;
        const uint32_t a0 = __ldg(w), a1 = __ldg(w + 1), a2 = __ldg(w + 2), a3 = __ldg(w + 3);
        const uint32_t a4 = sh ? __ldg(w + 4) : 0u;         // never touch a word without needed bytes
        b0 = shf_r(a0, a1, sh);
        b1 = shf_r(a1, a2, sh);
        b2 = shf_r(a2, a3, sh);
        b3 = shf_r(a3, a4, sh);
    }
};

// FindMatchLength (snappy-internal.h:95-143) over an accessor, 128 bytes per round.
template <class Src>
__device__ __forceinline__ uint32_t src_match_len(const Src& src, uint32_t a, uint32_t b,
                                                  uint32_t end, int lane) {
    uint32_t n = 0;
    for (;;) {
        const uint32_t pb = b + n + (lane << 2);
        uint32_t m = 0;
        if (pb < end) {
            const uint32_t x = src.ld32_upto(a + n + (lane << 2), end) ^ src.ld32_upto(pb, end);
            m = x ? (uint32_t)(__ffs(x) - 1) >> 3 : 4u;
            m = min(m, end - pb);
        }
        const uint32_t bal = __ballot_sync(kFull, m < 4);
        if (bal) {
            const int first = __ffs(bal) - 1;
            return n + (first << 2) + __shfl_sync(kFull, m, first);
        }
        n += 128;
    }
}

// The scan of a long run of misses (snappy.cc:587-608, entered with skip >= 64), 32 probes per warp step.
// The probe positions depend only on (ip, skip) -- step = skip >> 5, skip += step (:590-597) -- so lane k
// works out the k-th position of the schedule itself and all 32 probes are looked up and compared in
// parallel: two round trips per 32 probes instead of two per probe (a matchless 64 KiB fragment is 266
// probes, a 4 KiB page 155).  Exactness: probe k must see the inserts of probes 0..k-1; the table holds
// everything older, and among the 32 a lane's candidate is the highest LOWER lane with its hash
// (match.any), else the table entry.  Probes happen only up to the first hit, or up to the first position
// whose successor passes ip_limit (:598); exactly those lanes insert, the highest lane of a slot winning as
// in serial order.  Returns like the serial scan: true with (ip, cand) of the hit, false when the scan ran
// off the fragment.
//
// The schedule as a table.  skip and the position advance by the same step, so skip - 32 is the distance from
// the scan's first probe and the whole schedule is ONE sequence of distances, o(0) = 0, o(n+1) = o(n) + 1 +
// (o(n) >> 5): 267 entries cover a fragment.  Lane k reads o(n0 + k) -- one coalesced load -- instead of
// stepping through k probes (a 31-iteration loop per lane: 6 % of the page parser's instructions on text, most
// of them on incompressible data).
struct SkipSchedule {
    static constexpr int kN = 320;                           // o(266) = 66,672 > 65,536; n0 + 32 stays inside
    uint32_t off[kN];
    constexpr SkipSchedule() : off() {
        uint32_t o = 0;
        for (int n = 0; n < kN; ++n) { off[n] = o; o += 1u + (o >> 5); }
    }
};
static __device__ const SkipSchedule g_skip_schedule = SkipSchedule();

// n with o(n) == skip - 32 (skip is always a member: it only ever grows by this rule from 32)
__device__ __forceinline__ uint32_t skip_schedule_index(uint32_t skip, int lane) {
    const uint32_t o = skip - 32u;
    if (o <= 32u) return o;                                  // o(n) = n up to 32
    uint32_t below = 0;                                      // entries of my ten that are <= o
#pragma unroll
    for (int j = 0; j < SkipSchedule::kN / 32; ++j)
        below += __ldg(&g_skip_schedule.off[lane * (SkipSchedule::kN / 32) + j]) <= o ? 1u : 0u;
    for (int d = 16; d; d >>= 1) below += __shfl_xor_sync(kFull, below, d);
    return below - 1u;
}

template <class Src>
__device__ __forceinline__ bool src_serial_scan(const Src& src, uint16_t* table, int shift,
                                                uint32_t ip_limit, uint32_t& ip, uint32_t& skip,
                                                uint32_t& cand, int lane) {
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t n0 = skip_schedule_index(skip, lane);
    const uint32_t origin = ip - (skip - 32u);               // where this scan's first probe was
    for (;; n0 += 32u) {
        const uint32_t o = __ldg(&g_skip_schedule.off[n0 + lane]);   // my probe: the lane-th from here
        const uint32_t pos = origin + o, sk = 32u + o;
        const uint32_t step = sk >> 5, nxt = pos + step;     // :590-597 for my probe
        const bool valid = nxt <= ip_limit;                  // :598 (positions grow with the lane: a prefix is valid)
        uint32_t w = 0, h = 0x10000u + (uint32_t)lane;       // invalid lanes: a hash nobody shares
        if (valid) { w = src.ld32(pos); h = hash32(w, shift); }
        const uint32_t peers = __match_any_sync(kFull, h);
        const uint32_t lower = peers & lt;
        uint32_t c = valid ? table[h & 0xffffu] : 0u;
        const uint32_t peer_pos = __shfl_sync(kFull, pos, lower ? 31 - __clz(lower) : lane);
        if (lower) c = peer_pos;                             // an insert of this very batch shadows the table
        const bool hit = valid && src.ld32(c) == w;
        const uint32_t hits = __ballot_sync(kFull, hit), inval = __ballot_sync(kFull, !valid);
        const uint32_t fh = hits ? (uint32_t)__ffs(hits) - 1u : 32u, fi = inval ? (uint32_t)__ffs(inval) - 1u : 32u;
        const uint32_t n_probed = fh < fi ? fh + 1u : fi;    // lanes 0 .. n_probed-1 really probe (and insert, :606)
        const uint32_t probed = lanes_below(n_probed);
        if (((probed >> lane) & 1u) && (peers & probed & ~((2u << lane) - 1u)) == 0u) table[h] = (uint16_t)pos;
        __syncwarp();
        if (fh < fi) {
            ip = __shfl_sync(kFull, pos, fh);
            cand = __shfl_sync(kFull, c, fh);
            return true;
        }
        if (fi < 32) return false;
        ip = __shfl_sync(kFull, nxt, 31);                    // 32 misses: on with the schedule
        skip = __shfl_sync(kFull, sk + step, 31);
    }
}

// The same scan as a real call.  Measured (ms per parse): the page parser wants it inline (13.0 vs 14.1 per
// GiB of 4 KiB text pages: a call spills the window state), the pipelined resolver wants it out of its
// instruction stream (15.6 vs 16.1 per 512 MiB of text).
template <class Src>
__device__ __noinline__ bool src_serial_scan_call(const Src& src, uint16_t* table, int shift,
                                                  uint32_t ip_limit, uint32_t& ip, uint32_t& skip,
                                                  uint32_t& cand, int lane) {
    return src_serial_scan(src, table, shift, ip_limit, ip, skip, cand, lane);
}

// Write the pending inserts `I` of a window: table[h] = pos for every inserted lane, the highest
// lane of a hash slot winning.  Without peers in the window no two lanes share a slot.
__device__ __forceinline__ void window_commit(uint16_t* table, uint32_t h, uint32_t pos, uint32_t I,
                                              bool any_peer, int lane) {
    bool pending = (I >> lane) & 1u;
    if (pending) table[h] = (uint16_t)pos;
    __syncwarp();
    if (any_peer) {
        for (;;) {
            bool redo = pending && table[h] < pos;          // a lower lane's store landed last
            if (!__any_sync(kFull, redo)) break;
            pending = redo;
            if (pending) table[h] = (uint16_t)pos;
            __syncwarp();
        }
    }
}

template <class Src>
struct WindowParser {
    Src s;
    uint32_t len, ip_limit;
    uint16_t* table;
    int shift, lane;
    SeqWriter* out;
    uint32_t next_emit;

    // per-window lane registers
    uint32_t pos, h, T, ml, ek, nx, Sk;

    // The window's pending inserts I.  Shared-memory table: store, then settle same-slot lanes by reading back
    // (window_commit).  Global-memory table: a read-back is an L2 round trip, so the winner of a slot -- its
    // highest inserted lane, as in serial order -- is decided in registers from the same-hash mask.
    __device__ __forceinline__ void commit(uint32_t I, uint32_t ph, bool any_peer) {
        if constexpr (Src::kInPlace) {
            if (((I >> lane) & 1u) && (ph & I & ~((2u << lane) - 1u)) == 0u) table[h] = (uint16_t)pos;
            __syncwarp();
        } else {
            window_commit(table, h, pos, I, any_peer, lane);
        }
    }

    // Store the sequences of the fast copies in `chosen` (lane order = stream order).
    __device__ __forceinline__ void flush_chosen(uint32_t chosen, uint32_t lt) {
        const uint32_t prevmask = chosen & lt;
        const uint32_t prev = prevmask ? 31u - __clz(prevmask) : 0u;
        const uint32_t e_prev = __shfl_sync(kFull, ek, prev);
        if ((chosen >> lane) & 1u) {
            const uint32_t lit = prevmask ? (uint32_t)lane - e_prev : pos - next_emit;
            out->seqs[out->n + __popc(prevmask)] = seq_pack(lit, 4 + ml, pos - T);
        }
        out->n += __popc(chosen);
    }

    __device__ __forceinline__ void run() {
        const uint32_t lt = (1u << lane) - 1u;
        uint32_t win = 0, sl = 1, I = 0;                    // first probe is position 1 (:560)
        int scan0 = 1;                                      // lane of probe #0 of the current scan
        uint32_t tq = 0;                                    // in-place tables: entries looked up ahead (see below)
        for (;;) {                                          // one iteration per window
            // ---- parallel set-up -----------------------------------------------------------
            const uint32_t wbase = win << 5;
            pos = wbase + lane;
            uint32_t d0 = 0, d1 = 0, d2 = 0, d3 = 0;
            if (pos < ip_limit) s.ld16(pos, d0, d1, d2, d3);     // lanes past ip_limit never probe
            h = hash32(d0, shift);
            T = table[h];
            uint32_t c0, c1, c2, c3;
            s.ld16(T, c0, c1, c2, c3);       // (in place: 4 bytes for all + 12 only for hitting lanes was measured and
                                             // lost, 24.2 -> 26.1 ms per GiB: the second load is a dependent round trip)
            const uint32_t ph = __match_any_sync(kFull, h);             // lanes with my hash
            const bool peer = (ph & lt) != 0;                           // a lower lane has my hash
            const bool hit = (c0 == d0);
            {
                uint32_t x1 = c1 ^ d1, x2 = c2 ^ d2, x3 = c3 ^ d3, m = 12;
                if (x3) m = 8 + ((uint32_t)(__ffs(x3) - 1) >> 3);
                if (x2) m = 4 + ((uint32_t)(__ffs(x2) - 1) >> 3);
                if (x1) m = (uint32_t)(__ffs(x1) - 1) >> 3;
                ml = m;
            }
            const uint32_t inval = ~__ballot_sync(kFull, pos < ip_limit);
            const uint32_t vpre_end = inval ? (uint32_t)__ffs(inval) - 1u : 32u;   // first lane past ip_limit
            const uint32_t VPRE = lanes_below(vpre_end);
            const uint32_t L = __ballot_sync(kFull, hit || peer) & VPRE;           // lanes whose probe may hit
            const uint32_t SLOWM = __ballot_sync(kFull, peer || ml == 12);
            const bool any_peer = __any_sync(kFull, peer);
            // In-place input and a table in global memory (parse_gtab_kernel): every window would start with two
            // dependent L2 round trips (table entries, then candidate bytes).  Windows mostly follow one another,
            // so both are taken off the critical path by speculation, two deep: at window w the entries of window
            // w + 2 are looked up (the lines land in this SM's L1, which stays coherent for a table only this
            // warp writes), and the entries looked up one window ago -- they have long arrived -- say which
            // sectors to touch for the candidates of window w + 1.  The real look-up and the candidate read at
            // the start of a window are then L1 hits; an entry changed by an insert in between is seen by the
            // re-read and costs only that lane a miss, and after a jump the speculation is simply wasted.
            if constexpr (Src::kInPlace) {
                s.touch(tq); s.touch(tq + 15u);              // candidates of the next window (tq: table values < len)
                const uint32_t p2 = pos + 64u;
                tq = p2 < ip_limit ? (uint32_t)table[hash32(s.ld32(p2), shift)] : 0u;
            }
            if constexpr (Src::kInputAhead) {
                if (lane == 0 && wbase + 256u < len) s.prefetch(wbase + 256u);   // the input itself, two lines ahead
            }
            ek = lane + 4 + ml;                              // lane of ip after taking my copy
            {
                const uint32_t ipk = pos + 4 + ml;
                uint32_t v;
                if (ipk >= ip_limit) v = 256;                // parse ends after my copy (:643-645)
                else if (ek >= 32) v = 64u | (ek - 32);      // my copy leaves the window
                else {
                    const uint32_t m = L & (kFull << ek);
                    if (!m) v = 32;                          // no further hit in this window
                    else { v = __ffs(m) - 1; if ((SLOWM >> v) & 1u) v |= 128u; }
                }
                nx = v;
                const uint32_t last_skipped = ek - 1 < 32 ? ek - 1 : 32;   // lanes lane+1 .. ek-2 are not inserted
                Sk = lanes_below(last_skipped) & ~((2u << lane) - 1u);
            }
            // ---- sequential part: chase the copies of this window ------------------------------
            bool done_window = false;
            while (!done_window) {
                const uint32_t vend0 = scan0 < 0 ? (uint32_t)(scan0 + 32) : 32u;  // step-1 regime ends here
                const uint32_t m0 = L & lanes_from(sl) & lanes_below(vend0);
                uint32_t f = 32, last = 32, chosen = 0;
                if (m0) { f = __ffs(m0) - 1; if ((SLOWM >> f) & 1u) f |= 128u; }
                while (f < 32) {
                    chosen |= 1u << f;
                    last = f;
                    f = __shfl_sync(kFull, nx, f);
                }
                const uint32_t SK = __reduce_or_sync(kFull, ((chosen >> lane) & 1u) ? Sk : 0u);
                uint32_t e_last = 0;
                if (chosen) {
                    e_last = __shfl_sync(kFull, ek, last);
                    flush_chosen(chosen, lt);
                    next_emit = wbase + e_last;
                    scan0 = (int)e_last + 1;
                }
                if (f == 32) {                               // no more hits: the scan runs off its range
                    const uint32_t vend = chosen ? 32u : vend0;
                    const uint32_t upto = vend < vpre_end ? vend : vpre_end;
                    I |= lanes_from(sl) & lanes_below(upto) & ~SK;
                    if (vpre_end <= vend && inval) return;   // next probe would pass ip_limit (:598)
                    commit(I, ph, any_peer);
                    if (vend < 32) {                         // 32 misses in a row: skip reached 64
                        uint32_t ip = wbase + vend, skip = 64, cand;
                        if (!src_serial_scan(s, table, shift, ip_limit, ip, skip, cand, lane)) return;
                        const uint32_t matched = 4 + src_match_len(s, cand + 4, ip + 4, len, lane);
                        out->push(ip - next_emit, matched, ip - cand, lane);
                        ip += matched;
                        next_emit = ip;
                        if (ip >= ip_limit) return;
                        const uint32_t a = (ip - 1) & 31u;
                        win = (ip - 1) >> 5; I = 1u << a; sl = a + 1; scan0 = (int)a + 2;
                    } else {
                        win += 1; sl = 0; scan0 -= 32; I = 0;
                    }
                    done_window = true;
                } else if (f & 256u) {                       // last copy reached the end margin
                    return;
                } else if (f & 64u) {                        // last copy leaves the window
                    I |= lanes_from(sl) & ~SK;
                    commit(I, ph, any_peer);
                    const uint32_t ip = next_emit;
                    if (e_last == 32) { win += 1; sl = 0; scan0 = 1; I = 0; }
                    else {
                        const uint32_t a = (ip - 1) & 31u;
                        win = (ip - 1) >> 5; I = 1u << a; sl = a + 1; scan0 = (int)a + 2;
                    }
                    done_window = true;
                } else {                                     // exact evaluation of lane fs
                    const uint32_t fs = f & 31u;
                    I |= lanes_from(sl) & lanes_below(fs) & ~SK;
                    const uint32_t sameh = __ballot_sync(kFull, h == __shfl_sync(kFull, h, fs));
                    const uint32_t pk = sameh & I & lanes_below(fs);     // inserted peers, serial order
                    const uint32_t cand = pk ? wbase + (31u - __clz(pk)) : __shfl_sync(kFull, T, fs);
                    const uint32_t p = wbase + fs;
                    I |= 1u << fs;
                    if (s.ld32(cand) != s.ld32(p)) { sl = fs + 1; continue; }   // missed probe (:607)
                    const uint32_t matched = 4 + src_match_len(s, cand + 4, p + 4, len, lane);
                    out->push(p - next_emit, matched, p - cand, lane);
                    const uint32_t ip = p + matched;
                    next_emit = ip;
                    if (ip >= ip_limit) return;
                    const uint32_t e = ip - wbase;
                    if (e <= 32) I |= 1u << (e - 1);         // insert ip-1 (:649-651)
                    if (e < 32) { sl = e; scan0 = (int)e + 1; continue; }
                    commit(I, ph, any_peer);
                    if (e == 32) { win += 1; sl = 0; scan0 = 1; I = 0; }
                    else {
                        const uint32_t a = (ip - 1) & 31u;
                        win = (ip - 1) >> 5; I = 1u << a; sl = a + 1; scan0 = (int)a + 2;
                    }
                    done_window = true;
                }
            }
        }
    }
};

// Returns true when the parse found no copy at all: the job is one literal of `len` bytes (len > 0).
template <class Src>
__device__ __forceinline__ bool parse_fragment_window(const Src& s, uint32_t len,
                                                      uint16_t* table, int shift, SeqWriter& out,
                                                      int lane) {
    WindowParser<Src> wp;
    wp.next_emit = 0;
    if (len >= kMargin) {                                   // snappy.cc:557
        wp.s = s; wp.len = len; wp.ip_limit = len - kMargin;   // :558
        wp.table = table;
        wp.shift = shift; wp.lane = lane; wp.out = &out;
        wp.run();
    }
    if (wp.next_emit < len) out.push(len - wp.next_emit, 0, 0, lane);   // :665-668
    return wp.next_emit == 0 && len > 0 && out.n == 1;
}

}  // namespace sb
