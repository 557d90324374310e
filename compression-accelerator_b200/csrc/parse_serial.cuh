// parse_serial.cuh -- warp-uniform serial restatement of Snappy's greedy parse
// (tests/snappy/snappy.cc:539-671) producing sequences instead of bytes.  Every lane follows the
// same control flow; lanes only split the match extension.  Kept as the simple parser
// (SNAPPY_B200 mode 2, "exact_serial") and as the long-scan fallback of the window parser.
#pragma once
#include "kernels.h"
#include "snappy_device.cuh"

namespace sb {

struct SeqWriter {
    seq_t* seqs;
    uint32_t n;
    __device__ __forceinline__ void push(uint32_t lit, uint32_t len, uint32_t off, int lane) {
        if (lane == 0) seqs[n] = seq_pack(lit, len, off);
        ++n;
    }
};

// Serial scan from `ip` with skip counter `skip` until a 4-byte match (returns true, ip/cand
// set) or until the fragment's end margin is reached (returns false).  snappy.cc:587-608.
__device__ __forceinline__ bool serial_scan(const uint8_t* s, uint16_t* table, int shift,
                                            uint32_t ip_limit, uint32_t& ip, uint32_t& skip,
                                            uint32_t& cand) {
    uint32_t nxt = ip;
    for (;;) {
        ip = nxt;
        uint32_t step = skip >> 5;
        skip += step;
        nxt = ip + step;
        if (nxt > ip_limit) return false;                   // :598-600
        uint32_t w = lds32u(s, ip);
        uint32_t h = hash32(w, shift);
        cand = table[h];
        table[h] = (uint16_t)ip;                            // :606
        if (lds32u(s, cand) == w) return true;              // :607
    }
}

__device__ __forceinline__ void parse_fragment_serial(const uint8_t* s, uint32_t len,
                                                      uint16_t* table, int shift, SeqWriter& out,
                                                      int lane) {
    uint32_t ip = 0, next_emit = 0;
    if (len >= kMargin) {                                   // snappy.cc:557
        const uint32_t ip_limit = len - kMargin;            // :558
        ip = 1;                                             // :560
        for (;;) {
            uint32_t skip = 32, cand;
            if (!serial_scan(s, table, shift, ip_limit, ip, skip, cand)) break;
            uint32_t lit = ip - next_emit;                  // :614 literal before the match
            for (;;) {                                      // :627-656 copy loop
                uint32_t base = ip;
                uint32_t matched = 4 + warp_match_len(s, cand + 4, ip + 4, len, lane);
                ip += matched;
                out.push(lit, matched, base - cand, lane);
                lit = 0;
                next_emit = ip;
                if (ip >= ip_limit) goto remainder;         // :643-645
                uint32_t w0 = lds32u(s, ip - 1), w = lds32u(s, ip);
                table[hash32(w0, shift)] = (uint16_t)(ip - 1);      // :649-651
                uint32_t h = hash32(w, shift);
                cand = table[h];
                table[h] = (uint16_t)ip;                    // :652-655
                if (lds32u(s, cand) != w) break;            // :656
            }
            ip++;                                           // :658-659
        }
    }
remainder:
    if (next_emit < len) out.push(len - next_emit, 0, 0, lane);   // :665-668
}

}  // namespace sb
