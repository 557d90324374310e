// tests/host_lane/lane_host.cc -- TEST INFRASTRUCTURE ONLY.
//
// Compiles the per-fragment routine of the lane parser (compression-accelerator_b200/csrc/
// parse_lane.cuh, the same source the CUDA kernel instantiates) for the host with a plain-memory
// policy, and turns its sequence records into a Snappy stream with a small encoder written here
// (EmitLiteral / EmitCopy rules of tests/snappy/snappy.cc:337-431).  tests/test_lane_host.py compares
// that stream with the reference Snappy's byte for byte, so the parser's logic is pinned without a
// GPU.  The product never loads this file.
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

#include "../../compression-accelerator_b200/csrc/parse_lane.cuh"

namespace {

struct Host {
    static constexpr uint32_t kAhead = 192;
    static uint32_t ld_in(const uint32_t* p) { uint32_t v; memcpy(&v, p, 4); return v; }
    static uint32_t ld_byte(const uint8_t* p) { return *p; }
    static void* slot(void* table, uint32_t h) { return (sb::lane::Entry*)table + h; }
    static sb::lane::Entry ld_tab(const void* p) { return *(const sb::lane::Entry*)p; }
    static void st_tab(void* p, const sb::lane::Entry& e, uint32_t) { *(sb::lane::Entry*)p = e; }
    static sb::lane::Entry xchg_tab(void* p, const sb::lane::Entry& n) {
        sb::lane::Entry o = *(sb::lane::Entry*)p; *(sb::lane::Entry*)p = n; return o;
    }
    static void st_seq(sb::lane::seq64* p, sb::lane::seq64 v) { *p = v; }
    static void prefetch_tab(const void* p) { (void)*(volatile const uint32_t*)p; }   // must be a mapped slot
    static void prefetch_in(const uint8_t*, uint32_t, uint32_t) {}
};

size_t put_varint(uint8_t* d, uint32_t v) {
    size_t k = 0;
    while (v >= 128) { d[k++] = (uint8_t)(v | 128); v >>= 7; }
    d[k++] = (uint8_t)v;
    return k;
}

size_t put_copy64(uint8_t* o, uint32_t off, uint32_t len, bool lt12) {
    if (lt12 && off < 2048) { o[0] = (uint8_t)(1 | ((len - 4) << 2) | ((off >> 8) << 5)); o[1] = (uint8_t)off; return 2; }
    o[0] = (uint8_t)(2 | ((len - 1) << 2)); o[1] = (uint8_t)off; o[2] = (uint8_t)(off >> 8);
    return 3;
}

}  // namespace

extern "C" {

// Whole stream: varint32(n) + fragments.  One table reused for every fragment with a fresh epoch each
// time (starting at epoch0), i.e. stale entries of earlier fragments are always present.
// pf selects the look-ahead variant (kVar bits 0-1 of parse_fragment).  Returns the stream length, or 0
// if a fragment's running size differs from the encoded size / the records do not cover the fragment.
size_t lane_host_compress(const uint8_t* in, size_t n, uint8_t* out, int pf, uint32_t epoch0) {
    using namespace sb::lane;
    std::vector<Entry> table(kLaneMaxTable);
    memset(table.data(), 0, table.size() * sizeof(Entry));
    std::vector<seq64> seqs(65536 / 4 + 8);
    // padded private copy: the routine reads whole aligned words around the bytes it needs
    std::vector<uint8_t> buf(n + 64, 0xA5);
    uint8_t* src0 = buf.data() + 16 + ((uintptr_t)in & 3);      // keep the caller's alignment mod 4
    memcpy(src0, in, n);
    size_t op = put_varint(out, (uint32_t)n);
    uint32_t epoch = epoch0;
    for (size_t lo = 0; lo < n; lo += 65536) {
        const uint32_t len = (uint32_t)(n - lo < 65536 ? n - lo : 65536);
        const uint8_t* src = src0 + lo;
        Out o; o.seqs = seqs.data(); o.n = 0; o.bytes = 0;
        const uint32_t entries = entries_for(len), shift = shift_for(entries);
        if (pf == 0) parse_fragment<Host, 0>(src, len, table.data(), shift, epoch, o);
        else if (pf == 1) parse_fragment<Host, 1>(src, len, table.data(), shift, epoch, o);
        else if (pf == 2) parse_fragment<Host, 2>(src, len, table.data(), shift, epoch, o);
        else if (pf == 32) parse_fragment<Host, 32>(src, len, table.data(), shift, epoch, o);
        else parse_fragment<Host, 34>(src, len, table.data(), shift, epoch, o);
        epoch = epoch == 65535 ? 1 : epoch + 1;
        if (epoch == epoch0) memset(table.data(), 0, table.size() * sizeof(Entry));
        const size_t op0 = op;
        uint32_t ip = 0;
        for (uint32_t i = 0; i < o.n; ++i) {
            const seq64 s = seqs[i];
            const uint32_t lit = (uint32_t)s & 0x1ffff, cl = (uint32_t)(s >> 17) & 0xffff, off = (uint32_t)(s >> 33) & 0xffff;
            if (lit) {
                const uint32_t m = lit - 1;
                if (m < 60) out[op++] = (uint8_t)(m << 2);
                else if (m < 256) { out[op++] = 60 << 2; out[op++] = (uint8_t)m; }
                else { out[op++] = 61 << 2; out[op++] = (uint8_t)m; out[op++] = (uint8_t)(m >> 8); }
                memcpy(out + op, src + ip, lit);
                op += lit; ip += lit;
            }
            if (cl) {
                if (cl < 4 || off == 0 || off > ip) return 0;
                uint32_t rem = cl;
                if (rem >= 12) {
                    while (rem >= 68) { op += put_copy64(out + op, off, 64, false); rem -= 64; }
                    if (rem > 64) { op += put_copy64(out + op, off, 60, false); rem -= 60; }
                }
                op += put_copy64(out + op, off, rem, rem < 12);
                ip += cl;
            } else if (i + 1 != o.n) return 0;
        }
        if (ip != len || op - op0 != o.bytes) return 0;
    }
    return op;
}

}  // extern "C"
