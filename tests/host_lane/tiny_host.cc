// tests/host_lane/tiny_host.cc -- TEST INFRASTRUCTURE ONLY.
// Compiles the per-buffer routine of the small-buffer kernel (compression-accelerator_b200/csrc/
// parse_tiny.cuh, the same source the CUDA kernel instantiates) for the host; tests/test_tiny_host.py
// compares its streams with the reference Snappy's byte for byte.  The product never loads this file.
#include <stdint.h>
#include <string.h>
#include <vector>

#include "../../compression-accelerator_b200/csrc/parse_tiny.cuh"

namespace {
struct Host {
    static uint32_t ld32(const uint32_t* p) { uint32_t v; memcpy(&v, p, 4); return v; }
    static uint32_t ld8(const uint8_t* p) { return *p; }
    static void st8(uint8_t* p, uint8_t v) { *p = v; }
};
struct Tab {
    uint16_t* base;
    uint32_t ld(uint32_t h) const { return base[h]; }
    void st(uint32_t h, uint16_t v) const { base[h] = v; }
};
}  // namespace

extern "C" size_t tiny_host_compress(const uint8_t* in, size_t n, uint8_t* out) {
    using namespace sb::tiny;
    if (n > 65536) return 0;
    std::vector<uint8_t> buf(n + 64, 0xA5);                       // the routine reads whole aligned words
    uint8_t* src = buf.data() + 16 + ((uintptr_t)in & 3);
    memcpy(src, in, n);
    const uint32_t entries = entries_for((uint32_t)n);
    std::vector<uint16_t> table(entries, 0);
    return compress_one<Host>(src, (uint32_t)n, Tab{table.data()}, entries, out, true);
}
