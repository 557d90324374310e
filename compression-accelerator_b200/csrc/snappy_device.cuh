// snappy_device.cuh -- device-side building blocks shared by the Snappy kernels (sm_100a).
//
// Element encoders, unaligned shared-memory access, TMA bulk load of a fragment, warp-wide
// match extension.  Reference behaviour restated here is cited as tests/snappy/<file>:<line>
// (paths relative to the reference repository); nothing is copied from it.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace sb {

constexpr uint32_t kBlock = 65536;      // tests/snappy/snappy.h:196 kBlockSize
constexpr uint32_t kMaxTable = 16384;   // tests/snappy/snappy.h:199-200 kMaxHashTableSize
constexpr uint32_t kMargin = 15;        // tests/snappy/snappy.cc:556 kInputMarginBytes
constexpr uint32_t kHashMul = 0x1e35a7bdu;  // tests/snappy/snappy.cc:92 (same constant as HashTable.scala:23)
constexpr uint32_t kFull = 0xffffffffu;

// Funnel shifts as plain (not volatile) asm.  The toolkit's __funnelshift_* are `asm volatile`: never removed when
// their result is unused, never merged when repeated, and kept in program order with every other volatile asm --
// all of which costs issue slots and scheduling freedom in the one-warp parsers (7 % of the page parser's
// instructions were funnel shifts).
__device__ __forceinline__ uint32_t shf_r(uint32_t lo, uint32_t hi, uint32_t s) {      // (hi:lo >> (s & 31)), low word
    uint32_t r;
    asm("shf.r.wrap.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(lo), "r"(hi), "r"(s));
    return r;
}
__device__ __forceinline__ uint32_t shf_lc(uint32_t lo, uint32_t hi, uint32_t s) {     // (hi:lo << min(s, 32)), high word
    uint32_t r;
    asm("shf.l.clamp.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(lo), "r"(hi), "r"(s));
    return r;
}

// tests/snappy/snappy.cc:445-456 CalculateTableSize: 256 <= next_pow2(len) <= 16384
__host__ __device__ inline uint32_t table_entries(uint32_t len) {
    if (len > kMaxTable) return kMaxTable;
    uint32_t e = 256;
    while (e < len) e <<= 1;
    return e;
}

// tests/snappy/snappy-stubs-internal.h:552-578 Varint::Encode32
__host__ __device__ inline uint32_t varint32_len(uint32_t v) {
    return v < (1u << 7) ? 1 : v < (1u << 14) ? 2 : v < (1u << 21) ? 3 : v < (1u << 28) ? 4 : 5;
}
__host__ __device__ inline uint32_t put_varint32(uint8_t* dst, uint32_t v) {
    uint32_t k = 0;
    while (v >= 128) { dst[k++] = (uint8_t)(v | 128); v >>= 7; }
    dst[k++] = (uint8_t)v;
    return k;
}

#ifdef __CUDACC__

// Unaligned little-endian 32-bit read from a 4-byte-aligned shared buffer (two aligned words
// and a funnel shift; the GPU analogue of MemoryReadAligner.scala).  Reads up to 7 bytes past
// `pos`, so buffers carry padding.
__device__ __forceinline__ uint32_t lds32u(const uint8_t* s, uint32_t pos) {
    const uint32_t* w = reinterpret_cast<const uint32_t*>(s);
    uint32_t i = pos >> 2;
    return shf_r(w[i], w[i + 1], (pos & 3u) << 3);
}

__device__ __forceinline__ uint32_t hash32(uint32_t w, int shift) {  // snappy.cc:91-97
    return (w * kHashMul) >> shift;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}

// ---- mbarrier + 1-D TMA bulk copy (cp.async.bulk; SASS UBLKCP) --------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// One probe of the barrier's phase `parity`; may suspend the thread in hardware for a bounded
// time while the phase is incomplete.  True once the phase has completed.
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// dst (shared) and src (global) 16-byte aligned, bytes a multiple of 16.
__device__ __forceinline__ void tma_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes,
                                            uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
            "r"(smem_u32(smem_dst)),
        "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// ---- element encoders ---------------------------------------------------------------------
// Copy element of <=64 bytes (snappy.cc:380-400 EmitCopyAtMost64); returns bytes written.
__device__ __forceinline__ uint32_t put_copy_le64(uint8_t* out, uint32_t offset, uint32_t len,
                                                  bool total_lt12) {
    if (total_lt12 && offset < 2048) {
        out[0] = (uint8_t)(1u | ((len - 4) << 2) | ((offset >> 8) << 5));
        out[1] = (uint8_t)offset;
        return 2;
    }
    out[0] = (uint8_t)(2u | ((len - 1) << 2));
    out[1] = (uint8_t)offset;
    out[2] = (uint8_t)(offset >> 8);
    return 3;
}

// Number of output bytes EmitCopy (snappy.cc:402-431) produces for (offset, len).
__device__ __forceinline__ uint32_t copy_encoded_size(uint32_t offset, uint32_t len) {
    if (len < 12) return offset < 2048 ? 2 : 3;
    uint32_t k = (len >= 68) ? ((len - 4) >> 6) : 0;   // 64-byte copies
    uint32_t rest = len - (k << 6), n = 3 * k;
    if (rest > 64) { n += 3; rest -= 60; }
    return n + ((rest < 12 && offset < 2048) ? 2 : 3);
}

// Warp-cooperative EmitCopy: every lane passes the same (offset, len); the run of 64-byte
// copies of a long match (snappy.cc:412-420) is spread over the lanes.
__device__ __forceinline__ uint32_t warp_emit_copy(uint8_t* out, uint32_t op, uint32_t offset,
                                                   uint32_t len, int lane) {
    if (len < 12) {
        if (lane == 0) put_copy_le64(out + op, offset, len, true);
        return op + (offset < 2048 ? 2 : 3);
    }
    if (len >= 68) {
        uint32_t k = (len - 4) >> 6;
        for (uint32_t i = lane; i < k; i += 32) put_copy_le64(out + op + 3 * i, offset, 64, false);
        op += 3 * k;
        len -= k << 6;
    }
    if (len > 64) {
        if (lane == 0) put_copy_le64(out + op, offset, 60, false);
        op += 3;
        len -= 60;
    }
    if (lane == 0) put_copy_le64(out + op, offset, len, len < 12);
    return op + ((len < 12 && offset < 2048) ? 2 : 3);
}

// Literal header size for a literal of `len` bytes inside one fragment (len <= 65536).
__device__ __forceinline__ uint32_t literal_header_size(uint32_t len) {
    uint32_t n = len - 1;
    return n < 60 ? 1 : n < 256 ? 2 : 3;
}

// Warp-cooperative EmitLiteral (snappy.cc:337-378): tag (+1..2 length bytes, since len <= 64 Ki)
// then the bytes s[start, start+len) from shared memory to global memory.
__device__ __forceinline__ uint32_t warp_emit_literal(uint8_t* out, uint32_t op, const uint8_t* s,
                                                      uint32_t start, uint32_t len, int lane) {
    uint32_t n = len - 1;
    if (n < 60) {
        if (lane == 0) out[op] = (uint8_t)(n << 2);
        op += 1;
    } else if (n < 256) {
        if (lane == 0) { out[op] = 60u << 2; out[op + 1] = (uint8_t)n; }
        op += 2;
    } else {
        if (lane == 0) { out[op] = 61u << 2; out[op + 1] = (uint8_t)n; out[op + 2] = (uint8_t)(n >> 8); }
        op += 3;
    }
    uint8_t* d = out + op;
    if (len <= 64) {
        for (uint32_t i = lane; i < len; i += 32) d[i] = s[start + i];
        return op + len;
    }
    // long literal: bytes up to a 16-byte boundary of the destination, then 16-byte stores
    // assembled from unaligned shared reads, then the tail.
    uint32_t head = (uint32_t)((16u - ((uintptr_t)d & 15u)) & 15u);
    if ((uint32_t)lane < head) d[lane] = s[start + lane];
    uint32_t body = (len - head) >> 4;
    uint4* d4 = reinterpret_cast<uint4*>(d + head);
    uint32_t sp = start + head;
    for (uint32_t j = lane; j < body; j += 32) {
        uint32_t q = sp + (j << 4);
        const uint32_t* w = reinterpret_cast<const uint32_t*>(s) + (q >> 2);
        uint32_t sh = (q & 3u) << 3;
        uint32_t a0 = w[0], a1 = w[1], a2 = w[2], a3 = w[3], a4 = w[4];
        uint4 v;
        v.x = shf_r(a0, a1, sh);
        v.y = shf_r(a1, a2, sh);
        v.z = shf_r(a2, a3, sh);
        v.w = shf_r(a3, a4, sh);
        d4[j] = v;
    }
    uint32_t done = head + (body << 4);
    if (done + lane < len) d[done + lane] = s[start + done + lane];
    return op + len;
}

// Warp-cooperative FindMatchLength (snappy-internal.h:95-143): length of the common prefix of
// s[a..] and s[b..], with b bounded by `end`; 128 bytes per round (4 per lane + ballot).
__device__ __forceinline__ uint32_t warp_match_len(const uint8_t* s, uint32_t a, uint32_t b,
                                                   uint32_t end, int lane) {
    uint32_t n = 0;
    for (;;) {
        uint32_t pb = b + n + (lane << 2);
        uint32_t m = 0;
        if (pb < end) {
            uint32_t x = lds32u(s, a + n + (lane << 2)) ^ lds32u(s, pb);
            m = x ? (uint32_t)(__ffs(x) - 1) >> 3 : 4u;
            m = min(m, end - pb);
        }
        uint32_t bal = __ballot_sync(kFull, m < 4);
        if (bal) {
            int first = __ffs(bal) - 1;
            return n + (first << 2) + __shfl_sync(kFull, m, first);
        }
        n += 128;
    }
}

#endif  // __CUDACC__
}  // namespace sb
