/*
 * oracle/snappy_port.c -- CPU restatement of the reference's software Snappy
 * block compressor (and a minimal decompressor for round-trip checks).
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the shipped library
 * (compression-accelerator_b200/csrc, libsnappy_b200.so) includes, links or
 * calls this file.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may use it, and only as the checker
 * or as the timed CPU baseline ("port") when oracle/_ref is not available.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks this port byte-for-byte
 * against (a) oracle/_ref/libsnappy_ref.so, which is the reference's own
 * tests/snappy sources compiled unmodified by oracle/Makefile, (b) the
 * committed golden vectors in tests/golden/ (generated from that library by
 * tests/golden/make_golden.py), and (c) the known-answer compressed lengths
 * of benchmark/sw_results.csv carried in tests/golden/sw_results_lengths.json.
 *
 * Written offset-based (all positions are uint32 offsets from the start of
 * the fragment) rather than pointer-based like the reference; every function
 * cites the reference lines whose behaviour it restates.  Paths are relative
 * to /root/reference.
 */
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#define PORT_BLOCK 65536u          /* tests/snappy/snappy.h:196  kBlockSize   */
#define PORT_MAX_TABLE 16384u      /* tests/snappy/snappy.h:199-200           */
#define PORT_MARGIN 15u            /* tests/snappy/snappy.cc:556 kInputMarginBytes */

static inline uint32_t rd32(const uint8_t *p) { uint32_t v; memcpy(&v, p, 4); return v; }
static inline uint64_t rd64(const uint8_t *p) { uint64_t v; memcpy(&v, p, 8); return v; }

/* tests/snappy/snappy.cc:99-121  MaxCompressedLength */
size_t port_max_compressed_length(size_t n) { return 32 + n + n / 6; }

/* tests/snappy/snappy-stubs-internal.h:552-578  Varint::Encode32 */
size_t port_varint32(uint8_t *dst, uint32_t v) {
    size_t k = 0;
    while (v >= 128) { dst[k++] = (uint8_t)(v | 128); v >>= 7; }
    dst[k++] = (uint8_t)v;
    return k;
}

/* tests/snappy/snappy.cc:445-456  CalculateTableSize */
uint32_t port_table_entries(uint32_t fragment_len) {
    if (fragment_len > PORT_MAX_TABLE) return PORT_MAX_TABLE;
    if (fragment_len < 256) return 256;
    uint32_t e = 256;
    while (e < fragment_len) e <<= 1;      /* == 2 << Log2Floor(len-1) */
    return e;
}

/* tests/snappy/snappy.cc:91-97  HashBytes / Hash */
static inline uint32_t hash32(uint32_t word, int shift) { return (word * 0x1e35a7bdu) >> shift; }

/* tests/snappy/snappy.cc:337-378  EmitLiteral (both template arms write the
 * same bytes; the fast arm only over-copies slop past the literal) */
static size_t put_literal(uint8_t *out, size_t op, const uint8_t *src, uint32_t len) {
    uint32_t n = len - 1;
    if (n < 60) {
        out[op++] = (uint8_t)(n << 2);
    } else {
        int extra = 1;
        if (n >= (1u << 8)) extra = 2;
        if (n >= (1u << 16)) extra = 3;
        if (n >= (1u << 24)) extra = 4;
        out[op++] = (uint8_t)((59 + extra) << 2);
        for (int i = 0; i < extra; i++) out[op++] = (uint8_t)(n >> (8 * i));
    }
    memcpy(out + op, src, len);
    return op + len;
}

/* tests/snappy/snappy.cc:380-400  EmitCopyAtMost64 */
static size_t put_copy_le64(uint8_t *out, size_t op, uint32_t offset, uint32_t len, int total_lt12) {
    if (total_lt12 && offset < 2048) {
        out[op++] = (uint8_t)(1 | ((len - 4) << 2) | ((offset >> 8) << 5));
        out[op++] = (uint8_t)(offset & 0xff);
    } else {
        out[op++] = (uint8_t)(2 | ((len - 1) << 2));
        out[op++] = (uint8_t)(offset & 0xff);
        out[op++] = (uint8_t)(offset >> 8);
    }
    return op;
}

/* tests/snappy/snappy.cc:402-431  EmitCopy */
static size_t put_copy(uint8_t *out, size_t op, uint32_t offset, uint32_t len) {
    if (len < 12) return put_copy_le64(out, op, offset, len, 1);
    while (len >= 68) { op = put_copy_le64(out, op, offset, 64, 0); len -= 64; }
    if (len > 64)     { op = put_copy_le64(out, op, offset, 60, 0); len -= 60; }
    return put_copy_le64(out, op, offset, len, len < 12);
}

/* tests/snappy/snappy-internal.h:95-143  FindMatchLength: length of the
 * common prefix of in[a..] and in[b..], b bounded by end.  Only the length
 * matters to the caller (the "<8" flag equals n<8 on every path). */
static uint32_t common_prefix(const uint8_t *in, uint32_t a, uint32_t b, uint32_t end) {
    uint32_t n = 0;
    while (b + 8 <= end) {
        uint64_t x = rd64(in + a + n) ^ rd64(in + b);
        if (x) return n + (uint32_t)(__builtin_ctzll(x) >> 3);
        n += 8; b += 8;
    }
    while (b < end && in[a + n] == in[b]) { n++; b++; }
    return n;
}

/* Counters of loop-carried steps (SURVEY.md section 8a statistics). */
typedef struct { uint64_t probes, copies, chained, literals; } port_stats;

/* tests/snappy/snappy.cc:539-671  CompressFragment.
 * `table` must hold port_table_entries(len) zeroed uint16 (snappy.cc:475-481). */
size_t port_compress_fragment(const uint8_t *in, uint32_t len, uint8_t *out,
                              uint16_t *table, uint32_t entries, port_stats *st) {
    size_t op = 0;
    uint32_t ip = 0, next_emit = 0;
    int shift = 32 - __builtin_ctz(entries);            /* :548 */
    if (len >= PORT_MARGIN) {                            /* :557 */
        const uint32_t ip_limit = len - PORT_MARGIN;     /* :558 */
        ip = 1;                                          /* :560  Hash(++ip) */
        for (;;) {
            /* scan for a 4-byte match, accelerating skip (:587-608) */
            uint32_t skip = 32, cand, nxt = ip;
            do {
                ip = nxt;
                uint32_t step = skip >> 5;
                skip += step;
                nxt = ip + step;
                if (nxt > ip_limit) goto remainder;       /* :598-600 */
                uint32_t h = hash32(rd32(in + ip), shift);
                cand = table[h];
                table[h] = (uint16_t)ip;                  /* :606 */
                if (st) st->probes++;
            } while (rd32(in + ip) != rd32(in + cand));
            /* literal run before the match (:614) */
            if (st) st->literals++;
            op = put_literal(out, op, in + next_emit, ip - next_emit);
            /* copy loop (:627-656) */
            int first = 1;
            for (;;) {
                uint32_t base = ip;
                uint32_t matched = 4 + common_prefix(in, cand + 4, ip + 4, len);
                ip += matched;
                op = put_copy(out, op, base - cand, matched);
                if (st) { st->copies++; if (!first) st->chained++; }
                first = 0;
                next_emit = ip;
                if (ip >= ip_limit) goto remainder;       /* :643-645 */
                table[hash32(rd32(in + ip - 1), shift)] = (uint16_t)(ip - 1);   /* :649-651 */
                uint32_t h = hash32(rd32(in + ip), shift);
                cand = table[h];
                table[h] = (uint16_t)ip;                  /* :652-655 */
                if (rd32(in + ip) != rd32(in + cand)) break;   /* :656 */
            }
            ip++;                                         /* :658-659 */
        }
    }
remainder:
    if (next_emit < len) {                                /* :665-668 */
        if (st) st->literals++;
        op = put_literal(out, op, in + next_emit, len - next_emit);
    }
    return op;
}

/* tests/snappy/snappy.cc:1002-1070 Compress(Source*,Sink*) as driven by
 * RawCompress (:1404-1414): varint32 header, then independent <=64 KiB
 * fragments, each with a freshly zeroed table sized from its own length. */
size_t port_compress_stats(const uint8_t *in, size_t n, uint8_t *out, port_stats *st) {
    static __thread uint16_t table[PORT_MAX_TABLE];
    size_t op = port_varint32(out, (uint32_t)n);          /* :1006-1009 */
    size_t done = 0;
    while (done < n) {
        uint32_t len = (uint32_t)((n - done < PORT_BLOCK) ? (n - done) : PORT_BLOCK);   /* :1018 */
        uint32_t entries = port_table_entries(len);
        memset(table, 0, entries * sizeof(uint16_t));    /* :478 */
        op += port_compress_fragment(in + done, len, out + op, table, entries, st);
        done += len;
    }
    return op;
}

size_t port_compress(const uint8_t *in, size_t n, uint8_t *out) {
    return port_compress_stats(in, n, out, NULL);
}

/* Compress `count` independent buffers of `len` bytes at stride `in_stride`,
 * each into its own stream at out + i*out_stride; sizes[i] receives the
 * stream length.  This is RawCompress called once per buffer (config 4). */
void port_compress_batch(const uint8_t *in, size_t len, size_t in_stride, size_t count,
                         uint8_t *out, size_t out_stride, uint32_t *sizes) {
    for (size_t i = 0; i < count; i++)
        sizes[i] = (uint32_t)port_compress(in + i * in_stride, len, out + i * out_stride);
}

/* Minimal decompressor following tests/snappy/format_description.txt and the
 * tag semantics of snappy.cc:811-911 (DecompressAllTags).  Returns 0 on
 * success and writes *out_len; nonzero on malformed input. */
int port_uncompress(const uint8_t *src, size_t n, uint8_t *dst, size_t dst_cap, size_t *out_len) {
    size_t ip = 0; uint32_t ulen = 0; int sh = 0;
    for (;;) {
        if (ip >= n || sh > 28) return 1;
        uint8_t b = src[ip++];
        ulen |= (uint32_t)(b & 127) << sh;
        if (b < 128) break;
        sh += 7;
    }
    if (ulen > dst_cap) return 2;
    size_t op = 0;
    while (ip < n) {
        uint8_t tag = src[ip++];
        uint32_t kind = tag & 3, len, off;
        if (kind == 0) {
            len = (tag >> 2) + 1;
            if (len > 60) {
                uint32_t extra = len - 60;
                if (ip + extra > n) return 1;
                len = 0;
                for (uint32_t i = 0; i < extra; i++) len |= (uint32_t)src[ip + i] << (8 * i);
                len += 1; ip += extra;
            }
            if (ip + len > n || op + len > ulen) return 1;
            memcpy(dst + op, src + ip, len);
            ip += len; op += len;
            continue;
        }
        if (kind == 1) {
            if (ip + 1 > n) return 1;
            len = 4 + ((tag >> 2) & 7);
            off = ((uint32_t)(tag >> 5) << 8) | src[ip]; ip += 1;
        } else if (kind == 2) {
            if (ip + 2 > n) return 1;
            len = (tag >> 2) + 1;
            off = src[ip] | ((uint32_t)src[ip + 1] << 8); ip += 2;
        } else {
            if (ip + 4 > n) return 1;
            len = (tag >> 2) + 1;
            off = rd32(src + ip); ip += 4;
        }
        if (off == 0 || off > op || op + len > ulen) return 1;
        for (uint32_t i = 0; i < len; i++) dst[op + i] = dst[op - off + i];
        op += len;
    }
    if (op != ulen) return 1;
    *out_len = op;
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * Model of libsnappy_b200's FAST mode (SNAPPY_B200_FAST) -- NOT a restatement of the reference.
 * The fast mode is a parallel parse whose bytes may differ from the reference's; it is fully
 * deterministic, and this sequential model defines it so that tests can pin the CUDA kernel's
 * stream byte for byte (and measure its ratio against the reference within the stated 2 %).
 *
 *   candidates  the fragment is cut into aligned windows of 32 positions.  Every position p with
 *               p + 4 <= len is looked up in the same uint16 table / hash as the reference
 *               (snappy.cc:91-97, :445-456) as it stands after the PREVIOUS window, then all the
 *               window's positions are inserted (ascending, so the highest position of a slot
 *               stays): cand(p) = most recent position with p's hash in an earlier window (0 if none).
 *               Exception: if the 4 bytes at p - 1 equal the 4 bytes at p (a run), cand(p) = p - 1.
 *   hit(p)      p > 0 and the 4 bytes at cand(p) equal those at p; mlen(p) = common prefix,
 *               capped at 1024 and at len - p; long copies are split by EmitCopy (snappy.cc:402-431).
 *   parse       greedy: at p, if hit(p) emit literal [next_emit, p) + copy(p - cand, mlen) and
 *               continue at p + mlen, else p + 1.  No skip heuristic, no end margin.
 */
size_t port_fast_compress_fragment(const uint8_t *in, uint32_t len, uint8_t *out,
                                   uint16_t *table, uint32_t entries) {
    static __thread uint16_t cand[PORT_BLOCK];
    int shift = 32 - __builtin_ctz(entries);
    for (uint32_t w0 = 0; w0 < len; w0 += 32) {
        uint32_t w1 = (w0 + 32 < len) ? w0 + 32 : len;
        for (uint32_t p = w0; p < w1; p++)
            cand[p] = (p + 4 <= len) ? table[hash32(rd32(in + p), shift)] : 0xffff;
        for (uint32_t p = w0; p < w1; p++)
            if (p + 4 <= len) table[hash32(rd32(in + p), shift)] = (uint16_t)p;
    }
    size_t op = 0;
    uint32_t p = 1, next_emit = 0;
    while (p + 4 <= len) {
        uint32_t c = cand[p];
        if (rd32(in + p - 1) == rd32(in + p)) c = p - 1;      /* a run beats the table candidate */
        if (rd32(in + c) == rd32(in + p)) {
            uint32_t m = 4 + common_prefix(in, c + 4, p + 4, len);
            if (m > 1024) m = 1024;
            if (p > next_emit) op = put_literal(out, op, in + next_emit, p - next_emit);
            op = put_copy(out, op, p - c, m);
            p += m;
            next_emit = p;
        } else {
            p++;
        }
    }
    if (next_emit < len) op = put_literal(out, op, in + next_emit, len - next_emit);
    return op;
}

size_t port_fast_compress(const uint8_t *in, size_t n, uint8_t *out) {
    static __thread uint16_t table[PORT_MAX_TABLE];
    size_t op = port_varint32(out, (uint32_t)n);
    size_t done = 0;
    while (done < n) {
        uint32_t len = (uint32_t)((n - done < PORT_BLOCK) ? (n - done) : PORT_BLOCK);
        uint32_t entries = port_table_entries(len);
        memset(table, 0, entries * sizeof(uint16_t));
        op += port_fast_compress_fragment(in + done, len, out + op, table, entries);
        done += len;
    }
    return op;
}

/* ---- CRC-32C for the framing format (tests/snappy/framing_format.txt:38-57) -------------------------
 * RFC 3720 12.1: polynomial 0x1EDC6F41, reflected 0x82F63B78, initial value and final xor all ones.
 * Bit-at-a-time restatement (the checker need not be fast); pinned by the RFC's B.4 vectors in
 * tests/test_framing.py. */
uint32_t port_crc32c(const uint8_t *p, size_t n) {
    uint32_t crc = 0xffffffffu;
    for (size_t i = 0; i < n; ++i) {
        crc ^= p[i];
        for (int b = 0; b < 8; ++b) crc = (crc >> 1) ^ ((crc & 1u) ? 0x82F63B78u : 0u);
    }
    return ~crc;
}

uint32_t port_crc32c_masked(const uint8_t *p, size_t n) {       /* framing_format.txt:46-53 */
    const uint32_t x = port_crc32c(p, n);
    return ((x >> 15) | (x << 17)) + 0xa282ead8u;
}
