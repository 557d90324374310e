/*
 * oracle/ref_shim.cc -- thin extern "C" entry points over the REFERENCE's own
 * Snappy (tests/snappy, compiled unmodified from /root/reference by
 * oracle/Makefile into oracle/_ref/libsnappy_ref.so).  This file is ours; it
 * contains no reference code, only calls into its public API
 * (tests/snappy/snappy.h: RawCompress, RawUncompress, MaxCompressedLength,
 * GetUncompressedLength, IsValidCompressedBuffer) in the same way
 * tests/swcompression.cc:43-45 drives it.
 *
 * TEST / BENCH INFRASTRUCTURE ONLY: the shipped library never links this.
 */
#include <stddef.h>
#include "snappy.h"

extern "C" {

size_t ref_max_compressed_length(size_t n) { return snappy::MaxCompressedLength(n); }

size_t ref_compress(const unsigned char* in, size_t n, unsigned char* out) {
    size_t out_n = 0;
    snappy::RawCompress(reinterpret_cast<const char*>(in), n, reinterpret_cast<char*>(out), &out_n);
    return out_n;
}

/* 0 on success */
int ref_uncompress(const unsigned char* in, size_t n, unsigned char* out) {
    return snappy::RawUncompress(reinterpret_cast<const char*>(in), n, reinterpret_cast<char*>(out)) ? 0 : 1;
}

int ref_uncompressed_length(const unsigned char* in, size_t n, size_t* result) {
    return snappy::GetUncompressedLength(reinterpret_cast<const char*>(in), n, result) ? 0 : 1;
}

int ref_is_valid(const unsigned char* in, size_t n) {
    return snappy::IsValidCompressedBuffer(reinterpret_cast<const char*>(in), n) ? 1 : 0;
}

void ref_compress_batch(const unsigned char* in, size_t len, size_t in_stride, size_t count,
                        unsigned char* out, size_t out_stride, unsigned int* sizes) {
    for (size_t i = 0; i < count; i++)
        sizes[i] = (unsigned int)ref_compress(in + i * in_stride, len, out + i * out_stride);
}

}  // extern "C"

#define CPU_COMPRESS(in, n, out) ref_compress((in), (n), (out))
#define CPU_PREFIX(name) ref_##name
#include "cpu_bench.inc"
