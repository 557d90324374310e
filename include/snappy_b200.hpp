// snappy_b200.hpp -- the std::string call shapes of the reference's software Snappy
// (tests/snappy/snappy.h:76 Compress, :84 Uncompress, :163 MaxCompressedLength, :169
// GetUncompressedLength, as called by tests/swcompression.cc:45 and tests/snappy/snappy_unittest.cc),
// header-only over the C-ABI of snappy_b200.h.  A call site switches by changing its namespace:
//
//     size_t n = snappy::Compress(in.data(), in.size(), &out);       // reference
//     size_t n = snappy_b200::Compress(in.data(), in.size(), &out);  // B200, same bytes (exact mode)
//
// There is no CPU path behind these: without a CUDA device Compress returns 0 and Uncompress false.
#ifndef SNAPPY_B200_HPP_
#define SNAPPY_B200_HPP_

#include <string>

#include "snappy_b200.h"

namespace snappy_b200 {

inline size_t MaxCompressedLength(size_t source_bytes) { return snappy_b200_max_compressed_length(source_bytes); }

// snappy.cc:1416-1425: resizes *output to the compressed length and returns it (0 on failure; a valid
// stream is never empty).
inline size_t Compress(const char* input, size_t input_length, std::string* output,
                       int mode = SNAPPY_B200_EXACT, snappy_b200_ctx* ctx = nullptr) {
    output->resize(MaxCompressedLength(input_length));
    size_t n = 0;
    if (snappy_b200_compress(ctx, input, input_length, &(*output)[0], output->size(), &n, mode) != SNAPPY_B200_OK) n = 0;
    output->resize(n);
    return n;
}

// snappy.cc:433: false on a malformed preamble.
inline bool GetUncompressedLength(const char* compressed, size_t compressed_length, size_t* result) {
    return snappy_b200_uncompressed_length(compressed, compressed_length, result) == SNAPPY_B200_OK;
}

// snappy.cc:1348-1363: false on malformed input; *uncompressed holds the data on success.
inline bool Uncompress(const char* compressed, size_t compressed_length, std::string* uncompressed,
                       snappy_b200_ctx* ctx = nullptr) {
    size_t ulen = 0;
    if (!GetUncompressedLength(compressed, compressed_length, &ulen)) return false;
    uncompressed->resize(ulen);
    size_t n = 0;
    char dummy = 0;
    char* dst = ulen ? &(*uncompressed)[0] : &dummy;
    if (snappy_b200_uncompress(ctx, compressed, compressed_length, dst, ulen, &n) != SNAPPY_B200_OK) return false;
    return n == ulen;
}

}  // namespace snappy_b200

#endif  // SNAPPY_B200_HPP_
