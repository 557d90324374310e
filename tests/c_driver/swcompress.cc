// A C++ host program in the shape of the reference's tests/swcompression.cc: a std::string in, a
// std::string out, through the reference's own call shape with the namespace switched to the B200
// library (include/snappy_b200.hpp).  With a file argument it compresses that file instead of the test
// sentence (BASELINE config "data/fox-in-socks.txt compressed ... via tests/swcompression.cc").
//
//   g++ -std=c++11 -I include tests/c_driver/swcompress.cc -L compression-accelerator_b200 -lsnappy_b200
//
// Prints "<input bytes> <compressed bytes> <fnv1a32 of the stream> <round trip ok>"; exit 0 on success.
#include <cstdint>
#include <cstdio>
#include <fstream>
#include <iterator>
#include <string>

#include "snappy_b200.hpp"

int main(int argc, char* argv[]) {
    std::string input = "If Tid the thistle sifter lifts his sieve of sifted thistles, then he'll shift the "
                        "thistles in his sifting sieve to live unsifted.";
    if (argc > 1) {
        std::ifstream f(argv[1], std::ios::binary);
        if (!f) { std::fprintf(stderr, "cannot read %s\n", argv[1]); return 2; }
        input.assign(std::istreambuf_iterator<char>(f), std::istreambuf_iterator<char>());
    }
    std::string output, back;
    const size_t compressed_length = snappy_b200::Compress(input.data(), input.size(), &output);
    if (compressed_length == 0) {
        std::fprintf(stderr, "Compress failed: %s\n", snappy_b200_last_error());
        return 1;
    }
    std::uint32_t h = 2166136261u;                    // FNV-1a, the checksum SURVEY.md 8c quotes
    for (size_t i = 0; i < output.size(); ++i) h = (h ^ (unsigned char)output[i]) * 16777619u;
    const bool ok = snappy_b200::Uncompress(output.data(), output.size(), &back) && back == input;
    std::printf("%zu %zu %08x %s\n", input.size(), compressed_length, h, ok ? "ok" : "MISMATCH");
    return ok ? 0 : 1;
}
