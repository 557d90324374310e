/* A C host program in the shape of the reference's RoCC driver (tests/compression.c:41-61 of the
 * reference): one buffer through compress(source, uncompressed_length, dest) and back through
 * uncompress(source, compressed_length, dest), both bound to libsnappy_b200 instead of custom3
 * instructions.  Written for this repository's tests; the text is the
 * reference's 130-byte test sentence because its 112-byte stream is a known answer.
 *
 *   gcc -std=c99 -I include tests/c_driver/rocc_driver.c -L compression-accelerator_b200 -lsnappy_b200
 *
 * Prints one line per stage and the stream in hex; exit code 0 when the round trip matches. */
#include <stdio.h>
#include <string.h>

#include "snappy_b200.h"

static unsigned long compress(char* source, int uncompressed_length, char* dest)
{
    return snappy_b200_rocc_compress(source, uncompressed_length, dest);
}

static unsigned long uncompress(char* source, int compressed_length, char* dest)
{
    return snappy_b200_rocc_uncompress(source, compressed_length, dest);
}

int main(void)
{
    static const char text[] =
        "If Tid the thistle sifter lifts his sieve of sifted thistles, then he'll shift the "
        "thistles in his sifting sieve to live unsifted.";
    char source[256], stream[256], back[256];
    const int n = (int)strlen(text);
    unsigned long got, failed;
    int i;

    memset(source, 0, sizeof source);
    memset(stream, 0, sizeof stream);
    memset(back, 0, sizeof back);
    memcpy(source, text, (size_t)n);
    if ((size_t)n > sizeof stream - 64 || snappy_b200_max_compressed_length((size_t)n) > sizeof stream) {
        fprintf(stderr, "buffers too small\n");
        return 2;
    }
    got = compress(source, n, stream);
    printf("compress: %d -> %lu bytes\n", n, got);
    if (got == 0) {
        fprintf(stderr, "compress failed: %s\n", snappy_b200_last_error());
        return 1;
    }
    printf("stream:");
    for (i = 0; i < (int)got; ++i) printf(" %02x", (unsigned char)stream[i]);
    printf("\n");
    failed = uncompress(stream, (int)got, back);
    printf("uncompress: result %lu, %s\n", failed, memcmp(back, source, (size_t)n) == 0 ? "identical" : "DIFFERENT");
    return failed == 0 && memcmp(back, source, (size_t)n) == 0 ? 0 : 1;
}
