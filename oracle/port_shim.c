/*
 * oracle/port_shim.c -- multi-threaded timing driver over the restatement in
 * snappy_port.c.  TEST / BENCH INFRASTRUCTURE ONLY.
 */
#include <stddef.h>
size_t port_compress(const unsigned char *in, size_t n, unsigned char *out);
#define CPU_COMPRESS(in, n, out) port_compress((in), (n), (out))
#define CPU_PREFIX(name) port_##name
#include "cpu_bench.inc"
