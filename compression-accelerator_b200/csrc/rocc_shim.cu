// rocc_shim.cu -- include/b200_rocc.h: the accelerator's cmd/resp/busy/busyCycles IO and its
// test RAM as a host-side object over libsnappy_b200 (ScratchpadTest.scala:23-51, Test.scala:17-53).
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <new>

#include "../../include/b200_rocc.h"
#include "../../include/snappy_b200.h"

struct b200_rocc {
    snappy_b200_ctx* ctx = nullptr;
    uint8_t* mem = nullptr;          // pinned host memory: the TLTestRAM
    size_t mem_bytes = 0;
    // poked command wires
    uint32_t funct = 0;
    uint64_t rs1 = 0, rs2 = 0;
    int cmd_valid = 0;
    // architectural state (CompressionAccelerator.scala:63-67 length/src/dst registers)
    uint64_t length = 0;
    int resp_valid = 0;
    uint64_t resp_data = 0;
    uint64_t busy_cycles = 0;
    uint64_t steps = 0;
};

extern "C" {

b200_rocc* b200_rocc_new(size_t mem_bytes, int cuda_device) {
    b200_rocc* d = new (std::nothrow) b200_rocc;
    if (!d) return nullptr;
    if (snappy_b200_create(&d->ctx, cuda_device) != SNAPPY_B200_OK) { delete d; return nullptr; }
    d->mem_bytes = mem_bytes ? mem_bytes : (size_t)0x1000000;   // AddressSet(0, 0xFFFFFF)
    d->mem = (uint8_t*)snappy_b200_host_alloc(d->mem_bytes);
    if (!d->mem) { snappy_b200_destroy(d->ctx); delete d; return nullptr; }
    memset(d->mem, 0, d->mem_bytes);
    return d;
}

void b200_rocc_free(b200_rocc* d) {
    if (!d) return;
    if (d->mem) snappy_b200_host_free(d->mem);
    if (d->ctx) snappy_b200_destroy(d->ctx);
    delete d;
}

long b200_rocc_load_mem(b200_rocc* d, const void* bytes, size_t n, uint64_t addr) {
    if (!d || (!bytes && n) || addr > d->mem_bytes || n > d->mem_bytes - addr) return -1;
    memcpy(d->mem + addr, bytes, n);
    return (long)n;
}

long b200_rocc_load_mem_file(b200_rocc* d, const char* path, uint64_t addr) {
    if (!d || !path || addr > d->mem_bytes) return -1;
    FILE* f = fopen(path, "rb");
    if (!f) return -1;
    size_t n = fread(d->mem + addr, 1, d->mem_bytes - addr, f);
    fclose(f);
    return (long)n;
}

uint8_t* b200_rocc_mem(b200_rocc* d) { return d ? d->mem : nullptr; }
size_t b200_rocc_mem_bytes(b200_rocc* d) { return d ? d->mem_bytes : 0; }

void b200_rocc_poke_cmd(b200_rocc* d, uint32_t funct, uint64_t rs1, uint64_t rs2, int valid) {
    if (!d) return;
    d->funct = funct; d->rs1 = rs1; d->rs2 = rs2; d->cmd_valid = valid;
}

static void accept(b200_rocc* d) {
    if (d->funct == B200_ROCC_SET_LENGTH) {            // CompressionAccelerator.scala:362-366
        d->length = d->rs1;
        return;
    }
    d->resp_valid = 0;
    if (d->funct == B200_ROCC_COMPRESS) {              // :367-384: src = rs1, dst = rs2
        const uint64_t src = d->rs1, dst = d->rs2, n = d->length;
        size_t out_n = 0;
        uint64_t result = 0;
        if (src <= d->mem_bytes && n <= d->mem_bytes - src && dst <= d->mem_bytes) {
            const size_t cap = d->mem_bytes - dst;
            if (cap >= snappy_b200_max_compressed_length(n) &&
                snappy_b200_compress(d->ctx, d->mem + src, n, d->mem + dst, cap, &out_n,
                                     SNAPPY_B200_EXACT) == SNAPPY_B200_OK)
                result = out_n;
        }
        d->resp_data = result;                          // 0 = failure (a stream is never empty)
        d->resp_valid = 1;
        d->busy_cycles += (uint64_t)(snappy_b200_last_kernel_ms(d->ctx) * 1e6f);   // 1 cycle = 1 ns
    } else if (d->funct == B200_ROCC_UNCOMPRESS) {     // :385-388 (a stub in the reference hardware)
        // Same operand checks as COMPRESS: source, length, destination and the length the stream's own
        // preamble announces must all lie inside the device's memory (a poked address or a crafted stream
        // must not reach host memory outside it).  Non-zero = failure (tests/compression.c:32-39).
        const uint64_t src = d->rs1, dst = d->rs2, n = d->length;
        uint64_t result = 1;
        size_t ulen = 0, out_n = 0;
        if (src <= d->mem_bytes && n >= 1 && n <= d->mem_bytes - src && dst <= d->mem_bytes &&
            snappy_b200_uncompressed_length(d->mem + src, (size_t)n, &ulen) == SNAPPY_B200_OK &&
            ulen <= d->mem_bytes - dst &&
            snappy_b200_uncompress(d->ctx, d->mem + src, (size_t)n, d->mem + dst, d->mem_bytes - dst, &out_n) ==
                SNAPPY_B200_OK)
            result = 0;
        d->resp_data = result;
        d->resp_valid = 1;
    }
}

void b200_rocc_step(b200_rocc* d, uint64_t n) {
    if (!d) return;
    for (uint64_t i = 0; i < n; ++i) {
        ++d->steps;
        if (d->cmd_valid) accept(d);                    // executes to completion: never busy across steps
        if (!d->cmd_valid) break;                       // nothing changes on idle cycles
    }
    d->steps += 0;
}

int b200_rocc_peek_cmd_ready(b200_rocc* d) { return d ? 1 : 0; }
int b200_rocc_peek_busy(b200_rocc* d) { (void)d; return 0; }
int b200_rocc_peek_resp_valid(b200_rocc* d) { return d ? d->resp_valid : 0; }
uint64_t b200_rocc_peek_resp_data(b200_rocc* d) { return d ? d->resp_data : 0; }
uint64_t b200_rocc_peek_busy_cycles(b200_rocc* d) { return d ? d->busy_cycles : 0; }
int b200_rocc_peek_interrupt(b200_rocc* d) { (void)d; return 0; }   // io.interrupt := false (:410)

}  // extern "C"
