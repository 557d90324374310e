/*
 * b200_rocc.h -- the RoCC accelerator's top-level IO as a C-ABI "device" object, so that the
 * reference's Scala test harness can drive libsnappy_b200 the way it drives the Chisel design.
 *
 * What it mirrors (paths relative to the reference repository):
 *   src/main/scala/example/ScratchpadTest.scala:23-51  ScratchpadTestModule.io =
 *        { cmd: Flipped(Decoupled(RoCCCommand)), resp: Decoupled(RoCCResponse), busy, interrupt,
 *          exception, busyCycles: UInt(64) }  plus a TLTestRAM(AddressSet(0, 0xFFFFFF)) preloaded
 *        from a file (loadMemoryFromFile, :38).
 *   src/test/scala/example/Test.scala:17-53  the tester's protocol: poke funct=2 (SET_LENGTH,
 *        rs1=length), valid, step; poke funct=0 (COMPRESS, rs1=src, rs2=dst), step; valid=0;
 *        step until peek(resp.valid); peek(busyCycles), peek(resp.bits.data).
 *   tests/compression.c:13-27  the same two commands issued by the bare-metal driver.
 *
 * A "cycle" here is a bookkeeping unit: the command executes (H2D, kernels, D2H, blocking) on
 * the first step() after it was accepted; busy_cycles reports the device time of its kernels
 * at a nominal 1 GHz (1 cycle = 1 ns), which is what lands in the `cycles` column of the
 * benchmark CSVs (benchmark/hw_results.csv:1).
 */
#ifndef B200_ROCC_H_
#define B200_ROCC_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200_rocc b200_rocc;

/* funct values, tests/compression.c:13-15 and CompressionAccelerator.scala:362-389 */
enum { B200_ROCC_COMPRESS = 0, B200_ROCC_UNCOMPRESS = 1, B200_ROCC_SET_LENGTH = 2 };

/* A device with `mem_bytes` of backing memory at addresses [0, mem_bytes) (the TLTestRAM of
 * ScratchpadTest.scala:16; pass 0 for its 16 MiB).  NULL on failure. */
b200_rocc* b200_rocc_new(size_t mem_bytes, int cuda_device);
void b200_rocc_free(b200_rocc* dev);

/* loadMemoryFromFile (ScratchpadTest.scala:38): copy bytes / a raw data file to address addr.
 * Returns the number of bytes loaded, or -1. */
long b200_rocc_load_mem(b200_rocc* dev, const void* bytes, size_t n, uint64_t addr);
long b200_rocc_load_mem_file(b200_rocc* dev, const char* path, uint64_t addr);
/* Direct view of the backing memory (to read the compressed stream back). */
uint8_t* b200_rocc_mem(b200_rocc* dev);
size_t b200_rocc_mem_bytes(b200_rocc* dev);

/* poke(c.io.cmd.bits.inst.funct / rs1 / rs2) and poke(c.io.cmd.valid) (Test.scala:27-36). */
void b200_rocc_poke_cmd(b200_rocc* dev, uint32_t funct, uint64_t rs1, uint64_t rs2, int valid);
/* step(n): a valid command is accepted on a step when the unit is not busy
 * (cmd.ready := !busy, CompressionAccelerator.scala:72). */
void b200_rocc_step(b200_rocc* dev, uint64_t n);
/* peek(c.io.cmd.ready), peek(c.io.busy), peek(c.io.resp.valid), peek(c.io.resp.bits.data),
 * peek(c.io.busyCycles), peek(c.io.interrupt). */
int b200_rocc_peek_cmd_ready(b200_rocc* dev);
int b200_rocc_peek_busy(b200_rocc* dev);
int b200_rocc_peek_resp_valid(b200_rocc* dev);
uint64_t b200_rocc_peek_resp_data(b200_rocc* dev);
uint64_t b200_rocc_peek_busy_cycles(b200_rocc* dev);
int b200_rocc_peek_interrupt(b200_rocc* dev);

#ifdef __cplusplus
}
#endif
#endif /* B200_ROCC_H_ */
