/*
 * snappy_b200.h -- C-ABI of libsnappy_b200: a B200 (sm_100a) Snappy block compressor that
 * stands in for the RoCC compression accelerator of nullromo/compression-accelerator.
 *
 * Every entry point cites the reference interface it replaces (paths relative to the
 * reference repository).  Plain pointers and sizes only; no torch / C++ types.
 *
 * Streams produced are standard Snappy raw streams (tests/snappy/format_description.txt):
 * varint32(uncompressed length) followed by the elements of independent <=64 KiB fragments.
 * In SNAPPY_B200_EXACT mode the bytes are identical to the reference's software Snappy
 * (tests/snappy/snappy.cc:1002-1070 as driven by tests/swcompression.cc:45).
 */
#ifndef SNAPPY_B200_H_
#define SNAPPY_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Status codes: first three mirror tests/snappy/snappy-c.h:46-50 (snappy_status). */
typedef enum {
    SNAPPY_B200_OK = 0,
    SNAPPY_B200_INVALID_INPUT = 1,
    SNAPPY_B200_BUFFER_TOO_SMALL = 2,
    SNAPPY_B200_CUDA_ERROR = 3,
    SNAPPY_B200_NO_DEVICE = 4
} snappy_b200_status;

/* Parse mode.  EXACT keeps Snappy's greedy sequential parse and skip heuristic per fragment
 * (tests/snappy/snappy.cc:539-671) and is byte-identical to the reference.  FAST is a
 * parallel parse: always a valid Snappy stream, bytes may differ. */
typedef enum { SNAPPY_B200_EXACT = 0, SNAPPY_B200_FAST = 1 } snappy_b200_mode;
/* The `mode` arguments also accept diagnostic selectors that force ONE of the exact engines (tests, A/B
 * runs; every one of them produces the reference's bytes; not part of the contract and subject to change):
 *   2 serial warp parser, 3 single-warp window parser, 4 pipelined parser with the fragment staged by TMA,
 *   5 lane parser (one thread per fragment, fat table entries in global memory),
 *   6 pipelined parser (one CTA per fragment, table in shared memory),
 *   7 window parser with its table in global memory (one warp per fragment).
 * SNAPPY_B200_EXACT picks among 5, 6, 7 (and the page / small-buffer kernels of the batch entry) by shape. */

typedef struct snappy_b200_ctx snappy_b200_ctx; /* owns one CUDA stream + device/pinned scratch */

/* tests/snappy/snappy.cc:99-121 MaxCompressedLength / snappy-c.h:107 = 32 + n + n/6 */
size_t snappy_b200_max_compressed_length(size_t n);

/* Human-readable text of the last failure on the calling thread ("" if none). */
const char* snappy_b200_last_error(void);

/* Context lifetime.  device = CUDA ordinal.  One context may be used by one host thread at a
 * time; independent contexts are independent (the RoCC unit accepts one command at a time,
 * CompressionAccelerator.scala:72; software Snappy is re-entrant, snappy.cc:1011). */
snappy_b200_status snappy_b200_create(snappy_b200_ctx** ctx, int device);
void snappy_b200_destroy(snappy_b200_ctx* ctx);
/* The cudaStream_t the context launches on (as void*), for callers that order work around it. */
void* snappy_b200_stream(snappy_b200_ctx* ctx);

/* ---- host-pointer entry: replaces snappy::Compress / RawCompress (snappy.cc:1404-1425) and
 * the RoCC compress command (tests/compression.c:20-27).  Synchronous: H2D, kernels, D2H.
 * ctx may be NULL (a per-thread default context on device 0 is used).  n must be < 2^32
 * (Snappy's header is a varint32, snappy.cc:1006-1007). */
snappy_b200_status snappy_b200_compress(snappy_b200_ctx* ctx, const void* src, size_t n,
                                        void* dst, size_t dst_cap, size_t* out_n, int mode);

/* ---- device-pointer entry: same stream, input and output already in HBM.  Asynchronous on
 * the context's stream; *d_out_n (device uint64) receives the stream length.  dst_cap must be
 * >= snappy_b200_max_compressed_length(n). */
snappy_b200_status snappy_b200_compress_device(snappy_b200_ctx* ctx, const void* d_src, size_t n,
                                               void* d_dst, size_t dst_cap,
                                               unsigned long long* d_out_n, int mode);

/* ---- no-allocation entry.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Exact mode, one stream of n <= SNAPPY_B200_NOALLOC_MAX_INPUT
 * bytes (1 MiB), byte-identical to snappy_b200_compress_device; asynchronous on `cuda_stream` (a
 * cudaStream_t, may be 0).  Needs no context and allocates nothing: its scratch (sequence lists,
 * counts, sizes, offsets; about 2 MiB) is a static device array that every GPU loading the library
 * holds until exit.  All calls on a device share that array: calls that overlap on the GPU race --
 * serialise them (one stream, or events).  n beyond the maximum -> SNAPPY_B200_INVALID_INPUT with
 * snappy_b200_last_error() naming the limit.  dst_cap >= snappy_b200_max_compressed_length(n). */
#define SNAPPY_B200_NOALLOC_MAX_INPUT ((size_t)1 << 20)
size_t snappy_b200_noalloc_max_input(void);
snappy_b200_status snappy_b200_compress_device_noalloc(const void* d_src, size_t n, void* d_dst,
                                                       size_t dst_cap, unsigned long long* d_out_n,
                                                       void* cuda_stream);

/* ---- fragment-range entry (multi-GPU building block): compress n bytes that are a run of
 * whole 64 KiB fragments of a larger stream (only the last rank's run may end short) WITHOUT
 * the varint header; the payload lands contiguous at d_dst, its size in *d_out_n.  Rank r's
 * payload is placed at header_len + sum of earlier ranks' sizes (snappy.cc:1013-1061). */
snappy_b200_status snappy_b200_compress_fragments_device(snappy_b200_ctx* ctx, const void* d_src,
                                                         size_t n, void* d_dst, size_t dst_cap,
                                                         unsigned long long* d_out_n, int mode);

/* ---- batch entry: `count` independent buffers of `len` bytes (len <= 65536) at
 * d_src + i*in_stride, each compressed into ITS OWN Snappy stream at d_dst + i*out_stride
 * (out_stride >= max_compressed_length(len)); d_sizes[i] (device uint32) = stream length.
 * Equivalent to calling RawCompress once per buffer (snappy.cc:1404-1414); the page-shaped
 * workload of benchmark/hw_results.csv.  Asynchronous on the context's stream. */
snappy_b200_status snappy_b200_compress_batch_device(snappy_b200_ctx* ctx, const void* d_src,
                                                     size_t len, size_t in_stride, size_t count,
                                                     void* d_dst, size_t out_stride,
                                                     uint32_t* d_sizes, int mode);
/* Host-pointer flavour of the batch entry (synchronous). */
snappy_b200_status snappy_b200_compress_batch(snappy_b200_ctx* ctx, const void* src, size_t len,
                                              size_t in_stride, size_t count, void* dst,
                                              size_t out_stride, uint32_t* sizes, int mode);

/* ---- decompression: replaces snappy::RawUncompress (snappy.cc:1338-1346) and the RoCC uncompress
 * command (tests/compression.c:32-39).  A stream is decoded by one warp (element boundaries can
 * only be found by walking the stream), so throughput comes from many streams per call.  Malformed
 * input is detected, never dereferenced out of bounds (as snappy_unittest.cc:536- demands). */
/* Parse the varint32 preamble (snappy.cc GetUncompressedLength); host memory. */
snappy_b200_status snappy_b200_uncompressed_length(const void* src, size_t n, size_t* result);
/* Host pointers, synchronous.  dst_cap >= the stream's uncompressed length; *out_n = bytes written.
 * SNAPPY_B200_INVALID_INPUT for a malformed stream. */
snappy_b200_status snappy_b200_uncompress(snappy_b200_ctx* ctx, const void* src, size_t n, void* dst,
                                          size_t dst_cap, size_t* out_n);
/* Device pointers, asynchronous on the context's stream: *d_out_n = length, or ~0 if malformed.
 * Streams of 128 KiB and more are decoded block-parallel when they are block structured (an element
 * starts at every multiple of 65,536 output bytes and no copy reaches across one -- what the
 * reference compressor and this library produce); any other stream takes the serial walk, decided on
 * the device.  The result is the same either way. */
snappy_b200_status snappy_b200_uncompress_device(snappy_b200_ctx* ctx, const void* d_src, size_t n,
                                                 void* d_dst, size_t dst_cap, unsigned long long* d_out_n);
/* `count` streams: stream i = d_src + i*in_stride, d_sizes[i] bytes, decoded to d_dst + i*out_stride
 * (capacity out_stride each); d_out_lens[i] = decoded length or 0xffffffff if malformed. */
snappy_b200_status snappy_b200_uncompress_batch_device(snappy_b200_ctx* ctx, const void* d_src,
                                                       size_t in_stride, const uint32_t* d_sizes,
                                                       size_t count, void* d_dst, size_t out_stride,
                                                       uint32_t* d_out_lens);

/* ---- Snappy framing format ------------------------------------------------------------------------
 * tests/snappy/framing_format.txt: stream identifier (:68-75), then chunks of at most 65,536
 * uncompressed bytes (:88-91): type, 3-byte length, masked CRC-32C of the uncompressed bytes (:38-57),
 * an ordinary Snappy stream of the chunk (:80-86).  The reference vendors the specification only (no
 * writer or reader), so these entries have no reference function to mirror; they are SURVEY.md 8(f)
 * rank 2.  Chunks are independent, so -- unlike a raw stream -- a framed stream of any size decodes
 * in parallel, and every chunk is checked against its CRC.  The writer emits compressed chunks
 * (type 0x00) of 65,536 bytes (the last one shorter); the reader also accepts uncompressed (0x01),
 * padding / skippable (0x80-0xfe) and repeated identifier chunks, and rejects 0x02-0x7f (:122-128). */
size_t snappy_b200_frame_max_length(size_t n);
/* host pointers, synchronous */
snappy_b200_status snappy_b200_frame_compress(snappy_b200_ctx* ctx, const void* src, size_t n, void* dst,
                                              size_t dst_cap, size_t* out_n, int mode);
snappy_b200_status snappy_b200_frame_uncompress(snappy_b200_ctx* ctx, const void* src, size_t n, void* dst,
                                                size_t dst_cap, size_t* out_n);
/* total uncompressed bytes of a framed stream in host memory (header walk only) */
snappy_b200_status snappy_b200_frame_uncompressed_length(const void* src, size_t n, size_t* result);
/* device pointers.  compress: asynchronous on the context's stream, *d_out_n (device) = stream length.
 * uncompress: synchronous (the status depends on every chunk's CRC), *out_n is a host word. */
snappy_b200_status snappy_b200_frame_compress_device(snappy_b200_ctx* ctx, const void* d_src, size_t n,
                                                     void* d_dst, size_t dst_cap,
                                                     unsigned long long* d_out_n, int mode);
snappy_b200_status snappy_b200_frame_uncompress_device(snappy_b200_ctx* ctx, const void* d_src, size_t n,
                                                       void* d_dst, size_t dst_cap, size_t* out_n);
/* masked CRC-32C (framing_format.txt:38-57; RFC 3720 12.1) of consecutive chunk_len-byte chunks
 * (chunk_len <= 65536, last one shorter) of a device buffer; d_masked[i] for chunk i; asynchronous. */
snappy_b200_status snappy_b200_crc32c_chunks_device(snappy_b200_ctx* ctx, const void* d_src, size_t n,
                                                    size_t chunk_len, uint32_t* d_masked);

/* Write varint32(n) at dst (host memory), return its length 1..5
 * (tests/snappy/snappy-stubs-internal.h:552-578). */
size_t snappy_b200_varint32(void* dst, uint32_t n);

/* Block until everything queued on the context's stream is done. */
snappy_b200_status snappy_b200_sync(snappy_b200_ctx* ctx);

/* Pinned host memory helpers (so callers can hand the host entry DMA-able buffers). */
void* snappy_b200_host_alloc(size_t n);
void snappy_b200_host_free(void* p);

/* Kernel launch counters of the context (for bench.py's gpu_launches claim). */
unsigned long long snappy_b200_launch_count(snappy_b200_ctx* ctx);

/* Device time in ms of the last host-entry call's kernels (CUDA events), 0 if none. */
float snappy_b200_last_kernel_ms(snappy_b200_ctx* ctx);

/* Device time in ms (CUDA events on the context's stream) of the stages of the most recent
 * device/batch call: stage 0 = fragment parse kernel, stage 1 = size scan + gather.  Valid
 * after snappy_b200_sync(). */
float snappy_b200_stage_ms(snappy_b200_ctx* ctx, int stage);

/* Diagnostics: 16 counters of the pipelined parser since the last reset (all zero unless the
 * library was built with -DSNAPPY_B200_STATS).  Returns 0 on success. */
int snappy_b200_debug_stats(unsigned long long* out16, int reset);

/* ---- RoCC driver mirror: tests/compression.c:20-27 and :32-39.  Host pointers, blocking.
 * compress returns the compressed length (0 on failure: a valid stream is never empty);
 * uncompress returns 0 on success, non-zero on failure, as the reference documents. */
unsigned long snappy_b200_rocc_compress(char* source, int uncompressed_length, char* dest);
unsigned long snappy_b200_rocc_uncompress(char* source, int compressed_length, char* dest);

/* ---- multi-GPU exchange over NVLink peer memory (one process per GPU).  The only
 * datum that crosses a fragment or stream boundary is the running output offset
 * (tests/snappy/snappy.cc:1061), so a step exchanges every rank's payload size and lands the
 * payloads back to back in an archive on the root.  Mailboxes and the archive (two slots, steps
 * alternate) live in device memory shared through CUDA IPC; sizes, offsets and payload bytes move by
 * kernels of this library over peer pointers -- no host round trip, no collective library.
 *   create   allocates this rank's side and writes snappy_b200_peer_handle_bytes() bytes of IPC
 *            handles to handles_out; the caller exchanges them by any means (rank order, concatenated)
 *   connect  maps the peers' mailboxes and the root's archive
 *   gather   queues one step behind the work already on ctx's stream: payload d_payload[0, *d_size)
 *            -> root archive slot (step & 1), behind `base` bytes the root keeps for itself (the
 *            varint preamble when the ranks hold runs of fragments of ONE stream, else 0) and the
 *            payloads of the ranks below.  Steps are numbered 0, 1, 2, ... on every rank.
 *   reuse    makes ctx's stream wait until step's payload has left its buffer (before overwriting it)
 *   wait     queues a wait for every rank's step to be complete (then the slot may be read)
 *   result   blocks the host on the last wait; total bytes in the slot and each rank's offset
 * Every device-side wait is bounded (about 20 s) and reports through result() instead of hanging. */
typedef struct snappy_b200_peer snappy_b200_peer;
size_t snappy_b200_peer_handle_bytes(void);
const char* snappy_b200_peer_last_error(void);
snappy_b200_status snappy_b200_peer_create(snappy_b200_ctx* ctx, int device, int rank, int world, int root,
                                           size_t archive_slot_bytes, void* handles_out,
                                           snappy_b200_peer** out);
snappy_b200_status snappy_b200_peer_connect(snappy_b200_peer* peer, const void* all_handles);
snappy_b200_status snappy_b200_peer_gather(snappy_b200_peer* peer, const void* d_payload,
                                           const unsigned long long* d_size, unsigned long long step,
                                           unsigned long long base);
snappy_b200_status snappy_b200_peer_reuse(snappy_b200_peer* peer, unsigned long long step);
snappy_b200_status snappy_b200_peer_wait(snappy_b200_peer* peer, unsigned long long step);
snappy_b200_status snappy_b200_peer_result(snappy_b200_peer* peer, unsigned long long step,
                                           unsigned long long base, unsigned long long* total,
                                           unsigned long long* offsets);
void* snappy_b200_peer_archive(snappy_b200_peer* peer, int slot);   /* device pointer (mapped on non-root ranks) */
void* snappy_b200_peer_stream(snappy_b200_peer* peer);              /* the exchange cudaStream_t */
unsigned long long snappy_b200_peer_launch_count(snappy_b200_peer* peer);
void snappy_b200_peer_destroy(snappy_b200_peer* peer);

#ifdef __cplusplus
}
#endif
#endif /* SNAPPY_B200_H_ */
