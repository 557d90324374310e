// kernels.h -- host-side launchers of the Snappy kernels (internal to libsnappy_b200).
//
// Pipeline of one call (all on one stream):
//   parse   one warp-group per fragment: greedy parse -> list of sequences (literal run, copy)
//   size    encoded size of every fragment from its sequence list
//   scan    exclusive scan of sizes (+ varint header) -> output offset of every fragment
//   emit    sequences + input -> Snappy elements written ONCE at their final position
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace sb {

// A sequence = literal run followed by one copy (the unit EmitLiteral + EmitCopy produce per
// match in tests/snappy/snappy.cc:614-641).  Packed in a uint64:
//   bits  0..16  lit_len   (0..65536)   bytes of literal before the copy
//   bits 17..32  copy_len  (0, 4..65535) 0 = no copy (the fragment's trailing literal)
//   bits 33..48  offset    (1..65535)
typedef unsigned long long seq_t;
__host__ __device__ inline seq_t seq_pack(uint32_t lit, uint32_t len, uint32_t off) {
    return (seq_t)lit | ((seq_t)len << 17) | ((seq_t)off << 33);
}
__host__ __device__ inline uint32_t seq_lit(seq_t s) { return (uint32_t)s & 0x1ffffu; }
__host__ __device__ inline uint32_t seq_len(seq_t s) { return (uint32_t)(s >> 17) & 0xffffu; }
__host__ __device__ inline uint32_t seq_off(seq_t s) { return (uint32_t)(s >> 33) & 0xffffu; }

// nseq[job] of a job whose stream the parse kernel has already written itself (batch path: a page that
// parsed to ONE literal is emitted from the staged copy in shared memory, kernels.cu parse_kernel); the
// size and emit kernels leave such a job alone (sizes[job] is final).
constexpr uint32_t kSeqEmitted = 0xffffffffu;

// Sequences a job of `job_len` input bytes can produce: every copy consumes >= 4 bytes.
__host__ __device__ inline uint32_t seq_stride(uint32_t job_len) { return job_len / 4 + 8; }

// One "job" = one fragment of a stream, or one small buffer of a batch.
struct JobParams {
    const uint8_t* in;     // job j reads in + j*in_stride
    uint64_t n_total;      // bytes reachable from `in` (job length = min(job_len, n_total - j*in_stride))
    uint64_t in_stride;
    uint32_t job_len;      // 65536 in the stream path, buffer length in the batch path
    uint32_t n_jobs;
    seq_t* seqs;           // job j's sequences at seqs + j*seq_stride(job_len)
    uint32_t* nseq;        // number of sequences of job j
    uint32_t* sizes;       // encoded bytes of job j (header included when with_header)
    uint32_t with_header;  // 1: each job is its own stream and starts with varint32(len)
    uint32_t sparse;       // 1: seqs holds the fast mode's position-indexed copy slots (parse_fast.cuh)
    // emit destination: job j writes at out + (offsets ? offsets[j] : j*out_stride)
    uint8_t* out;
    uint64_t out_stride;
    const unsigned long long* offsets;
    unsigned int* err = nullptr;   // the context's error word (mapped host memory): a parser watchdog stores 1 here
    uint32_t fuse_emit = 0;        // batch path, page parser: the warp that parsed a job writes its elements too
                                   // (sizes[j] final, nseq[j] = kSeqEmitted; no size / emit launch follows)
};

cudaError_t launch_parse(const JobParams& p, int mode, cudaStream_t st, unsigned long long* launches);
bool page_parser_takes(uint32_t job_len);   // exact mode: jobs of this length go to the one-warp page parser (parse_kernel<true>)

// Lane parser (parse_lane.cuh): one thread per job, all jobs in flight, "fat" hash-table entries in a
// pool in global memory that is never cleared between launches (entries carry the launch's epoch).
// Writes seqs, nseq AND sizes (no size pass).  pool must hold lane_pool_bytes(n_jobs, job_len) bytes.
constexpr uint32_t kLaneEntryBytes = 16;
size_t lane_pool_bytes(uint64_t n_jobs, uint32_t job_len);
// beside_pipe: the launch shares the GPU with the pipelined parser's CTAs, so it asks for their (maximum)
// shared-memory carve-out -- an SM cannot change its split while CTAs are resident.
cudaError_t launch_parse_lane(const JobParams& p, void* pool, uint32_t epoch, cudaStream_t st,
                              unsigned long long* launches, bool beside_pipe = false);
// Window parser with its uint16 tables in global memory (parse_gtab_kernel), input read in place: a persistent
// grid of min(n_jobs, n_tables) one-warp CTAs, CTA b using table b of `tables` (gtab_table_bytes() each) and
// taking jobs from *ticket (a device word, zeroed by the launcher).  Concurrent launches need disjoint tables
// and tickets.  gtab_tables() = tables that fill one GPU.  Writes seqs and nseq (sizes: launch_size).
uint32_t gtab_tables();
size_t gtab_table_bytes();
cudaError_t launch_parse_gtab(const JobParams& p, void* tables, uint32_t n_tables, unsigned int* ticket, cudaStream_t st,
                              unsigned long long* launches);

cudaError_t launch_size(const JobParams& p, cudaStream_t st, unsigned long long* launches);

// Small buffers of a batch (exact mode, job_len <= 512): one thread per buffer parses AND emits it
// (parse_tiny.cuh); needs p.in / p.out / p.sizes only (no sequence lists, no size or emit pass).
bool tiny_takes(uint32_t job_len);
cudaError_t launch_tiny(const JobParams& p, cudaStream_t st, unsigned long long* launches);
cudaError_t launch_emit(const JobParams& p, cudaStream_t st, unsigned long long* launches);

// Exclusive scan of sizes[0..n_frag) (+hdr_len, + *carry_in when given) into offsets (uint64),
// running stream length into *d_total (and into *h_mirror, a mapped pinned host word, when
// given); writes varint32(n_input) at dst when hdr_len > 0.
cudaError_t launch_scan(const uint32_t* sizes, uint32_t n_frag, unsigned long long* offsets,
                        uint32_t hdr_len, uint32_t n_input, uint8_t* dst,
                        unsigned long long* d_total, cudaStream_t st, unsigned long long* launches,
                        const unsigned long long* carry_in = nullptr,
                        unsigned long long* h_mirror = nullptr);

// Debug: validate every job's sequence list (lengths add up, offsets in range).
cudaError_t launch_check(const JobParams& p, unsigned long long* report, cudaStream_t st);

// Diagnostics counters of the pipelined parser (all zero unless built with -DSNAPPY_B200_STATS).
cudaError_t read_stats(unsigned long long* out16, bool reset);

// Decompression (decode.cuh): `count` streams, stream j = in + j*in_stride with in_sizes[j] (or in_size)
// bytes, decoded to out + j*out_stride (capacity out_cap each); out_lens[j] = decoded length or
// 0xffffffff for a malformed stream; out_len64 (single stream) likewise with ~0.
cudaError_t launch_decode(const uint8_t* in, uint64_t in_stride, const uint32_t* in_sizes, uint32_t in_size,
                          uint32_t count, uint8_t* out, uint64_t out_stride, uint64_t out_cap,
                          uint32_t* out_lens, unsigned long long* out_len64, cudaStream_t st,
                          unsigned long long* launches);

// One large raw stream decoded block-parallel when it is block structured, serially otherwise
// (decode_par.cuh).  `scratch` must hold raw_decode_scratch_bytes(n, out_cap) bytes.  *out_len64 =
// decoded length or ~0 (malformed).  Asynchronous; the fast/slow decision is made on the device.
size_t raw_decode_scratch_bytes(uint64_t n, uint64_t out_cap);
cudaError_t launch_decode_large(const uint8_t* in, uint64_t n, uint8_t* out, uint64_t out_cap, void* scratch,
                                unsigned long long* out_len64, cudaStream_t st, unsigned long long* launches);

// ---- Snappy framing format (frame.cuh; tests/snappy/framing_format.txt) -------------------------
// Masked CRC-32C of `n_chunks` chunks of at most 65,536 bytes: chunk c = base + (offs ? offs[c] :
// c*stride), lens[c] bytes (lens == nullptr: min(fixed_len, n_total - c*stride)).
cudaError_t launch_crc32c_chunks(const uint8_t* base, const unsigned long long* offs, const uint32_t* lens,
                                 uint64_t stride, uint64_t n_total, uint32_t fixed_len, uint32_t n_chunks,
                                 uint32_t* out_masked, cudaStream_t st, unsigned long long* launches);
// writer: sizes[j] += 8 (chunk header + CRC); then, after the scan, identifier + chunk headers
cudaError_t launch_frame_bump(uint32_t* sizes, uint32_t n, cudaStream_t st, unsigned long long* launches);
cudaError_t launch_frame_headers(uint8_t* dst, const unsigned long long* offsets, const uint32_t* sizes,
                                 const uint32_t* crcs, uint32_t n, cudaStream_t st, unsigned long long* launches);
// reader: chunk index (kFrameRecordBytes per record at index_base), decode + CRC check of all chunks
constexpr size_t kFrameRecordBytes = 64;
struct FrameWalkResult {
    unsigned long long n_chunks, out_bytes;   // data chunks, bytes they decode to
    unsigned long long n_all;                 // every chunk of the stream (exact also when the index was too small)
    uint32_t error;                           // 0 ok, 1 malformed, 2 index too small: retry with max_records = n_all
    uint32_t pad;                             // scratch of the walk kernels (a chunk's parse raises it)
};
cudaError_t launch_frame_walk(const uint8_t* src, uint64_t n, void* index_base, uint64_t max_records,
                              FrameWalkResult* d_result, cudaStream_t st, unsigned long long* launches);
cudaError_t launch_frame_decode(const uint8_t* src, uint8_t* dst, void* index_base, uint64_t max_records,
                                uint32_t n_chunks, uint32_t* d_bad, cudaStream_t st,
                                unsigned long long* launches, bool decoded_already = false);
// reader, fused: chain walk + chunk decode in one kernel (frame.cuh); d_ctl = 16 bytes of device scratch.
// Afterwards d_result is as after launch_frame_walk, and launch_frame_decode(..., decoded_already = true)
// runs the CRC pass and the verification.
cudaError_t launch_frame_fused(const uint8_t* src, uint64_t n, uint8_t* dst, uint64_t dst_cap, void* index_base,
                               uint64_t max_records, FrameWalkResult* d_result, void* d_ctl, cudaStream_t st,
                               unsigned long long* launches);

cudaError_t kernels_init();  // one-time function attributes (dynamic shared memory opt-in)


// ---- multi-GPU exchange over peer memory (peer.cuh) ------------------------------------------
constexpr uint32_t kPeerMaxWorld = 64;
constexpr uint64_t kPeerSizeBits = 40;                       // payload sizes < 1 TiB
constexpr long long kPeerSpinCycles = 20000000000ll;         // ~10 s at 1.9 GHz

struct PeerMailbox {
    unsigned long long sizes[2][kPeerMaxWorld];              // (step + 1) << 40 | size, slot = step & 1
    unsigned long long finished[kPeerMaxWorld];              // steps completed by each rank
    unsigned long long error;                                // sticky: a wait timed out / archive too small
};

struct PeerTable {
    PeerMailbox* box[kPeerMaxWorld];                         // box[r] = rank r's mailbox as mapped here
};

cudaError_t launch_peer_step(const PeerTable& t, uint32_t rank, uint32_t world, uint32_t root, unsigned long long step,
                             unsigned long long base, const uint8_t* payload, const unsigned long long* d_size, uint8_t* archive_slot,
                             unsigned long long slot_bytes, unsigned long long* d_offset,
                             cudaStream_t st, cudaEvent_t copied,
                             unsigned long long* launches);
cudaError_t launch_peer_wait(PeerMailbox* mine, uint32_t world, unsigned long long step, cudaStream_t st,
                             unsigned long long* launches);

}  // namespace sb
