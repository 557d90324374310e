// capi.cu -- the C-ABI of libsnappy_b200 (include/snappy_b200.h): contexts, scratch, the
// stream / fragment-range / batch entry points and the RoCC driver mirror.
#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>
#include <stdlib.h>

#include <mutex>
#include <new>
#include <vector>

#include "../../include/snappy_b200.h"
#include "kernels.h"
#include "snappy_device.cuh"

namespace {

thread_local char g_err[256] = "";

snappy_b200_status fail(snappy_b200_status st, const char* what, cudaError_t e = cudaSuccess) {
    if (e != cudaSuccess)
        snprintf(g_err, sizeof g_err, "%s: %s", what, cudaGetErrorString(e));
    else
        snprintf(g_err, sizeof g_err, "%s", what);
    return st;
}

#define CU(call)                                                                  \
    do {                                                                          \
        cudaError_t e_ = (call);                                                  \
        if (e_ != cudaSuccess) return fail(SNAPPY_B200_CUDA_ERROR, #call, e_);    \
    } while (0)

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = n + n / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

}  // namespace

// Parse streams of the chunked host entry.  A stream starts its next chunk only when ALL fragments of
// its current one are done, so with S streams about S-1 chunks' worth of fragments are in flight: it
// takes 8 x 16 MiB (2048 fragments) to keep the 888 resident slots of the exact parser full.
constexpr int kParseStreams = 32;
constexpr int kPipeParseStreams = 8;
inline int pipe_streams() {
    static const int n = [] { const char* e = getenv("SNAPPY_B200_PARSE_STREAMS"); int v = e ? atoi(e) : kPipeParseStreams;
                              return v < 1 ? 1 : v > kParseStreams / 2 ? kParseStreams / 2 : v; }();
    return n;
}

constexpr size_t kMaxChunks = 1024;

struct snappy_b200_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    DevBuf seqs, nseq, sizes, offsets, d_in, d_out;
    DevBuf crcs, frame_ix, frame_res;          // framing: per-chunk CRCs, chunk index, walk result + flag
    DevBuf raw_ix;                             // segment / block index of one large raw stream (decode_par.cuh)
    DevBuf gtab_pool;                          // uint16 tables of the global-table window parser (parse_gtab_kernel)
    DevBuf lane_pool;                          // fat hash tables of the lane parser (parse_lane.cuh), never cleared per call
    uint32_t lane_epoch = 0;                   // epoch of the pool's most recent launch (0 = pool all-empty)
    unsigned long long* d_total = nullptr;     // device
    unsigned long long* h_total = nullptr;     // pinned
    unsigned int* h_err = nullptr;             // pinned + mapped: a parser watchdog stores 1 (surfaces as CUDA_ERROR)
    unsigned int* h_err_dev = nullptr;         // its device alias
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaEvent_t evp0 = nullptr, evp1 = nullptr, evp2 = nullptr;   // parse start / parse end / compact end
    bool staged = false;
    unsigned long long launches = 0;
    float last_ms = 0.f;
    // chunk pipeline of the host entry: H2D stream, parse streams, D2H stream; per-chunk events;
    // running stream length mirrored into mapped pinned host words by the scan kernel
    cudaStream_t s_h2d = nullptr, s_d2h = nullptr, s_parse[kParseStreams] = {};
    cudaStream_t s_emit = nullptr;             // high priority: size/scan/emit of the exact mode's chunks
    cudaEvent_t ev_done = nullptr;
    cudaEvent_t ev_start = nullptr;
    std::vector<cudaEvent_t> ev_h, ev_p, ev_e;
    unsigned long long* h_run = nullptr;       // pinned + mapped, kMaxChunks words
    unsigned long long* h_run_dev = nullptr;   // its device alias
};

namespace {

// SNAPPY_B200_DEBUG=1: after the parse, validate the sequence lists on the device and fail the
// call (before anything is emitted) if a job's lengths do not add up.
snappy_b200_status debug_check(snappy_b200_ctx* c, const sb::JobParams& p);

std::once_flag g_init_once;
cudaError_t g_init_err = cudaSuccess;

snappy_b200_status bind(snappy_b200_ctx* c) {
    CU(cudaSetDevice(c->device));
    return SNAPPY_B200_OK;
}

thread_local snappy_b200_ctx* g_default_ctx = nullptr;

// After a synchronisation point: did a parser watchdog give up on a fragment?  (Its output is then a well
// formed stream, but not the reference's bytes, so the call fails instead of returning it.)
snappy_b200_status watchdog_status(snappy_b200_ctx* c) {
    if (c->h_err && *(volatile unsigned int*)c->h_err) {
        *(volatile unsigned int*)c->h_err = 0;
        return fail(SNAPPY_B200_CUDA_ERROR, "parser watchdog: a fragment's pipeline stalled (result discarded)");
    }
    return SNAPPY_B200_OK;
}

snappy_b200_status default_ctx(snappy_b200_ctx** out) {
    if (!g_default_ctx) {
        snappy_b200_status st = snappy_b200_create(&g_default_ctx, 0);
        if (st != SNAPPY_B200_OK) return st;
    }
    *out = g_default_ctx;
    return SNAPPY_B200_OK;
}

// ---- exact parse: which kernel --------------------------------------------------------------------
// mode 0 (SNAPPY_B200_EXACT) picks by shape: the lane parser (one thread per fragment, all fragments in
// flight, parse_lane.cuh) when the call has enough fragments to hide DRAM latency behind their number,
// the pipelined parser (one CTA per fragment, parse_pipe.cuh) otherwise.  Modes 5 / 6 force one or the
// other (tests, A/B runs); both produce the reference's bytes.  Measured crossover on text (ms per
// parse, pipelined vs lane): 512 MiB 15.5 vs 18.5, 1 GiB 31.3 vs 22.4, 2 GiB 62 vs 40.7.  Running both
// side by side on disjoint fragment ranges was measured and lost (22.1 ms per GiB: the pipelined CTAs
// need the maximum shared-memory carve-out, which leaves the lane parser 28 KB of L1).
constexpr int kModeLane = 5, kModePipe = 6, kModeGtab = 7;
// Cap of the lane parser's table pool: calls that need more run in waves (lane_parse_all).  20 GiB by default;
// SNAPPY_B200_LANE_POOL_MIB lowers it (tests force waves on small streams with it).
inline size_t lane_pool_max() {
    static const size_t v = [] { const char* e = getenv("SNAPPY_B200_LANE_POOL_MIB"); return e ? (size_t)atol(e) << 20 : (size_t)20 << 30; }();
    return v;
}
inline uint32_t lane_min_frags() {
    static const uint32_t v = [] { const char* e = getenv("SNAPPY_B200_LANE_MIN_FRAGS"); return e ? (uint32_t)atol(e) : 13312u; }();
    return v;
}
inline bool lane_chosen(int mode, uint64_t n_jobs, uint32_t job_len) {
    if (mode == kModeLane) return true;
    // calls whose tables exceed the pool cap (streams over 2.5 GiB) go to the global-table parser: the lane parser
    // in waves was measured slower (3 GiB of text: two waves 78.2 ms, global-table parser 72.1 ms)
    return mode == SNAPPY_B200_EXACT && job_len == sb::kBlock && n_jobs >= lane_min_frags() &&
           sb::lane_pool_bytes(n_jobs, job_len) <= lane_pool_max();
}
inline int kernel_mode(int mode) { return mode == kModePipe ? (int)SNAPPY_B200_EXACT : mode; }

// The third exact engine: the window parser with its uint16 tables in global memory (parse_gtab_kernel; one
// warp per fragment, 28 per SM, input read in place).  6 ms per fragment and ~42 GB/s: between the pipelined
// parser (1.6 ms, 34 GB/s, six fragments per SM) and the lane parser (17 ms floor, 46 GB/s).  Measured on
// text, ms per parse (pipelined / global-table / lane): 64 MiB 2.9 / 3.2 / -, 128 MiB 4.6 / 4.2 / -, 256 MiB
// 7.9 / 6.1 / -, 512 MiB 15.5 / 12.1 / 18.5, 768 MiB - / 18.1 / 19.8, 896 MiB - / 22.3 / 21.0, 1 GiB 31.4 /
// 24.2 / 22.2.  Mode 7 forces it.
inline uint32_t gtab_min_frags() {
    static const uint32_t v = [] { const char* e = getenv("SNAPPY_B200_GTAB_MIN_FRAGS"); return e ? (uint32_t)atol(e) : 1792u; }();
    return v;
}
inline bool gtab_chosen(int mode, uint64_t n_jobs, uint32_t job_len) {
    if (mode == kModeGtab) return true;
    // batch buffers: over 4 KiB (8 KiB text buffers 72 vs 49 GB/s for the shared-memory page parser, which fits
    // only nine of them per SM; at 4 KiB the two tie and the page parser stays)
    return mode == SNAPPY_B200_EXACT && job_len > 4096 && n_jobs >= gtab_min_frags() && !lane_chosen(mode, n_jobs, job_len);
}
// Tables + tickets of the global-table parser: `regions` disjoint sets of `tables` tables (one per concurrent
// launch), tickets (one 256-byte slot per region) behind them.
snappy_b200_status gtab_reserve(snappy_b200_ctx* c, uint32_t regions, uint32_t tables) {
    CU(c->gtab_pool.reserve((size_t)regions * tables * sb::gtab_table_bytes() + (size_t)regions * 256));
    return SNAPPY_B200_OK;
}
inline void* gtab_region(snappy_b200_ctx* c, uint32_t r, uint32_t tables) {
    return (uint8_t*)c->gtab_pool.p + (size_t)r * tables * sb::gtab_table_bytes();
}
inline unsigned int* gtab_ticket(snappy_b200_ctx* c, uint32_t regions, uint32_t tables, uint32_t r) {
    return (unsigned int*)((uint8_t*)c->gtab_pool.p + (size_t)regions * tables * sb::gtab_table_bytes() + (size_t)r * 256);
}
// Reserve the pool for `n_jobs` tables and open a new epoch (stream-ordered on c->stream).
snappy_b200_status lane_begin(snappy_b200_ctx* c, uint64_t n_jobs, uint32_t job_len, uint32_t* epoch) {
    const size_t need = sb::lane_pool_bytes(n_jobs, job_len);
    const bool fresh = need > c->lane_pool.cap;
    CU(c->lane_pool.reserve(need));
    if (fresh || c->lane_epoch >= 65535u) {                  // new memory, or the 16-bit epoch would repeat
        CU(cudaMemsetAsync(c->lane_pool.p, 0, c->lane_pool.cap, c->stream));
        c->lane_epoch = 0;
    }
    *epoch = ++c->lane_epoch;
    return SNAPPY_B200_OK;
}

// Parse + sizes of all jobs of `p` on c->stream, in waves if the pool cap is smaller than the batch.
snappy_b200_status lane_parse_all(snappy_b200_ctx* c, const sb::JobParams& p) {
    if (p.n_jobs == 0) return SNAPPY_B200_OK;
    const size_t per_job = sb::lane_pool_bytes(1, p.job_len);
    uint64_t per_wave = lane_pool_max() / per_job;
    if (per_wave == 0) per_wave = 1;
    {   // equal waves: a short last wave would pay the kernel's latency floor for a handful of fragments
        const uint64_t waves = (p.n_jobs + per_wave - 1) / per_wave;
        per_wave = (p.n_jobs + waves - 1) / waves;
    }
    for (uint64_t j0 = 0; j0 < p.n_jobs; j0 += per_wave) {
        sb::JobParams q = p;
        q.in += j0 * p.in_stride;
        q.n_total -= j0 * p.in_stride;
        q.seqs += j0 * sb::seq_stride(p.job_len);
        q.nseq += j0;
        q.sizes += j0;
        q.n_jobs = (uint32_t)(p.n_jobs - j0 < per_wave ? p.n_jobs - j0 : per_wave);
        uint32_t epoch = 0;
        if (lane_begin(c, q.n_jobs, p.job_len, &epoch) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
        CU(sb::launch_parse_lane(q, c->lane_pool.p, epoch, c->stream, &c->launches));
    }
    return SNAPPY_B200_OK;
}

// Stream / fragment-range compression with input and output on the device.
snappy_b200_status compress_dev(snappy_b200_ctx* c, const void* d_src, size_t n, void* d_dst,
                                size_t dst_cap, unsigned long long* d_out_n, int mode,
                                int framing /* 0 fragments only, 1 raw stream, 2 framed stream */) {
    const bool with_header = framing == 1, framed = framing == 2;
    if (n >= (1ull << 32)) return fail(SNAPPY_B200_INVALID_INPUT, "input of 2^32 bytes or more");
    if (mode != SNAPPY_B200_EXACT && mode != SNAPPY_B200_FAST && mode != 2 /* exact, serial parser */ &&
        mode != 3 /* exact, single-warp window parser */ && mode != 4 /* exact, pipelined, input in smem */ &&
        mode != kModeLane && mode != kModePipe && mode != kModeGtab)
        return fail(SNAPPY_B200_INVALID_INPUT, "unknown mode");
    if (dst_cap < (framed ? snappy_b200_frame_max_length(n) : snappy_b200_max_compressed_length(n)))
        return fail(SNAPPY_B200_BUFFER_TOO_SMALL, "dst_cap < max length of the output");
    if (bind(c) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    const uint32_t n_frag = (uint32_t)((n + sb::kBlock - 1) / sb::kBlock);
    const uint32_t hdr = framed ? 10u : with_header ? sb::varint32_len((uint32_t)n) : 0;
    if (framed) CU(c->crcs.reserve((size_t)(n_frag + 1) * sizeof(uint32_t)));
    CU(c->seqs.reserve((size_t)n_frag * sb::seq_stride(sb::kBlock) * sizeof(sb::seq_t) + 256));
    CU(c->nseq.reserve((size_t)(n_frag + 1) * sizeof(uint32_t)));
    CU(c->sizes.reserve((size_t)(n_frag + 1) * sizeof(uint32_t)));
    CU(c->offsets.reserve((size_t)(n_frag + 1) * sizeof(unsigned long long)));
    sb::JobParams p;
    p.in = (const uint8_t*)d_src;
    p.n_total = n;
    p.in_stride = sb::kBlock;
    p.job_len = sb::kBlock;
    p.n_jobs = n_frag;
    p.seqs = (sb::seq_t*)c->seqs.p;
    p.nseq = (uint32_t*)c->nseq.p;
    p.sizes = (uint32_t*)c->sizes.p;
    // framed: every 64 KiB chunk is its own Snappy stream (varint preamble) behind an 8-byte chunk
    // header + CRC, so the stream is emitted 8 bytes into the chunk's slot (frame.cuh)
    p.with_header = framed ? 1 : 0;
    p.sparse = (mode == SNAPPY_B200_FAST);
    p.out = (uint8_t*)d_dst + (framed ? 8 : 0);
    p.out_stride = 0;
    p.offsets = (const unsigned long long*)c->offsets.p;
    p.err = c->h_err_dev;
    const bool lane = lane_chosen(mode, n_frag, sb::kBlock);
    const bool gtab = gtab_chosen(mode, n_frag, sb::kBlock);
    if (gtab && gtab_reserve(c, 1, sb::gtab_tables()) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    CU(cudaEventRecord(c->evp0, c->stream));
    if (lane) { if (lane_parse_all(c, p) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR; }
    else if (gtab) CU(sb::launch_parse_gtab(p, gtab_region(c, 0, sb::gtab_tables()), sb::gtab_tables(),
                                            gtab_ticket(c, 1, sb::gtab_tables(), 0), c->stream, &c->launches));
    else CU(sb::launch_parse(p, kernel_mode(mode), c->stream, &c->launches));
    CU(cudaEventRecord(c->evp1, c->stream));
    if (debug_check(c, p) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    if (!lane) CU(sb::launch_size(p, c->stream, &c->launches));   // the lane parser adds up sizes as it goes
    if (framed) {
        CU(sb::launch_frame_bump((uint32_t*)c->sizes.p, n_frag, c->stream, &c->launches));
        CU(sb::launch_crc32c_chunks((const uint8_t*)d_src, nullptr, nullptr, sb::kBlock, n, sb::kBlock, n_frag,
                                    (uint32_t*)c->crcs.p, c->stream, &c->launches));
    }
    CU(sb::launch_scan((const uint32_t*)c->sizes.p, n_frag, (unsigned long long*)c->offsets.p, hdr,
                       framed ? 0u : (uint32_t)n, (uint8_t*)d_dst, d_out_n, c->stream, &c->launches));
    CU(sb::launch_emit(p, c->stream, &c->launches));
    if (framed)
        CU(sb::launch_frame_headers((uint8_t*)d_dst, (const unsigned long long*)c->offsets.p,
                                    (const uint32_t*)c->sizes.p, (const uint32_t*)c->crcs.p, n_frag,
                                    c->stream, &c->launches));
    CU(cudaEventRecord(c->evp2, c->stream));
    c->staged = true;
    return SNAPPY_B200_OK;
}

constexpr size_t kPipeMinBytes = 24u << 20;     // below this one H2D + one D2H is as good as it gets

// Chunk size of the host pipeline: whole fragments; large inputs use 64 MiB (1024 fragments, a
// little more than one resident wave of the pipelined parser), smaller ones an eighth.
size_t pipe_chunk_bytes(size_t n, bool fast, bool lane) {
    static const size_t env = [] { const char* e = getenv("SNAPPY_B200_CHUNK_MIB"); return e ? (size_t)atol(e) << 20 : (size_t)0; }();
    if (lane && !env) {
        // a lane-parser launch takes the same time whatever its size (every fragment is its own chain),
        // so chunks only set how early the first one starts: 32 MiB = 0.6 ms of PCIe each, one stream each
        size_t c = 32u << 20;
        while ((n + c - 1) / c > kMaxChunks) c *= 2;
        return c;
    }
    // measured on 1 GiB text (8 parse streams): exact 16 MiB 36.0 ms, 32 MiB 36.4, 8 MiB 44; fast 32 MiB 21.2 ms, 16 MiB 22.6
    size_t c = env ? env : (n >= (256u << 20) ? (fast ? (32u << 20) : (16u << 20)) : n / 8);
    c = (c + sb::kBlock - 1) / sb::kBlock * sb::kBlock;
    if (c < (4u << 20)) c = 4u << 20;
    while ((n + c - 1) / c > kMaxChunks) c *= 2;
    return c;
}

snappy_b200_status ensure_pipe(snappy_b200_ctx* c, size_t n_chunks) {
    if (!c->s_h2d) {
        CU(cudaStreamCreateWithFlags(&c->s_h2d, cudaStreamNonBlocking));
        CU(cudaStreamCreateWithFlags(&c->s_d2h, cudaStreamNonBlocking));
        for (auto& s : c->s_parse) CU(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
        int least = 0, greatest = 0;
        CU(cudaDeviceGetStreamPriorityRange(&least, &greatest));
        CU(cudaStreamCreateWithPriority(&c->s_emit, cudaStreamNonBlocking, greatest));
        CU(cudaEventCreateWithFlags(&c->ev_done, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&c->ev_start, cudaEventDisableTiming));
        CU(cudaHostAlloc((void**)&c->h_run, kMaxChunks * sizeof(unsigned long long), cudaHostAllocMapped));
        CU(cudaHostGetDevicePointer((void**)&c->h_run_dev, c->h_run, 0));
    }
    while (c->ev_h.size() < n_chunks) {
        cudaEvent_t a, b, d;
        CU(cudaEventCreateWithFlags(&a, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&b, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&d, cudaEventDisableTiming));
        c->ev_h.push_back(a); c->ev_p.push_back(b); c->ev_e.push_back(d);
    }
    return SNAPPY_B200_OK;
}

// Host entry for large inputs: the stream is cut into chunks of whole fragments and the chunks
// flow through H2D -> parse -> size/scan/emit -> D2H on separate CUDA streams, so the PCIe copies
// of chunk k+1 / k-1 overlap the kernels of chunk k, and the parse kernels of neighbouring
// chunks overlap each other's tails.  Fragment outputs land at their final offsets in one device
// buffer (the scan carries the running length across chunks), so the bytes are those of the
// one-shot path.
snappy_b200_status compress_host_pipelined(snappy_b200_ctx* c, const uint8_t* src, size_t n,
                                           uint8_t* dst, size_t dst_cap, size_t* out_n, int mode) {
    const size_t cap = snappy_b200_max_compressed_length(n);
    const uint32_t n_frag = (uint32_t)((n + sb::kBlock - 1) / sb::kBlock);
    // Which exact parser takes which chunk.  A lane-parser launch needs ~17 ms whatever its size (every
    // fragment is one serial chain of DRAM round trips), the pipelined parser ~1.6 ms per 888 fragments, so
    // behind a PCIe copy of ~20 ms per GiB the chunks go to the pipelined parser.  SNAPPY_B200_E2E_LANE_PCT
    // > 0 sends that share of the input FIRST (its tail) and gives it to the lane parser, whose latency
    // then hides under the rest of the transfer; elements are still emitted in stream order (head chunks
    // as they finish, the tail last), so bytes and drain do not change.  Measured on 1 GiB text: 0 % 37.4
    // ms, 20 % 38.1, 33 % 38.6, 45 % 39.4 -- the two parsers slow each other down (the lane parser loses
    // its L1 to the pipelined parser's shared-memory carve-out), so the default is 0.  mode 5 forces the
    // lane parser everywhere (tests).
    const bool lane_all = mode == kModeLane;
    // Exact mode, 112 MiB and more: the chunks go to the global-table window parser (42 GB/s and 6 ms per
    // fragment, against 34 GB/s for the pipelined parser): 32 MiB chunks, one table region per parse stream.
    const bool use_gtab = mode == kModeGtab || (mode == SNAPPY_B200_EXACT && n_frag >= gtab_min_frags());
    const size_t chunk = pipe_chunk_bytes(n, mode == SNAPPY_B200_FAST, lane_all || use_gtab);
    const size_t n_chunks = (n + chunk - 1) / chunk;
    static const int tail_pct = [] { const char* e = getenv("SNAPPY_B200_E2E_LANE_PCT"); return e ? atoi(e) : 0; }();
    size_t k_split = n_chunks;                               // chunks [k_split, n_chunks) go to the lane parser
    if (lane_all) k_split = 0;
    else if (mode == SNAPPY_B200_EXACT && n_frag >= 4096 && tail_pct > 0)
        k_split = n_chunks - (n_chunks * (size_t)(tail_pct > 100 ? 100 : tail_pct)) / 100;
    const bool any_lane = k_split < n_chunks;
    const uint32_t hdr = sb::varint32_len((uint32_t)n);
    if (ensure_pipe(c, n_chunks) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    CU(c->d_in.reserve(n + 64));
    CU(c->d_out.reserve(cap + 64));
    CU(c->seqs.reserve((size_t)n_frag * sb::seq_stride(sb::kBlock) * sizeof(sb::seq_t) + 256));
    CU(c->nseq.reserve((size_t)(n_frag + 1) * sizeof(uint32_t)));
    CU(c->sizes.reserve((size_t)(n_frag + 1) * sizeof(uint32_t)));
    CU(c->offsets.reserve((size_t)(n_frag + 1) * sizeof(unsigned long long)));
    uint8_t* d_in = (uint8_t*)c->d_in.p;
    uint8_t* d_out = (uint8_t*)c->d_out.p;
    const uint32_t f_split = (uint32_t)(k_split * chunk / sb::kBlock);   // first fragment of the lane part
    uint32_t lane_epoch = 0;
    if (any_lane && lane_begin(c, n_frag - f_split, sb::kBlock, &lane_epoch) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    // parse streams: 8 for the pipelined parser; 12 for the global-table parser, whose chunks have long tails
    // (a fragment takes ~6 ms): measured 8 -> 31.4 ms per GiB end to end, 12 -> 30.6, 16 -> 30.7
    static const bool streams_env = getenv("SNAPPY_B200_PARSE_STREAMS") != nullptr;
    const int n_pipe_s = use_gtab && !streams_env ? 12 : pipe_streams();
    const uint32_t g_regions = (uint32_t)n_pipe_s, g_tables = (uint32_t)(chunk / sb::kBlock);
    if (use_gtab && gtab_reserve(c, g_regions, g_tables) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    // nothing of this call may overtake earlier work on the context's stream (shared scratch)
    CU(cudaEventRecord(c->ev_start, c->stream));
    CU(cudaStreamWaitEvent(c->s_h2d, c->ev_start, 0));
    CU(cudaEventRecord(c->ev0, c->stream));
    // size/scan/emit stream.  Exact mode: a high-priority stream, so these short kernels take free SM
    // slots ahead of the queued parse CTAs and a chunk's bytes reach the D2H copy early (measured
    // 38.7 -> 36.0 ms per GiB).  Fast mode is PCIe-bound and its emit pass is heavier: priority only
    // delays the parse there (21.2 -> 22.7 ms), so it stays on the context's stream.
    static const bool prio_on = [] { const char* e = getenv("SNAPPY_B200_STREAM_PRIO"); return !e || atoi(e) != 0; }();
    cudaStream_t es = (mode == SNAPPY_B200_FAST || !prio_on) ? c->stream : c->s_emit;
    if (es != c->stream) CU(cudaStreamWaitEvent(es, c->ev_start, 0));
    auto job_params = [&](size_t k) {
        const size_t lo = k * chunk, len = (n - lo < chunk) ? n - lo : chunk;
        const uint32_t f0 = (uint32_t)(lo / sb::kBlock);
        sb::JobParams p;
        p.in = d_in + lo;
        p.n_total = len;
        p.in_stride = sb::kBlock;
        p.job_len = sb::kBlock;
        p.n_jobs = (uint32_t)((len + sb::kBlock - 1) / sb::kBlock);
        p.seqs = (sb::seq_t*)c->seqs.p + (size_t)f0 * sb::seq_stride(sb::kBlock);
        p.nseq = (uint32_t*)c->nseq.p + f0;
        p.sizes = (uint32_t*)c->sizes.p + f0;
        p.with_header = 0;
        p.sparse = (mode == SNAPPY_B200_FAST);
        p.out = d_out;
        p.out_stride = 0;
        p.offsets = (const unsigned long long*)c->offsets.p + f0;
        p.err = c->h_err_dev;
        return p;
    };
    // copy + parse of chunk k (the lane parser when k >= k_split), parse streams round robin per parser
    auto copy_and_parse = [&](size_t k) -> snappy_b200_status {
        const size_t lo = k * chunk, len = (n - lo < chunk) ? n - lo : chunk;
        const uint32_t f0 = (uint32_t)(lo / sb::kBlock);
        const bool lane = k >= k_split;
        CU(cudaMemcpyAsync(d_in + lo, src + lo, len, cudaMemcpyHostToDevice, c->s_h2d));
        CU(cudaEventRecord(c->ev_h[k], c->s_h2d));
        const sb::JobParams p = job_params(k);
        cudaStream_t ps = lane ? c->s_parse[lane_all ? k % kParseStreams : n_pipe_s + (k - k_split) % (kParseStreams - n_pipe_s)]
                               : c->s_parse[k % n_pipe_s];
        CU(cudaStreamWaitEvent(ps, c->ev_h[k], 0));
        if (lane)
            CU(sb::launch_parse_lane(p, (uint8_t*)c->lane_pool.p + sb::lane_pool_bytes(f0 - f_split, sb::kBlock), lane_epoch,
                                     ps, &c->launches, /*beside_pipe=*/!lane_all));
        else if (use_gtab)
            CU(sb::launch_parse_gtab(p, gtab_region(c, (uint32_t)(k % n_pipe_s), g_tables), g_tables,
                                     gtab_ticket(c, g_regions, g_tables, (uint32_t)(k % n_pipe_s)), ps, &c->launches));
        else CU(sb::launch_parse(p, kernel_mode(mode), ps, &c->launches));
        CU(cudaEventRecord(c->ev_p[k], ps));
        return SNAPPY_B200_OK;
    };
    // sizes -> offsets -> elements of chunk k, in stream order on the emit stream
    auto scan_and_emit = [&](size_t k) -> snappy_b200_status {
        const sb::JobParams p = job_params(k);
        const uint32_t f0 = (uint32_t)(k * chunk / sb::kBlock);
        CU(cudaStreamWaitEvent(es, c->ev_p[k], 0));
        if (k < k_split) CU(sb::launch_size(p, es, &c->launches));      // the lane parser adds up sizes itself
        CU(sb::launch_scan(p.sizes, p.n_jobs, (unsigned long long*)c->offsets.p + f0, k == 0 ? hdr : 0,
                           (uint32_t)n, d_out, c->d_total, es, &c->launches,
                           k == 0 ? nullptr : c->d_total, c->h_run_dev + k));
        CU(sb::launch_emit(p, es, &c->launches));
        CU(cudaEventRecord(c->ev_e[k], es));
        return SNAPPY_B200_OK;
    };
    for (size_t k = k_split; k < n_chunks; ++k)              // the tail crosses PCIe first
        if (copy_and_parse(k) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    for (size_t k = 0; k < k_split; ++k) {
        if (copy_and_parse(k) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
        if (scan_and_emit(k) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    }
    for (size_t k = k_split; k < n_chunks; ++k)
        if (scan_and_emit(k) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    if (es != c->stream) {                                   // the context's stream stays the call's timeline
        CU(cudaEventRecord(c->ev_done, es));
        CU(cudaStreamWaitEvent(c->stream, c->ev_done, 0));
    }
    CU(cudaEventRecord(c->ev1, c->stream));
    // drain: as soon as chunk k's elements are in place its byte range goes back to the host
    size_t done = 0;
    snappy_b200_status st = SNAPPY_B200_OK;
    for (size_t k = 0; k < n_chunks; ++k) {
        CU(cudaEventSynchronize(c->ev_e[k]));
        const size_t run = (size_t)*(volatile unsigned long long*)(c->h_run + k);
        if (run > dst_cap) { st = fail(SNAPPY_B200_BUFFER_TOO_SMALL, "stream larger than dst_cap"); break; }
        if (run > done) CU(cudaMemcpyAsync(dst + done, d_out + done, run - done, cudaMemcpyDeviceToHost, c->s_d2h));
        done = run;
    }
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaStreamSynchronize(c->s_d2h));
    if (st != SNAPPY_B200_OK) return st;
    if (watchdog_status(c) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    cudaEventElapsedTime(&c->last_ms, c->ev0, c->ev1);
    *out_n = done;
    return SNAPPY_B200_OK;
}

snappy_b200_status debug_check(snappy_b200_ctx* c, const sb::JobParams& p) {
    static const bool on = [] { const char* e = getenv("SNAPPY_B200_DEBUG"); return e && *e == '1'; }();
    if (!on) return SNAPPY_B200_OK;
    unsigned long long* d_rep = nullptr;
    unsigned long long h_rep[8] = {0};
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaMalloc(&d_rep, sizeof h_rep));
    CU(cudaMemsetAsync(d_rep, 0, sizeof h_rep, c->stream));
    CU(sb::launch_check(p, d_rep, c->stream));
    CU(cudaMemcpyAsync(h_rep, d_rep, sizeof h_rep, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    cudaFree(d_rep);
    if (h_rep[0]) {
        snprintf(g_err, sizeof g_err, "debug: %llu bad sequence lists; first job %llu nseq %llu sum %llu len %llu",
                 h_rep[0], h_rep[1], h_rep[2], h_rep[3], h_rep[4]);
        fprintf(stderr, "libsnappy_b200 %s\n", g_err);
        return SNAPPY_B200_CUDA_ERROR;
    }
    return SNAPPY_B200_OK;
}

}  // namespace

extern "C" {

size_t snappy_b200_max_compressed_length(size_t n) { return 32 + n + n / 6; }

const char* snappy_b200_last_error(void) { return g_err; }

size_t snappy_b200_varint32(void* dst, uint32_t n) { return sb::put_varint32((uint8_t*)dst, n); }

snappy_b200_status snappy_b200_create(snappy_b200_ctx** out, int device) {
    if (!out) return fail(SNAPPY_B200_INVALID_INPUT, "null ctx pointer");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(SNAPPY_B200_NO_DEVICE, "no CUDA device (libsnappy_b200 has no CPU path)", e);
    if (device < 0 || device >= count) return fail(SNAPPY_B200_INVALID_INPUT, "bad device ordinal");
    CU(cudaSetDevice(device));
    std::call_once(g_init_once, [] { g_init_err = sb::kernels_init(); });
    if (g_init_err != cudaSuccess) return fail(SNAPPY_B200_CUDA_ERROR, "kernels_init", g_init_err);
    snappy_b200_ctx* c = new (std::nothrow) snappy_b200_ctx;
    if (!c) return fail(SNAPPY_B200_CUDA_ERROR, "out of host memory");
    c->device = device;
    CU(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    CU(cudaMalloc(&c->d_total, sizeof(unsigned long long)));
    CU(cudaMallocHost(&c->h_total, sizeof(unsigned long long)));
    CU(cudaHostAlloc((void**)&c->h_err, sizeof(unsigned int), cudaHostAllocMapped));
    *c->h_err = 0;
    CU(cudaHostGetDevicePointer((void**)&c->h_err_dev, c->h_err, 0));
    CU(cudaEventCreate(&c->ev0));
    CU(cudaEventCreate(&c->ev1));
    CU(cudaEventCreate(&c->evp0));
    CU(cudaEventCreate(&c->evp1));
    CU(cudaEventCreate(&c->evp2));
    *out = c;
    return SNAPPY_B200_OK;
}

void snappy_b200_destroy(snappy_b200_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    c->seqs.release(); c->nseq.release(); c->sizes.release(); c->offsets.release();
    c->d_in.release(); c->d_out.release(); c->lane_pool.release(); c->gtab_pool.release();
    c->crcs.release(); c->frame_ix.release(); c->frame_res.release(); c->raw_ix.release();
    if (c->d_total) cudaFree(c->d_total);
    if (c->h_total) cudaFreeHost(c->h_total);
    if (c->h_err) cudaFreeHost(c->h_err);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->evp0) cudaEventDestroy(c->evp0);
    if (c->evp1) cudaEventDestroy(c->evp1);
    if (c->evp2) cudaEventDestroy(c->evp2);
    if (c->stream) cudaStreamDestroy(c->stream);
    if (c->s_h2d) cudaStreamDestroy(c->s_h2d);
    if (c->s_d2h) cudaStreamDestroy(c->s_d2h);
    for (auto s_ : c->s_parse) if (s_) cudaStreamDestroy(s_);
    if (c->s_emit) cudaStreamDestroy(c->s_emit);
    if (c->ev_done) cudaEventDestroy(c->ev_done);
    if (c->ev_start) cudaEventDestroy(c->ev_start);
    for (auto e_ : c->ev_h) cudaEventDestroy(e_);
    for (auto e_ : c->ev_p) cudaEventDestroy(e_);
    for (auto e_ : c->ev_e) cudaEventDestroy(e_);
    if (c->h_run) cudaFreeHost(c->h_run);
    if (g_default_ctx == c) g_default_ctx = nullptr;
    delete c;
}

void* snappy_b200_stream(snappy_b200_ctx* c) { return c ? (void*)c->stream : nullptr; }

unsigned long long snappy_b200_launch_count(snappy_b200_ctx* c) { return c ? c->launches : 0; }

float snappy_b200_last_kernel_ms(snappy_b200_ctx* c) { return c ? c->last_ms : 0.f; }

int snappy_b200_debug_stats(unsigned long long* out16, int reset) {
    return sb::read_stats(out16, reset != 0) == cudaSuccess ? 0 : 1;
}

float snappy_b200_stage_ms(snappy_b200_ctx* c, int stage) {
    if (!c || !c->staged) return 0.f;
    float ms = 0.f;
    cudaError_t e = stage == 0 ? cudaEventElapsedTime(&ms, c->evp0, c->evp1)
                               : cudaEventElapsedTime(&ms, c->evp1, c->evp2);
    return e == cudaSuccess ? ms : 0.f;
}

snappy_b200_status snappy_b200_sync(snappy_b200_ctx* c) {
    if (!c) return fail(SNAPPY_B200_INVALID_INPUT, "null ctx");
    if (bind(c) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    CU(cudaStreamSynchronize(c->stream));
    return watchdog_status(c);
}

void* snappy_b200_host_alloc(size_t n) {
    void* p = nullptr;
    if (cudaMallocHost(&p, n ? n : 1) != cudaSuccess) return nullptr;
    return p;
}

void snappy_b200_host_free(void* p) { if (p) cudaFreeHost(p); }

snappy_b200_status snappy_b200_compress_device(snappy_b200_ctx* c, const void* d_src, size_t n,
                                               void* d_dst, size_t dst_cap,
                                               unsigned long long* d_out_n, int mode) {
    if (!c || !d_dst || !d_out_n || (!d_src && n)) return fail(SNAPPY_B200_INVALID_INPUT, "null argument");
    return compress_dev(c, d_src, n, d_dst, dst_cap, d_out_n, mode, 1);
}

snappy_b200_status snappy_b200_compress_fragments_device(snappy_b200_ctx* c, const void* d_src,
                                                         size_t n, void* d_dst, size_t dst_cap,
                                                         unsigned long long* d_out_n, int mode) {
    if (!c || !d_dst || !d_out_n || (!d_src && n)) return fail(SNAPPY_B200_INVALID_INPUT, "null argument");
    return compress_dev(c, d_src, n, d_dst, dst_cap, d_out_n, mode, 0);
}

snappy_b200_status snappy_b200_compress_batch_device(snappy_b200_ctx* c, const void* d_src,
                                                     size_t len, size_t in_stride, size_t count,
                                                     void* d_dst, size_t out_stride,
                                                     uint32_t* d_sizes, int mode) {
    if (!c || !d_dst || !d_sizes || (!d_src && len && count))
        return fail(SNAPPY_B200_INVALID_INPUT, "null argument");
    if (len > sb::kBlock) return fail(SNAPPY_B200_INVALID_INPUT, "batch buffers are limited to 65536 bytes");
    if (count > 0xffffffffull) return fail(SNAPPY_B200_INVALID_INPUT, "too many buffers");
    if (in_stride < len) return fail(SNAPPY_B200_INVALID_INPUT, "in_stride < len");
    if (out_stride < snappy_b200_max_compressed_length(len))
        return fail(SNAPPY_B200_BUFFER_TOO_SMALL, "out_stride < max_compressed_length(len)");
    if (bind(c) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    if (count == 0) return SNAPPY_B200_OK;
    // exact mode, buffers up to 256 bytes: one thread per buffer, parse and emission fused (parse_tiny.cuh)
    static const bool tiny_on = [] { const char* e = getenv("SNAPPY_B200_TINY"); return !e || atoi(e) != 0; }();
    const bool tiny = mode == SNAPPY_B200_EXACT && tiny_on && sb::tiny_takes((uint32_t)len);
    if (!tiny) {
        CU(c->seqs.reserve(count * sb::seq_stride((uint32_t)len) * sizeof(sb::seq_t) + 256));
        CU(c->nseq.reserve((count + 1) * sizeof(uint32_t)));
    }
    sb::JobParams p;
    p.in = (const uint8_t*)d_src;
    p.n_total = (count - 1) * in_stride + len;
    p.in_stride = in_stride;
    p.job_len = (uint32_t)len;
    p.n_jobs = (uint32_t)count;
    p.seqs = (sb::seq_t*)c->seqs.p;
    p.nseq = (uint32_t*)c->nseq.p;
    p.sizes = d_sizes;
    p.with_header = 1;
    p.sparse = (mode == SNAPPY_B200_FAST);
    p.out = (uint8_t*)d_dst;
    p.out_stride = out_stride;
    p.offsets = nullptr;
    p.err = c->h_err_dev;
    const bool lane = lane_chosen(mode, count, (uint32_t)len);
    const bool gtab = gtab_chosen(mode, count, (uint32_t)len);
    if (gtab && gtab_reserve(c, 1, sb::gtab_tables()) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    // buffers of the page parser and of the global-table parser: the parsing warp writes the elements too
    // (kernels.cu warp_emit_job).  Text, GB/s with / without: 600 B 81.8 / 52.8, 1,000 B 85.3 / 63.0, 2,000 B 93.6 /
    // 78.6, 4 KiB 75.3 / 71.2.
    static const size_t fuse_max = [] { const char* e = getenv("SNAPPY_B200_FUSE_EMIT_MAX"); return e ? (size_t)atol(e) : (size_t)8192; }();
    static const size_t fuse_gtab_max = [] { const char* e = getenv("SNAPPY_B200_FUSE_EMIT_GTAB_MAX"); return e ? (size_t)atol(e) : (size_t)32768; }();
    const bool fused = (mode == SNAPPY_B200_EXACT || mode == kModeGtab) && !tiny && !lane &&
                       (gtab ? len <= fuse_gtab_max : len <= fuse_max && sb::page_parser_takes((uint32_t)len));
    p.fuse_emit = fused ? 1u : 0u;
    CU(cudaEventRecord(c->evp0, c->stream));
    if (tiny) {
        CU(sb::launch_tiny(p, c->stream, &c->launches));
        CU(cudaEventRecord(c->evp1, c->stream));
        CU(cudaEventRecord(c->evp2, c->stream));
        c->staged = true;
        return SNAPPY_B200_OK;
    }
    if (lane) { if (lane_parse_all(c, p) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR; }
    else if (gtab) CU(sb::launch_parse_gtab(p, gtab_region(c, 0, sb::gtab_tables()), sb::gtab_tables(),
                                            gtab_ticket(c, 1, sb::gtab_tables(), 0), c->stream, &c->launches));
    else CU(sb::launch_parse(p, kernel_mode(mode), c->stream, &c->launches));
    CU(cudaEventRecord(c->evp1, c->stream));
    if (debug_check(c, p) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    if (!fused) {
        if (!lane) CU(sb::launch_size(p, c->stream, &c->launches));
        CU(sb::launch_emit(p, c->stream, &c->launches));
    }
    CU(cudaEventRecord(c->evp2, c->stream));
    c->staged = true;
    return SNAPPY_B200_OK;
}

snappy_b200_status snappy_b200_compress(snappy_b200_ctx* c, const void* src, size_t n, void* dst,
                                        size_t dst_cap, size_t* out_n, int mode) {
    if (!dst || !out_n || (!src && n)) return fail(SNAPPY_B200_INVALID_INPUT, "null argument");
    if (n >= (1ull << 32)) return fail(SNAPPY_B200_INVALID_INPUT, "input of 2^32 bytes or more");
    if (dst_cap < snappy_b200_max_compressed_length(n))
        return fail(SNAPPY_B200_BUFFER_TOO_SMALL, "dst_cap < max_compressed_length(n)");
    if (!c) { snappy_b200_status st = default_ctx(&c); if (st != SNAPPY_B200_OK) return st; }
    if (bind(c) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    const size_t cap = snappy_b200_max_compressed_length(n);
    if (n >= kPipeMinBytes)
        return compress_host_pipelined(c, (const uint8_t*)src, n, (uint8_t*)dst, dst_cap, out_n, mode);
    CU(c->d_in.reserve(n + 64));
    CU(c->d_out.reserve(cap + 64));
    if (n) CU(cudaMemcpyAsync(c->d_in.p, src, n, cudaMemcpyHostToDevice, c->stream));
    CU(cudaEventRecord(c->ev0, c->stream));
    snappy_b200_status st = compress_dev(c, c->d_in.p, n, c->d_out.p, cap, c->d_total, mode, 1);
    if (st != SNAPPY_B200_OK) return st;
    CU(cudaEventRecord(c->ev1, c->stream));
    CU(cudaMemcpyAsync(c->h_total, c->d_total, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    const size_t total = (size_t)*c->h_total;
    if (total > dst_cap) return fail(SNAPPY_B200_BUFFER_TOO_SMALL, "stream larger than dst_cap");
    CU(cudaMemcpyAsync(dst, c->d_out.p, total, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (watchdog_status(c) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    cudaEventElapsedTime(&c->last_ms, c->ev0, c->ev1);
    *out_n = total;
    return SNAPPY_B200_OK;
}

snappy_b200_status snappy_b200_compress_batch(snappy_b200_ctx* c, const void* src, size_t len,
                                              size_t in_stride, size_t count, void* dst,
                                              size_t out_stride, uint32_t* sizes, int mode) {
    if (!dst || !sizes || (!src && len && count)) return fail(SNAPPY_B200_INVALID_INPUT, "null argument");
    if (!c) { snappy_b200_status st = default_ctx(&c); if (st != SNAPPY_B200_OK) return st; }
    if (bind(c) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    if (count == 0) return SNAPPY_B200_OK;
    if (in_stride < len) return fail(SNAPPY_B200_INVALID_INPUT, "in_stride < len");
    const size_t in_bytes = (count - 1) * in_stride + len, out_bytes = count * out_stride;
    CU(c->d_in.reserve(in_bytes + 64));
    CU(c->d_out.reserve(out_bytes + 64));
    CU(c->sizes.reserve((count + 1) * sizeof(uint32_t)));
    CU(cudaMemcpyAsync(c->d_in.p, src, in_bytes, cudaMemcpyHostToDevice, c->stream));
    CU(cudaEventRecord(c->ev0, c->stream));
    snappy_b200_status st = snappy_b200_compress_batch_device(c, c->d_in.p, len, in_stride, count, c->d_out.p,
                                                              out_stride, (uint32_t*)c->sizes.p, mode);
    if (st != SNAPPY_B200_OK) return st;
    CU(cudaEventRecord(c->ev1, c->stream));
    CU(cudaMemcpyAsync(sizes, c->sizes.p, count * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(dst, c->d_out.p, out_bytes, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (watchdog_status(c) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    cudaEventElapsedTime(&c->last_ms, c->ev0, c->ev1);
    return SNAPPY_B200_OK;
}

/* tests/compression.c:20-27: compress(source, uncompressed_length, dest) -> compressed length.
 * The caller sizes dest (no capacity argument in the reference contract). */
unsigned long snappy_b200_rocc_compress(char* source, int uncompressed_length, char* dest) {
    if (uncompressed_length < 0) { fail(SNAPPY_B200_INVALID_INPUT, "negative length"); return 0; }
    size_t out_n = 0;
    size_t n = (size_t)uncompressed_length;
    snappy_b200_status st = snappy_b200_compress(nullptr, source, n, dest,
                                                 snappy_b200_max_compressed_length(n), &out_n,
                                                 SNAPPY_B200_EXACT);
    return st == SNAPPY_B200_OK ? (unsigned long)out_n : 0ul;
}

snappy_b200_status snappy_b200_uncompressed_length(const void* src, size_t n, size_t* result) {
    if (!src || !result) return fail(SNAPPY_B200_INVALID_INPUT, "null argument");
    const uint8_t* p = (const uint8_t*)src;
    uint32_t v = 0, shift = 0;
    for (size_t i = 0; i < n && shift <= 28; ++i, shift += 7) {
        const uint32_t b = p[i];
        if (shift == 28 && (b & 0x7fu) > 15u) break;
        v |= (b & 0x7fu) << shift;
        if (b < 128) { *result = v; return SNAPPY_B200_OK; }
    }
    return fail(SNAPPY_B200_INVALID_INPUT, "malformed length preamble");
}

namespace {
// One raw stream on the device.  Large streams take the block-parallel path (which falls back to the
// serial walk on the device when the stream is not block structured); small ones go to one warp.
constexpr size_t kParDecodeMinBytes = 128u << 10;
snappy_b200_status decode_one(snappy_b200_ctx* c, const void* d_src, size_t n, void* d_dst, size_t dst_cap,
                              unsigned long long* d_out_n) {
    static const bool par_on = [] { const char* e = getenv("SNAPPY_B200_PAR_DECODE"); return !e || atoi(e) != 0; }();
    if (par_on && n >= kParDecodeMinBytes) {
        CU(c->raw_ix.reserve(sb::raw_decode_scratch_bytes(n, dst_cap) + 64));
        CU(sb::launch_decode_large((const uint8_t*)d_src, n, (uint8_t*)d_dst, dst_cap, c->raw_ix.p, d_out_n,
                                   c->stream, &c->launches));
    } else {
        CU(sb::launch_decode((const uint8_t*)d_src, 0, nullptr, (uint32_t)n, 1, (uint8_t*)d_dst, 0, dst_cap,
                             nullptr, d_out_n, c->stream, &c->launches));
    }
    return SNAPPY_B200_OK;
}
}  // namespace

snappy_b200_status snappy_b200_uncompress_device(snappy_b200_ctx* c, const void* d_src, size_t n,
                                                 void* d_dst, size_t dst_cap, unsigned long long* d_out_n) {
    if (!c || !d_src || !d_out_n || (!d_dst && dst_cap)) return fail(SNAPPY_B200_INVALID_INPUT, "null argument");
    if (n == 0 || n > 0xffffffffull) return fail(SNAPPY_B200_INVALID_INPUT, "stream length out of range");
    if (bind(c) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    return decode_one(c, d_src, n, d_dst, dst_cap, d_out_n);
}

snappy_b200_status snappy_b200_uncompress_batch_device(snappy_b200_ctx* c, const void* d_src,
                                                       size_t in_stride, const uint32_t* d_sizes,
                                                       size_t count, void* d_dst, size_t out_stride,
                                                       uint32_t* d_out_lens) {
    if (!c || !d_src || !d_sizes || !d_dst || !d_out_lens) return fail(SNAPPY_B200_INVALID_INPUT, "null argument");
    if (count > 0xffffffffull) return fail(SNAPPY_B200_INVALID_INPUT, "too many streams");
    if (bind(c) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    CU(sb::launch_decode((const uint8_t*)d_src, in_stride, d_sizes, 0, (uint32_t)count, (uint8_t*)d_dst,
                         out_stride, out_stride, d_out_lens, nullptr, c->stream, &c->launches));
    return SNAPPY_B200_OK;
}

snappy_b200_status snappy_b200_uncompress(snappy_b200_ctx* c, const void* src, size_t n, void* dst,
                                          size_t dst_cap, size_t* out_n) {
    if (!src || !out_n || (!dst && dst_cap)) return fail(SNAPPY_B200_INVALID_INPUT, "null argument");
    size_t ulen = 0;
    snappy_b200_status st = snappy_b200_uncompressed_length(src, n, &ulen);
    if (st != SNAPPY_B200_OK) return st;
    if (ulen > dst_cap) return fail(SNAPPY_B200_BUFFER_TOO_SMALL, "dst_cap < uncompressed length");
    if (n > 0xffffffffull) return fail(SNAPPY_B200_INVALID_INPUT, "stream length out of range");
    if (!c) { st = default_ctx(&c); if (st != SNAPPY_B200_OK) return st; }
    if (bind(c) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    CU(c->d_in.reserve(n + 64));
    CU(c->d_out.reserve(ulen + 64));
    CU(cudaMemcpyAsync(c->d_in.p, src, n, cudaMemcpyHostToDevice, c->stream));
    if (decode_one(c, c->d_in.p, n, c->d_out.p, ulen, c->d_total) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    CU(cudaMemcpyAsync(c->h_total, c->d_total, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (*c->h_total == ~0ull) return fail(SNAPPY_B200_INVALID_INPUT, "malformed Snappy stream");
    const size_t got = (size_t)*c->h_total;
    if (got) CU(cudaMemcpyAsync(dst, c->d_out.p, got, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    *out_n = got;
    return SNAPPY_B200_OK;
}

/* tests/compression.c:32-39: uncompress(source, compressed_length, dest) -> non-zero on failure.
 * The caller sizes dest (no capacity argument in the reference contract); the stream's own length
 * preamble says how much is written.  (The reference hardware left this command a stub,
 * CompressionAccelerator.scala:385-388.) */
unsigned long snappy_b200_rocc_uncompress(char* source, int compressed_length, char* dest) {
    if (compressed_length <= 0) { fail(SNAPPY_B200_INVALID_INPUT, "bad length"); return 1ul; }
    size_t ulen = 0, out_n = 0;
    if (snappy_b200_uncompressed_length(source, (size_t)compressed_length, &ulen) != SNAPPY_B200_OK) return 1ul;
    return snappy_b200_uncompress(nullptr, source, (size_t)compressed_length, dest, ulen, &out_n) == SNAPPY_B200_OK
               ? 0ul : 1ul;
}

/* ---- Snappy framing format (tests/snappy/framing_format.txt) ---------------------------------- */

size_t snappy_b200_frame_max_length(size_t n) {
    const size_t chunks = (n + sb::kBlock - 1) / sb::kBlock;
    return 10 + chunks * 40 + n + n / 6;                 /* id + per chunk: header, CRC, MaxCompressedLength slack */
}

snappy_b200_status snappy_b200_frame_compress_device(snappy_b200_ctx* c, const void* d_src, size_t n,
                                                     void* d_dst, size_t dst_cap,
                                                     unsigned long long* d_out_n, int mode) {
    if (!c || !d_dst || !d_out_n || (!d_src && n)) return fail(SNAPPY_B200_INVALID_INPUT, "null argument");
    return compress_dev(c, d_src, n, d_dst, dst_cap, d_out_n, mode, 2);
}

snappy_b200_status snappy_b200_frame_compress(snappy_b200_ctx* c, const void* src, size_t n, void* dst,
                                              size_t dst_cap, size_t* out_n, int mode) {
    if (!dst || !out_n || (!src && n)) return fail(SNAPPY_B200_INVALID_INPUT, "null argument");
    if (n >= (1ull << 32)) return fail(SNAPPY_B200_INVALID_INPUT, "input of 2^32 bytes or more");
    const size_t cap = snappy_b200_frame_max_length(n);
    if (dst_cap < cap) return fail(SNAPPY_B200_BUFFER_TOO_SMALL, "dst_cap < frame_max_length(n)");
    if (!c) { snappy_b200_status st = default_ctx(&c); if (st != SNAPPY_B200_OK) return st; }
    if (bind(c) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    CU(c->d_in.reserve(n + 64));
    CU(c->d_out.reserve(cap + 64));
    if (n) CU(cudaMemcpyAsync(c->d_in.p, src, n, cudaMemcpyHostToDevice, c->stream));
    CU(cudaEventRecord(c->ev0, c->stream));
    snappy_b200_status st = compress_dev(c, c->d_in.p, n, c->d_out.p, cap, c->d_total, mode, 2);
    if (st != SNAPPY_B200_OK) return st;
    CU(cudaEventRecord(c->ev1, c->stream));
    CU(cudaMemcpyAsync(c->h_total, c->d_total, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    const size_t total = (size_t)*c->h_total;
    CU(cudaMemcpyAsync(dst, c->d_out.p, total, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    cudaEventElapsedTime(&c->last_ms, c->ev0, c->ev1);
    *out_n = total;
    return SNAPPY_B200_OK;
}

snappy_b200_status snappy_b200_crc32c_chunks_device(snappy_b200_ctx* c, const void* d_src, size_t n,
                                                    size_t chunk_len, uint32_t* d_masked) {
    if (!c || !d_masked || (!d_src && n)) return fail(SNAPPY_B200_INVALID_INPUT, "null argument");
    if (chunk_len == 0 || chunk_len > sb::kBlock) return fail(SNAPPY_B200_INVALID_INPUT, "chunk_len must be 1..65536");
    if (bind(c) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    const size_t chunks = (n + chunk_len - 1) / chunk_len;
    if (chunks > 0xffffffffull) return fail(SNAPPY_B200_INVALID_INPUT, "too many chunks");
    CU(sb::launch_crc32c_chunks((const uint8_t*)d_src, nullptr, nullptr, chunk_len, n, (uint32_t)chunk_len,
                                (uint32_t)chunks, d_masked, c->stream, &c->launches));
    return SNAPPY_B200_OK;
}

/* Header walk on the host (metadata only, like reading the varint preamble of a raw stream). */
snappy_b200_status snappy_b200_frame_uncompressed_length(const void* src, size_t n, size_t* result) {
    if (!src || !result) return fail(SNAPPY_B200_INVALID_INPUT, "null argument");
    const uint8_t* s = (const uint8_t*)src;
    static const uint8_t id[10] = {0xff, 0x06, 0x00, 0x00, 0x73, 0x4e, 0x61, 0x50, 0x70, 0x59};
    if (n < 10 || memcmp(s, id, 10) != 0) return fail(SNAPPY_B200_INVALID_INPUT, "no stream identifier");
    size_t pos = 0, out = 0;
    while (pos < n) {
        if (pos + 4 > n) return fail(SNAPPY_B200_INVALID_INPUT, "truncated chunk header");
        const uint32_t type = s[pos], len = s[pos + 1] | (s[pos + 2] << 8) | ((uint32_t)s[pos + 3] << 16);
        if (pos + 4 + len > n) return fail(SNAPPY_B200_INVALID_INPUT, "truncated chunk");
        if (type == 0xff) {
            if (len != 6 || memcmp(s + pos, id, 10) != 0) return fail(SNAPPY_B200_INVALID_INPUT, "bad stream identifier");
        } else if (type == 0x00) {
            if (len < 4) return fail(SNAPPY_B200_INVALID_INPUT, "short chunk");
            size_t ulen = 0;
            if (snappy_b200_uncompressed_length(s + pos + 8, len - 4, &ulen) != SNAPPY_B200_OK) return SNAPPY_B200_INVALID_INPUT;
            if (ulen > sb::kBlock) return fail(SNAPPY_B200_INVALID_INPUT, "chunk larger than 65536 bytes");
            out += ulen;
        } else if (type == 0x01) {
            if (len < 4 || len - 4 > sb::kBlock) return fail(SNAPPY_B200_INVALID_INPUT, "bad uncompressed chunk");
            out += len - 4;
        } else if (type >= 0x02 && type <= 0x7f) {
            return fail(SNAPPY_B200_INVALID_INPUT, "reserved unskippable chunk");
        }
        pos += 4 + (size_t)len;
    }
    *result = out;
    return SNAPPY_B200_OK;
}

snappy_b200_status snappy_b200_frame_uncompress_device(snappy_b200_ctx* c, const void* d_src, size_t n,
                                                       void* d_dst, size_t dst_cap, size_t* out_n) {
    if (!c || !d_src || !out_n || (!d_dst && dst_cap)) return fail(SNAPPY_B200_INVALID_INPUT, "null argument");
    if (bind(c) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    CU(c->frame_res.reserve(256));
    sb::FrameWalkResult* d_res = (sb::FrameWalkResult*)c->frame_res.p;
    uint32_t* d_bad = (uint32_t*)((uint8_t*)c->frame_res.p + 64);
    sb::FrameWalkResult res;
    size_t max_records = 2 * (dst_cap >> 16) + 1024;
    // Fused reader (frame.cuh): the chain walk and the chunk decode in one kernel.  It pays when there are more
    // chunks than warp slots, so that decoding early chunks overlaps the rest of the walk; with fewer chunks
    // the separate walk is quicker (the hop is ~2x slower beside running decoders) and every chunk decodes at
    // once behind it.  Measured, ms fused / separate: 128 MiB 6.0 / 5.3, 256 MiB 8.6 / 7.1, 512 MiB 11.3 /
    // 12.3, 1 GiB 17.9 / 22.8.  SNAPPY_B200_FRAME_FUSED = 0 / 1 forces one or the other.
    static const int fused_env = [] { const char* e = getenv("SNAPPY_B200_FRAME_FUSED"); return e ? (atoi(e) != 0 ? 1 : 0) : -1; }();
    const bool fused = fused_env >= 0 ? fused_env == 1 : (dst_cap >> 16) >= 6144;
    void* d_ctl = (uint8_t*)c->frame_res.p + 128;
    for (int pass = 0;; ++pass) {
        CU(c->frame_ix.reserve(max_records * sb::kFrameRecordBytes + 64));
        if (fused)
            CU(sb::launch_frame_fused((const uint8_t*)d_src, n, (uint8_t*)d_dst, dst_cap, c->frame_ix.p, max_records, d_res,
                                      d_ctl, c->stream, &c->launches));
        else
            CU(sb::launch_frame_walk((const uint8_t*)d_src, n, c->frame_ix.p, max_records, d_res, c->stream, &c->launches));
        CU(cudaMemcpyAsync(&res, d_res, sizeof(res), cudaMemcpyDeviceToHost, c->stream));
        CU(cudaStreamSynchronize(c->stream));
        if (res.error != 2 || pass) break;
        max_records = (size_t)res.n_all;                 /* the index was too small: now it is exact */
    }
    if (res.error) return fail(SNAPPY_B200_INVALID_INPUT, "malformed framed stream");
    if (res.out_bytes > dst_cap) return fail(SNAPPY_B200_BUFFER_TOO_SMALL, "dst_cap < uncompressed length");
    if (res.n_chunks > 0xffffffffull) return fail(SNAPPY_B200_INVALID_INPUT, "too many chunks");
    CU(cudaMemsetAsync(d_bad, 0, sizeof(uint32_t), c->stream));
    CU(sb::launch_frame_decode((const uint8_t*)d_src, (uint8_t*)d_dst, c->frame_ix.p, max_records,
                               (uint32_t)res.n_chunks, d_bad, c->stream, &c->launches, fused));
    uint32_t bad = 0;
    CU(cudaMemcpyAsync(&bad, d_bad, sizeof(bad), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (bad) return fail(SNAPPY_B200_INVALID_INPUT, "a chunk is malformed or fails its CRC-32C");
    *out_n = (size_t)res.out_bytes;
    return SNAPPY_B200_OK;
}

snappy_b200_status snappy_b200_frame_uncompress(snappy_b200_ctx* c, const void* src, size_t n, void* dst,
                                                size_t dst_cap, size_t* out_n) {
    if (!src || !out_n || (!dst && dst_cap)) return fail(SNAPPY_B200_INVALID_INPUT, "null argument");
    size_t ulen = 0;
    snappy_b200_status st = snappy_b200_frame_uncompressed_length(src, n, &ulen);
    if (st != SNAPPY_B200_OK) return st;
    if (ulen > dst_cap) return fail(SNAPPY_B200_BUFFER_TOO_SMALL, "dst_cap < uncompressed length");
    if (!c) { st = default_ctx(&c); if (st != SNAPPY_B200_OK) return st; }
    if (bind(c) != SNAPPY_B200_OK) return SNAPPY_B200_CUDA_ERROR;
    CU(c->d_in.reserve(n + 64));
    CU(c->d_out.reserve(ulen + 64));
    CU(cudaMemcpyAsync(c->d_in.p, src, n, cudaMemcpyHostToDevice, c->stream));
    size_t got = 0;
    st = snappy_b200_frame_uncompress_device(c, c->d_in.p, n, c->d_out.p, ulen, &got);
    if (st != SNAPPY_B200_OK) return st;
    if (got) CU(cudaMemcpyAsync(dst, c->d_out.p, got, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    *out_n = got;
    return SNAPPY_B200_OK;
}

}  // extern "C"

/* ---- no-allocation entry point -----------------------------------------------------------------
 * No counterpart in the reference; added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *
 * snappy_b200_compress_device_noalloc is the exact-mode stream compressor of snappy_b200_compress_device
 * for inputs up to SNAPPY_B200_NOALLOC_MAX_INPUT bytes (1 MiB = 16 fragments), with every piece of scratch
 * the kernels need -- the fragments' sequence lists, their counts, sizes and output offsets -- in ONE
 * file-scope device array instead of the context's cudaMalloc'ed buffers.  It needs no context, creates no
 * stream or event, and calls no cudaMalloc / cudaFree / allocating routine: the caller passes input,
 * output, the length word and the stream.  Same kernels, same launch sequence (parse_pipe_kernel<true>,
 * size_block_kernel, scan_sizes_kernel, emit_kernel), so the stream is byte-identical to the existing
 * entry's.
 *
 * The array is shared by every call on a device and keeps the last call's contents; the words that are
 * read before this call's kernels have written them do not exist (every count, size and offset is written
 * by a kernel of the call before a later kernel reads it), but the small header is cleared on the stream
 * anyway so that a launch failure cannot leave a stale count behind.  Two calls that overlap on the GPU
 * (different streams, or graphs replayed concurrently) RACE on the array: serialise them.
 * Every GPU that loads the library holds the array (about 2 MiB) until the process exits. */
namespace {

constexpr size_t kNoAllocMaxInput = 1u << 20;                                  // bytes per call
constexpr size_t kNoAllocMaxFrags = kNoAllocMaxInput / sb::kBlock;             // 16
constexpr size_t kNoAllocSeqBytes = kNoAllocMaxFrags * (size_t)(sb::kBlock / 4 + 8) * sizeof(sb::seq_t);   // 2,098,176
constexpr size_t kNoAllocHdrBytes = 3 * 256;                                   // nseq | sizes | offsets, 256 bytes each
constexpr size_t kNoAllocScratchBytes = kNoAllocSeqBytes + kNoAllocHdrBytes;
static_assert(kNoAllocSeqBytes % 256 == 0, "header must stay 256-byte aligned");
static_assert((kNoAllocMaxFrags + 1) * sizeof(unsigned long long) <= 256, "offsets fit their slot");

// In host code this NAME is a host-side stand-in, not a device address: never hand it to a kernel.  The
// address comes from cudaGetSymbolAddress, once per device, with that device current (noalloc_scratch()).
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

constexpr int kNoAllocMaxDevices = 64;
std::mutex g_noalloc_mu;
unsigned char* g_noalloc_ptr[kNoAllocMaxDevices] = {};

// Device address of the scratch array on the CURRENT device (set-up on first use per device).
snappy_b200_status noalloc_scratch(unsigned char** out) {
    int dev = -1;
    CU(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kNoAllocMaxDevices) return fail(SNAPPY_B200_INVALID_INPUT, "noalloc: device ordinal out of range");
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {
        cudaError_t e = sb::kernels_init();                  // function attributes (no allocation); idempotent
        if (e != cudaSuccess) return fail(SNAPPY_B200_CUDA_ERROR, "noalloc: kernels_init", e);
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) return fail(SNAPPY_B200_CUDA_ERROR, "noalloc: cudaGetSymbolAddress(g_noalloc_scratch)", e);
        g_noalloc_ptr[dev] = (unsigned char*)p;
    }
    *out = g_noalloc_ptr[dev];
    return SNAPPY_B200_OK;
}

}  // namespace

extern "C" {

size_t snappy_b200_noalloc_max_input(void) { return kNoAllocMaxInput; }

snappy_b200_status snappy_b200_compress_device_noalloc(const void* d_src, size_t n, void* d_dst, size_t dst_cap,
                                                       unsigned long long* d_out_n, void* cuda_stream) {
    if (!d_dst || !d_out_n || (!d_src && n)) return fail(SNAPPY_B200_INVALID_INPUT, "null argument");
    if (n > kNoAllocMaxInput) {
        snprintf(g_err, sizeof g_err, "noalloc: input of %zu bytes exceeds the static scratch's maximum of %zu bytes "
                 "(SNAPPY_B200_NOALLOC_MAX_INPUT); use snappy_b200_compress_device", n, kNoAllocMaxInput);
        return SNAPPY_B200_INVALID_INPUT;
    }
    if (dst_cap < snappy_b200_max_compressed_length(n))
        return fail(SNAPPY_B200_BUFFER_TOO_SMALL, "dst_cap < max_compressed_length(n)");
    unsigned char* scratch = nullptr;
    snappy_b200_status st = noalloc_scratch(&scratch);
    if (st != SNAPPY_B200_OK) return st;
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    uint32_t* nseq = (uint32_t*)(scratch + kNoAllocSeqBytes);
    uint32_t* sizes = (uint32_t*)(scratch + kNoAllocSeqBytes + 256);
    unsigned long long* offsets = (unsigned long long*)(scratch + kNoAllocSeqBytes + 512);
    const uint32_t n_frag = (uint32_t)((n + sb::kBlock - 1) / sb::kBlock);
    sb::JobParams p;
    p.in = (const uint8_t*)d_src;
    p.n_total = n;
    p.in_stride = sb::kBlock;
    p.job_len = sb::kBlock;
    p.n_jobs = n_frag;
    p.seqs = (sb::seq_t*)scratch;
    p.nseq = nseq;
    p.sizes = sizes;
    p.with_header = 0;
    p.sparse = 0;
    p.out = (uint8_t*)d_dst;
    p.out_stride = 0;
    p.offsets = offsets;
    p.err = nullptr;
    CU(cudaMemsetAsync(scratch + kNoAllocSeqBytes, 0, kNoAllocHdrBytes, stream));   // the shared header starts at zero
    CU(sb::launch_parse(p, SNAPPY_B200_EXACT, stream, nullptr));
    CU(sb::launch_size(p, stream, nullptr));
    CU(sb::launch_scan(sizes, n_frag, offsets, sb::varint32_len((uint32_t)n), (uint32_t)n, (uint8_t*)d_dst, d_out_n,
                       stream, nullptr));
    CU(sb::launch_emit(p, stream, nullptr));
    return SNAPPY_B200_OK;
}

}  // extern "C"
