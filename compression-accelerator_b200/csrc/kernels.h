// kernels.h -- host-side launchers of the Snappy kernels (internal to libsnappy_b200).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace sb {

// Per-fragment staging stride in the stream path: MaxCompressedLength(65536) = 76490
// (tests/snappy/snappy.cc:99-121) rounded up to a multiple of 128.
constexpr uint32_t kStageStride = 76544;

// One "job" = one fragment / one small buffer, parsed by one warp.
struct ParseParams {
    const uint8_t* in;     // job j reads in + j*in_stride
    uint64_t n_total;      // bytes reachable from `in` (job length = min(job_len, n_total - j*in_stride))
    uint64_t in_stride;
    uint32_t job_len;      // 65536 in the stream path, buffer length in the batch path
    uint8_t* out;          // job j writes out + j*out_stride
    uint64_t out_stride;
    uint32_t* sizes;       // sizes[j] = bytes written by job j (header included when with_header)
    uint32_t n_jobs;
    uint32_t with_header;  // 1: each job is its own stream and starts with varint32(len)
};

cudaError_t launch_parse(const ParseParams& p, int mode, cudaStream_t st, unsigned long long* launches);

// Exclusive scan of sizes[0..n_frag) (+hdr_len) into offsets (uint64), total stream length into
// *d_total; writes varint32(n_input) at dst when hdr_len > 0.
cudaError_t launch_scan(const uint32_t* sizes, uint32_t n_frag, unsigned long long* offsets,
                        uint32_t hdr_len, uint32_t n_input, uint8_t* dst,
                        unsigned long long* d_total, cudaStream_t st, unsigned long long* launches);

// Gather staged fragment outputs into the final contiguous stream.
cudaError_t launch_compact(const uint8_t* staging, uint64_t stage_stride, const uint32_t* sizes,
                           const unsigned long long* offsets, uint32_t n_frag, uint8_t* dst,
                           cudaStream_t st, unsigned long long* launches);

cudaError_t kernels_init();  // one-time function attributes (dynamic shared memory opt-in)

}  // namespace sb
