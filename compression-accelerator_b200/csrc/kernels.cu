// kernels.cu -- Snappy block-compression kernels for B200 (sm_100a) and their launchers.
//
//   parse_kernel   one CTA per fragment (stream path) or per small buffer (batch path): the
//                  fragment is staged into shared memory by a 1-D TMA bulk copy, the uint16 hash
//                  table lives in shared memory, and Snappy's greedy parse with its skip
//                  heuristic (tests/snappy/snappy.cc:539-671) produces a sequence list.
//   size_kernel / scan_sizes_kernel / emit_kernel (emit.cuh): sizes, offsets (+ varint header,
//                  snappy.cc:1006-1061) and the final elements, written once.
#include "kernels.h"
#include "snappy_device.cuh"
#include "parse_serial.cuh"
#include "parse_window.cuh"
#include "emit.cuh"

namespace sb {

// Shared-memory layout of one parse CTA: [0,16) mbarrier | [128, 128+table) table | input.
constexpr uint32_t kSmemTableOff = 128;
constexpr uint32_t kInputPad = 32;   // slack so unaligned word reads past the end stay in bounds

// peer-detector scratch (parse_window.cuh): fewer slots for small jobs so more of them fit an SM
__host__ __device__ inline uint32_t peer_slots(uint32_t job_len) {
    return job_len > 16384 ? kPeerSlots : 512u;
}
__host__ __device__ inline uint32_t parse_smem_bytes(uint32_t job_len) {
    return kSmemTableOff + table_entries(job_len) * 2 + peer_slots(job_len) * 4 +
           ((job_len + 15u) & ~15u) + kInputPad;
}

// Stage src[0,len) into shared memory: TMA bulk copy for the 16-byte-aligned body, ordinary
// loads for the tail (or for everything when the source is not 16-byte aligned), while the
// warp zeroes the hash table (snappy.cc:475-481).
__device__ __forceinline__ void stage_fragment(const uint8_t* __restrict__ src, uint32_t len,
                                               uint8_t* s_in, uint16_t* table, uint32_t entries,
                                               uint64_t* bar, int lane) {
    const bool aligned = ((uintptr_t)src & 15u) == 0;
    const uint32_t bulk = aligned ? (len & ~15u) : 0u;
    if (bulk) {
        if (lane == 0) {
            mbar_init(bar, 1);
            mbar_expect_tx(bar, bulk);
            tma_load_1d(s_in, src, bulk, bar);
        }
    }
    uint4* t4 = reinterpret_cast<uint4*>(table);
    const uint4 z = make_uint4(0, 0, 0, 0);
    for (uint32_t i = lane; i < entries / 8; i += 32) t4[i] = z;
    for (uint32_t i = bulk + lane; i < len; i += 32) s_in[i] = src[i];
    uint32_t padded = ((len + 15u) & ~15u) + kInputPad;
    for (uint32_t i = len + lane; i < padded; i += 32) s_in[i] = 0;
    __syncwarp();
    if (bulk) mbar_wait(bar, 0);
}

template <bool kWindow>
__global__ void __launch_bounds__(32) parse_kernel(JobParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int lane = threadIdx.x;
    const uint32_t job = blockIdx.x;
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem);
    uint16_t* table = reinterpret_cast<uint16_t*>(smem + kSmemTableOff);
    const uint32_t len = job_len_of(p, job);
    const uint32_t entries = table_entries(len);            // from THIS fragment's length
    uint32_t* scr = reinterpret_cast<uint32_t*>(smem + kSmemTableOff + table_entries(p.job_len) * 2);
    const uint32_t nscr = peer_slots(p.job_len);
    uint8_t* s_in = reinterpret_cast<uint8_t*>(scr + nscr);
    if (kWindow) for (uint32_t i = lane; i < nscr; i += 32) scr[i] = kFull;
    stage_fragment(p.in + (uint64_t)job * p.in_stride, len, s_in, table, entries, bar, lane);
    SeqWriter out{p.seqs + (uint64_t)job * seq_stride(p.job_len), 0};
    const int shift = 32 - (31 - __clz(entries));           // snappy.cc:548
    if (kWindow) parse_fragment_window(s_in, len, table, scr, nscr - 1, shift, out, lane);
    else parse_fragment_serial(s_in, len, table, shift, out, lane);
    if (lane == 0) p.nseq[job] = out.n;
}

// ---- scan of fragment sizes + header ---------------------------------------------------------
__global__ void __launch_bounds__(1024) scan_sizes_kernel(const uint32_t* __restrict__ sizes,
                                                          uint32_t n, unsigned long long* offsets,
                                                          uint32_t hdr_len, uint32_t n_input,
                                                          uint8_t* dst,
                                                          unsigned long long* d_total) {
    __shared__ unsigned long long warp_sum[32];
    __shared__ unsigned long long carry_s;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) {
        carry_s = hdr_len;
        if (hdr_len) put_varint32(dst, n_input);             // snappy.cc:1006-1009
    }
    __syncthreads();
    for (uint32_t base = 0; base < n; base += 1024) {
        uint32_t i = base + tid;
        unsigned long long v = i < n ? sizes[i] : 0ull, x = v;
        for (int d = 1; d < 32; d <<= 1) {
            unsigned long long y = __shfl_up_sync(kFull, x, d);
            if (lane >= d) x += y;
        }
        if (lane == 31) warp_sum[wid] = x;
        __syncthreads();
        if (wid == 0) {
            unsigned long long s = warp_sum[lane], t = s;
            for (int d = 1; d < 32; d <<= 1) {
                unsigned long long y = __shfl_up_sync(kFull, t, d);
                if (lane >= d) t += y;
            }
            warp_sum[lane] = t - s;                           // exclusive over warps
        }
        __syncthreads();
        unsigned long long excl = carry_s + warp_sum[wid] + (x - v);
        if (i < n) offsets[i] = excl;
        __syncthreads();
        if (tid == 1023) carry_s = excl + v;
        __syncthreads();
    }
    if (tid == 0) *d_total = carry_s;
}

// ---- launchers ----------------------------------------------------------------------------------
cudaError_t kernels_init() {
    cudaError_t e = cudaFuncSetAttribute(parse_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)parse_smem_bytes(kBlock));
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(parse_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)parse_smem_bytes(kBlock));
}

cudaError_t launch_parse(const JobParams& p, int mode, cudaStream_t st,
                         unsigned long long* launches) {
    if (p.n_jobs == 0) return cudaSuccess;
    if (mode == 2) parse_kernel<false><<<p.n_jobs, 32, parse_smem_bytes(p.job_len), st>>>(p);
    else parse_kernel<true><<<p.n_jobs, 32, parse_smem_bytes(p.job_len), st>>>(p);
    if (launches) ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_size(const JobParams& p, cudaStream_t st, unsigned long long* launches) {
    if (p.n_jobs == 0) return cudaSuccess;
    size_kernel<<<(p.n_jobs + 3) / 4, 128, 0, st>>>(p);
    if (launches) ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_emit(const JobParams& p, cudaStream_t st, unsigned long long* launches) {
    if (p.n_jobs == 0) return cudaSuccess;
    emit_kernel<<<p.n_jobs, kEmitThreads, 0, st>>>(p);
    if (launches) ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_scan(const uint32_t* sizes, uint32_t n_frag, unsigned long long* offsets,
                        uint32_t hdr_len, uint32_t n_input, uint8_t* dst,
                        unsigned long long* d_total, cudaStream_t st,
                        unsigned long long* launches) {
    scan_sizes_kernel<<<1, 1024, 0, st>>>(sizes, n_frag, offsets, hdr_len, n_input, dst, d_total);
    if (launches) ++*launches;
    return cudaGetLastError();
}

}  // namespace sb
