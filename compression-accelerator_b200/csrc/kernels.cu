// kernels.cu -- Snappy block-compression kernels for B200 (sm_100a).
//
//   K1/K2 parse_exact_kernel  one warp per fragment (stream path) or per small buffer (batch
//                             path): fragment staged into shared memory by a 1-D TMA bulk copy,
//                             uint16 hash table in shared memory, Snappy's greedy parse with the
//                             skip heuristic restated from tests/snappy/snappy.cc:539-671.
//   K3   scan_sizes_kernel + compact_kernel: exclusive scan of fragment sizes, varint header,
//                             coalesced gather into the final stream (snappy.cc:1006-1061).
#include "kernels.h"
#include "snappy_device.cuh"

namespace sb {

// Shared-memory layout of one parse CTA: [0,16) mbarrier | [128, 128+table) table | input.
constexpr uint32_t kSmemTableOff = 128;
constexpr uint32_t kInputPad = 32;   // slack so unaligned word reads past the end stay in bounds

__host__ __device__ inline uint32_t parse_smem_bytes(uint32_t job_len) {
    return kSmemTableOff + table_entries(job_len) * 2 + ((job_len + 15u) & ~15u) + kInputPad;
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
    // tail / unaligned body, then zero the padding
    for (uint32_t i = bulk + lane; i < len; i += 32) s_in[i] = src[i];
    uint32_t padded = ((len + 15u) & ~15u) + kInputPad;
    for (uint32_t i = len + lane; i < padded; i += 32) s_in[i] = 0;
    __syncwarp();
    if (bulk) mbar_wait(bar, 0);
}

// Greedy parse of one fragment, executed uniformly by the whole warp (every lane follows the
// same control flow; the lanes split literal copies, long copies and match extension).
// Restates snappy.cc:539-671; positions are offsets from the fragment start.
__device__ __forceinline__ uint32_t parse_fragment_exact(const uint8_t* s, uint32_t len,
                                                         uint16_t* table, int shift,
                                                         uint8_t* out, uint32_t op, int lane) {
    uint32_t ip = 0, next_emit = 0;
    if (len >= kMargin) {                                   // snappy.cc:557
        const uint32_t ip_limit = len - kMargin;            // :558
        ip = 1;                                             // :560 Hash(++ip)
        for (;;) {
            uint32_t skip = 32, cand, nxt = ip, w;
            for (;;) {                                      // :587-608 scan with accelerating skip
                ip = nxt;
                uint32_t step = skip >> 5;
                skip += step;
                nxt = ip + step;
                if (nxt > ip_limit) goto remainder;         // :598-600
                w = lds32u(s, ip);
                uint32_t h = hash32(w, shift);
                cand = table[h];
                table[h] = (uint16_t)ip;                    // :606
                if (lds32u(s, cand) == w) break;            // :607
            }
            op = warp_emit_literal(out, op, s, next_emit, ip - next_emit, lane);   // :614
            for (;;) {                                      // :627-656 copy loop
                uint32_t base = ip;
                uint32_t matched = 4 + warp_match_len(s, cand + 4, ip + 4, len, lane);
                ip += matched;
                op = warp_emit_copy(out, op, base - cand, matched, lane);
                next_emit = ip;
                if (ip >= ip_limit) goto remainder;         // :643-645
                uint32_t w0 = lds32u(s, ip - 1);
                w = lds32u(s, ip);
                table[hash32(w0, shift)] = (uint16_t)(ip - 1);      // :649-651
                uint32_t h = hash32(w, shift);
                cand = table[h];
                table[h] = (uint16_t)ip;                    // :652-655
                if (lds32u(s, cand) != w) break;            // :656
            }
            ip++;                                           // :658-659
        }
    }
remainder:
    if (next_emit < len)                                    // :665-668
        op = warp_emit_literal(out, op, s, next_emit, len - next_emit, lane);
    return op;
}

__global__ void __launch_bounds__(32) parse_exact_kernel(ParseParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int lane = threadIdx.x;
    const uint32_t job = blockIdx.x;
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem);
    uint16_t* table = reinterpret_cast<uint16_t*>(smem + kSmemTableOff);

    const uint64_t in_off = (uint64_t)job * p.in_stride;
    const uint64_t left = p.n_total - in_off;
    const uint32_t len = left < p.job_len ? (uint32_t)left : p.job_len;
    const uint32_t entries = table_entries(len);            // from THIS fragment's length
    uint8_t* s_in = smem + kSmemTableOff + table_entries(p.job_len) * 2;
    const uint8_t* src = p.in + in_off;
    uint8_t* out = p.out + (uint64_t)job * p.out_stride;

    stage_fragment(src, len, s_in, table, entries, bar, lane);

    uint32_t op = 0;
    if (p.with_header) {
        if (lane == 0) put_varint32(out, len);
        op = varint32_len(len);
    }
    const int shift = 32 - (31 - __clz(entries));           // snappy.cc:548
    op = parse_fragment_exact(s_in, len, table, shift, out, op, lane);
    if (lane == 0) p.sizes[job] = op;
}

// ---- K3: scan of fragment sizes + header -----------------------------------------------------
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
        unsigned long long carry = carry_s;
        unsigned long long excl = carry + warp_sum[wid] + (x - v);
        if (i < n) offsets[i] = excl;
        __syncthreads();
        if (tid == 1023) carry_s = excl + v;
        __syncthreads();
    }
    if (tid == 0) *d_total = carry_s;
}

// ---- K3: gather staged fragment outputs (snappy.cc:1060-1061 "written += ...") --------------
// One CTA per fragment.  Destination offset is arbitrary, the staged source is 16-byte
// aligned: destination-aligned 16-byte stores assembled from aligned 4-byte source words.
__global__ void __launch_bounds__(256) compact_kernel(const uint8_t* __restrict__ staging,
                                                      uint64_t stage_stride,
                                                      const uint32_t* __restrict__ sizes,
                                                      const unsigned long long* __restrict__ offsets,
                                                      uint8_t* __restrict__ dst) {
    const uint32_t f = blockIdx.x;
    const uint32_t n = sizes[f];
    const uint8_t* src = staging + (uint64_t)f * stage_stride;
    uint8_t* d = dst + offsets[f];
    const uint32_t tid = threadIdx.x;
    uint32_t head = (uint32_t)((16u - ((uintptr_t)d & 15u)) & 15u);
    if (head > n) head = n;
    if (tid < head) d[tid] = src[tid];
    const uint32_t body = (n - head) >> 4;
    uint4* d4 = reinterpret_cast<uint4*>(d + head);
    const uint32_t* w = reinterpret_cast<const uint32_t*>(src);
    const uint32_t sh = (head & 3u) << 3;
    for (uint32_t j = tid; j < body; j += 256) {
        uint32_t q = (head + (j << 4)) >> 2;
        uint32_t a0 = __ldg(w + q), a1 = __ldg(w + q + 1), a2 = __ldg(w + q + 2),
                 a3 = __ldg(w + q + 3), a4 = __ldg(w + q + 4);
        uint4 v;
        v.x = __funnelshift_r(a0, a1, sh);
        v.y = __funnelshift_r(a1, a2, sh);
        v.z = __funnelshift_r(a2, a3, sh);
        v.w = __funnelshift_r(a3, a4, sh);
        d4[j] = v;
    }
    const uint32_t done = head + (body << 4);
    if (done + tid < n) d[done + tid] = src[done + tid];
}

// ---- launchers ----------------------------------------------------------------------------------
cudaError_t kernels_init() {
    return cudaFuncSetAttribute(parse_exact_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)parse_smem_bytes(kBlock));
}

cudaError_t launch_parse(const ParseParams& p, int mode, cudaStream_t st,
                         unsigned long long* launches) {
    (void)mode;
    if (p.n_jobs == 0) return cudaSuccess;
    parse_exact_kernel<<<p.n_jobs, 32, parse_smem_bytes(p.job_len), st>>>(p);
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

cudaError_t launch_compact(const uint8_t* staging, uint64_t stage_stride, const uint32_t* sizes,
                           const unsigned long long* offsets, uint32_t n_frag, uint8_t* dst,
                           cudaStream_t st, unsigned long long* launches) {
    if (n_frag == 0) return cudaSuccess;
    compact_kernel<<<n_frag, 256, 0, st>>>(staging, stage_stride, sizes, offsets, dst);
    if (launches) ++*launches;
    return cudaGetLastError();
}

}  // namespace sb
