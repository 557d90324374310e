// peer.cuh -- the multi-GPU exchange step over NVLink peer memory (include/snappy_b200.h,
// snappy_b200_peer_*): one process per GPU, no collective library on the data path.
//
// What is exchanged (reference tests/snappy/snappy.cc:1061, the running output offset
// is the only datum that crosses a fragment / stream boundary): every rank's payload size, and the
// payloads themselves, which land back to back in an archive buffer on the root.  Each rank owns a
// small MAILBOX in its HBM that every peer maps through CUDA IPC; the root also owns the archive
// (two slots, steps alternate).  A step is three small kernels on the rank's exchange stream, behind
// the compress kernels of that step:
//   peer_publish_kernel  store (step, my size) into every rank's mailbox; wait for the sizes of the
//                        ranks below me; my offset = their sum
//   peer_copy_kernel     my payload -> root archive + offset, 16-byte stores through NVLink
//   peer_finish_kernel   fence, then store "step done" into every rank's mailbox
// Nothing here needs the host to know a size, so the next step's compress kernels are already
// queued while the payload moves (the copy kernel is small enough to run beside the parser's CTAs).
// Back-pressure: a rank starts step k only when every rank finished step k-2, so two mailbox /
// archive slots suffice.  Every wait is bounded (about two seconds) and raises an error flag instead
// of hanging the GPU.
#pragma once
#include "kernels.h"
#include "snappy_device.cuh"
#include "emit.cuh"
#include <cstdlib>

namespace sb {

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// Errors are raised in the ROOT's mailbox as well as in the rank's own: the root is where
// snappy_b200_peer_result looks, so a rank that timed out or overflowed its slot fails the step for all.
__device__ __forceinline__ void peer_raise(const PeerTable& t, uint32_t rank, uint32_t root, unsigned long long code) {
    t.box[rank]->error = code;
    if (root != rank) st_release_sys(&t.box[root]->error, code);
}

__global__ void __launch_bounds__(kPeerMaxWorld) peer_publish_kernel(PeerTable t, uint32_t rank, uint32_t world, uint32_t root,
                                                                     unsigned long long step, unsigned long long base,
                                                                     const unsigned long long* __restrict__ d_size,
                                                                     unsigned long long* d_offset) {
    __shared__ unsigned long long part[kPeerMaxWorld];
    __shared__ int bad;
    const uint32_t tid = threadIdx.x;
    PeerMailbox* mine = t.box[rank];
    if (tid == 0) bad = 0;
    __syncthreads();
    const long long t0 = clock64();
    if (tid < world && step >= 2)                            // every rank is done with the slot I reuse
        while (ld_acquire_sys(&mine->finished[tid]) + 1 < step)
            if (clock64() - t0 > kPeerSpinCycles) { bad = 1; break; }
    __syncthreads();
    const unsigned long long size = *d_size;
    const unsigned long long tag = (step + 1) << kPeerSizeBits;
    if (tid < world) st_release_sys(&t.box[tid]->sizes[step & 1][rank], tag | size);
    unsigned long long below = 0;
    // Sizes of the ranks below me (my offset), and the ROOT's size word whatever its rank: the root
    // publishes step k only after its host is done with the slot's previous content (step k-2), so no
    // payload of step k reaches the slot before that -- also when the root is not rank 0.
    if (tid < rank || (tid == root && tid != rank && tid < world)) {
        unsigned long long v;
        while (((v = ld_acquire_sys(&mine->sizes[step & 1][tid])) >> kPeerSizeBits) != step + 1)
            if (clock64() - t0 > 2 * kPeerSpinCycles) { bad = 1; break; }
        if (tid < rank) below = v & ((1ull << kPeerSizeBits) - 1);
    }
    part[tid] = below;
    __syncthreads();
    if (tid == 0) {
        unsigned long long off = base;                       // bytes the root keeps in front (a stream header)
        for (uint32_t r = 0; r < rank; ++r) off += part[r];
        if (bad) { off = ~0ull; peer_raise(t, rank, root, 1); }
        *d_offset = off;
    }
}

// block_copy (emit.cuh) with four 16-byte chunks in flight per thread: the stores cross NVLink, the
// loads come from local HBM, and a thread that waits for one load per store leaves the link idle.
__device__ __forceinline__ void peer_tile_copy(uint8_t* __restrict__ d, const uint8_t* __restrict__ s,
                                               uint32_t n, uint32_t tid, uint32_t nthreads) {
    uint32_t head = (uint32_t)((16u - ((uintptr_t)d & 15u)) & 15u);
    if (head > n) head = n;
    if (tid < head) d[tid] = s[tid];
    const uint32_t body = (n - head) >> 4;
    if (body) {
        uint4* d4 = reinterpret_cast<uint4*>(d + head);
        const uint8_t* sb_ = s + head;
        const uint32_t* w = reinterpret_cast<const uint32_t*>((uintptr_t)sb_ & ~(uintptr_t)3);
        const uint32_t sh = ((uint32_t)(uintptr_t)sb_ & 3u) << 3;
        for (uint32_t j0 = tid; j0 < body; j0 += 4 * nthreads) {
            uint32_t a[4][5];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const uint32_t j = j0 + u * nthreads;
                if (j < body) {
                    const uint32_t* q = w + (j << 2);
                    a[u][0] = __ldg(q); a[u][1] = __ldg(q + 1); a[u][2] = __ldg(q + 2); a[u][3] = __ldg(q + 3);
                    a[u][4] = sh ? __ldg(q + 4) : 0u;  // never read past the source when aligned
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const uint32_t j = j0 + u * nthreads;
                if (j < body) {
                    uint4 v;
                    v.x = shf_r(a[u][0], a[u][1], sh);
                    v.y = shf_r(a[u][1], a[u][2], sh);
                    v.z = shf_r(a[u][2], a[u][3], sh);
                    v.w = shf_r(a[u][3], a[u][4], sh);
                    d4[j] = v;
                }
            }
        }
    }
    const uint32_t done = head + (body << 4);
    if (done + tid < n) d[done + tid] = s[done + tid];
}

__global__ void __launch_bounds__(256) peer_copy_kernel(PeerTable t, uint32_t rank, uint32_t root,
                                                        const uint8_t* __restrict__ src,
                                                        const unsigned long long* __restrict__ d_size,
                                                        uint8_t* archive, unsigned long long archive_cap,
                                                        const unsigned long long* __restrict__ d_offset, uint32_t tile) {
    const unsigned long long n = *d_size, off = *d_offset;
    if (off == ~0ull) return;
    if (off + n > archive_cap) { if (threadIdx.x == 0 && blockIdx.x == 0) peer_raise(t, rank, root, 2); return; }
    const unsigned long long tiles = (n + tile - 1) / tile;
    for (unsigned long long k = blockIdx.x; k < tiles; k += gridDim.x) {
        const unsigned long long lo = k * tile;
        const uint32_t len = (uint32_t)(n - lo < tile ? n - lo : tile);
        peer_tile_copy(archive + off + lo, src + lo, len, threadIdx.x, blockDim.x);
    }
}

__global__ void __launch_bounds__(kPeerMaxWorld) peer_finish_kernel(PeerTable t, uint32_t rank, uint32_t world,
                                                                    unsigned long long step) {
    __threadfence_system();                                  // the copy kernel's stores, then the flag
    if (threadIdx.x < world) st_release_sys(&t.box[threadIdx.x]->finished[rank], step + 1);
}

__global__ void __launch_bounds__(kPeerMaxWorld) peer_wait_kernel(PeerMailbox* mine, uint32_t world,
                                                                  unsigned long long step) {
    const long long t0 = clock64();
    if (threadIdx.x < world)
        while (ld_acquire_sys(&mine->finished[threadIdx.x]) < step + 1)
            if (clock64() - t0 > 2 * kPeerSpinCycles) { mine->error = 3; break; }
}


cudaError_t launch_peer_step(const PeerTable& t, uint32_t rank, uint32_t world, uint32_t root, unsigned long long step,
                             unsigned long long base, const uint8_t* payload, const unsigned long long* d_size, uint8_t* archive_slot,
                             unsigned long long slot_bytes, unsigned long long* d_offset,
                             cudaStream_t st, cudaEvent_t copied,
                             unsigned long long* launches) {
    peer_publish_kernel<<<1, kPeerMaxWorld, 0, st>>>(t, rank, world, root, step, base, d_size, d_offset);
    // Few CTAs on purpose: 32 reach ~500 GB/s over NVLink when alone, and remote stores back up the
    // load/store path of the SM they run on, which slows the parser's CTAs sharing it (2 GPUs, ms per
    // step: 296 CTAs 33.70, 32 CTAs 33.36, 12 CTAs 33.61).  SNAPPY_B200_PEER_CTAS / _TILE_KB for A/B.
    static const uint32_t ctas = [] { const char* e = getenv("SNAPPY_B200_PEER_CTAS"); return e ? (uint32_t)atoi(e) : 32u; }();
    static const uint32_t tile = [] { const char* e = getenv("SNAPPY_B200_PEER_TILE_KB"); return (e ? (uint32_t)atoi(e) : 256u) << 10; }();
    peer_copy_kernel<<<ctas, 256, 0, st>>>(t, rank, root, payload, d_size, archive_slot, slot_bytes, d_offset, tile);
    cudaError_t e = cudaEventRecord(copied, st);
    if (e != cudaSuccess) return e;
    peer_finish_kernel<<<1, kPeerMaxWorld, 0, st>>>(t, rank, world, step);
    if (launches) *launches += 3;
    return cudaGetLastError();
}

cudaError_t launch_peer_wait(PeerMailbox* mine, uint32_t world, unsigned long long step, cudaStream_t st,
                             unsigned long long* launches) {
    peer_wait_kernel<<<1, kPeerMaxWorld, 0, st>>>(mine, world, step);
    if (launches) ++*launches;
    return cudaGetLastError();
}

}  // namespace sb
