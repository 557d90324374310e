// randlat.cu -- memory-system probe behind the lane parser's design (scripts/probes, not product).
// J threads (fpw active lanes per warp), each walking ITS OWN region of `region` bytes with a chain of
// dependent random 16-byte accesses, like one fragment's hash table.  Modes:
//   0 load                      1 load + store same slot          2 = 1 + prefetch.L2 of another slot
//   3 = 1 + speculative load of another slot (consumed next iteration)   4 = 1 + store to another slot
//   5 load + store, regions packed 8 KB apart (small footprint control)
// Prints ns per iteration (whole kernel time / iters).
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <stdint.h>

__device__ __forceinline__ uint32_t mix(uint32_t x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }

template <int MODE>
__global__ void __launch_bounds__(32) walk(uint4* pool, size_t region_entries, int iters, uint32_t fpw, uint32_t jobs,
                                           unsigned long long* sink) {
    const uint32_t lane = threadIdx.x;
    if (lane >= fpw) return;
    const uint32_t job = blockIdx.x * fpw + lane;
    if (job >= jobs) return;
    uint4* t = pool + (size_t)job * region_entries;
    const uint32_t mask = (uint32_t)region_entries - 1;
    uint32_t x = mix(job + 1);
    uint4 spec = make_uint4(0, 0, 0, 0);
    unsigned long long acc = 0;
    for (int i = 0; i < iters; ++i) {
        const uint32_t a = mix(x) & mask, b = mix(x + 0x9e3779b9u) & mask;
        uint4 v = __ldcg(t + a);
        if (MODE == 2) asm volatile("prefetch.global.L2 [%0];" ::"l"(t + b));
        if (MODE == 3) { acc += spec.y; spec = __ldcg(t + b); }
        if (MODE >= 1) __stcg(t + a, make_uint4(x, i, 0, 0));
        if (MODE == 4) __stcg(t + b, make_uint4(x, i, 1, 1));
        x = x * 0x9e3779b1u + v.x + 1;                           // the next address depends on the load
    }
    sink[job] = acc + x;
}

int main(int argc, char** argv) {
    const int mode = argc > 1 ? atoi(argv[1]) : 0;
    const uint32_t jobs = argc > 2 ? atoi(argv[2]) : 16384;
    const uint32_t fpw = argc > 3 ? atoi(argv[3]) : 8;
    const int iters = argc > 4 ? atoi(argv[4]) : 2000;
    size_t region = argc > 5 ? (size_t)atol(argv[5]) : (256u << 10);
    const size_t entries = region / 16;
    uint4* pool; unsigned long long* sink;
    cudaMalloc(&pool, (size_t)jobs * region);
    cudaMalloc(&sink, jobs * 8);
    cudaMemset(pool, 0, (size_t)jobs * region);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const uint32_t grid = (jobs + fpw - 1) / fpw;
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0);
        switch (mode) {
            case 0: walk<0><<<grid, 32>>>(pool, entries, iters, fpw, jobs, sink); break;
            case 1: walk<1><<<grid, 32>>>(pool, entries, iters, fpw, jobs, sink); break;
            case 2: walk<2><<<grid, 32>>>(pool, entries, iters, fpw, jobs, sink); break;
            case 3: walk<3><<<grid, 32>>>(pool, entries, iters, fpw, jobs, sink); break;
            default: walk<4><<<grid, 32>>>(pool, entries, iters, fpw, jobs, sink); break;
        }
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    cudaError_t err = cudaGetLastError();
    printf("mode %d jobs %u fpw %u region %zu KB footprint %.0f MB: %.3f ms, %.1f ns/iter, %.2f G iter/s  %s\n", mode, jobs, fpw,
           region >> 10, (double)jobs * region / 1e6, best, best * 1e6 / iters, (double)jobs * iters / best / 1e6,
           err == cudaSuccess ? "" : cudaGetErrorString(err));
    return 0;
}
