// randlat2.cu -- second memory-system probe (scripts/probes, not product): which instruction flavours
// make "read a random 16-byte slot of my own region, then overwrite it" fast?  Same set-up as randlat.cu.
//   LD flavour: 0 ld.cg  1 ld (ca)  2 ld.L1::no_allocate  3 ld.volatile  4 ld.cv
//   ST flavour: 0 st.cg  1 st (wb)  2 st.cs  3 st.wt
//   PAT: 0 load a, store a                 1 load a, store a one iteration LATE (after the next load issued)
//        2 atom.exch.b128 a                3 atom.exch.b128 a + store b       4 load a only
//        5 load a, store b (different slot) 6 exch a + load b (speculative)
//        7 exch a + 256-bit store b (32-byte slots: a full sector, no fill)   8 64-bit exch a (8-byte use of the slot)
//        9 exch a + exch b (result unused)
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <stdint.h>

__device__ __forceinline__ uint32_t mix(uint32_t x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }

template <int LD> __device__ __forceinline__ uint4 ld(const uint4* p) {
    uint4 v;
    if (LD == 0) asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    if (LD == 1) asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    if (LD == 2) asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    if (LD == 3) asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    if (LD == 4) asm volatile("ld.global.cv.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
template <int ST> __device__ __forceinline__ void st(uint4* p, uint4 v) {
    if (ST == 0) asm volatile("st.global.cg.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    if (ST == 1) asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    if (ST == 2) asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    if (ST == 3) asm volatile("st.global.wt.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 exch128(uint4* p, uint4 v) {
    uint4 r;
    asm volatile("{\n .reg .b128 o, n;\n mov.b128 n, {%4,%5,%6,%7};\n atom.global.relaxed.gpu.exch.b128 o, [%8], n;\n mov.b128 {%0,%1,%2,%3}, o;\n}"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "l"(p) : "memory");
    return r;
}

template <int PAT, int LD, int ST>
__global__ void __launch_bounds__(32) walk(uint4* pool, size_t region_entries, int iters, uint32_t fpw, uint32_t jobs,
                                           unsigned long long* sink) {
    const uint32_t lane = threadIdx.x;
    if (lane >= fpw) return;
    const uint32_t job = blockIdx.x * fpw + lane;
    if (job >= jobs) return;
    uint4* t = pool + (size_t)job * region_entries;
    const uint32_t mask = (uint32_t)region_entries - 1;
    uint32_t x = mix(job + 1), a_prev = 0;
    uint4 pend = make_uint4(0, 0, 0, 0), spec = make_uint4(0, 0, 0, 0);
    bool have = false;
    unsigned long long acc = 0;
    for (int i = 0; i < iters; ++i) {
        const uint32_t a = mix(x) & mask, b = mix(x + 0x9e3779b9u) & mask;
        const uint4 mine = make_uint4(x, i, 0, 0);
        uint4 v;
        if (PAT == 0) { v = ld<LD>(t + a); st<ST>(t + a, mine); }
        if (PAT == 1) {
            v = ld<LD>(t + a);                                   // issued before the previous iteration's store
            if (have) { st<ST>(t + a_prev, pend); if (a == a_prev) v = pend; }
            pend = mine; a_prev = a; have = true;
        }
        if (PAT == 2) v = exch128(t + a, mine);
        if (PAT == 3) { v = exch128(t + a, mine); st<ST>(t + b, mine); }
        if (PAT == 4) v = ld<LD>(t + a);
        if (PAT == 5) { v = ld<LD>(t + a); st<ST>(t + b, mine); }
        if (PAT == 6) { acc += spec.y; spec = ld<LD>(t + b); v = exch128(t + a, mine); }
        if (PAT == 7) {
            v = exch128(t + (a & ~1u), mine);
            asm volatile("st.global.v8.b32 [%0], {%1,%2,%1,%2,%1,%2,%1,%2};" ::"l"(t + (b & ~1u)), "r"(x), "r"(i) : "memory");
        }
        if (PAT == 8) {
            unsigned long long o = atomicExch(reinterpret_cast<unsigned long long*>(t + a), ((unsigned long long)i << 32) | x);
            v.x = (uint32_t)o; v.y = (uint32_t)(o >> 32);
        }
        if (PAT == 9) { v = exch128(t + a, mine); uint4 u = exch128(t + b, mine); acc += (u.x == 0xdeadbeefu); }
        x = x * 0x9e3779b1u + v.x + 1;                           // the next address depends on the value read
    }
    if (PAT == 1 && have) st<ST>(t + a_prev, pend);
    sink[job] = acc + x;
}

typedef void (*kern_t)(uint4*, size_t, int, uint32_t, uint32_t, unsigned long long*);
template <int PAT> kern_t pick(int ldf, int stf) {
#define C(L, S) if (ldf == L && stf == S) return walk<PAT, L, S>;
    C(0,0) C(0,1) C(0,2) C(0,3) C(1,0) C(1,1) C(1,2) C(1,3) C(2,0) C(2,1) C(2,2) C(2,3) C(3,0) C(3,1) C(4,0) C(4,1)
#undef C
    return nullptr;
}

int main(int argc, char** argv) {
    const int pat = argc > 1 ? atoi(argv[1]) : 0, ldf = argc > 2 ? atoi(argv[2]) : 0, stf = argc > 3 ? atoi(argv[3]) : 0;
    const uint32_t jobs = argc > 4 ? atoi(argv[4]) : 16384;
    const uint32_t fpw = argc > 5 ? atoi(argv[5]) : 8;
    const int iters = argc > 6 ? atoi(argv[6]) : 2000;
    const size_t region = argc > 7 ? (size_t)atol(argv[7]) : (256u << 10);
    const size_t entries = region / 16;
    kern_t k = pat == 0 ? pick<0>(ldf, stf) : pat == 1 ? pick<1>(ldf, stf) : pat == 2 ? pick<2>(ldf, stf) : pat == 3 ? pick<3>(ldf, stf)
             : pat == 4 ? pick<4>(ldf, stf) : pat == 5 ? pick<5>(ldf, stf) : pat == 6 ? pick<6>(ldf, stf)
             : pat == 7 ? pick<7>(ldf, stf) : pat == 8 ? pick<8>(ldf, stf) : pick<9>(ldf, stf);
    if (!k) { printf("no such combination\n"); return 1; }
    uint4* pool; unsigned long long* sink;
    cudaMalloc(&pool, (size_t)jobs * region);
    cudaMalloc(&sink, jobs * 8);
    cudaMemset(pool, 0, (size_t)jobs * region);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const uint32_t grid = (jobs + fpw - 1) / fpw;
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0);
        k<<<grid, 32>>>(pool, entries, iters, fpw, jobs, sink);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    cudaError_t err = cudaGetLastError();
    printf("pat %d ld %d st %d jobs %u fpw %u region %zu KB: %.3f ms, %.1f ns/iter, %.2f G iter/s  %s\n", pat, ldf, stf, jobs, fpw,
           region >> 10, best, best * 1e6 / iters, (double)jobs * iters / best / 1e6, err == cudaSuccess ? "" : cudaGetErrorString(err));
    return 0;
}
