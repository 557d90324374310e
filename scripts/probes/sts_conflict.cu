// Probe: which lane wins when several lanes of one warp store to the same shared-memory address
// in one instruction (u16 and u32)?  CUDA leaves it undefined; the fast-mode table warp would like
// "highest lane".  nvcc -arch=sm_100a -o sts_conflict sts_conflict.cu
#include <cstdio>
#include <cstdint>
__global__ void probe(int* out) {
    __shared__ uint16_t t16[64];
    __shared__ uint32_t t32[64];
    int lane = threadIdx.x;
    int res = 0;
    for (int pat = 0; pat < 6; ++pat) {
        // pattern: lanes are grouped; group g = f(lane, pat); all lanes of a group hit one slot
        int g = pat == 0 ? 0 : pat == 1 ? (lane & 1) : pat == 2 ? (lane >> 4) : pat == 3 ? (lane % 3) : pat == 4 ? (lane ^ 5) >> 2 : (lane * 7 % 11);
        for (int i = lane; i < 64; i += 32) { t16[i] = 0xffff; t32[i] = 0xffffffff; }
        __syncwarp();
        t16[g] = (uint16_t)lane;
        t32[g] = (uint32_t)lane;
        __syncwarp();
        int w16 = t16[g], w32 = t32[g];
        // expected winner: highest lane of my group
        unsigned peers = __match_any_sync(0xffffffffu, g);
        int hi = 31 - __clz(peers), lo = __ffs(peers) - 1;
        int ok16 = __all_sync(0xffffffffu, w16 == hi), ok32 = __all_sync(0xffffffffu, w32 == hi);
        int lo16 = __all_sync(0xffffffffu, w16 == lo), lo32 = __all_sync(0xffffffffu, w32 == lo);
        if (lane == 0) out[pat] = ok16 | (ok32 << 1) | (lo16 << 2) | (lo32 << 3);
        (void)res;
    }
}
int main() {
    int* d; int h[6];
    cudaMalloc(&d, sizeof h);
    probe<<<1, 32>>>(d);
    cudaMemcpy(h, d, sizeof h, cudaMemcpyDeviceToHost);
    for (int i = 0; i < 6; ++i)
        printf("pattern %d: u16 highest-wins %d  u32 highest-wins %d  u16 lowest-wins %d  u32 lowest-wins %d\n", i, h[i] & 1, (h[i] >> 1) & 1, (h[i] >> 2) & 1, (h[i] >> 3) & 1);
    return 0;
}
