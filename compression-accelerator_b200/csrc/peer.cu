// peer.cu -- C-ABI of the multi-GPU exchange over NVLink peer memory (include/snappy_b200.h,
// snappy_b200_peer_*): mailboxes and the root's archive are cudaMalloc'ed here and shared between the
// per-GPU processes through CUDA IPC handles the host language passes around; the kernels are in
// peer.cuh.
#include "kernels.h"
#include "../../include/snappy_b200.h"
#include <cuda_runtime.h>
#include <cstdio>
#include <cstring>
#include <new>

struct snappy_b200_peer {
    snappy_b200_ctx* ctx = nullptr;
    int device = 0;
    uint32_t rank = 0, world = 1, root = 0;
    size_t slot_bytes = 0;
    sb::PeerMailbox* box = nullptr;                          // mine (cudaMalloc)
    uint8_t* archive = nullptr;                              // root: mine; others: the root's, mapped
    bool archive_mapped = false;
    sb::PeerTable table{};
    bool mapped[sb::kPeerMaxWorld] = {};
    unsigned long long* d_offset = nullptr;                  // two words (slot = step & 1)
    cudaStream_t xs = nullptr;                               // exchange stream (high priority)
    cudaEvent_t ev_in = nullptr, ev_copied[2] = {nullptr, nullptr}, ev_wait = nullptr;
    unsigned long long* h_box = nullptr;                     // pinned copy of my mailbox for results
    unsigned long long launches = 0;
};

namespace {
thread_local char g_peer_err[256] = "";
snappy_b200_status pfail(snappy_b200_status st, const char* what, cudaError_t e = cudaSuccess) {
    if (e != cudaSuccess) snprintf(g_peer_err, sizeof g_peer_err, "%s: %s", what, cudaGetErrorString(e));
    else snprintf(g_peer_err, sizeof g_peer_err, "%s", what);
    return st;
}
#define PCU(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return pfail(SNAPPY_B200_CUDA_ERROR, #x, e_); } while (0)
constexpr size_t kHandleBytes = 2 * sizeof(cudaIpcMemHandle_t);   // mailbox, archive
}  // namespace

extern "C" {

const char* snappy_b200_peer_last_error(void) { return g_peer_err; }

size_t snappy_b200_peer_handle_bytes(void) { return kHandleBytes; }

snappy_b200_status snappy_b200_peer_create(snappy_b200_ctx* ctx, int device, int rank, int world, int root,
                                           size_t archive_slot_bytes, void* handles_out,
                                           snappy_b200_peer** out) {
    if (!ctx || !out || !handles_out || world < 1 || world > (int)sb::kPeerMaxWorld || rank < 0 || rank >= world ||
        root < 0 || root >= world)
        return pfail(SNAPPY_B200_INVALID_INPUT, "peer_create: bad argument");
    *out = nullptr;
    PCU(cudaSetDevice(device));
    snappy_b200_peer* p = new (std::nothrow) snappy_b200_peer;
    if (!p) return pfail(SNAPPY_B200_CUDA_ERROR, "out of host memory");
    p->ctx = ctx; p->device = device; p->rank = rank; p->world = world; p->root = root;
    p->slot_bytes = archive_slot_bytes;
    PCU(cudaMalloc(&p->box, sizeof(sb::PeerMailbox)));
    PCU(cudaMemset(p->box, 0, sizeof(sb::PeerMailbox)));
    PCU(cudaMalloc(&p->d_offset, 2 * sizeof(unsigned long long)));
    PCU(cudaMallocHost(&p->h_box, sizeof(sb::PeerMailbox)));
    int least = 0, greatest = 0;
    PCU(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    PCU(cudaStreamCreateWithPriority(&p->xs, cudaStreamNonBlocking, greatest));
    PCU(cudaEventCreateWithFlags(&p->ev_in, cudaEventDisableTiming));
    PCU(cudaEventCreateWithFlags(&p->ev_copied[0], cudaEventDisableTiming));
    PCU(cudaEventCreateWithFlags(&p->ev_copied[1], cudaEventDisableTiming));
    PCU(cudaEventCreateWithFlags(&p->ev_wait, cudaEventDisableTiming));
    cudaIpcMemHandle_t h[2];
    memset(h, 0, sizeof h);
    PCU(cudaIpcGetMemHandle(&h[0], p->box));
    if (rank == root) {
        PCU(cudaMalloc(&p->archive, 2 * archive_slot_bytes + 64));
        PCU(cudaIpcGetMemHandle(&h[1], p->archive));
    }
    PCU(cudaDeviceSynchronize());
    memcpy(handles_out, h, sizeof h);
    p->table.box[rank] = p->box;
    *out = p;
    return SNAPPY_B200_OK;
}

snappy_b200_status snappy_b200_peer_connect(snappy_b200_peer* p, const void* all_handles) {
    if (!p || !all_handles) return pfail(SNAPPY_B200_INVALID_INPUT, "peer_connect: null argument");
    PCU(cudaSetDevice(p->device));
    const uint8_t* base = (const uint8_t*)all_handles;
    for (uint32_t r = 0; r < p->world; ++r) {
        if (r == p->rank) continue;
        cudaIpcMemHandle_t h[2];
        memcpy(h, base + (size_t)r * kHandleBytes, sizeof h);
        void* q = nullptr;
        PCU(cudaIpcOpenMemHandle(&q, h[0], cudaIpcMemLazyEnablePeerAccess));
        p->table.box[r] = (sb::PeerMailbox*)q;
        p->mapped[r] = true;
        if (r == p->root) {
            void* a = nullptr;
            PCU(cudaIpcOpenMemHandle(&a, h[1], cudaIpcMemLazyEnablePeerAccess));
            p->archive = (uint8_t*)a;
            p->archive_mapped = true;
        }
    }
    return SNAPPY_B200_OK;
}

snappy_b200_status snappy_b200_peer_gather(snappy_b200_peer* p, const void* d_payload,
                                           const unsigned long long* d_size, unsigned long long step,
                                           unsigned long long base) {
    if (!p || !d_payload || !d_size) return pfail(SNAPPY_B200_INVALID_INPUT, "peer_gather: null argument");
    PCU(cudaSetDevice(p->device));
    cudaStream_t cs = (cudaStream_t)snappy_b200_stream(p->ctx);
    PCU(cudaEventRecord(p->ev_in, cs));                      // behind the compress kernels queued so far
    PCU(cudaStreamWaitEvent(p->xs, p->ev_in, 0));
    PCU(sb::launch_peer_step(p->table, p->rank, p->world, p->root, step, base, (const uint8_t*)d_payload, d_size,
                             p->archive + (step & 1) * p->slot_bytes, p->slot_bytes, p->d_offset + (step & 1),
                             p->xs, p->ev_copied[step & 1], &p->launches));
    return SNAPPY_B200_OK;
}

snappy_b200_status snappy_b200_peer_reuse(snappy_b200_peer* p, unsigned long long step) {
    if (!p) return pfail(SNAPPY_B200_INVALID_INPUT, "peer_reuse: null argument");
    PCU(cudaSetDevice(p->device));
    PCU(cudaStreamWaitEvent((cudaStream_t)snappy_b200_stream(p->ctx), p->ev_copied[step & 1], 0));
    return SNAPPY_B200_OK;
}

snappy_b200_status snappy_b200_peer_wait(snappy_b200_peer* p, unsigned long long step) {
    if (!p) return pfail(SNAPPY_B200_INVALID_INPUT, "peer_wait: null argument");
    PCU(cudaSetDevice(p->device));
    PCU(sb::launch_peer_wait(p->box, p->world, step, p->xs, &p->launches));
    PCU(cudaMemcpyAsync(p->h_box, p->box, sizeof(sb::PeerMailbox), cudaMemcpyDeviceToHost, p->xs));
    PCU(cudaEventRecord(p->ev_wait, p->xs));
    PCU(cudaStreamWaitEvent((cudaStream_t)snappy_b200_stream(p->ctx), p->ev_wait, 0));
    return SNAPPY_B200_OK;
}

snappy_b200_status snappy_b200_peer_result(snappy_b200_peer* p, unsigned long long step,
                                           unsigned long long base, unsigned long long* total,
                                           unsigned long long* offsets) {
    if (!p || !total) return pfail(SNAPPY_B200_INVALID_INPUT, "peer_result: null argument");
    PCU(cudaSetDevice(p->device));
    PCU(cudaEventSynchronize(p->ev_wait));
    const sb::PeerMailbox* m = (const sb::PeerMailbox*)p->h_box;
    if (m->error) return pfail(SNAPPY_B200_CUDA_ERROR, m->error == 2 ? "peer exchange: archive slot too small"
                                                                     : "peer exchange: a rank did not arrive in time");
    unsigned long long run = base;
    for (uint32_t r = 0; r < p->world; ++r) {
        const unsigned long long v = m->sizes[step & 1][r];
        if ((v >> sb::kPeerSizeBits) != step + 1) return pfail(SNAPPY_B200_CUDA_ERROR, "peer exchange: stale size word");
        if (offsets) offsets[r] = run;
        run += v & ((1ull << sb::kPeerSizeBits) - 1);
    }
    *total = run;
    return SNAPPY_B200_OK;
}

void* snappy_b200_peer_archive(snappy_b200_peer* p, int slot) {
    return p && p->archive ? (void*)(p->archive + (size_t)(slot & 1) * p->slot_bytes) : nullptr;
}

void* snappy_b200_peer_stream(snappy_b200_peer* p) { return p ? (void*)p->xs : nullptr; }

unsigned long long snappy_b200_peer_launch_count(snappy_b200_peer* p) { return p ? p->launches : 0; }

void snappy_b200_peer_destroy(snappy_b200_peer* p) {
    if (!p) return;
    cudaSetDevice(p->device);
    if (p->xs) cudaStreamSynchronize(p->xs);
    for (uint32_t r = 0; r < p->world; ++r)
        if (p->mapped[r]) cudaIpcCloseMemHandle(p->table.box[r]);
    if (p->archive) { if (p->archive_mapped) cudaIpcCloseMemHandle(p->archive); else cudaFree(p->archive); }
    if (p->box) cudaFree(p->box);
    if (p->d_offset) cudaFree(p->d_offset);
    if (p->h_box) cudaFreeHost(p->h_box);
    if (p->ev_in) cudaEventDestroy(p->ev_in);
    if (p->ev_copied[0]) cudaEventDestroy(p->ev_copied[0]);
    if (p->ev_copied[1]) cudaEventDestroy(p->ev_copied[1]);
    if (p->ev_wait) cudaEventDestroy(p->ev_wait);
    if (p->xs) cudaStreamDestroy(p->xs);
    delete p;
}

}  // extern "C"
