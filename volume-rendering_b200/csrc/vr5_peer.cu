// vr5_peer.cu - fused render + gather over NVLink peer memory (include/vr5.h, "fused render + gather").
// The reference is single-GPU (VolR.cpp:141-172); this replaces the render -> NCCL gather -> unpack sequence of the
// multi-GPU bench with direct stores of every pixel into the gathering rank's frame, plus two flag words per frame.
#include <cuda_runtime.h>

#include <stdint.h>
#include <string.h>

#include "../../include/vr5.h"
#include "vr5_internal.h"

namespace {

constexpr size_t kHeaderBytes = 4096;          // flag words, one 128-byte line each
constexpr int kMaxSlots = 16;
constexpr unsigned long long kWaitTimeoutNs = 4000000000ull;  // a wait that lasts 4 s gives up and raises the status flag

// header layout (uint32 words): [0] consumed, [32] timed_out, [64 + 32 * slot] arrive[slot]
__host__ __device__ inline uint32_t *consumed_word(uint8_t *base) { return reinterpret_cast<uint32_t *>(base); }
__host__ __device__ inline uint32_t *timeout_word(uint8_t *base) { return reinterpret_cast<uint32_t *>(base) + 32; }
__host__ __device__ inline uint32_t *arrive_word(uint8_t *base, int slot) { return reinterpret_cast<uint32_t *>(base) + 64 + 32 * slot; }

__device__ __forceinline__ uint32_t load_acquire_sys(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// one thread: spin until *flag >= target (wrap-safe), bounded in time
__global__ void peer_wait_kernel(const uint32_t *flag, uint32_t target, uint32_t *timed_out) {
    const unsigned long long t0 = global_ns();
    while ((int32_t)(load_acquire_sys(flag) - target) < 0) {
        if (global_ns() - t0 > kWaitTimeoutNs) { *timed_out = 1u; break; }
        __nanosleep(100);
    }
}

// one thread, after the cast kernel on the same stream: publish this rank's pixels, then count the arrival
__global__ void peer_arrive_kernel(uint32_t *arrive) {
    __threadfence_system();
    atomicAdd_system(arrive, 1u);
}

__global__ void peer_release_kernel(uint32_t *consumed, uint32_t value) {
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(consumed), "r"(value) : "memory");
}

size_t frame_stride(const vr5_renderer *r) { return (r->frame_bytes + 4095) & ~(size_t)4095; }

uint8_t *slot_ptr(const vr5_renderer *r, uint64_t seq) { return r->peer_base + kHeaderBytes + (size_t)(seq % (uint64_t)r->peer_slots) * r->peer_frame_stride; }

}  // namespace

extern "C" void vr5_peer_close(vr5_renderer *r) {
    if (!r || !r->peer_base) return;
    cudaSetDevice(r->device);
    cudaStreamSynchronize(r->stream);
    if (r->peer_owner) cudaFree(r->peer_base);
    else if (r->peer_ipc_mapped) cudaIpcCloseMemHandle(r->peer_base);
    r->peer_base = nullptr; r->peer_owner = false; r->peer_ipc_mapped = false; r->peer_slots = 0; r->peer_released = 0;
}

extern "C" int vr5_peer_create(vr5_renderer *r, int slots) {
    if (!r) return 1;
    if (slots < 1 || slots > kMaxSlots || !r->frame_bytes) { vr5_fail(r, "vr5_peer_create: need 1..16 slots and vr5_set_window first", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    vr5_peer_close(r);
    r->peer_frame_stride = frame_stride(r);
    const size_t bytes = kHeaderBytes + (size_t)slots * r->peer_frame_stride;
    cudaError_t e = cudaMalloc(&r->peer_base, bytes);
    if (e != cudaSuccess) { r->peer_base = nullptr; vr5_fail(r, "cudaMalloc peer ring", e, __FILE__, __LINE__); cudaGetLastError(); return 1; }
    VR5_CUDA(r, cudaMemsetAsync(r->peer_base, 0, bytes, r->stream));
    VR5_CUDA(r, cudaStreamSynchronize(r->stream));
    r->peer_owner = true; r->peer_slots = slots; r->peer_released = 0;
    return 0;
}

extern "C" int vr5_peer_export(vr5_renderer *r, void *handle_out64) {
    if (!r) return 1;
    if (!r->peer_owner || !handle_out64) { vr5_fail(r, "vr5_peer_export: not the owner of a ring", cudaSuccess, __FILE__, __LINE__); return 1; }
    static_assert(sizeof(cudaIpcMemHandle_t) == VR5_PEER_HANDLE_BYTES, "IPC handle size");
    VR5_CUDA(r, cudaSetDevice(r->device));
    cudaIpcMemHandle_t h;
    VR5_CUDA(r, cudaIpcGetMemHandle(&h, r->peer_base));
    memcpy(handle_out64, &h, sizeof(h));
    return 0;
}

extern "C" int vr5_peer_open(vr5_renderer *r, const void *handle64, int slots) {
    if (!r) return 1;
    if (!handle64 || slots < 1 || slots > kMaxSlots || !r->frame_bytes) { vr5_fail(r, "vr5_peer_open: bad arguments (vr5_set_window first)", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    vr5_peer_close(r);
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, sizeof(h));
    void *p = nullptr;
    VR5_CUDA(r, cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    r->peer_base = static_cast<uint8_t *>(p);
    r->peer_ipc_mapped = true; r->peer_owner = false; r->peer_slots = slots; r->peer_frame_stride = frame_stride(r);
    return 0;
}

extern "C" int vr5_peer_open_local(vr5_renderer *r, vr5_renderer *owner) {
    if (!r) return 1;
    if (!owner || !owner->peer_owner || owner->frame_bytes != r->frame_bytes) { vr5_fail(r, "vr5_peer_open_local: owner has no ring for this window", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    vr5_peer_close(r);
    if (owner->device != r->device) {
        cudaError_t e = cudaDeviceEnablePeerAccess(owner->device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { vr5_fail(r, "cudaDeviceEnablePeerAccess", e, __FILE__, __LINE__); return 1; }
        cudaGetLastError();
    }
    r->peer_base = owner->peer_base;
    r->peer_ipc_mapped = false; r->peer_owner = false; r->peer_slots = owner->peer_slots; r->peer_frame_stride = owner->peer_frame_stride;
    return 0;
}

extern "C" int vr5_peer_render(vr5_renderer *r, const vr5_params *p, uint64_t seq) {
    if (!r) return 1;
    if (!r->peer_base) { vr5_fail(r, "vr5_peer_render: no ring (vr5_peer_create / vr5_peer_open)", cudaSuccess, __FILE__, __LINE__); return 1; }
    if (r->packed) { vr5_fail(r, "vr5_peer_render: the partition must be unpacked (pixels go to their final place)", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    const uint64_t slots = (uint64_t)r->peer_slots;
    if (r->peer_owner) {
        // stream order does the rest: vr5_peer_release(seq - slots) was enqueued before this frame's kernel
        if (seq >= slots && r->peer_released < seq - slots + 1) { vr5_fail(r, "vr5_peer_render: slot not released yet (call vr5_peer_release in frame order)", cudaSuccess, __FILE__, __LINE__); return 1; }
    } else if (seq >= slots) {
        // wait until the owner has released the frame that last used this slot
        peer_wait_kernel<<<1, 1, 0, r->stream>>>(consumed_word(r->peer_base), (uint32_t)(seq - slots + 1), timeout_word(r->peer_base));
        r->launches++;
    }
    if (vr5_render_into(r, p, slot_ptr(r, seq))) return 1;
    if (!r->peer_owner) {
        peer_arrive_kernel<<<1, 1, 0, r->stream>>>(arrive_word(r->peer_base, (int)(seq % slots)));
        r->launches++;
    }
    VR5_CUDA(r, cudaGetLastError());
    return 0;
}

extern "C" int vr5_peer_wait(vr5_renderer *r, uint64_t seq, void **frame_dev_out) {
    if (!r) return 1;
    if (!r->peer_owner) { vr5_fail(r, "vr5_peer_wait: only the owner waits for frames", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    const uint64_t slots = (uint64_t)r->peer_slots;
    if (r->world > 1) {
        // every use of a slot adds world - 1 arrivals to its counter
        const uint32_t target = (uint32_t)((seq / slots + 1) * (uint64_t)(r->world - 1));
        peer_wait_kernel<<<1, 1, 0, r->stream>>>(arrive_word(r->peer_base, (int)(seq % slots)), target, timeout_word(r->peer_base));
        r->launches++;
        VR5_CUDA(r, cudaGetLastError());
    }
    if (frame_dev_out) *frame_dev_out = slot_ptr(r, seq);
    return 0;
}

extern "C" int vr5_peer_release(vr5_renderer *r, uint64_t seq) {
    if (!r) return 1;
    if (!r->peer_owner) { vr5_fail(r, "vr5_peer_release: only the owner releases frames", cudaSuccess, __FILE__, __LINE__); return 1; }
    if (seq != r->peer_released) { vr5_fail(r, "vr5_peer_release: frames are released in order", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    peer_release_kernel<<<1, 1, 0, r->stream>>>(consumed_word(r->peer_base), (uint32_t)(seq + 1));
    r->launches++;
    VR5_CUDA(r, cudaGetLastError());
    r->peer_released = seq + 1;
    return 0;
}

extern "C" int vr5_peer_status(vr5_renderer *r) {
    if (!r) return 1;
    if (!r->peer_base) { vr5_fail(r, "vr5_peer_status: no ring", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    uint32_t flag = 0;
    VR5_CUDA(r, cudaMemcpyAsync(&flag, timeout_word(r->peer_base), sizeof(flag), cudaMemcpyDeviceToHost, r->stream));
    VR5_CUDA(r, cudaStreamSynchronize(r->stream));
    if (flag) { vr5_fail(r, "peer ring: a device-side wait timed out (a rank is missing or out of frame order)", cudaSuccess, __FILE__, __LINE__); return 1; }
    return 0;
}
