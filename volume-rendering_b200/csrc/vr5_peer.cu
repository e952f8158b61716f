// vr5_peer.cu - fused render + gather over NVLink peer memory (include/vr5.h, "fused render + gather").
// The reference is single-GPU (VolR.cpp:141-172); this replaces the render -> NCCL gather -> unpack sequence of the
// multi-GPU bench with direct stores of every pixel into the gathering rank's frame, plus two flag words per frame.
#include <cuda.h>
#include <cuda_runtime.h>

#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

#include "../../include/vr5.h"
#include "vr5_internal.h"

namespace {

constexpr size_t kHeaderBytes = 16384;          // flag words, one 128-byte line each
constexpr int kMaxSlots = 32;
constexpr unsigned long long kWaitTimeoutNs = 4000000000ull;  // a wait that lasts 4 s gives up and raises the status flag

// header layout (uint32 words, one 128-byte line per flag): [32] timed_out, [64 + 64 * slot] arrive[slot] = ranks that
// finished their strips of a frame in that slot (all uses; EVERY rank arrives, the owner included: a use adds `world`),
// [96 + 64 * slot] consumed[slot] = uses of that slot the owner has released.
// Per-slot counters keep the protocol independent of the order in which frames on different lanes complete.
__host__ __device__ inline uint32_t *timeout_word(uint8_t *base) { return reinterpret_cast<uint32_t *>(base) + 32; }
__host__ __device__ inline uint32_t *arrive_word(uint8_t *base, int slot) { return reinterpret_cast<uint32_t *>(base) + 64 + 64 * slot; }
__host__ __device__ inline uint32_t *consumed_word(uint8_t *base, int slot) { return reinterpret_cast<uint32_t *>(base) + 96 + 64 * slot; }

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

// Stream memory operations (executed by the stream's front end, no SM slot needed - the SMs are full of persistent cast CTAs):
// the consumer side of a ring driven by frame sequences waits for arrive[slot] and stores consumed[slot] with them. Resolved
// at run time; one-thread kernels are the fallback.
typedef CUresult (*WaitValue32)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
typedef CUresult (*WriteValue32)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
struct MemOps { WaitValue32 wait = nullptr; WriteValue32 write = nullptr; };
const MemOps &mem_ops() {
    static MemOps ops;
    static bool tried = false;
    if (!tried) {
        tried = true;
        cudaDriverEntryPointQueryResult q;
        void *w = nullptr, *x = nullptr;
        if (!getenv("VR5_PEER_FLAG_KERNELS") &&
            cudaGetDriverEntryPoint("cuStreamWaitValue32", &w, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess &&
            cudaGetDriverEntryPoint("cuStreamWriteValue32", &x, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess) {
            ops.wait = reinterpret_cast<WaitValue32>(w);
            ops.write = reinterpret_cast<WriteValue32>(x);
        } else {
            cudaGetLastError();
        }
    }
    return ops;
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
    vr5_sync(r);
    if (r->peer_owner) cudaFree(r->peer_base);
    else if (r->peer_ipc_mapped) cudaIpcCloseMemHandle(r->peer_base);
    r->peer_base = nullptr; r->peer_owner = false; r->peer_ipc_mapped = false; r->peer_slots = 0; memset(r->peer_released_slot, 0, sizeof(r->peer_released_slot));
}

extern "C" int vr5_peer_create(vr5_renderer *r, int slots) {
    if (!r) return 1;
    if (slots < 1 || slots > kMaxSlots || !r->frame_bytes) { vr5_fail(r, "vr5_peer_create: need 1..32 slots and vr5_set_window first", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    vr5_peer_close(r);
    r->peer_frame_stride = frame_stride(r);
    const size_t bytes = kHeaderBytes + (size_t)slots * r->peer_frame_stride;
    cudaError_t e = cudaMalloc(&r->peer_base, bytes);
    if (e != cudaSuccess) { r->peer_base = nullptr; vr5_fail(r, "cudaMalloc peer ring", e, __FILE__, __LINE__); cudaGetLastError(); return 1; }
    VR5_CUDA(r, cudaMemsetAsync(r->peer_base, 0, bytes, r->stream));
    VR5_CUDA(r, cudaStreamSynchronize(r->stream));
    r->peer_owner = true; r->peer_slots = slots; memset(r->peer_released_slot, 0, sizeof(r->peer_released_slot));
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

// frames of the ring use lane seq % n_lanes; with slots a multiple of the lane count a slot is always reused on the lane
// that released it, so the owner's own stores need no extra ordering
static int lane_of(const vr5_renderer *r, uint64_t seq) { return (int)(seq % (uint64_t)r->n_lanes); }

extern "C" int vr5_peer_render(vr5_renderer *r, const vr5_params *p, uint64_t seq) {
    if (!r) return 1;
    if (!r->peer_base) { vr5_fail(r, "vr5_peer_render: no ring (vr5_peer_create / vr5_peer_open)", cudaSuccess, __FILE__, __LINE__); return 1; }
    if (r->packed) { vr5_fail(r, "vr5_peer_render: the partition must be unpacked (pixels go to their final place)", cudaSuccess, __FILE__, __LINE__); return 1; }
    if (r->peer_slots % r->n_lanes) { vr5_fail(r, "vr5_peer_render: slots must be a multiple of the overlap lanes", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    const uint64_t slots = (uint64_t)r->peer_slots;
    const int slot = (int)(seq % slots), lane = lane_of(r, seq);
    const uint32_t use = (uint32_t)(seq / slots);  // how many frames used this slot before
    cudaStream_t st = vr5_lane_stream(r, lane);
    if (vr5_lane_begin(r, lane)) return 1;
    if (r->peer_owner) {
        // stream order does the rest: vr5_peer_release of the slot's previous frame was enqueued on this lane
        if (r->peer_released_slot[slot] < use) { vr5_fail(r, "vr5_peer_render: slot not released yet (vr5_peer_release every frame)", cudaSuccess, __FILE__, __LINE__); return 1; }
    } else if (use > 0) {
        // wait until the owner has released the frame that last used this slot
        peer_wait_kernel<<<1, 1, 0, st>>>(consumed_word(r->peer_base, slot), use, timeout_word(r->peer_base));
        r->launches++;
    }
    if (vr5_render_into(r, p, slot_ptr(r, seq), lane)) return 1;
    // every rank arrives, the owner included (its own wait is then one counter compare whichever API filled the slot)
    peer_arrive_kernel<<<1, 1, 0, st>>>(arrive_word(r->peer_base, slot));
    r->launches++;
    VR5_CUDA(r, cudaGetLastError());
    return vr5_lane_end(r, lane);
}

extern "C" int vr5_peer_wait(vr5_renderer *r, uint64_t seq, void **frame_dev_out) {
    if (!r) return 1;
    if (!r->peer_owner) { vr5_fail(r, "vr5_peer_wait: only the owner waits for frames", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    const uint64_t slots = (uint64_t)r->peer_slots;
    {
        // every use of a slot adds `world` arrivals to its counter
        const uint32_t target = (uint32_t)((seq / slots + 1) * (uint64_t)r->world);
        peer_wait_kernel<<<1, 1, 0, vr5_lane_stream(r, lane_of(r, seq))>>>(arrive_word(r->peer_base, (int)(seq % slots)), target, timeout_word(r->peer_base));
        r->launches++;
        VR5_CUDA(r, cudaGetLastError());
    }
    if (frame_dev_out) *frame_dev_out = slot_ptr(r, seq);
    return vr5_lane_end(r, lane_of(r, seq));
}

extern "C" int vr5_peer_read(vr5_renderer *r, uint64_t seq, void *host_dst) {
    if (!r) return 1;
    if (!r->peer_owner || !host_dst) { vr5_fail(r, "vr5_peer_read: owner only, host buffer needed", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    VR5_CUDA(r, cudaMemcpyAsync(host_dst, slot_ptr(r, seq), r->frame_bytes, cudaMemcpyDeviceToHost, vr5_lane_stream(r, lane_of(r, seq))));
    return vr5_lane_end(r, lane_of(r, seq));
}

extern "C" int vr5_peer_release(vr5_renderer *r, uint64_t seq) {
    if (!r) return 1;
    if (!r->peer_owner) { vr5_fail(r, "vr5_peer_release: only the owner releases frames", cudaSuccess, __FILE__, __LINE__); return 1; }
    const uint64_t slots = (uint64_t)r->peer_slots;
    const int slot = (int)(seq % slots);
    if ((uint32_t)(seq / slots) != r->peer_released_slot[slot]) { vr5_fail(r, "vr5_peer_release: every frame of a slot is released once, in order", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    peer_release_kernel<<<1, 1, 0, vr5_lane_stream(r, lane_of(r, seq))>>>(consumed_word(r->peer_base, slot), (uint32_t)(seq / slots + 1));
    r->launches++;
    VR5_CUDA(r, cudaGetLastError());
    r->peer_released_slot[slot]++;
    return vr5_lane_end(r, lane_of(r, seq));
}

extern "C" int vr5_peer_sync_frame(vr5_renderer *r, uint64_t seq) {
    if (!r) return 1;
    VR5_CUDA(r, cudaSetDevice(r->device));
    VR5_CUDA(r, cudaStreamSynchronize(vr5_lane_stream(r, lane_of(r, seq))));
    return 0;
}

extern "C" int vr5_peer_status(vr5_renderer *r) {
    if (!r) return 1;
    if (!r->peer_base) { vr5_fail(r, "vr5_peer_status: no ring", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    uint32_t flag = 0;
    if (vr5_sync(r)) return 1;
    VR5_CUDA(r, cudaMemcpyAsync(&flag, timeout_word(r->peer_base), sizeof(flag), cudaMemcpyDeviceToHost, r->stream));
    VR5_CUDA(r, cudaStreamSynchronize(r->stream));
    if (flag) { vr5_fail(r, "peer ring: a device-side wait timed out (a rank is missing or out of frame order)", cudaSuccess, __FILE__, __LINE__); return 1; }
    return 0;
}

// ---- frame sequences through the ring ---------------------------------------------------------------------------------------
// ONE persistent launch per rank renders frames seq0 .. seq0 + n - 1 (vr5_sequence_launch): the hand-shake moves into the
// kernel. Before a warp stores pixels of a frame it makes sure the frame that used the slot before has been released
// (consumed[slot], acquire load, cached per CTA); the warp that finishes a rank's last tile of a frame adds the rank's arrival
// (system fence + system-scope atomic on arrive[slot]). No wait / arrive kernels, no per-frame launches, one call per camera path.
extern "C" int vr5_peer_render_sequence(vr5_renderer *r, const vr5_params *frames, int n, uint64_t seq0) {
    if (!r) return 1;
    if (!r->peer_base) { vr5_fail(r, "vr5_peer_render_sequence: no ring (vr5_peer_create / vr5_peer_open)", cudaSuccess, __FILE__, __LINE__); return 1; }
    if (r->packed) { vr5_fail(r, "vr5_peer_render_sequence: the partition must be unpacked (pixels go to their final place)", cudaSuccess, __FILE__, __LINE__); return 1; }
    if (!frames || n < 1) { vr5_fail(r, "vr5_peer_render_sequence: bad arguments", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    const uint64_t slots = (uint64_t)r->peer_slots;
    std::vector<vr5_frame_target> t((size_t)n);
    for (int i = 0; i < n; i++) {
        const uint64_t seq = seq0 + (uint64_t)i;
        const int slot = (int)(seq % slots);
        const uint32_t use = (uint32_t)(seq / slots);
        t[(size_t)i] = vr5_frame_target{slot_ptr(r, seq), arrive_word(r->peer_base, slot), (use && !(r->opt.seq_debug & 1)) ? consumed_word(r->peer_base, slot) : nullptr,
                                        (r->opt.seq_debug & 4) ? 0u : use};
    }
    return vr5_sequence_launch(r, frames, n, t.data(), timeout_word(r->peer_base), r->peer_owner ? 0 : 1);
}

// owner: consume frames seq0 .. seq0 + n - 1 in order, on the renderer's consumer stream (asynchronous; vr5_join / vr5_sync
// wait for it): wait for all `world` arrivals of a frame, optionally copy it (device to device) to copy_dst_dev + i *
// width * height * 4 - what a consumer does with a frame - and release its slot.
extern "C" int vr5_peer_consume_sequence(vr5_renderer *r, uint64_t seq0, int n, void *copy_dst_dev) {
    if (!r) return 1;
    if (!r->peer_owner) { vr5_fail(r, "vr5_peer_consume_sequence: only the owner consumes frames", cudaSuccess, __FILE__, __LINE__); return 1; }
    if (n < 1) { vr5_fail(r, "vr5_peer_consume_sequence: bad arguments", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    const uint64_t slots = (uint64_t)r->peer_slots;
    const MemOps &ops = mem_ops();
    cudaStream_t st = r->consumer_stream;
    for (int i = 0; i < n; i++) {
        const uint64_t seq = seq0 + (uint64_t)i;
        const int slot = (int)(seq % slots);
        const uint32_t use = (uint32_t)(seq / slots);
        if (use != r->peer_released_slot[slot]) { vr5_fail(r, "vr5_peer_consume_sequence: every frame of a slot is released once, in order", cudaSuccess, __FILE__, __LINE__); return 1; }
        const uint32_t target = (use + 1u) * (uint32_t)r->world;
        uint32_t *arrive = arrive_word(r->peer_base, slot), *consumed = consumed_word(r->peer_base, slot);
        if (!ops.wait || ops.wait((CUstream)st, (CUdeviceptr)(uintptr_t)arrive, target, CU_STREAM_WAIT_VALUE_GEQ) != CUDA_SUCCESS) {
            peer_wait_kernel<<<1, 1, 0, st>>>(arrive, target, timeout_word(r->peer_base));
            r->launches++;
        }
        if (copy_dst_dev)
            VR5_CUDA(r, cudaMemcpyAsync(static_cast<uint8_t *>(copy_dst_dev) + (size_t)i * r->frame_bytes, slot_ptr(r, seq), r->frame_bytes, cudaMemcpyDeviceToDevice, st));
        if (!ops.write || ops.write((CUstream)st, (CUdeviceptr)(uintptr_t)consumed, use + 1u, CU_STREAM_WRITE_VALUE_DEFAULT) != CUDA_SUCCESS) {
            peer_release_kernel<<<1, 1, 0, st>>>(consumed, use + 1u);
            r->launches++;
        }
        r->peer_released_slot[slot]++;
    }
    VR5_CUDA(r, cudaGetLastError());
    VR5_CUDA(r, cudaEventRecord(r->consumer_done, st));
    r->consumer_pending = true;
    return 0;
}
