// vr5_hostring.cu - host frame ring: multi-GPU delivery of finished frames to HOST memory (the reference is
// single-GPU, its renderers 1-2 hand the frame to the caller in host memory, VolR.cpp:76-87).
//
// Gathering on one GPU and copying the whole frame from there makes that GPU's PCIe link the limit (C3 at 8 GPUs: 8.3 MB x
// 6 300 frames/s = the link). Here the frame ring lives in host memory shared by all ranks (POSIX shm between processes, or
// plain memory inside one process); every rank renders its strips packed and copies them - one strided cudaMemcpy2DAsync
// on the frame's lane - straight to their rows of the frame's slot, each over its own PCIe link. No device ever waits
// for another: the hand-shake is two host-visible counters per slot,
//     done[rank][slot]  = frames this rank has delivered into the slot  (stored by a one-thread kernel, stream-ordered
//                         behind the copy, st.release.sys into the page-locked header)
//     consumed[slot]    = frames the consumer has released from the slot (plain host store)
// and both waits are bounded host-side polls (vr5_hostring_render: "the slot's previous frame was released";
// vr5_hostring_wait: "all ranks delivered").
#include <algorithm>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include <cuda.h>

#include "../../include/vr5.h"
#include "vr5_internal.h"

namespace {

constexpr uint64_t kHeaderBytes = 8192;
constexpr uint32_t kMagic = 0x56523548u;  // "VR5H"
constexpr int kMaxSlots = 16, kMaxRanks = 64;
constexpr double kTimeoutS = 4.0;

struct Header {
    uint32_t magic, slots, width, height;
    uint32_t pad[12];
    uint32_t consumed[kMaxSlots];              // offset 64
    uint32_t done[kMaxRanks][kMaxSlots];       // offset 128
};
static_assert(sizeof(Header) <= kHeaderBytes, "header");

__global__ void hostring_done_kernel(uint32_t *word, uint32_t value) {
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(word), "r"(value) : "memory");
}

// The "delivered" counter is advanced by a stream memory operation (cuStreamWriteValue32: executed by the front end in
// stream order, flushes the copy before it): unlike a one-thread kernel it needs no SM slot, and the SMs are fully
// occupied by the persistent cast kernels of the other lanes until their tails. The kernel stays as the fallback.
typedef CUresult (*WriteValue32)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
WriteValue32 write_value32() {
    static WriteValue32 fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        cudaDriverEntryPointQueryResult q;
        void *p = nullptr;
        if (!getenv("VR5_HR_FLAG_KERNEL") && cudaGetDriverEntryPoint("cuStreamWriteValue32", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<WriteValue32>(p);
        else
            cudaGetLastError();
    }
    return fn;
}

double now_s() {
    timespec t;
    clock_gettime(CLOCK_MONOTONIC, &t);
    return (double)t.tv_sec + 1e-9 * (double)t.tv_nsec;
}

// bounded poll of a counter another process (or a GPU) advances
bool wait_at_least(const uint32_t *word, uint32_t target) {
    const double t0 = now_s();
    for (unsigned spin = 0;; spin++) {
        if (__atomic_load_n(word, __ATOMIC_ACQUIRE) >= target) return true;
        if ((spin & 1023u) == 1023u && now_s() - t0 > kTimeoutS) return false;
#if defined(__x86_64__)
        __builtin_ia32_pause();
#endif
    }
}

Header *header(vr5_renderer *r) { return reinterpret_cast<Header *>(r->hr_base); }
uint8_t *slot_ptr(vr5_renderer *r, uint64_t seq) { return r->hr_base + kHeaderBytes + (seq % (uint64_t)r->hr_slots) * r->frame_bytes; }

}  // namespace

extern "C" uint64_t vr5_hostring_bytes(uint16_t width, uint16_t height, int slots) {
    if (slots < 1 || slots > kMaxSlots) return 0;
    return kHeaderBytes + (uint64_t)slots * width * height * 4u;
}

extern "C" void vr5_hostring_detach(vr5_renderer *r) {
    if (!r || !r->hr_base) return;
    cudaSetDevice(r->device);
    cudaDeviceSynchronize();
    if (r->hr_registered) cudaHostUnregister(r->hr_base);
    cudaGetLastError();
    r->hr_base = nullptr; r->hr_dev = nullptr; r->hr_slots = 0; r->hr_registered = false;
}

extern "C" int vr5_hostring_attach(vr5_renderer *r, void *base, uint64_t bytes, int slots, int init) {
    if (!r) return 1;
    vr5_hostring_detach(r);
    if (!base || slots < 1 || slots > kMaxSlots) { vr5_fail(r, "vr5_hostring_attach: base and 1..16 slots", cudaSuccess, __FILE__, __LINE__); return 1; }
    if (!r->frame_bytes) { vr5_fail(r, "vr5_hostring_attach: set the window first", cudaSuccess, __FILE__, __LINE__); return 1; }
    if (bytes < vr5_hostring_bytes(r->width, r->height, slots)) { vr5_fail(r, "vr5_hostring_attach: memory smaller than vr5_hostring_bytes", cudaSuccess, __FILE__, __LINE__); return 1; }
    if (r->world > kMaxRanks) { vr5_fail(r, "vr5_hostring_attach: at most 64 ranks", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    // page-lock this process's mapping for the device (memory that is already pinned, e.g. cudaHostAlloc, is used as it is)
    cudaError_t e = cudaHostRegister(base, bytes, cudaHostRegisterPortable | cudaHostRegisterMapped);
    if (e == cudaSuccess) r->hr_registered = true;
    else if (e == cudaErrorHostMemoryAlreadyRegistered) cudaGetLastError();
    else { vr5_fail(r, "cudaHostRegister (host frame ring)", e, __FILE__, __LINE__); cudaGetLastError(); return 1; }
    void *dev = nullptr;
    e = cudaHostGetDevicePointer(&dev, base, 0);
    if (e != cudaSuccess) { if (r->hr_registered) cudaHostUnregister(base); r->hr_registered = false; vr5_fail(r, "cudaHostGetDevicePointer (host frame ring)", e, __FILE__, __LINE__); return 1; }
    r->hr_base = static_cast<uint8_t *>(base);
    r->hr_dev = static_cast<uint8_t *>(dev);
    r->hr_slots = slots;
    Header *h = header(r);
    if (init) {
        memset(h, 0, kHeaderBytes);
        h->slots = (uint32_t)slots; h->width = r->width; h->height = r->height;
        __atomic_store_n(&h->magic, kMagic, __ATOMIC_RELEASE);
    } else if (__atomic_load_n(&h->magic, __ATOMIC_ACQUIRE) != kMagic || h->slots != (uint32_t)slots || h->width != r->width || h->height != r->height) {
        vr5_hostring_detach(r);
        vr5_fail(r, "vr5_hostring_attach: the ring was not initialised for this window / slot count (attach the initialising rank first)", cudaSuccess, __FILE__, __LINE__);
        return 1;
    }
    return 0;
}

extern "C" int vr5_hostring_render(vr5_renderer *r, const vr5_params *p, uint64_t seq) {
    if (!r) return 1;
    if (!r->hr_base) { vr5_fail(r, "vr5_hostring_render: no ring (vr5_hostring_attach)", cudaSuccess, __FILE__, __LINE__); return 1; }
    const bool bands = r->band_count >= 0;   // contiguous band (vr5_set_partition_band / _bands): rendered in place, copied as one block of rows
    if (!r->packed && r->world > 1 && !bands) { vr5_fail(r, "vr5_hostring_render: the partition must be packed strips (vr5_set_partition(..., 1)) or bands", cudaSuccess, __FILE__, __LINE__); return 1; }
    if (r->hr_slots % r->n_lanes) { vr5_fail(r, "vr5_hostring_render: slots must be a multiple of the overlap lanes", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    const uint64_t slots = (uint64_t)r->hr_slots;
    const int slot = (int)(seq % slots), lane = (int)(seq % (uint64_t)r->n_lanes);
    const uint32_t use = (uint32_t)(seq / slots);
    Header *h = header(r);
    if (use > 0 && !wait_at_least(&h->consumed[slot], use)) { vr5_fail(r, "vr5_hostring_render: timed out waiting for the consumer to release the slot", cudaSuccess, __FILE__, __LINE__); return 1; }
    vr5_renderer::Lane &ln = r->lanes[lane];
    if (!ln.d_frame) { vr5_fail(r, "vr5_hostring_render: no lane framebuffer (set the window first)", cudaSuccess, __FILE__, __LINE__); return 1; }
    cudaStream_t st = vr5_lane_stream(r, lane);
    if (vr5_lane_begin(r, lane)) return 1;
    if (vr5_render_into(r, p, ln.d_frame, lane)) return 1;
    // strips of this rank: packed strip j (4 rows) -> rows 4 * (rank + j * world) of the frame
    const uint32_t rows = r->world > 1 ? vr5_strip_rows(r->height, r->rank, r->world) : r->height;
    const size_t row_bytes = (size_t)r->width * 4u, strip_bytes = row_bytes * 4u;
    uint8_t *dst = slot_ptr(r, seq) + (size_t)r->rank * strip_bytes;
    if (bands) {   // vr5_render_into has just cut this frame's band (band_first / band_count)
        const size_t row0 = (size_t)r->band_first * 4u, row1 = std::min<size_t>(r->height, ((size_t)r->band_first + (size_t)r->band_count) * 4u);
        if (row1 > row0) VR5_CUDA(r, cudaMemcpyAsync(slot_ptr(r, seq) + row0 * row_bytes, ln.d_frame + row0 * row_bytes, (row1 - row0) * row_bytes, cudaMemcpyDeviceToHost, st));
    } else if (r->world == 1) {
        VR5_CUDA(r, cudaMemcpyAsync(slot_ptr(r, seq), ln.d_frame, r->frame_bytes, cudaMemcpyDeviceToHost, st));
    } else if (rows) {
        const uint32_t full = rows / 4u, rest = rows % 4u;
        if (full) VR5_CUDA(r, cudaMemcpy2DAsync(dst, strip_bytes * (size_t)r->world, ln.d_frame, strip_bytes, strip_bytes, full, cudaMemcpyDeviceToHost, st));
        if (rest) VR5_CUDA(r, cudaMemcpyAsync(dst + (size_t)full * strip_bytes * (size_t)r->world, ln.d_frame + (size_t)full * strip_bytes, rest * row_bytes, cudaMemcpyDeviceToHost, st));
    }
    uint32_t *done_dev = reinterpret_cast<uint32_t *>(r->hr_dev + offsetof(Header, done)) + (size_t)r->rank * kMaxSlots + slot;
    bool written = false;
    if (WriteValue32 wv = write_value32()) written = wv(st, (CUdeviceptr)(uintptr_t)done_dev, use + 1u, CU_STREAM_WRITE_VALUE_DEFAULT) == CUDA_SUCCESS;
    if (!written) {
        hostring_done_kernel<<<1, 1, 0, st>>>(done_dev, use + 1u);
        r->launches++;
        VR5_CUDA(r, cudaGetLastError());
    }
    return vr5_lane_end(r, lane);
}

extern "C" int vr5_hostring_wait(vr5_renderer *r, uint64_t seq, const void **frame_out) {
    if (!r) return 1;
    if (!r->hr_base) { vr5_fail(r, "vr5_hostring_wait: no ring", cudaSuccess, __FILE__, __LINE__); return 1; }
    const uint64_t slots = (uint64_t)r->hr_slots;
    const int slot = (int)(seq % slots);
    const uint32_t target = (uint32_t)(seq / slots) + 1u;
    Header *h = header(r);
    for (int k = 0; k < r->world; k++)
        if (!wait_at_least(&h->done[k][slot], target)) { vr5_fail(r, "vr5_hostring_wait: timed out waiting for a rank's strips", cudaSuccess, __FILE__, __LINE__); return 1; }
    if (frame_out) *frame_out = slot_ptr(r, seq);
    return 0;
}

extern "C" int vr5_hostring_release(vr5_renderer *r, uint64_t seq) {
    if (!r) return 1;
    if (!r->hr_base) { vr5_fail(r, "vr5_hostring_release: no ring", cudaSuccess, __FILE__, __LINE__); return 1; }
    const uint64_t slots = (uint64_t)r->hr_slots;
    Header *h = header(r);
    const int slot = (int)(seq % slots);
    if (__atomic_load_n(&h->consumed[slot], __ATOMIC_RELAXED) != (uint32_t)(seq / slots)) { vr5_fail(r, "vr5_hostring_release: every frame of a slot is released once, in order", cudaSuccess, __FILE__, __LINE__); return 1; }
    __atomic_store_n(&h->consumed[slot], (uint32_t)(seq / slots) + 1u, __ATOMIC_RELEASE);
    return 0;
}
