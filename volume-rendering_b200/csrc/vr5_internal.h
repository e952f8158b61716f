// vr5_internal.h - the renderer handle shared by the .cu translation units of libvr5.so (not part of the ABI).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <string>

struct vr5_renderer {
    int device = 0;
    int num_sms = 0;
    size_t l2_bytes = 0;
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr;  // own_stream, or an external one (vr5_set_stream)
    bool external_stream = false;
    // window
    uint16_t width = 0, height = 0;
    size_t frame_bytes = 0;
    uint8_t *d_frame = nullptr;  // internal framebuffer used when the caller hands a host pointer
    // tables
    float *d_tf = nullptr;        // 128 x float4, pre-multiplied
    uint32_t *d_esl = nullptr;    // 1024 words
    bool have_tf = false;
    bool tf_can_shade = false;    // some TF entry has alpha > 0.05 (CPURenderer.cpp:32)
    // volume
    uint8_t *d_vol_linear = nullptr;  // always kept: source of every layout, of vr5_get_volume and of the min/max builder
    uint64_t vol_bytes = 0;
    uint32_t nx = 0, ny = 0, nz = 0;
    int sampler = 0;
    int last_variant = -1;  // SAMPLER template id of the last cast_kernel launch
    uint8_t *d_vol_brick = nullptr;
    uint8_t *d_vol_padded = nullptr;  // (nx+1)(ny+1)(nz+1), edge replicated: unclamped LDG path
    uint8_t *d_vol_window = nullptr;  // window layout (fetch mode 4), NULL when the volume does not fit one 4-GiB window
    uint32_t win_hi = 0, win_sx = 0, win_sxy = 0;
    float win_mx = 0, win_my = 0, win_mz = 0;
    // power-of-two window (driver VMM mapping inside a 4-GiB-aligned reservation) + its texture; p2_wl == 0: not built
    unsigned long long p2_va = 0, p2_map_addr = 0, p2_handle = 0;
    size_t p2_va_size = 0, p2_map_size = 0;
    int p2_wl = 0;                    // 1 = built
    uint32_t p2_hi = 0, p2_sx = 0, p2_sxy = 0;   // high address half, row / slice strides in bytes
    float p2_mx = 0, p2_my = 0, p2_mz = 0;
    cudaArray_t d_array_p2 = nullptr;
    cudaTextureObject_t tex_p2 = 0;
    uint32_t bxn = 0, byn = 0, bzn = 0;
    cudaArray_t d_array = nullptr;
    cudaTextureObject_t tex = 0;
    // tile queue: two counters used alternately; each frame's kernel re-arms the other one
    unsigned *d_queue = nullptr;
    int parity = 0;
    void *d_counters = nullptr;
    // partition
    int rank = 0, world = 1, packed = 0;
    // peer ring (vr5_peer.cu): owner allocates, writers map it
    uint8_t *peer_base = nullptr;     // header (flags) + slots frames
    bool peer_owner = false, peer_ipc_mapped = false;
    int peer_slots = 0;
    size_t peer_frame_stride = 0;
    uint64_t peer_released = 0;       // owner, host side: frames released so far
    // bookkeeping
    int ctas_per_sm[256] = {0};  // occupancy cache per kernel variant
    uint64_t launches = 0;
    void *d_flush = nullptr;
    size_t flush_bytes = 0;
    std::string err;
};

void vr5_fail(vr5_renderer *r, const char *what, cudaError_t e, const char *file, int line);
int vr5_render_into(vr5_renderer *r, const struct vr5_params *p, void *out_dev);  // async launch of the cast kernel, device buffer
int vr5_adopt_volume(vr5_renderer *r, uint8_t *dev_owned, uint64_t bytes, uint16_t dx, uint16_t dy, uint16_t dz);

#define VR5_CUDA(r, call)                                         \
    do {                                                          \
        cudaError_t vr5_e_ = (call);                              \
        if (vr5_e_ != cudaSuccess) {                              \
            vr5_fail((r), #call, vr5_e_, __FILE__, __LINE__);     \
            return 1;                                             \
        }                                                         \
    } while (0)
