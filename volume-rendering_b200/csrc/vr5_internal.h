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
    // tables
    float *d_tf = nullptr;        // 128 x float4, pre-multiplied
    uint32_t *d_esl = nullptr;    // 1024 words
    bool have_tf = false;
    bool tf_opaque = false;       // some TF entry has alpha >= 0.999 (a single sample can terminate a ray)
    uint8_t *d_brick_max = nullptr;   // maximum voxel value per brick (build_brick_max), or NULL
    uint32_t brick_lb = 0, brick_nx = 0, brick_ny = 0, brick_nz = 0;
    int tf_zero_prefix = 0;       // leading TF entries that are all-zero, when every alpha lies in [0, 1] and all values are finite (else 0)
    int last_zskip = 0;           // the last cast_kernel launch was a zero-skip (ZS) build
    bool tf_can_shade = false;    // some TF entry has alpha > 0.05 (CPURenderer.cpp:32)
    // volume
    uint8_t *d_vol_linear = nullptr;  // always kept: source of every layout, of vr5_get_volume and of the min/max builder
    uint64_t vol_bytes = 0;
    uint32_t nx = 0, ny = 0, nz = 0;
    int sampler = 0;
    int last_variant = -1;  // SAMPLER template id of the last cast_kernel launch
    uint8_t *d_vol_brick = nullptr;
    uint8_t *d_vol_padded = nullptr;  // (nx+1)(ny+1)(nz+1), edge replicated: unclamped LDG path
    uint8_t *d_cells = nullptr;       // trilinear policy: 8-byte cell records, (n+1+2G) per axis (vr5_tri.cuh)
    uint64_t cells_bytes = 0;
    int tri_fast = 0;                 // last trilinear launch used the unclamped 32-bit-address march
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
    // frame lanes (vr5_set_overlap): frames go round-robin over n_lanes streams so that the tail of one frame's kernel
    // (few long rays left) overlaps the start of the next. Each lane owns a tile-queue counter (monotonic, see
    // KArgs::queue_base) and an internal framebuffer for host outputs. With one lane, lane 0 runs on `stream`.
    struct Lane {
        cudaStream_t stream = nullptr;
        unsigned *d_queue = nullptr;
        uint32_t queue_base = 0;
        cudaEvent_t done = nullptr;
        bool pending = false;        // a frame was launched on this lane and `stream` has not been made to wait for it
        uint8_t *d_frame = nullptr;
    } lanes[4];
    int n_lanes = 1;
    uint64_t frame_no = 0;
    cudaEvent_t ev_fork = nullptr;
    cudaStream_t launch_stream = nullptr;  // set by launch() for launch_one()
    uint32_t last_grid = 0;
    void *d_counters = nullptr;
    void *d_leap_stats = nullptr;   // two 64-bit counters: leaps taken by the zero-skip kernels, samples they left out
    // partition
    int rank = 0, world = 1, packed = 0;
    int band_first = 0, band_count = -1;   // >= 0: contiguous band of 4-row strips instead of the interleaved ones (vr5_set_partition_band)
    bool band_auto = false;                // the band is re-balanced for every frame's view (vr5_set_partition_bands)
    // peer ring (vr5_peer.cu): owner allocates, writers map it
    uint8_t *peer_base = nullptr;     // header (flags) + slots frames
    bool peer_owner = false, peer_ipc_mapped = false;
    int peer_slots = 0;
    size_t peer_frame_stride = 0;
    uint32_t peer_released_slot[32] = {0};  // owner, host side: releases per slot so far
    // host frame ring (vr5_hostring.cu): shared host memory, every rank copies its strips there over its own PCIe link
    uint8_t *hr_base = nullptr, *hr_dev = nullptr;  // host address / device alias of the page-locked block
    int hr_slots = 0;
    bool hr_registered = false;       // this handle page-locked the block (and un-registers it on detach)
    // tile-queue group order (vr5_render.cu, build_group_order): device table and the geometry it was built for
    uint32_t *d_group_order = nullptr;
    uint32_t order_key[6] = {0, 0, 0, 0, 0, 0};  // width, height, rank, world, packed, mode
    // frame sequences (vr5_render_sequence, vr5_peer_render_sequence): ONE persistent launch renders all frames of a camera
    // path from one tile queue; own monotonic queue counter (the launches run on `stream`, in order)
    unsigned *d_seq_queue = nullptr;
    uint32_t seq_queue_base = 0;
    // consumer side of a frame ring driven by sequences (vr5_peer_consume_sequence): stream memory operations on their own stream
    cudaStream_t consumer_stream = nullptr;
    cudaEvent_t consumer_done = nullptr;
    bool consumer_pending = false;
    // experiment knobs: read from the environment ONCE in vr5_create, changed afterwards only through vr5_set_option
    struct Options {
        int queue_order = 0;        // VR5_QUEUE_ORDER: 0 rows (default), 1 cols, 2 rings
        bool warp_fetch = false;    // VR5_WARP_FETCH: every warp takes its own queue slot even when the TF has no opaque entry
        bool tri_clamped = false;   // VR5_TRI_CLAMPED: trilinear march with clamped record indices
        int trif_ctas = 6;          // VR5_TRIF_CTAS: register budget of the contracted trilinear march (CTAs per SM)
        int trif_depth = 2;         // VR5_TRIF_DEPTH: cell requests in flight per ray
        int seq_debug = 0;          // VR5_SEQ_DEBUG (measurement only, breaks the ring protocol): 1 = sequences do not wait for releases, 4 = they run the wait but every slot counts as released, 8 = per-CTA queue reads forced
        bool host_single = false;   // VR5_HOST_SINGLE: the blocking host-buffer call renders the frame with one launch + one copy (no chunks)
        bool esl_exact = false;     // VR5_ESL_EXACT: the leading empty-space loop divides in every iteration (no predicted leaps)
        bool no_p2 = false;         // VR5_NO_P2: never build the power-of-two window
        bool leap_stats = false;    // VR5_LEAP_STATS: the zero-skip kernels count their leaps (vr5_leap_stats; contended atomics: not for timed runs)
        bool no_leap = false;       // VR5_NO_LEAP: the zero-skip kernels march through transparent bricks sample by sample (no brick table)
        int leap_brick_log2 = 0;    // VR5_LEAP_BRICK: log2 of the brick edge of that table (0 = default: <= 32 bricks per axis, edge >= 8)
        bool no_zskip = false;      // VR5_NO_ZSKIP: samples of the TF's all-zero prefix are composited like any other (no ZS kernels)
        int p2_wide = 0;            // VR5_P2_WIDE: 1 = 64-bit window addressing (kP2Wide) also below 4 GiB, -1 = never (texture alone)
        uint32_t win_sx = 0;        // VR5_WIN_SX: row stride of the window layouts (0 = default)
    } opt;
    // bookkeeping
    int ctas_per_sm[520] = {0};  // occupancy cache per kernel variant
    uint64_t launches = 0;
    void *d_flush = nullptr;
    size_t flush_bytes = 0;
    std::string err;
};

void vr5_fail(vr5_renderer *r, const char *what, cudaError_t e, const char *file, int line);
// lane plumbing for vr5_peer.cu: stream of a lane, fork from the user stream before work on a lane, mark its end
cudaStream_t vr5_lane_stream(vr5_renderer *r, int lane);
int vr5_lane_begin(vr5_renderer *r, int lane);
int vr5_lane_end(vr5_renderer *r, int lane);
int vr5_render_into(vr5_renderer *r, const struct vr5_params *p, void *out_dev, int lane);  // async launch of the cast kernel, device buffer
int vr5_adopt_volume(vr5_renderer *r, uint8_t *dev_owned, uint64_t bytes, uint16_t dx, uint16_t dy, uint16_t dz);
// where a frame of a sequence goes and how its completion is signalled (frame-ring hand-shake words, vr5_peer.cu), both optional
struct vr5_frame_target {
    void *out_dev;
    uint32_t *arrive;          // +1 (system-scope atomic, after a system fence) when this rank's last tile of the frame is stored
    const uint32_t *consumed;  // must reach consumed_target before the frame's pixels are stored
    uint32_t consumed_target;
};
// async launch of ONE persistent kernel over frames[0..n) on the renderer's stream (vr5_render.cu)
int vr5_sequence_launch(vr5_renderer *r, const struct vr5_params *frames, int n, const vr5_frame_target *targets, uint32_t *timed_out_dev, int targets_remote);

#define VR5_CUDA(r, call)                                         \
    do {                                                          \
        cudaError_t vr5_e_ = (call);                              \
        if (vr5_e_ != cudaSuccess) {                              \
            vr5_fail((r), #call, vr5_e_, __FILE__, __LINE__);     \
            return 1;                                             \
        }                                                         \
    } while (0)
