/*
 * vr5.h - C ABI of the B200-native ray caster ("GPURenderer5"), the drop-in boundary for the one hot path
 *         volume + transfer function + view parameters in  ->  RGBA8 framebuffer out.
 *
 * Every entry point replaces one piece of the reference's Renderer plugin interface (Renderer.h:13-28) as the
 * reference's CUDA renderers implement it; the reference file:line each one stands in for is cited per function.
 * Plain pointers and sizes only: no C++, CUDA or torch types. The C++ adaptor `class GPURenderer5 : public
 * Renderer` (volume-rendering_b200/csrc/GPURenderer5.{h,cpp}) and the headless bench drivers (bench.py through
 * ctypes, tools/volr_headless.cpp) are thin callers of this layer. See INTEGRATION.md for the binding a VolR
 * maintainer adds.
 *
 * Conventions (reference cuda_utils.h:21-49, GPURenderer1.cu:86-112): functions return int, 0 = success,
 * 1 = failure (missing input, allocation failure, CUDA error). The library never calls exit(); the message of the
 * last failure is kept per handle (vr5_last_error). There is NO CPU fallback: without a CUDA device vr5_create fails.
 *
 * Threading: like the reference (single GLUT thread) a handle is not re-entrant. Rendering is blocking by default
 * (the frame is complete when vr5_render returns); vr5_render_async + vr5_sync expose the stream for overlap.
 */
#ifndef VR5_H
#define VR5_H

#include "vr5_pod.h"

#ifdef __cplusplus
extern "C" {
#endif

#define VR5_VERSION 1

typedef struct vr5_renderer vr5_renderer; /* opaque; owns every device allocation it makes */

/* Per-frame parameters: `Raycaster` (RaycasterBase.h:20-30) minus the three pointers, which the renderer owns as
 * device copies (the reference overwrites them with device pointers in render_volume, GPURenderer1.cu:103-105). */
typedef struct vr5_params {
    vr5_view view;           /* ViewBase.h:14-21 */
    float ray_step;          /* RaycasterBase.h:24 */
    float ray_threshold;     /* RaycasterBase.h:25 */
    uint32_t esl;            /* RaycasterBase.h:26 (bool) */
    uint32_t esl_block_dims; /* RaycasterBase.h:28 */
    vr5_f3 esl_block_size;   /* RaycasterBase.h:29 */
    float light_kd;          /* RaycasterBase.h:30 */
    vr5_f3 min_bound;        /* Model::min_bound (ModelBase.h:15); (-1,-1,-1) in every reference path */
} vr5_params;

/* Work counters of one frame (instrumented kernel variant; never the timed kernel). SURVEY 8(d). */
typedef struct vr5_counters {
    uint64_t rays_hit;   /* rays with a non-empty clip interval (intersect() true) */
    uint64_t rays_accum; /* rays that entered the accumulation loop */
    uint64_t esl_iters;  /* empty-space-leap iterations (CPURenderer.cpp:18-25) */
    uint64_t samples;    /* accumulation-loop iterations = classified voxel fetches: the bench unit */
    uint64_t shaded;     /* samples that took the shading path (second voxel fetch) */
} vr5_counters;

/* Voxel sampler policies. 0..6 are nearest-neighbour with bit-identical results (they differ in memory path only). */
enum { VR5_SAMPLER_LINEAR = 0,  /* x-fastest linear array, LDG.U8 (the reference's layout, ModelBase.h:18-22) */
       VR5_SAMPLER_BRICK = 1,   /* 8x4x4-voxel bricks (one 128-B line each, 4x4x2 per 32-B sector), LDG.U8 */
       VR5_SAMPLER_TEX3D = 2,   /* cudaArray 3-D texture, integer-coordinate point fetch (TLD) */
       VR5_SAMPLER_HYBRID = 3,  /* samples alternate between the 3-D texture (TEX pipe) and a padded linear array (LSU pipe) */
       VR5_SAMPLER_HYBRID_WINDOW = 4,  /* texture : window layout 1 : 1. Window = doubly padded linear array inside one 4-GiB
                                        * address window, indexed by float bit patterns (no F2I, 32-bit address arithmetic) */
       VR5_SAMPLER_WINDOW = 5,         /* window layout only */
       VR5_SAMPLER_HYBRID_WINDOW2 = 6, /* texture : window 1 : 2 */
       /* NOT a parity policy: GPURenderer4's semantics (GPURenderer4.cu:33-97) - trilinear volume fetch, linearly
        * interpolated transfer function - evaluated in software fp32; 8 bytes per sample from an 8-byte-per-cell layout.
        * Bit-equal to the restated CPU checker (trilinear section of the test oracle); within a stated tolerance of the reference's
        * own GPURenderer4 run in the texture unit (tests/test_trilinear_pin.py), SURVEY 8(f) rank 2. */
       VR5_SAMPLER_TRILINEAR = 16,
       /* The same policy and memory layout with CONTRACTED arithmetic: every a + b*c of the march is one fused multiply-add,
        * as nvcc's default -fmad=true builds the reference's GPU renderers (GPURenderer4.cu:75-86), and the filtering
        * formulas are evaluated in their fused form (position fma(p, n/2, n/2 - 0.5), lerp fma(w, b - a, a)). Checked bit for
        * bit against the "trilinear-fma" section of the CPU checker; agrees with VR5_SAMPLER_TRILINEAR within 1 LSB. */
       VR5_SAMPLER_TRILINEAR_FMA = 18 };

/* ---- lifecycle ------------------------------------------------------------------------------------------ */

/* GPURenderer1 constructor minus the three set_* calls (GPURenderer1.cu:17-21) and cuda_init's device pick
 * (VolR.cpp:128-198). `device` is a CUDA ordinal; -1 = current device. */
int vr5_create(int device, vr5_renderer **out);
/* ~GPURenderer1 / ~GPURenderer4 (GPURenderer1.cu:23-28, GPURenderer4.cu:24-31): frees all device state. */
void vr5_destroy(vr5_renderer *r);
/* Renderer::get_name (Renderer.h:16-18) */
const char *vr5_name(void);
/* message of the last failure on this handle ("" if none); with r == NULL, of the last failed vr5_create */
const char *vr5_last_error(const vr5_renderer *r);
int vr5_version(void);

/* ---- scene state (host data in, device copies owned by the renderer) -------------------------------------- */

/* Renderer::set_window_buffer(View) (GPURenderer1.cu:74-84): (re)allocates the internal framebuffer. */
int vr5_set_window(vr5_renderer *r, uint16_t width, uint16_t height);
/* Renderer::set_transfer_fn(Raycaster) (GPURenderer1.cu:65-72): uploads the 128 pre-multiplied RGBA entries AND the
 * 1024-word ESL bitfield (both change together, RaycasterBase.cpp:46-74). */
int vr5_set_transfer_fn(vr5_renderer *r, const float *tf_premult_rgba128, const uint32_t *esl_bits1024);
/* Renderer::set_volume(Model) (GPURenderer1.cu:86-98; GPURenderer4.cu:114-143): uploads the x-fastest uint8 grid and
 * builds the layout of the active sampler. `bytes` must equal dx*dy*dz (64 bit: SURVEY H3). 1 = no data / no memory. */
int vr5_set_volume(vr5_renderer *r, const uint8_t *host_data, uint64_t bytes, uint16_t dx, uint16_t dy, uint16_t dz);
/* Same, from a linear grid already in device memory (device generator, NCCL broadcast). The renderer copies it. */
int vr5_set_volume_device(vr5_renderer *r, const uint8_t *dev_data, uint64_t bytes, uint16_t dx, uint16_t dy, uint16_t dz);
/* Copies the linear grid back to the host (tests: hand the oracle the exact bytes the kernel sees). */
int vr5_get_volume(vr5_renderer *r, uint8_t *host_out, uint64_t bytes);
/* Select VR5_SAMPLER_*; re-lays out the current volume if one is set. Default: VR5_SAMPLER_HYBRID_WINDOW (fastest on B200) or $VR5_SAMPLER. */
int vr5_set_sampler(vr5_renderer *r, int sampler);
int vr5_get_sampler(const vr5_renderer *r);
/* Memory path the LAST launch actually took: the VR5_SAMPLER_* id of the kernel variant (a window sampler renders a
 * frame through the clamped texture path when the frame's step could leave the guard margin; ids 7..9 are the
 * power-of-two specialisations of 4..6; VR5_SAMPLER_TRILINEAR + 1 is the trilinear march with unclamped 32-bit record
 * addressing, VR5_SAMPLER_TRILINEAR itself the clamped one, likewise VR5_SAMPLER_TRILINEAR_FMA (+ 1)), -1 before any launch. */
int vr5_last_variant(const vr5_renderer *r);
/* 1 when the LAST nearest-neighbour launch was a zero-skip build: the transfer function set by vr5_set_transfer_fn starts with
 * all-zero entries (the reference's own tables do: RaycasterBase.cpp:76-84, alpha 0 below entry 13, colours pre-multiplied) and
 * samples that classify to them skip the compositing they cannot change (CPURenderer.cpp:29-36: acc + 0 * (1 - acc.a)). Same
 * pixels; VR5_NO_ZSKIP=1 turns it off. 0 otherwise, -1 for a NULL handle. */
int vr5_last_zero_skip(const vr5_renderer *r);

/* ---- the hot path ----------------------------------------------------------------------------------------- */

/* Renderer::render_volume(uchar4*, Raycaster) (GPURenderer1.cu:100-112, GPURenderer23.cu:60-81). Writes width*height
 * RGBA8 pixels, row 0 first (RaycasterBase.h:44-50); every pixel is written (misses as 0), so no separate clear pass.
 * out_is_device: 0 = host pointer (D2H copy inside, as renderers 1-2), 1 = device pointer (rendered in place, as
 * renderers 3-4 into the mapped PBO, VolR.cpp:76-87), -1 = detect with cudaPointerGetAttributes. Blocking.
 * The view dims must match vr5_set_window. */
int vr5_render(vr5_renderer *r, const vr5_params *p, void *out_rgba8, int out_is_device);
/* Same call taking the reference's 160-byte POD as is (its three pointers are ignored, GPURenderer1.cu:103-105). */
int vr5_render_raycaster(vr5_renderer *r, const vr5_raycaster *rc, void *out_rgba8, int out_is_device);
/* Non-blocking variant for device buffers (or pinned host buffers): enqueue on the renderer's stream and return. */
int vr5_render_async(vr5_renderer *r, const vr5_params *p, void *out_rgba8, int out_is_device);
/* cudaStreamSynchronize on the renderer's stream and on every frame lane. */
int vr5_sync(vr5_renderer *r);
/* A whole camera path with ONE launch (the reference's benchmark loop calls render_volume once per frame, VolR.cpp:225-321):
 * frames[0..n_frames) into the device buffers outs_dev[0..n_frames), asynchronously on the renderer's stream. The persistent
 * grid pulls the tiles of ALL frames from one queue, so the launch has one tail however many frames it renders and the host
 * pays one submission. The frames may differ in view.origin / direction / right_plane / up_plane only. Pixels are bit-equal
 * to n_frames vr5_render calls. */
int vr5_render_sequence(vr5_renderer *r, const vr5_params *frames, int n_frames, void *const *outs_dev);
/* vr5_render_sequence for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc). NO COUNTERPART IN THE REFERENCE. Same arguments, same
 * pixels; the per-frame records the kernel reads live in a file-scope device array instead of a block allocated per call, so
 * the call makes no cudaMalloc / cudaFree / stream-ordered allocation. Limits: n_frames <= VR5_NOALLOC_MAX_FRAMES (more is
 * refused with an error; the array costs 24 KiB on every GPU that loads the library, until exit); VR5_QUEUE_ORDER must be
 * "rows". The array is shared by every call on the device and keeps the last call's records (a launch rewrites in full what
 * it reads, nothing has to start at zero): calls that OVERLAP on the GPU - other handles, other streams - race; calls on one
 * handle are stream-ordered and safe. Volume, transfer function and window must have been set before (those calls allocate). */
#define VR5_NOALLOC_MAX_FRAMES 256
int vr5_render_sequence_noalloc(vr5_renderer *r, const vr5_params *frames, int n_frames, void *const *outs_dev);
/* Experiment knobs by name (the VR5_* environment variables vr5_create reads ONCE: VR5_QUEUE_ORDER = rows | cols | rings,
 * VR5_WARP_FETCH, VR5_TRI_CLAMPED, VR5_TRIF_CTAS, VR5_TRIF_DEPTH, VR5_NO_P2, VR5_NO_ZSKIP, VR5_NO_LEAP, VR5_LEAP_BRICK, VR5_LEAP_STATS (the zero-skip kernels count their leaps for vr5_leap_stats: contended atomics, not for timed runs), VR5_P2_WIDE = 1 | -1 (64-bit window addressing for every volume | never), VR5_WIN_SX, VR5_ESL_EXACT = the leading empty-space
 * loop divides in every iteration, VR5_HOST_SINGLE = the blocking host-buffer call uses one launch + one copy). No effect on
 * the pixels. (VR5_SEQ_DEBUG is a measurement knob of the ring hand-shake that BREAKS the protocol: never set it in production.) */
int vr5_set_option(vr5_renderer *r, const char *name, const char *value);
/* Use an external stream (a cudaStream_t passed as void*, e.g. torch's current stream); NULL restores the own stream. */
int vr5_set_stream(vr5_renderer *r, void *cuda_stream);
/* Frame overlap. lanes = 1 (default): every frame runs on the renderer's stream, in order - the reference's behaviour.
 * lanes = 2..4: successive frames (vr5_render*, vr5_peer_render) go round-robin over that many internal streams, so the
 * tail of one frame's kernel (a few long rays left, most SMs idle) overlaps the start of the next frame, and a host
 * frame's D2H copy overlaps the next kernel. Each lane is ordered AFTER what the caller enqueued on the renderer's stream
 * before the call; the renderer's stream waits for the lanes only in vr5_join / vr5_sync (a blocking vr5_render waits for
 * its own frame). Output buffers of frames in flight must differ. */
int vr5_set_overlap(vr5_renderer *r, int lanes);
int vr5_get_overlap(const vr5_renderer *r);
/* Make the renderer's stream wait (on the device) for every frame launched so far. No-op with one lane. */
int vr5_join(vr5_renderer *r);
/* Block the host until lane `lane` is idle. The k-th vr5_render* call since vr5_set_overlap runs on lane k % lanes. */
int vr5_sync_lane(vr5_renderer *r, int lane);

/* ---- measurement extras (headless bench driver; replace Profiler.cpp:46-78) -------------------------------- */

/* Render `n_frames` frames back to back into a device buffer, timing EACH kernel launch with CUDA events on the
 * launching stream; ms_out[n_frames]. l2_flush != 0 writes a >L2-sized scratch buffer between frames (untimed). */
int vr5_render_timed(vr5_renderer *r, const vr5_params *frames, int n_frames, void *out_dev_rgba8, int l2_flush, float *ms_out);
/* Run the instrumented variant of the kernel on one frame and return its work counters (no framebuffer needed). */
int vr5_count(vr5_renderer *r, const vr5_params *p, vr5_counters *out);
/* Leaps over transparent bricks taken by the zero-skip kernels since the last call (out[0]) and the samples they left out (out[1]):
 * the reference marches through empty space sample by sample after its leading loop (CPURenderer.cpp:28-39); here a run of samples
 * that provably read voxels of a brick whose maximum classifies to an all-zero TF entry is left out, and the ray parameter after
 * it is the one the reference's sequential k += step would hold. Waits for the handle's work; resets the counters.
 * Counted only while the option VR5_LEAP_STATS is set (two atomics on one address per leap: never in a timed run). VR5_NO_LEAP=1 (read at
 * the next volume / sampler change) turns the leaps off, VR5_LEAP_BRICK = log2 of the brick edge (default: 32 voxels, at most 32 bricks
 * per axis). Same pixels either way. */
int vr5_leap_stats(vr5_renderer *r, uint64_t out[2]);
/* Number of kernels this handle has launched so far (render, count, layout, synth). */
uint64_t vr5_launch_count(const vr5_renderer *r);
/* Pinned host memory for the end-to-end path (cudaHostAlloc / cudaFreeHost). */
void *vr5_host_alloc(uint64_t bytes);
void vr5_host_free(void *p);
/* Device scratch owned by the caller (plain cudaMalloc/cudaFree on the renderer's device). */
void *vr5_device_alloc(vr5_renderer *r, uint64_t bytes);
void vr5_device_free(vr5_renderer *r, void *p);
int vr5_memcpy_d2h(vr5_renderer *r, void *host_dst, const void *dev_src, uint64_t bytes);

/* ---- image-space partition for multi-GPU (the reference is single-GPU) ------------------------ */

/* Rank `rank` of `world` renders the 4-row strips s with s % world == rank (interleaved for load balance).
 * packed = 0: pixels go to their place in a full width*height frame (other ranks' rows untouched);
 * packed = 1: the rank's strips are stored contiguously, strip j of the rank at row 4*j (vr5_strip_rows rows). */
int vr5_set_partition(vr5_renderer *r, int rank, int world, int packed);
/* Contiguous alternative: this renderer renders the 4-row strips first_strip .. first_strip + n_strips - 1 (pixels at their
 * place in the full frame). A band touches only its slab of the volume - at high GPU counts the interleaved strips make every
 * GPU read the whole volume for 1/N of the samples - but its cost depends on the view: the caller balances the bands
 * (vr5_balance_bands). vr5_set_partition returns to interleaved strips. */
int vr5_set_partition_band(vr5_renderer *r, int first_strip, int n_strips);
/* Cost-balanced bands, re-balanced for every frame: rank `rank` of `world` renders band `rank` of vr5_balance_bands(frame's
 * parameters, world). Every rank computes the same bounds from the same parameter bytes. Unpacked (pixels at their place). */
int vr5_set_partition_bands(vr5_renderer *r, int rank, int world);
/* first_strip_out[0 .. world]: band i = strips first_strip_out[i] .. first_strip_out[i + 1] - 1 (4-row strips), cut so that the
 * estimated cost (clipped ray length of a sparse grid of rays + a per-ray constant) is equal per band. Host arithmetic only. */
int vr5_balance_bands(const vr5_params *p, int world, uint32_t *first_strip_out);
/* rows of a packed strip buffer for (height, rank, world): 4 * number of strips owned (last strip may be short) */
uint32_t vr5_strip_rows(uint16_t height, int rank, int world);
/* On the gathering rank: scatter `world` packed strip buffers (concatenated, each padded to `rows_padded` rows) into
 * a full frame. Both pointers are device pointers. */
int vr5_unpack_strips(vr5_renderer *r, const void *packed_all, uint32_t rows_padded, int world, void *frame_rgba8);

/* ---- fused render + gather over NVLink peer memory (replaces render -> NCCL gather -> unpack) ---- */

/* The gathering rank ("owner") holds a ring of `slots` full frames plus flag words in one device allocation; every other
 * rank maps it (CUDA IPC) and its ray-cast kernel stores each pixel straight into its final place in the owner's frame
 * over NVLink. Frames carry a sequence number seq = 0, 1, 2, ... agreed by all ranks; frame seq lives in slot seq % slots.
 * Flags: arrive[slot] counts ranks that finished their strips of a frame in that slot (the owner included); consumed[slot] =
 * frames of it released. A rank may store frame seq only after frame seq - slots was released; the owner may read frame seq
 * only after all `world` ranks arrived. With frame overlap (vr5_set_overlap) frame seq runs on lane seq % lanes - slots must be a
 * multiple of lanes - and vr5_peer_wait / read / release of a frame are enqueued on that frame's lane. Both conditions are enforced on the device by one-thread wait kernels (bounded: a wait
 * that exceeds the timeout sets a status flag instead of hanging). One process per GPU; never two ranks on one GPU. */
#define VR5_PEER_HANDLE_BYTES 64
/* owner: allocate the ring for the current window (vr5_set_window first); 1 <= slots <= 32 */
int vr5_peer_create(vr5_renderer *r, int slots);
/* owner: IPC handle of the ring (cudaIpcMemHandle_t, 64 bytes) to be sent to the other ranks by any host transport */
int vr5_peer_export(vr5_renderer *r, void *handle_out64);
/* writer: map the owner's ring; `slots` and the window must equal the owner's */
int vr5_peer_open(vr5_renderer *r, const void *handle64, int slots);
/* same-process variant (tests; two renderers of one process): share the owner's ring directly */
int vr5_peer_open_local(vr5_renderer *r, vr5_renderer *owner);
/* any rank: render this rank's strips (vr5_set_partition, packed = 0) of frame `seq` into the ring, then signal. Async. */
int vr5_peer_render(vr5_renderer *r, const vr5_params *p, uint64_t seq);
/* owner: enqueue the wait for all writers of frame `seq`; *frame_dev_out = device pointer of the complete frame, valid
 * for work enqueued on the renderer's stream after this call and before vr5_peer_release(seq) */
int vr5_peer_wait(vr5_renderer *r, uint64_t seq, void **frame_dev_out);
/* owner: async D2H copy of frame `seq` (after vr5_peer_wait(seq)) into a (pinned) host buffer, on the frame's lane */
int vr5_peer_read(vr5_renderer *r, uint64_t seq, void *host_dst);
/* owner: frame `seq` is consumed; its slot may be overwritten by frame seq + slots */
int vr5_peer_release(vr5_renderer *r, uint64_t seq);
/* any rank: block the host until everything enqueued for frame `seq` (render, wait, read, release) has finished */
int vr5_peer_sync_frame(vr5_renderer *r, uint64_t seq);
/* Frame sequences through the ring: every rank renders its strips of frames seq0 .. seq0 + n - 1 with ONE persistent launch
 * (see vr5_render_sequence) on its renderer's stream. The hand-shake runs inside the kernel: a warp stores pixels of a frame
 * only after the slot's previous frame was released, the warp that finishes the rank's last tile of a frame counts the
 * rank's arrival. The owner calls it too, for its own strips. Asynchronous. */
int vr5_peer_render_sequence(vr5_renderer *r, const vr5_params *frames, int n, uint64_t seq0);
/* owner: consume frames seq0 .. seq0 + n - 1 in order on an internal stream (stream memory operations, no SM needed): wait
 * for all ranks' arrivals of a frame, copy it device-to-device to copy_dst_dev + i * width * height * 4 if copy_dst_dev is
 * not NULL, release its slot. Asynchronous; vr5_join orders the renderer's stream behind it, vr5_sync waits for it. */
int vr5_peer_consume_sequence(vr5_renderer *r, uint64_t seq0, int n, void *copy_dst_dev);
/* 0 = ok, 1 = some device-side wait timed out since vr5_peer_create / vr5_peer_open (synchronises the stream) */
int vr5_peer_status(vr5_renderer *r);
void vr5_peer_close(vr5_renderer *r);

/* ---- host frame ring: every rank delivers its strips to HOST memory over its own PCIe link (vr5_hostring.cu) ---- */

/* The reference's renderers 1-2 hand the finished frame to the caller in host memory (VolR.cpp:76-87). With several GPUs,
 * gathering on one device and copying from there makes that device's PCIe link the limit; here the ring of `slots` frames
 * lives in host memory visible to all ranks (POSIX shared memory between processes; any memory inside one process). Every
 * rank renders its strips packed (vr5_set_partition(..., packed = 1)) and copies them straight to their rows of the
 * frame's slot (or, with a band partition - vr5_set_partition_bands - renders its band in place and copies that block of rows). Frame seq uses slot seq % slots and lane seq % lanes (slots a multiple of lanes). No device waits for
 * another one; both hand-shakes (slot released / all strips delivered) are bounded host-side polls of counters in the
 * ring's header (4 s, then an error). */
uint64_t vr5_hostring_bytes(uint16_t width, uint16_t height, int slots);   /* header + slots frames; slots 1..16 */
/* Page-lock `base` for this device (cudaHostRegister; already pinned memory is used as it is) and use it as the ring for the
 * current window and partition. init = 1 on exactly one rank, before the others attach: clears the header. */
int vr5_hostring_attach(vr5_renderer *r, void *base, uint64_t bytes, int slots, int init);
/* any rank: render this rank's strips of frame `seq`, copy them into the ring, signal. Asynchronous on the frame's lane;
 * blocks the host (bounded) only while frame seq - slots has not been released. */
int vr5_hostring_render(vr5_renderer *r, const vr5_params *p, uint64_t seq);
/* consumer: block (bounded) until all `world` ranks delivered frame `seq`; *frame_out = the frame inside the ring, valid
 * until vr5_hostring_release(seq) */
int vr5_hostring_wait(vr5_renderer *r, uint64_t seq, const void **frame_out);
int vr5_hostring_release(vr5_renderer *r, uint64_t seq);
void vr5_hostring_detach(vr5_renderer *r);

/* ---- synthetic volumes generated on the device (bench inputs) ------------------------------- */

/* kind 0 = synthA (dense Marschner-Lobb-like), 1 = synthB (sparse spheres); integer arithmetic, byte-identical to
 * the host generator the oracle uses (oracle/synth_volume.c). Result becomes the renderer's current volume. */
int vr5_synth_volume(vr5_renderer *r, int kind, uint16_t n, uint32_t seed);

/* ---- scene builders (host side, once per change): what RaycasterBase / ViewBase manufacture ----------------- */

/* RaycasterBase::reset_transfer_fn default ramp (RaycasterBase.cpp:76-84): base (non pre-multiplied) 128 x RGBA */
void vr5_scene_default_base_tf(float *base_rgba128);
/* RaycasterBase::update_transfer_fn part 1 (RaycasterBase.cpp:46-52) */
void vr5_scene_premultiply_tf(const float *base_rgba128, float *premult_rgba128);
/* RaycasterBase::set_volume (RaycasterBase.cpp:94-125): ESL block edge, 32^3 x {min,max}, block size in world units */
void vr5_scene_build_minmax(const uint8_t *data, uint16_t dx, uint16_t dy, uint16_t dz, uint16_t *esl_block_dims,
                            uint8_t *minmax_32768x2, vr5_f3 *esl_block_size);
/* RaycasterBase::update_transfer_fn part 2 (RaycasterBase.cpp:53-67) */
void vr5_scene_build_esl_bits(const float *premult_rgba128, const uint8_t *minmax_32768x2, uint32_t *esl_bits1024);
/* RaycasterBase::reset_ray_step (RaycasterBase.cpp:86-92) */
float vr5_scene_default_ray_step(uint16_t dx, uint16_t dy, uint16_t dz);
/* ViewBase::set_viewport_dims + toggle_perspective(1) + set_camera_position (ViewBase.cpp:49-113, VolR.cpp:232-247)
 * with an own matrix stack instead of the GL one; light at its default (0,0,3). */
void vr5_scene_camera(uint16_t width, uint16_t height, int perspective, float ax_deg, float ay_deg, float az_deg,
                      float distance, vr5_view *out);
/* Same as vr5_scene_build_minmax but computed on the device from the renderer's current volume */
int vr5_build_minmax_device(vr5_renderer *r, uint16_t *esl_block_dims, uint8_t *minmax_32768x2, vr5_f3 *esl_block_size);
/* ModelBase::compute_histogram (ModelBase.cpp:19-33) on the device from the current volume: 256 normalised bins */
int vr5_histogram_device(vr5_renderer *r, float *hist256);

/* ---- volume files (host side; SURVEY 8(f) rank 3: ddsbase.cpp, ModelBase.cpp:35-109) ------------------------ */

/* readPVMvolume (ddsbase.cpp:343-435): PVM / PVM2 / PVM3 container, DDS-compressed ("DDS v3d" / "DDS v3e") or plain.
 * *data (malloc, release with vr5_volume_free) receives width*height*depth*components bytes, x fastest. 0 ok / 1 error
 * (never exits, unlike the reference's ERRORMSG); message in vr5_volume_io_error(). scale_xyz may be NULL. */
int vr5_load_pvm(const char *path, uint8_t **data, uint32_t *width, uint32_t *height, uint32_t *depth, uint32_t *components,
                 float *scale_xyz);
/* ModelBase::load_model's RAW branch (ModelBase.cpp:66-84); the dimensions the reference prompts for are arguments */
int vr5_load_raw(const char *path, uint32_t width, uint32_t height, uint32_t depth, uint32_t components, uint8_t **data);
/* quantize() (ddsbase.cpp:475-558): 16-bit big-endian voxels to 8 bit, linear or the reference's non-linear mapping */
int vr5_quantize16(const uint8_t *data16_be, uint32_t width, uint32_t height, uint32_t depth, int linear, uint8_t **data8);
/* ModelBase::load_model for a .pvm file: decode, quantise 2-component data, refuse more (ModelBase.cpp:85-93) */
int vr5_load_volume(const char *path, uint8_t **data8, uint32_t *width, uint32_t *height, uint32_t *depth);
/* the Differential Data Stream itself (ddsbase.cpp:187-245), magic included; the encoder exists for tests and tools */
int vr5_dds_decode(const uint8_t *stream, uint64_t bytes, uint8_t **out, uint64_t *out_bytes);
int vr5_dds_encode(const uint8_t *data, uint64_t bytes, uint32_t skip, uint32_t strip, int v3e, uint8_t **out, uint64_t *out_bytes);
void vr5_volume_free(void *p);
const char *vr5_volume_io_error(void);

#ifdef __cplusplus
}
#endif
#endif /* VR5_H */
