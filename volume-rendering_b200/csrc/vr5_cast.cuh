// vr5_cast.cuh - device side of the B200 ray caster: one ray per lane, 8x4-pixel warp tiles pulled from an atomic
// tile queue by persistent warps, transfer-function LUT and ESL bitfield staged in shared memory.
//
// Parity rule: every fp32 operation the reference's CPURenderer performs is
// performed here in the same order with the same IEEE-754 round-to-nearest result. All arithmetic goes through the
// __f*_rn intrinsics below, which nvcc never contracts into FMA (the TU is also built with -fmad=false), division
// and square root are the correctly rounded variants, denormals are kept (-ftz=false), min/max are the reference's
// ternaries (common.h:93-101), float->index is truncation (common.h:105-110). Integer index arithmetic is free to
// differ (layouts, 64-bit offsets) because it is exact.
//
// Two shortcuts leave work out WITHOUT touching that rule (both only when the host has seen an all-zero prefix in the transfer
// function, cast_ray<.., ZS>): a sample that classifies to an all-zero entry skips the compositing it cannot change, and runs of
// such samples inside bricks that hold nothing else are leapt over, landing on the ray parameter the reference's sequential
// k += step would hold (try_leap). The pixels stay the reference's, bit for bit (tests/test_zero_skip.py, tests/test_leap.py).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace vr5 {

constexpr int kCtaThreads = 256;  // 8 warps; each warp owns one 8x4 pixel tile at a time
constexpr int kTileW = 8;
constexpr int kTileH = 4;         // = strip height of the multi-GPU partition
#ifndef VR5_GROUP_LOG2X
#define VR5_GROUP_LOG2X 2
#endif
#ifndef VR5_GROUP_LOG2Y
#define VR5_GROUP_LOG2Y 3
#endif
constexpr int kGroupLog2X = VR5_GROUP_LOG2X;    // default queue order: groups of 4x8 tiles (32x32 px) so neighbouring queue slots
constexpr int kGroupLog2Y = VR5_GROUP_LOG2Y;    //                      touch neighbouring voxels (L1/L2 locality)
// Software-pipeline depth of the march loop: the voxel of sample i + kPrefetch is requested before sample i is composited
// (1 or 2; the launcher bounds the speculative reach with it, reach_ok() in vr5_render.cu). Depth 2 - two slots in an unrolled
// loop, no register moves, 55 instead of 54 instructions per pair of samples, bit-identical frames - measured SLOWER on B200:
// C3 0.585 against 0.561 ms per frame, C4 1.359 against 1.293 (profiles/r01_sampler_choice.md), so 1 stays the default.
#ifndef VR5_PREFETCH
#define VR5_PREFETCH 1
#endif
constexpr int kPrefetch = VR5_PREFETCH;
constexpr int kWindowGuard = 8;   // guard voxels around the window layout (speculative prefetch reach, see window_ok)
// 6 resident CTAs per SM (40 registers) measured best for the window variants (44 registers / 5 CTAs: 0.631 ms per C3 frame,
// 40 / 6: 0.617, 36 / 7: 0.717 - profiles/r01_sampler_choice.md). 0 = no minimum for the others.
#ifndef VR5_MIN_CTAS
#define VR5_MIN_CTAS (SAMPLER >= 2 ? (SHADE ? 5 : 6) : 0)
#endif

// Kernel variant ids (template parameter SAMPLER): 0..6 are the public VR5_SAMPLER_* policies. 7..9 are the power-of-two
// specialisations of the three window policies, chosen by the launcher when every volume dimension is a power of two:
// 7 = texture : window 1 : 1, 8 = window only, 9 = texture : window 1 : 2. 10 = variant 7 for a window array larger than 4 GiB
// (2048^3): the slice term of the address is a 32 x 32 -> 64-bit multiply-add, everything else as in 7.
constexpr int kP2First = 7;
constexpr int kP2Wide = 10;
constexpr int kP2Origin = 8;  // voxel index i lives at window / texture coordinate i + kP2Origin on every axis
__host__ __device__ constexpr bool is_p2(int s) { return s >= kP2First; }
__host__ __device__ constexpr int p2_mode(int s) { return s - kP2First; }

struct Counters {
    unsigned long long rays_hit, rays_accum, esl_iters, samples, shaded;
};

// One frame of a frame SEQUENCE (vr5_render_sequence / vr5_peer_render_sequence): what changes from frame to frame. The
// persistent grid pulls the tiles of all frames of the sequence from ONE queue (frame = slot / n_queue), so a launch has one
// tail however many frames it renders, and the host submits a whole camera path with one call. 80 bytes, 16-byte aligned.
struct FrameView {
    // direction, origin, right_plane, up_plane of the frame's View. The direction comes first: the kernel reads the record as
    // three 16-byte vectors, and (dx, dy) must land in an even-aligned register pair - the march multiplies it as one packed
    // operand; at offsets 3 and 4 the pair straddled two vector loads and cost two register moves per march step
    float dx, dy, dz, ox, oy, oz, rx, ry, rz, ux, uy, uz;
    uchar4 *out;               // the frame's buffer (a slot of a frame ring, possibly in a peer GPU's memory)
    // frame-ring hand-shake (vr5_peer.cu), both optional:
    uint32_t *arrive;          // += 1 (system scope, behind a system fence) by the warp that finishes this rank's last tile of the frame
    const uint32_t *consumed;  // must reach `consumed_target` (acquire, system scope) before any pixel of the frame is stored
    uint32_t consumed_target;
    // band partition (KArgs::bands): this rank's contiguous band of 4-row strips in THIS frame - the bands are balanced per
    // view, so they move from frame to frame
    uint32_t first_strip, n_strips, pad[3];
};
static_assert(sizeof(FrameView) == 96 && sizeof(FrameView) % 16 == 0, "FrameView layout");

struct RayView { float ox, oy, oz, dx, dy, dz, rx, ry, rz, ux, uy, uz; };

struct KArgs {
    // frame sequence: NULL / 1 for a single frame, whose view and buffer are the fields below
    const FrameView *frames;
    uint32_t n_frames;
    uint32_t *frame_done;    // sequence with hand-shake: queue slots finished per frame (zeroed before the launch)
    uint32_t *timed_out;     // set when a wait for `consumed` gives up (bounded: a missing consumer must not hang the GPU)
    uint32_t *ring_sync;     // sequence with hand-shake: [0] highest frame of this launch known to be writable (0xffffffff: none), [1] lock of the lane that polls the peer
    int ring_remote;         // the frames' buffers lie in a peer GPU's memory (system-scope fences before a slot is counted)
    int bands;               // sequence with per-frame bands: strip offset and count come from the frame's record (rank / tile_rows_local
                             // hold the single-frame values; tile_rows_local = the tallest band of the sequence sizes the queue)
    // view (ViewBase.h:14-21)
    float ox, oy, oz;        // origin
    float dx, dy, dz;        // direction
    float rx, ry, rz;        // right_plane
    float ux, uy, uz;        // up_plane
    float lx, ly, lz;        // light_pos
    int width, height;
    int half_w, half_h;      // dims / 2 (integer halves, ViewBase.h:26-27)
    int perspective;
    // volume (ModelBase.h:11-16)
    const uint8_t *vol;      // linear or bricked grid
    cudaTextureObject_t tex; // 3-D texture (VR5_SAMPLER_TEX3D)
    cudaTextureObject_t tex_p2; // power-of-two variants: texture whose coordinate is voxel index + kP2Origin
    uint32_t nx1, ny1, nz1;  // dims - 1 (index clamp)
    float hx, hy, hz;        // dims * 0.5f: (p + 1) * 0.5f * n == (p + 1) * (0.5f * n) exactly (power-of-two scaling)
    uint32_t sx;             // linear: row stride; brick: bricks per row
    uint64_t sxy;            // linear: slice stride; brick: bricks per slab
    const uint8_t *vol_pad;  // x-fastest grid padded by one replicated voxel per axis (unclamped LDG path), or NULL
    uint32_t psx, psxy;      // its row / slice strides
    // "window" layout (fetch mode 4): x-fastest grid with one replicated voxel on BOTH sides of every axis plus guard
    // margins, inside one 4-GiB-aligned address window. The low 32 address bits are bz*wsxy + by*wsx + bx, where
    // b = float bits of (v + wm) rounded toward zero: no F2I, no 64-bit address arithmetic (vr5_render.cu, build_window).
    uint32_t whi;            // high 32 bits of every address in the window
    uint32_t wsx, wsxy;      // row / slice strides
    float wmx, wmy, wmz;     // 2^23 + per-axis integer offsets that make the additive address constant vanish
    float mbx, mby, mbz;     // min_bound
    // march (RaycasterBase.h:24-30)
    float step, threshold;
    float shade_alpha;       // 0.05f when light_kd > 0.01f else +inf: folds the uniform half of CPURenderer.cpp:32
    uint32_t zero_lim;       // ZS kernels: voxel values below it classify to an all-zero TF entry (host: fill_args); 0 = none
    const uint8_t *brick_max;// ZS kernels: maximum voxel value per brick of (1 << brick_lb)^3 voxels (x fastest), or NULL: no leaps
    uint32_t brick_lb, brick_nx, brick_ny, brick_nz;   // log2 of the brick edge, bricks per axis
    unsigned long long *leap_stats;   // [0] leaps taken, [1] samples left out by them; NULL unless VR5_LEAP_STATS is set (contended atomics)
    float leap_margin;       // voxels a leap keeps away from the faces of its brick: 4 x the bound of the reference's position error
    float kd;
    int esl;
    uint32_t bd_magic;       // ceil(2^32 / esl_block_dims): exact idx / bd for idx < 2^16 via umulhi
    uint32_t bd_is_one;
    float bsx, bsy, bsz;     // esl_block_size
    // tables
    const float4 *tf;        // 128 pre-multiplied entries
    const uint32_t *esl_bits;// 1024 words
    // output
    uchar4 *out;
    int packed;              // 1: strips of this rank stored contiguously
    int rank, world;
    // tile queue
    unsigned *queue;         // this frame's counter
    uint32_t queue_base;     // value of the counter when this launch starts: the counter is never reset. Every launch adds
                             // exactly n_queue + (number of warps in the grid) to it (each warp ends on one failed fetch),
                             // so the host knows the next base without touching the device (vr5_render.cu, launch)
    uint32_t n_queue;        // padded tile count
    uint32_t tiles_x, tile_rows_local, groups_x, groups_y;
    uint32_t groups_x_magic;  // ceil(2^32 / groups_x): group / groups_x == umulhi(group, magic) for group < 2^16; 0: divide
    const uint32_t *group_order;  // permutation of the groups (build_group_order) or NULL: row-major
    int cta_fetch;           // 1: a CTA takes 8 queue slots at a time (one barrier per block), 0: every warp takes its own
    Counters *counters;
    // host side only: largest per-step displacement of a ray in voxels over the frame(s) (+inf when a view is not finite or
    // degenerate); the launcher compares it with the guard margins of the unclamped variants (vr5_render.cu, reach_ok)
    float reach;
};

// ---- exact fp32 helpers ----------------------------------------------------------------------------------
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fdiv(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float sel_min(float a, float b) { return a < b ? a : b; }  // common.h:93-96
__device__ __forceinline__ float sel_max(float a, float b) { return a > b ? a : b; }  // common.h:98-101

// common.h:105-110 map_float_int on an already scaled value: trunc toward zero, clamp to [0, n-1].
// cvt.rzi.u32.f32 saturates: negatives and NaN -> 0, >= 2^32 -> 0xffffffff -> n-1. Matches the x86 (long) cast for
// every |v| < 2^63 (beyond that, and for +inf, x86 yields LONG_MIN -> 0; unreachable for finite scenes).
__device__ __forceinline__ uint32_t trunc_clamp(float v, uint32_t n_minus_1) { return min(__float2uint_rz(v), n_minus_1); }

// (p + 1) * 0.5f * n with the two scalings merged (exact, see KArgs::hx)
__device__ __forceinline__ uint32_t world_to_voxel(float p, float half_n, uint32_t n_minus_1) {
    return trunc_clamp(fmul(fadd(p, 1.0f), half_n), n_minus_1);
}

// ---- voxel samplers (ModelBase.h:17-23; nearest neighbour, identical bytes on every path) ---------------------
template <int SAMPLER>
__device__ __forceinline__ uint32_t fetch_voxel(const KArgs &a, uint32_t ix, uint32_t iy, uint32_t iz);

template <>
__device__ __forceinline__ uint32_t fetch_voxel<0>(const KArgs &a, uint32_t ix, uint32_t iy, uint32_t iz) {
    return __ldg(a.vol + ((uint64_t)iz * a.sxy + (iy * a.sx + ix)));
}

// x-fastest array padded by one replicated voxel per axis (row stride a.psx, slice stride a.psxy < 2^32): three integer
// instructions of addressing (IMAD, 2 x IMAD.WIDE) and no clamp. Used by the unclamped linear sampler and by the LDG
// lanes of the hybrid sampler.
__device__ __forceinline__ uint32_t fetch_padded(const KArgs &a, uint32_t ix, uint32_t iy, uint32_t iz) {
    uint32_t v;
    asm("{\n\t.reg .u32 t;\n\t.reg .u64 p;\n\t"
        "mad.lo.u32 t, %2, %5, %1;\n\t"
        "mad.wide.u32 p, %3, %6, %4;\n\t"
        "mad.wide.u32 p, t, 1, p;\n\t"
        "ld.global.nc.u8 %0, [p];\n\t}"
        : "=r"(v) : "r"(ix), "r"(iy), "r"(iz), "l"(a.vol_pad), "r"(a.psx), "r"(a.psxy));
    return v;
}

// Window layout: the voxel index is the integer part of v (>= -1: the replicated low border stands in for the lower clamp,
// the replicated high border for the upper one). FADD.RZ against a constant in [2^23, 2^24) leaves floor(v) + const in the
// mantissa (exact for -2^22 < v < 2^22), so the float's bit pattern is used as the biased index directly.
__device__ __forceinline__ uint32_t fetch_window(uint32_t whi, uint32_t wsx, uint32_t wsxy, uint32_t bx, uint32_t by, uint32_t bz) {
    uint32_t v;
    asm("{\n\t.reg .u32 t;\n\t.reg .u64 p;\n\t"
        "mad.lo.u32 t, %2, %5, %1;\n\t"
        "mad.lo.u32 t, %3, %6, t;\n\t"
        "mov.b64 p, {t, %4};\n\t"
        "ld.global.nc.u8 %0, [p];\n\t}"
        : "=r"(v) : "r"(bx), "r"(by), "r"(bz), "r"(whi), "r"(wsx), "r"(wsxy));
    return v;
}

// 8x4x4 bricks: byte offset inside a brick = x0 x1 | y0 y1 | z0 || x2 | z1  (32-B sector = 4x4x2 voxels)
template <>
__device__ __forceinline__ uint32_t fetch_voxel<1>(const KArgs &a, uint32_t ix, uint32_t iy, uint32_t iz) {
    uint32_t inner = (ix & 3u) | ((iy & 3u) << 2) | ((iz & 1u) << 4) | ((ix & 4u) << 3) | ((iz & 2u) << 5);
    uint64_t brick = (uint64_t)(iz >> 2) * a.sxy + ((iy >> 2) * a.sx + (ix >> 3));
    return __ldg(a.vol + (brick * 128u + inner));
}

template <>
__device__ __forceinline__ uint32_t fetch_voxel<2>(const KArgs &a, uint32_t ix, uint32_t iy, uint32_t iz) {
    uint32_t r0, r1, r2, r3;
    asm("tex.3d.v4.u32.s32 {%0, %1, %2, %3}, [%4, {%5, %6, %7, %8}];"
        : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
        : "l"(a.tex), "r"(ix), "r"(iy), "r"(iz), "r"(0));
    return r0;
}

// Voxel index triple of a world-space point. CLAMP=false drops the upper clamp: legal only for the 3-D texture, which is
// padded by one replicated voxel per axis (index n == index n-1) and cannot fault; rays clipped to [-1,1]^3 never reach
// index n+1 (p <= 1 + a few ulp). The lower clamp is the saturation of cvt.rzi.u32.f32 either way.
struct VolScale { float hx, hy, hz; };  // dims * 0.5f held in registers

template <int SAMPLER, bool CLAMP>
__device__ __forceinline__ uint32_t sample_volume(const KArgs &a, const VolScale &v, float px, float py, float pz) {
    uint32_t iz, iy, ix;
    if (CLAMP) {
        iz = world_to_voxel(pz, v.hz, a.nz1); iy = world_to_voxel(py, v.hy, a.ny1); ix = world_to_voxel(px, v.hx, a.nx1);
    } else {
        iz = __float2uint_rz(fmul(fadd(pz, 1.0f), v.hz));
        iy = __float2uint_rz(fmul(fadd(py, 1.0f), v.hy));
        ix = __float2uint_rz(fmul(fadd(px, 1.0f), v.hx));
    }
    return (SAMPLER == 0 && !CLAMP) ? fetch_padded(a, ix, iy, iz) : fetch_voxel<SAMPLER>(a, ix, iy, iz);
}

// Loop-invariant kernel parameters are bounced through shared memory once per CTA and read back with volatile loads:
// that pins them in general registers for the whole march. Left alone, ptxas re-reads them from the constant bank
// through the uniform datapath (LDCU) on every iteration - one issue slot each in an issue-bound loop (profiles/).
struct Pinned { float hx, hy, hz, step, threshold, wmx, wmy, wmz; uint32_t whi, wsx, wsxy, zero_lim; };
struct Window { float mx, my, mz; uint32_t hi, sx, sxy; };  // the window constants held in registers

// Packed fp32 pairs (sm_100 add/mul.rn.f32x2 -> SASS FADD2 / FMUL2): two IEEE round-to-nearest results per issue slot,
// bit-identical to the scalar ops. ptxas 12.9 contracts a packed mul feeding a packed add into FFMA2 even with
// -fmad=false (tools/probe notes in DESIGN.md), so a packed product is only ever consumed by SCALAR adds here.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void unpack2(f32x2 v, float &lo, float &hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) { f32x2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) { f32x2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }

// LUT fetch through a 32-bit shared-window address kept in a register (avoids recomputing the window base per sample)
__device__ __forceinline__ float4 lut_fetch(uint32_t lut_addr, uint32_t s) {
    float4 c;
    asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(c.x), "=f"(c.y), "=f"(c.z), "=f"(c.w) : "r"(lut_addr + (s << 4)));
    return c;
}

__device__ __forceinline__ uint32_t div_block(const KArgs &a, uint32_t idx) {
    return a.bd_is_one ? idx : __umulhi(idx, a.bd_magic);
}

template <int V> struct FetchTag { static constexpr int value = V; };

// ---- leading empty-space loop (CPURenderer.cpp:18-25; sample_data_esl RaycasterBase.h:52-65; leap RaycasterBase.h:67-85) ----
// One iteration of the reference costs four IEEE divisions, three float->int and three int->float conversions and a floor:
// eleven operations of the conversion / special-function unit (a fraction of the fp32 rate), on a chain that an opaque
// transfer function runs about twice per sample (BASELINE configs[3]). The leap only enters the ray parameter as
//     k += floor(max(min(kx, ky, kz), 0) / step) * step,          kx = ((min_bound + block_size * bx) - p.x) / d.x, ...
// i.e. through the INTEGER m = floor(.): the quotients themselves never reach k. PREDICT = true therefore predicts m with
// reciprocal multiplications (y = RN(1 / d) per ray, one rounding each, relative error of x~ = dk / step against the
// reference's correctly rounded chain below 2^-21: three roundings of at most 2^-24 on the quotient that attains the minimum -
// signs and zeros are exact, so max(., 0) agrees - one on 1 / step, one on the product, one on each reference quotient) and
// ACCEPTS the prediction only when x~ is farther than x~ * 2^-19 from both neighbouring integers - then no integer can lie
// between x~ and the reference's x, and floor agrees. Otherwise (and for NaN / inf, which fail the test by construction) the
// lane redoes the iteration with the reference's divisions. Voxel indices come out of the mantissa of (v + 2^23) rounded
// toward zero (exact for -2^22 < v < 2^22; a slightly negative v gives -1, clamped to 0 like the saturating conversion), block
// indices go back to float through the same constant (exact: they are < 2^16). The host enables PREDICT only for finite
// views clipped to [-1, 1]^3 (every point of the loop lies in the cube up to rounding, so |v| <= n + 1), the ray enables it
// only when its non-zero direction components and the step lie in [2^-60, 2^60] (no reciprocal under- or overflows).
// What the loop leaves in k, p is bit for bit what the reference's loop leaves: tests/test_gpu_parity.py, test_gpu_full_size.py.
__device__ __forceinline__ uint32_t magic_voxel(float p, float half_n, uint32_t n_minus_1) {
    const int i = (int)(__float_as_uint(__fadd_rz(fmul(fadd(p, 1.0f), half_n), 8388608.0f)) - 0x4B000000u);
    return (uint32_t)min(max(i, 0), (int)n_minus_1);
}
__device__ __forceinline__ float small_uint_to_float(uint32_t b) { return fsub(__uint_as_float(0x4B000000u + b), 8388608.0f); }  // exact for b < 2^23
__device__ __forceinline__ bool recip_safe(float d) {  // 0, or 2^-60 <= |d| <= 2^60
    const uint32_t e = (__float_as_uint(d) >> 23) & 0xffu;
    return d == 0.0f || (e >= 127u - 60u && e <= 127u + 60u);
}

// what the predicted loop reads per iteration, staged in shared memory by the kernel and read back through a volatile
// pointer at loop entry: that pins the values in registers for the loop (ptxas otherwise re-loads every one of them from the
// constant bank in every iteration, one issue slot each) without keeping them alive across the march
struct EslPin { float bsx, bsy, bsz, hx, hy, hz; uint32_t nx1, ny1, nz1, bd_magic; };
__device__ __forceinline__ EslPin make_esl_pin(const KArgs &a) { return EslPin{a.bsx, a.bsy, a.bsz, a.hx, a.hy, a.hz, a.nx1, a.ny1, a.nz1, a.bd_magic}; }

// returns false when the ray leaves its clip interval without meeting a non-empty block (the reference returns without writing)
#ifndef VR5_ESL_INLINE
#define VR5_ESL_INLINE __forceinline__
#endif
template <bool COUNT, bool FMA>
__device__ VR5_ESL_INLINE bool skip_leading_empty_space(const KArgs &a, const EslPin *pin_shared, const uint32_t *esl_bits, float step, float ox, float oy,
                                                         float oz, float dx, float dy, float dz, float &k, float k_far, float &px, float &py, float &pz,
                                                         Counters &cnt) {
    auto point = [&](float o, float d, float kk) { return FMA ? __fmaf_rn(d, kk, o) : fadd(o, fmul(d, kk)); };
    const float hx = a.hx, hy = a.hy, hz = a.hz;
#ifndef VR5_NO_ESL_EARLYOUT
    {   // most rays of a dense volume start in a non-empty block: test the first point before setting the loop up
        const uint32_t bx = div_block(a, world_to_voxel(px, hx, a.nx1)), by = div_block(a, world_to_voxel(py, hy, a.ny1)), bz = div_block(a, world_to_voxel(pz, hz, a.nz1));
        if (k <= k_far && ((esl_bits[(bz * 32u + by) & 1023u] >> (bx & 31u)) & 1u) == 0u) return true;
    }
#endif
#ifndef VR5_NO_ESL_PREDICT
    if (a.esl == 2 && recip_safe(dx) && recip_safe(dy) && recip_safe(dz)) {
        // (a.esl == 2: the host has checked the view, the step and the block edge, and min_bound == -1: vr5_render.cu fill_args)
        const volatile EslPin *pv = pin_shared;
        const float bsx = pv->bsx, bsy = pv->bsy, bsz = pv->bsz, phx = pv->hx, phy = pv->hy, phz = pv->hz;
        const int nx1 = (int)pv->nx1, ny1 = (int)pv->ny1, nz1 = (int)pv->nz1;
        const uint32_t bd_magic = pv->bd_magic, esl_addr = (uint32_t)__cvta_generic_to_shared(esl_bits);
        // prediction constants per axis: quotient ~ n * y + c with (y, c) = (RN(1 / d), 0), or (0, 100) for a zero component
        const float yx = dx == 0.0f ? 0.0f : __frcp_rn(dx), yy = dy == 0.0f ? 0.0f : __frcp_rn(dy), yz = dz == 0.0f ? 0.0f : __frcp_rn(dz);
        const float cx = dx == 0.0f ? 100.0f : 0.0f, cy = dy == 0.0f ? 100.0f : 0.0f, cz = dz == 0.0f ? 100.0f : 0.0f;
        const float ys = __frcp_rn(step);
        // block index + (d > 0) goes back to float through the 2^23 bias: one integer add, one exact subtraction
        const uint32_t ux = 0x4B000000u + (dx > 0.0f ? 1u : 0u), uy = 0x4B000000u + (dy > 0.0f ? 1u : 0u), uz = 0x4B000000u + (dz > 0.0f ? 1u : 0u);
        auto voxel = [](float p, float half_n, int n1) {   // magic_voxel with pinned operands
            return (uint32_t)min(max((int)(__float_as_uint(__fadd_rz(fmul(fadd(p, 1.0f), half_n), 8388608.0f)) - 0x4B000000u), 0), n1);
        };
        while (k <= k_far) {
            const uint32_t bx = __umulhi(voxel(px, phx, nx1), bd_magic), by = __umulhi(voxel(py, phy, ny1), bd_magic), bz = __umulhi(voxel(pz, phz, nz1), bd_magic);
            uint32_t word;
            asm("ld.shared.u32 %0, [%1];" : "=r"(word) : "r"(esl_addr + (((bz << 5) + by) & 1023u) * 4u));
            if (((word >> (bx & 31u)) & 1u) == 0u) break;
            const float nx = fsub(fadd(-1.0f, fmul(bsx, fsub(__uint_as_float(bx + ux), 8388608.0f))), px);
            const float ny = fsub(fadd(-1.0f, fmul(bsy, fsub(__uint_as_float(by + uy), 8388608.0f))), py);
            const float nz = fsub(fadd(-1.0f, fmul(bsz, fsub(__uint_as_float(bz + uz), 8388608.0f))), pz);
            // prediction (not part of the bit-exact chain: any arithmetic will do, only the error bound above matters)
            const float x = fmaxf(fminf(fminf(__fmaf_rn(nx, yx, cx), __fmaf_rn(ny, yy, cy)), __fmaf_rn(nz, yz, cz)), 0.0f) * ys;
            float mf = fsub(__fadd_rz(x, 8388608.0f), 8388608.0f);  // floor(x) for 0 <= x < 2^23
            if (!(__fmaf_rn(x, 0x1p-19f, fabsf(fsub(fsub(x, mf), 0.5f))) <= 0.5f)) {
                // too close to an integer (or not finite): the reference's own arithmetic decides
                float kx = fdiv(nx, dx), ky = fdiv(ny, dy), kz = fdiv(nz, dz);
                if (dx == 0.0f) kx = 100.0f;
                if (dy == 0.0f) ky = 100.0f;
                if (dz == 0.0f) kz = 100.0f;
                mf = floorf(fdiv(sel_max(sel_min(sel_min(kx, ky), kz), 0.0f), step));
            }
            k = fadd(k, fmul(mf, step));
            if (COUNT) cnt.esl_iters++;
            k = fadd(k, step);
            px = point(ox, dx, k); py = point(oy, dy, k); pz = point(oz, dz, k);
        }
    } else
#endif
    {
        while (k <= k_far) {
            uint32_t bx = div_block(a, world_to_voxel(px, hx, a.nx1));
            uint32_t by = div_block(a, world_to_voxel(py, hy, a.ny1));
            uint32_t bz = div_block(a, world_to_voxel(pz, hz, a.nz1));
            const uint32_t word = esl_bits[(bz * 32u + by) & 1023u];
            if (((word >> (bx & 31u)) & 1u) == 0u) break;
            if (dx > 0.0f) bx++;
            if (dy > 0.0f) by++;
            if (dz > 0.0f) bz++;
            float kx = fdiv(fsub(fadd(a.mbx, fmul(a.bsx, (float)bx)), px), dx);
            float ky = fdiv(fsub(fadd(a.mby, fmul(a.bsy, (float)by)), py), dy);
            float kz = fdiv(fsub(fadd(a.mbz, fmul(a.bsz, (float)bz)), pz), dz);
            if (dx == 0.0f) kx = 100.0f;
            if (dy == 0.0f) ky = 100.0f;
            if (dz == 0.0f) kz = 100.0f;
            float dk = sel_min(kx, ky);
            dk = sel_min(dk, kz);
            dk = sel_max(dk, 0.0f);
            dk = fmul(floorf(fdiv(dk, step)), step);
            k = fadd(k, dk);
            if (COUNT) cnt.esl_iters++;
            k = fadd(k, step);
            px = point(ox, dx, k); py = point(oy, dy, k); pz = point(oz, dz, k);
        }
    }
    return !(k > k_far);
}

// ---- one ray (CPURenderer.cpp:11-41). Returns the packed RGBA8 pixel; 0 when the reference leaves it cleared. ----
// ZS ("zero skip"): the host has seen that the transfer function starts with all-zero entries (every shipped TF of the reference
// does: RaycasterBase.cpp:76-84 gives alpha 0 to the first 13 entries and update_transfer_fn pre-multiplies the colours) and
// that alpha stays inside [0, 1]. A sample that classifies to such an entry leaves the ray exactly as it found it - acc + 0 * (1 -
// acc.a) == acc for every finite acc, 0 <= 0.05 never shades, acc.a did not exceed the threshold a sample earlier - so its
// LUT fetch, compositing chain and termination test are skipped: only k += step, the next voxel request and k <= k_far remain.
template <int SAMPLER, bool COUNT, bool SHADE, bool CLAMP, bool ZS = false>
__device__ __forceinline__ uint32_t cast_ray(const KArgs &a, const RayView &v, const VolScale &vs, const Window &win, uint32_t lut_addr, const uint32_t *esl_bits,
                                             const EslPin *esl_pin, float step, float threshold, uint32_t zero_lim, int col, int row, Counters &cnt) {
    // View::get_ray (ViewBase.h:23-35)
    const float cx = (float)(col - a.half_w);
    const float cy = (float)(row - a.half_h);
    float ox, oy, oz, dx, dy, dz;
    if (a.perspective) {
        ox = v.ox; oy = v.oy; oz = v.oz;
        dx = fadd(fadd(v.dx, fmul(v.rx, cx)), fmul(v.ux, cy));
        dy = fadd(fadd(v.dy, fmul(v.ry, cx)), fmul(v.uy, cy));
        dz = fadd(fadd(v.dz, fmul(v.rz, cx)), fmul(v.uz, cy));
    } else {
        dx = v.dx; dy = v.dy; dz = v.dz;
        ox = fadd(fadd(v.ox, fmul(v.rx, cx)), fmul(v.ux, cy));
        oy = fadd(fadd(v.oy, fmul(v.ry, cx)), fmul(v.uy, cy));
        oz = fadd(fadd(v.oz, fmul(v.rz, cx)), fmul(v.uz, cy));
    }
    // Raycaster::intersect (RaycasterBase.h:32-42); the zero fix-up applies to a local copy only
    float k, k_far;
    {
        const float ex = dx == 0.0f ? 0.00001f : dx;
        const float ey = dy == 0.0f ? 0.00001f : dy;
        const float ez = dz == 0.0f ? 0.00001f : dz;
        const float k1x = fdiv(fsub(a.mbx, ox), ex), k1y = fdiv(fsub(a.mby, oy), ey), k1z = fdiv(fsub(a.mbz, oz), ez);
        const float k2x = fdiv(fsub(-a.mbx, ox), ex), k2y = fdiv(fsub(-a.mby, oy), ey), k2z = fdiv(fsub(-a.mbz, oz), ez);
        k = sel_max(sel_max(sel_min(k1x, k2x), sel_min(k1y, k2y)), sel_min(k1z, k2z));
        k_far = sel_min(sel_min(sel_max(k1x, k2x), sel_max(k1y, k2y)), sel_max(k1z, k2z));
        k = sel_max(k, 0.0f);
        if (!((k < k_far) && (k_far > 0.0f))) return 0u;
    }
    if (COUNT) cnt.rays_hit++;
    float px = fadd(ox, fmul(dx, k)), py = fadd(oy, fmul(dy, k)), pz = fadd(oz, fmul(dz, k));

    // leading empty-space loop
    if (a.esl && !skip_leading_empty_space<COUNT, false>(a, esl_pin, esl_bits, step, ox, oy, oz, dx, dy, dz, k, k_far, px, py, pz, cnt)) return 0u;
    if (COUNT) cnt.rays_accum++;

    // accumulation loop (CPURenderer.cpp:28-39), software pipelined: the voxel of sample i+1 is requested before sample i
    // is classified and composited, so its latency overlaps the LUT fetch and the compositing chain. The prefetch is
    // speculative (it may lie beyond k_far; clamped index or padded texture make it harmless) and changes no result.
    if (!(k <= k_far)) return 0u;  // NaN interval: the reference writes an all-zero pixel
    float ar = 0.0f, ag = 0.0f, ab = 0.0f, aa = 0.0f;
    // (dx, dy) as ONE value the compiler cannot see through: left to itself the allocator keeps the two scalars of the ray set-up
    // in unrelated registers and re-pairs them with two moves in every march step
    f32x2 dxy = pack2(dx, dy);
    asm volatile("" : "+l"(dxy));
    const f32x2 hxy = pack2(vs.hx, vs.hy), one2 = pack2(1.0f, 1.0f);
    constexpr int S0 = SAMPLER >= 3 ? 2 : SAMPLER;  // hybrids / window: first sample and shading samples come from the texture
    // voxel of the sample at ray parameter kq; FETCH selects the memory path (2 = texture, 0 = linear / padded linear, 4 = window,
    // 5 / 6 = power-of-two window / texture). Speculative: kq may lie up to kPrefetch steps past k_far.
    auto fetch = [&](auto fetch_tag, const float kq) -> uint32_t {
        constexpr int FETCH = decltype(fetch_tag)::value;
        // point: o + d*kq, then (p + 1) * (n/2) per axis; x,y packed, z scalar; packed products feed scalar adds
        float mx, my;
        unpack2(mul2(dxy, pack2(kq, kq)), mx, my);
        const float nx = fadd(ox, mx), ny = fadd(oy, my), nz = fadd(oz, fmul(dz, kq));
        uint32_t sn;
        if (FETCH == 5 || FETCH == 6 || FETCH == 7) {
            // power-of-two dims: (p + 1) * 2^e is exact, so floor((p + 1) * 2^e) + kP2Origin is read straight out of the
            // mantissa of (p + 1) + (2^(23-e) + kP2Origin * 2^-e) rounded toward zero - no multiply, no F2I.
            f32x2 bxy;
            asm("add.rz.f32x2 %0, %1, %2;" : "=l"(bxy) : "l"(add2(pack2(nx, ny), one2)), "l"(pack2(win.mx, win.my)));
            float bxf, byf;
            unpack2(bxy, bxf, byf);
            const uint32_t bx = __float_as_uint(bxf), by = __float_as_uint(byf), bz = __float_as_uint(__fadd_rz(fadd(nz, 1.0f), win.mz));
            if (FETCH == 7) {  // wide window: low word by * sx + bx (no wrap: build_p2), then bz * sxy + {low, hi} mod 2^64
                asm("{\n\t.reg .u32 t;\n\t.reg .u64 p;\n\tmad.lo.u32 t, %2, %4, %1;\n\tmov.b64 p, {t, %6};\n\tmad.wide.u32 p, %3, %5, p;\n\t"
                    "ld.global.nc.u8 %0, [p];\n\t}"
                    : "=r"(sn) : "r"(bx), "r"(by), "r"(bz), "r"(win.sx), "r"(win.sxy), "r"(win.hi));
            } else if (FETCH == 5) {  // window: the biased indices ARE the low address bits (the array is placed so that the biases cancel)
                // (a power-of-two row stride would make this a shift-add, but measured 10 % slower: cache-set aliasing)
                uint32_t lo;
                asm("{\n\t.reg .u32 t;\n\tmad.lo.u32 t, %2, %4, %1;\n\tmad.lo.u32 %0, %3, %5, t;\n\t}"
                    : "=r"(lo) : "r"(bx), "r"(by), "r"(bz), "r"(win.sx), "r"(win.sxy));
                // 64-bit window base (uniform) + 32-bit offset: one LDG with a uniform-register base, no register pair to assemble
                sn = __ldg(reinterpret_cast<const uint8_t *>((uint64_t)win.hi << 32) + lo);
            } else {           // texture: strip the exponent field, coordinate = index + kP2Origin
                uint32_t r0, r1, r2, r3;
                asm("tex.3d.v4.u32.s32 {%0, %1, %2, %3}, [%4, {%5, %6, %7, %8}];"
                    : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
                    : "l"(a.tex_p2), "r"(bx & 0x7fffffu), "r"(by & 0x7fffffu), "r"(bz & 0x7fffffu), "r"(0));
                sn = r0;
            }
        } else if (FETCH == 4) {
            float vx, vy;
            unpack2(mul2(add2(pack2(nx, ny), one2), hxy), vx, vy);
            const float vz = fmul(fadd(nz, 1.0f), vs.hz);
#ifndef VR5_NO_PACK_RZ
            float bxf, byf;
            { f32x2 r; asm("add.rz.f32x2 %0, %1, %2;" : "=l"(r) : "l"(pack2(vx, vy)), "l"(pack2(win.mx, win.my))); unpack2(r, bxf, byf); }
            sn = fetch_window(win.hi, win.sx, win.sxy, __float_as_uint(bxf), __float_as_uint(byf), __float_as_uint(__fadd_rz(vz, win.mz)));
#else
            sn = fetch_window(win.hi, win.sx, win.sxy, __float_as_uint(__fadd_rz(vx, win.mx)), __float_as_uint(__fadd_rz(vy, win.my)),
                              __float_as_uint(__fadd_rz(vz, win.mz)));
#endif
        } else {
            float vx, vy;
            unpack2(mul2(add2(pack2(nx, ny), one2), hxy), vx, vy);
            const float vz = fmul(fadd(nz, 1.0f), vs.hz);
            uint32_t ix = __float2uint_rz(vx), iy = __float2uint_rz(vy), iz = __float2uint_rz(vz);
            if (CLAMP) { ix = min(ix, a.nx1); iy = min(iy, a.ny1); iz = min(iz, a.nz1); }
            sn = (FETCH == 0 && !CLAMP) ? fetch_padded(a, ix, iy, iz) : fetch_voxel<(FETCH >= 4 ? 0 : FETCH)>(a, ix, iy, iz);
        }
        return sn;
    };
    // kPrefetch voxels are in flight per ray: slot j holds the voxel value ss[j] of the sample at ks[j]. A stage classifies and
    // composites the sample of one slot and refills the slot with the sample kPrefetch steps ahead, whose k follows that of the
    // other slot (the same sequential k += step as the reference). The loop is unrolled so that every slot keeps its registers.
    uint32_t ss[kPrefetch];
    float ks[kPrefetch];
    ss[0] = sample_volume<S0, CLAMP>(a, vs, px, py, pz);
    ks[0] = k;
#pragma unroll
    for (int j = 1; j < kPrefetch; j++) {
        ks[j] = fadd(ks[j - 1], step);
        ss[j] = fetch(FetchTag<S0 == 2 ? (is_p2(SAMPLER) ? 6 : 2) : S0>(), ks[j]);
    }
    // ---- leaps over transparent bricks (ZS kernels, a.brick_max != NULL) --------------------------------------------------------
    // The reference marches through empty space sample by sample once it has met its first non-empty block (its leaps,
    // RaycasterBase.h:67-85, only run in the leading loop). Under ZS such samples are no-ops, so a run of them may be left out
    // altogether IF the ray parameter afterwards is the one the reference's sequential k += step would hold. It is computable:
    // while k stays inside one binade [2^e, 2^(e+1)), k is a multiple of u = 2^(e-23) and RN(k + step) adds the SAME multiple
    // M * u every time (step = q * u + r rounds the same way for every k; excluded: r == u / 2 exactly, where ties-to-even
    // alternates) - so n steps on, the bit pattern of k is bits(k) + n * M, exactly (tools/check_kleap.c: 1.6e10 steps compared).
    // Which samples may go: all that provably read a voxel of a brick whose maximum is below zero_lim. From the transparent sample at
    // kc (voxel position v, inside brick b shrunk by a margin on every side) the ray stays inside the shrunk brick for t more units of k; samples
    // 1 .. n = floor(t / inc) lie inside the shrunk brick in exact arithmetic, and the reference's own fp32 position of each of them
    // is off by less than a quarter of the margin (host: fill_args derives the margin from |origin| and the dims), so each reads a voxel of b. The march continues with
    // the sample at k + n * inc, fetched and classified like any other. Returns 0: brick not transparent / cannot tell, 1: too close
    // to the brick's exit (or the binade's end) to gain anything, 2: kn = the ray parameter n >= 2 steps on.
    auto try_leap = [&](const float kc, const float ka, float &kn) -> int {
        const uint32_t kb = __float_as_uint(kc), kab = __float_as_uint(ka);
        const uint32_t M = kab - kb;                                       // ka = RN(kc + step): M * u in the binade of kc
        if ((kb ^ kab) >= 0x00800000u || (kb >> 23) < 30u || M == 0u) return 1;   // other binade (or kc tiny / zero): march
        const float inc = ka - kc;                                         // exact
        if (fabsf(step - inc) == __uint_as_float((kb & 0x7f800000u) - (24u << 23))) return 0;   // r == u / 2: no constant increment
        const float vx = (ox + dx * kc + 1.0f) * vs.hx, vy = (oy + dy * kc + 1.0f) * vs.hy, vz = (oz + dz * kc + 1.0f) * vs.hz;
        // (a position rounding has pushed outside the grid - negative, NaN: the conversion gives 0; beyond the last brick: clamped -
        // fails the margin test below against the brick it is assigned to)
        const uint32_t lb = a.brick_lb;
        const uint32_t bx = min((uint32_t)vx >> lb, a.brick_nx - 1u), by = min((uint32_t)vy >> lb, a.brick_ny - 1u), bz = min((uint32_t)vz >> lb, a.brick_nz - 1u);
        if (__ldg(a.brick_max + ((size_t)bz * (a.brick_nx * a.brick_ny) + by * a.brick_nx + bx)) >= zero_lim) return 0;
        const float B = (float)(1u << lb), lx = (float)(bx << lb), ly = (float)(by << lb), lz = (float)(bz << lb);
        const float sx = dx * vs.hx, sy = dy * vs.hy, sz = dz * vs.hz;     // voxels per unit of k
        auto rcp_approx = [](float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; };
        // time (in k, from kc) at which the ray leaves the SHRUNK brick per axis. The sample itself must lie inside the shrunk brick on
        // every side: a ray that runs almost parallel to a face it has just crossed may, in the reference's arithmetic, stay on the
        // other side of it for many more samples. A component of zero never leaves.
        const float mg = a.leap_margin, inf = __uint_as_float(0x7f800000u);
        float gap_min = 1.0f;
        auto exit_k = [&](float v, float lo, float sv, float r) -> float {
            const float hi_gap = (lo + B - mg) - v, lo_gap = v - (lo + mg);
            gap_min = fminf(gap_min, fminf(hi_gap, lo_gap));
            return sv != 0.0f ? (sv > 0.0f ? hi_gap : lo_gap) * r : inf;
        };
        const float rx = rcp_approx(fabsf(sx)), ry = rcp_approx(fabsf(sy)), rz = rcp_approx(fabsf(sz));
        float tx = exit_k(vx, lx, sx, rx), ty = exit_k(vy, ly, sy, ry), tz = exit_k(vz, lz, sz, rz);
        if (!(gap_min >= 0.0f)) return 1;                                  // inside a margin (or NaN): march on
        // From brick to brick (3-D DDA on the shrunk exit times) while the next brick is transparent too and the crossing is CLEAN: the
        // ray must be back inside the next brick's shrunk interior (2 * mg further along the crossing axis) before it leaves the
        // shrunk range of either other axis - then every sample of the crossing lies within mg of that one face only and reads a
        // voxel of one of the two bricks. t1 = exit time of the last brick accepted.
        const uint32_t nbx = a.brick_nx, nby = a.brick_ny, nbz = a.brick_nz;
        const uint8_t *bm = a.brick_max;
        uint32_t ix = bx, iy = by, iz = bz;
        const float t_max = k_far - kc;
        float t1 = 0.0f;
#pragma unroll 1
        for (int hop = 0; hop < 48; hop++) {
            t1 = fminf(tx, fminf(ty, tz));
            if (!(t1 < t_max)) break;                                      // (also NaN)
            if (tx <= ty && tx <= tz) {
                if (!(t1 + 2.0f * mg * rx <= fminf(ty, tz))) break;
                ix += sx > 0.0f ? 1u : 0xffffffffu;
                if (ix >= nbx || __ldg(bm + ((size_t)iz * (nbx * nby) + iy * nbx + ix)) >= zero_lim) break;
                tx += B * rx;
            } else if (ty <= tz) {
                if (!(t1 + 2.0f * mg * ry <= fminf(tx, tz))) break;
                iy += sy > 0.0f ? 1u : 0xffffffffu;
                if (iy >= nby || __ldg(bm + ((size_t)iz * (nbx * nby) + iy * nbx + ix)) >= zero_lim) break;
                ty += B * ry;
            } else {
                if (!(t1 + 2.0f * mg * rz <= fminf(tx, ty))) break;
                iz += sz > 0.0f ? 1u : 0xffffffffu;
                if (iz >= nbz || __ldg(bm + ((size_t)iz * (nbx * nby) + iy * nbx + ix)) >= zero_lim) break;
                tz += B * rz;
            }
        }
        if (!(t1 >= 0.0f)) return 1;
        t1 = fminf(t1, t_max);   // never past the ray's own end (a brick may overhang the grid): the landing sample is at most one step beyond k_far
        // samples 1 .. n lie inside the bricks accepted; the march continues with sample n + 1 - inside the binade: bits(k) + (n + 1) * M
        // must stay below the next power of two (two steps of room)
        float nf = floorf(t1 * rcp_approx(inc) * 0.99999f);
        const float nb = floorf((float)(0x007fffffu - (kb & 0x007fffffu)) * rcp_approx((float)M) * 0.99999f) - 3.0f;
        nf = fminf(nf, nb);
        if (!(nf >= 2.0f)) return 1;                                        // (false for NaN)
        nf += 1.0f;
        const uint32_t knb = kb + (uint32_t)nf * M;
        if ((knb ^ kb) >= 0x00800000u) return 1;
        kn = __uint_as_float(knb);
        if (a.leap_stats) {   // VR5_LEAP_STATS (tests, bench): two atomics on one address per leap - never in a timed run
            atomicAdd(a.leap_stats, 1ull);
            atomicAdd(a.leap_stats + 1, (unsigned long long)nf - 1ull);   // samples 1 .. n are never fetched (nf = n + 1 here)
        }
        return 2;
    };
    // Leaps are tried in every SECOND stage of the unrolled march only, by all lanes of a warp whose sample is transparent at once
    // (the lanes of a warp run the same stage at the same time: a warp-wide attempt costs what one lane's does) - unless the lane is
    // backing off: `cool` = attempts to sit out (no brick table: all of them). A ray tries its first leap once it has marched through
    // eight transparent samples in a row (four attempts sat out; every composited sample starts the count again): in dense data, where
    // transparent samples come in short runs, an attempt costs more than the samples it could leave out (the reference's own benchmark
    // sweep with its default table measured 2-5 % slower with unconditional attempts), in empty space the eight samples are noise.
    uint32_t cool = (ZS && a.brick_max) ? 4u : 0x7fffffffu;
    // one march step: returns true when the ray is finished
    auto stage = [&](auto fetch_tag, auto leap_tag, uint32_t &s_cur, float &k_cur, const float k_last, const float &k_next) -> bool {
        float ka = fadd(k_last, step);
        uint32_t sn = fetch(fetch_tag, ka);
        if (COUNT) cnt.samples++;
        bool ended = false;
        if (ZS && s_cur < zero_lim) {     // ZS: a transparent sample has nothing to composite (see ZS above)
            if (kPrefetch == 1 && decltype(leap_tag)::value) {
                if (cool == 0u) {
                    float kn;
                    const int got = try_leap(k_cur, ka, kn);
                    if (got == 2) { ka = kn; sn = fetch(fetch_tag, kn); }
                    else if (got == 0) cool = 3u;   // not a transparent brick: sit out three attempts (six samples)
                } else cool--;
            }
        } else {
            float4 c = lut_fetch(lut_addr, s_cur);  // lut[s] == transfer_fn[s / TF_RATIO]
            if (SHADE && c.w > a.shade_alpha) {  // Raycaster::shade (RaycasterBase.h:87-98)
                const float px = fadd(ox, fmul(dx, k_cur)), py = fadd(oy, fmul(dy, k_cur)), pz = fadd(oz, fmul(dz, k_cur));  // the sample's point
                const float lvx = fsub(a.lx, px), lvy = fsub(a.ly, py), lvz = fsub(a.lz, pz);
                const float len2 = fadd(fadd(fmul(lvx, lvx), fmul(lvy, lvy)), fmul(lvz, lvz));
                const float inv = fdiv(1.0f, __fsqrt_rn(len2));
                const float qx = fadd(px, fmul(fmul(lvx, inv), 0.01f));
                const float qy = fadd(py, fmul(fmul(lvy, inv), 0.01f));
                const float qz = fadd(pz, fmul(fmul(lvz, inv), 0.01f));
                const uint32_t sl = sample_volume<S0, true>(a, vs, qx, qy, qz);  // may leave the cube: always clamped
                const float diffuse = fmul(fsub(fdiv((float)sl, 255.0f), fdiv((float)s_cur, 255.0f)), a.kd);
                c.x = fadd(c.x, diffuse); c.y = fadd(c.y, diffuse); c.z = fadd(c.z, diffuse);
                if (COUNT) cnt.shaded++;
            }
            // acc = acc + c * (1 - acc.a): packed products, scalar accumulation
            const float t = fsub(1.0f, aa);
            const f32x2 t2 = pack2(t, t);
            float qr, qg, qb, qa;
            unpack2(mul2(pack2(c.x, c.y), t2), qr, qg);
            unpack2(mul2(pack2(c.z, c.w), t2), qb, qa);
            ar = fadd(ar, qr); ag = fadd(ag, qg); ab = fadd(ab, qb); aa = fadd(aa, qa);
            ended = aa > threshold;                    // early ray termination, strict >
            if (ZS && kPrefetch == 1) cool = max(cool, 4u);   // a visible sample: the next leap waits for eight transparent ones again
        }
        s_cur = sn; k_cur = ka;
        return ended || !((kPrefetch > 1 ? k_next : ka) <= k_far);  // while (k <= k_far) for the NEXT sample
    };
    // stage u of the unrolled loop works on slot u % kPrefetch; its fetch path follows the sampler's pattern
#define VR5_STAGE(TAG, U) stage(FetchTag<TAG>(), FetchTag<(U) & 1>(), ss[(U) % kPrefetch], ks[(U) % kPrefetch], ks[((U) + kPrefetch - 1) % kPrefetch], ks[((U) + 1) % kPrefetch])
    if (is_p2(SAMPLER)) {
        if (p2_mode(SAMPLER) == 0) {
            for (;;) {
                if (VR5_STAGE(6, 0)) break;
                if (VR5_STAGE(5, 1)) break;
            }
        } else if (p2_mode(SAMPLER) == 1) {
            for (;;) {
                if (VR5_STAGE(5, 0)) break;
                if (VR5_STAGE(5, 1)) break;
            }
        } else if (SAMPLER == kP2Wide) {
            for (;;) {
                if (VR5_STAGE(6, 0)) break;
                if (VR5_STAGE(7, 1)) break;
            }
        } else {
            for (;;) {
                if (VR5_STAGE(6, 0)) break;
                if (VR5_STAGE(5, 1)) break;
                if (VR5_STAGE(5, 2)) break;
                if (VR5_STAGE(6, 3)) break;
                if (VR5_STAGE(5, 4)) break;
                if (VR5_STAGE(5, 5)) break;
            }
        }
    } else if (SAMPLER == 3) {
        // hybrid: prefetches alternate between the texture pipe and the LSU pipe (padded linear array). A warp-wide TLD
        // costs 8 TEX wavefronts, which made the texture pipe the limiter (79 % busy, profiles/r01_v3_*); 1:1 measured
        // best against 2:1 and 1:2 (0.647 / 0.706 / 0.716 ms per C3 frame).
        for (;;) {
            if (VR5_STAGE(2, 0)) break;
            if (VR5_STAGE(0, 1)) break;
        }
    } else if (SAMPLER == 4) {  // texture : window, 1 : 1
        for (;;) {
            if (VR5_STAGE(2, 0)) break;
            if (VR5_STAGE(4, 1)) break;
        }
    } else if (SAMPLER == 5) {  // window only
        for (;;) {
            if (VR5_STAGE(4, 0)) break;
            if (VR5_STAGE(4, 1)) break;
        }
    } else if (SAMPLER == 6) {  // texture : window, 1 : 2
        for (;;) {
            if (VR5_STAGE(2, 0)) break;
            if (VR5_STAGE(4, 1)) break;
            if (VR5_STAGE(4, 2)) break;
            if (VR5_STAGE(2, 3)) break;
            if (VR5_STAGE(4, 4)) break;
            if (VR5_STAGE(4, 5)) break;
        }
    } else {
        for (;;) {
            if (VR5_STAGE(S0, 0)) break;
            if (VR5_STAGE(S0, 1)) break;
        }
    }
#undef VR5_STAGE
    // Raycaster::write_color (RaycasterBase.h:44-50)
    return trunc_clamp(fmul(ar, 256.0f), 255u) | (trunc_clamp(fmul(ag, 256.0f), 255u) << 8) |
           (trunc_clamp(fmul(ab, 256.0f), 255u) << 16) | (trunc_clamp(fmul(aa, 256.0f), 255u) << 24);
}

// Queue slot -> tile. Slots are handed out in groups of 4x8 tiles (32x32 px; neighbouring slots touch neighbouring voxels).
// With a.group_order the groups are visited in the host's order (build_group_order in vr5_render.cu: opt-in orders that end
// the launch on short rays; the default row-major sweep passes NULL).
__device__ __forceinline__ void tile_of_slot(const KArgs &a, uint32_t q, uint32_t &tx, uint32_t &tyl) {
    uint32_t group = q >> (kGroupLog2X + kGroupLog2Y);
    const uint32_t within = q & ((1u << (kGroupLog2X + kGroupLog2Y)) - 1u);
    if (a.group_order) group = __ldg(a.group_order + group);
    const uint32_t gy = a.groups_x_magic ? __umulhi(group, a.groups_x_magic) : group / a.groups_x;   // exact for group < 2^16 (fill_args)
    tx = ((group - gy * a.groups_x) << kGroupLog2X) + (within & ((1u << kGroupLog2X) - 1u));
    tyl = (gy << kGroupLog2Y) + (within >> kGroupLog2X);
}

// ---- persistent warps + atomic tile queue (shared by cast_kernel and tri_kernel) -------------------------------------------
__device__ __forceinline__ uint32_t load_acquire_sys_u32(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// view of the tile's frame: kernel parameters for a single frame; for a sequence the warp's shared-memory copy of the frame's
// record (refreshed when the warp moves on to another frame: run_tile_queue). Once per tile; nothing of it stays live across
// the march.
__device__ __forceinline__ RayView load_view(const KArgs &a, const FrameView *cached) {
    if (a.frames) {
        const float4 *p = reinterpret_cast<const float4 *>(cached);
        const float4 x = p[0], y = p[1], z = p[2];
        return RayView{x.w, y.x, y.y, x.x, x.y, x.z, y.z, y.w, z.x, z.y, z.z, z.w};
    }
    return RayView{a.ox, a.oy, a.oz, a.dx, a.dy, a.dz, a.rx, a.ry, a.rz, a.ux, a.uy, a.uz};
}

// Memory ordering of the frame-ring hand-shake. A warp's pixel stores must be visible to whoever observes the rank's arrival:
//   lanes store pixels -> warp / CTA barrier -> ONE thread: fence, then atomicAdd on the frame's slot counter (device scope)
//   the thread whose add completes the frame: fence (system scope), then the system-scope arrival.
// The fences are cumulative, so the one thread's fence covers the other lanes' stores it is ordered behind by the barrier. For a
// ring in a PEER GPU's memory (a.ring_remote) the first fence is a system-scope one as well: the pixels travel over NVLink and a
// device-scope fence says nothing about their arrival there.
__device__ __forceinline__ void ring_fence(bool system_scope) {
    if (system_scope) asm volatile("fence.acq_rel.sys;" ::: "memory");
    else asm volatile("fence.acq_rel.gpu;" ::: "memory");
}

// `count` queue slots of frame f are finished (pixels stored, barrier passed): count them; the add that completes the frame
// publishes the rank's arrival in the frame ring. Called by one thread.
__device__ __forceinline__ void frame_slots_done(const KArgs &a, uint32_t f, uint32_t count) {
    ring_fence(a.ring_remote != 0);
    if (atomicAdd(a.frame_done + f, count) + count == a.n_queue) {
        const FrameView &fv = a.frames[f];
        if (fv.arrive) {
            ring_fence(true);
            atomicAdd_system(fv.arrive, 1u);
        }
    }
}

// The slot of frame f may be written once the consumer has released the frame that used it before. Frames are released in
// order, so "highest frame known to be writable" answers every later question about earlier frames. The thread that fetches
// from the queue also asks: with per-CTA reads the ONE fetching thread, between the two barriers of the fetch (the other 255
// threads sit in the barrier: no lock, nobody spinning); with per-warp reads lane 0 of the warp. The release word may live in
// a PEER GPU's memory, so the question is asked in two steps:
//   a.ring_sync[0]    <- device-memory copy of the answer for this launch; every CTA reads it (one L2 access per frame)
//   the peer's word   <- ONE thread per GPU at a time (a.ring_sync[1] is its lock), pause backing off to ~8 us between polls
// History: every warp polling the peer word gave NVLink errors with a two-slot ring; one lane per warp behind a shared-memory
// lock with the other lanes spinning cost ~40 us per frame although the slot was free every time. Loads are RELAXED: nothing
// is read from the slot behind the flag, only written, and a store is issued after the load has returned.
// Bounded: after 4 s the wait gives up, raises the ring's status flag and the frame is written anyway.
__device__ __forceinline__ uint32_t load_relaxed_sys_u32(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ bool frame_known(uint32_t known, uint32_t f) { return known != 0xffffffffu && known >= f; }

// called by one thread per CTA (per-CTA queue reads) or per warp (per-warp reads); *s_ok is the CTA's cached copy of the answer.
// Whoever holds the per-GPU lock does not ask about ITS frame but advances the common frontier by one: it polls the release
// word of the first frame not yet known to be writable (releases come in order: if that one is not released, nothing later
// is). So every poll works for the waiter that is furthest behind, and a thread whose frame has long been released cannot be
// starved by threads that poll for later frames - the failure mode of the first versions (a shared-memory lock whose holders
// each asked about their own frame: the one lane that could have made progress never got the lock, its tile never finished,
// its frame never arrived, nobody's release ever came - until the 4-s bound).
__device__ __forceinline__ void wait_frame_writable(const KArgs &a, uint32_t f, volatile uint32_t *s_ok) {
    if (frame_known(*s_ok, f)) return;
    unsigned long long t0 = 0;
    unsigned pause = 256;
    for (;;) {
        const uint32_t known = load_relaxed_sys_u32(a.ring_sync);
        if (frame_known(known, f)) break;
        bool advanced = false;
        if (atomicCAS(a.ring_sync + 1, 0u, 1u) == 0u) {                                    // this thread asks for the GPU
            const uint32_t now = load_relaxed_sys_u32(a.ring_sync);                         // (the frontier may have moved meanwhile)
            const uint32_t next = now == 0xffffffffu ? 0u : now + 1u;                       // <= f: f is not known yet
            if (next <= f) {
                const FrameView &fv = a.frames[next];
                const uint32_t *consumed = fv.consumed;                                     // NULL: first use of the slot
                if (!consumed || (int32_t)(load_relaxed_sys_u32(consumed) - fv.consumed_target) >= 0) { atomicExch(a.ring_sync, next); advanced = true; }
            } else {
                advanced = true;
            }
            __threadfence();
            atomicExch(a.ring_sync + 1, 0u);
        }
        if (advanced) { pause = 256; continue; }
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        if (t0 == 0) t0 = t;
        if (t - t0 > 4000000000ull) { *a.timed_out = 1u; break; }
        __nanosleep(pause);
        if (pause < 8192) pause <<= 1;
    }
    if (!frame_known(*s_ok, f)) *s_ok = f;
}

// ray(view, col, row) -> RGBA8 pixel. COUNT: instrumented pass, nothing is stored.
template <bool COUNT, class RayFn>
__device__ __forceinline__ void run_tile_queue(const KArgs &a, RayFn &&ray) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __shared__ unsigned s_q;
    __shared__ uint32_t s_ok;          // sequence hand-shake: highest frame whose slot is known to be writable (0xffffffff: none yet)
    // finished queue slots are counted per frame, and a warp (CTA with CTA-wide fetch) adds its share to the frame's counter
    // ONCE, when it moves on to another frame or leaves: [w][0] = frame being counted (0xffffffff: none), [w][1] = slots
    __shared__ uint32_t s_acc[kCtaThreads / 32][2];
    // sequence: the record of the frame a warp is working on (view, buffer, hand-shake words) and which frame that is
    __shared__ __align__(16) FrameView s_view[kCtaThreads / 32];
    __shared__ uint32_t s_view_f[kCtaThreads / 32];
    if (a.frames) {
        if (lane == 0) s_view_f[warp] = 0xffffffffu;
        __syncwarp();
    }
    const bool shake = !COUNT && a.frames && a.frame_done;
    if (shake) {
        if (threadIdx.x == 0) s_ok = 0xffffffffu;
        if (lane == 0) { s_acc[warp][0] = 0xffffffffu; s_acc[warp][1] = 0u; }
        __syncthreads();
    }
    const uint32_t n_total = a.n_queue * a.n_frames;
    // the slots this warp (lane 0) / CTA (thread 0) has just finished go to the account; the account is flushed when the next
    // slot belongs to another frame or the queue is exhausted
    auto account = [&](uint32_t finished, uint32_t q_next) {
        uint32_t f_acc = s_acc[warp][0], n_acc = s_acc[warp][1] + (f_acc != 0xffffffffu ? finished : 0u);
        uint32_t f_next = f_acc;   // same frame unless the slot lies outside it (then: one division per frame)
        if (q_next >= n_total) f_next = 0xffffffffu;
        else if (f_acc == 0xffffffffu || q_next - f_acc * a.n_queue >= a.n_queue) f_next = q_next / a.n_queue;
        if (f_acc != 0xffffffffu && f_acc != f_next) { frame_slots_done(a, f_acc, n_acc); n_acc = 0u; }
        s_acc[warp][0] = f_next; s_acc[warp][1] = n_acc;
    };
    for (;;) {
        unsigned q = 0;
        if (a.cta_fetch) {
            // the CTA takes 8 consecutive queue slots at once = a 32x8-pixel block (queue groups are 4 tiles wide): its 8
            // warps march through neighbouring voxels in step and share L1 lines (C3: 0.606 -> 0.596 ms). The price is a
            // barrier per block, so the host only asks for it when rays of neighbouring tiles end together (no opaque TF
            // entry: C4 measured 5 % slower with it). Same counter arithmetic as one slot per warp.
            __syncthreads();
            if (threadIdx.x == 0) {
                s_q = atomicAdd(a.queue, (unsigned)(kCtaThreads / 32)) - a.queue_base;
                if (shake) {
                    // FIRST count the block just finished (all warps have stored their pixels: they are in the barrier above) -
                    // a CTA that waited for a slot while holding back its share of an earlier frame would keep that frame from
                    // ever arriving, and with it the release it is waiting for
                    account((uint32_t)(kCtaThreads / 32), s_q);
                    if (s_q < n_total) {        // then: the new block's frame must be writable before anybody marches (almost always: long since)
                        uint32_t fb = s_ok;     // same frame as the last block unless the slot lies outside it
                        if (fb == 0xffffffffu || s_q - fb * a.n_queue >= a.n_queue) fb = s_q / a.n_queue;
                        wait_frame_writable(a, fb, &s_ok);
                    }
                }
            }
            __syncthreads();
            q = s_q + (threadIdx.x >> 5);  // s_q is a multiple of 8 and n_queue of 32: the exit below is uniform over the CTA
        } else {
            if (shake) __syncwarp();       // the lanes' stores of the previous tile are ordered before lane 0's accounting
            if (lane == 0) {
                q = atomicAdd(a.queue, 1u) - a.queue_base;  // wrap-safe
                if (shake) {               // count the finished tile FIRST, then make sure the new tile's frame may be written
                    account(1u, q);
                    if (q < n_total) {
                        uint32_t fw = s_acc[warp][0];   // account() has just stored the new slot's frame there
                        wait_frame_writable(a, fw, &s_ok);
                    }
                }
            }
            q = __shfl_sync(0xffffffffu, q, 0);
        }
        if (q >= n_total) break;
        uint32_t f = 0, qi = q;
        if (a.frames) {
            // the frame of the slot: almost always the one the warp is in already (one multiply and compare); a division once per frame
            f = s_view_f[warp];
            qi = q - f * a.n_queue;
            if (f == 0xffffffffu || qi >= a.n_queue) { f = q / a.n_queue; qi = q - f * a.n_queue; }   // a block of 8 never straddles frames (n_queue % 32 == 0)
            if (s_view_f[warp] != f) {                   // first tile of this warp in frame f: fetch the frame's record (80 bytes)
                __syncwarp();
                if (lane < (int)(sizeof(FrameView) / 4)) reinterpret_cast<uint32_t *>(&s_view[warp])[lane] = __ldg(reinterpret_cast<const uint32_t *>(a.frames + f) + lane);
                if (lane == 0) s_view_f[warp] = f;
                __syncwarp();
            }
        }
        uint32_t rgba = 0u;
        {
            uint32_t tx, tyl;
            tile_of_slot(a, qi, tx, tyl);
            const int col = (int)tx * kTileW + (lane & 7);
            const bool fb = a.frames && a.bands;   // the frame's own band (uniform)
            const uint32_t strip0 = fb ? s_view[warp].first_strip : (uint32_t)a.rank, strips_here = fb ? s_view[warp].n_strips : a.tile_rows_local;
            const int row = (int)(tyl * a.world + strip0) * kTileH + (lane >> 3);
            if (tx < a.tiles_x && tyl < strips_here && col < a.width && row < a.height)  // ragged edges: lanes outside the image idle
                rgba = ray(load_view(a, &s_view[warp]), col, row);
        }
        if (!COUNT) {
            // Only the queue slot survives the march: the pixel's place is worked out AGAIN from it (a few dozen instructions per
            // tile) instead of keeping tile coordinates, frame index and buffer address in registers through the march loop,
            // where every register counts (40 per thread for six resident CTAs per SM).
            asm volatile("" : "+r"(qi));
            uint32_t tx, tyl;
            tile_of_slot(a, qi, tx, tyl);
            const int lane2 = threadIdx.x & 31;
            const int col = (int)tx * kTileW + (lane2 & 7);
            const bool fb = a.frames && a.bands;
            const uint32_t strip0 = fb ? s_view[threadIdx.x >> 5].first_strip : (uint32_t)a.rank, strips_here = fb ? s_view[threadIdx.x >> 5].n_strips : a.tile_rows_local;
            const int row = (int)(tyl * a.world + strip0) * kTileH + (lane2 >> 3);
            if (tx < a.tiles_x && tyl < strips_here && col < a.width && row < a.height) {
                uchar4 *out = a.frames ? s_view[threadIdx.x >> 5].out : a.out;
                const int out_row = a.packed ? (int)tyl * kTileH + (lane2 >> 3) : row;
                reinterpret_cast<uint32_t *>(out)[(size_t)out_row * a.width + col] = rgba;
            }
        }
    }
}

// ---- the kernel ---------------------------------------------------------------------------------------------------------
// SHADE=false is chosen by the host when no sample can take the shading path (light_kd <= 0.01 or no TF entry with
// alpha > 0.05), CLAMP=false when the sampler is the padded 3-D texture and min_bound == (-1,-1,-1).
template <int SAMPLER, bool COUNT, bool SHADE, bool CLAMP, bool ZS = false>
__global__ void __launch_bounds__(kCtaThreads, VR5_MIN_CTAS) cast_kernel(const __grid_constant__ KArgs a) {
    __shared__ float4 s_lut[256];
    __shared__ uint32_t s_esl[1024];
    __shared__ Pinned s_pinned;
    __shared__ EslPin s_esl_pin;
    for (int i = threadIdx.x; i < 256; i += kCtaThreads) s_lut[i] = a.tf[i >> 1];
    for (int i = threadIdx.x; i < 1024; i += kCtaThreads) s_esl[i] = a.esl_bits[i];
    if (threadIdx.x == 32) s_esl_pin = make_esl_pin(a);
    if (threadIdx.x == 0) s_pinned = Pinned{a.hx, a.hy, a.hz, a.step, a.threshold, a.wmx, a.wmy, a.wmz, a.whi, a.wsx, a.wsxy, a.zero_lim};
    __syncthreads();

    const int lane = threadIdx.x & 31;
    const volatile Pinned *pinned = &s_pinned;
    const VolScale vs = {pinned->hx, pinned->hy, pinned->hz};
    const float step = pinned->step, threshold = pinned->threshold;
    const uint32_t zero_lim = ZS ? pinned->zero_lim : 0u;
    Window win = {0.0f, 0.0f, 0.0f, 0u, 0u, 0u};
    if (SAMPLER >= 4) win = Window{pinned->wmx, pinned->wmy, pinned->wmz, pinned->whi, pinned->wsx, pinned->wsxy};
    const uint32_t lut_addr = (uint32_t)__cvta_generic_to_shared(s_lut);
    Counters cnt = {0, 0, 0, 0, 0};
    run_tile_queue<COUNT>(a, [&](const RayView &v, int col, int row) {
        return cast_ray<SAMPLER, COUNT, SHADE, CLAMP, ZS>(a, v, vs, win, lut_addr, s_esl, &s_esl_pin, step, threshold, zero_lim, col, row, cnt);
    });
    if (COUNT) {
        unsigned long long v[5] = {cnt.rays_hit, cnt.rays_accum, cnt.esl_iters, cnt.samples, cnt.shaded};
        unsigned long long *dst = reinterpret_cast<unsigned long long *>(a.counters);
#pragma unroll
        for (int i = 0; i < 5; i++) {
            unsigned long long x = v[i];
            for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
            if (lane == 0 && x) atomicAdd(dst + i, x);
        }
    }
}

}  // namespace vr5
