// vr5_tri.cuh - trilinear sampler policy (GPURenderer4 semantics, GPURenderer4.cu:33-97) as software
// fp32 arithmetic: every operation of the CPU checker's trilinear restatement (test infrastructure, never linked here), in the same order, through the exact
// __f*_rn helpers of vr5_cast.cuh. Compared bit for bit with that restatement, never with the nearest-neighbour parity bar.
//
// Memory layout ("cells"): one 8-byte record per cell (i0x, i0y, i0z) in [-1, n-1]^3 holding the eight voxels the cell's
// trilinear footprint needs, neighbours outside the grid already clamped. A sample is ONE aligned 8-byte load - exactly the
// 8 algorithmic bytes per sample of SURVEY 8(d) - instead of eight byte gathers.
#pragma once

#include "vr5_cast.cuh"

namespace vr5 {

constexpr int kTriGuard = 8;  // guard cells on every side of the record array (unclamped march: host check tri_reach_ok)
constexpr float kTriMagic = 12582912.0f;      // 1.5 * 2^23: floor(t) == bits(RD(t + magic)) - bits(magic) for |t| < 2^22
constexpr uint32_t kTriMagicBits = 0x4B400000u;

struct TriArgs {
    const uint2 *cells;      // (nx+1+2G)(ny+1+2G)(nz+1+2G) records, x fastest; record j holds cell i0 = j - 1 - G (clamped)
    uint32_t cx, cy;         // records per row / rows per slice
    float hfx, hfy, hfz;     // dims * 0.5f: ((p + 1) * 0.5f) * n == (p + 1) * (0.5f * n) exactly (power-of-two scaling)
    uint32_t jx1, jy1, jz1;  // last record index per axis (n + 2G)
    // unclamped march (FAST): the array lies inside one 4-GiB address window and the low 32 address bits of the record of a
    // point are bz*s2 + by*s1 + bx*8 + c0 (mod 2^32), b = bit pattern of RD(t + magic) per axis
    uint32_t hi, c0, s1, s2;
    // ZS kernels: a sample whose biased TF position floor, bits(RD(u + magic)), lies below this classifies to a pair of all-zero TF
    // entries (both ends of its interpolation interval lie inside the table's all-zero prefix): it has nothing to composite
    uint32_t zero_um;
};

// the march loop's uniform operands. With VR5_TRI_PIN they are read once per thread through a volatile shared copy, which
// pins them in registers; otherwise ptxas re-loads them from the constant bank (LDCU) in every iteration.
#ifndef VR5_TRI_PIN
#define VR5_TRI_PIN 1
#endif
struct TriRegs { float step, threshold, hfx, hfy, hfz; uint32_t s1, s2, c0, hi, pair_c; float bfx, bfy, bfz; };  // bf = hf - 0.5 (contracted policy)

// one axis: texel position t = ((p + 1) * 0.5) * n - 0.5, cell floor(t), weight t - floor(t). floor() comes from an add
// rounded toward -inf against 1.5 * 2^23, whose bit pattern also carries the integer. Returns the biased bits.
__device__ __forceinline__ uint32_t tri_axis(float p, float hf, float &w) {
    const float t = fsub(fmul(fadd(p, 1.0f), hf), 0.5f);
    const float tm = __fadd_rd(t, kTriMagic);
    w = fsub(t, fsub(tm, kTriMagic));  // tm - magic is exact
    asm volatile("" : "+f"(w));        // keep the weight in its register: ptxas otherwise recomputes the whole axis after the loop body
    return __float_as_uint(tm);
}

// byte i of a record word as the float 2^23 + byte (one PRMT, no conversion unit: I2F runs at 1/8 rate)
__device__ __forceinline__ float byte_biased(uint32_t word, int i) {
    return __uint_as_float(__byte_perm(word, 0x4B000000u, 0x7650u + (uint32_t)i));
}

__device__ __forceinline__ float lerp_rn(float a, float b, float w) { return fadd(a, fmul(w, fsub(b, a))); }
// a + w * (b - a) with a and b still biased by 2^23: the difference of the biased values IS b - a (exact, both in
// [2^23, 2^23 + 255]), so only a needs the exact un-biasing subtraction
__device__ __forceinline__ float lerp_biased(float ab, float bb, float w) { return fadd(fsub(ab, 8388608.0f), fmul(w, fsub(bb, ab))); }

// x / 255 correctly rounded without the division sequence: q0 = RN(x * r), rem = x - q0 * 255 (exact in an FMA), q = RN(q0 +
// rem * r) with r = RN(1 / 255). Equal to __fdiv_rn(x, 255.0f) for EVERY float in [0, 255] - checked exhaustively over all
// 1 132 396 545 bit patterns (tools/check_div255.c; tests/test_trilinear.py runs a strided sample of it).
__device__ __forceinline__ float div255(float x) {
    const float r = 0x1.010102p-8f;
    const float q0 = fmul(x, r);
    return __fmaf_rn(__fmaf_rn(-q0, 255.0f, x), r, q0);
}

// request the cell record of a point (the load) and keep its three weights
struct TriTap { uint2 c; float wx, wy, wz; };
template <bool FAST>
__device__ __forceinline__ TriTap tri_request(const TriArgs &t, const TriRegs &g, float px, float py, float pz) {
    TriTap tap;
    const uint32_t bx = tri_axis(px, g.hfx, tap.wx), by = tri_axis(py, g.hfy, tap.wy), bz = tri_axis(pz, g.hfz, tap.wz);
    if (FAST) {
        uint32_t lo;
        asm("{\n\t.reg .u32 t;\n\tshl.b32 t, %1, 3;\n\tadd.u32 t, t, %6;\n\tmad.lo.u32 t, %2, %4, t;\n\tmad.lo.u32 %0, %3, %5, t;\n\t}"
            : "=r"(lo) : "r"(bx), "r"(by), "r"(bz), "r"(g.s1), "r"(g.s2), "r"(g.c0));
        asm("{\n\t.reg .u64 p;\n\tmov.b64 p, {%2, %3};\n\tld.global.nc.v2.u32 {%0, %1}, [p];\n\t}" : "=r"(tap.c.x), "=r"(tap.c.y) : "r"(lo), "r"(g.hi));
    } else {  // record index clamped: the guard records already hold the clamped neighbours, like the checker's index clamp
        const int bias = (int)kTriMagicBits - 1 - kTriGuard;
        const uint32_t jx = (uint32_t)min(max((int)bx - bias, 0), (int)t.jx1);
        const uint32_t jy = (uint32_t)min(max((int)by - bias, 0), (int)t.jy1);
        const uint32_t jz = (uint32_t)min(max((int)bz - bias, 0), (int)t.jz1);
        tap.c = __ldg(t.cells + (((uint64_t)jz * t.cy + jy) * t.cx + jx));
    }
    return tap;
}

// checker's fetch_trilinear: lerp x, then y, then z on the raw byte values, result / 255
__device__ __forceinline__ float tri_resolve(const TriTap &tap) {
    const uint2 c = tap.c;
    const float wx = tap.wx, wy = tap.wy, wz = tap.wz;
    const float c00 = lerp_biased(byte_biased(c.x, 0), byte_biased(c.x, 1), wx), c10 = lerp_biased(byte_biased(c.x, 2), byte_biased(c.x, 3), wx);
    const float c01 = lerp_biased(byte_biased(c.y, 0), byte_biased(c.y, 1), wx), c11 = lerp_biased(byte_biased(c.y, 2), byte_biased(c.y, 3), wx);
    const float c0 = lerp_rn(c00, c10, wy), c1 = lerp_rn(c01, c11, wy);
    return div255(lerp_rn(c0, c1, wz));
}

__device__ __forceinline__ float tri_fetch(const TriArgs &t, const TriRegs &g, float px, float py, float pz) { return tri_resolve(tri_request<false>(t, g, px, py, pz)); }

// checker's classify_linear: 128-entry table, position s * 128 - 0.5, clamp, per-channel lerp a + w * (b - a).
// The shared table holds, for every f = floor(position) in [-1, 127], the pair (tf[clamp(f)], tf[clamp(f+1)] - tf[clamp(f)]):
// no index clamps and no subtraction in the loop (the difference is the same RN subtraction, done once per CTA).
constexpr int kTriPairs = 129;
// Two float4 tables, entries first and differences behind them (16-byte stride each: a quarter-warp's LDS.128 hits 8 different
// bank groups for 8 consecutive entries; interleaved (entry, difference) pairs at a 32-byte stride gave 4 and ncu counted up to
// 2.4 shared-memory bank conflicts per warp-sample on gradient-rich frames).
constexpr int kTriDiffOffset = kTriPairs * 16;  // byte offset of the difference table
// pair_c = shared address of the table - ((bits(magic) - 1) << 4) (mod 2^32): the entry address is one shift-add of the bits
__device__ __forceinline__ float4 tri_classify_at(uint32_t pair_c, float t, float tm) {   // t = TF position, tm = RD(t + magic)
    const float w = fsub(t, fsub(tm, kTriMagic));
    const uint32_t addr = pair_c + (__float_as_uint(tm) << 4);
    float4 a, d;
    asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w) : "r"(addr));
    asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4+%5];" : "=f"(d.x), "=f"(d.y), "=f"(d.z), "=f"(d.w) : "r"(addr), "n"(kTriDiffOffset));
    return make_float4(fadd(a.x, fmul(w, d.x)), fadd(a.y, fmul(w, d.y)), fadd(a.z, fmul(w, d.z)), fadd(a.w, fmul(w, d.w)));
}
__device__ __forceinline__ float4 tri_classify(uint32_t pair_c, float s) {
    const float t = fsub(fmul(s, 128.0f), 0.5f);
    return tri_classify_at(pair_c, t, __fadd_rd(t, kTriMagic));
}

// ---- contracted arithmetic (VR5_SAMPLER_TRILINEAR_FMA) ----------------------------------------------------------------
// The same policy with every a + b*c of the march written as ONE fused multiply-add, as nvcc's default -fmad=true builds the
// reference's GPU renderers (GPURenderer4.cu:75-86), and the filtering formulas in their fused form (the CPU checker's
// "trilinear-fma" section spells the operations out). The TU is still compiled with -fmad=false: every fusion below is an
// explicit fma, and x/y of a point, the four x-lerps and the colour pairs go through the packed f32x2 forms (FFMA2 / FADD2).
constexpr float kTfScale255 = 0x1.010102p-1f;  // RN(128 / 255)
constexpr float kInv255 = 0x1.010102p-8f;      // RN(1 / 255)

// packed pairs as float2 through the sm_100 intrinsics (FFMA2 / FADD2, every rounding mode); negation is an operand modifier
__device__ __forceinline__ float2 neg2(float2 a) { return make_float2(-a.x, -a.y); }
__device__ __forceinline__ float2 sub2(float2 a, float2 b) { return __fadd2_rn(a, neg2(b)); }
__device__ __forceinline__ float2 both(float v) { return make_float2(v, v); }

template <bool FMA>
__device__ __forceinline__ float ray_point(float o, float d, float k) { return FMA ? __fmaf_rn(d, k, o) : fadd(o, fmul(d, k)); }

// one axis: cell floor(t) and weight t - floor(t) of texel position t; returns the biased bits of RD(t + magic)
__device__ __forceinline__ uint32_t trif_axis(float t, float &w) {
    const float tm = __fadd_rd(t, kTriMagic);
    w = fsub(t, fsub(tm, kTriMagic));
    return __float_as_uint(tm);
}

// the record load of the contracted march; experiment builds override it (L2 prefetch-size hints, tools/build_variant.sh)
#ifndef VR5_TRIF_PACKZU
#define VR5_TRIF_PACKZU 1
#endif
#ifndef VR5_TRIF_LDHINT
#define VR5_TRIF_LDHINT 0
#endif
#if VR5_TRIF_LDHINT == 1
#define VR5_TRIF_LD "ld.global.nc.L2::128B.v2.u32"
#elif VR5_TRIF_LDHINT == 2
#define VR5_TRIF_LD "ld.global.nc.L2::256B.v2.u32"
#elif VR5_TRIF_LDHINT == 3
#define VR5_TRIF_LD "ld.global.nc.L1::evict_last.v2.u32"
#elif VR5_TRIF_LDHINT == 4
#define VR5_TRIF_LD "ld.global.nc.L1::evict_last.L2::128B.v2.u32"
#else
#define VR5_TRIF_LD "ld.global.nc.v2.u32"
#endif
__device__ __forceinline__ void trif_load(const TriArgs &t, const TriRegs &g, uint32_t bx, uint32_t by, uint32_t bz, bool fast, uint2 &c) {
    if (fast) {
        // (bx * 8 + c0) + by * s1 + bz * s2, every step with ONE kernel-uniform operand (an IMAD takes one uniform register: with
        // c0 as the addend of the bz * s2 term one of the two had to be moved into a general register per sample) and bz - the
        // index that arrives last - in the last step
        uint32_t lo;
        asm("{\n\t.reg .u32 t;\n\tshl.b32 t, %1, 3;\n\tadd.u32 t, t, %6;\n\tmad.lo.u32 t, %2, %4, t;\n\tmad.lo.u32 %0, %3, %5, t;\n\t}"
            : "=r"(lo) : "r"(bx), "r"(by), "r"(bz), "r"(g.s1), "r"(g.s2), "r"(g.c0));
        // volatile: the request stays where it is written, ahead of the current sample's arithmetic (ptxas otherwise sinks it below
        // the early-termination branch, to the end of the iteration, and the next iteration starts on its latency)
        asm volatile("{\n\t.reg .u64 p;\n\tmov.b64 p, {%2, %3};\n\t" VR5_TRIF_LD " {%0, %1}, [p];\n\t}" : "=r"(c.x), "=r"(c.y) : "r"(lo), "r"(g.hi));
    } else {
        const int bias = (int)kTriMagicBits - 1 - kTriGuard;
        const uint32_t jx = (uint32_t)min(max((int)bx - bias, 0), (int)t.jx1);
        const uint32_t jy = (uint32_t)min(max((int)by - bias, 0), (int)t.jy1);
        const uint32_t jz = (uint32_t)min(max((int)bz - bias, 0), (int)t.jz1);
        c = __ldg(t.cells + (((uint64_t)jz * t.cy + jy) * t.cx + jx));
    }
}

// request the cell record of texel position (txy, tz) and keep the three weights: x and y packed, z scalar
struct TriTap2 { uint2 c; float2 wxy; float wz; };
template <bool FAST>
__device__ __forceinline__ TriTap2 trif_request(const TriArgs &t, const TriRegs &g, float2 txy, float tz) {
    TriTap2 tap;
    const float2 tm2 = __fadd2_rd(txy, both(kTriMagic));
    tap.wxy = sub2(txy, __fadd2_rn(tm2, both(-kTriMagic)));
    const uint32_t bz = trif_axis(tz, tap.wz);
    trif_load(t, g, __float_as_uint(tm2.x), __float_as_uint(tm2.y), bz, FAST, tap.c);
    return tap;
}
// texel position of a free world point (shading sample): fma(p, n/2, n/2 - 0.5) per axis; clamped request
__device__ __forceinline__ TriTap2 trif_request_world(const TriArgs &t, const TriRegs &g, float px, float py, float pz) {
    return trif_request<false>(t, g, make_float2(__fmaf_rn(px, g.hfx, g.bfx), __fmaf_rn(py, g.hfy, g.bfy)), __fmaf_rn(pz, g.hfz, g.bfz));
}

// trilinear value on the byte scale: x-lerps of the z0 / z1 face pairs packed, then y packed, then z
__device__ __forceinline__ float trif_resolve(const TriTap2 &tap) {
    const uint2 c = tap.c;
    const float2 a0 = make_float2(byte_biased(c.x, 0), byte_biased(c.y, 0)), b0 = make_float2(byte_biased(c.x, 1), byte_biased(c.y, 1));  // y0 row: (z0, z1)
    const float2 a1 = make_float2(byte_biased(c.x, 2), byte_biased(c.y, 2)), b1 = make_float2(byte_biased(c.x, 3), byte_biased(c.y, 3));  // y1 row
    const float2 wx2 = both(tap.wxy.x), wy2 = both(tap.wxy.y);
    const float2 p = __ffma2_rn(wx2, sub2(b0, a0), __fadd2_rn(a0, both(-8388608.0f)));  // (c00, c01); b - a of the biased values is b - a (exact)
    const float2 q = __ffma2_rn(wx2, sub2(b1, a1), __fadd2_rn(a1, both(-8388608.0f)));  // (c10, c11)
    const float2 r = __ffma2_rn(wy2, sub2(q, p), p);
    return __fmaf_rn(tap.wz, fsub(r.y, r.x), r.x);
}

// linear TF on the byte-scale value: position fma(v, 128/255, -0.5), same pair table as tri_classify
// entry bits(um) (biased floor of the TF position) and weight w -> the interpolated colour pairs
__device__ __forceinline__ void trif_classify_at(uint32_t pair_c, uint32_t um_bits, float w, float2 &crg, float2 &cba) {
    const uint32_t addr = pair_c + (um_bits << 4);
    float4 a, d;
    asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w) : "r"(addr));
    asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4+%5];" : "=f"(d.x), "=f"(d.y), "=f"(d.z), "=f"(d.w) : "r"(addr), "n"(kTriDiffOffset));
    crg = __ffma2_rn(both(w), make_float2(d.x, d.y), make_float2(a.x, a.y));
    cba = __ffma2_rn(both(w), make_float2(d.z, d.w), make_float2(a.z, a.w));
}
__device__ __forceinline__ void trif_classify(uint32_t pair_c, float v, float2 &crg, float2 &cba) {
    const float u = __fmaf_rn(v, kTfScale255, -0.5f);
    const float um = __fadd_rd(u, kTriMagic);
    const float w = fsub(u, fsub(um, kTriMagic));
    const uint32_t addr = pair_c + (__float_as_uint(um) << 4);
    float4 a, d;
    asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w) : "r"(addr));
    asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4+%5];" : "=f"(d.x), "=f"(d.y), "=f"(d.z), "=f"(d.w) : "r"(addr), "n"(kTriDiffOffset));
    crg = __ffma2_rn(both(w), make_float2(d.x, d.y), make_float2(a.x, a.y));
    cba = __ffma2_rn(both(w), make_float2(d.z, d.w), make_float2(a.z, a.w));
}

// ZS ("zero skip", as in vr5_cast.cuh): the host has seen that the transfer function starts with all-zero entries, that every alpha lies
// in [0, 1] and the threshold is not negative. A sample whose interpolation interval of the TF lies inside that prefix gets the colour
// 0 + w * 0 = 0 and leaves the accumulators as they are (acc + 0 * (1 - acc.a), or fma(0, 1 - acc.a, acc)): its table fetches, the
// compositing and the shading test are skipped.
template <bool COUNT, bool FAST, bool SHADE, bool FMA, int DEPTH, bool ZS = false>
__device__ __forceinline__ uint32_t cast_ray_tri(const KArgs &a, const RayView &rv, const TriArgs &t, const TriRegs &g, const uint32_t *esl_bits, const EslPin *esl_pin, int col, int row,
                                                 Counters &cnt) {
    // View::get_ray, Raycaster::intersect and the leading empty-space loop are those of the nearest path (vr5_cast.cuh)
    const float cx = (float)(col - a.half_w), cy = (float)(row - a.half_h);
    float ox, oy, oz, dx, dy, dz;
    if (a.perspective) {
        ox = rv.ox; oy = rv.oy; oz = rv.oz;
        dx = fadd(fadd(rv.dx, fmul(rv.rx, cx)), fmul(rv.ux, cy));
        dy = fadd(fadd(rv.dy, fmul(rv.ry, cx)), fmul(rv.uy, cy));
        dz = fadd(fadd(rv.dz, fmul(rv.rz, cx)), fmul(rv.uz, cy));
    } else {
        dx = rv.dx; dy = rv.dy; dz = rv.dz;
        ox = fadd(fadd(rv.ox, fmul(rv.rx, cx)), fmul(rv.ux, cy));
        oy = fadd(fadd(rv.oy, fmul(rv.ry, cx)), fmul(rv.uy, cy));
        oz = fadd(fadd(rv.oz, fmul(rv.rz, cx)), fmul(rv.uz, cy));
    }
    float k, k_far;
    {
        const float ex = dx == 0.0f ? 0.00001f : dx, ey = dy == 0.0f ? 0.00001f : dy, ez = dz == 0.0f ? 0.00001f : dz;
        const float k1x = fdiv(fsub(a.mbx, ox), ex), k1y = fdiv(fsub(a.mby, oy), ey), k1z = fdiv(fsub(a.mbz, oz), ez);
        const float k2x = fdiv(fsub(-a.mbx, ox), ex), k2y = fdiv(fsub(-a.mby, oy), ey), k2z = fdiv(fsub(-a.mbz, oz), ez);
        k = sel_max(sel_max(sel_min(k1x, k2x), sel_min(k1y, k2y)), sel_min(k1z, k2z));
        k_far = sel_min(sel_min(sel_max(k1x, k2x), sel_max(k1y, k2y)), sel_max(k1z, k2z));
        k = sel_max(k, 0.0f);
        if (!((k < k_far) && (k_far > 0.0f))) return 0u;
    }
    if (COUNT) cnt.rays_hit++;
    float px = ray_point<FMA>(ox, dx, k), py = ray_point<FMA>(oy, dy, k), pz = ray_point<FMA>(oz, dz, k);
    const float step = g.step;
    if (a.esl && !skip_leading_empty_space<COUNT, FMA>(a, esl_pin, esl_bits, step, ox, oy, oz, dx, dy, dz, k, k_far, px, py, pz, cnt)) return 0u;
    if (COUNT) cnt.rays_accum++;
    if (!(k <= k_far)) return 0u;
    if (FMA) {
        // contracted march: the ray is carried into texel space once (dv = d * n/2, ov = fma(o, n/2, n/2 - 0.5)), a sample sits
        // at texel position fma(dv, k, ov); accumulators and colours live in packed pairs (r,g) / (b,a). DEPTH cells are in flight
        // per ray: the record of sample i + DEPTH is requested while sample i is interpolated (the kernel is bound by the latency
        // of these loads - most warp-wide requests contain a DRAM miss - as much as by issue slots; profiles/r01_sampler_choice.md).
        // The loop is unrolled DEPTH times so that every tap and its k keep their registers; each stage ends in ONE exit test,
        // which keeps it a single basic block in which the scheduler places the request ahead of the arithmetic.
        float2 acc_rg = make_float2(0.0f, 0.0f), acc_ba = acc_rg;
        const float2 dvxy = make_float2(fmul(dx, g.hfx), fmul(dy, g.hfy)), ovxy = make_float2(__fmaf_rn(ox, g.hfx, g.bfx), __fmaf_rn(oy, g.hfy, g.bfy));
        const float dvz = fmul(dz, g.hfz), ovz = __fmaf_rn(oz, g.hfz, g.bfz);
        TriTap2 taps[DEPTH];
        float ks[DEPTH];  // ks[j] = k of the sample held in taps[j]: the same sequential k += step as the march itself
        ks[0] = k;
        taps[0] = trif_request<false>(t, g, __ffma2_rn(dvxy, both(k), ovxy), __fmaf_rn(dvz, k, ovz));
#pragma unroll
        for (int j = 1; j < DEPTH; j++) {
            ks[j] = fadd(ks[j - 1], step);
            taps[j] = trif_request<FAST>(t, g, __ffma2_rn(dvxy, both(ks[j]), ovxy), __fmaf_rn(dvz, ks[j], ovz));
        }
        // consumes sample i (tap, k = kc) and refills the tap with sample i + DEPTH, whose k follows that of sample i + DEPTH - 1
        auto stage = [&](TriTap2 &tap, float &kc, const float k_last) -> bool {
            const float ka = fadd(k_last, step);
            const float v = trif_resolve(tap);
            float2 crg, cba;
#if VR5_TRIF_PACKZU
            if (DEPTH > 1 && FAST) {
                // the z axis of the requested sample and the TF position of the current one run the same four operations: packed
                // (with DEPTH > 1 the request has a whole stage of slack, so waiting for v costs it nothing)
                const float2 zu = __ffma2_rn(make_float2(dvz, v), make_float2(ka, kTfScale255), make_float2(ovz, -0.5f));
                const float2 zum = __fadd2_rd(zu, both(kTriMagic));
                const float2 wzu = sub2(zu, __fadd2_rn(zum, both(-kTriMagic)));
                const float2 txy = __ffma2_rn(dvxy, both(ka), ovxy);
                const float2 tm2 = __fadd2_rd(txy, both(kTriMagic));
                tap.wxy = sub2(txy, __fadd2_rn(tm2, both(-kTriMagic)));
                tap.wz = wzu.x;
                trif_load(t, g, __float_as_uint(tm2.x), __float_as_uint(tm2.y), __float_as_uint(zum.x), true, tap.c);
                if (ZS && __float_as_uint(zum.y) < t.zero_um) { kc = ka; return true; }   // transparent: acc.a is what it was, below the threshold
                trif_classify_at(g.pair_c, __float_as_uint(zum.y), wzu.y, crg, cba);
            } else
#endif
            {
                tap = trif_request<FAST>(t, g, __ffma2_rn(dvxy, both(ka), ovxy), __fmaf_rn(dvz, ka, ovz));
                const float u = __fmaf_rn(v, kTfScale255, -0.5f);
                const float um = __fadd_rd(u, kTriMagic);
                if (ZS && __float_as_uint(um) < t.zero_um) { if (COUNT) cnt.samples++; kc = ka; return true; }
                trif_classify_at(g.pair_c, __float_as_uint(um), fsub(u, fsub(um, kTriMagic)), crg, cba);
            }
            if (COUNT) cnt.samples++;
            if (SHADE && cba.y > a.shade_alpha) {
                const float px = __fmaf_rn(dx, kc, ox), py = __fmaf_rn(dy, kc, oy), pz = __fmaf_rn(dz, kc, oz);  // the sample's world point
                const float lvx = fsub(a.lx, px), lvy = fsub(a.ly, py), lvz = fsub(a.lz, pz);
                const float len2 = fadd(fadd(fmul(lvx, lvx), fmul(lvy, lvy)), fmul(lvz, lvz));
                const float inv = fdiv(1.0f, __fsqrt_rn(len2));
                const float qx = fadd(px, fmul(fmul(lvx, inv), 0.01f));
                const float qy = fadd(py, fmul(fmul(lvy, inv), 0.01f));
                const float qz = fadd(pz, fmul(fmul(lvz, inv), 0.01f));
                const float vl = trif_resolve(trif_request_world(t, g, qx, qy, qz));
                const float diffuse = fmul(fmul(fsub(vl, v), kInv255), a.kd);
                crg = make_float2(fadd(crg.x, diffuse), fadd(crg.y, diffuse));
                cba.x = fadd(cba.x, diffuse);
                if (COUNT) cnt.shaded++;
            }
            const float2 tr2 = both(fsub(1.0f, acc_ba.y));
            acc_rg = __ffma2_rn(crg, tr2, acc_rg);
            acc_ba = __ffma2_rn(cba, tr2, acc_ba);
            kc = ka;
            return !(acc_ba.y > g.threshold);  // early ray termination (strict >)
        };
        for (;;) {  // k <= k_far holds on entry (checked above); sample i + 1 marches on while (k <= k_far)
            bool more = true;
#pragma unroll
            for (int j = 0; j < DEPTH; j++) {
                if (more) {
                    const bool open = stage(taps[j], ks[j], ks[(j + DEPTH - 1) % DEPTH]);
                    more = open & (ks[(j + 1) % DEPTH] <= k_far);
                }
            }
            if (!more) break;
        }
        return trunc_clamp(fmul(acc_rg.x, 256.0f), 255u) | (trunc_clamp(fmul(acc_rg.y, 256.0f), 255u) << 8) |
               (trunc_clamp(fmul(acc_ba.x, 256.0f), 255u) << 16) | (trunc_clamp(fmul(acc_ba.y, 256.0f), 255u) << 24);
    }
    float ar = 0.0f, ag = 0.0f, ab = 0.0f, aa = 0.0f;
    // software pipelined like the nearest path: the cell of sample i + 1 is requested before sample i is interpolated and
    // composited (speculative past k_far: the record index is clamped, or - FAST - the host has checked that one step stays
    // inside the guard cells, so the load is always in bounds)
    TriTap tap = tri_request<false>(t, g, px, py, pz);
    while (k <= k_far) {  // GPURenderer4.cu:75-86
        const float kn = fadd(k, step);
        const float nx = fadd(ox, fmul(dx, kn)), ny = fadd(oy, fmul(dy, kn)), nz = fadd(oz, fmul(dz, kn));
        const TriTap next = tri_request<FAST>(t, g, nx, ny, nz);
        const float s = tri_resolve(tap);
        const float tpos = fsub(fmul(s, 128.0f), 0.5f);   // checker's classify_linear position
        const float tpm = __fadd_rd(tpos, kTriMagic);
        if (COUNT) cnt.samples++;
        if (!(ZS && __float_as_uint(tpm) < t.zero_um)) {
            float4 c = tri_classify_at(g.pair_c, tpos, tpm);
            // SHADE=false: the host has seen that no TF entry exceeds the guard's alpha (an interpolated alpha never exceeds both
            // of its entries) or that light_kd <= 0.01
            if (SHADE && c.w > a.shade_alpha) {  // GPURenderer4.cu:78 guard, :41-51 shade_texture
                const float lvx = fsub(a.lx, px), lvy = fsub(a.ly, py), lvz = fsub(a.lz, pz);
                const float len2 = fadd(fadd(fmul(lvx, lvx), fmul(lvy, lvy)), fmul(lvz, lvz));
                const float inv = fdiv(1.0f, __fsqrt_rn(len2));
                const float qx = fadd(px, fmul(fmul(lvx, inv), 0.01f));
                const float qy = fadd(py, fmul(fmul(lvy, inv), 0.01f));
                const float qz = fadd(pz, fmul(fmul(lvz, inv), 0.01f));
                const float diffuse = fmul(fsub(tri_fetch(t, g, qx, qy, qz), s), a.kd);
                c.x = fadd(c.x, diffuse); c.y = fadd(c.y, diffuse); c.z = fadd(c.z, diffuse);
                if (COUNT) cnt.shaded++;
            }
            const float tr = fsub(1.0f, aa);
            ar = fadd(ar, fmul(c.x, tr)); ag = fadd(ag, fmul(c.y, tr)); ab = fadd(ab, fmul(c.z, tr)); aa = fadd(aa, fmul(c.w, tr));
            if (aa > g.threshold) break;
        }
        k = kn; px = nx; py = ny; pz = nz; tap = next;
    }
    return trunc_clamp(fmul(ar, 256.0f), 255u) | (trunc_clamp(fmul(ag, 256.0f), 255u) << 8) |
           (trunc_clamp(fmul(ab, 256.0f), 255u) << 16) | (trunc_clamp(fmul(aa, 256.0f), 255u) << 24);
}

// same persistent-warp / tile-queue skeleton as cast_kernel
// MINB = resident CTAs per SM the register budget is cut for (0: none; 6 = 40 registers, 5 = 48, 4 = 64); DEPTH = cell
// requests in flight per ray (contracted march only)
template <bool COUNT, bool FAST, bool SHADE, bool FMA, int MINB = (SHADE ? 0 : 6), int DEPTH = 1, bool ZS = false>
__global__ void __launch_bounds__(kCtaThreads, MINB) tri_kernel(const __grid_constant__ KArgs a, const __grid_constant__ TriArgs t) {
    __shared__ float4 s_pair[2][kTriPairs];  // [0]: entries, [1]: differences to the next entry
    __shared__ uint32_t s_esl[1024];
    __shared__ EslPin s_esl_pin;
    if (threadIdx.x == 32) s_esl_pin = make_esl_pin(a);
    for (int i = threadIdx.x; i < kTriPairs; i += kCtaThreads) {
        const float4 lo = a.tf[min(max(i - 1, 0), 127)], hi = a.tf[min(i, 127)];
        s_pair[0][i] = lo;
        s_pair[1][i] = make_float4(fsub(hi.x, lo.x), fsub(hi.y, lo.y), fsub(hi.z, lo.z), fsub(hi.w, lo.w));
    }
    for (int i = threadIdx.x; i < 1024; i += kCtaThreads) s_esl[i] = a.esl_bits[i];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const uint32_t pair_c = (uint32_t)__cvta_generic_to_shared(&s_pair[0][0]) - ((kTriMagicBits - 1u) << 4);
#if VR5_TRI_PIN
    __shared__ TriRegs s_regs;
    if (threadIdx.x == 0) s_regs = TriRegs{a.step, a.threshold, t.hfx, t.hfy, t.hfz, t.s1, t.s2, t.c0, t.hi, pair_c, fsub(t.hfx, 0.5f), fsub(t.hfy, 0.5f), fsub(t.hfz, 0.5f)};
    __syncthreads();
    const volatile TriRegs *vr = &s_regs;
    const TriRegs g = {vr->step, vr->threshold, vr->hfx, vr->hfy, vr->hfz, vr->s1, vr->s2, vr->c0, vr->hi, vr->pair_c, vr->bfx, vr->bfy, vr->bfz};
#else
    const TriRegs g = {a.step, a.threshold, t.hfx, t.hfy, t.hfz, t.s1, t.s2, t.c0, t.hi, pair_c, fsub(t.hfx, 0.5f), fsub(t.hfy, 0.5f), fsub(t.hfz, 0.5f)};
#endif
    Counters cnt = {0, 0, 0, 0, 0};
    run_tile_queue<COUNT>(a, [&](const RayView &v, int col, int row) {
        return cast_ray_tri<COUNT, FAST, SHADE, FMA, DEPTH, ZS>(a, v, t, g, s_esl, &s_esl_pin, col, row, cnt);
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
