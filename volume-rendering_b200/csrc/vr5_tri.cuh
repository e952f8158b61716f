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

struct TriArgs {
    const uint2 *cells;      // (nx+1)(ny+1)(nz+1) records, x fastest; record j holds cell i0 = j - 1
    uint32_t cx, cy;         // records per row / rows per slice (nx + 1, ny + 1)
    float nfx, nfy, nfz;     // dims as float
    uint32_t nx, ny, nz;
};

// one axis: texel position t = ((p + 1) * 0.5) * n - 0.5, cell floor(t), weight t - floor(t). floor() comes from an add
// rounded toward -inf against 1.5 * 2^23 (exact for |t| < 2^22), whose bit pattern also carries the integer.
__device__ __forceinline__ void tri_axis(float p, float nf, uint32_t n, uint32_t &j, float &w) {
    const float t = fsub(fmul(fmul(fadd(p, 1.0f), 0.5f), nf), 0.5f);
    const float magic = 12582912.0f;
    const float tm = __fadd_rd(t, magic);
    const float fl = fsub(tm, magic);  // exact
    w = fsub(t, fl);
    const int i0 = (int)(__float_as_uint(tm) - 0x4B400000u);  // floor(t) as an integer
    j = (uint32_t)min(max(i0 + 1, 0), (int)n);                 // record index; clamps both neighbours like the oracle
}

__device__ __forceinline__ float byte_to_float(uint32_t word, int i) {
    return fsub(__uint_as_float(__byte_perm(word, 0x4B000000u, 0x7650u + (uint32_t)i)), 8388608.0f);
}

__device__ __forceinline__ float lerp_rn(float a, float b, float w) { return fadd(a, fmul(w, fsub(b, a))); }

// request the cell record of a point (the load) and keep its three weights
struct TriTap { uint2 c; float wx, wy, wz; };
__device__ __forceinline__ TriTap tri_request(const TriArgs &t, float px, float py, float pz) {
    uint32_t jx, jy, jz;
    TriTap tap;
    tri_axis(px, t.nfx, t.nx, jx, tap.wx);
    tri_axis(py, t.nfy, t.ny, jy, tap.wy);
    tri_axis(pz, t.nfz, t.nz, jz, tap.wz);
    const uint64_t rec = ((uint64_t)jz * t.cy + jy) * t.cx + jx;
    tap.c = __ldg(t.cells + rec);
    return tap;
}

// oracle fetch_trilinear: lerp x, then y, then z on the raw byte values, result / 255
__device__ __forceinline__ float tri_resolve(const TriTap &tap) {
    const uint2 c = tap.c;
    const float wx = tap.wx, wy = tap.wy, wz = tap.wz;
    // byte -> float without the conversion unit (I2F runs at 1/8 rate): drop the byte into the mantissa of 2^23 with one
    // PRMT, subtract 2^23 (exact)
    const float v000 = byte_to_float(c.x, 0), v100 = byte_to_float(c.x, 1), v010 = byte_to_float(c.x, 2), v110 = byte_to_float(c.x, 3);
    const float v001 = byte_to_float(c.y, 0), v101 = byte_to_float(c.y, 1), v011 = byte_to_float(c.y, 2), v111 = byte_to_float(c.y, 3);
    const float c00 = lerp_rn(v000, v100, wx), c10 = lerp_rn(v010, v110, wx);
    const float c01 = lerp_rn(v001, v101, wx), c11 = lerp_rn(v011, v111, wx);
    const float c0 = lerp_rn(c00, c10, wy), c1 = lerp_rn(c01, c11, wy);
    return fdiv(lerp_rn(c0, c1, wz), 255.0f);
}

__device__ __forceinline__ float tri_fetch(const TriArgs &t, float px, float py, float pz) { return tri_resolve(tri_request(t, px, py, pz)); }

// oracle classify_linear: 128-entry table, position s * 128 - 0.5, clamp, per-channel lerp
__device__ __forceinline__ float4 tri_classify(const float4 *s_tf, float s) {
    const float t = fsub(fmul(s, 128.0f), 0.5f);
    const float fl = floorf(t);
    const float w = fsub(t, fl);
    const int f = (int)fl;
    const int j0 = min(max(f, 0), 127), j1 = min(max(f + 1, 0), 127);
    const float4 a = s_tf[j0], b = s_tf[j1];
    return make_float4(lerp_rn(a.x, b.x, w), lerp_rn(a.y, b.y, w), lerp_rn(a.z, b.z, w), lerp_rn(a.w, b.w, w));
}

template <bool COUNT>
__device__ __forceinline__ uint32_t cast_ray_tri(const KArgs &a, const TriArgs &t, const float4 *s_tf, const uint32_t *esl_bits, int col, int row,
                                                 Counters &cnt) {
    // View::get_ray, Raycaster::intersect and the leading empty-space loop are those of the nearest path (vr5_cast.cuh)
    const float cx = (float)(col - a.half_w), cy = (float)(row - a.half_h);
    float ox, oy, oz, dx, dy, dz;
    if (a.perspective) {
        ox = a.ox; oy = a.oy; oz = a.oz;
        dx = fadd(fadd(a.dx, fmul(a.rx, cx)), fmul(a.ux, cy));
        dy = fadd(fadd(a.dy, fmul(a.ry, cx)), fmul(a.uy, cy));
        dz = fadd(fadd(a.dz, fmul(a.rz, cx)), fmul(a.uz, cy));
    } else {
        dx = a.dx; dy = a.dy; dz = a.dz;
        ox = fadd(fadd(a.ox, fmul(a.rx, cx)), fmul(a.ux, cy));
        oy = fadd(fadd(a.oy, fmul(a.ry, cx)), fmul(a.uy, cy));
        oz = fadd(fadd(a.oz, fmul(a.rz, cx)), fmul(a.uz, cy));
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
    float px = fadd(ox, fmul(dx, k)), py = fadd(oy, fmul(dy, k)), pz = fadd(oz, fmul(dz, k));
    const float step = a.step;
    if (a.esl) {
        while (k <= k_far) {
            uint32_t bx = div_block(a, world_to_voxel(px, a.hx, a.nx1));
            uint32_t by = div_block(a, world_to_voxel(py, a.hy, a.ny1));
            uint32_t bz = div_block(a, world_to_voxel(pz, a.hz, a.nz1));
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
            float dk = sel_max(sel_min(sel_min(kx, ky), kz), 0.0f);
            dk = fmul(floorf(fdiv(dk, step)), step);
            k = fadd(k, dk);
            if (COUNT) cnt.esl_iters++;
            k = fadd(k, step);
            px = fadd(ox, fmul(dx, k)); py = fadd(oy, fmul(dy, k)); pz = fadd(oz, fmul(dz, k));
        }
        if (k > k_far) return 0u;
    }
    if (COUNT) cnt.rays_accum++;
    if (!(k <= k_far)) return 0u;
    float ar = 0.0f, ag = 0.0f, ab = 0.0f, aa = 0.0f;
    // software pipelined like the nearest path: the cell of sample i + 1 is requested before sample i is interpolated and
    // composited (speculative past k_far: the record index is clamped, so the load is always in bounds)
    TriTap tap = tri_request(t, px, py, pz);
    while (k <= k_far) {  // GPURenderer4.cu:75-86
        const float kn = fadd(k, step);
        const float nx = fadd(ox, fmul(dx, kn)), ny = fadd(oy, fmul(dy, kn)), nz = fadd(oz, fmul(dz, kn));
        const TriTap next = tri_request(t, nx, ny, nz);
        const float s = tri_resolve(tap);
        float4 c = tri_classify(s_tf, s);
        if (COUNT) cnt.samples++;
        if (c.w > a.shade_alpha) {  // GPURenderer4.cu:78 guard, :41-51 shade_texture
            const float lvx = fsub(a.lx, px), lvy = fsub(a.ly, py), lvz = fsub(a.lz, pz);
            const float len2 = fadd(fadd(fmul(lvx, lvx), fmul(lvy, lvy)), fmul(lvz, lvz));
            const float inv = fdiv(1.0f, __fsqrt_rn(len2));
            const float qx = fadd(px, fmul(fmul(lvx, inv), 0.01f));
            const float qy = fadd(py, fmul(fmul(lvy, inv), 0.01f));
            const float qz = fadd(pz, fmul(fmul(lvz, inv), 0.01f));
            const float diffuse = fmul(fsub(tri_fetch(t, qx, qy, qz), s), a.kd);
            c.x = fadd(c.x, diffuse); c.y = fadd(c.y, diffuse); c.z = fadd(c.z, diffuse);
            if (COUNT) cnt.shaded++;
        }
        const float tr = fsub(1.0f, aa);
        ar = fadd(ar, fmul(c.x, tr)); ag = fadd(ag, fmul(c.y, tr)); ab = fadd(ab, fmul(c.z, tr)); aa = fadd(aa, fmul(c.w, tr));
        if (aa > a.threshold) break;
        k = kn; px = nx; py = ny; pz = nz; tap = next;
    }
    return trunc_clamp(fmul(ar, 256.0f), 255u) | (trunc_clamp(fmul(ag, 256.0f), 255u) << 8) |
           (trunc_clamp(fmul(ab, 256.0f), 255u) << 16) | (trunc_clamp(fmul(aa, 256.0f), 255u) << 24);
}

// same persistent-warp / tile-queue skeleton as cast_kernel
template <bool COUNT>
__global__ void __launch_bounds__(kCtaThreads) tri_kernel(const __grid_constant__ KArgs a, const __grid_constant__ TriArgs t) {
    __shared__ float4 s_tf[128];
    __shared__ uint32_t s_esl[1024];
    for (int i = threadIdx.x; i < 128; i += kCtaThreads) s_tf[i] = a.tf[i];
    for (int i = threadIdx.x; i < 1024; i += kCtaThreads) s_esl[i] = a.esl_bits[i];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    Counters cnt = {0, 0, 0, 0, 0};
    for (;;) {
        unsigned q = 0;
        if (lane == 0) q = atomicAdd(a.queue, 1u) - a.queue_base;
        q = __shfl_sync(0xffffffffu, q, 0);
        if (q >= a.n_queue) break;
        const uint32_t group = q >> (kGroupLog2X + kGroupLog2Y), within = q & ((1u << (kGroupLog2X + kGroupLog2Y)) - 1u);
        const uint32_t tx = ((group % a.groups_x) << kGroupLog2X) + (within & ((1u << kGroupLog2X) - 1u));
        const uint32_t tyl = ((group / a.groups_x) << kGroupLog2Y) + (within >> kGroupLog2X);
        if (tx >= a.tiles_x || tyl >= a.tile_rows_local) continue;
        const int col = (int)tx * kTileW + (lane & 7);
        const int row = (int)(tyl * a.world + a.rank) * kTileH + (lane >> 3);
        if (col < a.width && row < a.height) {
            const uint32_t rgba = cast_ray_tri<COUNT>(a, t, s_tf, s_esl, col, row, cnt);
            if (!COUNT) {
                const int out_row = a.packed ? (int)tyl * kTileH + (lane >> 3) : row;
                reinterpret_cast<uint32_t *>(a.out)[(size_t)out_row * a.width + col] = rgba;
            }
        }
    }
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
