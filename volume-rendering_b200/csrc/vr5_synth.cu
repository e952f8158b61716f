// vr5_synth.cu - device generators of the two synthetic bench volumes; integer arithmetic only,
// so the bytes equal those of the host generator the oracle side uses (checked in tests/test_gpu_parity.py).
// The reference has no synthetic data: these stand in for the datasets its project names but does not ship
// (VolumeRendering.vcproj:225-247).
//
//  synthA(N, seed)  dense Marschner-Lobb-like field (Q15 fixed point, 1024-entry cosine table from the host):
//     c = ((2i+1)*32768)/N - 32768 ; r = isqrt(cx^2+cy^2) ; c1 = COS[(r>>7)&1023] ; c2 = COS[((fM*c1)>>5)&1023], fM=5+seed
//     s = COS[((cz>>7)-256)&1023] ; num = 32767 - s + ((32767+c2)>>2) ; value = min(255,(num*255+40958)/81917)
//  synthB(N, seed)  64 LCG-placed solid spheres (200..255) with a shell (100) of max(1,N/128) voxels on 0.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "../../include/vr5.h"
#include "vr5_internal.h"

namespace {

struct Sphere { int32_t cx, cy, cz, r, value; };
struct SphereSet { Sphere s[64]; };

__constant__ int16_t c_cos[1024];

__device__ __forceinline__ uint32_t isqrt_u32(uint32_t v) {
    uint64_t r = (uint64_t)__fsqrt_rn((float)v);
    while (r * r > v) r--;
    while ((r + 1) * (r + 1) <= v) r++;
    return (uint32_t)r;
}

__device__ __forceinline__ int32_t centre_q15(uint32_t i, uint32_t n) { return (int32_t)(((2u * i + 1u) * 32768u) / n) - 32768; }

// one thread = 4 consecutive x voxels (one 32-bit store)
__global__ void synth_a_kernel(uint8_t *out, uint32_t n, int32_t fm) {
    const uint32_t quads = (n + 3u) / 4u;
    const uint64_t total = (uint64_t)quads * n * n;
    for (uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t xq = (uint32_t)(t % quads), y = (uint32_t)((t / quads) % n), z = (uint32_t)(t / ((uint64_t)quads * n));
        const int32_t cy = centre_q15(y, n), cz = centre_q15(z, n);
        const int32_t s = c_cos[((cz >> 7) - 256) & 1023];
        uint8_t *dst = out + ((uint64_t)z * n + y) * n;
        for (uint32_t i = 0; i < 4; i++) {
            const uint32_t x = xq * 4u + i;
            if (x >= n) break;
            const int32_t cx = centre_q15(x, n);
            const uint32_t r = isqrt_u32((uint32_t)(cx * cx) + (uint32_t)(cy * cy));
            const int32_t c1 = c_cos[(r >> 7) & 1023];
            const int32_t c2 = c_cos[((fm * c1) >> 5) & 1023];
            const int32_t num = 32767 - s + ((32767 + c2) >> 2);
            int32_t v = (num * 255 + 40958) / 81917;
            v = v > 255 ? 255 : (v < 0 ? 0 : v);
            dst[x] = (uint8_t)v;
        }
    }
}

__global__ void synth_b_kernel(uint8_t *out, uint32_t n, int32_t shell, const __grid_constant__ SphereSet set) {
    const uint64_t total = (uint64_t)n * n * n;
    for (uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (uint64_t)gridDim.x * blockDim.x) {
        const int32_t x = (int32_t)(t % n), y = (int32_t)((t / n) % n), z = (int32_t)(t / ((uint64_t)n * n));
        int32_t best = 0;
#pragma unroll 4
        for (int i = 0; i < 64; i++) {
            const Sphere &sp = set.s[i];
            const int64_t ddx = x - sp.cx, ddy = y - sp.cy, ddz = z - sp.cz;
            const int64_t d2 = ddx * ddx + ddy * ddy + ddz * ddz;
            const int64_t ro = sp.r + shell;
            const int32_t v = d2 <= (int64_t)sp.r * sp.r ? sp.value : (d2 <= ro * ro ? 100 : 0);
            best = v > best ? v : best;
        }
        out[t] = (uint8_t)best;
    }
}

}  // namespace

extern "C" int vr5_synth_volume(vr5_renderer *r, int kind, uint16_t n, uint32_t seed) {
    if (!r) return 1;
    if (n == 0 || n > 4096 || (kind != 0 && kind != 1)) { vr5_fail(r, "vr5_synth_volume: kind must be 0|1 and 1 <= n <= 4096", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    const uint64_t bytes = (uint64_t)n * n * n;
    uint8_t *d = nullptr;
    cudaError_t e = cudaMalloc(&d, bytes);
    if (e != cudaSuccess) { vr5_fail(r, "cudaMalloc synthetic volume", e, __FILE__, __LINE__); cudaGetLastError(); return 1; }
    if (kind == 0) {
        int16_t table[1024];
        for (int i = 0; i < 1024; i++) table[i] = (int16_t)lround(32767.0 * cos(2.0 * 3.14159265358979323846 * i / 1024.0));
        e = cudaMemcpyToSymbolAsync(c_cos, table, sizeof(table), 0, cudaMemcpyHostToDevice, r->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(r->stream);  // `table` is a stack array
        if (e != cudaSuccess) { cudaFree(d); vr5_fail(r, "upload cosine table", e, __FILE__, __LINE__); return 1; }
        synth_a_kernel<<<r->num_sms * 8, 256, 0, r->stream>>>(d, n, 5 + (int32_t)seed);
    } else {
        SphereSet set;
        uint32_t state = seed;
        auto next16 = [&state]() { state = state * 1664525u + 1013904223u; return state >> 16; };
        const uint32_t span = (uint32_t)n * 70u / 100u, base = (uint32_t)n * 15u / 100u;
        for (int i = 0; i < 64; i++) {
            set.s[i].cx = (int32_t)(base + ((next16() * span) >> 16));
            set.s[i].cy = (int32_t)(base + ((next16() * span) >> 16));
            set.s[i].cz = (int32_t)(base + ((next16() * span) >> 16));
            set.s[i].r = (int32_t)(((uint32_t)n * (25u + ((next16() * 75u) >> 16))) / 1000u);
            set.s[i].value = (int32_t)(200u + ((next16() * 56u) >> 16));
        }
        const int32_t shell = n / 128 > 1 ? n / 128 : 1;
        synth_b_kernel<<<r->num_sms * 8, 256, 0, r->stream>>>(d, n, shell, set);
    }
    r->launches++;
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(r->stream);
    if (e != cudaSuccess) { cudaFree(d); vr5_fail(r, "synthetic volume kernel", e, __FILE__, __LINE__); return 1; }
    return vr5_adopt_volume(r, d, bytes, n, n, n);
}
