// vr5_prep.cu - device versions of the two serial O(N^3) host passes that precede the hot path:
//   RaycasterBase::set_volume's per-block min/max (RaycasterBase.cpp:94-125) and
//   ModelBase::compute_histogram (ModelBase.cpp:19-33),
// fused into ONE read of the volume: one CTA per ESL block, 16-byte vector loads where the block rows allow it,
// warp-shuffle min/max, a shared-memory histogram per CTA flushed with 64-bit atomics.
// Outputs are byte-identical to the host builders (vr5_scene.cpp) / the reference for volumes below 2^32 voxels;
// above that the reference's 32-bit bins and indices wrap and this code keeps exact 64-bit counts.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <string.h>

#include <algorithm>

#include "../../include/vr5.h"
#include "vr5_internal.h"

namespace {

constexpr int kPrepThreads = 256;

__global__ void __launch_bounds__(kPrepThreads) minmax_hist_kernel(const uint8_t *__restrict__ vol, uint32_t nx, uint32_t ny, uint32_t nz,
                                                                     uint32_t bd, uint32_t cells_x, uint32_t cells_y, uint8_t *__restrict__ minmax,
                                                                     unsigned long long *__restrict__ hist) {
    __shared__ unsigned int s_hist[256];
    __shared__ unsigned int s_lo[kPrepThreads / 32], s_hi[kPrepThreads / 32];
    for (int i = threadIdx.x; i < 256; i += kPrepThreads) s_hist[i] = 0;
    __syncthreads();
    const uint32_t cx = blockIdx.x % cells_x, cy = (blockIdx.x / cells_x) % cells_y, cz = blockIdx.x / (cells_x * cells_y);
    const uint32_t x0 = cx * bd, y0 = cy * bd, z0 = cz * bd;
    const uint32_t wx = min(bd, nx - x0), wy = min(bd, ny - y0), wz = min(bd, nz - z0);
    unsigned int lo = 255, hi = 0;
    const uint32_t rows = wy * wz;
    // a row of the block is wx contiguous bytes; lanes of a warp split rows, each lane walks its row in 4-byte words
    // when the row start is 4-byte aligned (x0 and nx multiples of 4), else byte-wise
    const bool vec = ((x0 | nx | wx) & 3u) == 0u;
    for (uint32_t row = threadIdx.x; row < rows; row += kPrepThreads) {
        const uint32_t y = y0 + row % wy, z = z0 + row / wy;
        const uint8_t *p = vol + ((uint64_t)z * ny + y) * nx + x0;
        if (vec) {
            const uint32_t *p4 = reinterpret_cast<const uint32_t *>(p);
            for (uint32_t i = 0; i < wx / 4; i++) {
                const uint32_t w = __ldg(p4 + i);
#pragma unroll
                for (int b = 0; b < 4; b++) {
                    const unsigned int v = (w >> (8 * b)) & 255u;
                    lo = min(lo, v); hi = max(hi, v);
                    atomicAdd(&s_hist[v], 1u);
                }
            }
        } else {
            for (uint32_t i = 0; i < wx; i++) {
                const unsigned int v = __ldg(p + i);
                lo = min(lo, v); hi = max(hi, v);
                atomicAdd(&s_hist[v], 1u);
            }
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if ((threadIdx.x & 31) == 0) { s_lo[threadIdx.x >> 5] = lo; s_hi[threadIdx.x >> 5] = hi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < kPrepThreads / 32; w++) { lo = min(lo, s_lo[w]); hi = max(hi, s_hi[w]); }
        const uint32_t cell = (cz * VR5_ESL_VOLUME_DIMS + cy) * VR5_ESL_VOLUME_DIMS + cx;
        minmax[2 * cell] = (uint8_t)lo;
        minmax[2 * cell + 1] = (uint8_t)hi;
    }
    for (int i = threadIdx.x; i < 256; i += kPrepThreads)
        if (s_hist[i]) atomicAdd(&hist[i], (unsigned long long)s_hist[i]);
}

int run_prep(vr5_renderer *r, uint16_t *bd_out, uint8_t *minmax_host, vr5_f3 *bs_out, unsigned long long *hist_host) {
    if (!r->d_vol_linear) { vr5_fail(r, "no volume set", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    const uint32_t longest = std::max(r->nx, std::max(r->ny, r->nz));
    const uint32_t bd = std::max<uint32_t>(VR5_ESL_MIN_BLOCK_SIZE, (longest + VR5_ESL_VOLUME_DIMS - 1) / VR5_ESL_VOLUME_DIMS);
    const uint32_t cxn = (r->nx + bd - 1) / bd, cyn = (r->ny + bd - 1) / bd, czn = (r->nz + bd - 1) / bd;
    const size_t mm_bytes = 2u * 32768u;
    uint8_t *d_mm = nullptr;
    unsigned long long *d_hist = nullptr;
    VR5_CUDA(r, cudaMalloc(&d_mm, mm_bytes + 256 * sizeof(unsigned long long)));
    d_hist = reinterpret_cast<unsigned long long *>(d_mm + mm_bytes);
    // untouched cells keep the reference's initial {min 255, max 0} (RaycasterBase.cpp:101-104): 0x00ff little endian
    cudaError_t e = cudaMemsetAsync(d_hist, 0, 256 * sizeof(unsigned long long), r->stream);
    if (e == cudaSuccess) e = cudaMemset2DAsync(d_mm, 2, 0xff, 1, 32768, r->stream);       // min bytes = 255
    if (e == cudaSuccess) e = cudaMemset2DAsync(d_mm + 1, 2, 0x00, 1, 32768, r->stream);   // max bytes = 0
    if (e == cudaSuccess) {
        minmax_hist_kernel<<<cxn * cyn * czn, kPrepThreads, 0, r->stream>>>(r->d_vol_linear, r->nx, r->ny, r->nz, bd, cxn, cyn, d_mm, d_hist);
        r->launches++;
        e = cudaGetLastError();
    }
    if (e == cudaSuccess && minmax_host) e = cudaMemcpyAsync(minmax_host, d_mm, mm_bytes, cudaMemcpyDeviceToHost, r->stream);
    if (e == cudaSuccess && hist_host) e = cudaMemcpyAsync(hist_host, d_hist, 256 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, r->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(r->stream);
    cudaFree(d_mm);
    if (e != cudaSuccess) { vr5_fail(r, "min/max + histogram pass", e, __FILE__, __LINE__); return 1; }
    if (bd_out) *bd_out = (uint16_t)bd;
    if (bs_out) {
        bs_out->x = 2.0f * (float)bd / (float)r->nx;
        bs_out->y = 2.0f * (float)bd / (float)r->ny;
        bs_out->z = 2.0f * (float)bd / (float)r->nz;
    }
    return 0;
}

}  // namespace

extern "C" int vr5_build_minmax_device(vr5_renderer *r, uint16_t *esl_block_dims, uint8_t *minmax, vr5_f3 *esl_block_size) {
    if (!r) return 1;
    if (!minmax) { vr5_fail(r, "vr5_build_minmax_device: NULL output", cudaSuccess, __FILE__, __LINE__); return 1; }
    return run_prep(r, esl_block_dims, minmax, esl_block_size, nullptr);
}

extern "C" int vr5_histogram_device(vr5_renderer *r, float *hist256) {
    if (!r) return 1;
    if (!hist256) { vr5_fail(r, "vr5_histogram_device: NULL output", cudaSuccess, __FILE__, __LINE__); return 1; }
    unsigned long long bins[256];
    if (run_prep(r, nullptr, nullptr, nullptr, bins)) return 1;
    float top = 0.0f;  // ModelBase.cpp:26-32: fourth root, normalised to the largest bin
    for (int i = 0; i < 256; i++) {
        hist256[i] = sqrtf(sqrtf((float)bins[i]));
        top = std::max(top, hist256[i]);
    }
    for (int i = 0; i < 256; i++) hist256[i] = hist256[i] / top;
    return 0;
}
