// vr5_render.cu - host side of the C ABI (include/vr5.h): device state, layouts, launches, timing.
// Replaces the frame drivers of the reference's CUDA renderers (GPURenderer1.cu:65-112, GPURenderer23.cu:55-81,
// GPURenderer4.cu:89-153) with one renderer that owns all of its device state.
#include <cuda.h>
#include <cuda_runtime.h>

#include <limits>
#include <math.h>
#include <cmath>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <string>
#include <vector>

#include "../../include/vr5.h"
#include "vr5_cast.cuh"
#include "vr5_tri.cuh"
#include "vr5_internal.h"

using namespace vr5;

static std::string g_create_error;

void vr5_fail(vr5_renderer *r, const char *what, cudaError_t e, const char *file, int line) {
    char buf[512];
    snprintf(buf, sizeof(buf), "%s: %s (%s:%d)", what, e == cudaSuccess ? "invalid argument" : cudaGetErrorString(e), file, line);
    if (r) r->err = buf; else g_create_error = buf;
}

// ---- layout kernels --------------------------------------------------------------------------------------------
// linear (x fastest) -> 8x4x4 bricks, one thread per 4 consecutive x voxels of the padded brick grid
__global__ void brick_layout_kernel(const uint8_t *__restrict__ lin, uint8_t *__restrict__ brick, uint32_t nx, uint32_t ny,
                                    uint32_t nz, uint32_t bxn, uint32_t byn, uint32_t bzn) {
    const uint64_t n_words = (uint64_t)bxn * byn * bzn * 32u;  // 32 x 4-byte words per brick
    for (uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; w < n_words; w += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t b = w >> 5;
        const uint32_t in = (uint32_t)(w & 31u) << 2;  // byte offset of the word inside the brick, bits: x0 x1 y0 y1 z0 x2 z1
        const uint32_t lx = ((in >> 5) & 1u) << 2, ly = (in >> 2) & 3u, lz = ((in >> 4) & 1u) | (((in >> 6) & 1u) << 1);
        const uint32_t bx = (uint32_t)(b % bxn), by = (uint32_t)((b / bxn) % byn), bz = (uint32_t)(b / ((uint64_t)bxn * byn));
        const uint32_t x = bx * 8u + lx, y = by * 4u + ly, z = bz * 4u + lz;
        uint32_t v = 0;
        if (y < ny && z < nz) {
            const uint8_t *src = lin + ((uint64_t)z * ny + y) * nx + x;
#pragma unroll
            for (int i = 0; i < 4; i++)
                if (x + i < nx) v |= (uint32_t)src[i] << (8 * i);
        }
        reinterpret_cast<uint32_t *>(brick)[w] = v;
    }
}

// linear -> maximum voxel value per brick of (1 << lb)^3 voxels (x fastest), one warp per brick: the table the zero-skip kernels
// consult before they leap over a brick in which every sample would be a transparent one (vr5_cast.cuh, try_leap)
__global__ void brick_max_kernel(const uint8_t *__restrict__ lin, uint8_t *__restrict__ bmax, uint32_t nx, uint32_t ny, uint32_t nz, uint32_t lb,
                                 uint32_t nbx, uint32_t nby, uint32_t nbz) {
    const uint32_t lane = threadIdx.x & 31u, B = 1u << lb;
    const uint64_t n_bricks = (uint64_t)nbx * nby * nbz;
    for (uint64_t b = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; b < n_bricks; b += ((uint64_t)gridDim.x * blockDim.x) >> 5) {
        const uint32_t bx = (uint32_t)(b % nbx), by = (uint32_t)((b / nbx) % nby), bz = (uint32_t)(b / ((uint64_t)nbx * nby));
        const uint32_t x0 = bx << lb, y0 = by << lb, z0 = bz << lb;
        uint32_t m = 0;
        for (uint32_t row = lane; row < B * B; row += 32u) {   // a lane takes whole x-rows of the brick
            const uint32_t y = y0 + (row & (B - 1u)), z = z0 + (row >> lb);
            if (y >= ny || z >= nz) continue;
            const uint8_t *src = lin + ((uint64_t)z * ny + y) * nx;
            for (uint32_t x = x0; x < min(x0 + B, nx); x++) m = max(m, (uint32_t)src[x]);
        }
        for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (lane == 0) bmax[b] = (uint8_t)m;
    }
}

// linear -> linear padded by one replicated voxel at the upper end of every axis (source of the 3-D texture: lets the
// march drop its upper index clamp, see sample_volume<.., CLAMP=false>)
__global__ void pad_edge_kernel(const uint8_t *__restrict__ lin, uint8_t *__restrict__ padded, uint32_t nx, uint32_t ny, uint32_t nz) {
    const uint32_t px = nx + 1, py = ny + 1, pz = nz + 1;
    const uint64_t n = (uint64_t)px * py * pz;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t x = (uint32_t)(i % px), y = (uint32_t)((i / px) % py), z = (uint32_t)(i / ((uint64_t)px * py));
        padded[i] = lin[((uint64_t)min(z, nz - 1) * ny + min(y, ny - 1)) * nx + min(x, nx - 1)];
    }
}

// linear -> window layout (fetch mode 4): padded grid sx*sy*(nz+2) with original voxel (x,y,z) at (x+1,y+1,z+1), borders
// replicated, preceded and followed by `margin` guard bytes that only speculative prefetches can touch
__global__ void window_layout_kernel(const uint8_t *__restrict__ lin, uint8_t *__restrict__ win, uint64_t margin, uint64_t core, uint32_t sx,
                                     uint32_t sxy, uint32_t nx, uint32_t ny, uint32_t nz) {
    const uint64_t n = core + 2 * margin;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        uint8_t v = 0;
        if (i >= margin && i < margin + core) {
            const uint64_t off = i - margin;
            const int z = (int)(off / sxy) - 1, y = (int)((off % sxy) / sx) - 1, x = (int)(off % sx) - 1;
            v = lin[((uint64_t)min(max(z, 0), (int)nz - 1) * ny + min(max(y, 0), (int)ny - 1)) * nx + min(max(x, 0), (int)nx - 1)];
        }
        win[i] = v;
    }
}

// linear -> window of the power-of-two variants: rows of `sx` bytes, slices of `rows` rows, `sz` slices, voxel i at
// coordinate i + origin on every axis, everything outside the grid replicated from the nearest voxel (clamp)
__global__ void p2_layout_kernel(const uint8_t *__restrict__ lin, uint8_t *__restrict__ win, uint32_t sx, uint32_t rows, uint32_t sz, int origin,
                                 uint32_t nx, uint32_t ny, uint32_t nz) {
    const uint64_t n = (uint64_t)sz * rows * sx;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t row = i / sx;
        const int x = (int)(i % sx) - origin, y = (int)(row % rows) - origin, z = (int)(row / rows) - origin;
        win[i] = lin[((uint64_t)min(max(z, 0), (int)nz - 1) * ny + min(max(y, 0), (int)ny - 1)) * nx + min(max(x, 0), (int)nx - 1)];
    }
}

// linear -> trilinear cells (vr5_tri.cuh): record (jx, jy, jz) = the eight voxels around cell i0 = j - 1 - G, clamped to the
// grid (G guard cells per side); byte b of the 8-byte record is the neighbour with offsets (b & 1, (b >> 1) & 1, b >> 2)
__global__ void cell_layout_kernel(const uint8_t *__restrict__ lin, uint2 *__restrict__ cells, uint32_t nx, uint32_t ny, uint32_t nz) {
    const uint32_t cx = nx + 1 + 2 * kTriGuard, cy = ny + 1 + 2 * kTriGuard, cz = nz + 1 + 2 * kTriGuard;
    const uint64_t n = (uint64_t)cx * cy * cz;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const int x0 = (int)(i % cx) - 1 - kTriGuard, y0 = (int)((i / cx) % cy) - 1 - kTriGuard, z0 = (int)(i / ((uint64_t)cx * cy)) - 1 - kTriGuard;
        uint32_t w[2] = {0u, 0u};
#pragma unroll
        for (int b = 0; b < 8; b++) {
            const int x = min(max(x0 + (b & 1), 0), (int)nx - 1), y = min(max(y0 + ((b >> 1) & 1), 0), (int)ny - 1), z = min(max(z0 + (b >> 2), 0), (int)nz - 1);
            w[b >> 2] |= (uint32_t)lin[((uint64_t)z * ny + y) * nx + x] << (8 * (b & 3));
        }
        cells[i] = make_uint2(w[0], w[1]);
    }
}

__global__ void unpack_strips_kernel(const uint32_t *__restrict__ packed, uint32_t rows_padded, int world, uint32_t width,
                                     uint32_t height, uint32_t *__restrict__ frame) {
    const uint64_t n = (uint64_t)width * height;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t row = (uint32_t)(i / width), col = (uint32_t)(i % width);
        const uint32_t strip = row / kTileH, rank = strip % world, local = strip / world;
        frame[i] = packed[((uint64_t)rank * rows_padded + local * kTileH + row % kTileH) * width + col];
    }
}

__global__ void l2_flush_kernel(uint32_t *buf, uint64_t n_words, uint32_t v) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_words; i += (uint64_t)gridDim.x * blockDim.x) buf[i] = v;
}

// ---- frame lanes ---------------------------------------------------------------------------------------------------
cudaStream_t vr5_lane_stream(vr5_renderer *r, int lane) { return r->n_lanes == 1 ? r->stream : r->lanes[lane].stream; }

// everything this handle has in flight, on the user stream and on every lane
static cudaError_t sync_all(vr5_renderer *r) {
    cudaError_t first = r->stream ? cudaStreamSynchronize(r->stream) : cudaSuccess;
    for (auto &ln : r->lanes) {
        if (!ln.stream) continue;
        const cudaError_t e = cudaStreamSynchronize(ln.stream);
        if (first == cudaSuccess) first = e;
        ln.pending = false;
    }
    if (r->consumer_stream) {
        const cudaError_t e = cudaStreamSynchronize(r->consumer_stream);
        if (first == cudaSuccess) first = e;
        r->consumer_pending = false;
    }
    return first;
}

// Work about to be enqueued on a lane is ordered after whatever the caller has put on the user stream so far (its own
// consumer of an earlier frame, uploads). The join in the other direction is deferred (vr5_join / vr5_sync / a blocking
// call): waiting on the user stream right away would chain every frame behind the previous one again.
int vr5_lane_begin(vr5_renderer *r, int lane) {
    if (r->n_lanes == 1) return 0;
    VR5_CUDA(r, cudaEventRecord(r->ev_fork, r->stream));
    VR5_CUDA(r, cudaStreamWaitEvent(r->lanes[lane].stream, r->ev_fork, 0));
    return 0;
}

int vr5_lane_end(vr5_renderer *r, int lane) {
    if (r->n_lanes == 1) return 0;
    VR5_CUDA(r, cudaEventRecord(r->lanes[lane].done, r->lanes[lane].stream));
    r->lanes[lane].pending = true;
    return 0;
}

static int ensure_lane_frames(vr5_renderer *r) {
    for (int i = 0; i < r->n_lanes; i++)
        if (!r->lanes[i].d_frame && r->frame_bytes) VR5_CUDA(r, cudaMalloc(&r->lanes[i].d_frame, r->frame_bytes));
    return 0;
}

extern "C" int vr5_set_overlap(vr5_renderer *r, int lanes) {
    if (!r) return 1;
    if (lanes < 1 || lanes > 4) { vr5_fail(r, "vr5_set_overlap: 1..4 lanes", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    VR5_CUDA(r, sync_all(r));
    r->n_lanes = lanes;
    r->frame_no = 0;
    return ensure_lane_frames(r);
}
extern "C" int vr5_get_overlap(const vr5_renderer *r) { return r ? r->n_lanes : -1; }

extern "C" int vr5_join(vr5_renderer *r) {
    if (!r) return 1;
    VR5_CUDA(r, cudaSetDevice(r->device));
    for (int i = 0; i < r->n_lanes && r->n_lanes > 1; i++)
        if (r->lanes[i].pending) { VR5_CUDA(r, cudaStreamWaitEvent(r->stream, r->lanes[i].done, 0)); r->lanes[i].pending = false; }
    if (r->consumer_pending) { VR5_CUDA(r, cudaStreamWaitEvent(r->stream, r->consumer_done, 0)); r->consumer_pending = false; }
    return 0;
}

extern "C" int vr5_sync_lane(vr5_renderer *r, int lane) {
    if (!r) return 1;
    if (lane < 0 || lane >= r->n_lanes) { vr5_fail(r, "vr5_sync_lane: no such lane", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    VR5_CUDA(r, cudaStreamSynchronize(vr5_lane_stream(r, lane)));
    return 0;
}

// ---- lifecycle ---------------------------------------------------------------------------------------------------
extern "C" int vr5_version(void) { return VR5_VERSION; }
extern "C" const char *vr5_name(void) { return "CUDA B200 persistent tile queue (GPURenderer5)"; }
extern "C" const char *vr5_last_error(const vr5_renderer *r) { return r ? r->err.c_str() : g_create_error.c_str(); }

extern "C" int vr5_create(int device, vr5_renderer **out) {
    if (!out) return 1;
    *out = nullptr;
    int n_dev = 0;
    cudaError_t e = cudaGetDeviceCount(&n_dev);
    if (e != cudaSuccess || n_dev == 0) {
        vr5_fail(nullptr, "vr5_create: no CUDA device (there is no CPU fallback)", e, __FILE__, __LINE__);
        return 1;
    }
    if (device < 0) { e = cudaGetDevice(&device); if (e != cudaSuccess) { vr5_fail(nullptr, "cudaGetDevice", e, __FILE__, __LINE__); return 1; } }
    if (device >= n_dev) { vr5_fail(nullptr, "vr5_create: device ordinal out of range", cudaSuccess, __FILE__, __LINE__); return 1; }
    vr5_renderer *r = new vr5_renderer();
    r->device = device;
    auto bail = [&](const char *what, cudaError_t err) { vr5_fail(nullptr, what, err, __FILE__, __LINE__); vr5_destroy(r); return 1; };
    if ((e = cudaSetDevice(device)) != cudaSuccess) return bail("cudaSetDevice", e);
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return bail("cudaGetDeviceProperties", e);
    r->num_sms = prop.multiProcessorCount;
    r->l2_bytes = (size_t)prop.l2CacheSize;
    if ((e = cudaStreamCreateWithFlags(&r->own_stream, cudaStreamNonBlocking)) != cudaSuccess) return bail("cudaStreamCreate", e);
    r->stream = r->own_stream;
    if ((e = cudaMalloc(&r->d_tf, VR5_TF_SIZE * sizeof(float4))) != cudaSuccess) return bail("cudaMalloc tf", e);
    if ((e = cudaMalloc(&r->d_esl, VR5_ESL_VOLUME_SIZE * sizeof(uint32_t))) != cudaSuccess) return bail("cudaMalloc esl", e);
    for (auto &ln : r->lanes) {  // four lanes' worth of streams, counters and events; vr5_set_overlap picks how many are used
        if ((e = cudaStreamCreateWithFlags(&ln.stream, cudaStreamNonBlocking)) != cudaSuccess) return bail("cudaStreamCreate lane", e);
        if ((e = cudaMalloc(&ln.d_queue, 128)) != cudaSuccess) return bail("cudaMalloc queue", e);
        if ((e = cudaMemset(ln.d_queue, 0, 128)) != cudaSuccess) return bail("cudaMemset queue", e);
        if ((e = cudaEventCreateWithFlags(&ln.done, cudaEventDisableTiming)) != cudaSuccess) return bail("cudaEventCreate", e);
    }
    if ((e = cudaEventCreateWithFlags(&r->ev_fork, cudaEventDisableTiming)) != cudaSuccess) return bail("cudaEventCreate", e);
    if ((e = cudaMalloc(&r->d_counters, sizeof(Counters))) != cudaSuccess) return bail("cudaMalloc counters", e);
    if ((e = cudaMalloc(&r->d_leap_stats, 16)) != cudaSuccess) return bail("cudaMalloc leap statistics", e);
    if ((e = cudaMemset(r->d_leap_stats, 0, 16)) != cudaSuccess) return bail("cudaMemset leap statistics", e);
    if ((e = cudaMalloc(&r->d_seq_queue, 128)) != cudaSuccess) return bail("cudaMalloc sequence queue", e);
    if ((e = cudaMemset(r->d_seq_queue, 0, 128)) != cudaSuccess) return bail("cudaMemset sequence queue", e);
    if ((e = cudaStreamCreateWithFlags(&r->consumer_stream, cudaStreamNonBlocking)) != cudaSuccess) return bail("cudaStreamCreate consumer", e);
    if ((e = cudaEventCreateWithFlags(&r->consumer_done, cudaEventDisableTiming)) != cudaSuccess) return bail("cudaEventCreate", e);
    {   // sequence records are allocated in stream order: keep freed blocks in the pool instead of returning them to the OS at
        // every synchronisation (the default release threshold is 0, which made every sequence launch pay a fresh allocation)
        cudaMemPool_t pool = nullptr;
        unsigned long long keep = ~0ull;
        if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess && pool) cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        cudaGetLastError();
    }
    // experiment knobs: the environment is read here, once; the launch path only looks at r->opt (vr5_set_option changes it)
    for (const char *name : {"VR5_QUEUE_ORDER", "VR5_ROW_MAJOR", "VR5_WARP_FETCH", "VR5_TRI_CLAMPED", "VR5_TRIF_CTAS", "VR5_TRIF_DEPTH", "VR5_NO_P2", "VR5_NO_ZSKIP", "VR5_NO_LEAP", "VR5_LEAP_BRICK", "VR5_LEAP_STATS", "VR5_P2_WIDE", "VR5_WIN_SX", "VR5_ESL_EXACT", "VR5_SEQ_DEBUG", "VR5_HOST_SINGLE"})
        if (const char *v = getenv(name)) vr5_set_option(r, name, v);
    r->err.clear();
    const char *env = getenv("VR5_SAMPLER");
    r->sampler = VR5_SAMPLER_HYBRID_WINDOW;  // fastest path on B200 (profiles/r01_sampler_choice.md); $VR5_SAMPLER overrides
    if (env) {
        if (!strcmp(env, "brick") || !strcmp(env, "1")) r->sampler = VR5_SAMPLER_BRICK;
        else if (!strcmp(env, "linear") || !strcmp(env, "0")) r->sampler = VR5_SAMPLER_LINEAR;
        else if (!strcmp(env, "tex3d") || !strcmp(env, "2")) r->sampler = VR5_SAMPLER_TEX3D;
        else if (!strcmp(env, "hybrid") || !strcmp(env, "3")) r->sampler = VR5_SAMPLER_HYBRID;
        else if (!strcmp(env, "hybridw") || !strcmp(env, "4")) r->sampler = VR5_SAMPLER_HYBRID_WINDOW;
        else if (!strcmp(env, "window") || !strcmp(env, "5")) r->sampler = VR5_SAMPLER_WINDOW;
        else if (!strcmp(env, "hybridw2") || !strcmp(env, "6")) r->sampler = VR5_SAMPLER_HYBRID_WINDOW2;
        else if (!strcmp(env, "trilinear") || !strcmp(env, "16")) r->sampler = VR5_SAMPLER_TRILINEAR;
        else if (!strcmp(env, "trilinear-fma") || !strcmp(env, "18")) r->sampler = VR5_SAMPLER_TRILINEAR_FMA;
    }
    *out = r;
    return 0;
}

// ---- power-of-two window (kernel variants >= kP2First) -----------------------------------------------------------------
// The window array must start at an address whose low 32 bits equal a given constant (so that the float-bit biases of the
// three indices cancel, see vr5_cast.cuh) and must not cross a 4-GiB boundary. cudaMalloc cannot promise either, so the
// array is mapped with the driver's virtual-memory API into a 4-GiB-aligned reservation. Entry points are resolved at run
// time (cudaGetDriverEntryPoint): libvr5.so keeps linking against the runtime only.
struct VmmApi {
    CUresult (*reserve)(CUdeviceptr *, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
    CUresult (*addr_free)(CUdeviceptr, size_t) = nullptr;
    CUresult (*create)(CUmemGenericAllocationHandle *, size_t, const CUmemAllocationProp *, unsigned long long) = nullptr;
    CUresult (*release)(CUmemGenericAllocationHandle) = nullptr;
    CUresult (*map)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
    CUresult (*unmap)(CUdeviceptr, size_t) = nullptr;
    CUresult (*set_access)(CUdeviceptr, size_t, const CUmemAccessDesc *, size_t) = nullptr;
    CUresult (*granularity)(size_t *, const CUmemAllocationProp *, CUmemAllocationGranularity_flags) = nullptr;
    bool ok = false;
};

static const VmmApi &vmm_api() {
    static VmmApi api;
    static bool tried = false;
    if (tried) return api;
    tried = true;
    auto get = [](const char *name, void **fn) {
        cudaDriverEntryPointQueryResult q;
        return cudaGetDriverEntryPoint(name, fn, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess && *fn;
    };
    api.ok = get("cuMemAddressReserve", (void **)&api.reserve) && get("cuMemAddressFree", (void **)&api.addr_free) &&
             get("cuMemCreate", (void **)&api.create) && get("cuMemRelease", (void **)&api.release) && get("cuMemMap", (void **)&api.map) &&
             get("cuMemUnmap", (void **)&api.unmap) && get("cuMemSetAccess", (void **)&api.set_access) &&
             get("cuMemGetAllocationGranularity", (void **)&api.granularity);
    if (!api.ok) cudaGetLastError();
    return api;
}

static void free_p2(vr5_renderer *r) {
    if (r->tex_p2) { cudaDestroyTextureObject(r->tex_p2); r->tex_p2 = 0; }
    if (r->d_array_p2) { cudaFreeArray(r->d_array_p2); r->d_array_p2 = nullptr; }
    const VmmApi &v = vmm_api();
    if (r->p2_map_addr) { v.unmap((CUdeviceptr)r->p2_map_addr, r->p2_map_size); r->p2_map_addr = 0; }
    if (r->p2_handle) { v.release((CUmemGenericAllocationHandle)r->p2_handle); r->p2_handle = 0; }
    if (r->p2_va) { v.addr_free((CUdeviceptr)r->p2_va, r->p2_va_size); r->p2_va = 0; }
    r->p2_wl = 0;
}

static void free_layouts(vr5_renderer *r) {
    if (r->tex) { cudaDestroyTextureObject(r->tex); r->tex = 0; }
    if (r->d_array) { cudaFreeArray(r->d_array); r->d_array = nullptr; }
    if (r->d_vol_brick) { cudaFree(r->d_vol_brick); r->d_vol_brick = nullptr; }
    if (r->d_vol_padded) { cudaFree(r->d_vol_padded); r->d_vol_padded = nullptr; }
    if (r->d_vol_window) { cudaFree(r->d_vol_window); r->d_vol_window = nullptr; }
    if (r->d_cells) { cudaFree(r->d_cells); r->d_cells = nullptr; }
    if (r->d_brick_max) { cudaFree(r->d_brick_max); r->d_brick_max = nullptr; }
    free_p2(r);
}

extern "C" void vr5_destroy(vr5_renderer *r) {
    if (!r) return;
    cudaSetDevice(r->device);
    sync_all(r);
    vr5_hostring_detach(r);
    vr5_peer_close(r);
    free_layouts(r);
    cudaFree(r->d_group_order);
    cudaFree(r->d_vol_linear);
    for (auto &ln : r->lanes) {
        cudaFree(ln.d_frame);
        cudaFree(ln.d_queue);
        if (ln.done) cudaEventDestroy(ln.done);
        if (ln.stream) cudaStreamDestroy(ln.stream);
    }
    if (r->ev_fork) cudaEventDestroy(r->ev_fork);
    if (r->consumer_done) cudaEventDestroy(r->consumer_done);
    if (r->consumer_stream) cudaStreamDestroy(r->consumer_stream);
    cudaFree(r->d_seq_queue);
    cudaFree(r->d_tf);
    cudaFree(r->d_esl);
    cudaFree(r->d_counters);
    cudaFree(r->d_leap_stats);
    cudaFree(r->d_flush);
    if (r->own_stream) cudaStreamDestroy(r->own_stream);
    delete r;
}

// ---- scene state -------------------------------------------------------------------------------------------------
extern "C" int vr5_set_window(vr5_renderer *r, uint16_t width, uint16_t height) {
    if (!r) return 1;
    VR5_CUDA(r, cudaSetDevice(r->device));
    if (width == 0 || height == 0) { vr5_fail(r, "vr5_set_window: empty viewport", cudaSuccess, __FILE__, __LINE__); return 1; }
    if ((r->peer_base || r->hr_base) && (width != r->width || height != r->height)) {
        // a ring's slots are sized for the window it was created with: a larger window would write past them
        vr5_fail(r, "vr5_set_window: a frame ring is attached (vr5_peer_close / vr5_hostring_detach first)", cudaSuccess, __FILE__, __LINE__);
        return 1;
    }
    VR5_CUDA(r, sync_all(r));
    for (auto &ln : r->lanes) if (ln.d_frame) { cudaFree(ln.d_frame); ln.d_frame = nullptr; }
    r->width = width; r->height = height;
    r->frame_bytes = (size_t)width * height * 4;
    return ensure_lane_frames(r);
}

extern "C" int vr5_set_transfer_fn(vr5_renderer *r, const float *tf, const uint32_t *esl_bits) {
    if (!r) return 1;
    if (!tf || !esl_bits) { vr5_fail(r, "vr5_set_transfer_fn: NULL table", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    VR5_CUDA(r, sync_all(r));
    VR5_CUDA(r, cudaMemcpyAsync(r->d_tf, tf, VR5_TF_SIZE * sizeof(float4), cudaMemcpyHostToDevice, r->stream));
    VR5_CUDA(r, cudaMemcpyAsync(r->d_esl, esl_bits, VR5_ESL_VOLUME_SIZE * sizeof(uint32_t), cudaMemcpyHostToDevice, r->stream));
    VR5_CUDA(r, cudaStreamSynchronize(r->stream));
    r->have_tf = true;
    r->tf_can_shade = false;  // CPURenderer.cpp:32: only entries with alpha > 0.05 ever reach Raycaster::shade
    r->tf_opaque = false;         // an entry that can end a ray on its own makes neighbouring rays end at different depths
    for (int i = 0; i < VR5_TF_SIZE; i++) { r->tf_can_shade |= tf[4 * i + 3] > 0.05f; r->tf_opaque |= tf[4 * i + 3] >= 0.999f; }
    // leading all-zero entries (the ZS kernels skip the compositing of samples that classify to them, vr5_cast.cuh): only for a
    // table whose values are all finite with every alpha inside [0, 1] - then acc.a stays finite and 0 * (1 - acc.a) is a zero
    bool tame = true;
    for (int i = 0; i < VR5_TF_SIZE * 4; i++) tame = tame && std::isfinite(tf[i]);
    for (int i = 0; i < VR5_TF_SIZE; i++) tame = tame && tf[4 * i + 3] >= 0.0f && tf[4 * i + 3] <= 1.0f;
    int zp = 0;
    while (zp < VR5_TF_SIZE && tf[4 * zp] == 0.0f && tf[4 * zp + 1] == 0.0f && tf[4 * zp + 2] == 0.0f && tf[4 * zp + 3] == 0.0f) zp++;
    r->tf_zero_prefix = tame ? zp : 0;
    return 0;
}

// (nx+1)(ny+1)(nz+1) copy with the last voxel of every axis replicated; 0 on success
static int build_padded(vr5_renderer *r) {
    // + kWindowGuard slices, rows and bytes of tail: the unclamped LDG path prefetches past the clip interval (reach_ok)
    const uint64_t bytes = (uint64_t)(r->nx + 1) * (r->ny + 1) * (r->nz + 1) + (uint64_t)kWindowGuard * ((uint64_t)(r->nx + 1) * (r->ny + 1) + (r->nx + 1) + 1);
    cudaError_t e = cudaMalloc(&r->d_vol_padded, bytes);
    if (e != cudaSuccess) { vr5_fail(r, "cudaMalloc padded volume", e, __FILE__, __LINE__); cudaGetLastError(); r->d_vol_padded = nullptr; return 1; }
    pad_edge_kernel<<<r->num_sms * 8, 256, 0, r->stream>>>(r->d_vol_linear, r->d_vol_padded, r->nx, r->ny, r->nz);
    r->launches++;
    VR5_CUDA(r, cudaGetLastError());
    return 0;
}

// Window layout for fetch mode 4 (vr5_cast.cuh, fetch_window). Leaves d_vol_window NULL (not an error) when the volume
// cannot live inside one 4-GiB address window; the launcher then falls back to the texture path.
static int build_window(vr5_renderer *r) {
    uint64_t sx = r->nx + 2 < 32 ? 32 : r->nx + 2, sy = r->ny + 2 < 32 ? 32 : r->ny + 2;  // sxy >= 1024 keeps wmz below 2^23
    if (r->opt.win_sx >= sx && r->opt.win_sx <= 65536) sx = r->opt.win_sx;  // experiment knob
    const uint64_t sxy = sx * sy, core = sxy * (r->nz + 2);
    const uint64_t margin = ((uint64_t)kWindowGuard * (sxy + sx + 1) + 255) & ~255ull;
    const uint64_t total = core + 2 * margin;
    if (total >= (1ull << 32)) return 0;
    uint8_t *held[4] = {nullptr, nullptr, nullptr, nullptr}, *got = nullptr;
    for (int tries = 0; tries < 4 && !got; tries++) {  // an allocation that straddles a 4-GiB boundary is retried
        uint8_t *p = nullptr;
        cudaError_t e = cudaMalloc(&p, total);
        if (e != cudaSuccess) { cudaGetLastError(); break; }
        const uint64_t a0 = (uint64_t)(uintptr_t)p;
        if ((a0 >> 32) == ((a0 + total - 1) >> 32)) got = p; else held[tries] = p;
    }
    for (uint8_t *h : held) if (h) cudaFree(h);
    if (!got) return 0;
    r->d_vol_window = got;
    window_layout_kernel<<<r->num_sms * 8, 256, 0, r->stream>>>(r->d_vol_linear, got, margin, core, (uint32_t)sx, (uint32_t)sxy, r->nx, r->ny, r->nz);
    r->launches++;
    VR5_CUDA(r, cudaGetLastError());
    // biased index b = 0x4B000000 + m + i (bits of (2^23 + m) + v rounded toward zero). Choose mz, my, mx so that
    // bz*sxy + by*sx + bx == low 32 bits of &window[(iz+1), (iy+1), (ix+1)] with no additive constant left over.
    const uint32_t B = 0x4B000000u, K = (uint32_t)(sxy + sx + 1), G = 64;  // G: lowest offset, keeps 2^23 + m + v >= 2^23 for guard indices
    const uint32_t w_lo = (uint32_t)((uint64_t)(uintptr_t)got + margin);
    const uint32_t t = w_lo + K - B * K - G * K;  // mod 2^32
    const uint32_t mz = t / (uint32_t)sxy + G, rem = t % (uint32_t)sxy, my = rem / (uint32_t)sx + G, mx = rem % (uint32_t)sx + G;
    r->win_hi = (uint32_t)(((uint64_t)(uintptr_t)got) >> 32);
    r->win_sx = (uint32_t)sx; r->win_sxy = (uint32_t)sxy;
    r->win_mx = (float)(8388608u + mx); r->win_my = (float)(8388608u + my); r->win_mz = (float)(8388608u + mz);
    // self-check of the modular identity on the eight corners of the index range (-1 .. n)
    for (int c = 0; c < 8; c++) {
        const int32_t ix = (c & 1) ? (int32_t)r->nx : -1, iy = (c & 2) ? (int32_t)r->ny : -1, iz = (c & 4) ? (int32_t)r->nz : -1;
        const uint32_t lo = (B + mz + (uint32_t)iz) * (uint32_t)sxy + (B + my + (uint32_t)iy) * (uint32_t)sx + (B + mx + (uint32_t)ix);
        const uint32_t want = w_lo + (uint32_t)(iz + 1) * (uint32_t)sxy + (uint32_t)(iy + 1) * (uint32_t)sx + (uint32_t)(ix + 1);
        if (lo != want || mz + r->nz + 2 * G >= (1u << 23)) {
            cudaFree(r->d_vol_window); r->d_vol_window = nullptr;
            vr5_fail(r, "window layout: address identity does not hold", cudaSuccess, __FILE__, __LINE__);
            return 1;
        }
    }
    return 0;
}

static int exact_log2(uint32_t n) {
    for (int e = 0; e < 32; e++) if (n == (1u << e)) return e;
    return -1;
}

// Power-of-two window + its texture. Leaves p2_wl == 0 (not an error) when the volume is not eligible: a dimension that is
// not a power of two, no VMM support, or the required start address too close to 4 GiB.
static int build_p2(vr5_renderer *r) {
    const int ex = exact_log2(r->nx), ey = exact_log2(r->ny), ez = exact_log2(r->nz);
    if (ex < 1 || ey < 1 || ez < 1) return 0;
    const VmmApi &v = vmm_api();
    if (!v.ok) return 0;
    const uint32_t G = kP2Origin;
    // strides are deliberately NOT powers of two: 1024-byte rows measured 10 % slower than 514-byte ones (cache-set aliasing)
    uint32_t sx = r->nx + 2 * G + 2;
    if (r->opt.win_sx >= r->nx + 2 * G && r->opt.win_sx <= 65536) sx = r->opt.win_sx;  // experiment knob
    const uint32_t rows = r->ny + 2 * G;
    const uint32_t sz = r->nz + 2 * G;
    // index scale per axis is h = n / 2 = 2^(e-1); magic = 2^(23-(e-1)) + G / h has bit pattern ((150-(e-1)) << 23) + G
    const uint32_t bx = (uint32_t)(150 - (ex - 1)) << 23, by = (uint32_t)(150 - (ey - 1)) << 23, bz = (uint32_t)(150 - (ez - 1)) << 23;
    uint64_t size = (uint64_t)sz * rows * sx;
    // An array of 4 GiB or more (2048^3) takes the WIDE variant (kP2Wide): address = bz * sxy + {by * sx + bx, hi} in 64-bit
    // arithmetic mod 2^64. Its low word by * sx + bx = c_xy + (jy * sx + jx) must not wrap for any in-array (jx, jy); a row
    // stride a few bytes longer moves c_xy out of the way. VR5_P2_WIDE=1 forces the variant for small volumes (tests).
    const bool wide = size >= (1ull << 32) || r->opt.p2_wide > 0;
    if (wide && r->opt.p2_wide < 0) return 0;   // VR5_P2_WIDE=-1: volumes beyond 4 GiB go through the texture alone
    uint32_t c_xy = 0;
    if (wide) {
        bool found = false;
        for (int tries = 0; tries < 64 && !found; tries++, sx += 2) {
            c_xy = by * sx + bx;                                   // mod 2^32
            found = (uint64_t)c_xy + (uint64_t)rows * sx < (1ull << 32) && (uint64_t)rows * sx < (1ull << 32);
            if (found) break;
        }
        if (!found) return 0;
        size = (uint64_t)sz * rows * sx;
    }
    const uint32_t sxy = rows * sx;
    // mod 2^32: where the array must start inside a 4-GiB window for the biases of the three float patterns to cancel
    const uint32_t a0_lo = wide ? (uint32_t)((uint64_t)c_xy + (uint64_t)bz * sxy) : bx + by * sx + bz * sxy;
    if (!wide && (uint64_t)a0_lo + size > (1ull << 32)) return 0;
    CUmemAllocationProp prop;
    memset(&prop, 0, sizeof(prop));
    prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    prop.location.id = r->device;
    size_t gran = 0;
    if (v.granularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_MINIMUM) != CUDA_SUCCESS || gran == 0) return 0;
    CUdeviceptr va = 0;
    const size_t four_gib = 1ull << 32;
    const size_t va_size = wide ? (((uint64_t)a0_lo + size + gran + four_gib - 1) / four_gib) * four_gib : four_gib;
    if (v.reserve(&va, va_size, four_gib, 0, 0) != CUDA_SUCCESS) return 0;
    r->p2_va = va; r->p2_va_size = va_size;
    if (va & (four_gib - 1)) { free_p2(r); return 0; }
    const uint64_t map_lo = a0_lo & ~(uint64_t)(gran - 1);
    uint64_t map_hi = ((uint64_t)a0_lo + size + gran - 1) & ~(uint64_t)(gran - 1);
    if (map_hi > va_size) map_hi = va_size;
    CUmemGenericAllocationHandle h = 0;
    if (v.create(&h, map_hi - map_lo, &prop, 0) != CUDA_SUCCESS) { free_p2(r); return 0; }
    r->p2_handle = h;
    if (v.map(va + map_lo, map_hi - map_lo, 0, h, 0) != CUDA_SUCCESS) { free_p2(r); return 0; }
    r->p2_map_addr = va + map_lo; r->p2_map_size = map_hi - map_lo;
    CUmemAccessDesc acc;
    memset(&acc, 0, sizeof(acc));
    acc.location = prop.location;
    acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    if (v.set_access(va + map_lo, map_hi - map_lo, &acc, 1) != CUDA_SUCCESS) { free_p2(r); return 0; }
    uint8_t *a0 = reinterpret_cast<uint8_t *>((uintptr_t)(va + a0_lo));
    p2_layout_kernel<<<r->num_sms * 8, 256, 0, r->stream>>>(r->d_vol_linear, a0, sx, rows, sz, (int)G, r->nx, r->ny, r->nz);
    r->launches++;
    VR5_CUDA(r, cudaGetLastError());
    r->p2_hi = (uint32_t)((uint64_t)va >> 32);
    if (wide) {
        // {low, hi} + bz * sxy must land on a0 + (jz * sxy + jy * sx + jx): hi = high word of a0 - c_xy - Bz * sxy (mod 2^64),
        // whose low word is zero by the choice of a0_lo
        const uint64_t k = (uint64_t)(va + a0_lo) - (uint64_t)c_xy - (uint64_t)bz * sxy;
        if ((uint32_t)k != 0u) { free_p2(r); vr5_fail(r, "p2 wide window: start address is not congruent", cudaSuccess, __FILE__, __LINE__); return 1; }
        r->p2_hi = (uint32_t)(k >> 32);
    }
    r->p2_mx = ldexpf(1.0f, 23 - (ex - 1)) + ldexpf((float)G, -(ex - 1));
    r->p2_my = ldexpf(1.0f, 23 - (ey - 1)) + ldexpf((float)G, -(ey - 1));
    r->p2_mz = ldexpf(1.0f, 23 - (ez - 1)) + ldexpf((float)G, -(ez - 1));
    uint32_t cx, cy, cz;
    memcpy(&cx, &r->p2_mx, 4); memcpy(&cy, &r->p2_my, 4); memcpy(&cz, &r->p2_mz, 4);
    if (cx != bx + G || cy != by + G || cz != bz + G) { free_p2(r); vr5_fail(r, "p2 window: magic constants are not exact", cudaSuccess, __FILE__, __LINE__); return 1; }
    // texture over the same coordinates (index + G), copied out of the window (pitch 2^wl)
    cudaChannelFormatDesc desc = cudaCreateChannelDesc(8, 0, 0, 0, cudaChannelFormatKindUnsigned);
    cudaExtent ext = make_cudaExtent(r->nx + 2 * G, r->ny + 2 * G, r->nz + 2 * G);
    cudaError_t e = cudaMalloc3DArray(&r->d_array_p2, &desc, ext, 0);
    if (e != cudaSuccess) { cudaGetLastError(); r->d_array_p2 = nullptr; free_p2(r); return 0; }
    cudaMemcpy3DParms cp;
    memset(&cp, 0, sizeof(cp));
    cp.srcPtr = make_cudaPitchedPtr(a0, sx, r->nx + 2 * G, rows);
    cp.dstArray = r->d_array_p2;
    cp.extent = ext;
    cp.kind = cudaMemcpyDeviceToDevice;
    e = cudaMemcpy3DAsync(&cp, r->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(r->stream);
    if (e != cudaSuccess) { free_p2(r); vr5_fail(r, "p2 texture upload", e, __FILE__, __LINE__); return 1; }
    cudaResourceDesc res;
    memset(&res, 0, sizeof(res));
    res.resType = cudaResourceTypeArray;
    res.res.array.array = r->d_array_p2;
    cudaTextureDesc td;
    memset(&td, 0, sizeof(td));
    td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
    td.filterMode = cudaFilterModePoint;
    td.readMode = cudaReadModeElementType;
    td.normalizedCoords = 0;
    VR5_CUDA(r, cudaCreateTextureObject(&r->tex_p2, &res, &td, nullptr));
    r->p2_wl = wide ? 2 : 1; r->p2_sx = sx; r->p2_sxy = sxy;   // 2: only kP2Wide can address it
    return 0;
}

static inline bool is_trilinear(int sampler) { return sampler == VR5_SAMPLER_TRILINEAR || sampler == VR5_SAMPLER_TRILINEAR_FMA; }

// per-brick maxima for the leaps of the zero-skip kernels: brick edge 2^lb voxels, lb >= 5 and at most 32 bricks per axis by default (a table
// of <= 32 KiB, L1-resident: an attempt costs ~100 instructions, so a leap must be worth a few dozen samples), VR5_LEAP_BRICK = log2 of the edge overrides it; VR5_NO_LEAP: no table, no leaps
static int build_brick_max(vr5_renderer *r) {
    if (r->opt.no_leap) return 0;
    const uint32_t nmax = std::max<uint32_t>(r->nx, std::max<uint32_t>(r->ny, r->nz));
    uint32_t lb = 5;   // edge 32 at least: a hop of the brick-to-brick walk costs ~25 instructions and a table look-up (C4: edge 16 gains nothing, 8 loses)
    while (((nmax + (1u << lb) - 1u) >> lb) > 32u) lb++;
    if (r->opt.leap_brick_log2 >= 2 && r->opt.leap_brick_log2 <= 10) lb = (uint32_t)r->opt.leap_brick_log2;
    r->brick_lb = lb;
    r->brick_nx = (r->nx + (1u << lb) - 1u) >> lb; r->brick_ny = (r->ny + (1u << lb) - 1u) >> lb; r->brick_nz = (r->nz + (1u << lb) - 1u) >> lb;
    const uint64_t bytes = (uint64_t)r->brick_nx * r->brick_ny * r->brick_nz;
    cudaError_t e = cudaMalloc(&r->d_brick_max, bytes);
    if (e != cudaSuccess) { vr5_fail(r, "cudaMalloc brick maxima", e, __FILE__, __LINE__); cudaGetLastError(); r->d_brick_max = nullptr; return 1; }
    brick_max_kernel<<<r->num_sms * 8, 256, 0, r->stream>>>(r->d_vol_linear, r->d_brick_max, r->nx, r->ny, r->nz, lb, r->brick_nx, r->brick_ny, r->brick_nz);
    r->launches++;
    VR5_CUDA(r, cudaGetLastError());
    VR5_CUDA(r, cudaStreamSynchronize(r->stream));
    return 0;
}

static int build_layout(vr5_renderer *r) {
    free_layouts(r);
    if (!r->d_vol_linear) return 0;
    if (!is_trilinear(r->sampler) && build_brick_max(r)) return 1;
    if (is_trilinear(r->sampler)) {
        const uint64_t bytes = (uint64_t)(r->nx + 1 + 2 * kTriGuard) * (r->ny + 1 + 2 * kTriGuard) * (r->nz + 1 + 2 * kTriGuard) * 8u;
        // the unclamped march wants the array inside one 4-GiB address window: an allocation that straddles one is retried
        uint8_t *held[4] = {nullptr, nullptr, nullptr, nullptr};
        cudaError_t e = cudaSuccess;
        for (int tries = 0; tries < 4; tries++) {
            e = cudaMalloc(&r->d_cells, bytes);
            if (e != cudaSuccess) { r->d_cells = nullptr; break; }
            const uint64_t a0 = (uint64_t)(uintptr_t)r->d_cells;
            if (bytes >= (1ull << 32) || (a0 >> 32) == ((a0 + bytes - 1) >> 32) || tries == 3) break;
            held[tries] = r->d_cells; r->d_cells = nullptr;
        }
        for (uint8_t *h : held) if (h) cudaFree(h);
        if (!r->d_cells) { vr5_fail(r, "cudaMalloc trilinear cells (8 bytes per voxel)", e, __FILE__, __LINE__); cudaGetLastError(); return 1; }
        r->cells_bytes = bytes;
        cell_layout_kernel<<<r->num_sms * 8, 256, 0, r->stream>>>(r->d_vol_linear, reinterpret_cast<uint2 *>(r->d_cells), r->nx, r->ny, r->nz);
        r->launches++;
        VR5_CUDA(r, cudaGetLastError());
        VR5_CUDA(r, cudaStreamSynchronize(r->stream));
        return 0;
    }
    if (r->sampler >= VR5_SAMPLER_HYBRID_WINDOW) {
        if (!r->opt.no_p2 && build_p2(r)) return 1;
        if (!r->p2_wl && build_window(r)) return 1;
        VR5_CUDA(r, cudaStreamSynchronize(r->stream));
    }
    // the unclamped LDG path needs the slice stride of the padded grid in 32 bits
    const bool pad_ok = (uint64_t)(r->nx + 1) * (r->ny + 1) < (1ull << 32);
    // (volumes too big for a 4-GiB window, 2048^3, run the window policies through the texture alone: the 64-bit hybrid
    // measured slower there, 51.6 against 48.5 ms per C5 frame)
    if ((r->sampler == VR5_SAMPLER_LINEAR || r->sampler == VR5_SAMPLER_HYBRID) && pad_ok) {
        if (build_padded(r)) return 1;
        VR5_CUDA(r, cudaStreamSynchronize(r->stream));
    }
    if (r->sampler == VR5_SAMPLER_BRICK) {
        r->bxn = (r->nx + 7u) / 8u; r->byn = (r->ny + 3u) / 4u; r->bzn = (r->nz + 3u) / 4u;
        const uint64_t bytes = (uint64_t)r->bxn * r->byn * r->bzn * 128u;
        cudaError_t e = cudaMalloc(&r->d_vol_brick, bytes);
        if (e != cudaSuccess) { vr5_fail(r, "cudaMalloc bricked volume", e, __FILE__, __LINE__); cudaGetLastError(); return 1; }
        brick_layout_kernel<<<r->num_sms * 8, 256, 0, r->stream>>>(r->d_vol_linear, r->d_vol_brick, r->nx, r->ny, r->nz, r->bxn, r->byn, r->bzn);
        r->launches++;
        VR5_CUDA(r, cudaGetLastError());
        VR5_CUDA(r, cudaStreamSynchronize(r->stream));
    } else if (r->sampler >= VR5_SAMPLER_TEX3D) {
        cudaChannelFormatDesc desc = cudaCreateChannelDesc(8, 0, 0, 0, cudaChannelFormatKindUnsigned);
        cudaExtent ext = make_cudaExtent(r->nx + 1, r->ny + 1, r->nz + 1);  // one replicated voxel of padding per axis
        cudaError_t e = cudaMalloc3DArray(&r->d_array, &desc, ext, 0);
        if (e != cudaSuccess) { vr5_fail(r, "cudaMalloc3DArray", e, __FILE__, __LINE__); cudaGetLastError(); return 1; }
        const bool keep_padded = r->d_vol_padded != nullptr;
        if (!keep_padded && build_padded(r)) return 1;
        uint8_t *d_padded = r->d_vol_padded;
        cudaMemcpy3DParms cp;
        memset(&cp, 0, sizeof(cp));
        cp.srcPtr = make_cudaPitchedPtr(d_padded, r->nx + 1, r->nx + 1, r->ny + 1);
        cp.dstArray = r->d_array;
        cp.extent = ext;
        cp.kind = cudaMemcpyDeviceToDevice;
        e = cudaMemcpy3DAsync(&cp, r->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(r->stream);
        if (!keep_padded) { cudaFree(r->d_vol_padded); r->d_vol_padded = nullptr; }
        if (e != cudaSuccess) { vr5_fail(r, "3-D texture upload", e, __FILE__, __LINE__); return 1; }
        cudaResourceDesc res;
        memset(&res, 0, sizeof(res));
        res.resType = cudaResourceTypeArray;
        res.res.array.array = r->d_array;
        cudaTextureDesc td;
        memset(&td, 0, sizeof(td));
        td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
        td.filterMode = cudaFilterModePoint;
        td.readMode = cudaReadModeElementType;
        td.normalizedCoords = 0;
        VR5_CUDA(r, cudaCreateTextureObject(&r->tex, &res, &td, nullptr));
    }
    return 0;
}

static int install_volume(vr5_renderer *r, const uint8_t *src, cudaMemcpyKind kind, uint64_t bytes, uint16_t dx, uint16_t dy, uint16_t dz) {
    VR5_CUDA(r, cudaSetDevice(r->device));
    VR5_CUDA(r, sync_all(r));
    free_layouts(r);
    if (r->d_vol_linear) { cudaFree(r->d_vol_linear); r->d_vol_linear = nullptr; }
    r->vol_bytes = 0;
    if (!src) { vr5_fail(r, "vr5_set_volume: no data", cudaSuccess, __FILE__, __LINE__); return 1; }  // GPURenderer1.cu:91-92
    if (dx == 0 || dy == 0 || dz == 0 || bytes != (uint64_t)dx * dy * dz) {
        vr5_fail(r, "vr5_set_volume: bytes != dx*dy*dz", cudaSuccess, __FILE__, __LINE__);
        return 1;
    }
    cudaError_t e = cudaMalloc(&r->d_vol_linear, bytes);
    if (e != cudaSuccess) {  // GPURenderer1.cu:93-95: allocation failure is a soft error
        vr5_fail(r, "cudaMalloc volume (data is probably too large for device memory)", e, __FILE__, __LINE__);
        cudaGetLastError();
        r->d_vol_linear = nullptr;
        return 1;
    }
    if (src != r->d_vol_linear) VR5_CUDA(r, cudaMemcpyAsync(r->d_vol_linear, src, bytes, kind, r->stream));
    VR5_CUDA(r, cudaStreamSynchronize(r->stream));
    r->vol_bytes = bytes; r->nx = dx; r->ny = dy; r->nz = dz;
    return build_layout(r);
}

extern "C" int vr5_set_volume(vr5_renderer *r, const uint8_t *host, uint64_t bytes, uint16_t dx, uint16_t dy, uint16_t dz) {
    if (!r) return 1;
    return install_volume(r, host, cudaMemcpyHostToDevice, bytes, dx, dy, dz);
}

extern "C" int vr5_set_volume_device(vr5_renderer *r, const uint8_t *dev, uint64_t bytes, uint16_t dx, uint16_t dy, uint16_t dz) {
    if (!r) return 1;
    return install_volume(r, dev, cudaMemcpyDeviceToDevice, bytes, dx, dy, dz);
}

// adopt a linear device buffer allocated by this library (device generator) without copying
int vr5_adopt_volume(vr5_renderer *r, uint8_t *dev_owned, uint64_t bytes, uint16_t dx, uint16_t dy, uint16_t dz) {
    VR5_CUDA(r, sync_all(r));
    free_layouts(r);
    if (r->d_vol_linear) cudaFree(r->d_vol_linear);
    r->d_vol_linear = dev_owned; r->vol_bytes = bytes; r->nx = dx; r->ny = dy; r->nz = dz;
    return build_layout(r);
}

extern "C" int vr5_get_volume(vr5_renderer *r, uint8_t *host_out, uint64_t bytes) {
    if (!r) return 1;
    if (!r->d_vol_linear || !host_out || bytes != r->vol_bytes) { vr5_fail(r, "vr5_get_volume: no volume / size mismatch", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    VR5_CUDA(r, cudaMemcpyAsync(host_out, r->d_vol_linear, bytes, cudaMemcpyDeviceToHost, r->stream));
    VR5_CUDA(r, cudaStreamSynchronize(r->stream));
    return 0;
}

extern "C" int vr5_set_sampler(vr5_renderer *r, int sampler) {
    if (!r) return 1;
    if (sampler < 0 || (sampler > VR5_SAMPLER_HYBRID_WINDOW2 && !is_trilinear(sampler))) { vr5_fail(r, "vr5_set_sampler: unknown sampler", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    VR5_CUDA(r, sync_all(r));
    if (sampler == r->sampler && (r->d_vol_padded || r->d_vol_brick || r->tex || r->d_vol_window || r->p2_wl || r->d_cells || !r->d_vol_linear)) return 0;
    r->sampler = sampler;
    return build_layout(r);
}
extern "C" int vr5_get_sampler(const vr5_renderer *r) { return r ? r->sampler : -1; }
extern "C" int vr5_last_variant(const vr5_renderer *r) { return r ? r->last_variant : -1; }
extern "C" int vr5_last_zero_skip(const vr5_renderer *r) { return r ? r->last_zskip : -1; }

extern "C" int vr5_set_partition(vr5_renderer *r, int rank, int world, int packed) {
    if (!r) return 1;
    if (world < 1 || rank < 0 || rank >= world) { vr5_fail(r, "vr5_set_partition: need 0 <= rank < world", cudaSuccess, __FILE__, __LINE__); return 1; }
    r->rank = rank; r->world = world; r->packed = packed ? 1 : 0;
    r->band_first = 0; r->band_count = -1; r->band_auto = false;
    return 0;
}

extern "C" int vr5_set_partition_band(vr5_renderer *r, int first_strip, int n_strips) {
    if (!r) return 1;
    if (first_strip < 0 || n_strips < 0) { vr5_fail(r, "vr5_set_partition_band: need first_strip >= 0 and n_strips >= 0", cudaSuccess, __FILE__, __LINE__); return 1; }
    r->packed = 0; r->band_first = first_strip; r->band_count = n_strips; r->band_auto = false;
    return 0;
}

// Cost-balanced contiguous bands (north_star: "each GPU renders a load-balanced row strip"). The cost of a 4-row strip is
// estimated from the clipped length of a sparse grid of its rays (Raycaster::intersect in plain fp32: k_far - k_near in steps,
// plus a per-ray constant for set-up), interpolated over the strips and cut into `world` equal parts. The estimate sees the
// geometry only - exact for transfer functions without early termination (the semi-transparent bench TF), a proxy otherwise
// (an opaque TF is better served by interleaved strips). Pure host arithmetic on the parameters: every rank computes the same
// bounds from the same bytes. ~400 rays per frame, a few microseconds.
extern "C" int vr5_balance_bands(const vr5_params *p, int world, uint32_t *first_strip_out) {
    if (!p || world < 1 || !first_strip_out) return 1;
    const vr5_view &v = p->view;
    const int w = v.dims.x, h = v.dims.y, strips = (h + kTileH - 1) / kTileH;
    if (w < 1 || h < 1) return 1;
    constexpr int kRows = 24, kCols = 12;   // 288 rays: the cost profile over the rows is piecewise quadratic, a coarse grid pins it to ~1 %
    const int nr = strips < kRows ? strips : kRows;
    float cost[kRows], pos[kRows];   // cost density per strip at strip position pos[i]
    const float inv_step = p->ray_step > 0.0f && std::isfinite(p->ray_step) ? 1.0f / p->ray_step : 1.0f;
    const float mb[3] = {p->min_bound.x, p->min_bound.y, p->min_bound.z};
    const float o0[3] = {v.origin.x, v.origin.y, v.origin.z}, d0[3] = {v.direction.x, v.direction.y, v.direction.z};
    const float rr[3] = {v.right_plane.x, v.right_plane.y, v.right_plane.z}, uu[3] = {v.up_plane.x, v.up_plane.y, v.up_plane.z};
    float inv0[3];   // orthographic: one direction for all rays
    for (int a = 0; a < 3; a++) inv0[a] = 1.0f / (d0[a] == 0.0f ? 0.00001f : d0[a]);
    float cxs[kCols];
    for (int j = 0; j < kCols; j++) cxs[j] = (float)((2 * j + 1) * w / (2 * kCols) - w / 2);
    for (int i = 0; i < nr; i++) {
        pos[i] = ((float)i + 0.5f) * (float)strips / (float)nr;
        const float cy = (float)((int)pos[i] * kTileH + kTileH / 2 - h / 2);
        float sum = 0.0f;
        for (int j = 0; j < kCols; j++) {
            const float cx = cxs[j];
            float kn = -INFINITY, kf = INFINITY;
            for (int a = 0; a < 3; a++) {
                const float off = rr[a] * cx + uu[a] * cy;
                float o = o0[a], inv = inv0[a];
                if (v.perspective) { const float d = d0[a] + off; inv = 1.0f / (d == 0.0f ? 0.00001f : d); } else o += off;
                const float k1 = (mb[a] - o) * inv, k2 = (-mb[a] - o) * inv;
                const float lo = k1 < k2 ? k1 : k2, hi = k1 < k2 ? k2 : k1;   // (ternaries: fminf / fmaxf are libm calls without -ffinite-math-only)
                kn = lo > kn ? lo : kn; kf = hi < kf ? hi : kf;
            }
            kn = kn > 0.0f ? kn : 0.0f;
            const float len = (kf - kn) * inv_step;
            sum += 16.0f + ((kn < kf && kf > 0.0f) ? (len < 1.0e6f ? len : 1.0e6f) : 0.0f);   // set-up + samples
        }
        cost[i] = std::isfinite(sum) && sum > 0.0f ? sum : 16.0f * kCols;
    }
    // The density is constant before the first and after the last sample and linear in between; its integral is cut into `world`
    // equal parts analytically (a quadratic per segment): no per-strip work. ~4 us per call.
    double seg_area[kRows + 1], total = 0.0;   // segment k: [x_{k-1}, x_k] with x_{-1} = 0, x_nr = strips
    for (int k = 0; k <= nr; k++) {
        const double x0 = k == 0 ? 0.0 : pos[k - 1], x1 = k == nr ? (double)strips : pos[k];
        const double c0 = cost[k == 0 ? 0 : k - 1], c1 = cost[k == nr ? nr - 1 : k];
        seg_area[k] = 0.5 * (c0 + c1) * (x1 - x0);
        total += seg_area[k];
    }
    first_strip_out[0] = 0;
    int k = 0;
    double acc = 0.0;
    for (int r = 1; r < world; r++) {
        const double target = total * r / world;
        while (k < nr && acc + seg_area[k] < target) acc += seg_area[k++];
        const double x0 = k == 0 ? 0.0 : pos[k - 1], x1 = k == nr ? (double)strips : pos[k];
        const double c0 = cost[k == 0 ? 0 : k - 1], c1 = cost[k == nr ? nr - 1 : k], len = x1 - x0, rem = target - acc;
        double t;   // c0 t + (c1 - c0) t^2 / (2 len) = rem
        const double slope = len > 0.0 ? (c1 - c0) / len : 0.0;
        if (fabs(slope) < 1e-12 * (c0 + c1)) t = c0 > 0.0 ? rem / c0 : 0.0;
        else t = (-c0 + sqrt(fmax(c0 * c0 + 2.0 * slope * rem, 0.0))) / slope;
        int cut = (int)floor(x0 + t + 0.5);
        const int lo = (int)first_strip_out[r - 1] + (strips >= world ? 1 : 0), hi = strips - (strips >= world ? world - r : 0);   // every band at least one strip when there are enough
        cut = cut < lo ? lo : cut > hi ? hi : cut;
        first_strip_out[r] = (uint32_t)cut;
    }
    first_strip_out[world] = (uint32_t)strips;
    return 0;
}

extern "C" int vr5_set_partition_bands(vr5_renderer *r, int rank, int world) {
    if (!r) return 1;
    if (world < 1 || world > 64 || rank < 0 || rank >= world) { vr5_fail(r, "vr5_set_partition_bands: need 0 <= rank < world <= 64", cudaSuccess, __FILE__, __LINE__); return 1; }
    r->rank = rank; r->world = world; r->packed = 0;
    r->band_first = 0; r->band_count = 0; r->band_auto = true;
    return 0;
}

extern "C" uint32_t vr5_strip_rows(uint16_t height, int rank, int world) {
    if (world < 1 || rank < 0 || rank >= world) return 0;
    const uint32_t strips = (height + kTileH - 1) / kTileH;
    if ((uint32_t)rank >= strips) return 0;
    const uint32_t owned = (strips - rank + world - 1) / world;
    const uint32_t last_strip = rank + (owned - 1) * world;
    const uint32_t last_rows = (last_strip == strips - 1 && height % kTileH) ? height % kTileH : kTileH;
    return (owned - 1) * kTileH + last_rows;
}

// ---- launch ------------------------------------------------------------------------------------------------------
template <int S, bool C, bool SH, bool CL, bool ZS = false>
static cudaError_t launch_one(vr5_renderer *r, const KArgs &a) {
    if constexpr (!ZS && !C) if (a.zero_lim) return launch_one<S, false, SH, CL, true>(r, a);   // transparent TF prefix: the zero-skip build of the same kernel
    int &per_sm = r->ctas_per_sm[((S * 2 + (C ? 1 : 0)) * 2 + (SH ? 1 : 0)) * 2 + (CL ? 1 : 0) + (ZS ? 128 : 0)];  // S <= 10: 0..87, ZS 128..215
    if (per_sm == 0) {
        cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, cast_kernel<S, C, SH, CL, ZS>, kCtaThreads, 0);
        if (e != cudaSuccess) return e;
        if (per_sm < 1) per_sm = 1;
    }
    uint32_t grid = (uint32_t)(r->num_sms * per_sm);  // persistent: every CTA resident, a multiple of the SM count
    const uint32_t needed = (a.n_queue * a.n_frames + (kCtaThreads / 32) - 1) / (kCtaThreads / 32);
    if (grid > needed) grid = needed ? needed : 1;
    cast_kernel<S, C, SH, CL, ZS><<<grid, kCtaThreads, 0, r->launch_stream>>>(a);
    r->last_zskip = ZS ? 1 : 0;
    r->last_grid = grid;
    r->launches++;
    r->last_variant = S;
    return cudaGetLastError();
}

template <int S, bool CL>
static cudaError_t launch_sc(vr5_renderer *r, const KArgs &a, bool count, bool shade) {
    if (count) return shade ? launch_one<S, true, true, CL>(r, a) : launch_one<S, true, false, CL>(r, a);
    return shade ? launch_one<S, false, true, CL>(r, a) : launch_one<S, false, false, CL>(r, a);
}

// Order in which the tile queue hands out its 32x32-pixel groups. A persistent grid idles at the end of a launch behind its
// last full-length tiles (a C3 tile is ~0.3 ms of one warp's time; ncu: SMs active 94 % of a lone launch), so a launch would
// like to END on cheap tiles. Two orders that do were measured against the row-major sweep (profiles/r01_sampler_choice.md):
// "cols" sweeps along the longer image side (its far end holds empty or short silhouette rays: the cube's projection never
// exceeds 0.87 x the shorter side from the centre, ViewBase.cpp:49-55, :100-105); "rings" walks rings around the image centre
// inside out. Both cut the tail of a LONE launch (trilinear-fma: 1.275 -> 1.224 ms with rings) and both lose in the overlapped
// stream of frames, where the next frame fills the tail anyway and their wider working set costs L1/L2 hits (nearest: 0.561 ->
// 0.577 / 0.582 ms per frame). Default therefore: row-major, no table; VR5_QUEUE_ORDER=cols|rings selects the others.
// Experiment knobs (names = the environment variables vr5_create reads once). Returns 1 for an unknown name or a bad value.
extern "C" int vr5_set_option(vr5_renderer *r, const char *name, const char *value) {
    if (!r || !name || !value) return 1;
    auto truthy = [&]() { return value[0] != '\0' && strcmp(value, "0") != 0; };
    if (!strcmp(name, "VR5_QUEUE_ORDER")) {
        if (!strcmp(value, "rows")) r->opt.queue_order = 0;
        else if (!strcmp(value, "cols")) r->opt.queue_order = 1;
        else if (!strcmp(value, "rings")) r->opt.queue_order = 2;
        else { vr5_fail(r, "vr5_set_option: VR5_QUEUE_ORDER is rows, cols or rings", cudaSuccess, __FILE__, __LINE__); return 1; }
    } else if (!strcmp(name, "VR5_ROW_MAJOR")) { if (truthy()) r->opt.queue_order = 0; }
    else if (!strcmp(name, "VR5_WARP_FETCH")) r->opt.warp_fetch = truthy();
    else if (!strcmp(name, "VR5_TRI_CLAMPED")) r->opt.tri_clamped = truthy();
    else if (!strcmp(name, "VR5_TRIF_CTAS")) r->opt.trif_ctas = atoi(value);
    else if (!strcmp(name, "VR5_TRIF_DEPTH")) r->opt.trif_depth = atoi(value);
    else if (!strcmp(name, "VR5_ESL_EXACT")) r->opt.esl_exact = truthy();
    else if (!strcmp(name, "VR5_SEQ_DEBUG")) r->opt.seq_debug = atoi(value);
    else if (!strcmp(name, "VR5_HOST_SINGLE")) r->opt.host_single = truthy();
    else if (!strcmp(name, "VR5_NO_ZSKIP")) r->opt.no_zskip = truthy();
    else if (!strcmp(name, "VR5_NO_LEAP")) r->opt.no_leap = truthy();           // takes effect at the next layout build
    else if (!strcmp(name, "VR5_LEAP_STATS")) r->opt.leap_stats = truthy();
    else if (!strcmp(name, "VR5_LEAP_BRICK")) r->opt.leap_brick_log2 = atoi(value);   // log2 of the brick edge; next layout build
    else if (!strcmp(name, "VR5_NO_P2")) r->opt.no_p2 = truthy();        // takes effect at the next layout build
    else if (!strcmp(name, "VR5_P2_WIDE")) r->opt.p2_wide = strtol(value, nullptr, 10);   // 1 = force, -1 = never; next layout build
    else if (!strcmp(name, "VR5_WIN_SX")) r->opt.win_sx = (uint32_t)strtoul(value, nullptr, 10);
    else { vr5_fail(r, "vr5_set_option: unknown option", cudaSuccess, __FILE__, __LINE__); return 1; }
    return 0;
}

static int build_group_order(vr5_renderer *r, const KArgs *a, int mode) {
    const uint32_t key[6] = {(uint32_t)a->width, (uint32_t)a->height, (uint32_t)r->rank, (uint32_t)r->world, (uint32_t)r->packed, (uint32_t)mode};
    if (r->d_group_order && !memcmp(key, r->order_key, sizeof(key))) return 0;
    const uint32_t n = a->groups_x * a->groups_y;
    std::vector<uint32_t> order(n);
    if (mode == 1) {
        for (uint32_t gx = 0, i = 0; gx < a->groups_x; gx++)
            for (uint32_t gy = 0; gy < a->groups_y; gy++) order[i++] = gy * a->groups_x + gx;
    } else {
        struct Item { int ring; float angle; uint32_t group; };
        std::vector<Item> items(n);
        const float gw = (float)(kTileW << kGroupLog2X);  // group width in pixels
        for (uint32_t g = 0; g < n; g++) {
            const uint32_t gx = g % a->groups_x, gy = g / a->groups_x;
            const float cx = ((float)gx + 0.5f) * gw - 0.5f * (float)a->width;
            // a group's local tile rows are strips of this rank: global row = (local tile row * world + rank) * kTileH
            const float cy = (((float)gy + 0.5f) * (float)(1u << kGroupLog2Y) * (float)r->world + (float)r->rank) * (float)kTileH - 0.5f * (float)a->height;
            items[g] = Item{(int)(sqrtf(cx * cx + cy * cy) / 64.0f), atan2f(cy, cx), g};
        }
        std::sort(items.begin(), items.end(), [](const Item &x, const Item &y) {
            return x.ring != y.ring ? x.ring < y.ring : x.angle != y.angle ? x.angle < y.angle : x.group < y.group;
        });
        for (uint32_t i = 0; i < n; i++) order[i] = items[i].group;
    }
    sync_all(r);  // frames in flight may still read the previous table
    if (r->d_group_order) { cudaFree(r->d_group_order); r->d_group_order = nullptr; }
    VR5_CUDA(r, cudaMalloc(&r->d_group_order, (size_t)n * sizeof(uint32_t)));
    VR5_CUDA(r, cudaMemcpy(r->d_group_order, order.data(), (size_t)n * sizeof(uint32_t), cudaMemcpyHostToDevice));
    memcpy(r->order_key, key, sizeof(key));
    return 0;
}

// Largest displacement of a march step in voxels, on any axis, over all rays of a frame (d is affine in the pixel position,
// so the four image corners bound it). The unclamped kernel variants fetch up to a few steps past a ray's clip interval and
// rely on the guard margins of their layouts covering that reach: the launcher compares this figure with the margin
// (reach_ok). +inf - which sends the frame through the clamped variants - for anything that is not finite and for DEGENERATE
// views: Raycaster::intersect replaces an exactly-zero direction component by 1e-5 on a local copy only (RaycasterBase.h:
// 33-35), so a ray whose origin lies outside [-1, 1] on that axis can receive a valid interval at k ~ 1e5 x distance while
// its points stay outside the cube on that axis - possible only when every other component is tiny too (far_j <= (1 +
// |o_j|) / |d_j| must reach such a k). The reference's cameras never produce that (|d| ~ 1), but the ABI accepts any View
// bytes, so frames in which |d|_inf could fall below the bound derived here take the clamped path.
static float frame_reach(const KArgs &a) {
    const float inf = std::numeric_limits<float>::infinity();
    if (!(a.step > 0.0f) || !std::isfinite(a.step)) return inf;
    const float o[12] = {a.ox, a.oy, a.oz, a.rx, a.ry, a.rz, a.ux, a.uy, a.uz, a.dx, a.dy, a.dz};
    for (float v : o) if (!std::isfinite(v)) return inf;
    float reach = 0.0f, omax = 0.0f;
    for (int c = 0; c < 4; c++) {
        const float cx = (c & 1) ? (float)(a.width - 1 - a.half_w) : (float)(-a.half_w);
        const float cy = (c & 2) ? (float)(a.height - 1 - a.half_h) : (float)(-a.half_h);
        float dx = a.dx, dy = a.dy, dz = a.dz, ox = a.ox, oy = a.oy, oz = a.oz;
        if (a.perspective) { dx += a.rx * cx + a.ux * cy; dy += a.ry * cx + a.uy * cy; dz += a.rz * cx + a.uz * cy; }
        else { ox += a.rx * cx + a.ux * cy; oy += a.ry * cx + a.uy * cy; oz += a.rz * cx + a.uz * cy; }
        reach = fmaxf(reach, fmaxf(fabsf(dx) * a.step * a.hx, fmaxf(fabsf(dy) * a.step * a.hy, fabsf(dz) * a.step * a.hz)));
        omax = fmaxf(omax, fmaxf(fabsf(ox), fmaxf(fabsf(oy), fabsf(oz))));
    }
    if (!std::isfinite(reach) || !std::isfinite(omax)) return inf;
    // lower bound of |d|_inf over the frame: the direction itself (orthographic), or the distance of the origin from the
    // plane { direction + right * cx + up * cy } divided by sqrt(3) (perspective)
    float dmin;
    if (a.perspective) {
        const float nx = a.ry * a.uz - a.rz * a.uy, ny = a.rz * a.ux - a.rx * a.uz, nz = a.rx * a.uy - a.ry * a.ux;
        const float nn = sqrtf(nx * nx + ny * ny + nz * nz);
        dmin = nn > 0.0f ? fabsf(a.dx * nx + a.dy * ny + a.dz * nz) / nn * 0.57735f : 0.0f;
    } else {
        dmin = fmaxf(fabsf(a.dx), fmaxf(fabsf(a.dy), fabsf(a.dz)));
    }
    // a degenerate hit needs k >= (|o_axis| - 1) / 1e-5 with the point more than (guard - 2) voxels outside the grid, i.e.
    // |o_axis| - 1 > 2 (guard - 2) / n; and far_j >= k for every other axis, i.e. |d_j| <= (1 + |o_j|) * 1e-5 * n / (2 (guard - 2))
    const float guard = (float)((kWindowGuard < kTriGuard ? kWindowGuard : kTriGuard) - 2);
    const float bound = (1.0f + omax) * 1e-5f * fmaxf(a.hx, fmaxf(a.hy, a.hz)) / guard;
    if (!(dmin > 2.0f * bound)) return inf;
    return reach;
}

// tile-queue geometry from tiles_x and tile_rows_local: 32x32-pixel groups, slots per frame, the divide-by-multiply constant
static void set_queue_geometry(KArgs *a) {
    a->groups_x = (a->tiles_x + (1u << kGroupLog2X) - 1) >> kGroupLog2X;
    if (a->groups_x < 2) a->groups_x = 2;   // (the slots of the padding group lie outside the image and are skipped) keeps the magic below in 32 bits
    a->groups_y = (a->tile_rows_local + (1u << kGroupLog2Y) - 1) >> kGroupLog2Y;
    a->n_queue = (a->groups_x * a->groups_y) << (kGroupLog2X + kGroupLog2Y);
    // group / groups_x = umulhi(group, magic), exact for group < 2^16; 0 = more groups than that (beyond 8k x 8k pixels): divide
    a->groups_x_magic = a->groups_x * a->groups_y < 65536u ? (uint32_t)((0x100000000ull + a->groups_x - 1) / a->groups_x) : 0u;
}

static int fill_args(vr5_renderer *r, const vr5_params *p, void *out_dev, KArgs *a) {
    if (!p) { vr5_fail(r, "vr5_render: NULL params", cudaSuccess, __FILE__, __LINE__); return 1; }
    // the reference's own guard (GPURenderer1.cu:101-102): volume, TF + ESL table, buffer
    if (!r->d_vol_linear || !r->have_tf) { vr5_fail(r, "vr5_render: volume or transfer function not set", cudaSuccess, __FILE__, __LINE__); return 1; }
    if (p->view.dims.x != r->width || p->view.dims.y != r->height) { vr5_fail(r, "vr5_render: view dims differ from vr5_set_window", cudaSuccess, __FILE__, __LINE__); return 1; }
    if (p->esl && (p->esl_block_dims == 0 || p->esl_block_dims > 65535u)) { vr5_fail(r, "vr5_render: esl_block_dims must be in 1..65535", cudaSuccess, __FILE__, __LINE__); return 1; }
    memset(a, 0, sizeof(*a));
    a->n_frames = 1;
    const vr5_view &v = p->view;
    a->ox = v.origin.x; a->oy = v.origin.y; a->oz = v.origin.z;
    a->dx = v.direction.x; a->dy = v.direction.y; a->dz = v.direction.z;
    a->rx = v.right_plane.x; a->ry = v.right_plane.y; a->rz = v.right_plane.z;
    a->ux = v.up_plane.x; a->uy = v.up_plane.y; a->uz = v.up_plane.z;
    a->lx = v.light_pos.x; a->ly = v.light_pos.y; a->lz = v.light_pos.z;
    a->width = v.dims.x; a->height = v.dims.y;
    a->half_w = v.dims.x / 2; a->half_h = v.dims.y / 2;
    a->perspective = v.perspective ? 1 : 0;
    a->nx1 = r->nx - 1; a->ny1 = r->ny - 1; a->nz1 = r->nz - 1;
    a->hx = (float)r->nx * 0.5f; a->hy = (float)r->ny * 0.5f; a->hz = (float)r->nz * 0.5f;
    if (r->sampler == VR5_SAMPLER_BRICK) {
        a->vol = r->d_vol_brick; a->sx = r->bxn; a->sxy = (uint64_t)r->bxn * r->byn;
    } else {
        a->vol = r->d_vol_linear; a->sx = r->nx; a->sxy = (uint64_t)r->nx * r->ny;
    }
    a->tex = r->tex;
    a->vol_pad = r->d_vol_padded; a->psx = r->nx + 1; a->psxy = (r->nx + 1) * (r->ny + 1);
    a->whi = r->win_hi; a->wsx = r->win_sx; a->wsxy = r->win_sxy;
    a->wmx = r->win_mx; a->wmy = r->win_my; a->wmz = r->win_mz;
    a->mbx = p->min_bound.x; a->mby = p->min_bound.y; a->mbz = p->min_bound.z;
    a->step = p->ray_step; a->threshold = p->ray_threshold;
    a->kd = p->light_kd;
    a->shade_alpha = p->light_kd > 0.01f ? 0.05f : std::numeric_limits<float>::infinity();
    // ZS kernels: voxel value s classifies to tf[s / 2]; a threshold below zero (or NaN) would end rays on transparent samples
    a->zero_lim = (!r->opt.no_zskip && p->ray_threshold >= 0.0f) ? 2u * (uint32_t)r->tf_zero_prefix : 0u;
    a->esl = p->esl ? 1 : 0;
    const uint32_t bd = p->esl_block_dims ? p->esl_block_dims : 1;
    a->bd_is_one = bd == 1;
    a->bd_magic = bd == 1 ? 0u : (uint32_t)((0x100000000ull + bd - 1) / bd);
    a->bsx = p->esl_block_size.x; a->bsy = p->esl_block_size.y; a->bsz = p->esl_block_size.z;
    a->tf = reinterpret_cast<const float4 *>(r->d_tf);
    a->esl_bits = r->d_esl;
    a->out = reinterpret_cast<uchar4 *>(out_dev);
    a->packed = r->packed; a->rank = r->rank; a->world = r->world;
    if (r->band_auto) {   // cost-balanced bands: this frame's band of this rank
        uint32_t bounds[65];
        if (vr5_balance_bands(p, r->world, bounds)) { vr5_fail(r, "vr5_balance_bands failed", cudaSuccess, __FILE__, __LINE__); return 1; }
        r->band_first = (int)bounds[r->rank]; r->band_count = (int)(bounds[r->rank + 1] - bounds[r->rank]);
    }
    if (r->band_count >= 0) { a->rank = r->band_first; a->world = 1; }   // contiguous band: tile row t of the rank = strip band_first + t
    a->tiles_x = (a->width + kTileW - 1) / kTileW;
    const uint32_t strips = (a->height + kTileH - 1) / kTileH;
    a->tile_rows_local = (uint32_t)r->rank < strips ? (strips - r->rank + r->world - 1) / r->world : 0;
    if (r->band_count >= 0) a->tile_rows_local = (uint32_t)r->band_first < strips ? std::min<uint32_t>((uint32_t)r->band_count, strips - (uint32_t)r->band_first) : 0;
    a->cta_fetch = (!r->tf_opaque && !r->opt.warp_fetch) ? 1 : 0;
    set_queue_geometry(a);
    a->group_order = nullptr;
    if (const int mode = a->n_queue ? r->opt.queue_order : 0) {
        if (build_group_order(r, a, mode)) return 1;
        a->group_order = r->d_group_order;
    }
    a->counters = reinterpret_cast<Counters *>(r->d_counters);
    a->reach = frame_reach(*a);
    // leaps over transparent bricks (zero-skip kernels, vr5_cast.cuh try_leap): the reference's fp32 position of a sample, o + d * k, + 1,
    // * dims / 2, lies within err = 3 * 2^-24 * (|origin| + 2) * dims / 2 voxels (+ half an ulp of the product) of the exact point of
    // the ray; a leap keeps 4 x that bound + 1/32 voxel (the rounding of its own brick-to-brick walk: < 0.01 voxel over 48 bricks) away
    // from the faces of its bricks. Views for which that margin would exceed 1/4 voxel (huge grids seen from far away) march sample by sample.
    a->brick_max = nullptr;
    if (a->zero_lim && r->d_brick_max && std::isfinite(a->reach)) {
        float ob = 0.0f;
        const float ov[3] = {a->ox, a->oy, a->oz}, rv[3] = {a->rx, a->ry, a->rz}, uv[3] = {a->ux, a->uy, a->uz};
        for (int i = 0; i < 3; i++)
            ob = fmaxf(ob, fabsf(ov[i]) + (a->perspective ? 0.0f : fabsf(rv[i]) * (float)(a->half_w + 1) + fabsf(uv[i]) * (float)(a->half_h + 1)));
        const float hmax = fmaxf(a->hx, fmaxf(a->hy, a->hz));
        const float margin = 4.0f * (3.0f * 5.9604645e-8f * (ob + 2.0f) * hmax + 5.9604645e-8f * 2.0f * hmax) + 1.0f / 32.0f;
        if (std::isfinite(ob) && margin <= 0.25f) {
            a->brick_max = r->d_brick_max; a->brick_lb = r->brick_lb; a->brick_nx = r->brick_nx; a->brick_ny = r->brick_ny; a->brick_nz = r->brick_nz;
            a->leap_margin = margin; a->leap_stats = r->opt.leap_stats ? reinterpret_cast<unsigned long long *>(r->d_leap_stats) : nullptr;
        }
    }
    // predicted leaps (vr5_cast.cuh, skip_leading_empty_space): finite, non-degenerate view, rays clipped to the cube [-1, 1]^3
    // the grid fills, step with a safe reciprocal; VR5_ESL_EXACT keeps the reference's divisions in every iteration
    int step_exp = 0;
    frexpf(p->ray_step, &step_exp);
    if (a->esl && !r->opt.esl_exact && std::isfinite(a->reach) && a->mbx == -1.0f && a->mby == -1.0f && a->mbz == -1.0f && step_exp > -60 && step_exp < 60 && bd >= 2 &&
        std::isfinite(a->bsx) && std::isfinite(a->bsy) && std::isfinite(a->bsz))
        a->esl = 2;
    return 0;
}

// The window path has no index clamp: its prefetch runs one step past the clip interval, so that step must stay inside the
// guard margin (kWindowGuard voxels) on every axis for every ray of the frame. d is affine in the pixel position, so the
// four image corners bound it. Anything non-finite or out of range renders through the clamped texture path instead.
static bool reach_ok(const KArgs &a, float lim);
static bool window_ok(const vr5_renderer *r, const KArgs &a) {
    if (!r->d_vol_window && !r->p2_wl) return false;
    return reach_ok(a, (float)((r->p2_wl ? kP2Origin : kWindowGuard) - 2) / (float)kPrefetch);
}

static bool reach_ok(const KArgs &a, float lim) { return a.reach <= lim; }  // false for +inf / NaN

template <int ID>
static cudaError_t launch_p2_id(vr5_renderer *r, const KArgs &a, bool shade) {
    return shade ? launch_one<ID, false, true, false>(r, a) : launch_one<ID, false, false, false>(r, a);
}

static cudaError_t launch_p2(vr5_renderer *r, const KArgs &a, bool shade, int mode) {
    switch (mode) {
        case 0: return launch_p2_id<kP2First + 0>(r, a, shade);
        case 1: return launch_p2_id<kP2First + 1>(r, a, shade);
        case 2: return launch_p2_id<kP2First + 2>(r, a, shade);
    }
    return cudaErrorInvalidValue;
}

static int launch_variant(vr5_renderer *r, KArgs &a, bool count);

// one cast-kernel launch on a lane: hand it the lane's queue counter and account for what the launch adds to it
static int launch(vr5_renderer *r, KArgs a, bool count, int lane) {
    vr5_renderer::Lane &ln = r->lanes[lane];
    a.queue = ln.d_queue;
    a.queue_base = ln.queue_base;
    r->launch_stream = vr5_lane_stream(r, lane);
    r->last_grid = 0;
    const int rc = launch_variant(r, a, count);
    if (r->last_grid) ln.queue_base += a.n_queue + r->last_grid * (uint32_t)(kCtaThreads / 32);  // last_grid == 0: nothing was launched
    return rc;
}

// ---- frame sequences: ONE persistent launch over all frames of a camera path ------------------------------------------------
// The tiles of every frame go through one queue (slot / n_queue = frame), so the grid idles behind its last long rays once per
// sequence instead of once per frame, and the host submits the whole path with one call (VolR's benchmark loop renders its
// frames one call at a time, VolR.cpp:98-113, :225-321). Everything but the camera must be the same for all frames.
int vr5_sequence_launch(vr5_renderer *r, const vr5_params *frames, int n, const vr5_frame_target *targets, uint32_t *timed_out_dev, int targets_remote) {
    if (!frames || !targets || n < 1) { vr5_fail(r, "vr5_render_sequence: bad arguments", cudaSuccess, __FILE__, __LINE__); return 1; }
    KArgs a;
    if (fill_args(r, frames, targets[0].out_dev, &a)) return 1;
    if ((uint64_t)a.n_queue * (uint64_t)n >= (1ull << 31)) { vr5_fail(r, "vr5_render_sequence: too many tiles for one queue (split the sequence)", cudaSuccess, __FILE__, __LINE__); return 1; }
    std::vector<FrameView> fv((size_t)n);
    bool shake = false;
    const bool bands = r->band_count >= 0;   // per-frame bands travel in the frame records
    uint32_t max_rows = a.tile_rows_local;
    for (int i = 0; i < n; i++) {
        const vr5_params &p = frames[i];
        uint32_t band0 = (uint32_t)a.rank, bandn = a.tile_rows_local;
        if (!targets[i].out_dev) { vr5_fail(r, "vr5_render_sequence: NULL frame buffer", cudaSuccess, __FILE__, __LINE__); return 1; }
        if (i > 0) {
            const vr5_params &q = frames[0];
            const bool same = p.view.dims.x == q.view.dims.x && p.view.dims.y == q.view.dims.y && !p.view.perspective == !q.view.perspective &&
                              !memcmp(&p.view.light_pos, &q.view.light_pos, sizeof(vr5_f3)) && p.ray_step == q.ray_step && p.ray_threshold == q.ray_threshold &&
                              !p.esl == !q.esl && p.esl_block_dims == q.esl_block_dims && !memcmp(&p.esl_block_size, &q.esl_block_size, sizeof(vr5_f3)) &&
                              p.light_kd == q.light_kd && !memcmp(&p.min_bound, &q.min_bound, sizeof(vr5_f3));
            if (!same) { vr5_fail(r, "vr5_render_sequence: frames of a sequence may differ in origin / direction / right_plane / up_plane only", cudaSuccess, __FILE__, __LINE__); return 1; }
            KArgs ai;   // the unclamped variants must hold for every frame: carry the largest reach
            if (fill_args(r, &p, targets[i].out_dev, &ai)) return 1;
            a.reach = ai.reach > a.reach || !(ai.reach == ai.reach) ? ai.reach : a.reach;
            if (!ai.brick_max) a.brick_max = nullptr;   // ... and so must the position bound of the leaps, with the widest margin any frame asks for
            else if (ai.leap_margin > a.leap_margin) a.leap_margin = ai.leap_margin;
            band0 = (uint32_t)ai.rank; bandn = ai.tile_rows_local;
            if (bandn > max_rows) max_rows = bandn;
        }
        const vr5_view &v = p.view;
        fv[i] = FrameView{v.direction.x, v.direction.y, v.direction.z, v.origin.x, v.origin.y, v.origin.z, v.right_plane.x, v.right_plane.y, v.right_plane.z,
                          v.up_plane.x, v.up_plane.y, v.up_plane.z, reinterpret_cast<uchar4 *>(targets[i].out_dev), targets[i].arrive, targets[i].consumed,
                          targets[i].consumed_target, band0, bandn, {0u, 0u, 0u}};
        shake |= targets[i].arrive || targets[i].consumed;
    }
    // per-sequence device block [views | done counters], allocated and freed in stream order: sequences may queue up
    cudaStream_t st = r->stream;
    // [views | per-frame completion counters | ring_sync: known-writable frame, poller lock]
    const size_t views_bytes = (size_t)n * sizeof(FrameView), done_bytes = shake ? (size_t)n * sizeof(uint32_t) : 0, sync_bytes = shake ? 2 * sizeof(uint32_t) : 0;
    uint8_t *blk = nullptr;
    VR5_CUDA(r, cudaMallocAsync(&blk, views_bytes + done_bytes + sync_bytes, st));
    VR5_CUDA(r, cudaMemcpyAsync(blk, fv.data(), views_bytes, cudaMemcpyHostToDevice, st));
    if (shake) {
        VR5_CUDA(r, cudaMemsetAsync(blk + views_bytes, 0, done_bytes + sync_bytes, st));
        VR5_CUDA(r, cudaMemsetAsync(blk + views_bytes + done_bytes, 0xff, sizeof(uint32_t), st));   // no frame known yet
    }
    if (bands) {   // the tallest band of the sequence sizes the per-frame share of the queue
        a.bands = 1;
        a.tile_rows_local = max_rows;
        set_queue_geometry(&a);
        if ((uint64_t)a.n_queue * (uint64_t)n >= (1ull << 31)) { cudaFreeAsync(blk, st); vr5_fail(r, "vr5_render_sequence: too many tiles for one queue (split the sequence)", cudaSuccess, __FILE__, __LINE__); return 1; }
    }
    if (shake && (r->opt.seq_debug & 8)) a.cta_fetch = 1;   // (measurement knob: per-CTA queue reads whatever the transfer function)
    a.frames = reinterpret_cast<const FrameView *>(blk);
    a.n_frames = (uint32_t)n;
    a.frame_done = shake ? reinterpret_cast<uint32_t *>(blk + views_bytes) : nullptr;
    a.ring_sync = shake ? reinterpret_cast<uint32_t *>(blk + views_bytes + done_bytes) : nullptr;
    a.timed_out = timed_out_dev;
    a.ring_remote = targets_remote ? 1 : 0;
    a.queue = r->d_seq_queue;
    a.queue_base = r->seq_queue_base;
    r->launch_stream = st;
    r->last_grid = 0;
    const int rc = launch_variant(r, a, false);
    if (r->last_grid) r->seq_queue_base += a.n_queue * (uint32_t)n + r->last_grid * (uint32_t)(kCtaThreads / 32);
    VR5_CUDA(r, cudaFreeAsync(blk, st));
    return rc;
}

// ---- vr5_render_sequence without any allocation (no counterpart in the reference) ----
// vr5_render_sequence keeps the per-frame records its kernel reads (views, buffers) in a block it allocates and frees in stream
// order - an allocation per call, which callers under CUDA graph capture, in tight serving loops or in harnesses that forbid
// cudaMalloc cannot make. This variant keeps the records in a file-scope device array of VR5_NOALLOC_MAX_FRAMES entries (24 KiB;
// every GPU that loads the library holds it until exit). ONE array per device: every call overwrites the records of the call
// before (nothing in it has to start at zero: the records a launch reads are rewritten in full), so two calls that overlap on
// the GPU - different handles or streams - race; calls on one handle's stream are ordered and safe.
static __device__ __align__(256) FrameView g_noalloc_frames[VR5_NOALLOC_MAX_FRAMES];

// In host code `g_noalloc_frames` is a host-side stand-in, not a device address: the kernel gets the address that
// cudaGetSymbolAddress returns, looked up once per device with that device current.
static FrameView *noalloc_frames_ptr(vr5_renderer *r) {
    static FrameView *ptrs[64] = {nullptr};
    if (r->device < 0 || r->device >= 64) { vr5_fail(r, "vr5_render_sequence_noalloc: device ordinal beyond 63", cudaSuccess, __FILE__, __LINE__); return nullptr; }
    if (!ptrs[r->device]) {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_frames);   // the caller has made r->device current
        if (e != cudaSuccess || !p) { vr5_fail(r, "cudaGetSymbolAddress(g_noalloc_frames)", e, __FILE__, __LINE__); cudaGetLastError(); return nullptr; }
        ptrs[r->device] = static_cast<FrameView *>(p);
    }
    return ptrs[r->device];
}

extern "C" int vr5_render_sequence_noalloc(vr5_renderer *r, const vr5_params *frames, int n_frames, void *const *outs_dev) {
    if (!r) return 1;
    if (!frames || !outs_dev || n_frames < 1) { vr5_fail(r, "vr5_render_sequence_noalloc: bad arguments", cudaSuccess, __FILE__, __LINE__); return 1; }
    if (n_frames > VR5_NOALLOC_MAX_FRAMES) {
        vr5_fail(r, "vr5_render_sequence_noalloc: more than VR5_NOALLOC_MAX_FRAMES (256) frames - split the sequence or use vr5_render_sequence", cudaSuccess, __FILE__, __LINE__);
        return 1;
    }
    if (r->opt.queue_order != 0) {   // the opt-in tile orders build a device table on first use
        vr5_fail(r, "vr5_render_sequence_noalloc: VR5_QUEUE_ORDER must be rows (the other orders allocate a table)", cudaSuccess, __FILE__, __LINE__);
        return 1;
    }
    VR5_CUDA(r, cudaSetDevice(r->device));
    FrameView *dev_records = noalloc_frames_ptr(r);
    if (!dev_records) return 1;
    KArgs a;
    if (fill_args(r, frames, outs_dev[0], &a)) return 1;
    static thread_local FrameView fv[VR5_NOALLOC_MAX_FRAMES];   // host staging without the heap
    const bool bands = r->band_count >= 0;
    uint32_t max_rows = a.tile_rows_local;
    for (int i = 0; i < n_frames; i++) {
        const vr5_params &p = frames[i];
        uint32_t band0 = (uint32_t)a.rank, bandn = a.tile_rows_local;
        if (!outs_dev[i]) { vr5_fail(r, "vr5_render_sequence_noalloc: NULL frame buffer", cudaSuccess, __FILE__, __LINE__); return 1; }
        if (i > 0) {
            const vr5_params &q = frames[0];
            const bool same = p.view.dims.x == q.view.dims.x && p.view.dims.y == q.view.dims.y && !p.view.perspective == !q.view.perspective &&
                              !memcmp(&p.view.light_pos, &q.view.light_pos, sizeof(vr5_f3)) && p.ray_step == q.ray_step && p.ray_threshold == q.ray_threshold &&
                              !p.esl == !q.esl && p.esl_block_dims == q.esl_block_dims && !memcmp(&p.esl_block_size, &q.esl_block_size, sizeof(vr5_f3)) &&
                              p.light_kd == q.light_kd && !memcmp(&p.min_bound, &q.min_bound, sizeof(vr5_f3));
            if (!same) { vr5_fail(r, "vr5_render_sequence_noalloc: frames of a sequence may differ in origin / direction / right_plane / up_plane only", cudaSuccess, __FILE__, __LINE__); return 1; }
            KArgs ai;
            if (fill_args(r, &p, outs_dev[i], &ai)) return 1;
            a.reach = ai.reach > a.reach || !(ai.reach == ai.reach) ? ai.reach : a.reach;
            if (!ai.brick_max) a.brick_max = nullptr;   // ... and so must the position bound of the leaps, with the widest margin any frame asks for
            else if (ai.leap_margin > a.leap_margin) a.leap_margin = ai.leap_margin;
            band0 = (uint32_t)ai.rank; bandn = ai.tile_rows_local;
            if (bandn > max_rows) max_rows = bandn;
        }
        const vr5_view &v = p.view;
        fv[i] = FrameView{v.direction.x, v.direction.y, v.direction.z, v.origin.x, v.origin.y, v.origin.z, v.right_plane.x, v.right_plane.y, v.right_plane.z,
                          v.up_plane.x, v.up_plane.y, v.up_plane.z, reinterpret_cast<uchar4 *>(outs_dev[i]), nullptr, nullptr, 0u, band0, bandn, {0u, 0u, 0u}};
    }
    if (bands) { a.bands = 1; a.tile_rows_local = max_rows; set_queue_geometry(&a); }
    if ((uint64_t)a.n_queue * (uint64_t)n_frames >= (1ull << 31)) { vr5_fail(r, "vr5_render_sequence_noalloc: too many tiles for one queue (split the sequence)", cudaSuccess, __FILE__, __LINE__); return 1; }
    cudaStream_t st = r->stream;
    VR5_CUDA(r, cudaMemcpyAsync(dev_records, fv, (size_t)n_frames * sizeof(FrameView), cudaMemcpyHostToDevice, st));   // pageable source: staged before the call returns
    a.frames = dev_records;
    a.n_frames = (uint32_t)n_frames;
    a.frame_done = nullptr; a.timed_out = nullptr; a.ring_remote = 0;
    a.queue = r->d_seq_queue;
    a.queue_base = r->seq_queue_base;
    r->launch_stream = st;
    r->last_grid = 0;
    const int rc = launch_variant(r, a, false);
    if (r->last_grid) r->seq_queue_base += a.n_queue * (uint32_t)n_frames + r->last_grid * (uint32_t)(kCtaThreads / 32);
    return rc;
}

extern "C" int vr5_render_sequence(vr5_renderer *r, const vr5_params *frames, int n_frames, void *const *outs_dev) {
    if (!r) return 1;
    if (!frames || !outs_dev || n_frames < 1) { vr5_fail(r, "vr5_render_sequence: bad arguments", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    std::vector<vr5_frame_target> t((size_t)n_frames);
    for (int i = 0; i < n_frames; i++) t[(size_t)i] = vr5_frame_target{outs_dev[i], nullptr, nullptr, 0u};
    return vr5_sequence_launch(r, frames, n_frames, t.data(), nullptr, 0);
}

static int launch_variant(vr5_renderer *r, KArgs &a, bool count) {
    // shading can only trigger when light_kd > 0.01 (folded into shade_alpha) and some TF entry has alpha > 0.05
    const bool shade = r->tf_can_shade && a.shade_alpha < std::numeric_limits<float>::infinity();
    // the padded layouts may be indexed without the upper clamp when rays are clipped to the cube they cover
    // (a frame whose reach is not finite - degenerate or non-finite view, frame_reach - keeps every clamp)
    const bool cube = a.mbx == -1.0f && a.mby == -1.0f && a.mbz == -1.0f && std::isfinite(a.reach);
    cudaError_t e;
    if (is_trilinear(r->sampler)) {  // GPURenderer4 semantics in software fp32 (vr5_tri.cuh); _FMA: contracted arithmetic
        TriArgs t;
        t.cells = reinterpret_cast<const uint2 *>(r->d_cells);
        t.cx = r->nx + 1 + 2 * kTriGuard; t.cy = r->ny + 1 + 2 * kTriGuard;
        t.hfx = a.hx; t.hfy = a.hy; t.hfz = a.hz;
        t.jx1 = r->nx + 2 * kTriGuard; t.jy1 = r->ny + 2 * kTriGuard; t.jz1 = r->nz + 2 * kTriGuard;
        // unclamped 32-bit addressing: record j = b - (bits(magic) - 1 - G) per axis, byte address = base + 8 * rec
        const uint64_t base = (uint64_t)(uintptr_t)r->d_cells;
        const bool one_window = r->cells_bytes < (1ull << 32) && (base >> 32) == ((base + r->cells_bytes - 1) >> 32);
        const uint32_t kb = kTriMagicBits - 1u - (uint32_t)kTriGuard;
        t.hi = (uint32_t)(base >> 32);
        t.s1 = 8u * t.cx; t.s2 = 8u * t.cx * t.cy;                       // mod 2^32 like every term of the address sum
        t.c0 = (uint32_t)base - kb * (t.s2 + t.s1 + 8u);
        // one march step may leave the cube by step * |d| * n/2 cells per axis (+1 for the cell's far corner, +1 rounding)
        const bool fast = one_window && cube && !count && !r->opt.tri_clamped && reach_ok(a, (float)(kTriGuard - 2));
        // variants: instrumented (clamped, shading code present), and {clamped, unclamped} x {shading code, none}
        const bool fused = r->sampler == VR5_SAMPLER_TRILINEAR_FMA;
        int v = (count ? 0 : 1 + (fast ? 2 : 0) + (shade ? 1 : 0)) + (fused ? 5 : 0);
        int depth = 1;
        if (v == 8) {  // contracted, unclamped, no shading code: register budget x requests in flight (profiles/r01_sampler_choice.md)
            const int ctas = r->opt.trif_ctas;
            depth = r->opt.trif_depth;  // default 6 CTAs per SM x 2 requests in flight measured best (profiles/r01_sampler_choice.md)
            if (depth < 1 || depth > 4) depth = 1;
            // DEPTH requests run DEPTH steps past the clip interval: all of them must stay inside the guard cells
            if (depth > 1 && !reach_ok(a, (float)(kTriGuard - 2) / (float)depth)) depth = 1;
            static const int table[3][4] = {{8, 12, 12, 12}, {10, 13, 13, 13}, {11, 14, 15, 16}};  // rows: 6 / 5 / 4 CTAs per SM, columns: depth 1..4
            v = table[ctas == 5 ? 1 : ctas == 4 ? 2 : 0][depth - 1];
        }
        void (*const kernels[17])(const KArgs, const TriArgs) = {
            tri_kernel<true, false, true, false>, tri_kernel<false, false, false, false>, tri_kernel<false, false, true, false>,
            tri_kernel<false, true, false, false>, tri_kernel<false, true, true, false>,
            tri_kernel<true, false, true, true>, tri_kernel<false, false, false, true>, tri_kernel<false, false, true, true>,
            tri_kernel<false, true, false, true>, tri_kernel<false, true, true, true>,
            tri_kernel<false, true, false, true, 5>, tri_kernel<false, true, false, true, 4>,
            tri_kernel<false, true, false, true, 6, 2>, tri_kernel<false, true, false, true, 5, 2>, tri_kernel<false, true, false, true, 4, 2>,
            tri_kernel<false, true, false, true, 4, 3>, tri_kernel<false, true, false, true, 4, 4>};
        // zero-skip builds of the variants that render frames (not the instrumented ones, not the register-budget experiments): chosen when
        // the transfer function starts with all-zero entries (fill_args: zero_lim = 2 x the prefix, under the same conditions as for cast_kernel)
        void (*const kernels_zs[17])(const KArgs, const TriArgs) = {
            nullptr, tri_kernel<false, false, false, false, 6, 1, true>, tri_kernel<false, false, true, false, 0, 1, true>,
            tri_kernel<false, true, false, false, 6, 1, true>, tri_kernel<false, true, true, false, 0, 1, true>,
            nullptr, tri_kernel<false, false, false, true, 6, 1, true>, tri_kernel<false, false, true, true, 0, 1, true>,
            tri_kernel<false, true, false, true, 6, 1, true>, tri_kernel<false, true, true, true, 0, 1, true>,
            nullptr, nullptr, tri_kernel<false, true, false, true, 6, 2, true>, nullptr, nullptr, nullptr, nullptr};
        const bool zs = a.zero_lim >= 2u && kernels_zs[v] != nullptr;
        t.zero_um = zs ? kTriMagicBits + a.zero_lim / 2u - 1u : 0u;
        void (*const kernel)(const KArgs, const TriArgs) = zs ? kernels_zs[v] : kernels[v];
        int &per_sm = r->ctas_per_sm[480 + v + (zs ? 17 : 0)];  // 480..513 (cast_kernel variants use 0..215)
        if (per_sm == 0) {
            e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kCtaThreads, 0);
            if (e != cudaSuccess || per_sm < 1) per_sm = 1;
        }
        uint32_t grid = (uint32_t)(r->num_sms * per_sm);
        const uint32_t needed = (a.n_queue * a.n_frames + (kCtaThreads / 32) - 1) / (kCtaThreads / 32);
        if (grid > needed) grid = needed ? needed : 1;
        kernel<<<grid, kCtaThreads, 0, r->launch_stream>>>(a, t);
        r->last_zskip = zs ? 1 : 0;
        r->tri_fast = fast ? 1 : 0;
        r->last_grid = grid;
        r->launches++;
        r->last_variant = r->sampler + (fast ? 1 : 0);
        e = cudaGetLastError();
        if (e != cudaSuccess) { vr5_fail(r, "tri_kernel launch", e, __FILE__, __LINE__); return 1; }
        return 0;
    }
    // power-of-two specialisation of the window policies (the instrumented variant counts through the generic kernels:
    // work counters do not depend on the memory path)
    if (r->sampler >= VR5_SAMPLER_HYBRID_WINDOW && r->p2_wl && (r->p2_wl == 1 || r->sampler == VR5_SAMPLER_HYBRID_WINDOW) && !count && cube && window_ok(r, a)) {
        a.whi = r->p2_hi; a.wmx = r->p2_mx; a.wmy = r->p2_my; a.wmz = r->p2_mz; a.tex_p2 = r->tex_p2; a.wsx = r->p2_sx; a.wsxy = r->p2_sxy;
        int mode = r->sampler - VR5_SAMPLER_HYBRID_WINDOW;
        // a transfer function whose first quarter (or more) is all-zero makes most samples transparent ones (zero-skip kernels):
        // their march is little more than the voxel request, and window requests (LSU) are cheaper than texture requests (a
        // warp-wide TLD is 8 TEX wavefronts). C4: 1.130 ms per frame as texture : window 1 : 1, 1.080 window only (round 2).
        if (mode == 0 && a.zero_lim >= 64u) mode = 1;
        e = r->p2_wl == 2 ? launch_p2_id<kP2Wide>(r, a, shade) : launch_p2(r, a, shade, mode);
        if (e != cudaSuccess) { vr5_fail(r, "cast_kernel launch", e, __FILE__, __LINE__); return 1; }
        return 0;
    }
    switch (r->sampler) {
        case VR5_SAMPLER_LINEAR: e = (cube && r->d_vol_padded && reach_ok(a, (float)(kWindowGuard - 2) / (float)kPrefetch)) ? launch_sc<0, false>(r, a, count, shade) : launch_sc<0, true>(r, a, count, shade); break;
        case VR5_SAMPLER_BRICK: e = launch_sc<1, true>(r, a, count, shade); break;
        case VR5_SAMPLER_HYBRID: e = (cube && r->d_vol_padded && reach_ok(a, (float)(kWindowGuard - 2) / (float)kPrefetch)) ? launch_sc<3, false>(r, a, count, shade) : launch_sc<2, true>(r, a, count, shade); break;
        case VR5_SAMPLER_HYBRID_WINDOW: e = (cube && r->d_vol_window && window_ok(r, a)) ? launch_sc<4, false>(r, a, count, shade) : cube ? launch_sc<2, false>(r, a, count, shade) : launch_sc<2, true>(r, a, count, shade); break;
        case VR5_SAMPLER_WINDOW: e = (cube && r->d_vol_window && window_ok(r, a)) ? launch_sc<5, false>(r, a, count, shade) : cube ? launch_sc<2, false>(r, a, count, shade) : launch_sc<2, true>(r, a, count, shade); break;
        case VR5_SAMPLER_HYBRID_WINDOW2: e = (cube && r->d_vol_window && window_ok(r, a)) ? launch_sc<6, false>(r, a, count, shade) : cube ? launch_sc<2, false>(r, a, count, shade) : launch_sc<2, true>(r, a, count, shade); break;
        default: e = cube ? launch_sc<2, false>(r, a, count, shade) : launch_sc<2, true>(r, a, count, shade); break;
    }
    if (e != cudaSuccess) { vr5_fail(r, "cast_kernel launch", e, __FILE__, __LINE__); return 1; }
    return 0;
}

static int resolve_is_device(vr5_renderer *r, const void *ptr, int hint) {
    if (hint == 0 || hint == 1) return hint;
    cudaPointerAttributes attr;
    cudaError_t e = cudaPointerGetAttributes(&attr, ptr);
    if (e != cudaSuccess) { cudaGetLastError(); return 0; }
    (void)r;
    return (attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged) ? 1 : 0;
}

// The plugin call as VolR makes it (VolR.cpp:110): blocking, whole frame, HOST buffer, one lane. Done as one launch + one copy
// the frame's 8 MB cross the bus after the kernel has finished (pageable memory: ~0.5 ms behind a 0.6 ms kernel at 1080p). Here the
// frame is cut into kHostChunks cost-balanced bands (vr5_balance_bands) rendered by one launch each on the lane streams; the
// copy of band i runs while bands i + 1 .. are being rendered. Same pixels (each band is the partition path of the same kernel).
constexpr int kHostChunks = 4;
static int render_host_chunked(vr5_renderer *r, const vr5_params *p, void *out) {
    vr5_renderer::Lane &l0 = r->lanes[0];
    if (!l0.d_frame) { vr5_fail(r, "vr5_render: vr5_set_window not called", cudaSuccess, __FILE__, __LINE__); return 1; }
    uint32_t bounds[kHostChunks + 1];
    if (vr5_balance_bands(p, kHostChunks, bounds)) { vr5_fail(r, "vr5_balance_bands failed", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaEventRecord(r->ev_fork, r->stream));   // every chunk is ordered behind what the caller enqueued before
    const int saved_first = r->band_first, saved_count = r->band_count;
    const bool saved_auto = r->band_auto;
    r->band_auto = false;
    int rc = 0;
    cudaStream_t st[kHostChunks];
    for (int c = 0; c < kHostChunks && !rc; c++) {
        st[c] = c == 0 ? r->stream : r->lanes[c].stream;
        if (c > 0 && cudaStreamWaitEvent(st[c], r->ev_fork, 0) != cudaSuccess) { rc = 1; break; }
        r->band_first = (int)bounds[c]; r->band_count = (int)(bounds[c + 1] - bounds[c]);
        KArgs a;
        if (fill_args(r, p, l0.d_frame, &a)) { rc = 1; break; }
        vr5_renderer::Lane &ln = r->lanes[c];
        a.queue = ln.d_queue;
        a.queue_base = ln.queue_base;
        r->launch_stream = st[c];
        r->last_grid = 0;
        rc = launch_variant(r, a, false);
        if (r->last_grid) ln.queue_base += a.n_queue + r->last_grid * (uint32_t)(kCtaThreads / 32);
    }
    r->band_first = saved_first; r->band_count = saved_count; r->band_auto = saved_auto;
    const size_t row_bytes = (size_t)r->width * 4u;
    for (int c = 0; c < kHostChunks && !rc; c++) {   // all kernels are enqueued: a copy into pageable memory may block the host from here on
        const size_t row0 = (size_t)bounds[c] * kTileH, row1 = std::min<size_t>(r->height, (size_t)bounds[c + 1] * kTileH);
        if (row1 > row0 && cudaMemcpyAsync(static_cast<uint8_t *>(out) + row0 * row_bytes, l0.d_frame + row0 * row_bytes, (row1 - row0) * row_bytes,
                                           cudaMemcpyDeviceToHost, st[c]) != cudaSuccess) rc = 1;
    }
    cudaError_t e = cudaSuccess;
    for (int c = kHostChunks - 1; c >= 0; c--) { const cudaError_t ec = cudaStreamSynchronize(c == 0 ? r->stream : r->lanes[c].stream); if (e == cudaSuccess) e = ec; }
    if (rc || e != cudaSuccess) { if (e != cudaSuccess || r->err.empty()) vr5_fail(r, "vr5_render (host buffer, chunked)", e, __FILE__, __LINE__); return 1; }
    return 0;
}

static int render_impl(vr5_renderer *r, const vr5_params *p, void *out, int out_is_device, bool blocking) {
    if (!r) return 1;
    if (!out) { vr5_fail(r, "vr5_render: NULL buffer", cudaSuccess, __FILE__, __LINE__); return 1; }  // GPURenderer1.cu:101
    VR5_CUDA(r, cudaSetDevice(r->device));
    const int is_dev = resolve_is_device(r, out, out_is_device);
    if (!is_dev && blocking && r->n_lanes == 1 && r->world == 1 && !r->packed && r->band_count < 0 && !r->opt.host_single && p &&
        p->view.dims.y >= 8 * kHostChunks * kTileH && p->view.dims.x == r->width && p->view.dims.y == r->height) {
        r->frame_no++;
        return render_host_chunked(r, p, out);
    }
    const int lane = (int)(r->frame_no++ % (uint64_t)r->n_lanes);
    vr5_renderer::Lane &ln = r->lanes[lane];
    if (!is_dev && !ln.d_frame) { vr5_fail(r, "vr5_render: vr5_set_window not called", cudaSuccess, __FILE__, __LINE__); return 1; }
    KArgs a;
    if (fill_args(r, p, is_dev ? out : (void *)ln.d_frame, &a)) return 1;
    if (vr5_lane_begin(r, lane)) return 1;
    if (launch(r, a, false, lane)) return 1;
    cudaStream_t st = vr5_lane_stream(r, lane);
    if (!is_dev) {  // the copy rides on the frame's lane: with two lanes it overlaps the next frame's kernel
        const size_t rows = r->packed ? vr5_strip_rows(r->height, r->rank, r->world) : r->height;
        VR5_CUDA(r, cudaMemcpyAsync(out, ln.d_frame, rows * r->width * 4, cudaMemcpyDeviceToHost, st));
    }
    if (vr5_lane_end(r, lane)) return 1;
    if (blocking) { VR5_CUDA(r, cudaStreamSynchronize(st)); ln.pending = false; }
    return 0;
}

int vr5_render_into(vr5_renderer *r, const vr5_params *p, void *out_dev, int lane) {
    KArgs a;
    if (fill_args(r, p, out_dev, &a)) return 1;
    return launch(r, a, false, lane);
}

extern "C" int vr5_render(vr5_renderer *r, const vr5_params *p, void *out, int out_is_device) { return render_impl(r, p, out, out_is_device, true); }
extern "C" int vr5_render_async(vr5_renderer *r, const vr5_params *p, void *out, int out_is_device) { return render_impl(r, p, out, out_is_device, false); }

extern "C" int vr5_render_raycaster(vr5_renderer *r, const vr5_raycaster *rc, void *out, int out_is_device) {
    if (!r) return 1;
    if (!rc) { vr5_fail(r, "vr5_render_raycaster: NULL raycaster", cudaSuccess, __FILE__, __LINE__); return 1; }
    vr5_params p;
    memset(&p, 0, sizeof(p));
    p.view = rc->view;
    p.ray_step = rc->ray_step; p.ray_threshold = rc->ray_threshold;
    p.esl = rc->esl ? 1 : 0; p.esl_block_dims = rc->esl_block_dims; p.esl_block_size = rc->esl_block_size;
    p.light_kd = rc->light_kd; p.min_bound = rc->volume.min_bound;
    return render_impl(r, &p, out, out_is_device, true);
}

extern "C" int vr5_sync(vr5_renderer *r) {
    if (!r) return 1;
    VR5_CUDA(r, cudaSetDevice(r->device));
    VR5_CUDA(r, sync_all(r));
    return 0;
}

extern "C" int vr5_set_stream(vr5_renderer *r, void *cuda_stream) {
    if (!r) return 1;
    VR5_CUDA(r, cudaSetDevice(r->device));
    VR5_CUDA(r, sync_all(r));
    if (cuda_stream) { r->stream = (cudaStream_t)cuda_stream; r->external_stream = true; }
    else { r->stream = r->own_stream; r->external_stream = false; }
    return 0;
}

// ---- measurement ---------------------------------------------------------------------------------------------------
extern "C" int vr5_render_timed(vr5_renderer *r, const vr5_params *frames, int n_frames, void *out_dev, int l2_flush, float *ms_out) {
    if (!r) return 1;
    if (!frames || n_frames <= 0 || !out_dev || !ms_out) { vr5_fail(r, "vr5_render_timed: bad arguments", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    VR5_CUDA(r, sync_all(r));
    cudaStream_t st = vr5_lane_stream(r, 0);  // serial on lane 0: per-launch times must not overlap
    if (l2_flush && !r->d_flush) {
        r->flush_bytes = r->l2_bytes * 2 > (256u << 20) ? r->l2_bytes * 2 : (256u << 20);
        VR5_CUDA(r, cudaMalloc(&r->d_flush, r->flush_bytes));
    }
    std::vector<cudaEvent_t> ev(2 * (size_t)n_frames);
    for (auto &e : ev) VR5_CUDA(r, cudaEventCreate(&e));
    int rc = 0;
    for (int f = 0; f < n_frames && !rc; f++) {
        if (l2_flush) {
            l2_flush_kernel<<<r->num_sms * 4, 512, 0, st>>>(reinterpret_cast<uint32_t *>(r->d_flush), r->flush_bytes / 4, (uint32_t)f);
            r->launches++;
        }
        KArgs a;
        if (fill_args(r, frames + f, out_dev, &a)) { rc = 1; break; }
        cudaEventRecord(ev[2 * f], st);
        rc = launch(r, a, false, 0);
        cudaEventRecord(ev[2 * f + 1], st);
    }
    cudaError_t e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) { vr5_fail(r, "vr5_render_timed sync", e, __FILE__, __LINE__); rc = 1; }
    if (!rc)
        for (int f = 0; f < n_frames; f++) cudaEventElapsedTime(ms_out + f, ev[2 * f], ev[2 * f + 1]);
    for (auto &x : ev) cudaEventDestroy(x);
    return rc;
}

extern "C" int vr5_count(vr5_renderer *r, const vr5_params *p, vr5_counters *out) {
    if (!r) return 1;
    if (!out) { vr5_fail(r, "vr5_count: NULL out", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    KArgs a;
    if (fill_args(r, p, nullptr, &a)) return 1;
    VR5_CUDA(r, sync_all(r));
    cudaStream_t st = vr5_lane_stream(r, 0);
    VR5_CUDA(r, cudaMemsetAsync(r->d_counters, 0, sizeof(Counters), st));
    if (launch(r, a, true, 0)) return 1;
    static_assert(sizeof(Counters) == sizeof(vr5_counters), "counter layout");
    VR5_CUDA(r, cudaMemcpyAsync(out, r->d_counters, sizeof(Counters), cudaMemcpyDeviceToHost, st));
    VR5_CUDA(r, cudaStreamSynchronize(st));
    return 0;
}

extern "C" int vr5_leap_stats(vr5_renderer *r, uint64_t out[2]) {
    if (!r) return 1;
    if (!out) { vr5_fail(r, "vr5_leap_stats: NULL out", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    VR5_CUDA(r, sync_all(r));
    VR5_CUDA(r, cudaMemcpy(out, r->d_leap_stats, 16, cudaMemcpyDeviceToHost));
    VR5_CUDA(r, cudaMemset(r->d_leap_stats, 0, 16));
    return 0;
}

extern "C" uint64_t vr5_launch_count(const vr5_renderer *r) { return r ? r->launches : 0; }

extern "C" void *vr5_host_alloc(uint64_t bytes) {
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}
extern "C" void vr5_host_free(void *p) { if (p) cudaFreeHost(p); }

extern "C" void *vr5_device_alloc(vr5_renderer *r, uint64_t bytes) {
    if (!r) return nullptr;
    void *p = nullptr;
    if (cudaSetDevice(r->device) != cudaSuccess) return nullptr;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) { vr5_fail(r, "vr5_device_alloc", e, __FILE__, __LINE__); cudaGetLastError(); return nullptr; }
    return p;
}
extern "C" void vr5_device_free(vr5_renderer *r, void *p) { if (r && p) { cudaSetDevice(r->device); cudaFree(p); } }

extern "C" int vr5_memcpy_d2h(vr5_renderer *r, void *host_dst, const void *dev_src, uint64_t bytes) {
    if (!r) return 1;
    VR5_CUDA(r, cudaSetDevice(r->device));
    VR5_CUDA(r, cudaMemcpyAsync(host_dst, dev_src, bytes, cudaMemcpyDeviceToHost, r->stream));
    VR5_CUDA(r, cudaStreamSynchronize(r->stream));
    return 0;
}

extern "C" int vr5_unpack_strips(vr5_renderer *r, const void *packed_all, uint32_t rows_padded, int world, void *frame) {
    if (!r) return 1;
    if (!packed_all || !frame || world < 1 || !r->width) { vr5_fail(r, "vr5_unpack_strips: bad arguments", cudaSuccess, __FILE__, __LINE__); return 1; }
    VR5_CUDA(r, cudaSetDevice(r->device));
    unpack_strips_kernel<<<r->num_sms * 4, 256, 0, r->stream>>>(reinterpret_cast<const uint32_t *>(packed_all), rows_padded, world, r->width, r->height,
                                                               reinterpret_cast<uint32_t *>(frame));
    r->launches++;
    VR5_CUDA(r, cudaGetLastError());
    return 0;
}
