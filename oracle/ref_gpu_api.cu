/*
 * TEST / MEASUREMENT INFRASTRUCTURE ONLY (oracle build). Not part of the product path.
 *
 * Thin extern "C" driver over the reference's OWN CUDA renderers 1-4 (GPURenderer1.cu, GPURenderer23.cu, GPURenderer4.cu),
 * compiled UNMODIFIED from where they lie under /root/reference/VolumeRendering by oracle/Makefile into
 * oracle/_ref/libvolr_ref_gpu.so for sm_100a with nvcc's defaults (-fmad=true, as the reference's project builds them).
 * Renderer 4 uses the legacy texture-reference API that CUDA 12 removed: ref_gpu4_texobj.cu supplies it over texture
 * objects and includes the reference's file as it is (hardware-filtered tex3D / tex1D, the trilinear policies' anchor).
 * Nothing of the reference is copied into this repository: this TU only includes its headers and calls its classes.
 *
 * Purpose ("optional GPU comparison row"): the recompiled reference kernels are the baseline a
 * B200-native renderer has to beat; tools/ref_gpu_bench.py times them headless on the bench workloads, with a plain
 * cudaMalloc buffer in place of the GL pixel buffer (CUDA compute only, no graphics).
 *
 * Call sequence = the application's: renderers constructed in order 1, 2, 3 (VolR.cpp:425-429: renderer 1 owns the
 * shared device buffers, renderer 2 uploads the __constant__ tables renderer 3 also reads), then per frame
 * render_volume(buffer, raycaster) (VolR.cpp:98-113) - host buffer for ids 1-2, device buffer for id 3 (VolR.cpp:76-87).
 */
#include <chrono>
#include <stdio.h>
#include <string.h>

#include "cuda_utils.h"
#include "Renderer.h"

bool NO_SAFE = true; /* declared cuda_utils.h:21, defined VolR.cpp:30; true = log CUDA errors instead of exit() */

struct RefGpu {
    GPURenderer1 *r1;
    GPURenderer2 *r2;
    GPURenderer3 *r3;
    GPURenderer4 *r4;
    uchar4 *dev_frame;
    size_t frame_bytes;
};

extern "C" {

int refgpu_sizeof_raycaster(void) { return (int)sizeof(Raycaster); }

/* rc holds HOST pointers (volume data, transfer function, ESL bitfield), like RaycasterBase::raycaster */
void *refgpu_create(const Raycaster *rc) {
    RefGpu *h = new RefGpu();
    h->r1 = new GPURenderer1(*rc);
    h->r2 = new GPURenderer2(*rc);
    h->r3 = new GPURenderer3(*rc);
    h->r4 = new GPURenderer4(*rc); /* VolR.cpp:429: after renderer 1, whose window set-up it shares (Renderer.h:37-44) */
    h->frame_bytes = (size_t)rc->view.dims.x * rc->view.dims.y * 4;
    h->dev_frame = NULL;
    if (cudaMalloc((void **)&h->dev_frame, h->frame_bytes) != cudaSuccess || cudaGetLastError() != cudaSuccess) {
        delete h->r1; delete h;
        return NULL;
    }
    return h;
}

void refgpu_destroy(void *hv) {
    RefGpu *h = (RefGpu *)hv;
    if (!h) return;
    cudaFree(h->dev_frame);
    delete h->r4; delete h->r3; delete h->r2; delete h->r1; /* reverse order, VolR.cpp:328-329 */
    delete h;
}

/* Renders frames[0..n) with renderer `which` (1, 2: into host_out, D2H inside every call as shipped; 3, 4: into a device
 * buffer). Returns 0 and the wall-clock milliseconds of the whole loop (device idle before and after); for which >= 3
 * the last frame is copied to host_out after the clock stops. */
int refgpu_run(void *hv, int which, const Raycaster *frames, int n, unsigned char *host_out, double *ms) {
    RefGpu *h = (RefGpu *)hv;
    if (!h || !frames || n < 1 || !host_out) return 1;
    if (which < 1 || which > 4) return 1;
    Renderer *r = which == 1 ? (Renderer *)h->r1 : which == 2 ? (Renderer *)h->r2 : which == 3 ? (Renderer *)h->r3 : (Renderer *)h->r4;
    int rc = 0;
    cudaDeviceSynchronize();
    const auto t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < n; i++) rc |= r->render_volume(which >= 3 ? h->dev_frame : (uchar4 *)host_out, frames[i]);
    cudaDeviceSynchronize();
    const auto t1 = std::chrono::steady_clock::now();
    if (ms) *ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    if (which >= 3) cudaMemcpy(host_out, h->dev_frame, h->frame_bytes, cudaMemcpyDeviceToHost);
    if (cudaGetLastError() != cudaSuccess) rc |= 2;
    return rc;
}

/* A new volume / transfer function + ESL bitfield for every renderer (Renderer::set_volume, set_transfer_fn as the
 * application calls them for all renderers on a change, VolR.cpp:261-265, UI.cpp:52-58). rc holds HOST pointers. */
int refgpu_set_scene(void *hv, const Raycaster *rc, int volume_changed) {
    RefGpu *h = (RefGpu *)hv;
    if (!h || !rc) return 1;
    int bad = 0;
    Renderer *all[4] = {h->r1, h->r2, h->r3, h->r4};
    for (int i = 0; i < 4; i++) {
        if (volume_changed) bad |= all[i]->set_volume(rc->volume);
        all[i]->set_transfer_fn(*rc);
    }
    cudaDeviceSynchronize();
    return bad;
}

/* Renderer::set_window_buffer as the application calls it on a viewport change (VolR.cpp:99-105; only renderer 1
 * overrides it, the others share its statics, Renderer.h:37-44) + this driver's own device frame */
int refgpu_set_window(void *hv, const View *view) {
    RefGpu *h = (RefGpu *)hv;
    if (!h || !view) return 1;
    cudaDeviceSynchronize();
    h->r1->set_window_buffer(*view);
    cudaFree(h->dev_frame);
    h->dev_frame = NULL;
    h->frame_bytes = (size_t)view->dims.x * view->dims.y * 4;
    if (cudaMalloc((void **)&h->dev_frame, h->frame_bytes) != cudaSuccess) { cudaGetLastError(); return 1; }
    return 0;
}

} /* extern "C" */
