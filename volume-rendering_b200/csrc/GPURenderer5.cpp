// GPURenderer5.cpp - adaptor from the reference's Renderer virtuals to the vr5 C ABI (include/vr5.h).
// Error convention of the reference (cuda_utils.h:21-49, GPURenderer1.cu:86-112): log through Logger-style printf,
// return 1 on a missing input or failed allocation; never throw. CUDA failures are reported, not fatal (the C layer never
// calls exit(); a VolR build may wrap the non-zero return in its own cuda_safe_call policy).
#include "GPURenderer5.h"

#include <stdio.h>
#include <string.h>

#include "../../include/vr5.h"

static_assert(sizeof(Raycaster) == sizeof(vr5_raycaster) && sizeof(View) == sizeof(vr5_view) && sizeof(Model) == sizeof(vr5_model),
              "vr5_pod.h mirrors must match the VolR structs");

static void report(vr5_renderer *h, const char *where) {
    fprintf(stderr, "GPURenderer5 %s: %s\n", where, vr5_last_error(h));
}

GPURenderer5::GPURenderer5(Raycaster r) : handle_(NULL) {
    if (vr5_create(-1, &handle_) != 0) {
        report(NULL, "create");
        handle_ = NULL;
        return;
    }
    set_window_buffer(r.view);
    set_transfer_fn(r);
    set_volume(r.volume);
}

GPURenderer5::~GPURenderer5() { vr5_destroy(handle_); }

const char *GPURenderer5::get_name() { return vr5_name(); }

const char *GPURenderer5::last_error() const { return vr5_last_error(handle_); }

void GPURenderer5::set_window_buffer(View view) {
    if (!handle_) return;
    if (vr5_set_window(handle_, view.dims.x, view.dims.y) != 0) report(handle_, "set_window_buffer");
}

void GPURenderer5::set_transfer_fn(Raycaster r) {
    if (!handle_ || r.transfer_fn == NULL || r.esl_volume == NULL) return;
    if (vr5_set_transfer_fn(handle_, reinterpret_cast<const float *>(r.transfer_fn), r.esl_volume) != 0) report(handle_, "set_transfer_fn");
}

int GPURenderer5::set_volume(Model volume) {
    if (!handle_) return 1;
    // Model::size is 32 bit in the reference and wraps at 2048^3: trust the dims
    const unsigned long long bytes = (unsigned long long)volume.dims.x * volume.dims.y * volume.dims.z;
    const int rc = vr5_set_volume(handle_, volume.data, bytes, volume.dims.x, volume.dims.y, volume.dims.z);
    if (rc != 0 && volume.data != NULL) report(handle_, "set_volume");
    return rc;
}

int GPURenderer5::render_volume(uchar4 *buffer, Raycaster r) {
    if (!handle_ || buffer == NULL) return 1;
    vr5_raycaster pod;
    memcpy(&pod, &r, sizeof(pod));
    const int rc = vr5_render_raycaster(handle_, &pod, buffer, -1);
    if (rc != 0) report(handle_, "render_volume");
    return rc;
}
