/*
 * TEST INFRASTRUCTURE ONLY (oracle build). Not part of the product path.
 *
 * Thin extern "C" driver over the UNMODIFIED reference sources, compiled where they lie under
 * /root/reference/VolumeRendering by oracle/Makefile into oracle/_ref/libvolr_ref.so.
 * Nothing of the reference is copied into this repository: this TU only *includes* and *calls* it.
 *
 * It reproduces, headless, the call sequences of the reference application:
 *   - start-up            VolR.cpp:412-417  (load_model, set_view, reset_transfer_fn, set_volume)
 *   - per-frame           VolR.cpp:98-113   (RaycasterBase::set_view(ViewBase::view); render_volume)
 *   - benchmark poses     VolR.cpp:225-253  (toggle_perspective(true); set_camera_position(a, 2))
 *
 * `#include "ddsbase.h"
#include "CPURenderer.cpp"` pulls the reference's own translation unit in verbatim so that the
 * file-static `render_ray` (CPURenderer.cpp:11-41) can also be driven over a row range
 * (ref_render_rows) - same per-pixel code, different loop bounds - which the CPU-baseline timing
 * needs for bounded samples.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "CPURenderer.cpp" /* the reference TU itself (brings Renderer.h and all scene headers) */

extern "C" {

int ref_sizeof_raycaster(void) { return (int)sizeof(Raycaster); }
int ref_sizeof_view(void) { return (int)sizeof(View); }
int ref_sizeof_model(void) { return (int)sizeof(Model); }

/* ModelBase::load_model (ModelBase.cpp:35-109) followed by RaycasterBase::set_volume, as VolR.cpp:412-417 */
int ref_load_model(const char *path) {
    int err = ModelBase::load_model(path);
    if (err) return err;
    RaycasterBase::set_volume(ModelBase::volume);
    return 0;
}

/* Install a caller-provided uint8 volume (x fastest) exactly as load_model leaves ModelBase::volume. */
int ref_set_volume(const unsigned char *data, unsigned dx, unsigned dy, unsigned dz) {
    size_t n = (size_t)dx * dy * dz;
    unsigned char *copy = (unsigned char *)malloc(n);
    if (!copy) return 1;
    memcpy(copy, data, n);
    if (ModelBase::volume.data != NULL) free(ModelBase::volume.data);
    ModelBase::volume.dims = make_ushort3(dx, dy, dz);
    ModelBase::volume.size = (unsigned int)n;
    ModelBase::volume.data = copy;
    RaycasterBase::set_volume(ModelBase::volume);
    return 0;
}

void ref_get_volume_dims(unsigned *dx, unsigned *dy, unsigned *dz) {
    *dx = ModelBase::volume.dims.x; *dy = ModelBase::volume.dims.y; *dz = ModelBase::volume.dims.z;
}
const unsigned char *ref_get_volume_data(void) { return ModelBase::volume.data; }

/* RaycasterBase.cpp:76-84 */
void ref_reset_transfer_fn(void) { RaycasterBase::reset_transfer_fn(); }

/* base (non pre-multiplied) TF, 128 x RGBA, then RaycasterBase::update_transfer_fn (RaycasterBase.cpp:46-74) */
void ref_set_base_transfer_fn(const float *rgba128) {
    memcpy(RaycasterBase::base_transfer_fn, rgba128, TF_SIZE * sizeof(float4));
    RaycasterBase::update_transfer_fn();
}
void ref_get_base_transfer_fn(float *rgba128) { memcpy(rgba128, RaycasterBase::base_transfer_fn, TF_SIZE * sizeof(float4)); }
void ref_get_transfer_fn(float *rgba128) { memcpy(rgba128, RaycasterBase::raycaster.transfer_fn, TF_SIZE * sizeof(float4)); }
void ref_get_esl_volume(unsigned int *bits1024) { memcpy(bits1024, RaycasterBase::raycaster.esl_volume, ESL_VOLUME_SIZE * sizeof(esl_type)); }

/* GLUI writes these fields directly (UI.cpp:516-535); the change_* setters only add clamps + logging */
void ref_set_ray_step(float s) { RaycasterBase::raycaster.ray_step = s; }
void ref_reset_ray_step(void) { RaycasterBase::reset_ray_step(); }
void ref_set_ray_threshold(float t) { RaycasterBase::raycaster.ray_threshold = t; }
void ref_set_esl(int on) { RaycasterBase::raycaster.esl = on != 0; }
void ref_set_light_kd(float kd) { RaycasterBase::raycaster.light_kd = kd; }
void ref_get_ray_step_limits(float *lo, float *hi) { *lo = RaycasterBase::ray_step_limits.x; *hi = RaycasterBase::ray_step_limits.y; }

/* ViewBase.cpp:107-113 */
void ref_set_viewport(unsigned w, unsigned h) {
    ushort2 d; d.x = (unsigned short)w; d.y = (unsigned short)h;
    ViewBase::set_viewport_dims(d);
}
/* VolR.cpp:232-234: view.perspective = p; toggle_perspective(true) */
void ref_set_perspective(int p) {
    ViewBase::view.perspective = p != 0;
    ViewBase::toggle_perspective(1);
}
/* ViewBase.cpp:81-85 */
void ref_set_camera_position(float ax, float ay, float az, float distance) {
    ViewBase::set_camera_position(make_float3(ax, ay, az), distance);
}

/* draw_volume's RaycasterBase::set_view(ViewBase::view) (VolR.cpp:107), then hand out the 160-byte POD */
void ref_get_raycaster(void *out160) {
    RaycasterBase::set_view(ViewBase::view);
    memcpy(out160, &RaycasterBase::raycaster, sizeof(Raycaster));
}

/* CPURenderer::render_volume as shipped (CPURenderer.cpp:43-53); `raycaster160` may carry patched pointers */
int ref_render(const void *raycaster160, unsigned char *rgba8) {
    Raycaster r;
    memcpy(&r, raycaster160, sizeof(r));
    static CPURenderer renderer(RaycasterBase::raycaster);
    return renderer.render_volume((uchar4 *)rgba8, r);
}

/* Same per-pixel code (the reference's own static render_ray), rows [row0,row1) only; other rows untouched. */
int ref_render_rows(const void *raycaster160, unsigned char *rgba8, int row0, int row1) {
    Raycaster r;
    memcpy(&r, raycaster160, sizeof(r));
    if (r.volume.data == NULL || r.transfer_fn == NULL || r.esl_volume == NULL || rgba8 == NULL) return 1;
    if (row0 < 0) row0 = 0;
    if (row1 > r.view.dims.y) row1 = r.view.dims.y;
    uchar4 *buffer = (uchar4 *)rgba8;
    if (row1 > row0)
        memset(buffer + (size_t)row0 * r.view.dims.x, 0, (size_t)(row1 - row0) * r.view.dims.x * 4);
    for (int row = row0; row < row1; row++)
        for (int col = 0; col < r.view.dims.x; col++)
            render_ray(r, buffer, make_short2(col, row));
    return 0;
}

/* ddsbase.cpp:343-435 readPVMvolume as is: voxels only are copied out (caller's buffer of w*h*d*components bytes; pass
 * NULL to query the dimensions). Returns 0 ok / 1 not readable. */
int ref_read_pvm(const char *path, unsigned char *out, unsigned *w, unsigned *h, unsigned *d, unsigned *components) {
    unsigned char *v = readPVMvolume(path, w, h, d, components, NULL, NULL, NULL, NULL, NULL, NULL, NULL);
    if (v == NULL) return 1;
    if (out != NULL) memcpy(out, v, (size_t)*w * *h * *d * *components);
    free(v);
    return 0;
}

/* ddsbase.cpp:475-558 quantize as is (nofree = TRUE: the caller keeps its 16-bit buffer) */
int ref_quantize16(const unsigned char *data16_be, unsigned w, unsigned h, unsigned d, int linear, unsigned char *out8) {
    unsigned char *q = quantize((unsigned char *)data16_be, w, h, d, linear ? TRUE : FALSE, TRUE);
    if (q == NULL) return 1;
    memcpy(out8, q, (size_t)w * h * d);
    free(q);
    return 0;
}

} /* extern "C" */
