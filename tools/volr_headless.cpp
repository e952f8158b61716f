// volr_headless.cpp - headless C++ driver: the reference's draw_volume / benchmark_config_loop (VolR.cpp:98-113,
// :225-253) without GLUT, GLUI or the PBO. Builds the scene with the vr5_scene_* builders, constructs a GPURenderer5
// exactly as VolR constructs its renderers (VolR.cpp:425-429), renders 2 projections x 4 benchmark poses into a host
// buffer through Renderer::render_volume, times each frame (wall clock around the blocking call, like Profiler method 0
// but with a monotonic clock) and optionally dumps PPM images.
//
//   volr_headless [-n N] [-synth A|B] [-raw file dx dy dz] [-s W H] [-tf default|semi|opaque] [-ppm prefix] [-dump prefix]
//                 [-device] [-reps R]
// -device: the frame buffer is a cudaMalloc block, as the GL application hands renderers >= 3 the mapped pixel buffer
//          (VolR.cpp:76-87); render_volume finds out by itself which kind of pointer it got (vr5_render, out_is_device = -1).
// -dump:   raw RGBA8 frames, row 0 first, one file per pose (tests compare their sha256 with the reference's frames).
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include <vector>

#include <cuda_runtime_api.h>

#include "../include/vr5.h"
#include "../volume-rendering_b200/csrc/GPURenderer5.h"

static double now_ms() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}

int main(int argc, char **argv) {
    int n = 256, w = 800, h = 600, reps = 3, synth = 0;
    const char *tf_name = "default", *ppm = NULL, *raw = NULL, *dump = NULL;
    int rdx = 0, rdy = 0, rdz = 0, device_buffer = 0;
    for (int i = 1; i < argc; i++) {
        if (!strcmp(argv[i], "-n") && i + 1 < argc) n = atoi(argv[++i]);
        else if (!strcmp(argv[i], "-synth") && i + 1 < argc) synth = argv[++i][0] == 'B';
        else if (!strcmp(argv[i], "-raw") && i + 4 < argc) { raw = argv[++i]; rdx = atoi(argv[++i]); rdy = atoi(argv[++i]); rdz = atoi(argv[++i]); }
        else if (!strcmp(argv[i], "-s") && i + 2 < argc) { w = atoi(argv[++i]); h = atoi(argv[++i]); }
        else if (!strcmp(argv[i], "-tf") && i + 1 < argc) tf_name = argv[++i];
        else if (!strcmp(argv[i], "-ppm") && i + 1 < argc) ppm = argv[++i];
        else if (!strcmp(argv[i], "-reps") && i + 1 < argc) reps = atoi(argv[++i]);
        else if (!strcmp(argv[i], "-dump") && i + 1 < argc) dump = argv[++i];
        else if (!strcmp(argv[i], "-device")) device_buffer = 1;
        else { fprintf(stderr, "usage: %s [-n N] [-synth A|B] [-raw file dx dy dz] [-s W H] [-tf default|semi|opaque] [-ppm prefix] [-dump prefix] [-device] [-reps R]\n", argv[0]); return 2; }
    }
    // ---- volume on the host (ModelBase::volume) ----
    std::vector<uint8_t> vol;
    uint16_t dx, dy, dz;
    vr5_renderer *gen = NULL;
    if (raw) {
        dx = (uint16_t)rdx; dy = (uint16_t)rdy; dz = (uint16_t)rdz;
        vol.resize((size_t)dx * dy * dz);
        FILE *f = fopen(raw, "rb");
        if (!f || fread(vol.data(), 1, vol.size(), f) != vol.size()) { fprintf(stderr, "cannot read %s\n", raw); return 1; }
        fclose(f);
    } else {  // synthetic volume from the device generator, copied back so the scene goes through the host interface
        dx = dy = dz = (uint16_t)n;
        vol.resize((size_t)n * n * n);
        if (vr5_create(-1, &gen) || vr5_synth_volume(gen, synth, (uint16_t)n, synth ? 2 : 1) || vr5_get_volume(gen, vol.data(), vol.size())) {
            fprintf(stderr, "synthetic volume: %s\n", vr5_last_error(gen));
            return 1;
        }
        vr5_destroy(gen);
    }
    // ---- RaycasterBase::raycaster equivalents ----
    std::vector<float> base(512), tf(512);
    std::vector<uint8_t> minmax(65536);
    std::vector<uint32_t> esl(1024);
    vr5_scene_default_base_tf(base.data());
    const int longest = dx > dy ? (dx > dz ? dx : dz) : (dy > dz ? dy : dz);
    if (!strcmp(tf_name, "semi")) for (int i = 0; i < 128; i++) base[4 * i + 3] = 1.0f / (float)longest;
    if (!strcmp(tf_name, "opaque")) for (int i = 0; i < 128; i++) base[4 * i + 3] = i >= 64 ? 1.0f : 0.0f;
    vr5_scene_premultiply_tf(base.data(), tf.data());
    Raycaster rc;
    memset(&rc, 0, sizeof(rc));
    vr5_f3 bs;
    vr5_scene_build_minmax(vol.data(), dx, dy, dz, &rc.esl_block_dims, minmax.data(), &bs);
    vr5_scene_build_esl_bits(tf.data(), minmax.data(), esl.data());
    rc.volume.data = vol.data(); rc.volume.size = (unsigned)vol.size();
    rc.volume.dims.x = dx; rc.volume.dims.y = dy; rc.volume.dims.z = dz;
    rc.volume.min_bound.x = rc.volume.min_bound.y = rc.volume.min_bound.z = -1.0f;
    rc.transfer_fn = reinterpret_cast<float4 *>(tf.data());
    rc.esl_volume = esl.data();
    rc.ray_step = vr5_scene_default_ray_step(dx, dy, dz);
    rc.ray_threshold = 0.95f; rc.esl = true; rc.light_kd = 0.6f;  // RaycasterBase.cpp:9-20
    rc.esl_block_size.x = bs.x; rc.esl_block_size.y = bs.y; rc.esl_block_size.z = bs.z;
    vr5_view v0;
    vr5_scene_camera((uint16_t)w, (uint16_t)h, 0, 0, 0, 0, 2.0f, &v0);
    memcpy(&rc.view, &v0, sizeof(v0));

    Renderer *renderer = new GPURenderer5(rc);  // VolR.cpp:425-429
    printf("renderer: %s\nvolume %dx%dx%d, viewport %dx%d, tf %s, ray_step %.6g\n", renderer->get_name(), dx, dy, dz, w, h, tf_name, rc.ray_step);
    std::vector<uint32_t> frame((size_t)w * h);
    uchar4 *dev_frame = NULL;
    if (device_buffer && cudaMalloc((void **)&dev_frame, frame.size() * 4) != cudaSuccess) { fprintf(stderr, "cudaMalloc frame buffer failed\n"); return 1; }
    uchar4 *target = device_buffer ? dev_frame : reinterpret_cast<uchar4 *>(frame.data());
    const float poses[4][3] = {{0, 0, 0}, {-45, -45, 0}, {90, 0, 0}, {180, 90, 0}};  // VolR.cpp:235-244
    double sum = 0, worst = 0;
    int samples = 0;
    for (int persp = 0; persp < 2; persp++)
        for (int p = 0; p < 4; p++) {
            vr5_view v;
            vr5_scene_camera((uint16_t)w, (uint16_t)h, persp, poses[p][0], poses[p][1], poses[p][2], 2.0f, &v);
            memcpy(&rc.view, &v, sizeof(v));  // RaycasterBase::set_view, VolR.cpp:107
            double best = 1e30;
            for (int r = 0; r < reps; r++) {
                const double t0 = now_ms();
                if (renderer->render_volume(target, rc) != 0) { fprintf(stderr, "render_volume failed\n"); return 1; }
                const double t = now_ms() - t0;
                if (t < best) best = t;
            }
            if (device_buffer && cudaMemcpy(frame.data(), dev_frame, frame.size() * 4, cudaMemcpyDeviceToHost) != cudaSuccess) { fprintf(stderr, "cudaMemcpy failed\n"); return 1; }
            if (dump) {
                char name[512];
                snprintf(name, sizeof(name), "%s_%s_%d.rgba", dump, persp ? "persp" : "ortho", p);
                FILE *f = fopen(name, "wb");
                if (!f || fwrite(frame.data(), 4, frame.size(), f) != frame.size()) { fprintf(stderr, "cannot write %s\n", name); return 1; }
                fclose(f);
            }
            size_t lit = 0;
            for (uint32_t px : frame) lit += px != 0;
            printf("%s pose (%4.0f,%4.0f,%4.0f): %8.3f ms  (%zu non-zero pixels)\n", persp ? "persp" : "ortho", poses[p][0], poses[p][1], poses[p][2], best, lit);
            sum += best; worst = best > worst ? best : worst; samples++;
            if (ppm) {
                char name[512];
                snprintf(name, sizeof(name), "%s_%s_%d.ppm", ppm, persp ? "persp" : "ortho", p);
                FILE *f = fopen(name, "wb");
                if (f) {
                    fprintf(f, "P6\n%d %d\n255\n", w, h);
                    for (int y = h - 1; y >= 0; y--)  // row 0 is the bottom row on screen (GL texture upload, UI.cpp:102-115)
                        for (int x = 0; x < w; x++) fwrite(&frame[(size_t)y * w + x], 1, 3, f);
                    fclose(f);
                }
            }
        }
    printf("avg %.3f ms, max %.3f ms over %d frames (%s)\n", sum / samples, worst, samples,
           device_buffer ? "device buffer, blocking" : "host buffer, blocking, D2H included");
    delete renderer;
    cudaFree(dev_frame);
    return 0;
}
