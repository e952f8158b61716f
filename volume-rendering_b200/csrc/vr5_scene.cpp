// vr5_scene.cpp - host-side scene builders of the headless driver: what the reference's static managers
// RaycasterBase (RaycasterBase.cpp:46-125) and ViewBase (ViewBase.cpp:26-113) manufacture for the hot path.
// Fresh code; the camera uses its own 3x3 rotation algebra instead of the OpenGL matrix stack the reference
// borrows as a math library (ViewBase.cpp:34-47). Outputs are checked against the reference's own managers in
// tests/test_scene_builders.py (tables bit-exact; camera vectors to fp32 rounding of sinf/cosf).
#include <math.h>
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "../../include/vr5.h"

// RaycasterBase.cpp:76-84: red ramp over the first third, green over the second, blue over the rest;
// opacity i/128 once the 8-bit sample exceeds 10 % of full scale (i > 12.75), else fully transparent.
extern "C" void vr5_scene_default_base_tf(float *base) {
    const int n = VR5_TF_SIZE, third = n / 3;
    const float alpha_cut = (255.0f * 0.1f) / (float)VR5_TF_RATIO;
    for (int i = 0; i < n; i++) {
        float r = 0.0f, g = 0.0f, b = 0.0f;
        if (i <= third) r = (float)(i * 3) / (float)n;
        else if (i <= 2 * third) g = (float)((i - third) * 3) / (float)n;
        else b = (float)((i - 2 * third) * 3) / (float)n;
        base[4 * i + 0] = r; base[4 * i + 1] = g; base[4 * i + 2] = b;
        base[4 * i + 3] = (float)i > alpha_cut ? (float)i / (float)n : 0.0f;
    }
}

// RaycasterBase.cpp:46-52
extern "C" void vr5_scene_premultiply_tf(const float *base, float *premult) {
    for (int i = 0; i < VR5_TF_SIZE; i++) {
        const float alpha = base[4 * i + 3];
        for (int c = 0; c < 3; c++) premult[4 * i + c] = base[4 * i + c] * alpha;
        premult[4 * i + 3] = alpha;
    }
}

// RaycasterBase.cpp:94-125. Row-wise running min/max per block column instead of a per-voxel table update.
extern "C" void vr5_scene_build_minmax(const uint8_t *data, uint16_t dx, uint16_t dy, uint16_t dz, uint16_t *esl_block_dims,
                                       uint8_t *minmax, vr5_f3 *esl_block_size) {
    const uint32_t longest = std::max<uint32_t>(dx, std::max<uint32_t>(dy, dz));
    const uint32_t bd = std::max<uint32_t>(VR5_ESL_MIN_BLOCK_SIZE, (longest + VR5_ESL_VOLUME_DIMS - 1) / VR5_ESL_VOLUME_DIMS);
    const uint32_t cells = VR5_ESL_VOLUME_DIMS * VR5_ESL_VOLUME_DIMS * VR5_ESL_VOLUME_DIMS;
    for (uint32_t c = 0; c < cells; c++) { minmax[2 * c] = 255; minmax[2 * c + 1] = 0; }
    for (uint32_t z = 0; z < dz; z++)
        for (uint32_t y = 0; y < dy; y++) {
            const uint8_t *line = data + ((uint64_t)z * dy + y) * dx;
            uint8_t *cell = minmax + 2 * (((z / bd) * VR5_ESL_VOLUME_DIMS + (y / bd)) * VR5_ESL_VOLUME_DIMS);
            for (uint32_t x0 = 0; x0 < dx; x0 += bd, cell += 2) {
                const uint32_t x1 = std::min<uint32_t>(dx, x0 + bd);
                uint8_t lo = cell[0], hi = cell[1];
                for (uint32_t x = x0; x < x1; x++) { lo = std::min(lo, line[x]); hi = std::max(hi, line[x]); }
                cell[0] = lo; cell[1] = hi;
            }
        }
    *esl_block_dims = (uint16_t)bd;
    esl_block_size->x = 2.0f * (float)bd / (float)dx;
    esl_block_size->y = 2.0f * (float)bd / (float)dy;
    esl_block_size->z = 2.0f * (float)bd / (float)dz;
}

// RaycasterBase.cpp:53-67: a block is empty when every TF entry its [min/2, max/2] range can reach has zero alpha.
extern "C" void vr5_scene_build_esl_bits(const float *premult, const uint8_t *minmax, uint32_t *bits) {
    int first_visible_from[VR5_TF_SIZE + 1];  // smallest j >= i with alpha[j] != 0, VR5_TF_SIZE if none
    first_visible_from[VR5_TF_SIZE] = VR5_TF_SIZE;
    for (int i = VR5_TF_SIZE - 1; i >= 0; i--) first_visible_from[i] = premult[4 * i + 3] != 0.0f ? i : first_visible_from[i + 1];
    for (int w = 0; w < VR5_ESL_VOLUME_SIZE; w++) {
        uint32_t word = 0;
        for (int b = 0; b < 32; b++) {
            const uint8_t *mm = minmax + 2 * (w * 32 + b);
            if (first_visible_from[mm[0] / VR5_TF_RATIO] > mm[1] / VR5_TF_RATIO) word |= 1u << b;
        }
        bits[w] = word;
    }
}

// RaycasterBase.cpp:86-92
extern "C" float vr5_scene_default_ray_step(uint16_t dx, uint16_t dy, uint16_t dz) {
    const float longest = (float)(int)std::max<uint32_t>(dx, std::max<uint32_t>(dy, dz));
    float step = 2.0f / longest;
    step -= step / longest;
    return step;
}

namespace {
struct Mat3 { float m[3][3]; };  // row-major, v' = M v

Mat3 axis_rotation(float deg, float x, float y, float z) {
    const float rad = deg * 3.14159265358979323846f / 180.0f;
    const float c = cosf(rad), s = sinf(rad), t = 1.0f - c;
    return Mat3{{{x * x * t + c, x * y * t - z * s, x * z * t + y * s},
                 {y * x * t + z * s, y * y * t + c, y * z * t - x * s},
                 {x * z * t - y * s, y * z * t + x * s, z * z * t + c}}};
}
Mat3 mul(const Mat3 &a, const Mat3 &b) {
    Mat3 o;
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            float acc = 0.0f;
            for (int k = 0; k < 3; k++) acc += a.m[i][k] * b.m[k][j];
            o.m[i][j] = acc;
        }
    return o;
}
}  // namespace

// ViewBase.cpp: set_viewport_dims (:107-113), toggle_perspective(update) (:100-105), set_camera_position (:81-85) ->
// camera_zoom (:74-79) + camera_rotate(reset) (:57-60) -> update_view (:49-55). After a reset the three glRotatef
// axes are the unit axes, so the camera matrix is Rx(ax) Ry(ay) Rz(az); vector_rotate (:26-32) dots a vector with
// the COLUMNS of that matrix, i.e. applies its transpose.
extern "C" void vr5_scene_camera(uint16_t width, uint16_t height, int perspective, float ax, float ay, float az, float distance,
                                 vr5_view *out) {
    memset(out, 0, sizeof(*out));
    out->dims.x = width; out->dims.y = height;
    out->perspective = perspective ? 1 : 0;
    out->light_pos = vr5_f3{0.0f, 0.0f, 3.0f};  // ViewBase.cpp:9-15 default, only the UI moves it
    const float cam_z = std::min(3.0f, std::max(0.1f, 0.0f + distance));  // distance_limits, ViewBase.cpp:17, :75
    const float view_size = perspective ? 1.5f : cam_z;
    const Mat3 m = mul(mul(axis_rotation(ax, 1, 0, 0), axis_rotation(ay, 0, 1, 0)), axis_rotation(az, 0, 0, 1));
    // transpose applied to (0, 0, cam_z), (1, 0, 0), (0, 1, 0)
    auto apply_t = [&m](float x, float y, float z) {
        return vr5_f3{x * m.m[0][0] + y * m.m[1][0] + z * m.m[2][0], x * m.m[0][1] + y * m.m[1][1] + z * m.m[2][1],
                      x * m.m[0][2] + y * m.m[1][2] + z * m.m[2][2]};
    };
    out->origin = apply_t(0.0f, 0.0f, cam_z);
    const vr5_f3 neg{-out->origin.x, -out->origin.y, -out->origin.z};
    const float inv_len = 1.0f / sqrtf(neg.x * neg.x + neg.y * neg.y + neg.z * neg.z);
    out->direction = vr5_f3{neg.x * inv_len, neg.y * inv_len, neg.z * inv_len};
    const float px = view_size / (float)std::min<int>(width, height);
    const vr5_f3 right = apply_t(1.0f, 0.0f, 0.0f), up = apply_t(0.0f, 1.0f, 0.0f);
    out->right_plane = vr5_f3{right.x * px, right.y * px, right.z * px};
    out->up_plane = vr5_f3{up.x * px, up.y * px, up.z * px};
}
