/*
 * vr5_pod.h - plain-C mirrors of the three POD structs that cross the reference's Renderer boundary.
 *
 * The reference passes the whole scene BY VALUE as `struct Raycaster` (RaycasterBase.h:20-30), which
 * embeds `struct Model` (ModelBase.h:11-16) and `struct View` (ViewBase.h:14-21). The mirrors below are
 * layout-identical on x86-64 SysV (sizes 32 / 68 / 160 bytes; offsets listed per field and enforced by the
 * static asserts at the bottom, and cross-checked against the reference's own headers in
 * tests/test_boundary.py), so a `Raycaster` can be reinterpreted as a `vr5_raycaster` and vice versa.
 *
 * No CUDA, C++ or torch types appear here: this header is shared by the C-ABI (vr5.h), the CPU oracle
 * (oracle/) and the C++ GPURenderer5 adaptor.
 */
#ifndef VR5_POD_H
#define VR5_POD_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VR5_TF_SIZE 128          /* RaycasterBase.h:15  TF_SIZE */
#define VR5_TF_RATIO 2           /* RaycasterBase.h:16  TF_RATIO = 256 / TF_SIZE */
#define VR5_ESL_VOLUME_DIMS 32   /* RaycasterBase.h:13  ESL_VOLUME_DIMS */
#define VR5_ESL_VOLUME_SIZE 1024 /* RaycasterBase.h:14  32^3 bits / 32 */
#define VR5_ESL_MIN_BLOCK_SIZE 8 /* RaycasterBase.h:12 */

typedef struct vr5_f3 { float x, y, z; } vr5_f3;
typedef struct vr5_f4 { float x, y, z, w; } vr5_f4;
typedef struct vr5_us2 { uint16_t x, y; } vr5_us2;
typedef struct vr5_us3 { uint16_t x, y, z; } vr5_us3;

/* ModelBase.h:11-16 */
typedef struct vr5_model {
    uint8_t *data;    /* @0   x fastest, then y, then z: index z*dx*dy + y*dx + x (ModelBase.h:18-22) */
    uint32_t size;    /* @8   bytes; wraps at 2048^3 in the reference - vr5.h carries 64 bit */
    vr5_us3 dims;     /* @12 */
    vr5_f3 min_bound; /* @20  always (-1,-1,-1) (ModelBase.cpp:10-14) */
} vr5_model;

/* ViewBase.h:14-21 */
typedef struct vr5_view {
    vr5_us2 dims;        /* @0 */
    vr5_f3 origin;       /* @4 */
    vr5_f3 direction;    /* @16 */
    vr5_f3 right_plane;  /* @28  pre-scaled by world units per pixel (ViewBase.cpp:52-54) */
    vr5_f3 up_plane;     /* @40 */
    vr5_f3 light_pos;    /* @52 */
    uint8_t perspective; /* @64  C++ bool */
    uint8_t _pad[3];
} vr5_view;

/* RaycasterBase.h:20-30 */
typedef struct vr5_raycaster {
    vr5_model volume;        /* @0 */
    vr5_view view;           /* @32 */
    uint32_t _pad0;          /* @100 */
    vr5_f4 *transfer_fn;     /* @104 128 pre-multiplied RGBA entries (RaycasterBase.cpp:46-52) */
    float ray_step;          /* @112 */
    float ray_threshold;     /* @116 */
    uint8_t esl;             /* @120 C++ bool */
    uint8_t _pad1[7];
    uint32_t *esl_volume;    /* @128 1024 words, bit set = block empty (RaycasterBase.cpp:62-67) */
    uint16_t esl_block_dims; /* @136 */
    uint16_t _pad2;
    vr5_f3 esl_block_size;   /* @140 */
    float light_kd;          /* @152 */
    uint32_t _pad3;
} vr5_raycaster;

#ifdef __cplusplus
#define VR5_STATIC_ASSERT(c, m) static_assert(c, m)
#else
#define VR5_STATIC_ASSERT(c, m) _Static_assert(c, m)
#endif
VR5_STATIC_ASSERT(sizeof(vr5_model) == 32, "Model is 32 bytes");
VR5_STATIC_ASSERT(offsetof(vr5_model, size) == 8 && offsetof(vr5_model, dims) == 12 && offsetof(vr5_model, min_bound) == 20, "Model offsets");
VR5_STATIC_ASSERT(sizeof(vr5_view) == 68, "View is 68 bytes");
VR5_STATIC_ASSERT(offsetof(vr5_view, origin) == 4 && offsetof(vr5_view, direction) == 16 && offsetof(vr5_view, right_plane) == 28 &&
                  offsetof(vr5_view, up_plane) == 40 && offsetof(vr5_view, light_pos) == 52 && offsetof(vr5_view, perspective) == 64, "View offsets");
VR5_STATIC_ASSERT(sizeof(vr5_raycaster) == 160, "Raycaster is 160 bytes");
VR5_STATIC_ASSERT(offsetof(vr5_raycaster, view) == 32 && offsetof(vr5_raycaster, transfer_fn) == 104 && offsetof(vr5_raycaster, ray_step) == 112 &&
                  offsetof(vr5_raycaster, ray_threshold) == 116 && offsetof(vr5_raycaster, esl) == 120 && offsetof(vr5_raycaster, esl_volume) == 128 &&
                  offsetof(vr5_raycaster, esl_block_dims) == 136 && offsetof(vr5_raycaster, esl_block_size) == 140 &&
                  offsetof(vr5_raycaster, light_kd) == 152, "Raycaster offsets");

#ifdef __cplusplus
}
#endif
#endif /* VR5_POD_H */
