// volr_compat.h - stand-alone declarations of the types that cross VolR's Renderer boundary, for building GPURenderer5
// and the headless driver WITHOUT the reference tree. Data members only, in the reference's order and types, so the
// layout is the reference's (Model 32 B, View 68 B, Raycaster 160 B on x86-64; static_asserts below and
// tests/test_boundary.py, which compiles the same asserts against the reference's real headers when they are present).
// The reference's inline ray math (RaycasterBase.h:32-98, ViewBase.h:23-35, ModelBase.h:17-23) is deliberately absent:
// the adaptor never evaluates it on the host.
#ifndef VR5_VOLR_COMPAT_H
#define VR5_VOLR_COMPAT_H

#include <stddef.h>
#include <vector_types.h>  // CUDA toolkit: float3, float4, ushort2, ushort3, uchar4 (as common.h:12-14)

typedef unsigned int esl_type;  // RaycasterBase.h:18

struct Model {  // ModelBase.h:11-16
    unsigned char *data;
    unsigned int size;
    ushort3 dims;
    float3 min_bound;
};

struct View {  // ViewBase.h:14-21
    ushort2 dims;
    float3 origin;
    float3 direction;
    float3 right_plane;
    float3 up_plane;
    float3 light_pos;
    bool perspective;
};

struct Raycaster {  // RaycasterBase.h:20-30
    Model volume;
    View view;
    float4 *transfer_fn;
    float ray_step;
    float ray_threshold;
    bool esl;
    esl_type *esl_volume;
    unsigned short esl_block_dims;
    float3 esl_block_size;
    float light_kd;
};

class Renderer {  // Renderer.h:13-28
  public:
    virtual ~Renderer() {}
    virtual const char *get_name() { return "Default"; }
    virtual void set_window_buffer(View) {}
    virtual void set_transfer_fn(Raycaster) {}
    virtual int set_volume(Model) { return 0; }
    virtual int render_volume(uchar4 *buffer, Raycaster r) = 0;
};

static_assert(sizeof(Model) == 32 && sizeof(View) == 68 && sizeof(Raycaster) == 160, "VolR POD layout");
static_assert(offsetof(Raycaster, view) == 32 && offsetof(Raycaster, transfer_fn) == 104 && offsetof(Raycaster, esl_volume) == 128 &&
              offsetof(Raycaster, esl_block_dims) == 136 && offsetof(Raycaster, esl_block_size) == 140 && offsetof(Raycaster, light_kd) == 152,
              "Raycaster offsets");
#endif
