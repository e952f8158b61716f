// GPURenderer5.h - `class GPURenderer5 : public Renderer`, the sixth entry of VolR's renderer list.
//
// Built INSIDE the reference tree it includes the reference's own "Renderer.h" (define VR5_WITH_VOLR_HEADERS, see
// INTEGRATION.md). Built standalone (this repository, the headless driver) it uses the layout-identical declarations
// of volr_compat.h instead. Either way it is a thin adaptor from the Renderer virtuals (Renderer.h:13-28) to the C ABI
// of include/vr5.h; every device resource lives behind the vr5_renderer handle it owns (unlike
// GPURenderer2-4 it does not lean on GPURenderer1's statics, Renderer.h:37-44).
#ifndef VR5_GPURENDERER5_H
#define VR5_GPURENDERER5_H

#ifdef VR5_WITH_VOLR_HEADERS
#include "Renderer.h"
#else
#include "volr_compat.h"
#endif

struct vr5_renderer;

class GPURenderer5 : public Renderer {
  public:
    GPURenderer5(Raycaster r);  // like GPURenderer1.cu:17-21: set_window_buffer, set_transfer_fn, set_volume
    virtual ~GPURenderer5();
    virtual const char *get_name();
    virtual void set_window_buffer(View view);
    virtual void set_transfer_fn(Raycaster r);
    virtual int set_volume(Model volume);
    // `buffer` may be a host pointer (glMapBuffer path, renderer ids < 3) or a device pointer (CUDA-mapped PBO, ids >= 3,
    // VolR.cpp:76-87): detected with cudaPointerGetAttributes inside vr5_render (out_is_device = -1).
    virtual int render_volume(uchar4 *buffer, Raycaster r);
    const char *last_error() const;
    vr5_renderer *handle() { return handle_; }

  private:
    vr5_renderer *handle_;
};

#endif
