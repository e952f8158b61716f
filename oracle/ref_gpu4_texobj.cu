/*
 * TEST / MEASUREMENT INFRASTRUCTURE ONLY (oracle build). Not part of the product path.
 *
 * The reference's GPURenderer4 ("CUDA CM + 3D Texture Memory + GLI", GPURenderer4.cu) compiled UNMODIFIED for sm_100a:
 * this TU supplies what CUDA 12 removed - the legacy texture REFERENCE API - on top of texture OBJECTS, then includes
 * the reference's file from where it lies under /root/reference/VolumeRendering (oracle/Makefile passes the -I). Nothing
 * of the reference is copied into this repository.
 *
 * What the shim provides (and nothing else):
 *   texture<T, D, readMode> name;      -> a __managed__ POD holding the reference's own settings (normalized, filterMode,
 *                                         addressMode[]) and the cudaTextureObject_t created from them at bind time
 *   cudaBindTextureToArray(ref, array, desc) / cudaUnbindTexture(ref)
 *   tex1D / tex2D / tex3D(ref, ...)    -> tex1D / tex2D / tex3D<T>(ref.obj, ...): the SAME hardware fetch
 * so the kernel body (GPURenderer4.cu:33-87) and its texture set-up (GPURenderer4.cu:89-143: volume = 3-D uchar array,
 * cudaFilterModeLinear, normalised coordinates, cudaReadModeNormalizedFloat, clamp; transfer function = 1-D float4 array,
 * linear, normalised, clamp; ESL bitfield = 2-D uint array, point) run exactly as the reference wrote them, in the B200's
 * texture unit (fixed-point filter weights and all). CUDA compute only - no graphics API is involved: render_volume()
 * takes any device pointer (ref_gpu_api.cu hands it a cudaMalloc block instead of the mapped GL pixel buffer).
 *
 * Purpose: the external parity anchor of the trilinear sampler policies (VR5_SAMPLER_TRILINEAR / _FMA): tests/
 * test_trilinear_pin.py compares their frames with this renderer's under a stated LSB tolerance.
 */
#include <cuda_runtime.h>
#include <string.h>

template <class T, int DIM, enum cudaTextureReadMode MODE>
struct Ref4TexRef {  /* POD: __managed__ variables cannot be dynamically initialised */
    cudaTextureObject_t obj;
    int normalized;
    enum cudaTextureFilterMode filterMode;
    enum cudaTextureAddressMode addressMode[3];
};

template <class T, enum cudaTextureReadMode MODE> struct Ref4Fetch { typedef T type; };
template <> struct Ref4Fetch<unsigned char, cudaReadModeNormalizedFloat> { typedef float type; };

template <class T, int DIM, enum cudaTextureReadMode MODE>
static cudaError_t cudaBindTextureToArray(Ref4TexRef<T, DIM, MODE> &ref, cudaArray_t array, const cudaChannelFormatDesc &) {
    if (ref.obj) { cudaDestroyTextureObject(ref.obj); ref.obj = 0; }
    cudaResourceDesc res;
    memset(&res, 0, sizeof(res));
    res.resType = cudaResourceTypeArray;
    res.res.array.array = array;
    cudaTextureDesc td;
    memset(&td, 0, sizeof(td));
    for (int i = 0; i < 3; i++) td.addressMode[i] = ref.addressMode[i];
    td.filterMode = ref.filterMode;
    td.readMode = MODE;
    td.normalizedCoords = ref.normalized ? 1 : 0;
    cudaTextureObject_t obj = 0;
    const cudaError_t e = cudaCreateTextureObject(&obj, &res, &td, NULL);
    if (e == cudaSuccess) { ref.obj = obj; cudaDeviceSynchronize(); }
    return e;
}

template <class T, int DIM, enum cudaTextureReadMode MODE>
static cudaError_t cudaUnbindTexture(Ref4TexRef<T, DIM, MODE> &ref) {
    cudaError_t e = cudaSuccess;
    if (ref.obj) { cudaDeviceSynchronize(); e = cudaDestroyTextureObject(ref.obj); ref.obj = 0; }
    return e;
}

template <class T, enum cudaTextureReadMode MODE>
static __device__ __forceinline__ typename Ref4Fetch<T, MODE>::type tex1D(const Ref4TexRef<T, 1, MODE> &ref, float x) {
    return tex1D<typename Ref4Fetch<T, MODE>::type>(ref.obj, x);
}
template <class T, enum cudaTextureReadMode MODE>
static __device__ __forceinline__ typename Ref4Fetch<T, MODE>::type tex2D(const Ref4TexRef<T, 2, MODE> &ref, float x, float y) {
    return tex2D<typename Ref4Fetch<T, MODE>::type>(ref.obj, x, y);
}
template <class T, enum cudaTextureReadMode MODE>
static __device__ __forceinline__ typename Ref4Fetch<T, MODE>::type tex3D(const Ref4TexRef<T, 3, MODE> &ref, float x, float y, float z) {
    return tex3D<typename Ref4Fetch<T, MODE>::type>(ref.obj, x, y, z);
}

/* the reference's three file-scope declarations `texture<...> name;` become managed PODs: host code sets their fields
 * (GPURenderer4.cu:94-97, :107-110, :136-140), device code reads .obj */
#define texture __managed__ Ref4TexRef
#include "GPURenderer4.cu"
#undef texture
