// probe_tld.cu - does an integer-coordinate texture fetch (tex.3d ... .s32 -> SASS TLD) clamp out-of-range coordinates
// to the edge when the texture object uses cudaAddressModeClamp? Decides whether the kernel needs its own upper clamp.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdint.h>
#include <string.h>
__global__ void probe(cudaTextureObject_t tex, const int *xyz, int n, uint32_t *out) {
    int i = threadIdx.x;
    if (i >= n) return;
    uint32_t r0, r1, r2, r3;
    asm("tex.3d.v4.u32.s32 {%0, %1, %2, %3}, [%4, {%5, %6, %7, %8}];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
        : "l"(tex), "r"(xyz[3 * i]), "r"(xyz[3 * i + 1]), "r"(xyz[3 * i + 2]), "r"(0));
    out[i] = r0;
}
int main() {
    const int nx = 5, ny = 4, nz = 3;
    uint8_t h[nx * ny * nz];
    for (int i = 0; i < nx * ny * nz; i++) h[i] = (uint8_t)(i + 1);
    cudaArray_t arr; cudaChannelFormatDesc d = cudaCreateChannelDesc(8, 0, 0, 0, cudaChannelFormatKindUnsigned);
    cudaMalloc3DArray(&arr, &d, make_cudaExtent(nx, ny, nz), 0);
    cudaMemcpy3DParms cp; memset(&cp, 0, sizeof(cp));
    cp.srcPtr = make_cudaPitchedPtr(h, nx, nx, ny); cp.dstArray = arr; cp.extent = make_cudaExtent(nx, ny, nz); cp.kind = cudaMemcpyHostToDevice;
    cudaMemcpy3D(&cp);
    cudaResourceDesc res; memset(&res, 0, sizeof(res)); res.resType = cudaResourceTypeArray; res.res.array.array = arr;
    cudaTextureDesc td; memset(&td, 0, sizeof(td));
    td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp; td.filterMode = cudaFilterModePoint; td.readMode = cudaReadModeElementType;
    cudaTextureObject_t tex; cudaCreateTextureObject(&tex, &res, &td, nullptr);
    int q[][3] = {{0,0,0},{4,3,2},{5,3,2},{4,4,2},{4,3,3},{6,5,4},{100,0,0},{-1,0,0},{0,-1,0},{0,0,-1},{-5,-5,-5},{1<<20,1,1},{2,2,1}};
    const int n = sizeof(q) / sizeof(q[0]);
    int *dq; uint32_t *dout, out[n];
    cudaMalloc(&dq, sizeof(q)); cudaMalloc(&dout, sizeof(out)); cudaMemcpy(dq, q, sizeof(q), cudaMemcpyHostToDevice);
    probe<<<1, 32>>>(tex, dq, n, dout);
    cudaError_t e = cudaMemcpy(out, dout, sizeof(out), cudaMemcpyDeviceToHost);
    printf("status %s\n", cudaGetErrorString(e));
    int ok = 1;
    for (int i = 0; i < n; i++) {
        int cx = q[i][0] < 0 ? 0 : (q[i][0] >= nx ? nx - 1 : q[i][0]), cy = q[i][1] < 0 ? 0 : (q[i][1] >= ny ? ny - 1 : q[i][1]), cz = q[i][2] < 0 ? 0 : (q[i][2] >= nz ? nz - 1 : q[i][2]);
        int want = h[(cz * ny + cy) * nx + cx];
        printf("(%d,%d,%d) -> %u  clamp-to-edge would give %d %s\n", q[i][0], q[i][1], q[i][2], out[i], want, (int)out[i] == want ? "" : "DIFFERENT");
        ok &= (int)out[i] == want;
    }
    printf("TLD_CLAMPS_TO_EDGE=%d\n", ok);
    return 0;
}
