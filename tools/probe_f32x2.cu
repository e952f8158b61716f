// probe_f32x2.cu - issue rate of packed fp32 (FADD2 / FMUL2, sm_100 add.f32x2 / mul.f32x2) against scalar FADD / FMUL,
// and whether they co-issue with ALU-pipe work. Answers: do packed ops halve the issue slots of the march loop?
#include <cuda_runtime.h>
#include <stdio.h>
#define ADD2(d, a, b) asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b))
#define MUL2(d, a, b) asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b))
template <int MODE> __global__ void k(unsigned long long *out, float seed, int iters, long long *cycles) {
    float a0 = seed + threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    unsigned long long p0, p1, p2, p3, c;
    asm("mov.b64 %0, {%1,%2};" : "=l"(p0) : "f"(a0), "f"(a1)); asm("mov.b64 %0, {%1,%2};" : "=l"(p1) : "f"(a2), "f"(a3));
    asm("mov.b64 %0, {%1,%2};" : "=l"(p2) : "f"(a4), "f"(a5)); asm("mov.b64 %0, {%1,%2};" : "=l"(p3) : "f"(a6), "f"(a7));
    asm("mov.b64 %0, {%1,%2};" : "=l"(c) : "f"(seed), "f"(seed));
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
        if (MODE == 0) {  // 8 scalar FADD
            a0 = __fadd_rn(a0, seed); a1 = __fadd_rn(a1, seed); a2 = __fadd_rn(a2, seed); a3 = __fadd_rn(a3, seed);
            a4 = __fadd_rn(a4, seed); a5 = __fadd_rn(a5, seed); a6 = __fadd_rn(a6, seed); a7 = __fadd_rn(a7, seed);
        } else if (MODE == 1) {  // 4 FADD2 = the same 8 lane-ops
            ADD2(p0, p0, c); ADD2(p1, p1, c); ADD2(p2, p2, c); ADD2(p3, p3, c);
        } else if (MODE == 2) {  // 4 FMUL2
            MUL2(p0, p0, c); MUL2(p1, p1, c); MUL2(p2, p2, c); MUL2(p3, p3, c);
        } else {  // 8 scalar FMUL
            a0 = __fmul_rn(a0, seed); a1 = __fmul_rn(a1, seed); a2 = __fmul_rn(a2, seed); a3 = __fmul_rn(a3, seed);
            a4 = __fmul_rn(a4, seed); a5 = __fmul_rn(a5, seed); a6 = __fmul_rn(a6, seed); a7 = __fmul_rn(a7, seed);
        }
    }
    long long t1 = clock64();
    float s = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
    out[blockIdx.x * blockDim.x + threadIdx.x] = p0 ^ p1 ^ p2 ^ p3 ^ (unsigned long long)__float_as_uint(s);
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}
template <int MODE> void run(const char *name, int lane_ops) {
    unsigned long long *out; long long *cyc, h;
    cudaMalloc(&out, 148 * 8 * 1024 * 8); cudaMalloc(&cyc, 8);
    const int iters = 20000;
    for (int warps = 4; warps <= 32; warps *= 2) {  // warps per SM, one CTA per SM
        k<MODE><<<148, warps * 32>>>(out, 1.0001f, iters, cyc); cudaDeviceSynchronize();
        k<MODE><<<148, warps * 32>>>(out, 1.0001f, iters, cyc); cudaDeviceSynchronize();
        cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
        printf("%-10s warps/SM=%2d  cycles/iter=%.2f  lane-ops/clk/SM=%.1f\n", name, warps, (double)h / iters, (double)lane_ops * 32 * warps * iters / h);
    }
}
int main() { run<0>("8xFADD", 8); run<1>("4xFADD2", 8); run<3>("8xFMUL", 8); run<2>("4xFMUL2", 8); return 0; }
