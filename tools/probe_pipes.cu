// probe_pipes.cu - issue rates of the instruction kinds in the march loop, alone and mixed, per SM sub-partition.
// Answers (profiles/r01_pipe_rates.md): is the fma pipe 1 or 0.5 warp-instructions per clock for FADD/FMUL/FMUL2/IMAD,
// and which ALU-pipe replacements co-issue with it.
#include <cuda_runtime.h>
#include <stdio.h>
#define REP8(X) X(0) X(1) X(2) X(3) X(4) X(5) X(6) X(7)
template <int MODE> __global__ void k(unsigned *out, float seed, unsigned iseed, int iters, long long *cycles) {
    float f0 = seed + threadIdx.x, f1 = f0 + 1, f2 = f0 + 2, f3 = f0 + 3, f4 = f0 + 4, f5 = f0 + 5, f6 = f0 + 6, f7 = f0 + 7;
    unsigned i0 = iseed + threadIdx.x, i1 = i0 + 1, i2 = i0 + 2, i3 = i0 + 3, i4 = i0 + 4, i5 = i0 + 5, i6 = i0 + 6, i7 = i0 + 7;
    unsigned long long p0, p1, p2, p3, c;
    asm("mov.b64 %0, {%1,%2};" : "=l"(p0) : "f"(f0), "f"(f1)); asm("mov.b64 %0, {%1,%2};" : "=l"(p1) : "f"(f2), "f"(f3));
    asm("mov.b64 %0, {%1,%2};" : "=l"(p2) : "f"(f4), "f"(f5)); asm("mov.b64 %0, {%1,%2};" : "=l"(p3) : "f"(f6), "f"(f7));
    asm("mov.b64 %0, {%1,%2};" : "=l"(c) : "f"(seed), "f"(seed));
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < iters; i++) {
#define FADD(n) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(f##n) : "f"(seed));
#define FMUL(n) asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(f##n) : "f"(seed));
#define FADDRZ(n) asm volatile("add.rz.f32 %0, %0, %1;" : "+f"(f##n) : "f"(seed));
#define FFMA(n) asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+f"(f##n) : "f"(seed));
#define IMAD(n) asm volatile("mad.lo.u32 %0, %0, %1, %1;" : "+r"(i##n) : "r"(iseed));
#define LEA(n) asm volatile("shf.l.wrap.b32 %0, %0, %1, 3;" : "+r"(i##n) : "r"(iseed));
#define IADD(n) asm volatile("add.u32 %0, %0, %1;" : "+r"(i##n) : "r"(iseed));
#define F2I(n) asm volatile("cvt.rzi.u32.f32 %0, %1;" : "=r"(i##n) : "f"(f##n));
#define FSETP(n) asm volatile("{.reg .pred p; setp.gt.f32 p, %1, %2; selp.u32 %0, %0, %3, p;}" : "+r"(i##n) : "f"(f##n), "f"(seed), "r"(iseed));
        if (MODE == 0) { REP8(FADD) }
        if (MODE == 1) { REP8(FMUL) }
        if (MODE == 2) { REP8(FFMA) }
        if (MODE == 3) { REP8(FADDRZ) }
        if (MODE == 4) { REP8(IMAD) }
        if (MODE == 5) { REP8(LEA) }
        if (MODE == 6) { REP8(IADD) }
        if (MODE == 7) { REP8(F2I) }
        if (MODE == 8) {  // 4 FADD2 + 4 FMUL2
            asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p0) : "l"(c)); asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p1) : "l"(c));
            asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p2) : "l"(c)); asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p3) : "l"(c));
            asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p0) : "l"(c)); asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p1) : "l"(c));
            asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p2) : "l"(c)); asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p3) : "l"(c));
        }
        if (MODE == 9) { FADD(0) IADD(0) FADD(1) IADD(1) FADD(2) IADD(2) FADD(3) IADD(3) FADD(4) IADD(4) FADD(5) IADD(5) FADD(6) IADD(6) FADD(7) IADD(7) }  // 8 fma + 8 alu
        if (MODE == 10) { FADD(0) IMAD(0) FADD(1) IMAD(1) FADD(2) IMAD(2) FADD(3) IMAD(3) FADD(4) IMAD(4) FADD(5) IMAD(5) FADD(6) IMAD(6) FADD(7) IMAD(7) }  // 16 fma-pipe
        if (MODE == 11) { REP8(FADD) REP8(FMUL) }  // 16 fp
        if (MODE == 12) { FADD(0) F2I(0) FADD(1) F2I(1) FADD(2) F2I(2) FADD(3) F2I(3) FADD(4) FADD(5) FADD(6) FADD(7) }  // 8 fma + 4 xu
    }
    long long t1 = clock64();
    float s = f0 + f1 + f2 + f3 + f4 + f5 + f6 + f7;
    out[blockIdx.x * blockDim.x + threadIdx.x] = (unsigned)(p0 ^ p1 ^ p2 ^ p3) ^ __float_as_uint(s) ^ i0 ^ i1 ^ i2 ^ i3 ^ i4 ^ i5 ^ i6 ^ i7;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}
template <int MODE> void run(const char *name, int n_instr) {
    unsigned *out; long long *cyc, h;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
    const int iters = 20000;
    for (int warps = 8; warps <= 32; warps *= 2) {  // warps per SM, one CTA per SM -> warps/4 per sub-partition
        k<MODE><<<148, warps * 32>>>(out, 1.0001f, 3u, iters, cyc); cudaDeviceSynchronize();
        k<MODE><<<148, warps * 32>>>(out, 1.0001f, 3u, iters, cyc); cudaDeviceSynchronize();
        cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
        printf("%-22s warps/SMSP=%d  warp-instr/clk/SMSP=%.3f\n", name, warps / 4, (double)n_instr * (warps / 4) * iters / h);
    }
    cudaFree(out); cudaFree(cyc);
}
int main() {
    run<0>("FADD r,r", 8); run<1>("FMUL r,r", 8); run<2>("FFMA r,r,r", 8); run<3>("FADD.RZ", 8); run<4>("IMAD", 8); run<5>("SHF (alu)", 8);
    run<6>("IADD (alu)", 8); run<7>("F2I (xu)", 8); run<8>("FADD2 packed", 8); run<9>("FADD+IADD mix", 16); run<10>("FADD+IMAD mix", 16);
    run<11>("FADD+FMUL", 16); run<12>("8 FADD + 4 F2I", 12);
    return 0;
}
