// tools/check_kleap.c - the arithmetic fact the leaps of the zero-skip kernels rest on (vr5_cast.cuh, try_leap), checked by brute force.
// Claim: while k stays inside one binade [2^e, 2^(e+1)), the reference's k += step (CPURenderer.cpp:37; fp32, round to nearest even)
// advances the bit pattern of k by a CONSTANT M = bits(RN(k0 + step)) - bits(k0): k is a multiple of u = 2^(e-23), step = q*u + r with
// 0 <= r < u rounds the same way from every such k - unless r == u/2 exactly, where ties-to-even alternates (the kernel refuses those).
// So n sequential steps land on bits(k0) + n*M exactly. Usage: check_kleap [cases]; prints "... mismatches 0" and returns 0 when it holds.
#include <stdio.h>
#include <stdlib.h>
#include <stdint.h>
#include <string.h>
#include <math.h>
static uint32_t fb(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static float bf(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static uint64_t rs = 88172645463325252ull;
static uint64_t rnd(void) { rs ^= rs << 13; rs ^= rs >> 7; rs ^= rs << 17; return rs; }
int main(int argc, char **argv) {
    long cases = argc > 1 ? atol(argv[1]) : 200000, bad = 0, steps = 0, ties = 0, refused = 0, walked = 0;
    for (long c = 0; c < cases; c++) {
        const int ek = (int)(rnd() % 40) - 20;
        const float k0 = ldexpf(1.0f + (float)(rnd() & 0x7fffff) / 8388608.0f, ek);
        const int es = ek - (int)(rnd() % 30) + 3;
        uint32_t mant = (uint32_t)(rnd() & 0x7fffff);
        if (rnd() % 4 == 0) mant &= ~((1u << (rnd() % 23)) - 1u);   // trailing zeros: exact sums and ties
        const float step = bf(((uint32_t)(es + 127) << 23) | mant);
        if (!(step > 0) || !isfinite(step)) continue;
        volatile float ka = k0 + step;
        const uint32_t kb = fb(k0), kab = fb(ka), eb = kb & 0xff800000u, M = kab - kb;
        if ((kb ^ kab) >= 0x00800000u || (kb >> 23) < 30u || M == 0u) { refused++; continue; }      // the kernel's own preconditions
        volatile float inc = ka - k0, res = step - inc;
        if (fabsf(res) == bf((kb & 0x7f800000u) - (24u << 23))) { ties++; continue; }
        volatile float ks = k0;
        const float top = bf(eb + 0x00800000u);
        long n = 0;
        walked++;
        for (;;) {
            volatile float nx = ks + step;
            if (!(nx < top)) break;
            ks = nx; n++;
            if (fb(ks) != kb + (uint32_t)n * M) { bad++; if (bad < 10) printf("MISMATCH k0=%a step=%a n=%ld seq=%a\n", k0, step, n, ks); break; }
            if (n > 100000) break;
        }
        steps += n;
    }
    printf("cases %ld, walked %ld, steps compared %ld, ties refused %ld, other refusals %ld, mismatches %ld\n", cases, walked, steps, ties, refused, bad);
    return bad != 0;
}
