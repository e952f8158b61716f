/* tools/check_div255.c - the trilinear kernel (vr5_tri.cuh, div255) replaces x / 255.0f by
 *     q0 = x * r;  q = fma(fma(-q0, 255, x), r, q0)      with r = RN(1/255) = 0x1.010102p-8
 * This program checks q == x / 255.0f bit for bit over EVERY float in [0, 255] (1 132 396 545 bit patterns, ~40 s), or over
 * a strided sample when given a stride argument (tests/test_trilinear.py uses 997).
 *     gcc -O2 -mfma -ffp-contract=off tools/check_div255.c -o /tmp/check_div255 -lm && /tmp/check_div255 [stride]
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

int main(int argc, char **argv) {
    const uint32_t stride = argc > 1 ? (uint32_t)strtoul(argv[1], NULL, 10) : 1u;
    const float r = 0x1.010102p-8f, top_f = 255.0f;
    uint32_t top;
    memcpy(&top, &top_f, 4);
    unsigned long long bad = 0, seen = 0;
    for (uint64_t b = 0; b <= top; b += stride ? stride : 1u) {
        const uint32_t bits = (uint32_t)b;
        float x;
        memcpy(&x, &bits, 4);
        const float want = x / 255.0f;
        const float q0 = x * r;
        const float q = fmaf(fmaf(-q0, 255.0f, x), r, q0);
        if (memcmp(&q, &want, 4)) {
            if (!bad) printf("first mismatch at bits %08x\n", bits);
            bad++;
        }
        seen++;
    }
    printf("checked %llu values, %llu mismatches\n", seen, bad);
    return bad ? 1 : 0;
}
