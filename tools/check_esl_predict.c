/* Host check of the predicted empty-space leap of vr5_cast.cuh (skip_leading_empty_space, predicted path).
 *
 * Reference (RaycasterBase.h:67-85): m = floor( RN( max(min3(RN(nx/dx), RN(ny/dy), RN(nz/dz)), 0) / step ) ).
 * Prediction: x~ = max(min3(RN(nx*yx), ...), 0) * ys with y = RN(1/d), mf = floor(x~), ACCEPTED only when
 * fma(x~, 2^-19, |(x~ - mf) - 0.5|) <= 0.5. The claim the kernel relies on: an accepted prediction equals the reference's m.
 * This program hammers the claim with random operands and with operands constructed so that dk / step lands within a few ulps
 * of an integer (where a sloppy acceptance test would fail), and reports the acceptance rate.
 *
 *   gcc -O2 -ffp-contract=off tools/check_esl_predict.c -o /tmp/check_esl -lm && /tmp/check_esl [millions of cases]
 */
#include <fenv.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

static uint64_t rng = 0x9E3779B97F4A7C15ull;
static uint64_t next(void) { rng ^= rng << 13; rng ^= rng >> 7; rng ^= rng << 17; return rng; }
static float uni(float lo, float hi) { return lo + (hi - lo) * (float)((next() >> 40) * (1.0 / 16777216.0)); }
static float sel_min(float a, float b) { return a < b ? a : b; }
static float sel_max(float a, float b) { return a > b ? a : b; }

static float add_rz(float a, float b) {  /* __fadd_rz */
    fesetround(FE_TOWARDZERO);
    volatile float r = a + b;
    fesetround(FE_TONEAREST);
    return r;
}

static long checked = 0, accepted = 0, wrong = 0;

static void one(float nx, float ny, float nz, float dx, float dy, float dz, float step) {
    /* reference */
    float kx = nx / dx, ky = ny / dy, kz = nz / dz;
    if (dx == 0.0f) kx = 100.0f;
    if (dy == 0.0f) ky = 100.0f;
    if (dz == 0.0f) kz = 100.0f;
    const float m_ref = floorf(sel_max(sel_min(sel_min(kx, ky), kz), 0.0f) / step);
    /* prediction */
    const float yx = 1.0f / dx, yy = 1.0f / dy, yz = 1.0f / dz, ys = 1.0f / step;
    const float qx = dx == 0.0f ? 100.0f : nx * yx, qy = dy == 0.0f ? 100.0f : ny * yy, qz = dz == 0.0f ? 100.0f : nz * yz;
    const float x = sel_max(sel_min(sel_min(qx, qy), qz), 0.0f) * ys;
    const float mf = add_rz(x, 8388608.0f) - 8388608.0f;
    const int ok = fmaf(x, 0x1p-19f, fabsf((x - mf) - 0.5f)) <= 0.5f;
    checked++;
    if (ok) {
        accepted++;
        if (mf != m_ref) {
            if (wrong++ < 10) printf("MISMATCH n=(%a %a %a) d=(%a %a %a) step=%a: predicted %g, reference %g\n", nx, ny, nz, dx, dy, dz, step, mf, m_ref);
        }
    }
}

int main(int argc, char **argv) {
    const long millions = argc > 1 ? atol(argv[1]) : 20;
    for (long i = 0; i < millions * 1000000L; i++) {
        const int mode = (int)(next() % 8);
        float step = ldexpf(uni(1.0f, 2.0f), -(int)(next() % 12) - 1);           /* 2^-13 .. 1 */
        float dx = uni(-1.5f, 1.5f), dy = uni(-1.5f, 1.5f), dz = uni(-1.5f, 1.5f);
        if (next() % 16 == 0) dx = 0.0f;
        if (next() % 16 == 0) dy = ldexpf(dy, -(int)(next() % 50));              /* tiny components */
        float nx = uni(-2.0f, 2.0f), ny = uni(-2.0f, 2.0f), nz = uni(-2.0f, 2.0f);
        if (next() % 32 == 0) nz = 0.0f;
        if (mode >= 4) {
            /* adversarial: make the minimum quotient sit within a few ulps of an integer multiple of the step */
            const float m = (float)(next() % (mode == 7 ? 4000 : 70));
            float target = m * step;                                            /* dk ~ m * step */
            const int ulps = (int)(next() % 9) - 4;
            for (int u = 0; u < (ulps < 0 ? -ulps : ulps); u++) target = nextafterf(target, ulps < 0 ? -INFINITY : INFINITY);
            nx = target * dx;                                                   /* RN(nx / dx) lands at or next to the target */
            ny = nx / dx * dy * uni(1.5f, 3.0f) + (dy > 0 ? 1.0f : -1.0f);       /* the other two stay larger */
            nz = nx / dx * dz * uni(1.5f, 3.0f) + (dz > 0 ? 1.0f : -1.0f);
            if (dx == 0.0f) { nx = 0.0f; }
        }
        one(nx, ny, nz, dx, dy, dz, step);
    }
    printf("%ld cases, %ld accepted (%.3f %%), %ld mismatches among the accepted\n", checked, accepted, 100.0 * accepted / checked, wrong);
    return wrong ? 1 : 0;
}
