// tools/check_leap_geometry.c - host replay of the leap rule of the zero-skip kernels (vr5_cast.cuh, try_leap) against the reference's
// own sample arithmetic. For random and adversarial rays (almost parallel to brick faces, grazing edges and corners, far origins, big
// grids, tiny and huge direction components) it takes a leap from a random sample of the march exactly as the kernel does - brick
// table look-up, shrunk exit times, brick-to-brick walk over clean crossings, floor(t / inc), binade cap - with a random subset of the
// bricks marked "not transparent", and then checks the two things the kernel's claim rests on:
//   (1) every sample the leap leaves out reads, in the REFERENCE's arithmetic (ModelBase.h:17-23 on pt = origin + direction * k,
//       CPURenderer.cpp:37-38), a voxel of one of the bricks the walk accepted;
//   (2) the ray parameter after the leap has the bit pattern the reference's sequential k += step arrives at.
// The reciprocals are perturbed by a few ulps (the kernel uses rcp.approx). Usage: check_leap_geometry [cases]; prints "... violations 0".
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static uint32_t fb(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static float bf(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static uint64_t rs = 0x9E3779B97F4A7C15ull;
static uint64_t rnd(void) { rs ^= rs << 13; rs ^= rs >> 7; rs ^= rs << 17; return rs; }
static float urand(float a, float b) { return a + (b - a) * (float)((rnd() >> 11) * (1.0 / 9007199254740992.0)); }
static float rcp_pert(float x) {   // a reciprocal that is a few ulps off, like rcp.approx (and inf for 0 / denormals: .ftz)
    if (fabsf(x) < 1.17549435e-38f) return INFINITY;
    float r = 1.0f / x;
    int d = (int)(rnd() % 5) - 2;
    return bf(fb(r) + (uint32_t)d);
}
// the reference's voxel index of a point coordinate: map_float_int((p + 1) * 0.5f, n) (common.h:105-110; vr5_cast.cuh shows the product form is the same float)
static uint32_t ref_voxel(float p, uint32_t n) {
    volatile float a = p + 1.0f;
    volatile float v = a * (0.5f * (float)n);
    long i = (long)v;
    if (i < 0) i = 0;
    if (i > (long)n - 1) i = (long)n - 1;
    return (uint32_t)i;
}

#define MAXB 64
typedef struct { uint32_t x, y, z; } B3;

int main(int argc, char **argv) {
    long cases = argc > 1 ? atol(argv[1]) : 200000;
    const float margin_scale = getenv("CLG_MARGIN_SCALE") ? (float)atof(getenv("CLG_MARGIN_SCALE")) : 1.0f;
    long leaps = 0, skipped = 0, hops_total = 0, viol_vox = 0, viol_k = 0, refused = 0;
    static const uint32_t dims_list[] = {64, 100, 256, 512, 1000, 1024, 2048, 4096};
    for (long c = 0; c < cases; c++) {
        const uint32_t n = dims_list[rnd() % 8];
        uint32_t lb = 3 + (uint32_t)(rnd() % 4);
        while ((n >> lb) > 64) lb++;
        const uint32_t nb = (n + (1u << lb) - 1) >> lb;
        const float h = 0.5f * (float)n, B = (float)(1u << lb);
        // ray: origin on a shell, direction towards a random point of the cube; adversarial modes flatten components
        float o[3], d[3];
        const float dist = (float[]){1.05f, 1.3f, 2.0f, 3.0f, 5.0f, 9.0f, 40.0f}[rnd() % 7];
        float dir[3] = {urand(-1, 1), urand(-1, 1), urand(-1, 1)};
        float len = sqrtf(dir[0] * dir[0] + dir[1] * dir[1] + dir[2] * dir[2]) + 1e-9f;
        float tgt[3] = {urand(-0.9f, 0.9f), urand(-0.9f, 0.9f), urand(-0.9f, 0.9f)};
        const int mode = (int)(rnd() % 8);
        for (int i = 0; i < 3; i++) { dir[i] /= len; }
        if (mode == 1) { dir[rnd() % 3] *= 1e-4f; }                       // almost parallel to one family of faces
        if (mode == 2) { int a = (int)(rnd() % 3); dir[a] *= 1e-7f; dir[(a + 1) % 3] *= 1e-3f; }
        if (mode == 3) { dir[rnd() % 3] = 0.0f; }                         // exactly parallel
        if (mode == 4) {                                                  // target ON a brick face / edge / corner (voxel-space multiples of B)
            for (int i = 0; i < 3; i++) if (rnd() % 2) tgt[i] = ((float)((rnd() % nb) << lb) + urand(-0.3f, 0.3f)) / h - 1.0f;
        }
        const float dscale = (mode == 5) ? urand(0.001f, 0.01f) : (mode == 6) ? urand(50.0f, 500.0f) : 1.0f;   // un-normalised directions (perspective)
        for (int i = 0; i < 3; i++) { o[i] = tgt[i] - dir[i] * dist; d[i] = dir[i] * dscale; }
        // clip to [-1, 1]^3 (Raycaster::intersect, RaycasterBase.h:32-42, in float)
        float kn = 0.0f, kf = 3.4e38f;
        for (int i = 0; i < 3; i++) {
            const float e = d[i] == 0.0f ? 0.00001f : d[i];
            const float k1 = (-1.0f - o[i]) / e, k2 = (1.0f - o[i]) / e;
            kn = fmaxf(kn, fminf(k1, k2)); kf = fminf(kf, fmaxf(k1, k2));
        }
        if (!(kn < kf) || !(kf > 0.0f)) { refused++; continue; }
        float step = (2.0f / (float)n - 2.0f / (float)n / (float)n) * urand(0.3f, 2.5f) / dscale;
        if (rnd() % 6 == 0) step = bf(fb(step) & ~((1u << (rnd() % 20)) - 1u));   // few mantissa bits: exact sums, ties
        if (!(step > 0.0f)) { refused++; continue; }
        // march to a random sample of the ray
        volatile float k = kn;
        long j0 = (long)(rnd() % 2000), jmax = (long)((kf - kn) / step);
        if (jmax < 4) { refused++; continue; }
        j0 %= jmax;
        for (long j = 0; j < j0; j++) k = k + step;
        const float kc = k;
        volatile float ka_v = kc + step;
        const float ka = ka_v;
        // ---- try_leap, as in vr5_cast.cuh -------------------------------------------------------------------------------------
        const uint32_t kb = fb(kc), kab = fb(ka), M = kab - kb;
        if ((kb ^ kab) >= 0x00800000u || (kb >> 23) < 30u || M == 0u) { refused++; continue; }
        volatile float inc_v = ka - kc; const float inc = inc_v;
        volatile float res_v = step - inc;
        if (fabsf(res_v) == bf((kb & 0x7f800000u) - (24u << 23))) { refused++; continue; }
        float v[3], s[3], r[3], t[3], lo[3];
        uint32_t b[3];
        float ob = fmaxf(fabsf(o[0]), fmaxf(fabsf(o[1]), fabsf(o[2])));
        float mg = 4.0f * (3.0f * 5.9604645e-8f * (ob + 2.0f) * h + 5.9604645e-8f * 2.0f * h) + 1.0f / 32.0f;   // fill_args
        if (!(mg <= 0.25f)) { refused++; continue; }
        mg *= margin_scale;   // (CLG_MARGIN_SCALE=0: the check must then FIND violations - a test of this tool)
        float gap_min = 1.0f;
        for (int i = 0; i < 3; i++) {
            volatile float p = o[i] + d[i] * kc; volatile float p1 = p + 1.0f; volatile float vv = p1 * h;
            v[i] = vv;
            uint32_t bi = (vv >= 0.0f) ? (uint32_t)vv >> lb : 0u;   // (conversion saturates at 0 on the device)
            if (bi > nb - 1) bi = nb - 1;
            b[i] = bi; lo[i] = (float)(bi << lb);
            s[i] = d[i] * h; r[i] = rcp_pert(fabsf(s[i]));
            const float hi_gap = (lo[i] + B - mg) - v[i], lo_gap = v[i] - (lo[i] + mg);
            gap_min = fminf(gap_min, fminf(hi_gap, lo_gap));
            t[i] = s[i] != 0.0f ? (s[i] > 0.0f ? hi_gap : lo_gap) * r[i] : INFINITY;
        }
        if (!(gap_min >= 0.0f)) { refused++; continue; }
        // random "not transparent" bricks: a hash of the brick coordinates
        const uint32_t salt = (uint32_t)rnd(), dens = (uint32_t)(rnd() % 4);   // 0: none ... 3: about a third
        #define OPAQUE(x, y, z) (dens && ((((x) * 73856093u) ^ ((y) * 19349663u) ^ ((z) * 83492791u) ^ salt) % 9u) < dens)
        if (OPAQUE(b[0], b[1], b[2])) { refused++; continue; }
        B3 acc[MAXB]; int nacc = 0;
        acc[nacc++] = (B3){b[0], b[1], b[2]};
        uint32_t ib[3] = {b[0], b[1], b[2]};
        const float t_max = kf - kc;
        float t1 = 0.0f;
        for (int hop = 0; hop < 48; hop++) {
            t1 = fminf(t[0], fminf(t[1], t[2]));
            if (!(t1 < t_max)) break;
            int a = (t[0] <= t[1] && t[0] <= t[2]) ? 0 : (t[1] <= t[2]) ? 1 : 2;
            const float others = fminf(t[(a + 1) % 3], t[(a + 2) % 3]);
            if (!(t1 + 2.0f * mg * r[a] <= others)) break;
            ib[a] += s[a] > 0.0f ? 1u : 0xffffffffu;
            if (ib[a] >= nb || OPAQUE(ib[0], ib[1], ib[2])) break;
            t[a] += B * r[a];
            acc[nacc++] = (B3){ib[0], ib[1], ib[2]};
            hops_total++;
        }
        if (!(t1 >= 0.0f)) { refused++; continue; }
        t1 = fminf(t1, t_max);
        float nf = floorf(t1 * rcp_pert(inc) * 0.99999f);
        const float nbin = floorf((float)(0x007fffffu - (kb & 0x007fffffu)) * rcp_pert((float)M) * 0.99999f) - 3.0f;
        nf = fminf(nf, nbin);
        if (!(nf >= 2.0f)) { refused++; continue; }
        const long nskip = (long)nf;          // samples 1 .. n are left out, the march goes on with sample n + 1
        const uint32_t knb = kb + (uint32_t)(nskip + 1) * M;
        if ((knb ^ kb) >= 0x00800000u) { refused++; continue; }
        // ---- the reference's march over the same samples ---------------------------------------------------------------------
        volatile float kr = kc;
        for (long j = 1; j <= nskip; j++) {
            kr = kr + step;
            uint32_t bx[3];
            for (int i = 0; i < 3; i++) { volatile float pm = d[i] * kr; volatile float p = o[i] + pm; bx[i] = ref_voxel(p, n) >> lb; }
            int ok = 0;
            for (int q = 0; q < nacc; q++) if (acc[q].x == bx[0] && acc[q].y == bx[1] && acc[q].z == bx[2]) { ok = 1; break; }
            if (!ok) {
                viol_vox++;
                if (viol_vox < 10) printf("VOXEL VIOLATION n=%u lb=%u mode=%d j=%ld/%ld brick (%u %u %u) start (%u %u %u) o=(%a %a %a) d=(%a %a %a) kc=%a step=%a mg=%g\n",
                                          n, lb, mode, j, nskip, bx[0], bx[1], bx[2], b[0], b[1], b[2], o[0], o[1], o[2], d[0], d[1], d[2], kc, step, mg);
                break;
            }
        }
        kr = kr + step;
        if (fb(kr) != knb && viol_vox == 0) { viol_k++; if (viol_k < 10) printf("K VIOLATION kc=%a step=%a n=%ld seq=%a leap=%a\n", kc, step, nskip, kr, bf(knb)); }
        leaps++; skipped += nskip;
    }
    printf("cases %ld: leaps %ld, samples left out %ld, hops %ld, refused %ld, violations %ld (voxel %ld, k %ld)\n", cases, leaps, skipped, hops_total, refused,
           viol_vox + viol_k, viol_vox, viol_k);
    return (viol_vox + viol_k) != 0;
}
