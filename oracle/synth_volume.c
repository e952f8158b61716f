/*
 * synth_volume.c - host generator of the deterministic synthetic volumes. TEST INFRASTRUCTURE ONLY.
 *
 * SURVEY section 8(d) names two synthetic inputs; both are defined here in INTEGER arithmetic only, so that
 * this host generator and the product's device generator (csrc/vr5_synth.cu, written independently)
 * produce identical bytes (checked by tests/test_gpu_synth.py). The reference has no synthetic volumes;
 * these stand in for the data files its project file names but does not ship (VolumeRendering.vcproj:225-247).
 *
 *  synthA(N, seed): dense band-limited Marschner-Lobb-like field, values span 0..255, no empty ESL block:
 *      c{x,y,z} = ((2i+1) * 32768) / N - 32768                     voxel centre, Q15 in [-1, 1)
 *      r        = isqrt(cx^2 + cy^2)                               Q15 radius
 *      c1       = COS[(r >> 7) & 1023]                             cos(pi r / 2), Q15
 *      c2       = COS[((fM * c1) >> 5) & 1023], fM = 5 + seed      cos(2 pi fM c1), Q15
 *      s        = COS[((cz >> 7) - 256) & 1023]                    sin(pi z / 2), Q15
 *      num      = 32767 - s + ((32767 + c2) >> 2)                  (1 - s + (1 + c2)/4), Q15
 *      value    = min(255, (num * 255 + 40958) / 81917)            / 2.5, scaled to 0..255
 *    COS[i] = lround(32767 cos(2 pi i / 1024)) is computed once on the host and handed to the device.
 *
 *  synthB(N, seed): sparse - 64 solid spheres (value 200..255) with a thin shell (value 100) on a 0 background;
 *    sphere parameters from a 32-bit LCG (x <- 1664525 x + 1013904223, top 16 bits used):
 *      centre_a = N*15/100 + ((rnd * (N*70/100)) >> 16)   for a = x, y, z
 *      radius   = (N * (25 + ((rnd * 75) >> 16))) / 1000
 *      value    = 200 + ((rnd * 56) >> 16)
 *    shell thickness max(1, N/128) voxels; a voxel takes the maximum over all spheres.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

void orc_synth_cos_table(int16_t *table1024) {
    for (int i = 0; i < 1024; i++) table1024[i] = (int16_t)lround(32767.0 * cos(2.0 * 3.14159265358979323846 * i / 1024.0));
}

static inline uint32_t isqrt_u32(uint32_t v) {
    uint64_t r = (uint64_t)sqrtf((float)v);
    while (r * r > v) r--;
    while ((r + 1) * (r + 1) <= v) r++;
    return (uint32_t)r;
}

/* fills slices z in [z0, z1) of the n^3 volume `out`; re-entrant, so callers may split z over threads */
void orc_synth_a(uint8_t *out, uint32_t n, uint32_t seed, uint32_t z0, uint32_t z1) {
    int16_t cs[1024];
    orc_synth_cos_table(cs);
    const int32_t fm = 5 + (int32_t)seed;
    int32_t *centre = (int32_t *)malloc(n * sizeof(int32_t));
    for (uint32_t i = 0; i < n; i++) centre[i] = (int32_t)(((2 * i + 1) * 32768u) / n) - 32768;
    if (z1 > n) z1 = n;
    for (uint32_t z = z0; z < z1; z++) {
        int32_t cz = centre[z];
        int32_t s = cs[((cz >> 7) - 256) & 1023];
        for (uint32_t y = 0; y < n; y++) {
            int32_t cy = centre[y];
            uint8_t *row = out + ((uint64_t)z * n + y) * n;
            for (uint32_t x = 0; x < n; x++) {
                int32_t cx = centre[x];
                uint32_t r = isqrt_u32((uint32_t)(cx * cx) + (uint32_t)(cy * cy));
                int32_t c1 = cs[(r >> 7) & 1023];
                int32_t c2 = cs[((fm * c1) >> 5) & 1023];
                int32_t num = 32767 - s + ((32767 + c2) >> 2);
                int32_t v = (num * 255 + 40958) / 81917;
                row[x] = (uint8_t)(v > 255 ? 255 : (v < 0 ? 0 : v));
            }
        }
    }
    free(centre);
}

typedef struct orc_sphere { int32_t cx, cy, cz, r, value; } orc_sphere;

void orc_synth_b_spheres(orc_sphere *spheres64, uint32_t n, uint32_t seed) {
    uint32_t state = seed;
#define NEXT16() (state = state * 1664525u + 1013904223u, (state >> 16))
    for (int i = 0; i < 64; i++) {
        uint32_t span = n * 70 / 100, base = n * 15 / 100;
        spheres64[i].cx = (int32_t)(base + ((NEXT16() * span) >> 16));
        spheres64[i].cy = (int32_t)(base + ((NEXT16() * span) >> 16));
        spheres64[i].cz = (int32_t)(base + ((NEXT16() * span) >> 16));
        spheres64[i].r = (int32_t)((n * (25 + ((NEXT16() * 75) >> 16))) / 1000);
        spheres64[i].value = (int32_t)(200 + ((NEXT16() * 56) >> 16));
    }
#undef NEXT16
}

void orc_synth_b(uint8_t *out, uint32_t n, uint32_t seed) {
    orc_sphere sp[64];
    orc_synth_b_spheres(sp, n, seed);
    memset(out, 0, (size_t)n * n * n);
    int32_t shell = (int32_t)(n / 128 > 1 ? n / 128 : 1);
    for (int i = 0; i < 64; i++) {
        int32_t ro = sp[i].r + shell;
        int64_t r2 = (int64_t)sp[i].r * sp[i].r, ro2 = (int64_t)ro * ro;
        int32_t z0 = sp[i].cz - ro < 0 ? 0 : sp[i].cz - ro, z1 = sp[i].cz + ro >= (int32_t)n ? (int32_t)n - 1 : sp[i].cz + ro;
        int32_t y0 = sp[i].cy - ro < 0 ? 0 : sp[i].cy - ro, y1 = sp[i].cy + ro >= (int32_t)n ? (int32_t)n - 1 : sp[i].cy + ro;
        int32_t x0 = sp[i].cx - ro < 0 ? 0 : sp[i].cx - ro, x1 = sp[i].cx + ro >= (int32_t)n ? (int32_t)n - 1 : sp[i].cx + ro;
        for (int32_t z = z0; z <= z1; z++)
            for (int32_t y = y0; y <= y1; y++) {
                uint8_t *row = out + ((uint64_t)z * n + y) * n;
                int64_t dyz = (int64_t)(z - sp[i].cz) * (z - sp[i].cz) + (int64_t)(y - sp[i].cy) * (y - sp[i].cy);
                for (int32_t x = x0; x <= x1; x++) {
                    int64_t d2 = dyz + (int64_t)(x - sp[i].cx) * (x - sp[i].cx);
                    int32_t v = d2 <= r2 ? sp[i].value : (d2 <= ro2 ? 100 : 0);
                    if (v > row[x]) row[x] = (uint8_t)v;
                }
            }
    }
}
