/*
 * oracle_port.c - CPU restatement of the reference ray caster. TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this.
 * The product path (libvr5.so) never links or calls it and has no CPU fallback.
 *
 * Parity status: PINNED. This restatement is checked bit-for-bit against the reference's own CPURenderer
 * (compiled unmodified into oracle/_ref/libvolr_ref.so by oracle/Makefile) on Bucky.pvm and on synthetic
 * volumes by tests/test_oracle_pin.py (where /root/reference exists), and against the committed golden
 * fixtures those runs produced (tests/golden/, made by tests/golden/make_golden.py).
 *
 * Build rules: -O2 -ffp-contract=off, never -march=native / -mfma / -ffast-math. x86-64 SSE
 * fp32 is then strict IEEE-754 single precision with one rounding per operation, which is what every
 * expression below relies on. Each fp32 operation is written as its own statement so that the evaluation
 * order of the reference expression it restates is explicit.
 *
 * Documented deviation: voxel indices are computed in 64 bit. The reference uses 32-bit unsigned
 * arithmetic (ModelBase.h:18-22), identical for every volume below 2^32 voxels.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/vr5_pod.h"

typedef struct orc_counters {
    uint64_t rays_hit;   /* rays for which intersect() returned true */
    uint64_t rays_accum; /* rays that entered the accumulation loop (and therefore wrote a pixel) */
    uint64_t esl_iters;  /* iterations of the empty-space loop that leapt */
    uint64_t samples;    /* iterations of the accumulation loop = classified voxel fetches (the bench unit) */
    uint64_t shaded;     /* samples that took the shading path (second voxel fetch) */
} orc_counters;

/* common.h:93-101 - ternary semantics (NaN and signed-zero behaviour differ from fminf/fmaxf) */
static inline float sel_min(float a, float b) { return a < b ? a : b; }
static inline float sel_max(float a, float b) { return a > b ? a : b; }

/* common.h:105-110 map_float_int: truncate f*n toward zero, clamp to [0, n-1] */
static inline uint32_t unit_to_index(float f, uint32_t n) {
    float scaled = f * (float)n;
    long i = (long)scaled;
    if (i >= (long)(int)n) i = (long)n - 1;
    if (i < 0) i = 0;
    return (uint32_t)i;
}

/* the (p + 1) * 0.5f world -> unit-cube mapping used by ModelBase.h:19-21 and RaycasterBase.h:57-62, 68-72 */
static inline uint32_t world_to_voxel(float p, uint32_t n) {
    float shifted = p + 1.0f;
    float unit = shifted * 0.5f;
    return unit_to_index(unit, n);
}

/* ModelBase.h:17-23 nearest-neighbour fetch (64-bit index: see header) */
static inline uint8_t fetch_voxel(const vr5_model *m, float px, float py, float pz) {
    uint64_t zi = world_to_voxel(pz, m->dims.z);
    uint64_t yi = world_to_voxel(py, m->dims.y);
    uint64_t xi = world_to_voxel(px, m->dims.x);
    return m->data[zi * m->dims.x * m->dims.y + yi * m->dims.x + xi];
}

/* ViewBase.h:23-35 get_ray */
static inline void primary_ray(const vr5_view *v, int col, int row, vr5_f3 *o, vr5_f3 *d) {
    float cx = (float)(col - (int)v->dims.x / 2);
    float cy = (float)(row - (int)v->dims.y / 2);
    float rx = v->right_plane.x * cx, ry = v->right_plane.y * cx, rz = v->right_plane.z * cx;
    float ux = v->up_plane.x * cy, uy = v->up_plane.y * cy, uz = v->up_plane.z * cy;
    if (v->perspective) {
        *o = v->origin;
        float tx = v->direction.x + rx, ty = v->direction.y + ry, tz = v->direction.z + rz;
        d->x = tx + ux; d->y = ty + uy; d->z = tz + uz;
    } else {
        *d = v->direction;
        float tx = v->origin.x + rx, ty = v->origin.y + ry, tz = v->origin.z + rz;
        o->x = tx + ux; o->y = ty + uy; o->z = tz + uz;
    }
}

/* RaycasterBase.h:32-42 intersect: slab test against [min_bound, -min_bound]; zero components replaced on a copy */
static inline int clip_ray(const vr5_model *m, vr5_f3 o, vr5_f3 d, float *k_near, float *k_far) {
    if (d.x == 0) d.x = 0.00001f;
    if (d.y == 0) d.y = 0.00001f;
    if (d.z == 0) d.z = 0.00001f;
    float lo_x = m->min_bound.x - o.x, lo_y = m->min_bound.y - o.y, lo_z = m->min_bound.z - o.z;
    float hi_x = -m->min_bound.x - o.x, hi_y = -m->min_bound.y - o.y, hi_z = -m->min_bound.z - o.z;
    float k1x = lo_x / d.x, k1y = lo_y / d.y, k1z = lo_z / d.z;
    float k2x = hi_x / d.x, k2y = hi_y / d.y, k2z = hi_z / d.z;
    float near = sel_max(sel_max(sel_min(k1x, k2x), sel_min(k1y, k2y)), sel_min(k1z, k2z));
    float far = sel_min(sel_min(sel_max(k1x, k2x), sel_max(k1y, k2y)), sel_max(k1z, k2z));
    near = sel_max(near, 0);
    *k_near = near;
    *k_far = far;
    return (near < far) && (far > 0);
}

/* RaycasterBase.h:52-65 sample_data_esl: word = (zblk*32 + yblk), bit = xblk; set bit = empty block */
static inline int block_is_empty(const vr5_raycaster *r, float px, float py, float pz) {
    uint32_t bd = r->esl_block_dims;
    uint16_t word = (uint16_t)((world_to_voxel(pz, r->volume.dims.z) / bd) * VR5_ESL_VOLUME_DIMS +
                               (world_to_voxel(py, r->volume.dims.y) / bd));
    uint32_t bits = r->esl_volume[word];
    uint16_t bit = (uint16_t)(world_to_voxel(px, r->volume.dims.x) / bd);
    return (bits & (1u << bit)) != 0;
}

/* RaycasterBase.h:67-85 leap_empty_space: distance to the nearest exit face of the current block, in whole steps */
static inline float leap_distance(const vr5_raycaster *r, float px, float py, float pz, vr5_f3 d) {
    uint32_t bd = r->esl_block_dims;
    uint16_t bx = (uint16_t)(world_to_voxel(px, r->volume.dims.x) / bd);
    uint16_t by = (uint16_t)(world_to_voxel(py, r->volume.dims.y) / bd);
    uint16_t bz = (uint16_t)(world_to_voxel(pz, r->volume.dims.z) / bd);
    if (d.x > 0) bx++;
    if (d.y > 0) by++;
    if (d.z > 0) bz++;
    float fx = r->esl_block_size.x * (float)bx, fy = r->esl_block_size.y * (float)by, fz = r->esl_block_size.z * (float)bz;
    float wx = r->volume.min_bound.x + fx, wy = r->volume.min_bound.y + fy, wz = r->volume.min_bound.z + fz;
    float ex = wx - px, ey = wy - py, ez = wz - pz;
    float kx = ex / d.x, ky = ey / d.y, kz = ez / d.z;
    if (d.x == 0) kx = 100;
    if (d.y == 0) ky = 100;
    if (d.z == 0) kz = 100;
    float dk = sel_min(kx, ky);
    dk = sel_min(dk, kz);
    dk = sel_max(dk, 0);
    float q = dk / r->ray_step;
    dk = floorf(q) * r->ray_step;
    return dk;
}

/* RaycasterBase.h:87-98 shade (the caller has already applied the alpha / light_kd guard of CPURenderer.cpp:32) */
static inline void shade_sample(const vr5_raycaster *r, vr5_f4 *c, float px, float py, float pz, uint8_t sample) {
    float lx = r->view.light_pos.x - px, ly = r->view.light_pos.y - py, lz = r->view.light_pos.z - pz;
    float xx = lx * lx, yy = ly * ly, zz = lz * lz;
    float s = xx + yy;
    s = s + zz;
    float inv = 1.0f / sqrtf(s);
    float nx = lx * inv, ny = ly * inv, nz = lz * inv;
    float sample_l;
    if (c->w > 0.05f && r->light_kd > 0.01f) {
        float ox = nx * 0.01f, oy = ny * 0.01f, oz = nz * 0.01f;
        float qx = px + ox, qy = py + oy, qz = pz + oz;
        sample_l = (float)fetch_voxel(&r->volume, qx, qy, qz) / 255.0f;
    } else {
        sample_l = (float)sample / 255.0f;
    }
    float here = (float)sample / 255.0f;
    float diff = sample_l - here;
    diff = diff * r->light_kd;
    c->x += diff;
    c->y += diff;
    c->z += diff;
}

/* CPURenderer.cpp:11-41 render_ray + RaycasterBase.h:44-50 write_color */
static void cast_pixel(const vr5_raycaster *r, uint8_t *rgba8, int col, int row, orc_counters *cnt) {
    vr5_f3 o, d;
    float k, k_far;
    primary_ray(&r->view, col, row, &o, &d);
    if (!clip_ray(&r->volume, o, d, &k, &k_far)) return;
    cnt->rays_hit++;
    float px, py, pz;
#define ADVANCE_POINT()                                   \
    do {                                                  \
        float mx = d.x * k, my = d.y * k, mz = d.z * k;   \
        px = o.x + mx; py = o.y + my; pz = o.z + mz;      \
    } while (0)
    ADVANCE_POINT();
    while (k <= k_far) { /* CPURenderer.cpp:18-25 leading empty-space loop */
        if (r->esl && block_is_empty(r, px, py, pz)) {
            k += leap_distance(r, px, py, pz, d);
            cnt->esl_iters++;
        } else {
            break;
        }
        k += r->ray_step;
        ADVANCE_POINT();
    }
    if (k > k_far) return;
    cnt->rays_accum++;
    vr5_f4 acc = {0, 0, 0, 0};
    while (k <= k_far) { /* CPURenderer.cpp:29-39 accumulation loop */
        uint8_t sample = fetch_voxel(&r->volume, px, py, pz);
        vr5_f4 cur = r->transfer_fn[sample / VR5_TF_RATIO];
        cnt->samples++;
        if (cur.w > 0.05f && r->light_kd > 0.01f) {
            shade_sample(r, &cur, px, py, pz, sample);
            cnt->shaded++;
        }
        float t = 1 - acc.w;
        float cx = cur.x * t, cy = cur.y * t, cz = cur.z * t, cw = cur.w * t;
        acc.x = acc.x + cx; acc.y = acc.y + cy; acc.z = acc.z + cz; acc.w = acc.w + cw;
        if (acc.w > r->ray_threshold) break;
        k += r->ray_step;
        ADVANCE_POINT();
    }
#undef ADVANCE_POINT
    uint8_t *px8 = rgba8 + 4 * ((size_t)row * r->view.dims.x + (size_t)col);
    px8[0] = (uint8_t)unit_to_index(acc.x, 256);
    px8[1] = (uint8_t)unit_to_index(acc.y, 256);
    px8[2] = (uint8_t)unit_to_index(acc.z, 256);
    px8[3] = (uint8_t)unit_to_index(acc.w, 256);
}

/*
 * CPURenderer.cpp:43-53 render_volume restricted to rows [row0, row1): clears those rows, casts one ray per
 * pixel in row-major order. Returns 1 on a missing input like the reference, else 0. `cnt` may be NULL.
 * Re-entrant: distinct row ranges may run on distinct threads.
 */
int orc_render_rows(const vr5_raycaster *r, uint8_t *rgba8, int row0, int row1, orc_counters *cnt) {
    if (r->volume.data == NULL || r->transfer_fn == NULL || r->esl_volume == NULL || rgba8 == NULL) return 1;
    orc_counters local = {0, 0, 0, 0, 0};
    int w = r->view.dims.x, h = r->view.dims.y;
    if (row0 < 0) row0 = 0;
    if (row1 > h) row1 = h;
    if (row1 > row0) memset(rgba8 + (size_t)row0 * w * 4, 0, (size_t)(row1 - row0) * w * 4);
    for (int row = row0; row < row1; row++)
        for (int col = 0; col < w; col++) cast_pixel(r, rgba8, col, row, &local);
    if (cnt) *cnt = local;
    return 0;
}

int orc_render(const vr5_raycaster *r, uint8_t *rgba8, orc_counters *cnt) {
    return orc_render_rows(r, rgba8, 0, r->view.dims.y, cnt);
}

/* ---------------------------------------------------------------------------------------------------------
 * Scene builders restated (host side, once per change)
 * ------------------------------------------------------------------------------------------------------- */

/* RaycasterBase.cpp:76-84 reset_transfer_fn: R ramp for i<=42, G for 43..84, B above, alpha i/128 above 12.75 */
void orc_default_base_tf(float *rgba128) {
    const int third = VR5_TF_SIZE / 3;
    for (int i = 0; i < VR5_TF_SIZE; i++) {
        float *e = rgba128 + 4 * i;
        e[0] = i <= third ? (float)(i * 3) / (float)VR5_TF_SIZE : 0.0f;
        e[1] = (i > third && i <= third * 2) ? (float)((i - third) * 3) / (float)VR5_TF_SIZE : 0.0f;
        e[2] = i > third * 2 ? (float)((i - third * 2) * 3) / (float)VR5_TF_SIZE : 0.0f;
        e[3] = (float)i > ((255.0f * 0.1f) / (float)VR5_TF_RATIO) ? (float)i / (float)VR5_TF_SIZE : 0.0f;
    }
}

/* RaycasterBase.cpp:46-52 pre-multiplication */
void orc_premultiply_tf(const float *base128, float *out128) {
    for (int i = 0; i < VR5_TF_SIZE; i++) {
        float a = base128[4 * i + 3];
        out128[4 * i + 0] = base128[4 * i + 0] * a;
        out128[4 * i + 1] = base128[4 * i + 1] * a;
        out128[4 * i + 2] = base128[4 * i + 2] * a;
        out128[4 * i + 3] = a;
    }
}

/* RaycasterBase.cpp:53-67: bit i set <=> no TF entry in [min/2, max/2] has non-zero alpha */
void orc_build_esl_bits(const float *premult128, const uint8_t *minmax /* 32768 x {min,max} */, uint32_t *bits1024) {
    uint16_t next_opaque[VR5_TF_SIZE];
    for (int x = 0; x < VR5_TF_SIZE; x++) {
        int y = x;
        while (y < VR5_TF_SIZE && premult128[4 * y + 3] == 0) y++;
        next_opaque[x] = (uint16_t)y;
    }
    memset(bits1024, 0, VR5_ESL_VOLUME_SIZE * sizeof(uint32_t));
    for (uint32_t i = 0; i < 32768u; i++) {
        uint8_t lo = minmax[2 * i], hi = minmax[2 * i + 1];
        if (next_opaque[lo / VR5_TF_RATIO] > hi / VR5_TF_RATIO) bits1024[i / 32] |= 1u << (i % 32);
    }
}

/* RaycasterBase.cpp:94-125 set_volume: ESL block edge, per-block min/max (64-bit index), block size in world units */
void orc_build_minmax(const uint8_t *data, uint32_t dx, uint32_t dy, uint32_t dz, uint16_t *block_dims_out,
                      uint8_t *minmax /* 32768 x 2 */, float *block_size3) {
    uint32_t max_dim = dx > dy ? (dx > dz ? dx : dz) : (dy > dz ? dy : dz);
    uint32_t bd = (max_dim + VR5_ESL_VOLUME_DIMS - 1) / VR5_ESL_VOLUME_DIMS;
    if (bd < VR5_ESL_MIN_BLOCK_SIZE) bd = VR5_ESL_MIN_BLOCK_SIZE;
    for (uint32_t i = 0; i < 32768u; i++) { minmax[2 * i] = 255; minmax[2 * i + 1] = 0; }
    for (uint32_t z = 0; z < dz; z++)
        for (uint32_t y = 0; y < dy; y++) {
            const uint8_t *rowp = data + ((uint64_t)z * dy + y) * dx;
            uint32_t base = (z / bd) * 1024u + (y / bd) * 32u;
            for (uint32_t x = 0; x < dx; x++) {
                uint8_t s = rowp[x];
                uint8_t *mm = minmax + 2 * (base + x / bd);
                if (mm[0] > s) mm[0] = s;
                if (mm[1] < s) mm[1] = s;
            }
        }
    *block_dims_out = (uint16_t)bd;
    block_size3[0] = 2.0f * (float)bd / (float)dx;
    block_size3[1] = 2.0f * (float)bd / (float)dy;
    block_size3[2] = 2.0f * (float)bd / (float)dz;
}

/* RaycasterBase.cpp:86-92 reset_ray_step */
float orc_default_ray_step(uint32_t max_dim) {
    float step = 2.0f / (float)(int)max_dim;
    step -= step / (float)(int)max_dim;
    return step;
}

/* ModelBase.cpp:19-33 compute_histogram: 4th root of bin counts, normalised to the largest */
void orc_histogram(const uint8_t *data, uint64_t n, float *hist256) {
    uint32_t bins[256];
    memset(bins, 0, sizeof(bins));
    for (uint64_t i = 0; i < n; i++) bins[data[i]]++;
    float top = 0;
    for (int i = 0; i < 256; i++) {
        hist256[i] = sqrtf(sqrtf((float)bins[i])); /* C++ float overloads of sqrt in the reference */
        if (hist256[i] > top) top = hist256[i];
    }
    for (int i = 0; i < 256; i++) hist256[i] = hist256[i] / top;
}

/* ---------------------------------------------------------------------------------------------------------
 * Trilinear sampler policy: a CPU restatement of GPURenderer4's semantics
 * (GPURenderer4.cu:33-97) with the texture unit's arithmetic written out in fp32.
 *
 * Parity status of THIS section: PINNED BY TOLERANCE (round 2). (a) The reference's GPURenderer4.cu, compiled
 * unmodified over a texture-object shim (oracle/ref_gpu4_texobj.cu), renders the same frames in the B200's texture
 * unit; the CUDA policies that are bit-equal to this section are held to it under the tolerance stated in
 * tests/test_trilinear_pin.py (histograms: profiles/r02_ref4_pin.md) - the texture unit filters with 8 fractional
 * weight bits, which no CPU code here reproduces, so equality is not expected. (b) An independent float64
 * restatement of the CUDA filtering formulas (tests/test_trilinear_independent.py) agrees with this section to
 * 0-1 LSB. What is restated is the documented filtering formula evaluated in fp32, one rounding per operation, in
 * the order written below; the CUDA trilinear kernel follows the same order and is compared bit for bit against
 * this function, never against the nearest-neighbour parity bar.
 *
 *   volume:  tex3D(volume_texture, (p+1)*0.5) - normalised coordinates, linear filter, clamp, normalised-float
 *            read mode (GPURenderer4.cu:76, :136-141): texel position t = u*n - 0.5, cell i0 = floor(t),
 *            weight w = t - i0, neighbours clamped to [0, n-1], lerp x then y then z as a + w*(b - a) on the raw
 *            byte values, result / 255.
 *   TF:      tex1D(transfer_fn_texture, sample) - 128 entries, normalised, linear, clamp (GPURenderer4.cu:77,
 *            :102-107): position s*128 - 0.5, lerp of the two entries per channel.
 *   shading: GPURenderer4.cu:41-51: second trilinear sample at p + L*0.01, (sample_l - sample) * kd.
 * ------------------------------------------------------------------------------------------------------- */
static inline float lerp_f(float a, float b, float w) {
    float d = b - a;
    float m = w * d;
    return a + m;
}

static inline void tri_axis(float p, uint32_t n, int64_t *i0, int64_t *i1, float *w) {
    float shifted = p + 1.0f;
    float unit = shifted * 0.5f;
    float scaled = unit * (float)n;
    float t = scaled - 0.5f;
    float fl = floorf(t);
    *w = t - fl;
    int64_t a = (int64_t)fl, b = a + 1;
    if (a < 0) a = 0;
    if (a > (int64_t)n - 1) a = (int64_t)n - 1;
    if (b < 0) b = 0;
    if (b > (int64_t)n - 1) b = (int64_t)n - 1;
    *i0 = a; *i1 = b;
}

static inline float fetch_trilinear(const vr5_model *m, float px, float py, float pz) {
    int64_t x0, x1, y0, y1, z0, z1;
    float wx, wy, wz;
    tri_axis(px, m->dims.x, &x0, &x1, &wx);
    tri_axis(py, m->dims.y, &y0, &y1, &wy);
    tri_axis(pz, m->dims.z, &z0, &z1, &wz);
    const uint64_t sx = m->dims.x, sxy = (uint64_t)m->dims.x * m->dims.y;
#define VOX(X, Y, Z) ((float)m->data[(uint64_t)(Z) * sxy + (uint64_t)(Y) * sx + (uint64_t)(X)])
    float c00 = lerp_f(VOX(x0, y0, z0), VOX(x1, y0, z0), wx);
    float c10 = lerp_f(VOX(x0, y1, z0), VOX(x1, y1, z0), wx);
    float c01 = lerp_f(VOX(x0, y0, z1), VOX(x1, y0, z1), wx);
    float c11 = lerp_f(VOX(x0, y1, z1), VOX(x1, y1, z1), wx);
#undef VOX
    float c0 = lerp_f(c00, c10, wy);
    float c1 = lerp_f(c01, c11, wy);
    float c = lerp_f(c0, c1, wz);
    return c / 255.0f;
}

static inline vr5_f4 classify_linear(const vr5_f4 *tf, float s) {
    float scaled = s * (float)VR5_TF_SIZE;
    float t = scaled - 0.5f;
    float fl = floorf(t);
    float w = t - fl;
    int j0 = (int)fl, j1 = j0 + 1;
    if (j0 < 0) j0 = 0;
    if (j0 > VR5_TF_SIZE - 1) j0 = VR5_TF_SIZE - 1;
    if (j1 < 0) j1 = 0;
    if (j1 > VR5_TF_SIZE - 1) j1 = VR5_TF_SIZE - 1;
    vr5_f4 a = tf[j0], b = tf[j1], c;
    c.x = lerp_f(a.x, b.x, w); c.y = lerp_f(a.y, b.y, w); c.z = lerp_f(a.z, b.z, w); c.w = lerp_f(a.w, b.w, w);
    return c;
}

static void cast_pixel_trilinear(const vr5_raycaster *r, uint8_t *rgba8, int col, int row, orc_counters *cnt) {
    vr5_f3 o, d;
    float k, k_far;
    primary_ray(&r->view, col, row, &o, &d);
    if (!clip_ray(&r->volume, o, d, &k, &k_far)) return;
    cnt->rays_hit++;
    float px, py, pz;
#define ADVANCE_POINT()                                   \
    do {                                                  \
        float mx = d.x * k, my = d.y * k, mz = d.z * k;   \
        px = o.x + mx; py = o.y + my; pz = o.z + mz;      \
    } while (0)
    ADVANCE_POINT();
    while (k <= k_far) { /* GPURenderer4.cu:63-71: same empty-space loop, bitfield read through a texture there */
        if (r->esl && block_is_empty(r, px, py, pz)) {
            k += leap_distance(r, px, py, pz, d);
            cnt->esl_iters++;
        } else {
            break;
        }
        k += r->ray_step;
        ADVANCE_POINT();
    }
    if (k > k_far) return;
    cnt->rays_accum++;
    vr5_f4 acc = {0, 0, 0, 0};
    while (k <= k_far) { /* GPURenderer4.cu:75-86 */
        float sample = fetch_trilinear(&r->volume, px, py, pz);
        vr5_f4 cur = classify_linear(r->transfer_fn, sample);
        cnt->samples++;
        if (cur.w > 0.05f && r->light_kd > 0.01f) { /* GPURenderer4.cu:41-51 shade_texture */
            float lx = r->view.light_pos.x - px, ly = r->view.light_pos.y - py, lz = r->view.light_pos.z - pz;
            float xx = lx * lx, yy = ly * ly, zz = lz * lz;
            float s = xx + yy;
            s = s + zz;
            float inv = 1.0f / sqrtf(s);
            float nx = lx * inv, ny = ly * inv, nz = lz * inv;
            float ox = nx * 0.01f, oy = ny * 0.01f, oz = nz * 0.01f;
            float qx = px + ox, qy = py + oy, qz = pz + oz;
            float sample_l = fetch_trilinear(&r->volume, qx, qy, qz);
            float diff = sample_l - sample;
            diff = diff * r->light_kd;
            cur.x += diff; cur.y += diff; cur.z += diff;
            cnt->shaded++;
        }
        float t = 1 - acc.w;
        float cx = cur.x * t, cy = cur.y * t, cz = cur.z * t, cw = cur.w * t;
        acc.x = acc.x + cx; acc.y = acc.y + cy; acc.z = acc.z + cz; acc.w = acc.w + cw;
        if (acc.w > r->ray_threshold) break;
        k += r->ray_step;
        ADVANCE_POINT();
    }
#undef ADVANCE_POINT
    uint8_t *px8 = rgba8 + 4 * ((size_t)row * r->view.dims.x + (size_t)col);
    px8[0] = (uint8_t)unit_to_index(acc.x, 256);
    px8[1] = (uint8_t)unit_to_index(acc.y, 256);
    px8[2] = (uint8_t)unit_to_index(acc.z, 256);
    px8[3] = (uint8_t)unit_to_index(acc.w, 256);
}

int orc_render_rows_trilinear(const vr5_raycaster *r, uint8_t *rgba8, int row0, int row1, orc_counters *cnt) {
    if (r->volume.data == NULL || r->transfer_fn == NULL || r->esl_volume == NULL || rgba8 == NULL) return 1;
    orc_counters local = {0, 0, 0, 0, 0};
    int w = r->view.dims.x, h = r->view.dims.y;
    if (row0 < 0) row0 = 0;
    if (row1 > h) row1 = h;
    if (row1 > row0) memset(rgba8 + (size_t)row0 * w * 4, 0, (size_t)(row1 - row0) * w * 4);
    for (int row = row0; row < row1; row++)
        for (int col = 0; col < w; col++) cast_pixel_trilinear(r, rgba8, col, row, &local);
    if (cnt) *cnt = local;
    return 0;
}

/* ---------------------------------------------------------------------------------------------------------
 * Trilinear sampler policy, CONTRACTED arithmetic ("trilinear-fma"). Parity status of THIS section: PINNED BY
 * TOLERANCE like the section above (same two anchors: GPURenderer4 on the GPU, the float64 restatement on the CPU).
 *
 * The reference's GPU renderers are built by nvcc with its default -fmad=true, which contracts every
 * `a + b*c` of their march loops (GPURenderer4.cu:75-86: `pt = origin + direction*k`, `acc = acc + cur*(1 -
 * acc.w)`) into one fused multiply-add; only the CPU renderer rounds the product separately. This section
 * restates the same policy with those contractions written out as fmaf() - one rounding per fused operation -
 * and with the filtering formulas in their fused form:
 *
 *   point:    p = fma(d, k, o) per axis (leading empty-space loop, shading); k += step sequentially.
 *   volume:   the ray is carried into texel space ONCE: per axis dv = d * n/2 and ov = fma(o, n/2, n/2 - 0.5)
 *             (texel position of a world point p is fma(p, n/2, n/2 - 0.5) = (p+1)*0.5*n - 0.5 with one rounding;
 *             n/2 and n/2 - 0.5 are exact). A march sample sits at texel position t = fma(dv, k, ov); the shading
 *             sample, a free world point, at fma(p, n/2, n/2 - 0.5). Cell floor(t), weight w = t - floor(t),
 *             neighbours clamped to [0, n-1], lerp x then y then z as fma(w, b - a, a) on the raw byte values ->
 *             v in [0, 255].
 *   TF:       position u = fma(v, RN(128/255), -0.5), entry floor(u), weight u - floor(u), entries clamped to
 *             [0, 127], per channel fma(w, b - a, a).
 *   shading:  guard and light vector as GPURenderer4.cu:41-51 in separate roundings (rare path), second fetch
 *             v_l at p + L*0.01, diffuse = ((v_l - v) * RN(1/255)) * kd.
 *   blend:    t = 1 - acc.w; acc = fma(cur, t, acc) per channel.
 *
 * fmaf() is correctly rounded on every platform (glibc uses the FMA unit where there is one and an exact
 * software path elsewhere), so this section does not depend on the build's -mfma setting.
 * ------------------------------------------------------------------------------------------------------- */
#define ORC_TF_SCALE_255 0x1.010102p-1f /* RN(128 / 255) */
#define ORC_INV_255 0x1.010102p-8f      /* RN(1 / 255)   */

static inline float lerp_fma(float a, float b, float w) {
    float d = b - a;
    return fmaf(w, d, a);
}

/* texel position of a world coordinate */
static inline float texel_of_world(float p, uint32_t n) {
    float half_n = (float)n * 0.5f;
    float bias = half_n - 0.5f;
    return fmaf(p, half_n, bias);
}

static inline void tri_axis_fma(float t, uint32_t n, int64_t *i0, int64_t *i1, float *w) {
    float fl = floorf(t);
    *w = t - fl;
    int64_t a = (int64_t)fl, b = a + 1;
    if (a < 0) a = 0;
    if (a > (int64_t)n - 1) a = (int64_t)n - 1;
    if (b < 0) b = 0;
    if (b > (int64_t)n - 1) b = (int64_t)n - 1;
    *i0 = a; *i1 = b;
}

/* trilinear value on the byte scale, [0, 255], at texel position (tx, ty, tz) */
static inline float fetch_trilinear_fma(const vr5_model *m, float tx, float ty, float tz) {
    int64_t x0, x1, y0, y1, z0, z1;
    float wx, wy, wz;
    tri_axis_fma(tx, m->dims.x, &x0, &x1, &wx);
    tri_axis_fma(ty, m->dims.y, &y0, &y1, &wy);
    tri_axis_fma(tz, m->dims.z, &z0, &z1, &wz);
    const uint64_t sx = m->dims.x, sxy = (uint64_t)m->dims.x * m->dims.y;
#define VOX(X, Y, Z) ((float)m->data[(uint64_t)(Z) * sxy + (uint64_t)(Y) * sx + (uint64_t)(X)])
    float c00 = lerp_fma(VOX(x0, y0, z0), VOX(x1, y0, z0), wx);
    float c10 = lerp_fma(VOX(x0, y1, z0), VOX(x1, y1, z0), wx);
    float c01 = lerp_fma(VOX(x0, y0, z1), VOX(x1, y0, z1), wx);
    float c11 = lerp_fma(VOX(x0, y1, z1), VOX(x1, y1, z1), wx);
#undef VOX
    float c0 = lerp_fma(c00, c10, wy);
    float c1 = lerp_fma(c01, c11, wy);
    return lerp_fma(c0, c1, wz);
}

static inline vr5_f4 classify_linear_fma(const vr5_f4 *tf, float v) {
    float u = fmaf(v, ORC_TF_SCALE_255, -0.5f);
    float fl = floorf(u);
    float w = u - fl;
    int j0 = (int)fl, j1 = j0 + 1;
    if (j0 < 0) j0 = 0;
    if (j0 > VR5_TF_SIZE - 1) j0 = VR5_TF_SIZE - 1;
    if (j1 < 0) j1 = 0;
    if (j1 > VR5_TF_SIZE - 1) j1 = VR5_TF_SIZE - 1;
    vr5_f4 a = tf[j0], b = tf[j1], c;
    c.x = lerp_fma(a.x, b.x, w); c.y = lerp_fma(a.y, b.y, w); c.z = lerp_fma(a.z, b.z, w); c.w = lerp_fma(a.w, b.w, w);
    return c;
}

static void cast_pixel_trilinear_fma(const vr5_raycaster *r, uint8_t *rgba8, int col, int row, orc_counters *cnt) {
    vr5_f3 o, d;
    float k, k_far;
    primary_ray(&r->view, col, row, &o, &d);
    if (!clip_ray(&r->volume, o, d, &k, &k_far)) return;
    cnt->rays_hit++;
    float px = fmaf(d.x, k, o.x), py = fmaf(d.y, k, o.y), pz = fmaf(d.z, k, o.z);
    while (k <= k_far) {
        if (r->esl && block_is_empty(r, px, py, pz)) {
            k += leap_distance(r, px, py, pz, d);
            cnt->esl_iters++;
        } else {
            break;
        }
        k += r->ray_step;
        px = fmaf(d.x, k, o.x); py = fmaf(d.y, k, o.y); pz = fmaf(d.z, k, o.z);
    }
    if (k > k_far) return;
    cnt->rays_accum++;
    /* the ray in texel space */
    const vr5_model *m = &r->volume;
    float hnx = (float)m->dims.x * 0.5f, hny = (float)m->dims.y * 0.5f, hnz = (float)m->dims.z * 0.5f;
    float dvx = d.x * hnx, dvy = d.y * hny, dvz = d.z * hnz;
    float ovx = texel_of_world(o.x, m->dims.x), ovy = texel_of_world(o.y, m->dims.y), ovz = texel_of_world(o.z, m->dims.z);
    vr5_f4 acc = {0, 0, 0, 0};
    while (k <= k_far) {
        float v = fetch_trilinear_fma(m, fmaf(dvx, k, ovx), fmaf(dvy, k, ovy), fmaf(dvz, k, ovz));
        vr5_f4 cur = classify_linear_fma(r->transfer_fn, v);
        cnt->samples++;
        if (cur.w > 0.05f && r->light_kd > 0.01f) {
            px = fmaf(d.x, k, o.x); py = fmaf(d.y, k, o.y); pz = fmaf(d.z, k, o.z);
            float lx = r->view.light_pos.x - px, ly = r->view.light_pos.y - py, lz = r->view.light_pos.z - pz;
            float xx = lx * lx, yy = ly * ly, zz = lz * lz;
            float s = xx + yy;
            s = s + zz;
            float inv = 1.0f / sqrtf(s);
            float nx = lx * inv, ny = ly * inv, nz = lz * inv;
            float ox = nx * 0.01f, oy = ny * 0.01f, oz = nz * 0.01f;
            float qx = px + ox, qy = py + oy, qz = pz + oz;
            float v_l = fetch_trilinear_fma(m, texel_of_world(qx, m->dims.x), texel_of_world(qy, m->dims.y), texel_of_world(qz, m->dims.z));
            float diff = v_l - v;
            diff = diff * ORC_INV_255;
            diff = diff * r->light_kd;
            cur.x += diff; cur.y += diff; cur.z += diff;
            cnt->shaded++;
        }
        float t = 1 - acc.w;
        acc.x = fmaf(cur.x, t, acc.x); acc.y = fmaf(cur.y, t, acc.y); acc.z = fmaf(cur.z, t, acc.z); acc.w = fmaf(cur.w, t, acc.w);
        if (acc.w > r->ray_threshold) break;
        k += r->ray_step;
    }
    uint8_t *px8 = rgba8 + 4 * ((size_t)row * r->view.dims.x + (size_t)col);
    px8[0] = (uint8_t)unit_to_index(acc.x, 256);
    px8[1] = (uint8_t)unit_to_index(acc.y, 256);
    px8[2] = (uint8_t)unit_to_index(acc.z, 256);
    px8[3] = (uint8_t)unit_to_index(acc.w, 256);
}

int orc_render_rows_trilinear_fma(const vr5_raycaster *r, uint8_t *rgba8, int row0, int row1, orc_counters *cnt) {
    if (r->volume.data == NULL || r->transfer_fn == NULL || r->esl_volume == NULL || rgba8 == NULL) return 1;
    orc_counters local = {0, 0, 0, 0, 0};
    int w = r->view.dims.x, h = r->view.dims.y;
    if (row0 < 0) row0 = 0;
    if (row1 > h) row1 = h;
    if (row1 > row0) memset(rgba8 + (size_t)row0 * w * 4, 0, (size_t)(row1 - row0) * w * 4);
    for (int row = row0; row < row1; row++)
        for (int col = 0; col < w; col++) cast_pixel_trilinear_fma(r, rgba8, col, row, &local);
    if (cnt) *cnt = local;
    return 0;
}
