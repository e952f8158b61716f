"""Independent check of the trilinear restatement: a float64 numpy ray march written from the CUDA texture
filtering formulas (CUDA C Programming Guide, "Texture Fetching", linear filtering with normalised coordinates), NOT from
oracle_port.c:

    tex(x, y, z), N texels per axis, clamp addressing:
        xB = x * N - 0.5,  i = floor(xB),  alpha = xB - i            (same for y -> j, beta and z -> k, gamma)
        T = sum over the 8 corners of weight * texel[clamp(i + di), clamp(j + dj), clamp(k + dk)]
    cudaReadModeNormalizedFloat on uint8: texel / 255.
    tex1D on the 128-entry transfer function: the same formula in one dimension, per channel.

GPURenderer4's march (GPURenderer4.cu:53-87) around it: sample = tex3D(volume, (pt + 1) / 2), colour = tex1D(tf, sample),
shade_texture when colour.w > 0.05 and light_kd > 0.01, front-to-back blend, early termination, k += step in fp32.
Everything except the k sequence (which decides WHERE samples sit, and is the reference's fp32 accumulation) runs in
float64, so agreement with the fp32 restatements is by tolerance: the BASELINE tolerance (<= 1 LSB on >= 99.9 % of the
pixels, <= 2 LSB max) on scenes without early termination, and a stated looser bound where a threshold decision can flip
a whole sample (early ray termination, shading guard)."""
import numpy as np
import pytest

import oracle
from common import parity_report


def tex_linear_3d(vol, x, y, z):
    """vol [D, H, W] float64 texel values; x, y, z normalised coordinates (arrays). Clamp addressing."""
    d, h, w = vol.shape

    def axis(c, n):
        b = c * n - 0.5
        i = np.floor(b)
        return i.astype(np.int64), b - i

    i, a = axis(x, w)
    j, b = axis(y, h)
    k, g = axis(z, d)

    def at(kk, jj, ii):
        return vol[np.clip(kk, 0, d - 1), np.clip(jj, 0, h - 1), np.clip(ii, 0, w - 1)]

    return ((1 - a) * (1 - b) * (1 - g) * at(k, j, i) + a * (1 - b) * (1 - g) * at(k, j, i + 1) +
            (1 - a) * b * (1 - g) * at(k, j + 1, i) + a * b * (1 - g) * at(k, j + 1, i + 1) +
            (1 - a) * (1 - b) * g * at(k + 1, j, i) + a * (1 - b) * g * at(k + 1, j, i + 1) +
            (1 - a) * b * g * at(k + 1, j + 1, i) + a * b * g * at(k + 1, j + 1, i + 1))


def tex_linear_1d(table, x):
    """table [N, 4]; x normalised. Clamp addressing."""
    n = table.shape[0]
    b = x * n - 0.5
    i = np.floor(b)
    a = (b - i)[:, None]
    i = i.astype(np.int64)
    return (1 - a) * table[np.clip(i, 0, n - 1)] + a * table[np.clip(i + 1, 0, n - 1)]


def march_float64(sc: oracle.Scene) -> np.ndarray:
    rc = sc.rc
    assert not rc.esl, "the independent march has no empty-space loop (it never changes a lit pixel's k sequence start otherwise)"
    w, h = sc.width, sc.height
    v = rc.view
    f3 = lambda p: np.array([p.x, p.y, p.z], dtype=np.float64)
    cols, rows = np.meshgrid(np.arange(w), np.arange(h))
    cx = (cols - w // 2).astype(np.float64).ravel()[:, None]
    cy = (rows - h // 2).astype(np.float64).ravel()[:, None]
    n_px = w * h
    # rays in fp32 like View::get_ray / Raycaster::intersect (ViewBase.h:23-35, RaycasterBase.h:32-42): they fix the k interval
    f = np.float32
    if v.perspective:
        o = np.tile(f3(v.origin).astype(f), (n_px, 1))
        d = (f3(v.direction).astype(f) + f3(v.right_plane).astype(f) * cx.astype(f)).astype(f)
        d = (d + f3(v.up_plane).astype(f) * cy.astype(f)).astype(f)
    else:
        d = np.tile(f3(v.direction).astype(f), (n_px, 1))
        o = (f3(v.origin).astype(f) + f3(v.right_plane).astype(f) * cx.astype(f)).astype(f)
        o = (o + f3(v.up_plane).astype(f) * cy.astype(f)).astype(f)
    e = np.where(d == 0, f(0.00001), d)
    with np.errstate(all="ignore"):
        k1 = ((f(-1.0) - o) / e).astype(f)
        k2 = ((f(1.0) - o) / e).astype(f)
    near = np.minimum(k1, k2).max(axis=1)
    far = np.maximum(k1, k2).min(axis=1)
    near = np.maximum(near, f(0))
    alive = (near < far) & (far > 0)
    k = near.copy()
    step, thr, kd = f(rc.ray_step), np.float64(f(rc.ray_threshold)), np.float64(f(rc.light_kd))
    vol = sc.volume.astype(np.float64) / 255.0
    tf = sc.tf.astype(np.float64)
    light = f3(v.light_pos)
    acc = np.zeros((n_px, 4))
    wrote = alive.copy()
    o64, d64 = o.astype(np.float64), d.astype(np.float64)
    alive &= k <= far
    while alive.any():
        idx = np.nonzero(alive)[0]
        pt = o64[idx] + d64[idx] * k[idx].astype(np.float64)[:, None]
        s = tex_linear_3d(vol, (pt[:, 0] + 1) * 0.5, (pt[:, 1] + 1) * 0.5, (pt[:, 2] + 1) * 0.5)
        c = tex_linear_1d(tf, s)
        sh = (c[:, 3] > 0.05) & (kd > 0.01)
        if sh.any():
            p = pt[sh]
            ld = light - p
            ld /= np.linalg.norm(ld, axis=1)[:, None]
            q = p + ld * 0.01
            sl = tex_linear_3d(vol, (q[:, 0] + 1) * 0.5, (q[:, 1] + 1) * 0.5, (q[:, 2] + 1) * 0.5)
            c[sh, :3] += ((sl - s[sh]) * kd)[:, None]
        acc[idx] += c * (1 - acc[idx, 3])[:, None]
        done = acc[idx, 3] > thr
        k[idx] = (k[idx] + step).astype(f)
        alive[idx] = ~done & (k[idx] <= far[idx])
    out = np.clip(np.trunc(acc * 256.0), 0, 255).astype(np.uint8)   # write_color: map_float_int(c, 256)
    out[~wrote] = 0
    return out.reshape(h, w, 4)


def scene(port, R, tf_name, persp, angles, wh, **kw):
    rng = np.random.default_rng(11)
    # smooth-ish but non-trivial data: random field blurred once, so neighbouring voxels differ by tens, not hundreds
    raw = rng.integers(0, 256, size=(22, 26, 30)).astype(np.float64)
    vol = ((raw + np.roll(raw, 1, 0) + np.roll(raw, 1, 1) + np.roll(raw, 1, 2)) / 4).astype(np.uint8)
    base = port.default_base_tf()
    if tf_name == "semi":
        base = oracle.tf_semi(base, 30)
    view = oracle.View.from_buffer_copy(bytes(R.camera(wh[0], wh[1], persp, angles, 2.0)))
    return port.make_scene(vol, base, view, esl=False, **kw)


@pytest.mark.parametrize("mode", [True, "fma"], ids=["one-rounding", "contracted"])
@pytest.mark.parametrize("persp,angles,wh", [(0, (20.0, 35.0, 0.0), (96, 72)), (1, (-45.0, -45.0, 0.0), (83, 61))])
def test_restated_trilinear_matches_float64_texture_formula_no_termination(mode, persp, angles, wh, port, R):
    """Semi-transparent TF, threshold 1, no light: no decision can flip, so the only difference is fp32 rounding of the
    filtering and blending - held to the BASELINE tolerance."""
    sc = scene(port, R, "semi", persp, angles, wh, threshold=1.0, light_kd=0.0)
    want = march_float64(sc)
    got, cnt = port.render(sc, trilinear=mode)
    rep = parity_report(got, want)
    assert cnt["samples"] > 50_000 and (want.reshape(-1, 4).max(axis=1) > 0).mean() > 0.2
    assert rep["within1"] >= 0.999 and rep["max"] <= 2, rep


@pytest.mark.parametrize("mode", [True, "fma"], ids=["one-rounding", "contracted"])
def test_restated_trilinear_matches_float64_texture_formula_default_tf(mode, port, R):
    """Default TF with shading and early termination: a sample whose alpha sits on the 0.05 shading guard or a ray whose
    opacity sits on the 0.95 threshold can take the other branch in fp32, which changes one sample's contribution. Stated
    bound for that: <= 1 LSB on >= 99 % of the pixels, <= 8 LSB max (one extra sample of weight <= 0.05 at alpha <= 0.5
    is <= 0.025 * 256 = 6.4 LSB)."""
    sc = scene(port, R, "default", 1, (30.0, 60.0, 0.0), (90, 70))
    want = march_float64(sc)
    got, cnt = port.render(sc, trilinear=mode)
    rep = parity_report(got, want)
    assert cnt["shaded"] > 1000
    assert rep["within1"] >= 0.99 and rep["max"] <= 8, rep
