"""Leaps over transparent bricks (vr5_cast.cuh, try_leap; zero-skip kernels only). The reference marches through empty space sample
by sample after its leading loop (CPURenderer.cpp:28-39); the kernel leaves out runs of samples that provably read voxels of a brick
whose maximum classifies to an all-zero TF entry and continues at the ray parameter the reference's sequential k += step would hold
(bits(k) + n * M inside one binade, tools/check_kleap.c). Pixels must stay bit-equal to the CPU checker - for every brick size, for
steps that are exact / inexact / ties in a binade, for ray parameters in several binades, both projections, ragged volumes - and the
leap counters must show that leaps were actually taken (vr5_leap_stats).
"""
import os

import numpy as np
import pytest

import oracle
from common import POSES, assert_parity, product_scene, sha

pytestmark = pytest.mark.gpu


def _sparse_volume(shape, seed):
    """Background below 100 (transparent under the tables below), a few dense blobs and solid slabs above it."""
    rng = np.random.default_rng(seed)
    vol = rng.integers(0, 60, size=shape, dtype=np.uint8)
    dz, dy, dx = shape
    zz, yy, xx = np.meshgrid(np.arange(dz), np.arange(dy), np.arange(dx), indexing="ij")
    for _ in range(5):
        c = [rng.uniform(0.15, 0.85) * n for n in (dz, dy, dx)]
        rad = rng.uniform(0.06, 0.16) * min(shape)
        m = (zz - c[0]) ** 2 + (yy - c[1]) ** 2 + (xx - c[2]) ** 2 < rad * rad
        vol[m] = rng.integers(130, 256, size=int(m.sum()), dtype=np.uint8)
    vol[: max(1, dz // 9)] = 0
    return vol


def _tables(port):
    base = port.default_base_tf()
    opaque = oracle.tf_opaque(base)                      # zero below entry 64 (voxel < 128), alpha 1 above: first hit ends the ray
    thin = base.copy(); thin[:50] = 0.0; thin[50:, 3] = 0.03   # zero below entry 50 (voxel < 100), thin above: rays go on through blobs and gaps
    return {"opaque": opaque, "thin": thin}


_WANT = {}   # the CPU checker's frames, shared by the parametrisations


def _render_both(g, R, port, vol, base, view, kw):
    sc = port.make_scene(vol, base, view, **kw)
    scene, params = product_scene(R, sc)
    key = (vol.shape, sha(vol), sha(base), bytes(view), repr(sorted(kw.items())))
    if key not in _WANT:
        _WANT[key] = port.render(sc)[0]
    want = _WANT[key]
    g.set_scene(scene)
    g.leap_stats()
    frame = g.render_volume(params)
    return frame, want, g.leap_stats()


@pytest.mark.parametrize("brick_log2", [2, 3, 4, 5])
@pytest.mark.parametrize("sampler", ["hybridw", "window", "tex3d", "linear"])
def test_leaps_are_invisible(sampler, brick_log2, port, R):
    g = R.GPURenderer5(0)
    try:
        g.set_option("VR5_LEAP_BRICK", str(brick_log2))
        g.set_option("VR5_LEAP_STATS", "1")
        g.set_sampler(sampler)
        total = [0, 0]
        for shape, seed in [((64, 64, 64), 21), ((37, 53, 70), 22)]:
            vol = _sparse_volume(shape, seed)
            n = max(shape)
            step0 = float(np.float32(2.0) / np.float32(n) - np.float32(2.0) / np.float32(n) / np.float32(n))
            for name, base in _tables(port).items():
                for persp, ang, wh, dist, step in [
                        (0, POSES[1], (163, 97), 2.0, step0), (1, (20, 130, 0), (129, 131), 2.0, step0 * 1.7),
                        (0, POSES[0], (96, 64), 6.0, step0 / 3), (1, POSES[3], (64, 48), 1.2, step0),
                        (0, (20, 200, 0), (80, 56), 2.0, 2.0 ** -6),                   # exact in every binade that matters: M * u == step
                        (0, (35, 40, 0), (80, 56), 2.0, 2.0 ** -5 + 2.0 ** -24),        # a tie (r == u / 2) in [1, 2): no leaps there
                        (1, (10, 300, 0), (80, 56), 3.0, step0 * 0.77)]:
                    view = oracle.View.from_buffer_copy(bytes(R.camera(wh[0], wh[1], persp, ang, dist)))
                    g.set_window_buffer(*wh)
                    frame, want, st = _render_both(g, R, port, vol, base, view, {"ray_step": step})
                    assert g.last_zero_skip() == 1
                    assert_parity(frame, want, exact=True, what=(sampler, brick_log2, shape, name, persp, ang, dist, step, st))
                    total[0] += st[0]; total[1] += st[1]
        assert total[0] > 0 and total[1] > total[0], total      # leaps were taken, and each left out at least one sample
    finally:
        g.close()


def test_leaps_off_and_in_sequences(port, R):
    """VR5_NO_LEAP (read at the next layout build) marches sample by sample: same frames, no leaps counted; a frame sequence leaps too."""
    vol = _sparse_volume((96, 96, 96), 31)
    base = _tables(port)["thin"]
    w, h = 160, 96
    # (camera distances differ: the frames of a sequence ask for different leap margins, the launch takes the widest)
    views = [oracle.View.from_buffer_copy(bytes(R.camera(w, h, 0, a, dist))) for a, dist in [(POSES[0], 2.0), (POSES[1], 2.0), ((20, 200, 0), 8.0), (POSES[3], 3.0)]]
    scs = [port.make_scene(vol, base, v) for v in views]
    want = [port.render(sc)[0] for sc in scs]
    for off in ("0", "1"):
        g = R.GPURenderer5(0)
        try:
            g.set_option("VR5_NO_LEAP", off)
            g.set_option("VR5_LEAP_STATS", "1")
            g.set_option("VR5_LEAP_BRICK", "3")   # (the default edge, 32, leaves a 96^3 volume 27 bricks: too few to count on an empty one)
            g.set_window_buffer(w, h)
            scene, _ = product_scene(R, scs[0])
            g.set_scene(scene)
            params = [product_scene(R, sc)[1] for sc in scs]
            g.leap_stats()
            for p, ref in zip(params, want):
                assert sha(g.render_volume(p)) == sha(ref)
            st = g.leap_stats()
            assert (st[0] > 0) == (off == "0"), (off, st)
            ptrs = [g.device_alloc(w * h * 4) for _ in params]
            g.render_sequence(params, ptrs)
            g.sync()
            st = g.leap_stats()
            assert (st[0] > 0) == (off == "0"), (off, st)
            for p, ref in zip(ptrs, want):
                out = np.zeros((h, w, 4), dtype=np.uint8)
                g.memcpy_d2h(out, p)
                g.device_free(p)
                assert sha(out) == sha(ref)
        finally:
            g.close()


def test_leaps_random_views(port, R):
    """150 random views of a sparse volume - any angle, distance 1.05 .. 9 (ray parameters from ~0 up to 12: many binades), step factor
    0.3 .. 2.5 (plus steps with few mantissa bits: exact sums and ties), both projections, two brick edges - bit-equal to the CPU checker."""
    rng = np.random.default_rng(77)
    vol = _sparse_volume((64, 80, 56), 41)
    n = max(vol.shape)
    step0 = float(np.float32(2.0) / np.float32(n) - np.float32(2.0) / np.float32(n) / np.float32(n))
    tables = _tables(port)
    w, h = 72, 56
    leaps = 0
    for lb in ("3", "4"):
        g = R.GPURenderer5(0)
        try:
            g.set_option("VR5_LEAP_BRICK", lb)
            g.set_option("VR5_LEAP_STATS", "1")
            g.set_window_buffer(w, h)
            for i in range(int(os.environ.get("VR5_LEAP_SOAK", "75"))):   # (a soak run: VR5_LEAP_SOAK=2000)
                ang = (float(rng.uniform(-90, 90)), float(rng.uniform(0, 360)), float(rng.uniform(-30, 30)))
                dist = float(rng.choice([1.05, 1.3, 2.0, 2.0, 3.1, 4.0, 6.5, 9.0]))
                step = step0 * float(rng.uniform(0.3, 2.5))
                if i % 5 == 0:   # few mantissa bits: exact in some binades, a tie in others
                    step = float(np.float32(2.0 ** -6 + 2.0 ** -int(rng.integers(20, 27))))
                persp = int(rng.integers(0, 2))
                base = tables["opaque" if i % 2 else "thin"]
                view = oracle.View.from_buffer_copy(bytes(R.camera(w, h, persp, ang, dist)))
                frame, want, st = _render_both(g, R, port, vol, base, view, {"ray_step": step})
                assert_parity(frame, want, exact=True, what=(lb, i, ang, dist, step, persp, st))
                leaps += st[0]
        finally:
            g.close()
    assert leaps > 1000, leaps
