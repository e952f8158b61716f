"""Zero-skip kernels (vr5_cast.cuh, ZS): when the transfer function starts with all-zero entries - the reference's own tables
do, RaycasterBase.cpp:76-84 - a sample that classifies to one of them leaves the ray as it found it (CPURenderer.cpp:29-36:
acc + 0 * (1 - acc.a)), so its LUT fetch, compositing and termination test are skipped. The pixels must stay those of the
reference's loop, which composites every sample: checked bit for bit against the CPU oracle, with the skip on and off, and on
tables / thresholds for which the host must NOT enable it (alpha outside [0, 1], non-finite entries, a negative threshold).
"""
import numpy as np
import pytest

import oracle
from common import POSES, assert_parity, product_scene, sha

pytestmark = pytest.mark.gpu
SAMPLERS = ["linear", "brick", "tex3d", "hybrid", "hybridw", "window", "hybridw2"]


def _volume(shape, seed):
    """Mostly low values (inside the transparent prefix of the default TF: voxel < 26) with denser blobs, so that transparent and
    visible samples alternate along a ray and differ between the lanes of a warp."""
    rng = np.random.default_rng(seed)
    vol = rng.integers(0, 40, size=shape, dtype=np.uint8)
    blobs = rng.integers(0, 256, size=shape, dtype=np.uint8)
    mask = rng.random(shape) < 0.08
    vol[mask] = blobs[mask]
    vol[: shape[0] // 3] = 0
    return np.minimum(vol, 253)   # TF entry 127 is never sampled (one table below puts an infinity there)


def _tables(port):
    base = port.default_base_tf()
    out = {"default": (base.copy(), True), "opaque": (oracle.tf_opaque(base), True)}
    t = base.copy(); t[:, 3] = np.float32(1.0) / np.float32(64)
    out["semi"] = (t, False)                           # no zero entry at all
    t = base.copy(); t[40, 3] = 1.5
    out["alpha_above_one"] = (t, False)                # acc.a may leave [0, 1]: the host keeps the plain kernels
    t = base.copy(); t[50, 3] = -0.25
    out["negative_alpha"] = (t, False)
    t = base.copy(); t[127, 0] = np.inf
    out["infinite_colour"] = (t, False)                # never sampled (voxels <= 253), but the table is not finite: plain kernels
    t = base.copy(); t[:5, 3] = -0.0
    out["minus_zero_prefix"] = (t, True)               # -0.0 == 0.0: still a zero entry
    t = base.copy(); t[:30] = 0.0; t[30:, 3] = 0.02
    out["long_prefix_thin"] = (t, True)
    return out


@pytest.mark.parametrize("sampler", SAMPLERS)
def test_zero_skip_is_invisible(sampler, port, R):
    g = R.GPURenderer5(0)
    try:
        g.set_sampler(sampler)
        for shape, seed in [((64, 64, 64), 5), ((37, 53, 70), 6)]:
            vol = _volume(shape, seed)
            for name, (base, expect_skip) in _tables(port).items():
                for persp, ang, wh, kw in [(0, POSES[1], (163, 97), {}), (1, (20, 130, 0), (129, 131), {"esl": False}),
                                           (0, POSES[0], (96, 64), {"threshold": 0.0}), (1, POSES[3], (64, 48), {"threshold": -0.5})]:
                    view = oracle.View.from_buffer_copy(bytes(R.camera(wh[0], wh[1], persp, ang, 2.0)))
                    sc = port.make_scene(vol, base, view, **kw)
                    scene, params = product_scene(R, sc)
                    want, cnt = port.render(sc)
                    g.set_window_buffer(*wh)
                    g.set_scene(scene)
                    frames = {}
                    for off in ("0", "1"):
                        g.set_option("VR5_NO_ZSKIP", off)
                        frames[off] = g.render_volume(params)
                        skip = g.last_zero_skip()
                        assert skip == (1 if (off == "0" and expect_skip and kw.get("threshold", 0.95) >= 0.0) else 0), (name, kw, off, skip)
                        assert_parity(frames[off], want, exact=True, what=(sampler, shape, name, kw, "no_zskip=" + off))
                    g.set_option("VR5_NO_ZSKIP", "0")
                    assert g.count(params) == cnt
    finally:
        g.close()


def test_zero_skip_in_sequences_and_lanes(port, R):
    """The same frames through vr5_render_sequence (one launch for the camera path) and through frame lanes."""
    vol = _volume((64, 64, 64), 9)
    base = port.default_base_tf()
    w, h = 160, 96
    g = R.GPURenderer5(0)
    try:
        g.set_window_buffer(w, h)
        scs = [port.make_scene(vol, base, oracle.View.from_buffer_copy(bytes(R.camera(w, h, p, a, 2.0)))) for p, a in
               [(0, POSES[0]), (0, POSES[1]), (0, (20, 200, 0)), (0, POSES[3])]]   # one projection per sequence (vr5_render_sequence)
        scene, _ = product_scene(R, scs[0])
        g.set_scene(scene)
        params = [product_scene(R, sc)[1] for sc in scs]
        want = [port.render(sc)[0] for sc in scs]
        ptrs = [g.device_alloc(w * h * 4) for _ in params]
        g.render_sequence(params, ptrs)
        g.sync()
        assert g.last_zero_skip() == 1
        for p, ref in zip(ptrs, want):
            out = np.zeros((h, w, 4), dtype=np.uint8)
            g.memcpy_d2h(out, p)
            g.device_free(p)
            assert sha(out) == sha(ref)
    finally:
        g.close()


@pytest.mark.parametrize("policy", [("trilinear", True), ("trilinear-fma", "fma")], ids=["trilinear", "trilinear-fma"])
@pytest.mark.parametrize("clamped", [False, True])
def test_zero_skip_trilinear_policies(policy, clamped, port, R, monkeypatch):
    """The trilinear policies (vr5_tri.cuh, GPURenderer4 semantics) skip the classification and compositing of samples whose TF
    interpolation interval lies inside the table's all-zero prefix: bit-equal to the restated trilinear checker, skip on and off."""
    if clamped:
        monkeypatch.setenv("VR5_TRI_CLAMPED", "1")
    sampler, mode = policy
    g = R.GPURenderer5(0, sampler=sampler)
    try:
        for shape, seed in [((64, 64, 64), 15), ((37, 53, 70), 16)]:
            vol = _volume(shape, seed)
            vol[:, : shape[1] // 2] //= 3          # a region whose interpolated values stay inside the default table's zero prefix
            for name, (base, expect_skip) in _tables(port).items():
                for persp, ang, wh, kw in [(0, POSES[1], (163, 97), {}), (1, (20, 130, 0), (129, 131), {"esl": False}),
                                           (0, POSES[0], (96, 64), {"threshold": 0.0}), (1, POSES[3], (64, 48), {"threshold": -0.5})]:
                    view = oracle.View.from_buffer_copy(bytes(R.camera(wh[0], wh[1], persp, ang, 2.0)))
                    sc = port.make_scene(vol, base, view, **kw)
                    scene, params = product_scene(R, sc)
                    want, cnt = port.render(sc, trilinear=mode)
                    g.set_window_buffer(*wh)
                    g.set_scene(scene)
                    for off in ("0", "1"):
                        g.set_option("VR5_NO_ZSKIP", off)
                        frame = g.render_volume(params)
                        skip = g.last_zero_skip()
                        assert skip == (1 if (off == "0" and expect_skip and kw.get("threshold", 0.95) >= 0.0) else 0), (name, kw, off, skip)
                        assert_parity(frame, want, exact=True, what=(sampler, clamped, shape, name, kw, "no_zskip=" + off))
                    g.set_option("VR5_NO_ZSKIP", "0")
                    assert g.count(params) == cnt
    finally:
        g.close()
