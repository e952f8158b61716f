"""GPU parity tests (run on the B200 box: pytest -m gpu). Every call goes through the C ABI of libvr5.so
(include/vr5.h) and is checked against the CPU oracle on identical inputs.

Tolerance (BASELINE.json north_star): <= 1 LSB per RGBA8 channel on >= 99.9 % of pixels and <= 2 LSB max -
asserted by common.assert_parity. The kernel is designed to be bit-exact, so the tests additionally assert 0 LSB.
"""
import json
import os

import numpy as np
import pytest

import oracle
from common import POSES, assert_parity, golden_scene, product_scene, sha, tf_by_name

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# kernel variant ids a sampler policy may run as: the window policies switch to their power-of-two specialisations (7..9)
# when every volume dimension is a power of two (vr5_cast.cuh)
# (8 for hybridw: a transfer function with a long all-zero prefix takes the window-only march, vr5_render.cu launch_variant)
VARIANT_OF = {"linear": {0}, "brick": {1}, "tex3d": {2}, "hybrid": {3}, "hybridw": {4, 7, 8}, "window": {5, 8}, "hybridw2": {6, 9}}
SAMPLERS = ["linear", "brick", "tex3d", "hybrid", "hybridw", "window", "hybridw2"]


def _cases():
    with open(os.path.join(os.path.dirname(__file__), "golden", "golden.json")) as f:
        return [c["name"] for c in json.load(f)["cases"]]


@pytest.fixture(scope="module")
def gpu(R):
    r = R.GPURenderer5(0)
    yield r
    r.close()


@pytest.mark.parametrize("sampler", SAMPLERS)
@pytest.mark.parametrize("name", _cases())
def test_golden_frames(name, sampler, golden, port, volumes, R, gpu):
    """The reference's own frames (sha256 committed in golden.json) reproduced by the CUDA path, all samplers."""
    case = next(c for c in golden["cases"] if c["name"] == name)
    sc = golden_scene(port, case, volumes(case["volume"]))
    scene, params = product_scene(R, sc)
    gpu.set_sampler(sampler)
    gpu.set_window_buffer(case["width"], case["height"])
    gpu.set_scene(scene)
    frame = gpu.render_volume(params)
    if sha(frame) != case["frame_sha256"]:  # diagnose through the oracle before failing
        want, _ = port.render(sc)
        assert_parity(frame, want, exact=True, what=name)
    assert sha(frame) == case["frame_sha256"]
    assert gpu.last_variant() in VARIANT_OF[sampler]  # the requested memory path ran, not a fallback
    assert gpu.count(params) == case["counters"]


@pytest.mark.parametrize("name", [n for n in _cases() if "96" not in n])
def test_golden_frames_wide_window(name, golden, port, volumes, R):
    """Kernel variant 10 (window arrays beyond 4 GiB: 64-bit slice term of the address) forced onto the small power-of-two
    goldens with VR5_P2_WIDE=1; full size: tests/test_gpu_full_size.py (2048^3)."""
    case = next(c for c in golden["cases"] if c["name"] == name)
    vol = volumes(case["volume"])
    if any(d & (d - 1) for d in vol.shape):
        pytest.skip("not a power-of-two volume")
    sc = golden_scene(port, case, vol)
    scene, params = product_scene(R, sc)
    g = R.GPURenderer5(0)
    try:
        g.set_option("VR5_P2_WIDE", "1")
        g.set_sampler("hybridw")
        g.set_window_buffer(case["width"], case["height"])
        g.set_scene(scene)
        frame = g.render_volume(params)
        assert g.last_variant() == 10, g.last_variant()
        assert sha(frame) == case["frame_sha256"]
        w, h = case["width"], case["height"]
        ptrs = [g.device_alloc(w * h * 4) for _ in range(2)]   # ... and as a frame sequence
        g.render_sequence([params, params], ptrs)
        g.sync()
        assert g.last_variant() == 10
        for p in ptrs:
            out = np.zeros((h, w, 4), dtype=np.uint8)
            g.memcpy_d2h(out, p)
            g.device_free(p)
            assert sha(out) == case["frame_sha256"]
    finally:
        g.close()


@pytest.mark.parametrize("sampler", SAMPLERS)
def test_random_volume_ragged(sampler, port, R, gpu):
    """Non-cubic volume, odd viewport (ragged 8x4 tiles), both projections, product-side scene builders."""
    rng = np.random.default_rng(11)
    vol = rng.integers(0, 256, size=(37, 53, 70), dtype=np.uint8)
    vol[:, :17] = 0
    vol[9:25] = 0
    gpu.set_sampler(sampler)
    for tf_name, kw in [("default", {}), ("opaque", {}), ("semi", {"ray_threshold": 1.0, "light_kd": 0.0}), ("default", {"esl": False})]:
        base = {"default": R.default_base_tf(), "opaque": R.tf_opaque(), "semi": R.tf_semi(70)}[tf_name]
        bd, mm, bs = R.build_minmax(vol)
        tf = R.premultiply_tf(base)
        scene = R.Scene(vol, tf, R.build_esl_bits(tf, mm), bd, bs, R.default_ray_step(70, 53, 37), **kw)
        gpu.set_scene(scene)
        for persp, ang, wh in [(0, POSES[1], (163, 97)), (1, (20, 130, 0), (129, 131)), (0, POSES[0], (61, 33)), (1, POSES[3], (8, 4))]:
            view = R.camera(wh[0], wh[1], persp, ang, 2.0)
            gpu.set_window_buffer(*wh)
            params = scene.params(view)
            frame = gpu.render_volume(params)
            osc = port.make_scene(vol, tf_by_name(port, tf_name, 70), oracle.View.from_buffer_copy(bytes(view)),
                                  threshold=scene.ray_threshold, esl=scene.esl, light_kd=scene.light_kd)
            want, cnt = port.render(osc)
            assert_parity(frame, want, exact=True, what=(sampler, tf_name, persp, ang))
            assert gpu.count(params) == cnt


def test_device_synth_equals_host_synth(port, gpu):
    for kind, n, seed, host in [(0, 64, 1, lambda: port.synth_a(64, 1)), (1, 64, 2, lambda: port.synth_b(64, 2)),
                                (0, 100, 3, lambda: port.synth_a(100, 3)), (1, 129, 5, lambda: port.synth_b(129, 5))]:
        gpu.synth_volume(kind, n, seed)
        assert np.array_equal(gpu.get_volume((n, n, n)), host()), (kind, n, seed)


def test_device_minmax_histogram(port, R, gpu):
    rng = np.random.default_rng(3)
    for shape in [(64, 64, 64), (37, 53, 70), (8, 300, 12)]:
        vol = rng.integers(0, 256, size=shape, dtype=np.uint8)
        vol[: shape[0] // 2] //= 4
        assert gpu.set_volume(vol) == 0
        bd, mm, bs = gpu.build_minmax_device()
        bd2, mm2, bs2 = port.minmax(vol)
        assert bd == bd2 and np.array_equal(mm, mm2) and tuple(np.float32(bs)) == tuple(bs2)
        assert np.array_equal(gpu.histogram_device(), port.histogram(vol))


def test_size_independent_properties_256(port, R, gpu):
    """Config-2 sized volume (256^3 @1920x1080): full-size checks that need no full oracle frame."""
    n, w, h = 256, 1920, 1080
    gpu.set_sampler("linear")
    gpu.synth_volume(0, n, 1)
    vol = gpu.get_volume((n, n, n))
    bd, mm, bs = gpu.build_minmax_device()
    tf = R.premultiply_tf(R.tf_semi(n))
    scene = R.Scene(None, tf, R.build_esl_bits(tf, mm), bd, bs, R.default_ray_step(n, n, n))
    gpu.set_transfer_fn(scene.tf, scene.esl_bits)
    gpu.set_window_buffer(w, h)
    view = R.camera(w, h, 0, (20, 40, 0), 2.0)
    params = scene.params(view)
    frames = {}
    for s in SAMPLERS:  # all memory paths give the same bytes; rendering twice is idempotent
        gpu.set_sampler(s)
        frames[s] = gpu.render_volume(params)
        assert np.array_equal(frames[s], gpu.render_volume(params))
    assert all(np.array_equal(frames["linear"], frames[s]) for s in SAMPLERS)
    # oracle on a band of rows through the middle and the ragged bottom rows
    osc = port.make_scene(vol, tf_by_name(port, "semi", n), oracle.View.from_buffer_copy(bytes(view)))
    for rows in [(536, 548), (0, 4), (1076, 1080)]:
        want, _ = port.render(osc, rows=rows)
        assert_parity(frames["linear"][rows[0]:rows[1]], want[rows[0]:rows[1]], exact=True, what=rows)
    # semi TF: alpha stays below the 0.95 threshold -> no early termination, no shading: samples depend on geometry only
    cnt = gpu.count(params)
    assert cnt["shaded"] == 0 and cnt["esl_iters"] == 0 and cnt["rays_accum"] == cnt["rays_hit"]
    assert cnt["rays_hit"] == int((frames["linear"][..., 3] > 0).sum())


def test_partition_pack_unpack(R, gpu, port, golden, volumes):
    """Image-space partition: every rank's strips (packed or in place) reassemble to the single-GPU frame."""
    case = next(c for c in golden["cases"] if c["name"] == "synthA64_default_persp")  # 317 x 203: ragged
    sc = golden_scene(port, case, volumes(case["volume"]))
    scene, params = product_scene(R, sc)
    w, h = case["width"], case["height"]
    gpu.set_sampler("linear")
    gpu.set_window_buffer(w, h)
    gpu.set_scene(scene)
    gpu.set_partition(0, 1, False)
    whole = gpu.render_volume(params)
    L = R.lib()
    for world in (2, 3, 8):
        rows_padded = max(L.vr5_strip_rows(h, r, world) for r in range(world))
        packed = gpu.device_alloc(world * rows_padded * w * 4)
        frame_dev = gpu.device_alloc(w * h * 4)
        in_place = np.zeros((h, w, 4), dtype=np.uint8)
        for rank in range(world):
            gpu.set_partition(rank, world, True)
            gpu.render_device(params, packed + rank * rows_padded * w * 4)
            gpu.set_partition(rank, world, False)
            part = gpu.render_volume(params, out=in_place.copy())
            own = (np.arange(h) // 4) % world == rank
            assert np.array_equal(part[own], whole[own])
            in_place[own] = part[own]
        assert np.array_equal(in_place, whole)
        gpu.unpack_strips(packed, rows_padded, world, frame_dev)
        gpu.sync()
        got = np.zeros((h, w, 4), dtype=np.uint8)
        gpu.memcpy_d2h(got, frame_dev)
        assert np.array_equal(got, whole), world
        gpu.device_free(packed)
        gpu.device_free(frame_dev)
    gpu.set_partition(0, 1, False)


@pytest.mark.parametrize("order", ["cols", "rings"])
def test_queue_orders_cover_every_tile(order, R, gpu, port, golden, volumes, monkeypatch):
    """The opt-in tile-queue orders (VR5_QUEUE_ORDER, build_group_order) are permutations of the row-major sweep: every pixel
    is still rendered exactly once, whole frame and per-rank strips, on ragged window sizes."""
    case = next(c for c in golden["cases"] if c["name"] == "synthA64_default_persp")  # 317 x 203: ragged
    sc = golden_scene(port, case, volumes(case["volume"]))
    scene, params = product_scene(R, sc)
    w, h = case["width"], case["height"]
    gpu.set_sampler("hybridw")
    gpu.set_window_buffer(w, h)
    gpu.set_scene(scene)
    gpu.set_partition(0, 1, False)
    whole = gpu.render_volume(params)
    assert sha(whole) == case["frame_sha256"]
    gpu.set_option("VR5_QUEUE_ORDER", order)   # the environment is read once, at vr5_create; later changes go through the ABI
    try:
        assert np.array_equal(gpu.render_volume(params), whole)
        for world in (2, 3):
            in_place = np.zeros((h, w, 4), dtype=np.uint8)
            for rank in range(world):
                gpu.set_partition(rank, world, False)
                part = gpu.render_volume(params, out=in_place.copy())
                own = (np.arange(h) // 4) % world == rank
                in_place[own] = part[own]
            assert np.array_equal(in_place, whole), (order, world)
    finally:
        gpu.set_option("VR5_QUEUE_ORDER", "rows")
        gpu.set_partition(0, 1, False)


@pytest.mark.parametrize("sampler", ["hybridw", "window", "linear", "hybrid", "tex3d", "trilinear", "trilinear-fma"])
@pytest.mark.parametrize("perspective", [0, 1])
def test_degenerate_view_keeps_the_clamps(sampler, perspective, port, R):
    """the C ABI accepts any View bytes. A direction with an exactly-zero x component is fixed up to 1e-5 only
    inside Raycaster::intersect (RaycasterBase.h:33-35); with the origin outside [-1, 1] on that axis and the other components
    tiny, rays get a valid interval at k ~ 1e5 while every sample lies OUTSIDE the cube on x. The reference clamps the voxel
    index (common.h:105-110); the unclamped kernel variants would form addresses far outside their guard margins. The launcher
    must route such frames through the clamped variants: frames bit-equal to the CPU checker, no fault."""
    rng = np.random.default_rng(3)
    vol = rng.integers(1, 256, size=(64, 64, 64), dtype=np.uint8)
    w, h = 64, 48
    view = R.View()
    view.dims[0], view.dims[1] = w, h
    view.origin = R.F3(-2.0, 0.0, 0.0)
    view.direction = R.F3(0.0, 1e-6, -1e-6)
    view.right_plane = R.F3(0.004, 0.0, 0.0) if not perspective else R.F3(0.0, 1e-9, 0.0)
    view.up_plane = R.F3(0.0, 0.01, 0.01) if not perspective else R.F3(0.0, 0.0, 1e-9)
    view.light_pos = R.F3(0.0, 0.0, 3.0)
    view.perspective = perspective
    base = R.tf_semi(8)
    tf = R.premultiply_tf(base)
    bd, mm, bs = R.build_minmax(vol)
    scene = R.Scene(vol, tf, R.build_esl_bits(tf, mm), bd, bs, 2.0e4, esl=False)
    params = scene.params(view)
    osc = port.make_scene(vol, oracle.tf_semi(port.default_base_tf(), 8), oracle.View.from_buffer_copy(bytes(view)), ray_step=2.0e4, esl=False)
    mode = {"trilinear": True, "trilinear-fma": "fma"}.get(sampler, False)
    want, cnt = port.render(osc, trilinear=mode)
    assert cnt["samples"] > 1000 and want.max() > 0     # the degenerate rays do hit and do accumulate colour
    g = R.GPURenderer5(0, sampler=sampler)
    try:
        g.set_window_buffer(w, h)
        g.set_scene(scene)
        frame = g.render_volume(params)
        assert_parity(frame, want, exact=True, what=(sampler, perspective))
        assert g.last_variant() in (0, 2, 16, 18), g.last_variant()   # linear / texture with clamps, clamped trilinear marches
        assert g.count(params) == cnt
    finally:
        g.close()


def test_error_behaviour(R):
    """Reference conventions: 1 = failure, never exit (GPURenderer1.cu:86-112, cuda_utils.h:21-49)."""
    r = R.GPURenderer5(0)
    assert "GPURenderer5" in r.get_name()
    assert r.set_volume(None) == 1 and "no data" in r.last_error()
    p = R.Params()
    p.view.dims[0], p.view.dims[1] = 16, 16
    out = np.zeros((16, 16, 4), dtype=np.uint8)
    assert r._L.vr5_render(r._h, p, out.ctypes.data, 0) == 1  # nothing set yet
    r.set_window_buffer(16, 16)
    assert r.set_volume(np.zeros((4, 4, 4), dtype=np.uint8)) == 0
    assert r._L.vr5_render(r._h, p, out.ctypes.data, 0) == 1 and "transfer function" in r.last_error()
    r.set_transfer_fn(np.zeros((128, 4), np.float32), np.zeros(1024, np.uint32))
    assert r._L.vr5_render(r._h, p, None, 0) == 1  # NULL buffer
    p.view.dims[0] = 32
    assert r._L.vr5_render(r._h, p, out.ctypes.data, 0) == 1 and "dims" in r.last_error()
    r.close()


def test_reference_benchmark_sweep_matches_oracle():
    """tools/volr_suite.py --check: the reference's 24-configuration benchmark matrix (VolR.cpp:225-321: datasets, ERT / ESL
    toggles, viewport scales 0.9..0.3, ray step x1.1..x1.7; 2 projections x 4 poses each), every frame equal to the oracle."""
    import subprocess
    import sys
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "volr_suite.py"), "--n", "48", "--window", "200", "150", "--check", "--reps", "1"],
                         capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "24 configurations, 192 frames, all equal to the CPU oracle" in out.stdout
