"""Cost-balanced band partition (include/vr5.h vr5_balance_bands / vr5_set_partition_bands; north_star: "each GPU renders a
load-balanced row strip"). With interleaved strips every GPU reads the whole volume for 1/N of the samples (measured on one
B200 playing rank 0 of 8: 726-752 Gsamples/s against 967-1004 in a contiguous band, tools/sim_rank.py); bands keep a GPU in its
slab of the volume and are re-balanced for every view from the clipped ray lengths."""
import numpy as np
import pytest

from common import golden_scene, product_scene


def _params(R, w, h, persp, angles, n=64):
    tf = R.premultiply_tf(R.tf_semi(n))
    scene = R.Scene(None, tf, np.zeros(1024, np.uint32), 8, (0.25, 0.25, 0.25), R.default_ray_step(n, n, n))
    return scene.params(R.camera(w, h, persp, angles, 2.0))


@pytest.mark.parametrize("persp", [0, 1])
@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_balance_bands_properties(world, persp, R, port):
    """Bounds are monotone, start at 0, end at the strip count, every band has a strip; the bands' true costs (the CPU checker's
    sample counts of their rows, semi-transparent TF: cost = samples) are within 12 % of the mean for orbit views."""
    import oracle
    w, h = 320, 544     # 136 strips: the cut granularity (one strip) is 6 % of a band at world 8
    strips = (h + 3) // 4
    vol = np.full((64, 64, 64), 90, dtype=np.uint8)
    worst = 0.0
    for ang in [(20.0, 0.0, 0.0), (20.0, 37.0, 0.0), (-45.0, -45.0, 0.0), (90.0, 10.0, 5.0)]:
        p = _params(R, w, h, persp, ang)
        b = R.balance_bands(p, world)
        assert b[0] == 0 and b[-1] == strips and np.all(np.diff(b.astype(np.int64)) >= 1)
        if world > 1:
            sc = port.make_scene(vol, oracle.tf_semi(port.default_base_tf(), 64), oracle.View.from_buffer_copy(bytes(p.view)), threshold=1.0, light_kd=0.0, esl=False)
            costs = []
            for i in range(world):
                _, cnt = port.render(sc, rows=(int(b[i]) * 4, min(h, int(b[i + 1]) * 4)))
                costs.append(cnt["samples"] + 16 * (min(h, int(b[i + 1]) * 4) - int(b[i]) * 4) * w)
            worst = max(worst, max(abs(c - np.mean(costs)) / np.mean(costs) for c in costs))
    assert worst < 0.12, worst


def test_balance_bands_is_a_pure_function_of_the_parameters(R):
    """Every rank must cut the same bands: same bytes in, same bounds out (and no hidden state between calls)."""
    p = _params(R, 1920, 1080, 0, (20.0, 123.0, 0.0))
    a = R.balance_bands(p, 8)
    R.balance_bands(_params(R, 640, 480, 1, (1.0, 2.0, 3.0)), 5)
    assert np.array_equal(a, R.balance_bands(p, 8))
    q = R.Params.from_buffer_copy(bytes(p))
    assert np.array_equal(a, R.balance_bands(q, 8))


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["synthA64_semi_persp", "synthB64_opaque_ortho", "bucky_ortho_-45_-45_0"])
def test_bands_cover_the_frame_exactly(name, golden, port, volumes, R):
    """The bands of all ranks, rendered in place into one frame (each rank touches only its rows), give the whole frame - the
    reference's frame - for per-frame calls and for sequences, whose bands move from frame to frame."""
    case = next(c for c in golden["cases"] if c["name"] == name)
    sc = golden_scene(port, case, volumes(case["volume"]))
    scene, params = product_scene(R, sc)
    w, h = case["width"], case["height"]
    persp = 1 if "persp" in name else 0
    gpu = R.GPURenderer5(0)
    try:
        gpu.set_window_buffer(w, h)
        gpu.set_scene(scene)
        whole = gpu.render_volume(params)
        assert __import__("common").sha(whole) == case["frame_sha256"]
        angles = [(20.0, 0.0, 0.0), (20.0, 72.0, 0.0), (-45.0, -45.0, 0.0), (33.0, 200.0, 0.0)]
        frames = [scene.params(R.camera(w, h, persp, a, 2.0)) for a in angles]
        want = [gpu.render_volume(p) for p in frames]
        for world in (2, 5):
            # per-frame calls
            dev = gpu.device_alloc(w * h * 4)
            bounds = R.balance_bands(params, world)
            for rank in range(world):
                gpu.set_partition_bands(rank, world)
                gpu.render_device(params, dev)
            got = np.zeros((h, w, 4), dtype=np.uint8)
            gpu.memcpy_d2h(got, dev)
            assert np.array_equal(got, whole), (name, world, bounds)
            # sequences: every rank's launch into the same buffers
            ptrs = [gpu.device_alloc(w * h * 4) for _ in frames]
            for rank in range(world):
                gpu.set_partition_bands(rank, world)
                gpu.render_sequence(frames, ptrs)
            gpu.sync()
            for i, ptr in enumerate(ptrs):
                gpu.memcpy_d2h(got, ptr)
                assert np.array_equal(got, want[i]), (name, world, i)
                gpu.device_free(ptr)
            gpu.device_free(dev)
            if world == 5 and persp:
                assert len({tuple(R.balance_bands(p, world)) for p in frames}) > 1   # the bands do move with the view
    finally:
        gpu.set_partition(0, 1, False)
        gpu.close()
