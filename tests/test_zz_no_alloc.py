"""vr5_render_sequence_noalloc (include/vr5.h): the frame-sequence launch for callers that cannot allocate - its per-frame
records live in a file-scope `static __device__` array of VR5_NOALLOC_MAX_FRAMES entries instead of a block allocated per call.
No counterpart in the reference. Collected last on purpose.
Called through the C ABI (ctypes) and compared with the existing entry point vr5_render_sequence on TWO calls in a
row (the second overwrites the first call's records in the shared array); pixels must be bit-equal (0 LSB, the tolerance of the
nearest-path tests). Passes alone in a fresh process and after the other tests: it uses its own renderer handle."""
import ctypes as C

import numpy as np
import pytest

ANGLES_A = [(20.0, 0.0, 0.0), (20.0, 72.0, 0.0), (-45.0, -45.0, 0.0), (90.0, 10.0, 5.0), (180.0, 90.0, 0.0)]
ANGLES_B = [(33.0, 200.0, 0.0), (5.0, 300.0, 0.0), (61.0, 17.0, 3.0)]


def _render(R, gpu, fn, frames, w, h):
    """frames through entry point `fn` (same signature as vr5_render_sequence) into fresh device buffers -> host arrays"""
    L = R.lib()
    ptrs = [gpu.device_alloc(w * h * 4) for _ in frames]
    try:
        arr = (R.Params * len(frames))(*frames)
        outs = (C.c_void_p * len(frames))(*ptrs)
        rc = fn(gpu._h, arr, len(frames), outs)
        assert rc == 0, gpu.last_error()
        gpu.sync()
        out = []
        for p in ptrs:
            a = np.zeros((h, w, 4), dtype=np.uint8)
            gpu.memcpy_d2h(a, p)
            out.append(a)
        return out
    finally:
        for p in ptrs:
            gpu.device_free(p)


@pytest.mark.gpu
@pytest.mark.parametrize("sampler", ["hybridw", "trilinear-fma"])
def test_no_alloc_sequence_equals_the_allocating_entry_point(sampler, R):
    L = R.lib()
    n, w, h = 64, 317, 203   # ragged tiles
    gpu = R.GPURenderer5(0, sampler=sampler)
    try:
        gpu.synth_volume(0, n, 1)
        bd, mm, bs = gpu.build_minmax_device()
        tf = R.premultiply_tf(R.default_base_tf())
        scene = R.Scene(None, tf, R.build_esl_bits(tf, mm), bd, bs, R.default_ray_step(n, n, n))
        gpu.set_transfer_fn(scene.tf, scene.esl_bits)
        gpu.set_window_buffer(w, h)
        first = [scene.params(R.camera(w, h, 1, a, 2.0)) for a in ANGLES_A]
        second = [scene.params(R.camera(w, h, 1, a, 2.0)) for a in ANGLES_B]
        want_first = _render(R, gpu, L.vr5_render_sequence, first, w, h)
        want_second = _render(R, gpu, L.vr5_render_sequence, second, w, h)
        assert len({a.tobytes() for a in want_first + want_second}) == len(first) + len(second) and want_first[0].max() > 0
        # two calls in a row through the new entry point: the second reuses (and overwrites) the shared device array
        got_first = _render(R, gpu, L.vr5_render_sequence_noalloc, first, w, h)
        got_second = _render(R, gpu, L.vr5_render_sequence_noalloc, second, w, h)
        for a, b in zip(got_first + got_second, want_first + want_second):
            assert int(np.abs(a.astype(np.int16) - b.astype(np.int16)).max()) == 0
        # ... and back to back without a host synchronisation in between (stream-ordered on one handle: no race)
        ptrs = [gpu.device_alloc(w * h * 4) for _ in range(len(first) + len(second))]
        for frames, base in ((first, 0), (second, len(first))):
            arr = (R.Params * len(frames))(*frames)
            outs = (C.c_void_p * len(frames))(*ptrs[base:base + len(frames)])
            assert L.vr5_render_sequence_noalloc(gpu._h, arr, len(frames), outs) == 0, gpu.last_error()
        gpu.sync()
        for p, b in zip(ptrs, want_first + want_second):
            a = np.zeros((h, w, 4), dtype=np.uint8)
            gpu.memcpy_d2h(a, p)
            assert np.array_equal(a, b)
            gpu.device_free(p)
    finally:
        gpu.close()


@pytest.mark.gpu
def test_no_alloc_sequence_limits(R):
    """More frames than the array holds, or a tile order that needs a device table, are refused with a clear message."""
    L = R.lib()
    n, w, h = 32, 64, 48
    gpu = R.GPURenderer5(0)
    try:
        gpu.synth_volume(0, n, 1)
        bd, mm, bs = gpu.build_minmax_device()
        tf = R.premultiply_tf(R.tf_semi(n))
        scene = R.Scene(None, tf, R.build_esl_bits(tf, mm), bd, bs, R.default_ray_step(n, n, n))
        gpu.set_transfer_fn(scene.tf, scene.esl_bits)
        gpu.set_window_buffer(w, h)
        p = scene.params(R.camera(w, h, 0, (20.0, 30.0, 0.0), 2.0))
        dev = gpu.device_alloc(w * h * 4)
        limit = 256
        arr = (R.Params * (limit + 1))(*([p] * (limit + 1)))
        outs = (C.c_void_p * (limit + 1))(*([dev] * (limit + 1)))
        assert L.vr5_render_sequence_noalloc(gpu._h, arr, limit + 1, outs) == 1 and "VR5_NOALLOC_MAX_FRAMES" in gpu.last_error()
        assert L.vr5_render_sequence_noalloc(gpu._h, arr, limit, outs) == 0, gpu.last_error()   # the maximum itself is fine
        gpu.sync()
        want = gpu.render_volume(p)
        got = np.zeros((h, w, 4), dtype=np.uint8)
        gpu.memcpy_d2h(got, dev)
        assert np.array_equal(got, want)
        gpu.set_option("VR5_QUEUE_ORDER", "rings")
        assert L.vr5_render_sequence_noalloc(gpu._h, arr, 2, outs) == 1 and "VR5_QUEUE_ORDER" in gpu.last_error()
        gpu.set_option("VR5_QUEUE_ORDER", "rows")
        gpu.device_free(dev)
    finally:
        gpu.close()
