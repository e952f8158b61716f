"""BASELINE.json's full-size configurations on the GPU: oracle parity on a stated subset of rows (a whole frame costs the
CPU oracle minutes to hours) plus size-independent properties - every memory path gives the same bytes,
strips of a partition reassemble to the whole frame, work counters obey the transfer function's invariants."""
import os

import numpy as np
import pytest

import oracle
from common import assert_parity, sha

pytestmark = pytest.mark.gpu


def _mem_available_gib():
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable:"):
                return int(line.split()[1]) / 2 ** 20
    except OSError:
        pass
    return 0.0


def _rows_subset(h):
    """48 rows: top, bottom, centre and two quarter bands (centre rows carry the longest rays)."""
    bands = [0, h // 4 - 4, h // 2 - 8, h // 2, 3 * h // 4, h - 8]
    return [(b, b + 8) for b in bands]


def _setup(R, gpu, kind, n, seed, tf_name, w, h):
    gpu.synth_volume(kind, n, seed)
    bd, mm, bs = gpu.build_minmax_device()
    base = {"semi": lambda: R.tf_semi(n), "opaque": R.tf_opaque, "default": R.default_base_tf}[tf_name]()
    tf = R.premultiply_tf(base)
    scene = R.Scene(None, tf, R.build_esl_bits(tf, mm), bd, bs, R.default_ray_step(n, n, n))
    gpu.set_transfer_fn(scene.tf, scene.esl_bits)
    gpu.set_window_buffer(w, h)
    return scene, base


def _oracle_rows(port, vol, base, view, spans):
    osc = port.make_scene(vol, base, oracle.View.from_buffer_copy(bytes(view)))
    out = np.zeros((view.dims[1], view.dims[0], 4), dtype=np.uint8)
    for s in spans:
        part, _ = port.render(osc, rows=s)
        out[s[0]:s[1]] = part[s[0]:s[1]]
    return out


@pytest.mark.parametrize("cfg", ["c3", "c4"])
def test_full_size_parity_and_properties(cfg, R, port):
    kind, n, seed, tf_name, w, h = {"c3": (0, 512, 1, "semi", 1920, 1080), "c4": (1, 1024, 2, "opaque", 3840, 2160)}[cfg]
    gpu = R.GPURenderer5(0)
    try:
        scene, base = _setup(R, gpu, kind, n, seed, tf_name, w, h)
        vol = gpu.get_volume((n, n, n))
        for persp, ang in ((0, (20.0, 37.0, 0.0)), (1, (-45.0, -45.0, 0.0))):
            view = R.camera(w, h, persp, ang, 2.0)
            params = scene.params(view)
            frames = {}
            for s in ("hybridw", "tex3d", "linear", "window"):
                gpu.set_sampler(s)
                frames[s] = gpu.render_volume(params)
            assert len({sha(f) for f in frames.values()}) == 1               # every memory path, same bytes
            assert sha(gpu.render_volume(params)) == sha(frames["window"])   # idempotent
            spans = _rows_subset(h)
            want = _oracle_rows(port, vol, base, view, spans)
            for a, b in spans:
                assert_parity(frames["hybridw"][a:b], want[a:b], exact=True, what=(cfg, persp, a))
            cnt = gpu.count(params)
            assert cnt["rays_accum"] <= cnt["rays_hit"] <= w * h
            if tf_name == "semi":   # alpha 1/N: never shaded, never terminated early, no empty block in synthA
                assert cnt["shaded"] == 0 and cnt["esl_iters"] == 0 and cnt["rays_accum"] == cnt["rays_hit"]
                lit = int((frames["hybridw"][..., 3] > 0).sum())    # a ray needs n/256 samples before its alpha byte leaves 0
                assert 0.99 * cnt["rays_hit"] <= lit <= cnt["rays_hit"]
            else:                   # opaque: alpha is 0 until the first opaque sample ends the ray at 1; synthB is mostly empty space
                assert cnt["esl_iters"] > 0 and cnt["shaded"] <= cnt["rays_accum"]
                assert set(np.unique(frames["hybridw"][..., 3]).tolist()) <= {0, 255}
                assert cnt["samples"] <= cnt["rays_accum"] * (2 * 3 ** 0.5 / scene.ray_step + 2)
            # strips of a 3-rank partition reassemble to the frame
            gpu.set_sampler("hybridw")
            acc = np.zeros_like(frames["hybridw"])
            for rank in range(3):
                gpu.set_partition(rank, 3, False)
                part = gpu.render_volume(params, out=np.zeros_like(acc))
                own = (np.arange(h) // 4) % 3 == rank
                acc[own] = part[own]
            gpu.set_partition(0, 1, False)
            assert sha(acc) == sha(frames["hybridw"])
    finally:
        gpu.close()


def test_c5_2048_cubed_properties_and_oracle_rows(R, port, record_property):
    """synthA 2048^3 (8 GiB, 64-bit voxel offsets: SURVEY H3) at 7680x4320. The window array is larger than 4 GiB: the default
    policy runs as kernel variant 10 (64-bit slice term, vr5_cast.cuh kP2Wide), compared with the texture-only fallback
    (VR5_P2_WIDE=-1) and with the plain linear array. Oracle rows only where the host has room for the 8 GiB copy."""
    n, w, h = 2048, 7680, 4320
    gpu = R.GPURenderer5(0)
    try:
        try:
            scene, base = _setup(R, gpu, 0, n, 1, "semi", w, h)
        except R.Vr5Error as e:
            pytest.skip("no room for the 8 GiB volume: %s" % e)
        view = R.camera(w, h, 0, (20.0, 37.0, 0.0), 2.0)
        params = scene.params(view)
        frames = {}
        for s in ("hybridw", "linear"):
            gpu.set_sampler(s)
            frames[s] = gpu.render_volume(params)
            if s == "hybridw":
                assert gpu.last_variant() == 10, gpu.last_variant()
        assert sha(frames["hybridw"]) == sha(frames["linear"])
        gpu.set_option("VR5_P2_WIDE", "-1")
        gpu.set_sampler("hybridw")
        assert sha(gpu.render_volume(params)) == sha(frames["linear"]) and gpu.last_variant() == 2
        gpu.set_option("VR5_P2_WIDE", "0")
        cnt = gpu.count(params)
        lit = int((frames["linear"][..., 3] > 0).sum())
        assert cnt["shaded"] == 0 and cnt["esl_iters"] == 0 and 0.99 * cnt["rays_hit"] <= lit <= cnt["rays_hit"]
        assert cnt["samples"] > 20e9                                         # about 35 G samples per frame
        # which branch ran is part of the record (junit property + a printed line)
        oracle_branch = _mem_available_gib() > 40 and os.environ.get("VR5_C5_ORACLE", "1") != "0"
        record_property("c5_oracle_rows_checked", 6 if oracle_branch else 0)
        record_property("c5_mem_available_gib", round(_mem_available_gib(), 1))
        print("C5 ORACLE ROWS: %s (MemAvailable %.1f GiB)" % ("6 rows compared with the restated CPU checker, 0 LSB required" if oracle_branch else "SKIPPED - properties only", _mem_available_gib()))
        if oracle_branch:
            vol = gpu.get_volume((n, n, n))
            spans = [(0, 2), (h // 2, h // 2 + 2), (h - 2, h)]
            want = _oracle_rows(port, vol, base, view, spans)
            for a, b in spans:
                assert_parity(frames["linear"][a:b], want[a:b], exact=True, what=("c5", a))
    finally:
        gpu.close()
