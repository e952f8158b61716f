"""Trilinear sampler policy: GPURenderer4's semantics evaluated in software fp32.

The CPU side is oracle_port.c's trilinear sections. Their external anchors are elsewhere: the reference's GPURenderer4 itself,
compiled unmodified over a texture-object shim and compared under a stated tolerance (tests/test_trilinear_pin.py), and an
independent float64 restatement of the CUDA filtering formulas (tests/test_trilinear_independent.py). Here the CUDA policy
VR5_SAMPLER_TRILINEAR, which performs the same fp32 operations in the same order, is held to BIT equality with the restatement;
VR5_SAMPLER_TRILINEAR_FMA - the same policy with the contractions nvcc applies to the reference's GPU renderers written out as
fused multiply-adds - to bit equality with the checker's "trilinear-fma" section and to the stated LSB tolerance against the
one-rounding-per-operation policy."""
import os
import subprocess

import numpy as np
import pytest

import oracle
from common import POSES, assert_parity, golden_scene, product_scene, tf_by_name


def test_oracle_trilinear_reduces_to_nearest_on_constant_data(port, R):
    """Property pin of the restatement: with a constant volume every trilinear fetch is v/255 exactly, and with a constant
    transfer function classification cannot depend on the sampler - so both policies must give the same frame."""
    vol = np.full((24, 24, 24), 137, dtype=np.uint8)
    base = np.tile(np.array([0.3, 0.6, 0.9, 0.02], dtype=np.float32), (128, 1))
    ref_view = oracle.View.from_buffer_copy(bytes(R.camera(96, 64, 0, (10.0, 25.0, 0.0), 2.0)))  # host-side camera builder
    sc = port.make_scene(vol, base, ref_view, light_kd=0.0, esl=False)
    near, cn = port.render(sc)
    tri, ct = port.render(sc, trilinear=True)
    assert np.array_equal(near, tri) and cn == ct and cn["samples"] > 0


def test_oracle_trilinear_is_smooth_where_nearest_steps(port, bucky, golden):
    """The trilinear frame differs from the nearest one on real data, but by little: the same scene, same counters class."""
    case = next(c for c in golden["cases"] if c["name"] == "bucky_ortho_0_0_0")
    sc = golden_scene(port, case, bucky)
    near, cn = port.render(sc)
    tri, ct = port.render(sc, trilinear=True)
    assert cn["rays_hit"] == ct["rays_hit"] and cn["esl_iters"] == ct["esl_iters"]
    d = np.abs(near.astype(np.int16) - tri.astype(np.int16))
    assert 0 < d.max() and d.mean() < 12


def test_oracle_trilinear_fma_agrees_with_separate_roundings(port, bucky, golden):
    """The contracted restatement differs from the one-rounding-per-operation restatement only by rounding: same rays, same
    sample counts, frames within the stated tolerance (<= 1 LSB on >= 99.9 % of pixels, <= 2 LSB max), and not identical."""
    differing = 0
    for name in ["bucky_ortho_0_0_0", "bucky_persp_-45_-45_0"]:
        case = next(c for c in golden["cases"] if c["name"] == name)
        sc = golden_scene(port, case, bucky)
        tri, ct = port.render(sc, trilinear=True)
        fma, cf = port.render(sc, trilinear="fma")
        assert ct == cf and cf["samples"] > 0
        differing += assert_parity(fma, tri, exact=False, what=name)["differing"]
    assert differing > 0


def test_oracle_trilinear_fma_constant_volume(port, R):
    """Constant volume, constant transfer function: every fused lerp returns its operand exactly (w * 0 + a), so the frame
    is the closed-form blend acc <- fma(c, 1 - acc.a, acc) repeated `samples` times per ray."""
    vol = np.full((24, 24, 24), 137, dtype=np.uint8)
    rgba = np.array([0.3, 0.6, 0.9, 0.02], dtype=np.float32)
    base = np.tile(rgba, (128, 1))
    ref_view = oracle.View.from_buffer_copy(bytes(R.camera(96, 64, 0, (10.0, 25.0, 0.0), 2.0)))
    sc = port.make_scene(vol, base, ref_view, light_kd=0.0, esl=False)
    frame, cnt = port.render(sc, trilinear="fma")
    near, cn = port.render(sc)
    assert cnt["rays_accum"] == cn["rays_accum"] > 0
    from fractions import Fraction

    def round32(x: Fraction) -> np.float32:  # correctly rounded (nearest, ties to even) fp32 of an exact positive rational
        if x == 0:
            return np.float32(0.0)
        e = x.numerator.bit_length() - x.denominator.bit_length()
        if Fraction(2) ** e > x:
            e -= 1
        scaled = x / Fraction(2) ** (e - 23)
        n = scaled.numerator // scaled.denominator
        rem = scaled - n
        if rem > Fraction(1, 2) or (rem == Fraction(1, 2) and n % 2 == 1):
            n += 1
        return np.float32(float(Fraction(n) * Fraction(2) ** (e - 23)))

    tf0 = np.asarray(sc.tf, dtype=np.float32).reshape(128, 4)[64]
    acc = [np.float32(0.0)] * 4
    seen = set()
    for _ in range(200):  # more samples than any ray through a 24^3 grid takes
        t = np.float32(1.0) - acc[3]
        acc = [round32(Fraction(float(c)) * Fraction(float(t)) + Fraction(float(a))) for c, a in zip(tf0, acc)]
        seen.add(tuple(min(int(np.float32(a) * np.float32(256.0)), 255) for a in acc))
    lit = frame.reshape(-1, 4)
    lit = lit[lit[:, 3] > 0]
    assert len(lit) > 0 and all(tuple(int(v) for v in px) in seen for px in np.unique(lit, axis=0))


def test_division_by_255_identity(tmp_path):
    """vr5_tri.cuh computes sample / 255 as two FMAs around a multiplication by RN(1/255). The identity with the IEEE
    division holds for every float in [0, 255] (tools/check_div255.c without an argument checks all 1.13e9 of them); here a
    strided sample of them."""
    exe = tmp_path / "check_div255"
    src = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools", "check_div255.c")
    subprocess.run(["gcc", "-O2", "-mfma", "-ffp-contract=off", src, "-o", str(exe), "-lm"], check=True)
    out = subprocess.run([str(exe), "997"], check=True, capture_output=True, text=True).stdout
    assert "0 mismatches" in out, out


POLICIES = [("trilinear", True, 16), ("trilinear-fma", "fma", 18)]  # sampler name, checker mode, base variant id


@pytest.mark.gpu
@pytest.mark.parametrize("policy", POLICIES, ids=[p[0] for p in POLICIES])
@pytest.mark.parametrize("clamped", [False, True])
@pytest.mark.parametrize("name", ["bucky_ortho_0_0_0", "bucky_persp_-45_-45_0", "synthA64_default_persp"])
def test_gpu_trilinear_matches_restated_oracle(name, clamped, policy, golden, port, volumes, R, monkeypatch):
    """Both marches - unclamped 32-bit record addressing (cubic volumes whose step stays inside the guard cells; variant 17)
    and the clamped one (variant 16, forced here through its knob) - against the restatement, bit for bit."""
    if clamped:
        monkeypatch.setenv("VR5_TRI_CLAMPED", "1")
    case = next(c for c in golden["cases"] if c["name"] == name)
    sc = golden_scene(port, case, volumes(case["volume"]))
    scene, params = product_scene(R, sc)
    sampler, mode, base_variant = policy
    gpu = R.GPURenderer5(0, sampler=sampler)
    try:
        gpu.set_window_buffer(case["width"], case["height"])
        gpu.set_scene(scene)
        frame = gpu.render_volume(params)
        want, cnt = port.render(sc, trilinear=mode)
        assert_parity(frame, want, exact=True, what=name)
        assert gpu.last_variant() == base_variant + (0 if clamped else 1)
        assert gpu.count(params) == cnt
    finally:
        gpu.close()


@pytest.mark.gpu
@pytest.mark.parametrize("warp_fetch", [False, True])
@pytest.mark.parametrize("ctas,depth", [("6", "1"), ("5", "1"), ("4", "1"), ("6", "2"), ("5", "2"), ("4", "2"), ("4", "3"), ("4", "4")])
def test_gpu_trilinear_fma_budget_and_depth_variants(ctas, depth, warp_fetch, golden, port, volumes, R, monkeypatch):
    """The contracted unclamped march is built for three register budgets (VR5_TRIF_CTAS = 6, 5, 4 CTAs per SM) and one to four
    cell requests in flight per ray (VR5_TRIF_DEPTH), with the tile queue read per CTA or per warp: same bits every way."""
    monkeypatch.setenv("VR5_TRIF_CTAS", ctas)
    monkeypatch.setenv("VR5_TRIF_DEPTH", depth)
    if warp_fetch:
        monkeypatch.setenv("VR5_WARP_FETCH", "1")
    for name in ["synthA64_semi_ortho", "synthA64_semi_persp"]:
        case = next(c for c in golden["cases"] if c["name"] == name)
        sc = golden_scene(port, case, volumes(case["volume"]))
        scene, params = product_scene(R, sc)
        gpu = R.GPURenderer5(0, sampler="trilinear-fma")
        try:
            gpu.set_window_buffer(case["width"], case["height"])
            gpu.set_scene(scene)
            frame = gpu.render_volume(params)
            want, _ = port.render(sc, trilinear="fma")
            assert_parity(frame, want, exact=True, what=(name, ctas, depth))
            assert gpu.last_variant() == 19
        finally:
            gpu.close()


@pytest.mark.gpu
@pytest.mark.parametrize("policy", POLICIES, ids=[p[0] for p in POLICIES])
def test_gpu_trilinear_ragged_random(policy, port, R):
    """Non-cubic random volume (worst case for interpolation: neighbours uncorrelated), ragged viewports, every TF preset,
    both projections, shading on and off, partition strips."""
    rng = np.random.default_rng(5)
    vol = rng.integers(0, 256, size=(29, 41, 66), dtype=np.uint8)
    vol[:, :13] = 0
    gpu = R.GPURenderer5(0, sampler=policy[0])
    try:
        for tf_name, kw in [("default", {}), ("opaque", {}), ("semi", {"ray_threshold": 1.0, "light_kd": 0.0}), ("default", {"esl": False})]:
            base = {"default": R.default_base_tf(), "opaque": R.tf_opaque(), "semi": R.tf_semi(66)}[tf_name]
            bd, mm, bs = R.build_minmax(vol)
            tf = R.premultiply_tf(base)
            scene = R.Scene(vol, tf, R.build_esl_bits(tf, mm), bd, bs, R.default_ray_step(66, 41, 29), **kw)
            gpu.set_scene(scene)
            for persp, ang, wh in [(0, POSES[1], (163, 97)), (1, (20, 130, 0), (129, 131)), (1, POSES[3], (8, 4))]:
                view = R.camera(wh[0], wh[1], persp, ang, 2.0)
                gpu.set_window_buffer(*wh)
                params = scene.params(view)
                frame = gpu.render_volume(params)
                osc = port.make_scene(vol, tf_by_name(port, tf_name, 66), oracle.View.from_buffer_copy(bytes(view)),
                                      threshold=scene.ray_threshold, esl=scene.esl, light_kd=scene.light_kd)
                want, cnt = port.render(osc, trilinear=policy[1])
                assert_parity(frame, want, exact=True, what=(tf_name, persp, ang))
                assert gpu.count(params) == cnt
    finally:
        gpu.close()
