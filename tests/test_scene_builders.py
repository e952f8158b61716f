"""CPU tests of the product's host-side scene builders (vr5_scene.cpp, reached through the C ABI) against the restated
oracle and, where present, the reference's own managers (RaycasterBase / ViewBase)."""
import numpy as np
import pytest

import oracle
from common import POSES, sha


def test_tables_match_oracle(R, port, bucky):
    assert np.array_equal(R.default_base_tf(), port.default_base_tf())
    rng = np.random.default_rng(5)
    for vol in (bucky, rng.integers(0, 256, size=(40, 56, 72), dtype=np.uint8), port.synth_b(64, 2), np.zeros((9, 300, 17), np.uint8)):
        bd, mm, bs = R.build_minmax(vol)
        bd2, mm2, bs2 = port.minmax(vol)
        assert bd == bd2 and np.array_equal(mm, mm2) and tuple(np.float32(bs)) == tuple(bs2)
        for base in (R.default_base_tf(), R.tf_opaque(), R.tf_semi(max(vol.shape))):
            tf = R.premultiply_tf(base)
            assert np.array_equal(tf, port.premultiply(base))
            assert np.array_equal(R.build_esl_bits(tf, mm), port.esl_bits(tf, mm2))
        dz, dy, dx = vol.shape
        assert R.default_ray_step(dx, dy, dz) == port.ray_step(max(vol.shape))


def test_golden_tables(R, golden, bucky):
    case = next(c for c in golden["cases"] if c["name"] == "bucky_ortho_0_0_0")
    tf = R.premultiply_tf(R.default_base_tf())
    bd, mm, bs = R.build_minmax(bucky)
    assert sha(tf) == case["tf_sha256"] == golden["default_tf_premult_sha256"]
    assert sha(R.build_esl_bits(tf, mm)) == case["esl_bits_sha256"]
    assert bd == case["esl_block_dims"] and list(bs) == case["esl_block_size"]
    assert R.default_ray_step(32, 32, 32) == case["ray_step"] == 0.060546875


def test_tf_presets_agree_with_oracle_presets(R, port):
    base = port.default_base_tf()
    assert np.array_equal(R.tf_semi(512), oracle.tf_semi(base, 512))
    assert np.array_equal(R.tf_opaque(), oracle.tf_opaque(base))


def test_camera_matches_golden_views(R, golden):
    """The reference's own View bytes (ViewBase + GL-matrix shim) for the benchmark poses are reproduced."""
    for case in golden["cases"]:
        if not case["name"].startswith("bucky_"):
            continue
        _, proj, ax, ay, az = case["name"].split("_")
        v = R.camera(case["width"], case["height"], proj == "persp", (float(ax), float(ay), float(az)), 2.0)
        want = np.frombuffer(bytes.fromhex(case["view_hex"])[4:64], dtype=np.float32)
        got = np.frombuffer(bytes(v)[4:64], dtype=np.float32)
        assert np.abs(want - got).max() <= 1e-6
        assert bytes(v)[:4] == bytes.fromhex(case["view_hex"])[:4] and bytes(v)[64] == bytes.fromhex(case["view_hex"])[64]


def test_camera_matches_reference_live(R, ref):
    for persp in (0, 1):
        for a in POSES + [(20, 133, 7), (20, 271, 0)]:
            ref.camera(1920, 1080, persp, a, 2.0)
            rv = ref.raycaster().view
            v = R.camera(1920, 1080, persp, a, 2.0)
            A = np.frombuffer(bytes(rv)[4:64], dtype=np.float32)
            B = np.frombuffer(bytes(v)[4:64], dtype=np.float32)
            assert np.abs(A - B).max() <= 1e-6, (persp, a)


def test_camera_distance_clamp(R):
    far = R.camera(64, 64, 0, (0, 0, 0), 10.0)  # distance_limits (0.1, 3), ViewBase.cpp:17
    assert abs(far.origin.z - 3.0) < 1e-6
    near = R.camera(64, 64, 1, (0, 0, 0), 0.0)
    assert abs(near.origin.z - 0.1) < 1e-6
