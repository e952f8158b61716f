"""CPU tests: the restated oracle (oracle/oracle_port.c) against the golden vectors the REFERENCE produced
(tests/golden/golden.json, made by tests/golden/make_golden.py), and - where the compiled reference is present -
against the reference itself, live. This is what "parity pinned" rests on."""
import numpy as np
import pytest

import oracle
from common import POSES, golden_scene, sha, tf_by_name


def test_golden_inputs(golden, bucky, port):
    # Bucky decodes to 32^3, byte sum 1 174 726
    assert list(bucky.shape) == golden["bucky"]["shape"] and int(bucky.sum()) == golden["bucky"]["sum"] == 1174726
    assert sha(bucky) == golden["bucky"]["sha256"]
    assert sha(port.premultiply(port.default_base_tf())) == golden["default_tf_premult_sha256"]
    assert sha(port.cos_table()) == golden["synth"]["cos_table"]
    assert sha(port.synth_a(64, 1)) == golden["synth"]["A64"]
    assert sha(port.synth_b(64, 2)) == golden["synth"]["B64"]
    assert sha(port.synth_a(96, 1)) == golden["synth"]["A96"]


def test_default_tf_known_answers(port):
    # SURVEY 8(c) known answers of the reference's default pre-multiplied TF
    tf = port.premultiply(port.default_base_tf())
    assert tf[12, 3] == 0.0
    assert abs(tf[13, 3] - 0.1015625) < 1e-9
    assert tuple(tf[16]) == (0.046875, 0.0, 0.0, 0.125)
    assert tuple(tf[64]) == (0.0, 0.2578125, 0.0, 0.5)
    assert np.allclose(tf[127], (0, 0, 0.99993896, 0.9921875), atol=1e-8)
    assert port.ray_step(32) == 0.060546875


def _case_ids(golden_path="tests/golden/golden.json"):
    import json, os
    with open(os.path.join(os.path.dirname(__file__), "golden", "golden.json")) as f:
        return [c["name"] for c in json.load(f)["cases"]]


@pytest.mark.parametrize("name", _case_ids())
def test_port_matches_reference_golden(name, golden, port, volumes):
    case = next(c for c in golden["cases"] if c["name"] == name)
    vol = volumes(case["volume"])
    sc = golden_scene(port, case, vol)
    # the restated scene builders must reproduce the reference's tables and parameters
    assert sha(sc.tf) == case["tf_sha256"]
    assert sha(sc.esl_bits) == case["esl_bits_sha256"]
    assert int(sc.rc.esl_block_dims) == case["esl_block_dims"]
    assert list(sc.rc.esl_block_size.tup()) == case["esl_block_size"]
    frame, cnt = port.render(sc)
    assert sha(frame) == case["frame_sha256"], "restated oracle differs from the reference frame"
    assert int((frame.reshape(-1, 4).max(1) > 0).sum()) == case["nonzero_pixels"]
    assert cnt == case["counters"]


def test_survey_sample_counts(golden):
    # accumulation samples per frame measured on the as-shipped reference
    want = {"bucky_ortho_0_0_0": 8175245, "bucky_ortho_-45_-45_0": 6917693, "bucky_ortho_90_0_0": 8122231,
            "bucky_ortho_180_90_0": 8072700, "bucky_persp_0_0_0": 4768145, "bucky_persp_-45_-45_0": 5389506,
            "bucky_persp_90_0_0": 4806025, "bucky_persp_180_90_0": 4792244}
    got = {c["name"]: c["counters"]["samples"] for c in golden["cases"]}
    for k, v in want.items():
        assert got[k] == v
    assert next(c for c in golden["cases"] if c["name"] == "bucky_ortho_0_0_0")["esl_popcount"] == 32704


def test_row_ranges_compose(port, golden, volumes):
    """orc_render_rows over disjoint row ranges == whole frame (the multi-threaded / sampled baseline relies on it)."""
    case = next(c for c in golden["cases"] if c["name"] == "synthB64_opaque_ortho")
    sc = golden_scene(port, case, volumes(case["volume"]))
    whole, cnt = port.render(sc, threads=1)
    a, ca = port.render(sc, rows=(0, 100), threads=3)
    b, cb = port.render(sc, rows=(100, sc.height), threads=2)
    assert np.array_equal(whole[:100], a[:100]) and np.array_equal(whole[100:], b[100:])
    assert not a[100:].any() and not b[:100].any()
    assert {k: ca[k] + cb[k] for k in cnt} == cnt


def test_missing_inputs_return_1(port):
    rc = oracle.Raycaster()
    out = np.zeros((4, 4, 4), dtype=np.uint8)
    import ctypes as C
    assert port.lib.orc_render_rows(C.byref(rc), out.ctypes.data_as(C.POINTER(C.c_uint8)), 0, 4, None) == 1


# ---- live pin against the compiled reference (only where oracle/_ref exists) ------------------------------------------
def test_port_equals_reference_live(ref, port):
    rng = np.random.default_rng(7)
    vol = rng.integers(0, 256, size=(24, 40, 56), dtype=np.uint8)  # non-cubic, ragged
    vol[:, :14] = 0
    vol[6:18] = 0
    base = port.default_base_tf()
    for tf_name, opts in [("default", {}), ("opaque", {"esl": 0}), ("semi", {"thr": 1.0, "kd": 0.0})]:
        tfb = tf_by_name(port, tf_name, 56)
        ref.set_volume(vol)
        ref.set_base_tf(tfb)
        ref.lib.ref_set_esl(opts.get("esl", 1))
        ref.lib.ref_set_ray_threshold(opts.get("thr", 0.95))
        ref.lib.ref_set_light_kd(opts.get("kd", 0.6))
        for persp, a, wh in [(0, POSES[1], (161, 97)), (1, (20, 130, 0), (128, 130)), (0, POSES[0], (64, 64))]:
            ref.camera(wh[0], wh[1], persp, a, 2.0)
            sc = ref.scene()
            want = ref.render(sc.rc)
            got, _ = port.render(sc)
            assert np.array_equal(want, got), (tf_name, persp, a)
            # restated builders == the reference's managers
            sc2 = port.make_scene(vol, tfb, sc.rc.view, threshold=opts.get("thr", 0.95), esl=opts.get("esl", 1), light_kd=opts.get("kd", 0.6))
            assert np.array_equal(sc2.tf, sc.tf) and np.array_equal(sc2.esl_bits, sc.esl_bits)
            assert sc2.rc.ray_step == sc.rc.ray_step and sc2.rc.esl_block_dims == sc.rc.esl_block_dims
    ref.lib.ref_set_esl(1); ref.lib.ref_set_ray_threshold(0.95); ref.lib.ref_set_light_kd(0.6)


def test_reference_rows_equal_reference_frame(ref, port, bucky):
    ref.set_volume(bucky)
    ref.reset_transfer_fn()
    ref.camera(200, 150, 1, POSES[1], 2.0)
    rc = ref.raycaster()
    whole = ref.render(rc)
    part = np.zeros_like(whole)
    ref.render_rows(rc, part, 0, 70)
    ref.render_rows(rc, part, 70, 150)
    assert np.array_equal(whole, part)


def test_histogram_matches_reference_formula(port, bucky):
    h = port.histogram(bucky)
    bins = np.bincount(bucky.ravel(), minlength=256).astype(np.float32)
    want = np.sqrt(np.sqrt(bins)).astype(np.float32)
    want = want / want.max()
    assert np.allclose(h, want, rtol=0, atol=1e-7) and h.max() == 1.0
