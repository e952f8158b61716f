"""Shared helpers of the parity tests."""
import hashlib

import numpy as np

import oracle

# BASELINE.json north_star tolerance (stated here once, used by every GPU parity test):
#   |new - oracle| <= 1 LSB per RGBA8 channel on >= 99.9 % of pixels, and <= 2 LSB everywhere.
# The design goal of the kernel is 0 LSB (bit-exact); tests that assert exact equality say so.
TOL_FRACTION_WITHIN_1 = 0.999
TOL_MAX_LSB = 2

POSES = [(0, 0, 0), (-45, -45, 0), (90, 0, 0), (180, 90, 0)]  # reference VolR.cpp:235-244


def sha(a) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def tf_by_name(port, name, n):
    base = port.default_base_tf()
    if name == "default":
        return base
    if name == "semi":
        return oracle.tf_semi(base, n)
    if name == "opaque":
        return oracle.tf_opaque(base)
    raise KeyError(name)


def golden_scene(port, case, volume) -> oracle.Scene:
    """oracle.Scene for a golden case: the reference's exact View bytes + restated builders for the tables."""
    view = oracle.View.from_buffer_copy(bytes.fromhex(case["view_hex"]))
    n = max(volume.shape)
    sc = port.make_scene(volume, tf_by_name(port, case["tf"], n), view, ray_step=case["ray_step"],
                         threshold=case["ray_threshold"], esl=case["esl"], light_kd=case["light_kd"])
    return sc


def parity_report(new: np.ndarray, old: np.ndarray) -> dict:
    d = np.abs(new.astype(np.int16) - old.astype(np.int16)).reshape(-1, 4).max(axis=1)
    return {"pixels": int(d.size), "exact": float((d == 0).mean()), "within1": float((d <= 1).mean()),
            "max": int(d.max()) if d.size else 0, "differing": int((d > 0).sum())}


def assert_parity(new, old, exact=True, what=""):
    rep = parity_report(new, old)
    assert rep["within1"] >= TOL_FRACTION_WITHIN_1 and rep["max"] <= TOL_MAX_LSB, (what, rep)
    if exact:
        assert rep["differing"] == 0, (what, rep)
    return rep


def product_scene(R, sc: oracle.Scene):
    """The same scene expressed through the product's own Scene/Params types (tables taken from the oracle scene
    so that the kernel and the oracle consume identical bytes)."""
    rc = sc.rc
    scene = R.Scene(sc.volume, sc.tf, sc.esl_bits, rc.esl_block_dims, rc.esl_block_size.tup(), rc.ray_step,
                    rc.ray_threshold, bool(rc.esl), rc.light_kd)
    view = R.View.from_buffer_copy(bytes(rc.view))
    return scene, scene.params(view)
