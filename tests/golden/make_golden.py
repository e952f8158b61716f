#!/usr/bin/env python
"""Generates tests/golden/* by RUNNING THE REFERENCE ITSELF (oracle/_ref/libvolr_ref.so = the reference's own
CPURenderer / RaycasterBase / ViewBase / ModelBase / ddsbase compiled unmodified from /root/reference).

Run in the build container only (needs /root/reference):   python tests/golden/make_golden.py

Outputs (small, committed):
  bucky32.u8.gz   Bucky.pvm decoded by the reference's ddsbase (32^3 uint8) - the config-1 dataset
  golden.json     for every case: the exact View bytes the reference produced (hex), march parameters, the sha256 of
                  the reference frame, its non-zero pixel count, and the reference's pre-multiplied TF / ESL bitfield
                  hashes. Work counters come from the restated oracle (the reference does not count) and are only
                  recorded after its frame matched the reference frame bit for bit.
"""
import gzip
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle  # noqa: E402

POSES = [(0, 0, 0), (-45, -45, 0), (90, 0, 0), (180, 90, 0)]  # VolR.cpp:235-244


def sha(a) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    ref, port = oracle.Ref(), oracle.Port()
    cases = []

    def add_case(name, volume_spec, volume, base_tf, tf_name, w, h, persp, angles, esl=1, thr=0.95, kd=0.6, step_scale=1.0):
        ref.set_volume(volume)
        ref.set_base_tf(base_tf)
        ref.lib.ref_reset_ray_step()
        rc0 = ref.raycaster()
        ref.lib.ref_set_ray_step(np.float32(rc0.ray_step) * np.float32(step_scale))
        ref.lib.ref_set_esl(esl)
        ref.lib.ref_set_ray_threshold(thr)
        ref.lib.ref_set_light_kd(kd)
        ref.camera(w, h, persp, angles, 2.0)
        sc = ref.scene()
        frame = ref.render(sc.rc)
        mine, cnt = port.render(sc)
        assert np.array_equal(frame, mine), name + ": restated oracle differs from the reference"
        cases.append({
            "name": name, "volume": volume_spec, "tf": tf_name, "width": w, "height": h,
            "view_hex": bytes(sc.rc.view).hex(), "ray_step": float(sc.rc.ray_step), "ray_threshold": float(sc.rc.ray_threshold),
            "esl": int(sc.rc.esl), "esl_block_dims": int(sc.rc.esl_block_dims), "esl_block_size": list(sc.rc.esl_block_size.tup()),
            "light_kd": float(sc.rc.light_kd), "tf_sha256": sha(sc.tf), "esl_bits_sha256": sha(sc.esl_bits),
            "esl_popcount": int(sum(bin(int(x)).count("1") for x in sc.esl_bits)),
            "frame_sha256": sha(frame), "nonzero_pixels": int((frame.reshape(-1, 4).max(1) > 0).sum()),
            "counters": cnt,
        })
        print(name, cnt, flush=True)

    # config 1: Bucky.pvm through the reference's own loader, default TF, 800x600, 4 poses x 2 projections
    assert ref.load_model("/root/reference/VolumeRendering/Bucky.pvm") == 0
    bucky = ref.volume()
    with gzip.GzipFile(os.path.join(HERE, "bucky32.u8.gz"), "wb", mtime=0) as f:
        f.write(bucky.tobytes())
    base = port.default_base_tf()
    ref.reset_transfer_fn()
    assert np.array_equal(ref.base_tf(), base)
    for persp in (0, 1):
        for a in POSES:
            add_case("bucky_%s_%d_%d_%d" % ("persp" if persp else "ortho", *a), {"kind": "bucky"}, bucky, base, "default", 800, 600, persp, a)

    # synthetic volumes (integer-defined generators, oracle/synth_volume.c): exercise ESL, ERT, shading, ragged sizes
    a64, b64 = port.synth_a(64, 1), port.synth_b(64, 2)
    a96 = port.synth_a(96, 1)
    semi64 = oracle.tf_semi(base, 64)
    opaque = oracle.tf_opaque(base)
    add_case("synthA64_default_ortho", {"kind": "synthA", "n": 64, "seed": 1}, a64, base, "default", 320, 240, 0, (-45, -45, 0))
    add_case("synthA64_default_persp", {"kind": "synthA", "n": 64, "seed": 1}, a64, base, "default", 317, 203, 1, (20, 130, 0))
    add_case("synthA64_semi_ortho", {"kind": "synthA", "n": 64, "seed": 1}, a64, semi64, "semi", 320, 240, 0, (20, 40, 0))
    add_case("synthA64_semi_persp", {"kind": "synthA", "n": 64, "seed": 1}, a64, semi64, "semi", 322, 241, 1, (20, 220, 0))
    add_case("synthB64_opaque_ortho", {"kind": "synthB", "n": 64, "seed": 2}, b64, opaque, "opaque", 320, 240, 0, (-45, -45, 0))
    add_case("synthB64_opaque_persp", {"kind": "synthB", "n": 64, "seed": 2}, b64, opaque, "opaque", 320, 240, 1, (90, 0, 0))
    add_case("synthB64_default_noesl", {"kind": "synthB", "n": 64, "seed": 2}, b64, base, "default", 256, 256, 0, (180, 90, 0), esl=0)
    add_case("synthB64_default_thr1_kd0", {"kind": "synthB", "n": 64, "seed": 2}, b64, base, "default", 256, 256, 1, (0, 0, 0), thr=1.0, kd=0.0)
    add_case("synthA96_default_step1.7", {"kind": "synthA", "n": 96, "seed": 1}, a96, base, "default", 299, 301, 0, (33, 77, 5), step_scale=1.7)
    add_case("synthA96_default_axis_aligned", {"kind": "synthA", "n": 96, "seed": 1}, a96, base, "default", 128, 128, 0, (0, 0, 0))

    meta = {
        "generator": "tests/golden/make_golden.py", "source": "reference CPURenderer via oracle/_ref/libvolr_ref.so",
        "bucky": {"sha256": sha(bucky), "sum": int(bucky.sum()), "shape": list(bucky.shape)},
        "default_tf_premult_sha256": sha(port.premultiply(base)),
        "synth": {"A64": sha(a64), "B64": sha(b64), "A96": sha(a96), "cos_table": sha(port.cos_table())},
        "cases": cases,
    }
    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(meta, f, indent=1)
    print("wrote", len(cases), "cases")


if __name__ == "__main__":
    main()
