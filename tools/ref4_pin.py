#!/usr/bin/env python
"""Parity pin of the trilinear sampler policies against the reference's GPURenderer4 (run on the GPU box).

oracle/_ref/libvolr_ref_gpu.so holds the reference's GPURenderer4.cu compiled UNMODIFIED for sm_100a (its legacy texture
references mapped onto texture objects by oracle/ref_gpu4_texobj.cu): hardware-filtered tex3D on the uint8 volume
(normalised coordinates, cudaFilterModeLinear, cudaReadModeNormalizedFloat, clamp), hardware-filtered tex1D on the
128-entry float4 transfer function, nvcc's default FMA contraction. This tool renders the same frames with it and with
VR5_SAMPLER_TRILINEAR / VR5_SAMPLER_TRILINEAR_FMA (software fp32 filtering) and prints ONE JSON line with, per case, the
histogram of the per-pixel maximum channel difference in LSB - the measured basis of the tolerance that
tests/test_trilinear_pin.py states. The texture unit filters with 8 fractional weight bits, so the frames are NOT expected
to be bit-equal: the histogram says how far apart they are.

    python tools/ref4_pin.py [--full] > profiles/r02_ref4_pin.json
"""
import argparse
import gzip
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

BINS = [0, 1, 2, 3, 4, 8, 16, 32, 256]  # histogram edges: [0], [1], [2], [3], [4..7], [8..15], [16..31], [32..255]


def lsb_report(a: np.ndarray, b: np.ndarray) -> dict:
    d = np.abs(a.astype(np.int16) - b.astype(np.int16)).reshape(-1, 4)
    m = d.max(axis=1)
    hist = [int(((m >= lo) & (m < hi)).sum()) for lo, hi in zip([0, 1, 2, 3, 4, 8, 16, 32], BINS[1:])]
    lit = (a.reshape(-1, 4).max(axis=1) > 0) | (b.reshape(-1, 4).max(axis=1) > 0)
    return {"pixels": int(m.size), "lit_pixels": int(lit.sum()), "hist_0_1_2_3_4to7_8to15_16to31_32up": hist,
            "within1": float((m <= 1).mean()), "within2": float((m <= 2).mean()), "within4": float((m <= 4).mean()),
            "max_lsb": int(m.max()), "mean_abs_per_channel": [float(x) for x in d.mean(axis=0)]}


def compare_case(R, oracle, sc, ref_gpu=None, samplers=("trilinear", "trilinear-fma", "hybridw"), rcs=None):
    """Frames of one oracle.Scene (or of the Raycaster list `rcs` sharing its tables) through GPURenderer4 and through
    the product's samplers; returns {sampler: lsb_report vs GPURenderer4} accumulated over the frames."""
    from common import product_scene
    rcs = rcs or [sc.rc]
    ref_gpu = ref_gpu or oracle.RefGpu.get(sc)
    want = [ref_gpu.render(4, rc)[0] for rc in rcs]
    scene5, _ = product_scene(R, sc)
    out = {}
    for s in samplers:
        gpu = R.GPURenderer5(0, sampler=s)
        try:
            gpu.set_window_buffer(sc.width, sc.height)
            gpu.set_scene(scene5)
            got = []
            for rc in rcs:
                got.append(gpu.render_volume(scene5.params(R.View.from_buffer_copy(bytes(rc.view)))))
            out[s] = lsb_report(np.stack(got), np.stack(want))
        finally:
            gpu.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--full", action="store_true", help="also C2 (256^3) and C3 (512^3) orbit frames at 1920x1080")
    args = ap.parse_args()
    import bench
    bench.claim_stdout()  # points fd 1 at stderr; emit() writes the line to the real stdout
    import oracle
    import vr5_loader
    from common import golden_scene

    if not oracle.RefGpu.available():
        bench.emit({"tool": "ref4_pin", "unavailable": "oracle/_ref/libvolr_ref_gpu.so not built (needs /root/reference at build time)"})
        return
    vr5_loader.load()
    R = vr5_loader.renderer_module()
    port = oracle.Port()
    with open(os.path.join(ROOT, "tests", "golden", "golden.json")) as f:
        golden = json.load(f)
    with gzip.open(os.path.join(ROOT, "tests", "golden", "bucky32.u8.gz"), "rb") as f:
        bucky = np.frombuffer(f.read(), dtype=np.uint8).reshape(32, 32, 32).copy()
    vols = {}

    def volume(spec):
        key = json.dumps(spec, sort_keys=True)
        if key not in vols:
            vols[key] = bucky if spec["kind"] == "bucky" else port.synth_a(spec["n"], spec["seed"]) if spec["kind"] == "synthA" else port.synth_b(spec["n"], spec["seed"])
        return vols[key]

    cases = {}
    for case in golden["cases"]:
        sc = golden_scene(port, case, volume(case["volume"]))
        cases[case["name"]] = compare_case(R, oracle, sc)
    if args.full:
        for wl_name, n_frames in (("c2", 4), ("c3", 3)):
            wl = bench.WORKLOADS[wl_name]
            cpu = bench.CpuReference(wl)
            scs = [cpu.scene(bench.orbit_angles(i, n_frames)[1]) for i in range(n_frames)]
            cases[wl_name + "_orbit"] = compare_case(R, oracle, scs[0], rcs=[s.rc for s in scs])
            cases[wl_name + "_orbit"]["frames"] = n_frames
            # timing row: the recompiled GPURenderer4 (hardware filtering) next to the contracted software policy
            ref_gpu = oracle.RefGpu.get(scs[0])
            ref_gpu.render(4, [s.rc for s in scs])
            _, ms4 = ref_gpu.render(4, [s.rc for s in scs] * 4)
            cases[wl_name + "_orbit"]["gpurenderer4_ms_per_frame"] = ms4 / (4 * n_frames)
    bench.emit({"tool": "ref4_pin", "reference": "GPURenderer4.cu unmodified, texture references over texture objects (oracle/ref_gpu4_texobj.cu)",
                "metric": "per-pixel max channel |product - GPURenderer4| in LSB; hist bins 0,1,2,3,4-7,8-15,16-31,32+", "cases": cases})


if __name__ == "__main__":
    main()
