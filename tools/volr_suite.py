#!/usr/bin/env python
"""Headless reproduction of VolR's benchmark sweep (VolR.cpp:225-321, config names VolR.cpp:34-38) for GPURenderer5.

Every configuration renders 2 projections x 4 poses at distance 2 (benchmark_config_loop, VolR.cpp:225-253) and reports the
average frame time like Profiler::print_avg (Profiler.cpp). Configurations, in the reference's order:

  datasets            Bucky (the .pvm if given, else the committed fixture) + stand-ins for the six datasets the reference
                      names but does not ship (Daisy ... Porsche): synthetic volumes, said so in the table
  optimisation steps  on the "Foot" stand-in: no optimisations (ESL off, threshold 1), ERT on, ERT + ESL on (VolR.cpp:281-293)
  viewport scale      0.9 ... 0.3 of the window, float accumulation as in VolR.cpp:297-305, dims truncated like
                      ViewBase::set_viewport_dims (ViewBase.cpp:106-112)
  ray step            x1.1 ... x1.7 of the default step, clamped to the reference's limits step/3 .. step*1.666
                      (RaycasterBase.cpp:26-29, :86-92)

--check renders every frame on the CPU oracle too (tests' restatement, pinned to the reference) and requires bit equality:
the sweep then doubles as a parity run over the reference's own benchmark matrix. Needs a GPU; there is no CPU fallback.

    python tools/volr_suite.py [--window 800 600] [--n 128] [--pvm path/Bucky.pvm] [--check] [--reps 3] [--json out.json]
"""
import argparse
import gzip
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

POSES = [(0, 0, 0), (-45, -45, 0), (90, 0, 0), (180, 90, 0)]  # VolR.cpp:235-244


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--window", type=int, nargs=2, default=[800, 600])
    ap.add_argument("--n", type=int, default=128, help="edge of the synthetic stand-in volumes")
    ap.add_argument("--pvm", default=None, help="a real dataset (e.g. the reference's Bucky.pvm)")
    ap.add_argument("--check", action="store_true", help="compare every frame with the CPU oracle, bit for bit")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--sampler", default=None)
    ap.add_argument("--json", default=None)
    args = ap.parse_args()

    import vr5_loader
    vr5_loader.load()
    R = vr5_loader.renderer_module()
    port = None
    if args.check:
        import oracle
        port = oracle.Port()

    W0, H0 = args.window
    gpu = R.GPURenderer5(0, sampler=args.sampler) if args.sampler else R.GPURenderer5(0)
    rows = []

    def dataset(name):
        if name == "Bucky":
            if args.pvm:
                return R.load_volume(args.pvm), "file " + os.path.basename(args.pvm)
            with gzip.open(os.path.join(ROOT, "tests", "golden", "bucky32.u8.gz"), "rb") as f:
                return np.frombuffer(f.read(), dtype=np.uint8).reshape(32, 32, 32).copy(), "fixture decoded by the reference"
        kind = {"Daisy": 0, "VisMale": 0, "Engine": 1, "Foot": 1, "Pig": 0, "Porsche": 1}[name]
        n = args.n if name not in ("VisMale", "Pig") else max(32, args.n // 2)
        gpu.synth_volume(kind, n, 1 + kind)
        return gpu.get_volume((n, n, n)), "stand-in: synth%s %d^3 (dataset not shipped with the reference)" % ("AB"[kind], n)

    def run_config(label, vol, note, w, h, esl=True, threshold=0.95, step_factor=1.0):
        dz, dy, dx = vol.shape
        bd, mm, bs = R.build_minmax(vol)
        tf = R.premultiply_tf(R.default_base_tf())
        step0 = np.float32(R.default_ray_step(dx, dy, dz))
        step = np.float32(step0 * np.float32(step_factor))
        step = min(max(step, np.float32(step0 / np.float32(3))), np.float32(step0 * np.float32(1.666)))  # change_ray_step's CLAMP
        scene = R.Scene(vol, tf, R.build_esl_bits(tf, mm), bd, bs, float(step), ray_threshold=threshold, esl=esl)
        gpu.set_window_buffer(w, h)
        gpu.set_scene(scene)
        times, samples, checked = [], 0, 0
        for persp in (0, 1):
            for pose in POSES:
                view = R.camera(w, h, persp, pose, 2.0)
                params = scene.params(view)
                best = 1e30
                for _ in range(args.reps):
                    t0 = time.perf_counter()
                    frame = gpu.render_volume(params)      # blocking, host buffer: what Profiler method 0 brackets
                    best = min(best, (time.perf_counter() - t0) * 1e3)
                times.append(best)
                samples += gpu.count(params)["samples"]
                if port is not None:
                    import oracle
                    osc = port.make_scene(vol, port.default_base_tf(), oracle.View.from_buffer_copy(bytes(view)), ray_step=float(step),
                                          threshold=threshold, esl=esl, light_kd=scene.light_kd)
                    want, cnt = port.render(osc)
                    if not np.array_equal(frame, want):
                        d = np.abs(frame.astype(np.int16) - want.astype(np.int16))
                        raise SystemExit("PARITY FAILURE in %s pose %s persp %d: max %d LSB on %d pixels" % (label, pose, persp, d.max(), (d.max(axis=2) > 0).sum()))
                    checked += 1
        avg = float(np.mean(times))
        rows.append({"config": label, "viewport": "%dx%d" % (w, h), "volume": "%dx%dx%d" % (dx, dy, dz), "avg_ms": avg, "fps": 1e3 / avg,
                     "msamples_per_s": samples / (sum(times) * 1e-3) / 1e6, "frames": len(times), "oracle_checked": checked, "note": note})
        print("%-22s %-9s %-12s avg %8.3f ms  %8.1f fps  %9.1f Msamples/s  %s%s" % (label, rows[-1]["viewport"], rows[-1]["volume"], avg, 1e3 / avg,
              rows[-1]["msamples_per_s"], "[%d frames == oracle] " % checked if checked else "", note), flush=True)

    for name in ("Bucky", "Daisy", "VisMale", "Engine", "Foot", "Pig", "Porsche"):   # VolR.cpp:271-279
        vol, note = dataset(name)
        run_config(name, vol, note, W0, H0)
    foot, note = dataset("Foot")
    run_config("Foot: No optims", foot, note, W0, H0, esl=False, threshold=1.0)      # VolR.cpp:281-286
    run_config("F: ERT on", foot, note, W0, H0, esl=False, threshold=0.95)
    run_config("F: ERT+ESL on", foot, note, W0, H0, esl=True, threshold=0.95)
    scale = np.float32(1.0)
    while scale > np.float32(0.3):                                                   # VolR.cpp:297-305
        scale = np.float32(scale - np.float32(0.1))
        w, h = int(np.float32(W0) * scale), int(np.float32(H0) * scale)              # (unsigned short)(dims * scale)
        run_config("Scale %.1f" % scale, foot, note, w, h)
    factor = np.float32(1.0)
    while factor <= np.float32(1.7):                                                 # VolR.cpp:309-315
        factor = np.float32(factor + np.float32(0.1))
        run_config("Ray step *%.1f" % factor, foot, note, W0, H0, step_factor=float(factor))
    gpu.close()
    if args.json:
        with open(args.json, "w") as f:
            json.dump({"renderer": "GPURenderer5", "window": [W0, H0], "rows": rows}, f, indent=1)
    print("%d configurations, %d frames%s" % (len(rows), sum(r["frames"] for r in rows),
          ", all equal to the CPU oracle" if args.check else ""))


if __name__ == "__main__":
    main()
