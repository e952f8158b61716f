#!/usr/bin/env python
"""The reference's OWN CUDA renderers, recompiled unmodified for sm_100a, timed headless on a bench workload next to
GPURenderer5 (SURVEY 8(d) "optional GPU comparison row"; run on the GPU box).

oracle/_ref/libvolr_ref_gpu.so = GPURenderer1.cu + GPURenderer23.cu of the reference, compiled where they lie with nvcc's
defaults (oracle/Makefile, driver oracle/ref_gpu_api.cu). Measurement infrastructure: nothing of it is on the product path.
Renderer 3 ("CUDA CM + GL interop") writes into a device buffer - here a plain cudaMalloc block instead of the GL pixel
buffer - and is the like-for-like row for the device-timed `value` of bench.py; renderer 1 ("CUDA Straightforward")
copies every frame to pageable host memory as shipped and is the row for `e2e`.

Prints one JSON line: ms per frame and FPS of both, Gsamples/s with the samples counted by GPURenderer5's instrumented
pass on the same frames, GPURenderer5's own ms per frame in the same process (serial blocking frames, device buffer),
and the LSB difference of the last frame (the reference's kernels are built with -fmad=true, so a few pixels differ from
the CPU renderer's - and therefore from GPURenderer5's - frame: SURVEY P5)."""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c3", choices=["c2", "c3", "c4"])  # 2048^3 overflows the reference's 32-bit voxel index
    ap.add_argument("--steps", type=int, default=24)
    ap.add_argument("--warmup", type=int, default=3)
    args = ap.parse_args()

    import bench
    bench.claim_stdout()
    import oracle
    import vr5_loader
    from common import parity_report, product_scene

    wl = bench.WORKLOADS[args.workload]
    so = os.path.join(ROOT, "oracle", "_ref", "libvolr_ref_gpu.so")
    if not os.path.exists(so):
        bench.emit({"impl": "reference-gpu", "unavailable": "oracle/_ref/libvolr_ref_gpu.so not built (needs /root/reference at build time)"})
        return
    lib = C.CDLL(so)
    lib.refgpu_create.restype = C.c_void_p
    lib.refgpu_create.argtypes = [C.POINTER(oracle.Raycaster)]
    lib.refgpu_destroy.argtypes = [C.c_void_p]
    lib.refgpu_run.argtypes = [C.c_void_p, C.c_int, C.POINTER(oracle.Raycaster), C.c_int, C.POINTER(C.c_uint8), C.POINTER(C.c_double)]
    assert lib.refgpu_sizeof_raycaster() == C.sizeof(oracle.Raycaster)

    cpu = bench.CpuReference(wl)  # host volume + the reference's own tables and Raycaster POD
    scenes = [cpu.scene(bench.orbit_angles(i, args.steps)[1]) for i in range(args.steps)]
    frames = (oracle.Raycaster * len(scenes))(*[sc.rc for sc in scenes])
    w, h = wl["w"], wl["h"]
    out = np.zeros((h, w, 4), dtype=np.uint8)
    outp = out.ctypes.data_as(C.POINTER(C.c_uint8))
    devnull, saved = os.open(os.devnull, os.O_WRONLY), os.dup(1)
    os.dup2(devnull, 1)  # the reference's Logger prints to stdout
    try:
        handle = lib.refgpu_create(C.byref(scenes[0].rc))
        assert handle, "refgpu_create failed"
        ms = C.c_double()
        res = {}
        for which, name in ((3, "GPURenderer3 (device buffer)"), (1, "GPURenderer1 (host buffer, D2H per frame)")):
            rc = lib.refgpu_run(handle, which, frames, min(args.warmup, len(scenes)), outp, C.byref(ms))
            assert rc == 0, rc
            rc = lib.refgpu_run(handle, which, frames, len(scenes), outp, C.byref(ms))
            assert rc == 0, rc
            res[which] = (name, ms.value / len(scenes), out.copy())
        lib.refgpu_destroy(handle)
    finally:
        os.dup2(saved, 1)
        os.close(devnull)
        os.close(saved)

    # GPURenderer5 on the same frames: samples per frame, its own serial device-buffer time, last-frame difference
    vr5_loader.load()
    R = vr5_loader.renderer_module()
    gpu = R.GPURenderer5(0)
    gpu.set_window_buffer(w, h)
    scene5, _ = product_scene(R, scenes[0])
    gpu.set_scene(scene5)
    params = [product_scene(R, sc)[1] for sc in scenes]
    samples = sum(gpu.count(p)["samples"] for p in params) / len(params)
    dev = gpu.device_alloc(w * h * 4)
    for p in params[: args.warmup]:
        gpu.render_device(p, dev)
    gpu.sync()
    t0 = time.perf_counter()
    for p in params:
        gpu.render_device(p, dev, blocking=False)
    gpu.sync()
    ours_ms = (time.perf_counter() - t0) * 1e3 / len(params)
    ours_last = gpu.render_volume(params[-1])
    gpu.device_free(dev)
    gpu.close()

    line = {"impl": "reference-gpu", "workload": wl["name"], "steps": args.steps, "samples_per_frame": samples,
            "note": "the reference's CUDA renderers recompiled unmodified for sm_100a (nvcc defaults), wall clock around the frame loop, device idle before and after"}
    for which, key in ((3, "renderer3_device"), (1, "renderer1_host")):
        name, per, frame = res[which]
        rep = parity_report(frame, ours_last)
        line[key] = {"name": name, "ms_per_frame": per, "fps": 1e3 / per, "gsamples_per_s": samples / (per * 1e-3) / 1e9,
                     "vs_gpurenderer5_last_frame": {"exact": rep["exact"], "within1": rep["within1"], "max_lsb": rep["max"], "differing_pixels": rep["differing"]}}
    line["gpurenderer5_serial_device"] = {"ms_per_frame": ours_ms, "fps": 1e3 / ours_ms, "gsamples_per_s": samples / (ours_ms * 1e-3) / 1e9,
                                          "speedup_over_renderer3": res[3][1] / ours_ms}
    bench.emit(line)  # importing bench points fd 1 at stderr (library banners); emit() writes to the real stdout


if __name__ == "__main__":
    main()
