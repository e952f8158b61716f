#!/usr/bin/env python
"""debug: one GPU, ring of S slots, N frames in one sequence launch with a concurrent consumer; time + ring status"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import vr5_loader
vr5_loader.load(); R = vr5_loader.renderer_module()
def run(slots, nframes, copy, tfname, wh=(320, 240), n=64):
    gpu = R.GPURenderer5(0)
    gpu.synth_volume(1 if tfname == "opaque" else 0, n, 2 if tfname == "opaque" else 1)
    bd, mm, bs = gpu.build_minmax_device()
    tf = R.premultiply_tf(R.tf_opaque() if tfname == "opaque" else R.tf_semi(n))
    scene = R.Scene(None, tf, R.build_esl_bits(tf, mm), bd, bs, R.default_ray_step(n, n, n))
    gpu.set_transfer_fn(scene.tf, scene.esl_bits); gpu.set_window_buffer(*wh)
    frames = [scene.params(R.camera(wh[0], wh[1], 0, (20.0, 31.0 * i, 0.0), 2.0)) for i in range(nframes)]
    gpu.peer_create(slots)
    dst = gpu.device_alloc(nframes * wh[0] * wh[1] * 4) if copy else None
    t0 = time.perf_counter()
    gpu.peer_render_sequence(frames, 0); gpu.peer_consume_sequence(0, nframes, dst); gpu.join(); gpu.sync()
    dt = time.perf_counter() - t0
    try:
        gpu.peer_status(); st = "ok"
    except Exception as e:
        st = "TIMED OUT"
    print("slots %2d frames %2d copy %d tf %-6s warp_fetch %s: %.3f s, %s" % (slots, nframes, copy, tfname, os.environ.get("VR5_WARP_FETCH", "-"), dt, st), flush=True)
    gpu.close()
for tfname in ("semi", "opaque"):
    for slots, nframes, copy in ((2, 7, 1), (2, 7, 0), (8, 7, 1), (2, 3, 1), (2, 3, 0)):
        run(slots, nframes, copy, tfname)
