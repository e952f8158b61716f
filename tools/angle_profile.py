#!/usr/bin/env python
"""Per-frame kernel time over the C3 orbit for each sampler policy (which view directions does a memory path like?)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import vr5_loader
vr5_loader.load()
R = vr5_loader.renderer_module()
n, w, h = int(os.environ.get("N", 512)), 1920, 1080
samplers = sys.argv[1:] or ["tex3d", "window", "hybridw"]
gpu = R.GPURenderer5(0)
gpu.synth_volume(0, n, 1)
bd, mm, bs = gpu.build_minmax_device()
tf = R.premultiply_tf(R.tf_semi(n))
scene = R.Scene(None, tf, R.build_esl_bits(tf, mm), bd, bs, R.default_ray_step(n, n, n))
gpu.set_transfer_fn(scene.tf, scene.esl_bits)
gpu.set_window_buffer(w, h)
out = gpu.device_alloc(w * h * 4)
frames = [scene.params(R.camera(w, h, 0, (20.0, float(a), 0.0), 2.0)) for a in range(0, 360, 5)]
samples = np.array([gpu.count(p)["samples"] for p in frames], dtype=np.float64)
print("angle  Msamples " + " ".join("%9s" % s for s in samplers))
res = {}
for s in samplers:
    gpu.set_sampler(s)
    gpu.render_timed(frames[:8], out)
    res[s] = gpu.render_timed(frames, out)
    res[s] = np.minimum(res[s], gpu.render_timed(frames, out))
for i in range(0, len(frames), 3):
    print("%5d %9.1f " % (i * 5, samples[i] / 1e6) + " ".join("%9.4f" % res[s][i] for s in samplers))
print("mean ms         " + " ".join("%9.4f" % res[s].mean() for s in samplers))
print("Gsamples/s      " + " ".join("%9.1f" % (samples.sum() / res[s].sum() / 1e6) for s in samplers))
