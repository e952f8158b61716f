#!/usr/bin/env python
"""Headless bench driver for the ray-casting hot path (replaces the reference's GLUT/PBO benchmark loop,
VolR.cpp:225-321, and its Profiler, Profiler.cpp:46-78).

    python bench.py --gpus N --steps K --warmup W              # this repo's CUDA path (libvr5.so through the C ABI)
    python bench.py --impl reference --gpus N --steps K ...    # the reference's own CPURenderer on the host cores

A step = one frame of the 360-frame orbit over one synthetic volume. Default workload = BASELINE.json
configs[2]: synthA 512^3 uint8, 1920x1080, semi-transparent TF. Prints ONE JSON line on rank 0.
Multi-GPU (torchrun, one process per GPU): image-space partition in interleaved 4-row strips, volume replicated,
strips gathered to rank 0 with NCCL; total work is fixed, so scaling is "strong".
"""
import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# stdout carries exactly one JSON line. Libraries write there too (NCCL's version banner at communicator creation, the
# reference's Logger), so file descriptor 1 is pointed at stderr for the life of the process and the line goes to the saved one.
if os.environ.get("NCCL_DEBUG", "").upper() in ("VERSION", "INFO", "TRACE") and not os.environ.get("NCCL_DEBUG_FILE"):
    os.environ["NCCL_DEBUG"] = "WARN"
sys.stdout.flush()
_JSON_FD = os.dup(1)
os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    os.write(_JSON_FD, (json.dumps(line) + "\n").encode())

WORKLOADS = {  # BASELINE.json configs[1..4]
    "c2": dict(kind=0, n=256, seed=1, tf="semi", w=1920, h=1080, name="synthA 256^3 uint8, semi TF, 1920x1080, 360-frame orbit"),
    "c3": dict(kind=0, n=512, seed=1, tf="semi", w=1920, h=1080, name="synthA 512^3 uint8, semi TF, 1920x1080, 360-frame orbit"),
    "c4": dict(kind=1, n=1024, seed=2, tf="opaque", w=3840, h=2160, name="synthB 1024^3 uint8, opaque TF, 3840x2160, 360-frame orbit"),
    "c5": dict(kind=0, n=2048, seed=1, tf="semi", w=7680, h=4320, name="synthA 2048^3 uint8 (8 GiB, replicated), semi TF, 7680x4320, 360-frame orbit"),
}
ORBIT_FRAMES = 360
UNIT = "Gsamples/s"
METRIC = "Gsamples/s and FPS, 512^3 uint8 @1920x1080, 1/2/4/8 B200 vs CPURenderer"


def orbit_angles(i, steps):
    """Frame f of the 360-frame orbit: set_camera_position((20, 360 f / 360, 0), 2); K steps are spread over it."""
    stride = max(1, ORBIT_FRAMES // max(1, steps)) if steps < ORBIT_FRAMES else 1
    f = (i * stride) % ORBIT_FRAMES
    return f, (20.0, 360.0 * f / ORBIT_FRAMES, 0.0)


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU through NVML (what nvidia-smi reads) while a region runs."""
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
               0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting"}

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz, self.ok = [], set(), None, False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # NVML missing: report that instead of inventing numbers
            self.err = str(e)
        self._stop = threading.Event()
        self._t = None

    def _run(self):
        while not self._stop.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                mask = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(self.nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in self.REASONS.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.005)

    def start(self):
        if self.ok:
            self._stop.clear()
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()

    def stop(self):
        if self._t:
            self._stop.set()
            self._t.join()
            self._t = None

    def summary(self):
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "note": "NVML sampling unavailable"}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "n_samples": len(self.samples)}


# ---- the reference's CPU path (oracle/_ref when present, else the restated port): cpu_baseline and --impl reference --
class CpuReference:
    """Drives the reference's own CPURenderer (kind "reference") or, if oracle/_ref was not prebuilt, the restated
    oracle (kind "port") on host volumes. The only place bench.py executes anything under oracle/."""

    def __init__(self, wl):
        import oracle
        self.oracle, self.wl = oracle, wl
        self.port = oracle.Port()
        n = wl["n"]
        self.volume = self.port.synth_a(n, wl["seed"]) if wl["kind"] == 0 else self.port.synth_b(n, wl["seed"])
        base = self.port.default_base_tf()
        self.base_tf = oracle.tf_semi(base, n) if wl["tf"] == "semi" else (oracle.tf_opaque(base) if wl["tf"] == "opaque" else base)
        self.kind = "reference" if os.path.exists(oracle.REF_SO) and n <= 1024 else "port"  # 32-bit index limit, SURVEY H3
        if self.kind == "reference":
            self.ref = oracle.Ref()
            devnull, saved = os.open(os.devnull, os.O_WRONLY), os.dup(1)
            os.dup2(devnull, 1)  # the reference's Logger prints to stdout; keep the JSON line clean
            try:
                self.ref.set_volume(self.volume)
                self.ref.set_base_tf(self.base_tf)
            finally:
                os.dup2(saved, 1)
                os.close(devnull)
                os.close(saved)
        self._scene_cache = {}

    def scene(self, angles):
        """Scene with the reference's own tables; the View comes from the same camera code as the CUDA arm uses
        (vr5_scene_camera, bit-identical to ViewBase on these poses: tests/test_scene_builders.py)."""
        if self.kind == "reference":
            self.ref.camera(self.wl["w"], self.wl["h"], 0, angles, 2.0)
            return self._ref_scene()
        view = self._port_view(angles)
        return self.port.make_scene(self.volume, self.base_tf, view)

    def _ref_scene(self):
        o = self.oracle
        return o.Scene(self.volume, self.ref.tf(), self.ref.esl_bits(), self.ref.raycaster())

    def _port_view(self, angles):
        import vr5_loader
        R = vr5_loader.renderer_module()
        v = R.camera(self.wl["w"], self.wl["h"], 0, angles, 2.0)
        return self.oracle.View.from_buffer_copy(bytes(v))

    def render_rows(self, scene, out, spans, threads):
        """Cast the rows of `spans` with `threads` host threads (ctypes releases the GIL); returns seconds."""
        from concurrent.futures import ThreadPoolExecutor
        if self.kind == "reference":
            fn = lambda s: self.ref.render_rows(scene.rc, out, s[0], s[1])
        else:
            fn = lambda s: self.port.lib.orc_render_rows(C.byref(scene.rc), out.ctypes.data_as(C.POINTER(C.c_uint8)), s[0], s[1], None)
        t0 = time.perf_counter()
        if threads == 1:
            for s in spans:
                fn(s)
        else:
            with ThreadPoolExecutor(max_workers=threads) as ex:
                list(ex.map(fn, spans))
        return time.perf_counter() - t0

    def count_rows(self, scene, spans, threads):
        """Samples in the given rows, from the restated oracle's counters (untimed; the reference does not count)."""
        from concurrent.futures import ThreadPoolExecutor
        scratch = np.zeros((scene.height, scene.width, 4), dtype=np.uint8)

        def one(s):
            cnt = self.oracle.Counters()
            self.port.lib.orc_render_rows(C.byref(scene.rc), scratch.ctypes.data_as(C.POINTER(C.c_uint8)), s[0], s[1], C.byref(cnt))
            return cnt.samples
        with ThreadPoolExecutor(max_workers=threads) as ex:
            return int(sum(ex.map(one, spans)))


def strip_spans(h, every=1, chunk_rows=8):
    """Row spans covering every `every`-th 4-row strip group; chunked finely so host threads stay balanced."""
    spans = []
    for r in range(0, h, chunk_rows * every):
        spans.append((r, min(h, r + chunk_rows)))
    return spans


def run_reference_arm(args, wl):
    """--impl reference: the reference's CPURenderer code on all host threads, on this arm's config and metric.
    A step = one orbit frame, sampled (every m-th 8-row band) when few cores are available so a run ends in minutes."""
    threads = os.cpu_count() or 1
    cpu = CpuReference(wl)
    est_frame_core_s = 11.0 * (wl["n"] / 512.0) * (wl["w"] * wl["h"]) / (1920.0 * 1080.0)
    # rows are rendered twice per step (timed, then counted): keep the whole --steps/--warmup run near 90 s of wall clock
    every = max(1, int(round(est_frame_core_s * 2.0 * (args.steps + args.warmup) / threads / 90.0)))
    spans = strip_spans(wl["h"], every)
    frac = sum(b - a for a, b in spans) / wl["h"]
    out = np.zeros((wl["h"], wl["w"], 4), dtype=np.uint8)
    total_samples, total_t = 0, 0.0
    for i in range(args.warmup + args.steps):
        _, ang = orbit_angles(max(0, i - args.warmup), args.steps)
        sc = cpu.scene(ang)
        t = cpu.render_rows(sc, out, spans, threads)
        if i >= args.warmup:
            total_t += t
            total_samples += cpu.count_rows(sc, spans, threads)
    value = total_samples / total_t / 1e9
    sample = "every %d-th 8-row band of each orbit frame (%.1f %% of the rows), row-parallel over %d threads" % (every, 100 * frac, threads)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total_t / args.steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "fps": args.steps * frac / total_t,
        "config": {"workload": wl["name"], "volume": "%d^3" % wl["n"], "viewport": "%dx%d" % (wl["w"], wl["h"]), "tf": wl["tf"],
                   "projection": "orthographic", "note": "reference render_ray (CPURenderer.cpp:11-41) driven row-parallel; as shipped it is single-threaded"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": cpu.kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ---- this repo's arm ---------------------------------------------------------------------------------------------------
def run_b200_arm(args, wl):
    import torch
    import torch.distributed as dist

    import vr5_loader
    vr5_loader.load()
    R = vr5_loader.renderer_module()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d" % (args.gpus, world))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    n, w, h = wl["n"], wl["w"], wl["h"]
    gpu = R.GPURenderer5(local_rank, sampler=args.sampler)
    stream = torch.cuda.Stream(device=dev)  # a real (non-default) stream shared by the kernels, NCCL and the timing events
    torch.cuda.set_stream(stream)
    gpu.set_stream(stream.cuda_stream)
    gpu.synth_volume(wl["kind"], n, wl["seed"])          # replicated: every rank generates the same bytes
    bd, mm, bs = gpu.build_minmax_device()
    base = {"semi": lambda: R.tf_semi(n), "opaque": R.tf_opaque, "default": R.default_base_tf}[wl["tf"]]()
    tf = R.premultiply_tf(base)
    scene = R.Scene(None, tf, R.build_esl_bits(tf, mm), bd, bs, R.default_ray_step(n, n, n))
    gpu.set_transfer_fn(scene.tf, scene.esl_bits)
    gpu.set_window_buffer(w, h)

    def params_for(i, steps):
        _, ang = orbit_angles(i, steps)
        return scene.params(R.camera(w, h, 0, ang, 2.0))

    frames = [params_for(i, args.steps) for i in range(args.steps)]
    warm = [params_for(i * 7 + 3, ORBIT_FRAMES) for i in range(max(args.warmup, 3))]

    # N > 1: "peer" = every rank's kernel stores its pixels straight into rank 0's frame ring over NVLink (fused render +
    # gather, csrc/vr5_peer.cu); "nccl" = packed strips + dist.gather + unpack kernel (the plain-collective baseline)
    gather = args.gather if world > 1 else "none"
    lanes = 1 if gather == "nccl" else args.overlap   # the NCCL gather consumes each frame on the one torch stream
    gpu.set_overlap(lanes)
    gpu.set_partition(rank, world, world > 1 and gather == "nccl")
    PEER_SLOTS = 2 * max(1, args.overlap)   # a multiple of the lanes; two uses of headroom per lane
    if gather == "peer":
        box = [gpu.peer_create(PEER_SLOTS) if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        if rank != 0:
            gpu.peer_open(box[0], PEER_SLOTS)
        dist.barrier()
    seq = [0]
    # end-to-end leg at N > 1: "host" = every rank copies its own strips into a frame ring in shared HOST memory over its
    # own PCIe link (csrc/vr5_hostring.cu); "peer" = rank 0 copies the gathered frame from its HBM (one link for all)
    e2e_gather = args.e2e_gather if (world > 1 and gather == "peer") else "same"
    hseq, hrel, ring_mem, ring_name = [0], [0], None, None
    hprof = [0.0, 0.0]
    # The consumer learns of a delivery through host memory, later than a stream would, and a frame is only delivered after
    # kernel + copy + flag: it trails the producers by RING_LAG frames (measured at N=8: lag 3 -> 6 300 FPS, 4 -> 8 100,
    # >= 7 -> the device rate), in a ring of 3 x lanes slots (producer of frame j needs frame j - slots released).
    RING_SLOTS = 3 * max(1, args.overlap)
    RING_LAG = RING_SLOTS - max(1, args.overlap) - 1
    if os.environ.get("VR5_RING_SLOTS"):   # experiment knobs
        RING_SLOTS = int(os.environ["VR5_RING_SLOTS"])
    if os.environ.get("VR5_RING_LAG"):
        RING_LAG = int(os.environ["VR5_RING_LAG"])
    if e2e_gather == "host":     # the ring must fit the node's shared memory, else fall back to rank 0's copy
        fits = [False]
        if rank == 0:
            try:
                st = os.statvfs("/dev/shm")
                fits[0] = st.f_bavail * st.f_frsize > 1.25 * R.hostring_bytes(w, h, RING_SLOTS)
            except OSError:
                pass
        dist.broadcast_object_list(fits, src=0)
        if not fits[0]:
            e2e_gather = "peer"
    if e2e_gather == "host":
        nbytes = R.hostring_bytes(w, h, RING_SLOTS)
        box = ["vr5_ring_%d_%d" % (os.getpid(), int(time.time())) if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        ring_name = box[0]
        if rank == 0:
            ring_mem = R.ring_memory(ring_name, nbytes, create=True)
        dist.barrier()
        if rank != 0:
            ring_mem = R.ring_memory(ring_name, nbytes, create=False)

    L = R.lib()
    rows_padded = max(L.vr5_strip_rows(h, r, world) for r in range(world))
    frame_devs = [torch.empty((h, w, 4), dtype=torch.uint8, device=dev) for _ in range(lanes)]  # frames in flight differ
    frame_dev = frame_devs[0]
    packed = torch.empty((rows_padded, w, 4), dtype=torch.uint8, device=dev) if gather == "nccl" else None
    gathered = torch.empty((world, rows_padded, w, 4), dtype=torch.uint8, device=dev) if (gather == "nccl" and rank == 0) else None
    pinneds = [torch.empty((h, w, 4), dtype=torch.uint8).pin_memory() for _ in range(lanes)] if rank == 0 else None
    pinned = pinneds[0] if rank == 0 else None

    def step_device(p):
        """One frame, device resident: N=1 renders in place; N>1 renders this rank's strips, gathers them to rank 0
        over NCCL and scatters them into the frame there. Everything is enqueued on one stream, no host sync."""
        if world == 1:
            gpu.render_device(p, frame_devs[seq[0] % lanes].data_ptr(), blocking=False)
            seq[0] += 1
        elif gather == "peer":
            gpu.peer_render(p, seq[0])
            if rank == 0:
                gpu.peer_wait(seq[0])       # frame complete in the ring (a consumer would read it here)
                gpu.peer_release(seq[0])
            seq[0] += 1
        else:
            gpu.render_device(p, packed.data_ptr(), blocking=False)
            dist.gather(packed, list(gathered.unbind(0)) if rank == 0 else None, dst=0)
            if rank == 0:
                gpu.unpack_strips(gathered.data_ptr(), rows_padded, world, frame_dev.data_ptr())

    def step_e2e(p):
        """The call a user makes: the frame lands in HOST memory (pinned), like renderers 1-2 of the reference. With one
        lane the call blocks per frame; with frame overlap frame i is enqueued (kernel + D2H on its lane) and the host
        then waits for frame i - lanes + 1, i.e. it receives every frame, in order, lanes - 1 frames behind."""
        i = seq[0]
        if world == 1 and lanes == 1:
            gpu.render_to_pinned(p, pinned.data_ptr())
        elif world == 1:
            gpu.render_async_host(p, pinneds[i % lanes].data_ptr())
            seq[0] += 1
            if i >= lanes - 1:
                gpu.sync_lane((i - (lanes - 1)) % lanes)
        elif e2e_gather == "host":
            j = hseq[0]
            ta = time.perf_counter()
            gpu.hostring_render(p, j)          # kernel + strided D2H of this rank's strips into the shared host frame
            tb = time.perf_counter()
            hseq[0] += 1
            if rank == 0 and j >= RING_LAG:    # the consumer receives every frame, in order, RING_LAG frames behind
                gpu.hostring_wait(hrel[0])     # all ranks' strips of that frame are in host memory
                gpu.hostring_release(hrel[0])
                hrel[0] += 1
            tc = time.perf_counter()
            hprof[0] += tb - ta; hprof[1] += tc - tb
        elif gather == "peer":
            gpu.peer_render(p, i)
            if rank == 0:
                gpu.peer_wait(i)
                gpu.peer_read(i, pinneds[i % lanes].data_ptr())   # D2H of the complete frame, on the frame's lane
                gpu.peer_release(i)
            seq[0] += 1
            if i >= lanes - 1:
                gpu.peer_sync_frame(i - (lanes - 1))
        else:
            step_device(p)
            if rank == 0:
                pinned.copy_(frame_dev, non_blocking=True)
            stream.synchronize()

    def barrier():
        if world == 1:
            gpu.set_overlap(lanes)   # every lane idle, next vr5_render* call on lane 0 ...
            seq[0] += (-seq[0]) % lanes   # ... and this script's buffer index with it
        else:
            gpu.sync()               # peer frames keep their sequence: lane = seq % lanes, slot = seq % slots on every rank
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # -- work counters per frame (instrumented kernel, untimed) ------------------------------------------------------
    counts = [gpu.count(p) for p in frames]
    samples_local = sum(c["samples"] for c in counts)
    shaded_local = sum(c["shaded"] for c in counts)
    tot = torch.tensor([samples_local, shaded_local], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(tot)
    samples_total, shaded_total = int(tot[0]), int(tot[1])

    # -- warm-up -------------------------------------------------------------------------------------------------
    for p in warm:
        step_device(p)
    barrier()

    # -- timed region: exactly K steps, device timed, max over ranks -----------------------------------------------
    clocks = ClockSampler(local_rank)
    launches0 = gpu.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    clocks.start()
    barrier()
    e0.record(stream)
    for p in frames:
        step_device(p)
    gpu.join()           # the stream waits (on the device) for the last frame of every lane
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    launches = gpu.launch_count() - launches0
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    lt = torch.tensor([launches], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(lt)
    ms = float(t[0])
    launches = int(lt[0])
    # keep the same load running (untimed) until the clock sampler has seen >= 1 s of it; the number of extra steps is
    # derived from the reduced time, so every rank runs the same count (the peer ring needs equal frame sequences)
    soak_steps = int(max(0.0, 1.0 - ms / 1e3) / max(ms / 1e3 / args.steps, 1e-6)) + 1
    for i in range(soak_steps):
        step_device(frames[i % len(frames)])
    torch.cuda.synchronize(dev)
    clocks.stop()

    # -- end to end: host buffers, blocking call per frame, wall clock, max over ranks --------------------------------
    def drain_host():
        while rank == 0 and hrel[0] < hseq[0]:   # the last lanes - 1 frames
            gpu.hostring_wait(hrel[0])
            gpu.hostring_release(hrel[0])
            hrel[0] += 1

    if e2e_gather == "host":
        gpu.sync()
        gpu.set_partition(rank, world, True)     # packed strips for the strided copy
        if rank == 0:
            gpu.hostring_attach(ring_mem, RING_SLOTS, init=True)    # clears the header before anyone else looks at it
        dist.barrier()
        if rank != 0:
            gpu.hostring_attach(ring_mem, RING_SLOTS, init=False)
        dist.barrier()
    for p in warm[:3]:
        step_e2e(p)
    drain_host()
    barrier()
    t0 = time.perf_counter()
    for p in frames:
        step_e2e(p)
    drain_host()
    barrier()
    e2e_s = time.perf_counter() - t0
    if os.environ.get("VR5_BENCH_DEBUG"):
        print("rank %d: e2e %.1f ms, in render calls %.1f ms, in wait/release %.1f ms" % (rank, e2e_s * 1e3, hprof[0] * 1e3, hprof[1] * 1e3), file=sys.stderr)
    if e2e_gather == "host":
        gpu.hostring_detach()
        gpu.set_partition(rank, world, False)
        ring_mem.close()
        dist.barrier()
        if rank == 0:
            R.ring_memory_unlink(ring_name)
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t[0])

    # -- dominant kernel alone: per-launch CUDA events on its stream (this rank's share of each frame) ---------------
    out_ptr = packed.data_ptr() if gather == "nccl" else frame_dev.data_ptr()
    if gather == "peer":
        gpu.peer_status()   # raises if any device-side wait timed out
    k_ms = gpu.render_timed(frames, out_ptr, l2_flush=False)
    k_ms_flushed = gpu.render_timed(frames[: min(len(frames), 24)], out_ptr, l2_flush=True)
    strip_px = (L.vr5_strip_rows(h, rank, world) if world > 1 else h) * w
    voxel_bytes = 8.0 if args.sampler.startswith("trilinear") else 1.0   # nearest 1 B per fetch, trilinear 8 B
    alg_bytes = voxel_bytes * (samples_local + shaded_local) / len(frames) + 4.0 * strip_px  # + 4 B per pixel written
    peak, peak_src = measured_peak()
    achieved = alg_bytes / (float(k_ms.mean()) * 1e-3) / 1e9
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get("%s_%s" % (args.workload, args.sampler))
    except Exception:
        pass

    line = None
    if rank == 0:
        value = samples_total / (ms * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": len(warm),
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "fps": args.steps / (ms * 1e-3),
            "config": {"workload": wl["name"], "volume": "%d^3" % n, "viewport": "%dx%d" % (w, h), "tf": wl["tf"],
                       "projection": "orthographic", "sampler": args.sampler, "frame_overlap_lanes": lanes, "samples_per_frame": samples_total / args.steps,
                       "parallelism": ("image-space strips x%d, volume replicated, %s" % (world, "pixels stored straight into rank 0's frame ring over NVLink peer memory (fused render + gather)" if gather == "peer" else "NCCL gather to rank 0")) if world > 1 else "single GPU",
                       "l2": "no flush in the timed loop: per frame the kernel reads the %.0f MB volume and writes the %.1f MB frame (L2 is 126 MB); "
                             "the flushed per-kernel time is reported in roofline.kernel_ms_l2_flushed" % (n ** 3 / 1e6, w * h * 4 / 1e6)},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                         "peak_source": peak_src, "kernel": ("vr5::tri_kernel<variant %d>" % gpu.last_variant() if args.sampler.startswith("trilinear") else "vr5::cast_kernel<%s, variant %d>" % (args.sampler, gpu.last_variant())),
                         "bytes_per_sample": voxel_bytes, "kernel_ms": float(k_ms.mean()),
                         # the same algorithmic bytes over the time a frame takes in the timed loop, where the tail of one launch
                         # overlaps the head of the next (frame lanes): the rate the workload is actually served at
                         "achieved_in_stream": alg_bytes / (ms / args.steps * 1e-3) / 1e9, "frac_in_stream": alg_bytes / (ms / args.steps * 1e-3) / 1e9 / peak,
                         "kernel_ms_l2_flushed": float(k_ms_flushed.mean()), "algorithmic_bytes_per_launch": alg_bytes,
                         "note": ("trilinear policy, contracted arithmetic (GPURenderer4 semantics as nvcc -fmad=true builds them, NOT the parity mode): 8 B per sample" if args.sampler == "trilinear-fma" else
                                  "trilinear policy (GPURenderer4 semantics, NOT the parity mode): 8 B per sample" if args.sampler == "trilinear" else
                                  "parity mode fetches 1 B per sample; the kernel is bound by the fp32 pipe and issue slots, not by HBM (DESIGN.md)")},
            "e2e": {"value": samples_total / e2e_s / 1e9, "unit": UNIT, "fps": args.steps / e2e_s, "h2d_bytes_per_step": C.sizeof(R.Params),
                    "d2h_bytes_per_step": w * h * 4,
                    "path": {"host": "every rank copies its strips into a frame ring in shared host memory over its own PCIe link (vr5_hostring_*)",
                             "peer": "rank 0 copies the frame gathered in its HBM (vr5_peer_read)", "same": "kernel + D2H of the frame on its lane"}[e2e_gather]},
            "gpu_launches": launches, "clocks": clocks.summary(),
        }
    if world > 1:
        dist.barrier()
    # -- CPU baseline next to it (rank 0, N=1 only): the reference's CPURenderer as shipped = 1 thread, one orbit frame ----
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(args, wl)
    if rank == 0:
        emit(line)
    gpu.close()
    if world > 1:
        dist.destroy_process_group()


def cpu_baseline(args, wl):
    cpu = CpuReference(wl)
    _, ang = orbit_angles(0, args.steps)
    sc = cpu.scene(ang)
    # bounded sample: ~10-30 s of single-thread work. One 512^3/1080p frame costs ~10 s; larger configs take a row band.
    est = 11.0 * (wl["n"] / 512.0) * (wl["w"] * wl["h"]) / (1920.0 * 1080.0)
    every = max(1, int(round(est / 20.0)))
    spans = strip_spans(wl["h"], every)
    out = np.zeros((wl["h"], wl["w"], 4), dtype=np.uint8)
    secs = cpu.render_rows(sc, out, spans, 1)
    samples = cpu.count_rows(sc, spans, os.cpu_count() or 1)
    frac = sum(b - a for a, b in spans) / wl["h"]
    return {"value": samples / secs / 1e9, "unit": UNIT, "cores": 1, "kind": cpu.kind, "fps": frac / secs,
            "host_cores_available": os.cpu_count(),
            "sample": "orbit frame 0, %s, render_ray loop of CPURenderer as shipped (single thread), %.1f s" %
                      ("all rows" if every == 1 else "every %d-th 8-row band (%.1f %% of rows)" % (every, 100 * frac), secs)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=360)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--sampler", default=os.environ.get("VR5_SAMPLER", "hybridw"), choices=["linear", "brick", "tex3d", "hybrid", "hybridw", "window", "hybridw2", "trilinear", "trilinear-fma"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--overlap", type=int, default=4, help="frame lanes (vr5_set_overlap): 1 = strictly serial frames")
    ap.add_argument("--gather", default="peer", choices=["peer", "nccl"], help="N > 1: how strips reach rank 0")
    ap.add_argument("--e2e-gather", default="host", choices=["host", "peer"],
                    help="N > 1, end-to-end leg: host = every rank copies its strips into a shared host frame ring; peer = rank 0 copies the gathered frame")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) != 0:
            return  # under torchrun only rank 0 runs the CPU arm
        run_reference_arm(args, wl)
    else:
        run_b200_arm(args, wl)


if __name__ == "__main__":
    main()
