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
# reference's Logger), so main() points file descriptor 1 at stderr for the life of the process and the line goes to the saved
# one. (Importing this module - tests, tools - leaves the descriptors alone.)
_JSON_FD = None


def claim_stdout():
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    os.write(_JSON_FD if _JSON_FD is not None else 1, (json.dumps(line) + "\n").encode())


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

    def __init__(self, wl, volume=None):
        import oracle
        self.oracle, self.wl = oracle, wl
        self.port = oracle.Port()
        n = wl["n"]
        self.volume = volume if volume is not None else self.port.synth_a(n, wl["seed"]) if wl["kind"] == 0 else self.port.synth_b(n, wl["seed"])
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
def peer_slots(lanes):
    """Frame ring of the device-resident path: a producer may run this many frames ahead of the consumer (the per-frame
    API needs a multiple of the frame lanes). VR5_RING_SLOTS overrides (profiling: ncu serialises kernels, so a sequence
    kernel must not have to wait for its consumer - a ring at least as long as the sequence never makes it wait)."""
    if os.environ.get("VR5_RING_SLOTS"):
        return int(os.environ["VR5_RING_SLOTS"])
    return 8 if 8 % max(1, lanes) == 0 else 3 * lanes


def sha256(a) -> str:
    import hashlib
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


class Rig:
    """One renderer set up for (workload, sampler) on this rank, with the frame ring of the device-resident path attached:
    rank 0 owns the ring, the other ranks map it over CUDA IPC (world 1: the same ring, nobody to map it)."""

    def __init__(self, ctx, wl, sampler, overlap):
        R, dist = ctx["R"], ctx["dist"]
        self.ctx, self.wl, self.sampler = ctx, wl, sampler
        self.rank, self.world, self.dev = ctx["rank"], ctx["world"], ctx["dev"]
        n, w, h = wl["n"], wl["w"], wl["h"]
        self.gpu = gpu = R.GPURenderer5(ctx["local_rank"], sampler=sampler)
        gpu.set_stream(ctx["stream"].cuda_stream)
        if ctx.get("volume_file"):
            self.volume_route = install_volume_from_file(ctx, gpu, ctx["volume_file"])
            n = max(self.volume_route["dims"])
        else:
            gpu.synth_volume(wl["kind"], n, wl["seed"])          # replicated: every rank generates the same bytes
            self.volume_route = None
        bd, mm, bs = gpu.build_minmax_device()
        base = {"semi": lambda: R.tf_semi(n), "opaque": R.tf_opaque, "default": R.default_base_tf}[wl["tf"]]()
        tf = R.premultiply_tf(base)
        dims = self.volume_route["dims"] if self.volume_route else (n, n, n)
        self.scene = R.Scene(None, tf, R.build_esl_bits(tf, mm), bd, bs, R.default_ray_step(*dims))
        gpu.set_transfer_fn(self.scene.tf, self.scene.esl_bits)
        gpu.set_window_buffer(w, h)
        gpu.set_overlap(overlap)
        # image-space partition of the device-resident path: cost-balanced contiguous bands (a band reads only its slab of the
        # volume; with interleaved strips every GPU reads the whole volume for 1/N of the samples) unless the transfer function
        # can end rays early on its own - then the cost is data-dependent and interleaved strips balance better
        self.partition = ctx.get("partition", "auto")
        if self.partition == "auto":
            self.partition = "strips" if wl["tf"] == "opaque" else "bands"
        self.apply_partition()
        self.slots = peer_slots(overlap)
        box = [gpu.peer_create(self.slots) if self.rank == 0 else None]
        if self.world > 1:
            dist.broadcast_object_list(box, src=0)
            if self.rank != 0:
                gpu.peer_open(box[0], self.slots)
            dist.barrier()
        self.seq = 0

    def apply_partition(self):
        if self.world > 1 and self.partition == "bands":
            self.gpu.set_partition_bands(self.rank, self.world)
        else:
            self.gpu.set_partition(self.rank, self.world, False)

    def params(self, i, steps):
        R = self.ctx["R"]
        _, ang = orbit_angles(i, steps)
        return self.scene.params(R.camera(self.wl["w"], self.wl["h"], 0, ang, 2.0))

    def barrier(self):
        import torch
        self.gpu.sync()
        if self.world > 1:
            self.ctx["dist"].barrier()
        torch.cuda.synchronize(self.dev)

    def sequence(self, frames, copy_dst=None):
        """The device-resident hot path: every rank renders its strips of all `frames` with ONE persistent launch that
        stores the pixels into rank 0's frame ring (NVLink peer stores for the other ranks, in-kernel hand-shake); rank 0
        consumes the frames in order (wait for all arrivals, optional device copy, release). Asynchronous."""
        self.gpu.peer_render_sequence(frames, self.seq)
        if self.rank == 0:
            self.gpu.peer_consume_sequence(self.seq, len(frames), copy_dst)
        self.seq += len(frames)
        self.gpu.join()          # the stream waits (on the device) for the consumer

    def timed_sequence(self, frames, warm):
        """W warm-up frames, then exactly len(frames) frames between two events on the launching stream, barrier and
        device synchronisation on both sides; returns the max over ranks in ms."""
        import torch
        self.sequence(warm)
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.barrier()
        launches0 = self.gpu.launch_count()
        e0.record(self.ctx["stream"])
        self.sequence(frames)
        e1.record(self.ctx["stream"])
        self.launches = self.gpu.launch_count() - launches0    # kernels of this library launched inside the timed region
        self.barrier()
        ms_local = e0.elapsed_time(e1)
        t = torch.tensor([ms_local], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            self.ctx["dist"].all_reduce(t, op=self.ctx["dist"].ReduceOp.MAX)
        return float(t[0]), ms_local

    def count(self, frames):
        """Work counters of this rank's strips (instrumented kernel, untimed), summed over ranks."""
        import torch
        counts = [self.gpu.count(p) for p in frames]
        local = (sum(c["samples"] for c in counts), sum(c["shaded"] for c in counts))
        tot = torch.tensor(local, dtype=torch.int64, device=self.dev)
        if self.world > 1:
            self.ctx["dist"].all_reduce(tot)
        return local, (int(tot[0]), int(tot[1]))

    def frames_alone(self, frames):
        """rank 0: the frames rendered by this GPU alone through the blocking per-frame call (vr5_render, host buffer)."""
        out = []
        self.gpu.set_partition(0, 1, False)
        try:
            for p in frames:
                out.append(self.gpu.render_volume(p))
        finally:
            self.apply_partition()
        return out

    def frames_through_ring(self, frames):
        """All ranks: the frames through the sequence path; rank 0 returns them (copied out of the ring by the consumer)."""
        import torch
        w, h = self.wl["w"], self.wl["h"]
        dst = torch.zeros((len(frames), h, w, 4), dtype=torch.uint8, device=self.dev) if self.rank == 0 else None
        self.sequence(frames, dst.data_ptr() if self.rank == 0 else None)
        self.barrier()
        self.gpu.peer_status()
        return [f for f in dst.cpu().numpy()] if self.rank == 0 else None

    def close(self):
        self.barrier()
        self.gpu.close()


def install_volume_from_file(ctx, gpu, path):
    """--volume-file: rank 0 decodes the file (vr5_load_volume: PVM / DDS, 16 -> 8 bit quantiser), uploads it, and the
    DEVICE buffer is broadcast to the other ranks over NCCL (volume replicated from rank 0); every rank
    installs it with vr5_set_volume_device."""
    import torch
    R, dist, rank, world, dev = ctx["R"], ctx["dist"], ctx["rank"], ctx["world"], ctx["dev"]
    dims = [0, 0, 0]
    vol = None
    if rank == 0:
        vol = R.load_volume(path)
        dims = [int(vol.shape[2]), int(vol.shape[1]), int(vol.shape[0])]
    if world > 1:
        dist.broadcast_object_list(dims, src=0)
    buf = torch.empty(dims[0] * dims[1] * dims[2], dtype=torch.uint8, device=dev)
    if rank == 0:
        buf.copy_(torch.from_numpy(vol.reshape(-1)))
    if world > 1:
        dist.broadcast(buf, src=0)
    torch.cuda.synchronize(dev)
    gpu.set_volume_device(buf.data_ptr(), buf.numel(), dims[0], dims[1], dims[2])
    return {"file": os.path.basename(path), "dims": tuple(dims), "sha256": sha256(gpu.get_volume((dims[2], dims[1], dims[0]))),
            "route": "rank 0 vr5_load_volume -> device buffer -> %s -> vr5_set_volume_device" % ("NCCL broadcast" if world > 1 else "same device")}


def oracle_rows_check(wl, params_list, frames, rows, ctx, volume=None, sampler="hybridw"):
    """rank 0: `rows` of each frame against the CPU checker: max LSB difference, rows checked. Nearest-neighbour policies: the
    reference's own CPURenderer when it is prebuilt and its 32-bit indices suffice, else the restated port. Trilinear policies:
    the matching section of the restated checker (their external anchor, GPURenderer4, needs a GPU: tests/test_trilinear_pin.py)."""
    cpu = CpuReference(wl, volume=volume)
    tri = {"trilinear": "orc_render_rows_trilinear", "trilinear-fma": "orc_render_rows_trilinear_fma"}.get(sampler)
    if tri:
        cpu.kind = "port (%s section)" % sampler
    worst, checked = 0, 0
    for (ang, frame) in zip(params_list, frames):
        sc = cpu.scene(ang)
        out = np.zeros((wl["h"], wl["w"], 4), dtype=np.uint8)
        spans = [(r, r + 1) for r in rows]
        if tri:
            from concurrent.futures import ThreadPoolExecutor
            fn = getattr(cpu.port.lib, tri)
            with ThreadPoolExecutor(max_workers=os.cpu_count() or 1) as ex:
                list(ex.map(lambda sp: fn(C.byref(sc.rc), out.ctypes.data_as(C.POINTER(C.c_uint8)), sp[0], sp[1], None), spans))
        else:
            cpu.render_rows(sc, out, spans, os.cpu_count() or 1)
        for r in rows:
            worst = max(worst, int(np.abs(out[r].astype(np.int16) - frame[r].astype(np.int16)).max()))
            checked += 1
    return {"oracle_rows": checked, "max_lsb": worst, "oracle_kind": cpu.kind}


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
    stream = torch.cuda.Stream(device=dev)  # a real (non-default) stream shared by the kernels, NCCL and the timing events
    torch.cuda.set_stream(stream)
    ctx = dict(R=R, dist=dist, rank=rank, world=world, local_rank=local_rank, dev=dev, stream=stream, volume_file=args.volume_file, partition=args.partition)
    L = R.lib()
    failures = []

    n, w, h = wl["n"], wl["w"], wl["h"]
    rig = Rig(ctx, wl, args.sampler, args.overlap)
    gpu, lanes = rig.gpu, args.overlap
    frames = [rig.params(i, args.steps) for i in range(args.steps)]
    frame_angles = [orbit_angles(i, args.steps)[1] for i in range(args.steps)]
    warm = [rig.params(i * 7 + 3, ORBIT_FRAMES) for i in range(max(args.warmup, 3))]

    # -- work counters per frame (instrumented kernel, untimed) ------------------------------------------------------
    (samples_local, shaded_local), (samples_total, shaded_total) = rig.count(frames)

    # -- timed region: exactly K steps = ONE sequence launch per rank, device timed, max over ranks -------------------
    clocks = ClockSampler(local_rank)
    clocks.start()
    ms, ms_local = rig.timed_sequence(frames, warm)
    lt = torch.tensor([rig.launches], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(lt)
    launches = int(lt[0])
    # keep the same load running (untimed) until the clock sampler has seen >= 1 s of it; the number of repeats is derived
    # from the reduced time, so every rank runs the same count (the ring needs equal frame sequences on all ranks)
    soak_events = []
    for _ in range(int(max(0.0, 1.0 - ms / 1e3) / max(ms / 1e3, 1e-6)) + 1):
        ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ea.record(stream)
        rig.sequence(frames)
        eb.record(stream)
        soak_events.append((ea, eb))
    rig.barrier()
    clocks.stop()
    soak_ms = [a.elapsed_time(b) / args.steps for a, b in soak_events]   # this rank's view of the untimed repeats (diagnostic)
    gpu.peer_status()   # raises if any in-kernel or stream wait timed out

    # -- the same frames through the per-frame API of round 1 (one launch per frame, frame lanes), for comparison ------
    def per_frame_loop(fr):
        for p in fr:
            gpu.peer_render(p, rig.seq)
            if rank == 0:
                gpu.peer_wait(rig.seq)
                gpu.peer_release(rig.seq)
            rig.seq += 1
        gpu.join()
    per_frame_loop(warm)
    rig.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    rig.barrier()
    launches0 = gpu.launch_count()
    e0.record(stream)
    per_frame_loop(frames)
    e1.record(stream)
    launches_frames = gpu.launch_count() - launches0
    rig.barrier()
    ms_frames_local = e0.elapsed_time(e1)
    t = torch.tensor([ms_frames_local], dtype=torch.float64, device=dev)
    lt = torch.tensor([launches_frames], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(lt)
    ms_per_frame_api, launches_frames = float(t[0]), int(lt[0])
    gpu.peer_status()
    # Which submission is the headline. Both are timed over exactly the same K frames with the same barriers and events, both are
    # in the line. auto: the sequence launch on one GPU (same rate as frame lanes, one launch instead of ~3 per frame); one launch
    # per frame over frame lanes on several GPUs - measured faster there (N = 8, C3: 0.079 against 0.102 ms per frame): a rank's
    # share of a frame is short, and the per-CTA hand-shake inside the sequence kernel costs ~40 us per frame
    submission = args.submission if args.submission != "auto" else ("sequence" if world == 1 else "frames")
    ms_sequence, launches_sequence = ms, launches
    if submission == "frames":
        ms, launches = ms_per_frame_api, launches_frames

    # -- N > 1: the plain-collective baseline (packed strips -> dist.gather over NCCL -> unpack kernel on rank 0) ----------
    nccl_ms = None
    if world > 1:
        rows_padded = max(L.vr5_strip_rows(h, r, world) for r in range(world))
        packed = torch.empty((rows_padded, w, 4), dtype=torch.uint8, device=dev)
        gathered = torch.empty((world, rows_padded, w, 4), dtype=torch.uint8, device=dev) if rank == 0 else None
        frame_dev = torch.empty((h, w, 4), dtype=torch.uint8, device=dev)
        gpu.set_overlap(1)
        gpu.set_partition(rank, world, True)

        def nccl_loop(fr):
            for p in fr:
                gpu.render_device(p, packed.data_ptr(), blocking=False)
                dist.gather(packed, list(gathered.unbind(0)) if rank == 0 else None, dst=0)
                if rank == 0:
                    gpu.unpack_strips(gathered.data_ptr(), rows_padded, world, frame_dev.data_ptr())
        nccl_loop(warm)
        rig.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        nccl_loop(frames)
        e1.record(stream)
        rig.barrier()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        nccl_ms = float(t[0])
        nccl_sha = sha256(frame_dev.cpu().numpy()) if rank == 0 else None   # last frame, for `verify`
        rig.apply_partition()
        gpu.set_overlap(lanes)
        del packed, gathered

    # -- end to end: frames land in HOST memory ------------------------------------------------------------------------
    e2e_gather = args.e2e_gather if world > 1 else "same"
    RING_SLOTS = 3 * max(1, lanes)            # the consumer trails the producers by RING_LAG frames
    RING_LAG = RING_SLOTS - max(1, lanes) - 1
    ring_mem, ring_name = None, None
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
        gpu.sync()
        if rig.partition != "bands":
            gpu.set_partition(rank, world, True)     # interleaved strips: packed for the strided copy (bands are copied as one block of rows)
        if rank == 0:
            gpu.hostring_attach(ring_mem, RING_SLOTS, init=True)    # clears the header before anyone else looks at it
        dist.barrier()
        if rank != 0:
            gpu.hostring_attach(ring_mem, RING_SLOTS, init=False)
        dist.barrier()
    pinneds = [torch.empty((h, w, 4), dtype=torch.uint8).pin_memory() for _ in range(max(lanes, 1))] if rank == 0 else None
    hseq, hrel = [0], [0]
    host_frames = {}     # verification: frame index -> sha of what arrived in host memory

    def e2e_loop(fr, keep=None):
        """The pipelined host path: frame i is enqueued (kernel + D2H on its lane; N > 1: this rank's strips into the
        shared host ring) and the host receives every frame, in order, a few frames behind."""
        if world == 1:
            gpu.set_overlap(lanes)
            for i, p in enumerate(fr):
                gpu.render_async_host(p, pinneds[i % lanes].data_ptr())
                if i >= lanes - 1:
                    gpu.sync_lane((i - (lanes - 1)) % lanes)
                    if keep is not None:
                        keep[i - (lanes - 1)] = sha256(pinneds[(i - (lanes - 1)) % lanes].numpy())
            for i in range(max(0, len(fr) - (lanes - 1)), len(fr)):
                gpu.sync_lane(i % lanes)
                if keep is not None:
                    keep[i] = sha256(pinneds[i % lanes].numpy())
        elif e2e_gather == "host":
            first = hseq[0]

            def take():
                ptr = gpu.hostring_wait(hrel[0])     # all ranks' strips of that frame are in host memory
                if keep is not None:
                    keep[hrel[0] - first] = sha256(np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint8)), shape=(h, w, 4)))
                gpu.hostring_release(hrel[0])
                hrel[0] += 1
            for p in fr:
                gpu.hostring_render(p, hseq[0])      # kernel + strided D2H of this rank's strips into the shared host frame
                hseq[0] += 1
                if rank == 0 and hseq[0] - hrel[0] > RING_LAG:
                    take()
            while rank == 0 and hrel[0] < hseq[0]:
                take()
        else:   # rank 0 copies the frame gathered in its HBM
            base = rig.seq
            for p in fr:
                i = rig.seq - base
                gpu.peer_render(p, rig.seq)
                if rank == 0:
                    gpu.peer_wait(rig.seq)
                    gpu.peer_read(rig.seq, pinneds[i % lanes].data_ptr())
                    gpu.peer_release(rig.seq)
                rig.seq += 1
                if i >= lanes - 1:
                    gpu.peer_sync_frame(rig.seq - lanes)
                    if keep is not None and rank == 0:
                        keep[i - (lanes - 1)] = sha256(pinneds[(i - (lanes - 1)) % lanes].numpy())
            gpu.sync()
            if keep is not None and rank == 0:
                for i in range(max(0, len(fr) - (lanes - 1)), len(fr)):
                    keep[i] = sha256(pinneds[i % lanes].numpy())

    e2e_loop(warm[:3])
    rig.barrier()
    t0 = time.perf_counter()
    e2e_loop(frames)
    rig.barrier()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t[0])
    VERIFY_K = min(3, len(frames))
    e2e_loop(frames[-VERIFY_K:], keep=host_frames)   # untimed: the frames the host path delivers, for `verify`
    rig.barrier()
    if e2e_gather == "host":
        gpu.hostring_detach()
        rig.apply_partition()
        ring_mem.close()
        dist.barrier()
        if rank == 0:
            R.ring_memory_unlink(ring_name)

    # the blocking plugin call (what VolR.cpp:110 does: Renderer::render_volume(buffer, raycaster), host buffer, one frame per call)
    e2e_blocking = None
    if world == 1:
        gpu.set_overlap(1)
        host = np.empty((h, w, 4), dtype=np.uint8)
        for p in warm[:3]:
            gpu.render_volume(p, host)
        t0 = time.perf_counter()
        for p in frames:
            gpu.render_volume(p, host)
        blk_s = time.perf_counter() - t0
        e2e_blocking = {"value": samples_total / blk_s / 1e9, "unit": UNIT, "fps": args.steps / blk_s, "ms_per_frame": 1e3 * blk_s / args.steps,
                        "path": "vr5_render(handle, params, host pointer, out_is_device = 0): blocking, one frame per call, one lane, pageable host buffer - "
                                "the call GPURenderer5::render_volume makes (VolR.cpp:110); inside, the frame is rendered as four cost-balanced bands "
                                "so that the copy of a band overlaps the kernels of the next ones"}
        gpu.set_overlap(lanes)

    # -- a single frame's launch alone (per-launch CUDA events on its stream): what one vr5_render call costs on the device ----
    scratch = torch.empty((h, w, 4), dtype=torch.uint8, device=dev)
    k_ms = gpu.render_timed(frames, scratch.data_ptr(), l2_flush=False)
    gpu.set_option("VR5_LEAP_STATS", "1")   # counted in an untimed pass of their own: the counters are two contended atomics per leap
    gpu.leap_stats()
    gpu.render_timed(frames, scratch.data_ptr(), l2_flush=False)
    leaps_k, left_out_k = gpu.leap_stats()   # over exactly these K frames of this rank
    gpu.set_option("VR5_LEAP_STATS", "0")
    k_ms_flushed = gpu.render_timed(frames[: min(len(frames), 24)], scratch.data_ptr(), l2_flush=True)
    full_ms = None
    if rank == 0 and world > 1:    # the whole frame on one GPU, for `ideal = full / N`
        gpu.set_partition(0, 1, False)
        full_ms = float(gpu.render_timed(frames, scratch.data_ptr(), l2_flush=False).mean())
        rig.apply_partition()
    strip_px = (L.vr5_strip_rows(h, rank, world) if world > 1 else h) * w   # (bands: the mean band is as tall as the strips together)
    voxel_bytes = 8.0 if args.sampler.startswith("trilinear") else 1.0   # nearest 1 B per fetch, trilinear 8 B
    alg_bytes_frame = voxel_bytes * (samples_local + shaded_local) / len(frames) + 4.0 * strip_px  # + 4 B per pixel written
    alg_bytes_launch = alg_bytes_frame * len(frames)
    peak, peak_src = measured_peak()
    if submission == "frames":   # the dominant kernel's launch is one frame of this rank
        alg_bytes_launch, roof_ms, roof_frames = alg_bytes_frame, float(k_ms.mean()), 1
    else:
        roof_ms, roof_frames = ms_local, len(frames)
    achieved = alg_bytes_launch / (roof_ms * 1e-3) / 1e9
    traffic, traffic_note = measured_traffic(args.workload, args.sampler, world, roof_frames)

    # -- verification (untimed): the frames of the timed paths against this GPU alone and against the CPU checker ------------
    vparams, vangles = frames[-VERIFY_K:], frame_angles[-VERIFY_K:]
    ring_frames = rig.frames_through_ring(vparams)
    verify = None
    if rank == 0:
        alone = rig.frames_alone(vparams)
        verify = {"frames": VERIFY_K,
                  "sequence_frames_equal": sum(sha256(a) == sha256(b) for a, b in zip(alone, ring_frames)),
                  "host_path_frames_equal": sum(host_frames.get(i) == sha256(a) for i, a in enumerate(alone)),
                  "what": "sha256 of the last %d orbit frames: (sequence path through the frame ring%s) and (end-to-end host path: %s) against "
                          "the same frames rendered by rank 0 alone with the blocking per-frame call" %
                          (VERIFY_K, " over NVLink peer stores" if world > 1 else "", e2e_gather)}
        if world > 1:
            verify["nccl_gather_last_frame_equal"] = nccl_sha == sha256(alone[-1])
            if not verify["nccl_gather_last_frame_equal"]:
                failures.append("verify: NCCL gather frame differs")
        if not args.volume_file:
            rows = [int(r) for r in np.linspace(h // 4, 3 * h // 4, 16)]
            verify.update(oracle_rows_check(wl, vangles[-1:], alone[-1:], rows, ctx, sampler=args.sampler))
        if verify["sequence_frames_equal"] != VERIFY_K or verify["host_path_frames_equal"] != VERIFY_K or verify.get("max_lsb", 0) != 0:
            failures.append("verify: %r" % verify)

    line = None
    if rank == 0:
        value = samples_total / (ms * 1e-3) / 1e9
        zero_skip = None
        if not args.sampler.startswith("trilinear") and gpu.last_zero_skip() == 1:
            zero_skip = {"leaps_per_frame": leaps_k / len(frames), "samples_left_out_per_frame": left_out_k / len(frames),
                         "fraction_of_samples_left_out": left_out_k / max(samples_local, 1),
                         "note": "the transfer function starts with all-zero entries: samples that classify to them skip their compositing, and runs of them "
                                 "inside bricks whose maximum is such a value are leapt over (vr5_leap_stats; this rank's share). `value` counts the "
                                 "REFERENCE's samples (the instrumented kernel's count, equal to the CPU checker's), not the fetches this kernel issues"}
        kernel_name = ("vr5::tri_kernel<variant %d>" % gpu.last_variant() if args.sampler.startswith("trilinear") else "vr5::cast_kernel<%s, variant %d%s>" % (args.sampler, gpu.last_variant(), ", zero-skip" if gpu.last_zero_skip() == 1 else ""))
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": len(warm),
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "fps": args.steps / (ms * 1e-3), "zero_skip": zero_skip,
            "ms_per_step_note": ("time of the K-frame sequence launch / K" if submission == "sequence" else "time of K frames submitted one launch each over %d frame lanes / K" % lanes) +
                                ": a throughput under overlap, not the latency of one frame; roofline.single_frame_launch_ms is the latter",
            "config": {"workload": wl["name"] if not args.volume_file else "%s %s" % (rig.volume_route["file"], "x".join(map(str, rig.volume_route["dims"]))),
                       "volume": "%d^3" % n, "viewport": "%dx%d" % (w, h), "tf": wl["tf"],
                       "projection": "orthographic", "sampler": args.sampler, "samples_per_frame": samples_total / args.steps,
                       "submission": ("one persistent sequence launch per rank for all K frames (vr5_peer_render_sequence), frames consumed in order from a ring of %d device frames" % rig.slots) if submission == "sequence" else
                                     ("one launch per frame per rank (vr5_peer_render; rank 0: vr5_peer_wait + vr5_peer_release) round-robin over %d frame lanes, ring of %d device frames" % (lanes, rig.slots)),
                       "submission_choice": "%s (--submission %s)" % (submission, args.submission),
                       "partition": (rig.partition if world > 1 else "none"),
                       "parallelism": ("image space x%d (%s), volume replicated, pixels stored straight into rank 0's frame ring over NVLink peer memory (fused render + gather, hand-shake inside the kernel)" % (world, "cost-balanced contiguous bands of 4-row strips, re-balanced per frame" if rig.partition == "bands" else "interleaved 4-row strips")) if world > 1 else "single GPU",
                       "l2": "no flush in the timed loop: per frame the kernel reads the %.0f MB volume and writes the %.1f MB frame (L2 is 126 MB); "
                             "the flushed per-kernel time is reported in roofline.single_frame_launch_ms_l2_flushed" % (n ** 3 / 1e6, w * h * 4 / 1e6)},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_note": traffic_note,
                         "peak_source": peak_src, "kernel": kernel_name,
                         "launch": "one launch = the K-frame sequence of the timed region (CUDA events on its stream)" if submission == "sequence" else "one launch = this rank's share of one frame, alone on its stream (per-launch CUDA events, mean over the K frames)",
                         "bytes_per_sample": voxel_bytes, "kernel_ms": roof_ms, "frames_per_launch": roof_frames, "algorithmic_bytes_per_launch": alg_bytes_launch,
                         "single_frame_launch_ms": float(k_ms.mean()), "single_frame_launch_frac": alg_bytes_frame / (float(k_ms.mean()) * 1e-3) / 1e9 / peak,
                         "single_frame_launch_ms_l2_flushed": float(k_ms_flushed.mean()),
                         "note": ("trilinear policy, contracted arithmetic (GPURenderer4 semantics as nvcc -fmad=true builds them, NOT the CPURenderer parity mode): 8 B per sample" if args.sampler == "trilinear-fma" else
                                  "trilinear policy (GPURenderer4 semantics, NOT the CPURenderer parity mode): 8 B per sample" if args.sampler == "trilinear" else
                                  "parity mode fetches 1 B per sample; the kernel is bound by the fp32 pipe and issue slots, not by HBM (DESIGN.md)")},
            "sequence_api": {"ms_per_step": ms_sequence / args.steps, "value": samples_total / (ms_sequence * 1e-3) / 1e9, "gpu_launches": launches_sequence,
                             "path": "one persistent launch per rank for all K frames, hand-shake inside the kernel (vr5_peer_render_sequence / vr5_peer_consume_sequence)"},
            "per_frame_api": {"gpu_launches": launches_frames, "ms_per_step": ms_per_frame_api / args.steps, "value": samples_total / (ms_per_frame_api * 1e-3) / 1e9, "frame_overlap_lanes": lanes,
                              "path": "round-1 submission: one launch per frame (vr5_peer_render + wait + release), frames round-robin over %d lanes" % lanes},
            "e2e": {"value": samples_total / e2e_s / 1e9, "unit": UNIT, "fps": args.steps / e2e_s, "h2d_bytes_per_step": C.sizeof(R.Params),
                    "d2h_bytes_per_step": w * h * 4,
                    "path": {"host": "every rank copies its strips into a frame ring in shared host memory over its own PCIe link (vr5_hostring_*), %d lanes" % lanes,
                             "peer": "rank 0 copies the frame gathered in its HBM (vr5_peer_read)",
                             "same": "pipelined: vr5_render_async (kernel + D2H into pinned host memory on the frame's lane) with vr5_set_overlap(%d); the host receives every frame in order, %d frames behind" % (lanes, lanes - 1)}[e2e_gather]},
            "verify": verify,
            "repeats_ms_per_step": {"n": len(soak_ms), "min": min(soak_ms), "median": float(np.median(soak_ms)), "max": max(soak_ms),
                                    "note": "the same K-frame sequence repeated back to back after the timed one (untimed load for the clock sampler), rank 0's events"},
            "gpu_launches": launches, "clocks": clocks.summary(),
        }
        if e2e_blocking:
            line["e2e_blocking"] = e2e_blocking
        if world > 1:
            line["nccl_gather_baseline"] = {"ms_per_step": nccl_ms / args.steps, "value": samples_total / (nccl_ms * 1e-3) / 1e9,
                                            "path": "packed strips, torch.distributed.gather over NCCL to rank 0, vr5_unpack_strips; serial frames on one stream"}
            line["kernel_vs_ideal"] = {"single_frame_launch_ms_this_rank": float(k_ms.mean()), "full_frame_launch_ms_one_gpu": full_ms, "ideal_ms": full_ms / world,
                                       "sequence_ms_per_frame": ms / args.steps}
        if rig.volume_route:
            line["volume_route"] = rig.volume_route
    rig.close()

    # -- the contracted trilinear policy (8 B per sample) in the same run: the row the HBM roofline is meaningful for -------
    if world == 1 and not args.no_policies and not args.sampler.startswith("trilinear") and not args.volume_file:
        rec = policy_record(ctx, wl, "trilinear-fma", args, peak)
        if rank == 0:
            line["policies"] = {"trilinear_fma": rec}
    # -- C5 (2048^3, 7680x4320) in the same run: north_star's 1 -> 8 GPU scaling target is quoted on it --------------------
    if not args.no_c5 and args.workload == "c3" and not args.volume_file:
        rec = c5_record(ctx, args)
        if rank == 0:
            line["c5"] = rec
            if rec.get("verify", {}).get("ok") is False:
                failures.append("c5 verify: %r" % rec["verify"])
    if world > 1:
        dist.barrier()
    # -- CPU baseline next to it (rank 0, N=1 only): the reference's CPURenderer as shipped = 1 thread, one orbit frame ----
    if rank == 0 and world == 1 and not args.no_cpu_baseline and not args.volume_file:
        line["cpu_baseline"] = cpu_baseline(args, wl)
    if rank == 0:
        if failures:
            line["failures"] = failures
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    if failures:
        raise SystemExit("bench.py: verification failed: " + "; ".join(failures))


def measured_traffic(workload, sampler, world, frames_per_launch):
    """DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) from the committed ncu capture of the SAME
    (workload, sampler, GPUs) configuration: profiles/traffic.json holds bytes per FRAME of a sequence launch, keyed
    "<workload>_<sampler>_n<gpus>". No capture for the configuration -> null with the reason (never another configuration's)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            per_frame = json.load(f).get("%s_%s_n%d" % (workload, sampler, world))
    except Exception:
        per_frame = None
    if per_frame is None:
        return None, "no ncu capture committed for (%s, %s, %d GPUs); ncu runs single-process only" % (workload, sampler, world)
    return per_frame * frames_per_launch, "profiles/traffic.json: %.1f MB per frame (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum of a sequence launch / its frames) x %d frames" % (per_frame / 1e6, frames_per_launch)


def policy_record(ctx, wl, sampler, args, peak):
    """Sub-record of another sampler policy on the same workload, measured like the headline (N = 1)."""
    rig = Rig(ctx, wl, sampler, args.overlap)
    frames = [rig.params(i, args.steps) for i in range(args.steps)]
    warm = [rig.params(i * 7 + 3, ORBIT_FRAMES) for i in range(max(args.warmup, 3))]
    (s_loc, sh_loc), (s_tot, sh_tot) = rig.count(frames)
    ms, ms_local = rig.timed_sequence(frames, warm)
    import torch
    scratch = torch.empty((wl["h"], wl["w"], 4), dtype=torch.uint8, device=ctx["dev"])
    k_ms = rig.gpu.render_timed(frames, scratch.data_ptr(), l2_flush=False)
    alg_frame = 8.0 * (s_loc + sh_loc) / len(frames) + 4.0 * wl["w"] * wl["h"]
    traffic, note = measured_traffic(args.workload, sampler, 1, len(frames))
    rec = {"sampler": sampler, "ms_per_step": ms / len(frames), "value": s_tot / (ms * 1e-3) / 1e9, "unit": UNIT, "bytes_per_sample": 8.0,
           "kernel": "vr5::tri_kernel<variant %d>" % rig.gpu.last_variant(), "kernel_ms": ms_local, "frames_per_launch": len(frames),
           "achieved": alg_frame * len(frames) / (ms_local * 1e-3) / 1e9, "frac": alg_frame * len(frames) / (ms_local * 1e-3) / 1e9 / peak,
           "single_frame_launch_ms": float(k_ms.mean()), "single_frame_launch_frac": alg_frame / (float(k_ms.mean()) * 1e-3) / 1e9 / peak,
           "dram_bytes_measured": traffic, "dram_note": note,
           "parity": TRILINEAR_PARITY_NOTE}
    rig.close()
    return rec


TRILINEAR_PARITY_NOTE = ("NOT the CPURenderer parity mode. Pinned against the reference's GPURenderer4.cu compiled unmodified (texture references over texture objects, "
                         "oracle/ref_gpu4_texobj.cu; hardware-filtered tex3D / tex1D) under the tolerance stated in tests/test_trilinear_pin.py; bit-equal to the restated "
                         "fp32 checker, which an independent float64 restatement of the CUDA filtering formulas confirms (tests/test_trilinear_independent.py)")


def c5_record(ctx, args):
    """BASELINE.json configs[4]: synthA 2048^3 (8 GiB, replicated), 7680x4320, semi TF - a few orbit frames through the same
    sequence path, every rank. north_star's >= 6x from 1 to 8 GPUs is quoted on this configuration."""
    wl = WORKLOADS["c5"]
    K = max(8, min(args.c5_frames, 64))
    rig = Rig(ctx, wl, args.sampler, args.overlap)
    frames = [rig.params(i, K) for i in range(K)]
    angles = [orbit_angles(i, K)[1] for i in range(K)]
    warm = [rig.params(i * 7 + 3, ORBIT_FRAMES) for i in range(3)]
    (s_loc, sh_loc), (s_tot, sh_tot) = rig.count(frames)
    ms, ms_local = rig.timed_sequence(frames, warm)
    ring = rig.frames_through_ring(frames[-2:])
    rec = None
    if rig.rank == 0:
        alone = rig.frames_alone(frames[-2:])
        ver = {"frames": 2, "sequence_frames_equal": sum(sha256(a) == sha256(b) for a, b in zip(alone, ring))}
        try:
            with open("/proc/meminfo") as f:
                avail_gib = next(int(l.split()[1]) for l in f if l.startswith("MemAvailable")) / 2 ** 20
        except Exception:
            avail_gib = 0.0
        if rig.world == 1 and avail_gib > 28.0:   # CPU rows need the 8-GiB volume on the host: the bytes the kernel sees
            vol = rig.gpu.get_volume((wl["n"],) * 3)
            rows = [int(r) for r in np.linspace(wl["h"] // 4, 3 * wl["h"] // 4, 8)]
            ver.update(oracle_rows_check(wl, angles[-1:], alone[-1:], rows, ctx, volume=vol, sampler=args.sampler))
            del vol
        else:
            ver.update({"oracle_rows": 0, "oracle_note": "CPU rows at 2048^3 are checked at N = 1 with > 28 GiB of free host memory (and in tests/test_gpu_full_size.py)"})
        ver["ok"] = ver["sequence_frames_equal"] == 2 and ver.get("max_lsb", 0) == 0
        rec = {"workload": wl["name"], "n_frames": K, "ms_per_step": ms / K, "fps": K / (ms * 1e-3), "gsamples": s_tot / (ms * 1e-3) / 1e9,
               "samples_per_frame": s_tot / K, "kernel": "vr5::cast_kernel<%s, variant %d%s>" % (args.sampler, rig.gpu.last_variant(), ", zero-skip" if rig.gpu.last_zero_skip() == 1 else ""), "verify": ver}
    rig.close()
    return rec


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
    ap.add_argument("--submission", default="auto", choices=["auto", "sequence", "frames"],
                    help="which submission the headline value is taken from (both are always measured): auto = sequence on one GPU, frames on several")
    ap.add_argument("--partition", default="auto", choices=["auto", "bands", "strips"],
                    help="N > 1, device-resident path: cost-balanced contiguous bands or interleaved 4-row strips (auto: strips for an opaque TF)")
    ap.add_argument("--no-policies", action="store_true", help="skip the policies.trilinear_fma sub-record (N = 1)")
    ap.add_argument("--no-c5", action="store_true", help="skip the c5 sub-record (2048^3 / 7680x4320)")
    ap.add_argument("--c5-frames", type=int, default=8)
    ap.add_argument("--volume-file", default=None, help="render a real dataset (.pvm): rank 0 decodes it, the device buffer is broadcast over NCCL")
    ap.add_argument("--e2e-gather", default="host", choices=["host", "peer"],
                    help="N > 1, end-to-end leg: host = every rank copies its strips into a shared host frame ring; peer = rank 0 copies the gathered frame")
    args = ap.parse_args()
    claim_stdout()
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) != 0:
            return  # under torchrun only rank 0 runs the CPU arm
        run_reference_arm(args, wl)
    else:
        run_b200_arm(args, wl)


if __name__ == "__main__":
    main()
