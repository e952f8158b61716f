"""ctypes bindings of the two CPU checkers. TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may
import this package. The product (``libvr5.so`` and everything under ``volume-rendering_b200/``) never does.

* :class:`Port`  - ``liboracle.so``: our C restatement of the reference algorithm (``oracle_port.c``) and the
  host generator of the synthetic volumes (``synth_volume.c``).
* :class:`Ref`   - ``_ref/libvolr_ref.so``: the reference's own CPURenderer / RaycasterBase / ViewBase /
  ModelBase / ddsbase compiled unmodified (``oracle/Makefile``). Present wherever it was prebuilt.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "liboracle.so")
REF_SO = os.path.join(HERE, "_ref", "libvolr_ref.so")
REF_GPU_SO = os.path.join(HERE, "_ref", "libvolr_ref_gpu.so")


def build(quiet: bool = True) -> None:
    """Compile liboracle.so and, where /root/reference exists, _ref/libvolr_ref.so (oracle/Makefile)."""
    subprocess.run(["make", "-C", HERE], check=True, stdout=subprocess.DEVNULL if quiet else None)


class F3(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("z", C.c_float)]

    def tup(self):
        return (self.x, self.y, self.z)


class Model(C.Structure):  # include/vr5_pod.h vr5_model == reference ModelBase.h:11-16
    _fields_ = [("data", C.c_void_p), ("size", C.c_uint32), ("dims", C.c_uint16 * 3), ("min_bound", F3)]


class View(C.Structure):  # vr5_view == reference ViewBase.h:14-21
    _fields_ = [("dims", C.c_uint16 * 2), ("origin", F3), ("direction", F3), ("right_plane", F3),
                ("up_plane", F3), ("light_pos", F3), ("perspective", C.c_uint8), ("_pad", C.c_uint8 * 3)]


class Raycaster(C.Structure):  # vr5_raycaster == reference RaycasterBase.h:20-30
    _fields_ = [("volume", Model), ("view", View), ("_pad0", C.c_uint32), ("transfer_fn", C.c_void_p),
                ("ray_step", C.c_float), ("ray_threshold", C.c_float), ("esl", C.c_uint8), ("_pad1", C.c_uint8 * 7),
                ("esl_volume", C.c_void_p), ("esl_block_dims", C.c_uint16), ("_pad2", C.c_uint16),
                ("esl_block_size", F3), ("light_kd", C.c_float), ("_pad3", C.c_uint32)]


assert C.sizeof(Model) == 32 and C.sizeof(View) == 68 and C.sizeof(Raycaster) == 160


class Counters(C.Structure):
    _fields_ = [("rays_hit", C.c_uint64), ("rays_accum", C.c_uint64), ("esl_iters", C.c_uint64),
                ("samples", C.c_uint64), ("shaded", C.c_uint64)]

    def as_dict(self):
        return {k: int(getattr(self, k)) for k, _ in self._fields_}


def _u8p(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint8))


class Scene:
    """Owns the host arrays a Raycaster points at (volume, pre-multiplied TF, ESL bits) and the POD itself."""

    def __init__(self, volume: np.ndarray, tf_premult: np.ndarray, esl_bits: np.ndarray, rc: Raycaster):
        self.volume = np.ascontiguousarray(volume, dtype=np.uint8)
        self.tf = np.ascontiguousarray(tf_premult, dtype=np.float32).reshape(128, 4)
        self.esl_bits = np.ascontiguousarray(esl_bits, dtype=np.uint32).reshape(1024)
        self.rc = Raycaster.from_buffer_copy(bytes(rc))
        self.rc.volume.data = self.volume.ctypes.data
        self.rc.transfer_fn = self.tf.ctypes.data
        self.rc.esl_volume = self.esl_bits.ctypes.data

    @property
    def width(self):
        return int(self.rc.view.dims[0])

    @property
    def height(self):
        return int(self.rc.view.dims[1])

    def with_view(self, view: View) -> "Scene":
        rc = Raycaster.from_buffer_copy(bytes(self.rc))
        rc.view = View.from_buffer_copy(bytes(view))
        return Scene(self.volume, self.tf, self.esl_bits, rc)


class Port:
    """liboracle.so - the restated algorithm."""

    def __init__(self):
        if not os.path.exists(PORT_SO):
            build()
        self.lib = C.CDLL(PORT_SO)
        L = self.lib
        L.orc_render_rows.argtypes = [C.POINTER(Raycaster), C.POINTER(C.c_uint8), C.c_int, C.c_int, C.POINTER(Counters)]
        L.orc_render_rows.restype = C.c_int
        L.orc_render_rows_trilinear.argtypes = L.orc_render_rows.argtypes
        L.orc_render_rows_trilinear.restype = C.c_int
        L.orc_render_rows_trilinear_fma.argtypes = L.orc_render_rows.argtypes
        L.orc_render_rows_trilinear_fma.restype = C.c_int
        L.orc_default_base_tf.argtypes = [C.POINTER(C.c_float)]
        L.orc_premultiply_tf.argtypes = [C.POINTER(C.c_float), C.POINTER(C.c_float)]
        L.orc_build_esl_bits.argtypes = [C.POINTER(C.c_float), C.POINTER(C.c_uint8), C.POINTER(C.c_uint32)]
        L.orc_build_minmax.argtypes = [C.POINTER(C.c_uint8), C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint16),
                                       C.POINTER(C.c_uint8), C.POINTER(C.c_float)]
        L.orc_default_ray_step.argtypes = [C.c_uint32]
        L.orc_default_ray_step.restype = C.c_float
        L.orc_histogram.argtypes = [C.POINTER(C.c_uint8), C.c_uint64, C.POINTER(C.c_float)]
        L.orc_synth_a.argtypes = [C.POINTER(C.c_uint8), C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
        L.orc_synth_b.argtypes = [C.POINTER(C.c_uint8), C.c_uint32, C.c_uint32]
        L.orc_synth_cos_table.argtypes = [C.POINTER(C.c_int16)]

    # -- rendering ---------------------------------------------------------------------------------------
    def render(self, scene: Scene, rows=None, threads: int | None = None, trilinear=False):
        """Frame (or row range) as uint8 [H, W, 4] plus summed counters; rows are split over host threads.
        trilinear=True: the restated GPURenderer4 sampler policy (fp32 trilinear volume fetch, linear TF), one rounding
        per operation; trilinear="fma": the same policy with the contractions nvcc applies to the reference's GPU
        renderers written out as fused multiply-adds."""
        fn = (self.lib.orc_render_rows_trilinear_fma if trilinear == "fma" else
              self.lib.orc_render_rows_trilinear if trilinear else self.lib.orc_render_rows)
        w, h = scene.width, scene.height
        row0, row1 = (0, h) if rows is None else rows
        out = np.zeros((h, w, 4), dtype=np.uint8)
        threads = threads or min(os.cpu_count() or 1, 64)
        chunk = max(1, (row1 - row0 + threads * 4 - 1) // (threads * 4))
        spans = [(r, min(r + chunk, row1)) for r in range(row0, row1, chunk)]

        def work(span):
            cnt = Counters()
            rc = fn(C.byref(scene.rc), _u8p(out), span[0], span[1], C.byref(cnt))
            if rc != 0:
                raise RuntimeError("orc_render_rows returned %d" % rc)
            return cnt.as_dict()

        with ThreadPoolExecutor(max_workers=threads) as ex:
            parts = list(ex.map(work, spans))
        total = {k: sum(p[k] for p in parts) for k in parts[0]} if parts else Counters().as_dict()
        return out, total

    # -- scene builders ----------------------------------------------------------------------------------
    def default_base_tf(self) -> np.ndarray:
        tf = np.zeros((128, 4), dtype=np.float32)
        self.lib.orc_default_base_tf(tf.ctypes.data_as(C.POINTER(C.c_float)))
        return tf

    def premultiply(self, base: np.ndarray) -> np.ndarray:
        base = np.ascontiguousarray(base, dtype=np.float32)
        out = np.zeros((128, 4), dtype=np.float32)
        self.lib.orc_premultiply_tf(base.ctypes.data_as(C.POINTER(C.c_float)), out.ctypes.data_as(C.POINTER(C.c_float)))
        return out

    def minmax(self, volume: np.ndarray):
        dz, dy, dx = volume.shape
        bd = C.c_uint16()
        mm = np.zeros((32768, 2), dtype=np.uint8)
        bs = np.zeros(3, dtype=np.float32)
        self.lib.orc_build_minmax(_u8p(volume), dx, dy, dz, C.byref(bd), _u8p(mm), bs.ctypes.data_as(C.POINTER(C.c_float)))
        return int(bd.value), mm, bs

    def esl_bits(self, premult: np.ndarray, minmax: np.ndarray) -> np.ndarray:
        bits = np.zeros(1024, dtype=np.uint32)
        premult = np.ascontiguousarray(premult, dtype=np.float32)
        self.lib.orc_build_esl_bits(premult.ctypes.data_as(C.POINTER(C.c_float)), _u8p(minmax),
                                    bits.ctypes.data_as(C.POINTER(C.c_uint32)))
        return bits

    def ray_step(self, max_dim: int) -> float:
        return float(self.lib.orc_default_ray_step(max_dim))

    def histogram(self, volume: np.ndarray) -> np.ndarray:
        h = np.zeros(256, dtype=np.float32)
        self.lib.orc_histogram(_u8p(volume), volume.size, h.ctypes.data_as(C.POINTER(C.c_float)))
        return h

    def make_scene(self, volume: np.ndarray, base_tf: np.ndarray, view: View, *, ray_step=None, threshold=0.95,
                   esl=True, light_kd=0.6) -> Scene:
        """RaycasterBase::set_volume + update_transfer_fn + set_view restated (RaycasterBase.cpp:46-129)."""
        volume = np.ascontiguousarray(volume, dtype=np.uint8)
        dz, dy, dx = volume.shape
        bd, mm, bs = self.minmax(volume)
        tf = self.premultiply(base_tf)
        bits = self.esl_bits(tf, mm)
        rc = Raycaster()
        rc.volume.size = volume.size & 0xFFFFFFFF
        rc.volume.dims[0], rc.volume.dims[1], rc.volume.dims[2] = dx, dy, dz
        rc.volume.min_bound = F3(-1, -1, -1)
        rc.view = View.from_buffer_copy(bytes(view))
        rc.ray_step = self.ray_step(max(dx, dy, dz)) if ray_step is None else ray_step
        rc.ray_threshold = threshold
        rc.esl = 1 if esl else 0
        rc.esl_block_dims = bd
        rc.esl_block_size = F3(*[float(v) for v in bs])
        rc.light_kd = light_kd
        return Scene(volume, tf, bits, rc)

    # -- synthetic volumes -------------------------------------------------------------------------------
    def synth_a(self, n: int, seed: int = 1, threads: int | None = None) -> np.ndarray:
        out = np.empty((n, n, n), dtype=np.uint8)
        threads = threads or min(os.cpu_count() or 1, 64)
        step = max(1, (n + threads - 1) // threads)
        with ThreadPoolExecutor(max_workers=threads) as ex:
            list(ex.map(lambda z0: self.lib.orc_synth_a(_u8p(out), n, seed, z0, min(z0 + step, n)), range(0, n, step)))
        return out

    def synth_b(self, n: int, seed: int = 2) -> np.ndarray:
        out = np.empty((n, n, n), dtype=np.uint8)
        self.lib.orc_synth_b(_u8p(out), n, seed)
        return out

    def cos_table(self) -> np.ndarray:
        t = np.zeros(1024, dtype=np.int16)
        self.lib.orc_synth_cos_table(t.ctypes.data_as(C.POINTER(C.c_int16)))
        return t


class Ref:
    """_ref/libvolr_ref.so - the reference's own code. Process-global state (the reference uses statics)."""

    def __init__(self):
        if not os.path.exists(REF_SO):
            if os.path.isdir("/root/reference/VolumeRendering"):
                build()
            if not os.path.exists(REF_SO):
                raise FileNotFoundError(REF_SO + " (built only where /root/reference exists)")
        self.lib = C.CDLL(REF_SO)
        L = self.lib
        L.ref_load_model.argtypes = [C.c_char_p]
        L.ref_set_volume.argtypes = [C.POINTER(C.c_uint8), C.c_uint, C.c_uint, C.c_uint]
        L.ref_get_volume_dims.argtypes = [C.POINTER(C.c_uint)] * 3
        L.ref_get_volume_data.restype = C.c_void_p
        for name in ("ref_set_base_transfer_fn", "ref_get_base_transfer_fn", "ref_get_transfer_fn"):
            getattr(L, name).argtypes = [C.POINTER(C.c_float)]
        L.ref_get_esl_volume.argtypes = [C.POINTER(C.c_uint32)]
        L.ref_set_ray_step.argtypes = [C.c_float]
        L.ref_set_ray_threshold.argtypes = [C.c_float]
        L.ref_set_esl.argtypes = [C.c_int]
        L.ref_set_light_kd.argtypes = [C.c_float]
        L.ref_get_ray_step_limits.argtypes = [C.POINTER(C.c_float)] * 2
        L.ref_set_viewport.argtypes = [C.c_uint, C.c_uint]
        L.ref_set_perspective.argtypes = [C.c_int]
        L.ref_set_camera_position.argtypes = [C.c_float] * 4
        L.ref_get_raycaster.argtypes = [C.POINTER(Raycaster)]
        L.ref_render.argtypes = [C.POINTER(Raycaster), C.POINTER(C.c_uint8)]
        L.ref_render.restype = C.c_int
        L.ref_render_rows.argtypes = [C.POINTER(Raycaster), C.POINTER(C.c_uint8), C.c_int, C.c_int]
        L.ref_render_rows.restype = C.c_int
        assert L.ref_sizeof_raycaster() == 160 and L.ref_sizeof_view() == 68 and L.ref_sizeof_model() == 32

    def read_pvm(self, path: str) -> np.ndarray:
        """The reference's own readPVMvolume: voxels [D, H, W, components]."""
        w, h, d, nc = C.c_uint(), C.c_uint(), C.c_uint(), C.c_uint()
        self.lib.ref_read_pvm.argtypes = [C.c_char_p, C.POINTER(C.c_uint8), C.POINTER(C.c_uint), C.POINTER(C.c_uint), C.POINTER(C.c_uint), C.POINTER(C.c_uint)]
        if self.lib.ref_read_pvm(path.encode(), None, C.byref(w), C.byref(h), C.byref(d), C.byref(nc)) != 0:
            raise RuntimeError("reference readPVMvolume failed on " + path)
        out = np.zeros((d.value, h.value, w.value, nc.value), dtype=np.uint8)
        self.lib.ref_read_pvm(path.encode(), _u8p(out), C.byref(w), C.byref(h), C.byref(d), C.byref(nc))
        return out

    def quantize16(self, data16_be: np.ndarray, w: int, h: int, d: int, linear: bool = False) -> np.ndarray:
        """The reference's own quantize()."""
        src = np.ascontiguousarray(data16_be, dtype=np.uint8)
        out = np.zeros((d, h, w), dtype=np.uint8)
        self.lib.ref_quantize16.argtypes = [C.POINTER(C.c_uint8), C.c_uint, C.c_uint, C.c_uint, C.c_int, C.POINTER(C.c_uint8)]
        if self.lib.ref_quantize16(_u8p(src), w, h, d, 1 if linear else 0, _u8p(out)) != 0:
            raise RuntimeError("reference quantize failed")
        return out

    def load_model(self, path: str) -> int:
        return self.lib.ref_load_model(path.encode())

    def set_volume(self, volume: np.ndarray) -> None:
        volume = np.ascontiguousarray(volume, dtype=np.uint8)
        dz, dy, dx = volume.shape
        if self.lib.ref_set_volume(_u8p(volume), dx, dy, dz) != 0:
            raise MemoryError

    def volume(self) -> np.ndarray:
        d = [C.c_uint() for _ in range(3)]
        self.lib.ref_get_volume_dims(*[C.byref(v) for v in d])
        dx, dy, dz = [v.value for v in d]
        buf = (C.c_uint8 * (dx * dy * dz)).from_address(self.lib.ref_get_volume_data())
        return np.frombuffer(buf, dtype=np.uint8).reshape(dz, dy, dx).copy()

    def reset_transfer_fn(self):
        self.lib.ref_reset_transfer_fn()

    def set_base_tf(self, base: np.ndarray):
        base = np.ascontiguousarray(base, dtype=np.float32)
        self.lib.ref_set_base_transfer_fn(base.ctypes.data_as(C.POINTER(C.c_float)))

    def base_tf(self) -> np.ndarray:
        t = np.zeros((128, 4), dtype=np.float32)
        self.lib.ref_get_base_transfer_fn(t.ctypes.data_as(C.POINTER(C.c_float)))
        return t

    def tf(self) -> np.ndarray:
        t = np.zeros((128, 4), dtype=np.float32)
        self.lib.ref_get_transfer_fn(t.ctypes.data_as(C.POINTER(C.c_float)))
        return t

    def esl_bits(self) -> np.ndarray:
        b = np.zeros(1024, dtype=np.uint32)
        self.lib.ref_get_esl_volume(b.ctypes.data_as(C.POINTER(C.c_uint32)))
        return b

    def camera(self, w, h, perspective, angles, distance=2.0):
        """benchmark_config_loop's sequence (VolR.cpp:232-247) for one pose."""
        self.lib.ref_set_viewport(w, h)
        self.lib.ref_set_perspective(1 if perspective else 0)
        self.lib.ref_set_camera_position(angles[0], angles[1], angles[2], distance)

    def raycaster(self) -> Raycaster:
        rc = Raycaster()
        self.lib.ref_get_raycaster(C.byref(rc))
        return rc

    def scene(self) -> Scene:
        """Snapshot of the reference's current state as a self-contained Scene (arrays copied)."""
        return Scene(self.volume(), self.tf(), self.esl_bits(), self.raycaster())

    def render(self, rc: Raycaster | None = None) -> np.ndarray:
        rc = rc if rc is not None else self.raycaster()
        out = np.zeros((rc.view.dims[1], rc.view.dims[0], 4), dtype=np.uint8)
        if self.lib.ref_render(C.byref(rc), _u8p(out)) != 0:
            raise RuntimeError("reference render_volume returned 1")
        return out

    def render_rows(self, rc: Raycaster, out: np.ndarray, row0: int, row1: int) -> None:
        if self.lib.ref_render_rows(C.byref(rc), _u8p(out), row0, row1) != 0:
            raise RuntimeError("ref_render_rows returned 1")


class RefGpu:
    """_ref/libvolr_ref_gpu.so - the reference's own CUDA renderers 1-4 recompiled unmodified for sm_100a (renderer 4
    through the texture-object shim of ref_gpu4_texobj.cu). Needs a GPU. The reference keeps its device state in statics
    that its destructors free without clearing, so ONE instance lives per process (RefGpu.get) and scenes / windows are
    changed through the reference's own set_* calls."""
    _instance = None

    @classmethod
    def available(cls) -> bool:
        return os.path.exists(REF_GPU_SO)

    @classmethod
    def get(cls, scene: Scene) -> "RefGpu":
        if cls._instance is None:
            cls._instance = cls(scene)
        else:
            cls._instance.set_scene(scene)
        return cls._instance

    def __init__(self, scene: Scene):
        self.lib = C.CDLL(REF_GPU_SO)
        L = self.lib
        L.refgpu_create.restype = C.c_void_p
        L.refgpu_create.argtypes = [C.POINTER(Raycaster)]
        L.refgpu_run.argtypes = [C.c_void_p, C.c_int, C.POINTER(Raycaster), C.c_int, C.POINTER(C.c_uint8), C.POINTER(C.c_double)]
        L.refgpu_set_scene.argtypes = [C.c_void_p, C.POINTER(Raycaster), C.c_int]
        L.refgpu_set_window.argtypes = [C.c_void_p, C.POINTER(View)]
        assert L.refgpu_sizeof_raycaster() == C.sizeof(Raycaster)
        with _quiet_stdout():
            self.h = L.refgpu_create(C.byref(scene.rc))
        if not self.h:
            raise RuntimeError("refgpu_create failed")
        self.dims = (scene.width, scene.height)
        self._scene = scene   # keeps the host arrays the POD points at alive

    def set_scene(self, scene: Scene, volume_changed: bool = True):
        with _quiet_stdout():
            if (scene.width, scene.height) != self.dims:
                if self.lib.refgpu_set_window(self.h, C.byref(scene.rc.view)) != 0:
                    raise RuntimeError("refgpu_set_window failed")
                self.dims = (scene.width, scene.height)
            if self.lib.refgpu_set_scene(self.h, C.byref(scene.rc), 1 if volume_changed else 0) != 0:
                raise RuntimeError("refgpu_set_scene failed")
        self._scene = scene

    def render(self, which: int, rcs) -> tuple[np.ndarray, float]:
        """Frames rcs (one Raycaster or a list; all with this instance's window) through renderer `which` (1..4);
        returns the LAST frame [H, W, 4] and the wall-clock ms of the loop."""
        rcs = rcs if isinstance(rcs, (list, tuple)) else [rcs]
        arr = (Raycaster * len(rcs))(*rcs)
        out = np.zeros((self.dims[1], self.dims[0], 4), dtype=np.uint8)
        ms = C.c_double()
        with _quiet_stdout():
            rc = self.lib.refgpu_run(self.h, which, arr, len(rcs), _u8p(out), C.byref(ms))
        if rc != 0:
            raise RuntimeError("refgpu_run(%d) returned %d" % (which, rc))
        return out, float(ms.value)


class _quiet_stdout:
    """The reference's Logger prints to stdout; keep it out of JSON lines and test output."""

    def __enter__(self):
        import sys
        sys.stdout.flush()
        self.saved, self.null = os.dup(1), os.open(os.devnull, os.O_WRONLY)
        os.dup2(self.null, 1)

    def __exit__(self, *exc):
        os.dup2(self.saved, 1)
        os.close(self.null)
        os.close(self.saved)


def ref_available() -> bool:
    return os.path.exists(REF_SO) or os.path.isdir("/root/reference/VolumeRendering")


def tf_semi(base_default: np.ndarray, n: int) -> np.ndarray:
    """SURVEY 8(d) `semi`: default colours, alpha == 1/N everywhere (no early termination, no shading)."""
    t = base_default.copy()
    t[:, 3] = np.float32(1.0) / np.float32(n)
    return t


def tf_opaque(base_default: np.ndarray) -> np.ndarray:
    """SURVEY 8(d) `opaque`: alpha = 1 for index >= 64 else 0 (first hit terminates; heavy ESL)."""
    t = base_default.copy()
    t[:, 3] = 0.0
    t[64:, 3] = 1.0
    return t
