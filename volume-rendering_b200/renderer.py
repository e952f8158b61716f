"""Python mirror of the reference's Renderer interface over the vr5 C ABI, plus the scene-builder wrappers.

``GPURenderer5`` keeps the reference's method names and return conventions (Renderer.h:13-28; 0 = ok, 1 = failure
for set_volume / render_volume) so the parity tests read like calls on a VolR renderer. Everything here is a thin
ctypes call into libvr5.so; no arithmetic of the hot path lives in Python.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import (Counters, F3, Params, SAMPLERS, View, Vr5Error, _f32p, _u32p, _u8p, lib)


class Scene:
    """Host-side bundle of what RaycasterBase::raycaster holds: volume, pre-multiplied TF, ESL bits, march params."""

    def __init__(self, volume, tf_premult, esl_bits, esl_block_dims, esl_block_size, ray_step, ray_threshold=0.95,
                 esl=True, light_kd=0.6):
        self.volume = volume  # numpy [dz, dy, dx] uint8, or None when the volume lives on the device only
        self.tf = np.ascontiguousarray(tf_premult, dtype=np.float32).reshape(128, 4)
        self.esl_bits = np.ascontiguousarray(esl_bits, dtype=np.uint32).reshape(1024)
        self.esl_block_dims = int(esl_block_dims)
        self.esl_block_size = tuple(float(v) for v in esl_block_size)
        self.ray_step = float(ray_step)
        self.ray_threshold = float(ray_threshold)
        self.esl = bool(esl)
        self.light_kd = float(light_kd)

    def params(self, view: View) -> Params:
        p = Params()
        p.view = View.from_buffer_copy(bytes(view))
        p.ray_step, p.ray_threshold = self.ray_step, self.ray_threshold
        p.esl, p.esl_block_dims = (1 if self.esl else 0), self.esl_block_dims
        p.esl_block_size = F3(*self.esl_block_size)
        p.light_kd = self.light_kd
        p.min_bound = F3(-1.0, -1.0, -1.0)
        return p


# ---- scene builders (vr5_scene.cpp; RaycasterBase / ViewBase equivalents) ------------------------------------------
def default_base_tf() -> np.ndarray:
    t = np.zeros((128, 4), dtype=np.float32)
    lib().vr5_scene_default_base_tf(_f32p(t))
    return t


def premultiply_tf(base) -> np.ndarray:
    base = np.ascontiguousarray(base, dtype=np.float32)
    out = np.zeros((128, 4), dtype=np.float32)
    lib().vr5_scene_premultiply_tf(_f32p(base), _f32p(out))
    return out


def build_minmax(volume: np.ndarray):
    dz, dy, dx = volume.shape
    bd, bs = C.c_uint16(), F3()
    mm = np.zeros((32768, 2), dtype=np.uint8)
    lib().vr5_scene_build_minmax(_u8p(volume), dx, dy, dz, C.byref(bd), _u8p(mm), C.byref(bs))
    return int(bd.value), mm, (bs.x, bs.y, bs.z)


def build_esl_bits(tf_premult, minmax) -> np.ndarray:
    bits = np.zeros(1024, dtype=np.uint32)
    tf_premult = np.ascontiguousarray(tf_premult, dtype=np.float32)
    lib().vr5_scene_build_esl_bits(_f32p(tf_premult), _u8p(minmax), _u32p(bits))
    return bits


def default_ray_step(dx, dy, dz) -> float:
    return float(lib().vr5_scene_default_ray_step(dx, dy, dz))


def camera(width, height, perspective, angles, distance=2.0) -> View:
    v = View()
    lib().vr5_scene_camera(width, height, 1 if perspective else 0, angles[0], angles[1], angles[2], distance, C.byref(v))
    return v


def tf_semi(n: int) -> np.ndarray:
    """SURVEY 8(d) `semi`: default colours, alpha == 1/N everywhere (no early termination, no shading)."""
    t = default_base_tf()
    t[:, 3] = np.float32(1.0) / np.float32(n)
    return t


def tf_opaque() -> np.ndarray:
    """SURVEY 8(d) `opaque`: alpha 1 for TF index >= 64, else 0."""
    t = default_base_tf()
    t[:, 3] = 0.0
    t[64:, 3] = 1.0
    return t


class GPURenderer5:
    """``class GPURenderer5 : public Renderer`` as seen from Python (same calls as GPURenderer1.cu:17-112)."""

    def __init__(self, device: int = -1, sampler: str | int | None = None):
        self._L = lib()
        h = C.c_void_p()
        if self._L.vr5_create(device, C.byref(h)) != 0:
            raise Vr5Error(self._L.vr5_last_error(None).decode())
        self._h = h
        self._keep = []
        if sampler is not None:
            self.set_sampler(sampler)

    def close(self):
        if getattr(self, "_h", None):
            self._L.vr5_destroy(self._h)
            self._h = None

    __del__ = close

    def _check(self, rc, soft=False):
        if rc != 0 and not soft:
            raise Vr5Error(self.last_error())
        return rc

    def last_error(self) -> str:
        return self._L.vr5_last_error(self._h).decode()

    def get_name(self) -> str:
        return self._L.vr5_name().decode()

    # -- Renderer interface -----------------------------------------------------------------------------------------
    def set_window_buffer(self, width: int, height: int):
        self._check(self._L.vr5_set_window(self._h, width, height))
        self.width, self.height = width, height

    def set_transfer_fn(self, tf_premult, esl_bits):
        tf = np.ascontiguousarray(tf_premult, dtype=np.float32)
        bits = np.ascontiguousarray(esl_bits, dtype=np.uint32)
        self._check(self._L.vr5_set_transfer_fn(self._h, _f32p(tf), _u32p(bits)))

    def set_volume(self, volume) -> int:
        """0 ok / 1 failure like the reference (GPURenderer1.cu:86-98); ``None`` = no data."""
        if volume is None:
            return self._L.vr5_set_volume(self._h, None, 0, 0, 0, 0)
        volume = np.ascontiguousarray(volume, dtype=np.uint8)
        dz, dy, dx = volume.shape
        return self._L.vr5_set_volume(self._h, _u8p(volume), volume.size, dx, dy, dz)

    def set_volume_device(self, dev_ptr: int, nbytes: int, dx: int, dy: int, dz: int):
        """Install a linear uint8 grid that already lives in device memory (device generator, NCCL broadcast)."""
        self._check(self._L.vr5_set_volume_device(self._h, dev_ptr, nbytes, dx, dy, dz))

    def render_volume(self, params: Params, out: np.ndarray | None = None) -> np.ndarray:
        """Blocking render into a host array [H, W, 4] uint8 (host-pointer path of render_volume)."""
        w, h = params.view.dims[0], params.view.dims[1]
        if out is None:
            out = np.empty((h, w, 4), dtype=np.uint8)
        self._check(self._L.vr5_render(self._h, C.byref(params), out.ctypes.data, 0))
        return out

    def render_raycaster(self, rc_bytes_ptr, out: np.ndarray) -> int:
        return self._L.vr5_render_raycaster(self._h, rc_bytes_ptr, out.ctypes.data, 0)

    # -- extras -----------------------------------------------------------------------------------------------------
    def set_scene(self, scene: Scene):
        if scene.volume is not None:
            self._check(self.set_volume(scene.volume))
        self.set_transfer_fn(scene.tf, scene.esl_bits)

    def set_sampler(self, sampler):
        s = SAMPLERS[sampler] if isinstance(sampler, str) else int(sampler)
        self._check(self._L.vr5_set_sampler(self._h, s))

    def last_variant(self) -> int:
        """VR5_SAMPLER_* id of the kernel variant the last launch used (see vr5_last_variant)."""
        return self._L.vr5_last_variant(self._h)

    def last_zero_skip(self) -> int:
        """1 when the last nearest-neighbour launch skipped the compositing of samples in the TF's all-zero prefix."""
        return self._L.vr5_last_zero_skip(self._h)

    def leap_stats(self):
        """(leaps taken, samples left out) by the zero-skip kernels since the last call (vr5_leap_stats); resets the counters."""
        out = (C.c_uint64 * 2)()
        self._check(self._L.vr5_leap_stats(self._h, out))
        return int(out[0]), int(out[1])

    def synth_volume(self, kind: int, n: int, seed: int):
        self._check(self._L.vr5_synth_volume(self._h, kind, n, seed))

    def get_volume(self, shape) -> np.ndarray:
        out = np.empty(shape, dtype=np.uint8)
        self._check(self._L.vr5_get_volume(self._h, _u8p(out), out.size))
        return out

    def count(self, params: Params) -> dict:
        c = Counters()
        self._check(self._L.vr5_count(self._h, C.byref(params), C.byref(c)))
        return c.as_dict()

    def render_device(self, params: Params, dev_ptr: int, blocking=True):
        fn = self._L.vr5_render if blocking else self._L.vr5_render_async
        self._check(fn(self._h, C.byref(params), dev_ptr, 1))

    def render_to_pinned(self, params: Params, host_ptr: int):
        self._check(self._L.vr5_render(self._h, C.byref(params), host_ptr, 0))

    def render_timed(self, frames, dev_ptr: int, l2_flush=False) -> np.ndarray:
        arr = (Params * len(frames))(*frames)
        ms = np.zeros(len(frames), dtype=np.float32)
        self._check(self._L.vr5_render_timed(self._h, arr, len(frames), dev_ptr, 1 if l2_flush else 0, _f32p(ms)))
        return ms

    def render_sequence(self, frames, dev_ptrs):
        """All `frames` with ONE persistent launch into the device buffers `dev_ptrs` (vr5_render_sequence); asynchronous."""
        arr = (Params * len(frames))(*frames)
        outs = (C.c_void_p * len(frames))(*dev_ptrs)
        self._check(self._L.vr5_render_sequence(self._h, arr, len(frames), outs))

    def set_option(self, name: str, value) -> None:
        """Experiment knob by the name of its environment variable (vr5_set_option)."""
        self._check(self._L.vr5_set_option(self._h, name.encode(), str(value).encode()))

    def sync(self):
        self._check(self._L.vr5_sync(self._h))

    def set_stream(self, stream_ptr):
        self._check(self._L.vr5_set_stream(self._h, stream_ptr))

    def set_overlap(self, lanes: int):
        """Frames round-robin over `lanes` internal streams (vr5_set_overlap)."""
        self._check(self._L.vr5_set_overlap(self._h, lanes))

    def join(self):
        self._check(self._L.vr5_join(self._h))

    def sync_lane(self, lane: int):
        self._check(self._L.vr5_sync_lane(self._h, lane))

    def render_async_host(self, params: Params, host_ptr: int):
        """Non-blocking frame into a pinned host buffer (kernel + D2H on the frame's lane)."""
        self._check(self._L.vr5_render_async(self._h, C.byref(params), host_ptr, 0))

    def set_partition(self, rank, world, packed):
        self._check(self._L.vr5_set_partition(self._h, rank, world, 1 if packed else 0))

    def set_partition_band(self, first_strip: int, n_strips: int):
        """Contiguous band of 4-row strips instead of interleaved ones (vr5_set_partition_band)."""
        self._check(self._L.vr5_set_partition_band(self._h, first_strip, n_strips))

    def set_partition_bands(self, rank: int, world: int):
        """Cost-balanced contiguous bands, re-balanced per frame (vr5_set_partition_bands)."""
        self._check(self._L.vr5_set_partition_bands(self._h, rank, world))

    def unpack_strips(self, packed_ptr, rows_padded, world, frame_ptr):
        self._check(self._L.vr5_unpack_strips(self._h, packed_ptr, rows_padded, world, frame_ptr))

    # -- fused render + gather over peer memory (include/vr5.h) -----------------------------------------------------
    def peer_create(self, slots: int) -> bytes:
        """Owner: allocate the frame ring, return its 64-byte IPC handle."""
        self._check(self._L.vr5_peer_create(self._h, slots))
        buf = C.create_string_buffer(64)
        self._check(self._L.vr5_peer_export(self._h, buf))
        return buf.raw

    def peer_open(self, handle: bytes, slots: int):
        buf = C.create_string_buffer(handle, 64)
        self._check(self._L.vr5_peer_open(self._h, buf, slots))

    def peer_open_local(self, owner: "GPURenderer5"):
        self._check(self._L.vr5_peer_open_local(self._h, owner._h))

    def peer_render(self, params: Params, seq: int):
        self._check(self._L.vr5_peer_render(self._h, C.byref(params), seq))

    def peer_wait(self, seq: int) -> int:
        p = C.c_void_p()
        self._check(self._L.vr5_peer_wait(self._h, seq, C.byref(p)))
        return p.value

    def peer_read(self, seq: int, host_ptr: int):
        self._check(self._L.vr5_peer_read(self._h, seq, host_ptr))

    def peer_sync_frame(self, seq: int):
        self._check(self._L.vr5_peer_sync_frame(self._h, seq))

    def peer_release(self, seq: int):
        self._check(self._L.vr5_peer_release(self._h, seq))

    def peer_render_sequence(self, frames, seq0: int):
        """This rank's strips of frames seq0 .. seq0 + len(frames) - 1 into the ring with ONE launch; asynchronous."""
        arr = (Params * len(frames))(*frames)
        self._check(self._L.vr5_peer_render_sequence(self._h, arr, len(frames), seq0))

    def peer_consume_sequence(self, seq0: int, n: int, copy_dst_dev: int | None = None):
        """Owner: wait for / optionally copy out / release frames seq0 .. seq0 + n - 1 on the consumer stream."""
        self._check(self._L.vr5_peer_consume_sequence(self._h, seq0, n, copy_dst_dev))

    def peer_status(self):
        self._check(self._L.vr5_peer_status(self._h))

    def peer_close(self):
        self._L.vr5_peer_close(self._h)

    # -- host frame ring: every rank copies its strips to shared HOST memory over its own PCIe link (include/vr5.h) --
    def hostring_attach(self, buffer, slots: int, init: bool):
        """`buffer`: a writable buffer object (multiprocessing.shared_memory .buf, numpy array, ...) of at least
        hostring_bytes(width, height, slots) bytes; it must outlive the attachment (hostring_detach / close)."""
        view = (C.c_char * len(buffer)).from_buffer(buffer)
        self._hostring_view = view   # keeps the export alive
        self._check(self._L.vr5_hostring_attach(self._h, C.addressof(view), len(buffer), slots, 1 if init else 0))

    def hostring_render(self, params: Params, seq: int):
        self._check(self._L.vr5_hostring_render(self._h, C.byref(params), seq))

    def hostring_wait(self, seq: int) -> int:
        """Blocks until every rank's strips of frame `seq` are in host memory; returns the frame's address in the ring."""
        p = C.c_void_p()
        self._check(self._L.vr5_hostring_wait(self._h, seq, C.byref(p)))
        return p.value

    def hostring_release(self, seq: int):
        self._check(self._L.vr5_hostring_release(self._h, seq))

    def hostring_detach(self):
        self._L.vr5_hostring_detach(self._h)
        self._hostring_view = None

    def device_alloc(self, nbytes) -> int:
        p = self._L.vr5_device_alloc(self._h, nbytes)
        if not p:
            raise Vr5Error(self.last_error())
        return p

    def device_free(self, p):
        self._L.vr5_device_free(self._h, p)

    def memcpy_d2h(self, host: np.ndarray, dev_ptr: int):
        self._check(self._L.vr5_memcpy_d2h(self._h, host.ctypes.data, dev_ptr, host.nbytes))

    def launch_count(self) -> int:
        return int(self._L.vr5_launch_count(self._h))

    def build_minmax_device(self):
        bd, bs = C.c_uint16(), F3()
        mm = np.zeros((32768, 2), dtype=np.uint8)
        self._check(self._L.vr5_build_minmax_device(self._h, C.byref(bd), _u8p(mm), C.byref(bs)))
        return int(bd.value), mm, (bs.x, bs.y, bs.z)

    def histogram_device(self) -> np.ndarray:
        h = np.zeros(256, dtype=np.float32)
        self._check(self._L.vr5_histogram_device(self._h, _f32p(h)))
        return h


# ---- volume files (include/vr5.h "volume files"; host side) ---------------------------------------------------------------
def balance_bands(params: Params, world: int) -> np.ndarray:
    """Band bounds in 4-row strips for one frame's parameters: band i = strips [b[i], b[i + 1]) (vr5_balance_bands)."""
    b = np.zeros(world + 1, dtype=np.uint32)
    if lib().vr5_balance_bands(C.byref(params), world, _u32p(b)) != 0:
        raise Vr5Error("vr5_balance_bands failed")
    return b


def hostring_bytes(width: int, height: int, slots: int) -> int:
    """Size of a host frame ring (header + slots frames), 0 for an invalid slot count (vr5_hostring_bytes)."""
    return int(lib().vr5_hostring_bytes(width, height, slots))


def ring_memory(name: str, nbytes: int, create: bool):
    """Host memory for a frame ring shared by the ranks of one node: a file in /dev/shm mapped read-write (POSIX shared
    memory without Python's resource tracker). The creator sizes it; remove it with ring_memory_unlink when all ranks are done."""
    import mmap
    path = os.path.join("/dev/shm", name)
    fd = os.open(path, os.O_RDWR | (os.O_CREAT | os.O_EXCL if create else 0), 0o600)
    try:
        if create:
            os.ftruncate(fd, nbytes)
        return mmap.mmap(fd, nbytes, mmap.MAP_SHARED, mmap.PROT_READ | mmap.PROT_WRITE)
    finally:
        os.close(fd)


def ring_memory_unlink(name: str):
    try:
        os.unlink(os.path.join("/dev/shm", name))
    except FileNotFoundError:
        pass


def _take(ptr, nbytes) -> np.ndarray:
    out = np.ctypeslib.as_array(ptr, shape=(nbytes,)).copy() if nbytes else np.zeros(0, dtype=np.uint8)
    lib().vr5_volume_free(C.cast(ptr, C.c_void_p))
    return out


def _io_check(rc):
    if rc != 0:
        raise Vr5Error(lib().vr5_volume_io_error().decode())


def load_pvm(path: str):
    """readPVMvolume: (voxels [D, H, W, components] uint8, (sx, sy, sz))."""
    p = C.POINTER(C.c_uint8)()
    w, h, d, nc = C.c_uint32(), C.c_uint32(), C.c_uint32(), C.c_uint32()
    scale = (C.c_float * 3)()
    _io_check(lib().vr5_load_pvm(path.encode(), C.byref(p), C.byref(w), C.byref(h), C.byref(d), C.byref(nc), scale))
    vox = _take(p, w.value * h.value * d.value * nc.value)
    return vox.reshape(d.value, h.value, w.value, nc.value), tuple(scale)


def load_volume(path: str) -> np.ndarray:
    """ModelBase::load_model for .pvm: 8-bit volume [D, H, W] (two-component data quantised, more refused)."""
    p = C.POINTER(C.c_uint8)()
    w, h, d = C.c_uint32(), C.c_uint32(), C.c_uint32()
    _io_check(lib().vr5_load_volume(path.encode(), C.byref(p), C.byref(w), C.byref(h), C.byref(d)))
    return _take(p, w.value * h.value * d.value).reshape(d.value, h.value, w.value)


def load_raw(path: str, w: int, h: int, d: int, components: int = 1) -> np.ndarray:
    p = C.POINTER(C.c_uint8)()
    _io_check(lib().vr5_load_raw(path.encode(), w, h, d, components, C.byref(p)))
    return _take(p, w * h * d * components).reshape(d, h, w, components)


def quantize16(data16_be: np.ndarray, w: int, h: int, d: int, linear: bool = False) -> np.ndarray:
    src = np.ascontiguousarray(data16_be, dtype=np.uint8)
    assert src.size == 2 * w * h * d
    p = C.POINTER(C.c_uint8)()
    _io_check(lib().vr5_quantize16(_u8p(src), w, h, d, 1 if linear else 0, C.byref(p)))
    return _take(p, w * h * d).reshape(d, h, w)


def dds_decode(stream: bytes) -> bytes:
    src = np.frombuffer(stream, dtype=np.uint8)
    p, n = C.POINTER(C.c_uint8)(), C.c_uint64()
    _io_check(lib().vr5_dds_decode(_u8p(src), src.size, C.byref(p), C.byref(n)))
    return _take(p, n.value).tobytes()


def dds_encode(data: bytes, skip: int = 1, strip: int = 1, v3e: bool = False) -> bytes:
    src = np.frombuffer(data, dtype=np.uint8)
    p, n = C.POINTER(C.c_uint8)(), C.c_uint64()
    _io_check(lib().vr5_dds_encode(_u8p(src), src.size, skip, strip, 1 if v3e else 0, C.byref(p), C.byref(n)))
    return _take(p, n.value).tobytes()
