"""volume-rendering_b200 - ctypes binding of libvr5.so, the C ABI of the B200-native ray caster (include/vr5.h).

The product is the CUDA library; this module only loads it and mirrors the reference's ``Renderer`` plugin
interface (reference Renderer.h:13-28: ``get_name / set_window_buffer / set_transfer_fn / set_volume /
render_volume``) so the Python tests and ``bench.py`` read like calls on a VolR renderer. There is no CPU fallback:
if ``libvr5.so`` is missing or no CUDA device exists, construction raises.

The directory name carries a hyphen (it follows the upstream repository name); import it through
``vr5_loader.load()`` at the repository root, which registers it as ``volume_rendering_b200``.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libvr5.so")
CSRC = os.path.join(HERE, "csrc")

SAMPLER_LINEAR, SAMPLER_BRICK, SAMPLER_TEX3D, SAMPLER_HYBRID = 0, 1, 2, 3
SAMPLERS = {"linear": 0, "brick": 1, "tex3d": 2, "hybrid": 3, "hybridw": 4, "window": 5, "hybridw2": 6, "trilinear": 16, "trilinear-fma": 18}


class Vr5Error(RuntimeError):
    pass


class F3(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("z", C.c_float)]


class View(C.Structure):  # include/vr5_pod.h vr5_view == reference ViewBase.h:14-21
    _fields_ = [("dims", C.c_uint16 * 2), ("origin", F3), ("direction", F3), ("right_plane", F3),
                ("up_plane", F3), ("light_pos", F3), ("perspective", C.c_uint8), ("_pad", C.c_uint8 * 3)]


class Params(C.Structure):  # include/vr5.h vr5_params
    _fields_ = [("view", View), ("ray_step", C.c_float), ("ray_threshold", C.c_float), ("esl", C.c_uint32),
                ("esl_block_dims", C.c_uint32), ("esl_block_size", F3), ("light_kd", C.c_float), ("min_bound", F3)]


class Counters(C.Structure):  # include/vr5.h vr5_counters
    _fields_ = [("rays_hit", C.c_uint64), ("rays_accum", C.c_uint64), ("esl_iters", C.c_uint64),
                ("samples", C.c_uint64), ("shaded", C.c_uint64)]

    def as_dict(self):
        return {k: int(getattr(self, k)) for k, _ in self._fields_}


assert C.sizeof(View) == 68


def build(verbose: bool = False) -> str:
    """Compile libvr5.so in-tree for sm_100a (csrc/Makefile). nvcc cross-compiles without a GPU."""
    subprocess.run(["make", "-C", CSRC], check=True, stdout=None if verbose else subprocess.DEVNULL)
    return LIB_PATH


_lib = None


def lib() -> C.CDLL:
    """Load libvr5.so (fails loudly when it has not been built: there is no other implementation)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise Vr5Error("libvr5.so is not built (run `make -C volume-rendering_b200/csrc` or __graft_entry__.build()); "
                       "there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    vp, u8p, f32p, u32p = C.c_void_p, C.POINTER(C.c_uint8), C.POINTER(C.c_float), C.POINTER(C.c_uint32)
    sig = {
        "vr5_create": (C.c_int, [C.c_int, C.POINTER(vp)]),
        "vr5_destroy": (None, [vp]),
        "vr5_name": (C.c_char_p, []),
        "vr5_last_error": (C.c_char_p, [vp]),
        "vr5_version": (C.c_int, []),
        "vr5_set_window": (C.c_int, [vp, C.c_uint16, C.c_uint16]),
        "vr5_set_transfer_fn": (C.c_int, [vp, f32p, u32p]),
        "vr5_set_volume": (C.c_int, [vp, u8p, C.c_uint64, C.c_uint16, C.c_uint16, C.c_uint16]),
        "vr5_set_volume_device": (C.c_int, [vp, vp, C.c_uint64, C.c_uint16, C.c_uint16, C.c_uint16]),
        "vr5_get_volume": (C.c_int, [vp, u8p, C.c_uint64]),
        "vr5_set_sampler": (C.c_int, [vp, C.c_int]),
        "vr5_get_sampler": (C.c_int, [vp]),
        "vr5_last_variant": (C.c_int, [vp]),
        "vr5_last_zero_skip": (C.c_int, [vp]),
        "vr5_leap_stats": (C.c_int, [vp, C.POINTER(C.c_uint64)]),
        "vr5_render": (C.c_int, [vp, C.POINTER(Params), vp, C.c_int]),
        "vr5_render_raycaster": (C.c_int, [vp, vp, vp, C.c_int]),
        "vr5_render_async": (C.c_int, [vp, C.POINTER(Params), vp, C.c_int]),
        "vr5_render_sequence": (C.c_int, [vp, C.POINTER(Params), C.c_int, C.POINTER(vp)]),
        "vr5_render_sequence_noalloc": (C.c_int, [vp, C.POINTER(Params), C.c_int, C.POINTER(vp)]),
        "vr5_set_option": (C.c_int, [vp, C.c_char_p, C.c_char_p]),
        "vr5_sync": (C.c_int, [vp]),
        "vr5_set_stream": (C.c_int, [vp, vp]),
        "vr5_set_overlap": (C.c_int, [vp, C.c_int]),
        "vr5_get_overlap": (C.c_int, [vp]),
        "vr5_join": (C.c_int, [vp]),
        "vr5_sync_lane": (C.c_int, [vp, C.c_int]),
        "vr5_render_timed": (C.c_int, [vp, C.POINTER(Params), C.c_int, vp, C.c_int, f32p]),
        "vr5_count": (C.c_int, [vp, C.POINTER(Params), C.POINTER(Counters)]),
        "vr5_launch_count": (C.c_uint64, [vp]),
        "vr5_host_alloc": (vp, [C.c_uint64]),
        "vr5_host_free": (None, [vp]),
        "vr5_device_alloc": (vp, [vp, C.c_uint64]),
        "vr5_device_free": (None, [vp, vp]),
        "vr5_memcpy_d2h": (C.c_int, [vp, vp, vp, C.c_uint64]),
        "vr5_set_partition": (C.c_int, [vp, C.c_int, C.c_int, C.c_int]),
        "vr5_set_partition_band": (C.c_int, [vp, C.c_int, C.c_int]),
        "vr5_set_partition_bands": (C.c_int, [vp, C.c_int, C.c_int]),
        "vr5_balance_bands": (C.c_int, [C.POINTER(Params), C.c_int, u32p]),
        "vr5_strip_rows": (C.c_uint32, [C.c_uint16, C.c_int, C.c_int]),
        "vr5_unpack_strips": (C.c_int, [vp, vp, C.c_uint32, C.c_int, vp]),
        "vr5_peer_create": (C.c_int, [vp, C.c_int]),
        "vr5_peer_export": (C.c_int, [vp, vp]),
        "vr5_peer_open": (C.c_int, [vp, vp, C.c_int]),
        "vr5_peer_open_local": (C.c_int, [vp, vp]),
        "vr5_peer_render": (C.c_int, [vp, C.POINTER(Params), C.c_uint64]),
        "vr5_peer_wait": (C.c_int, [vp, C.c_uint64, C.POINTER(vp)]),
        "vr5_peer_release": (C.c_int, [vp, C.c_uint64]),
        "vr5_peer_read": (C.c_int, [vp, C.c_uint64, vp]),
        "vr5_peer_sync_frame": (C.c_int, [vp, C.c_uint64]),
        "vr5_peer_render_sequence": (C.c_int, [vp, C.POINTER(Params), C.c_int, C.c_uint64]),
        "vr5_peer_consume_sequence": (C.c_int, [vp, C.c_uint64, C.c_int, vp]),
        "vr5_peer_status": (C.c_int, [vp]),
        "vr5_peer_close": (None, [vp]),
        "vr5_hostring_bytes": (C.c_uint64, [C.c_uint16, C.c_uint16, C.c_int]),
        "vr5_hostring_attach": (C.c_int, [vp, vp, C.c_uint64, C.c_int, C.c_int]),
        "vr5_hostring_render": (C.c_int, [vp, C.POINTER(Params), C.c_uint64]),
        "vr5_hostring_wait": (C.c_int, [vp, C.c_uint64, C.POINTER(vp)]),
        "vr5_hostring_release": (C.c_int, [vp, C.c_uint64]),
        "vr5_hostring_detach": (None, [vp]),
        "vr5_synth_volume": (C.c_int, [vp, C.c_int, C.c_uint16, C.c_uint32]),
        "vr5_load_pvm": (C.c_int, [C.c_char_p, C.POINTER(u8p), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), f32p]),
        "vr5_load_raw": (C.c_int, [C.c_char_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(u8p)]),
        "vr5_quantize16": (C.c_int, [u8p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.POINTER(u8p)]),
        "vr5_load_volume": (C.c_int, [C.c_char_p, C.POINTER(u8p), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
        "vr5_dds_decode": (C.c_int, [u8p, C.c_uint64, C.POINTER(u8p), C.POINTER(C.c_uint64)]),
        "vr5_dds_encode": (C.c_int, [u8p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_int, C.POINTER(u8p), C.POINTER(C.c_uint64)]),
        "vr5_volume_free": (None, [vp]),
        "vr5_volume_io_error": (C.c_char_p, []),
        "vr5_scene_default_base_tf": (None, [f32p]),
        "vr5_scene_premultiply_tf": (None, [f32p, f32p]),
        "vr5_scene_build_minmax": (None, [u8p, C.c_uint16, C.c_uint16, C.c_uint16, C.POINTER(C.c_uint16), u8p, C.POINTER(F3)]),
        "vr5_scene_build_esl_bits": (None, [f32p, u8p, u32p]),
        "vr5_scene_default_ray_step": (C.c_float, [C.c_uint16, C.c_uint16, C.c_uint16]),
        "vr5_scene_camera": (None, [C.c_uint16, C.c_uint16, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.POINTER(View)]),
        "vr5_build_minmax_device": (C.c_int, [vp, C.POINTER(C.c_uint16), u8p, C.POINTER(F3)]),
        "vr5_histogram_device": (C.c_int, [vp, f32p]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)  # AttributeError here = the library does not export what include/vr5.h declares
        fn.restype, fn.argtypes = res, args
    _lib = L
    return L


EXPORTED_SYMBOLS = None  # filled by tests from include/vr5.h


def _f32p(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _u8p(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint8))


def _u32p(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint32))
