"""Frame sequences (include/vr5.h vr5_render_sequence, vr5_peer_render_sequence / vr5_peer_consume_sequence): ONE persistent
launch renders a whole camera path - the tiles of all frames come out of one queue - and, through the frame ring, the
arrive / release hand-shake runs inside the kernel. Contract: the pixels are bit-equal to one vr5_render call per frame
(which the parity tests hold to the reference), whatever the sampler, the partition or the queue-fetch mode."""
import os
import sys

import numpy as np
import pytest

from common import golden_scene, product_scene

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

ANGLES = [(20.0, 0.0, 0.0), (20.0, 72.0, 0.0), (-45.0, -45.0, 0.0), (90.0, 10.0, 5.0), (180.0, 90.0, 0.0), (33.0, 200.0, 0.0), (5.0, 300.0, 0.0)]


def _frames(R, scene, w, h, persp, n=len(ANGLES)):
    return [scene.params(R.camera(w, h, persp, a, 2.0)) for a in ANGLES[:n]]


def _sequence_host(R, gpu, frames, w, h, rows=None):
    """render_sequence into fresh device buffers, copied back: [n, rows, w, 4]"""
    rows = rows or h
    ptrs = [gpu.device_alloc(rows * w * 4) for _ in frames]
    try:
        gpu.render_sequence(frames, ptrs)
        gpu.sync()
        out = []
        for p in ptrs:
            a = np.zeros((rows, w, 4), dtype=np.uint8)
            gpu.memcpy_d2h(a, p)
            out.append(a)
        return out
    finally:
        for p in ptrs:
            gpu.device_free(p)


@pytest.mark.gpu
@pytest.mark.parametrize("sampler", ["hybridw", "tex3d", "linear", "brick", "window", "trilinear", "trilinear-fma"])
@pytest.mark.parametrize("name,persp", [("synthA64_semi_ortho", 0), ("synthB64_opaque_persp", 1), ("bucky_persp_-45_-45_0", 1)])
def test_sequence_equals_per_frame_calls(name, persp, sampler, golden, port, volumes, R):
    """semi TF = CTA-wide queue fetch, opaque TF = per-warp fetch + leading empty-space loop, default TF on Bucky = shading."""
    case = next(c for c in golden["cases"] if c["name"] == name)
    sc = golden_scene(port, case, volumes(case["volume"]))
    scene, _ = product_scene(R, sc)
    w, h = case["width"], case["height"]
    gpu = R.GPURenderer5(0, sampler=sampler)
    try:
        gpu.set_window_buffer(w, h)
        gpu.set_scene(scene)
        frames = _frames(R, scene, w, h, persp)
        want = [gpu.render_volume(p) for p in frames]
        variant = gpu.last_variant()
        got = _sequence_host(R, gpu, frames, w, h)
        assert gpu.last_variant() == variant   # the same kernel variant serves the sequence
        for i, (a, b) in enumerate(zip(got, want)):
            assert np.array_equal(a, b), (name, sampler, i)
        assert len({a.tobytes() for a in want}) == len(want)   # the frames really differ
    finally:
        gpu.close()


@pytest.mark.gpu
@pytest.mark.parametrize("packed", [False, True])
def test_sequence_partition_strips(packed, golden, port, volumes, R):
    """Rank r of 3 renders its strips of every frame of the sequence: in place (other rows untouched) or packed."""
    case = next(c for c in golden["cases"] if c["name"] == "synthA64_default_persp")  # 317 x 203: ragged tiles
    sc = golden_scene(port, case, volumes(case["volume"]))
    scene, _ = product_scene(R, sc)
    w, h = case["width"], case["height"]
    gpu = R.GPURenderer5(0)
    try:
        gpu.set_window_buffer(w, h)
        gpu.set_scene(scene)
        frames = _frames(R, scene, w, h, 1, 4)
        whole = [gpu.render_volume(p) for p in frames]
        L = R.lib()
        for rank in range(3):
            gpu.set_partition(rank, 3, packed)
            rows = L.vr5_strip_rows(h, rank, 3) if packed else h
            got = _sequence_host(R, gpu, frames, w, h, rows)
            own = np.nonzero((np.arange(h) // 4) % 3 == rank)[0]
            for a, b in zip(got, whole):
                if packed:
                    assert np.array_equal(a, b[own])
                else:
                    assert np.array_equal(a[own], b[own])
    finally:
        gpu.set_partition(0, 1, False)
        gpu.close()


@pytest.mark.gpu
def test_sequence_refuses_frames_that_differ_in_more_than_the_camera(golden, port, volumes, R):
    case = next(c for c in golden["cases"] if c["name"] == "synthA64_semi_ortho")
    sc = golden_scene(port, case, volumes(case["volume"]))
    scene, _ = product_scene(R, sc)
    w, h = case["width"], case["height"]
    gpu = R.GPURenderer5(0)
    try:
        gpu.set_window_buffer(w, h)
        gpu.set_scene(scene)
        frames = _frames(R, scene, w, h, 0, 3)
        frames[2].ray_step *= 1.5
        ptrs = [gpu.device_alloc(w * h * 4) for _ in frames]
        with pytest.raises(R.Vr5Error, match="differ"):
            gpu.render_sequence(frames, ptrs)
        frames = _frames(R, scene, w, h, 0, 2) + _frames(R, scene, w, h, 1, 1)   # mixed projections
        with pytest.raises(R.Vr5Error, match="differ"):
            gpu.render_sequence(frames, ptrs)
        for p in ptrs:
            gpu.device_free(p)
    finally:
        gpu.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["synthA64_semi_ortho", "synthB64_opaque_ortho"])
def test_ring_sequence_one_gpu_concurrent_consumer(name, golden, port, volumes, R):
    """World 1, ring of 2 slots, 7 frames in ONE launch: the kernel has to wait - inside the kernel - for the consumer to
    release slots, while the consumer (stream memory operations on its own stream) waits for the kernel's arrivals. Both
    directions of the hand-shake run concurrently on one GPU; the copied-out frames equal the per-frame calls."""
    case = next(c for c in golden["cases"] if c["name"] == name)
    sc = golden_scene(port, case, volumes(case["volume"]))
    scene, _ = product_scene(R, sc)
    w, h = case["width"], case["height"]
    gpu = R.GPURenderer5(0)
    try:
        gpu.set_window_buffer(w, h)
        gpu.set_scene(scene)
        frames = _frames(R, scene, w, h, 0)
        want = [gpu.render_volume(p) for p in frames]
        gpu.peer_create(2)
        dst = gpu.device_alloc(len(frames) * w * h * 4)
        seq = 0
        for rounds in range(2):   # the second round starts in the middle of the ring's use counts
            gpu.peer_render_sequence(frames, seq)
            gpu.peer_consume_sequence(seq, len(frames), dst)
            seq += len(frames)
            gpu.join()
            gpu.sync()
            gpu.peer_status()     # no wait timed out
            got = np.zeros((len(frames), h, w, 4), dtype=np.uint8)
            gpu.memcpy_d2h(got, dst)
            for i in range(len(frames)):
                assert np.array_equal(got[i], want[i]), (rounds, i)
        # the per-frame API continues on the same ring (same flag words, same sequence numbers)
        for p in frames[:3]:
            gpu.peer_render(p, seq)
            gpu.peer_wait(seq)
            one = np.zeros((h, w, 4), dtype=np.uint8)
            gpu.peer_read(seq, one.ctypes.data)
            gpu.peer_release(seq)
            gpu.peer_sync_frame(seq)
            assert np.array_equal(one, want[seq - 2 * len(frames)])
            seq += 1
        gpu.device_free(dst)
    finally:
        gpu.close()


@pytest.mark.gpu
def test_ring_sequence_two_ranks_one_gpu(golden, port, volumes, R):
    """Owner and writer renderers of one process on one device, world 2: sequences no longer than the ring, enqueued in an
    order in which no kernel has to wait for another kernel of the same GPU; checks slot arithmetic, arrival counting
    (both ranks arrive) and release across both ranks."""
    case = next(c for c in golden["cases"] if c["name"] == "synthA64_default_persp")
    sc = golden_scene(port, case, volumes(case["volume"]))
    scene, _ = product_scene(R, sc)
    w, h = case["width"], case["height"]
    owner, writer = R.GPURenderer5(0), R.GPURenderer5(0)
    try:
        for g, rank in ((owner, 0), (writer, 1)):
            g.set_window_buffer(w, h)
            g.set_scene(scene)
        frames = _frames(R, scene, w, h, 1, 6)
        want = [owner.render_volume(p) for p in frames]
        owner.set_partition(0, 2, False)
        writer.set_partition(1, 2, False)
        slots = 3
        owner.peer_create(slots)
        writer.peer_open_local(owner)
        dst = owner.device_alloc(slots * w * h * 4)
        for seq0 in (0, 3):
            batch = frames[seq0:seq0 + slots]
            writer.peer_render_sequence(batch, seq0)   # its slots are free: the owner released their previous frames
            writer.sync()
            owner.peer_render_sequence(batch, seq0)
            owner.peer_consume_sequence(seq0, slots, dst)
            owner.sync()
            got = np.zeros((slots, h, w, 4), dtype=np.uint8)
            owner.memcpy_d2h(got, dst)
            for i in range(slots):
                assert np.array_equal(got[i], want[seq0 + i]), (seq0, i)
        owner.peer_status()
        owner.device_free(dst)
    finally:
        writer.peer_close()
        owner.close()
        writer.close()


def _two_gpu_sequence_worker(rank, world, port_no, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port_no), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    import vr5_loader
    vr5_loader.load()
    R = vr5_loader.renderer_module()
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n, w, h = 64, 317, 203
    gpu = R.GPURenderer5(rank)
    gpu.synth_volume(0, n, 1)
    bd, mm, bs = gpu.build_minmax_device()
    tf = R.premultiply_tf(R.tf_semi(n))
    scene = R.Scene(None, tf, R.build_esl_bits(tf, mm), bd, bs, R.default_ray_step(n, n, n))
    gpu.set_transfer_fn(scene.tf, scene.esl_bits)
    gpu.set_window_buffer(w, h)
    frames = [scene.params(R.camera(w, h, 0, (20.0, 37.0 * i, 0.0), 2.0)) for i in range(11)]
    want = None
    if rank == 0:
        gpu.set_partition(0, 1, False)
        want = [gpu.render_volume(p) for p in frames]
    gpu.set_partition(rank, world, False)
    slots = 2
    box = [gpu.peer_create(slots) if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    if rank != 0:
        gpu.peer_open(box[0], slots)
    dist.barrier()
    ok = True
    dst = gpu.device_alloc(len(frames) * w * h * 4) if rank == 0 else None
    seq = 0
    for rounds in range(2):   # 11 frames through 2 slots, twice: every slot is reused many times, across the NVLink
        gpu.peer_render_sequence(frames, seq)
        if rank == 0:
            gpu.peer_consume_sequence(seq, len(frames), dst)
        seq += len(frames)
        gpu.join()
        gpu.sync()
        if rank == 0:
            got = np.zeros((len(frames), h, w, 4), dtype=np.uint8)
            gpu.memcpy_d2h(got, dst)
            ok = ok and all(np.array_equal(got[i], want[i]) for i in range(len(frames)))
        dist.barrier()
    gpu.peer_status()
    dist.barrier()
    gpu.peer_close()
    gpu.close()
    ret[rank] = bool(ok)
    dist.destroy_process_group()


@pytest.mark.gpu
def test_ring_sequence_two_gpus_ipc():
    """Two processes, two GPUs, CUDA IPC: the writer's kernel stores into the owner's ring over the peer link and both
    kernels and the consumer run concurrently; no host synchronisation inside a sequence."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    import socket

    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port_no = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    procs = [ctx.Process(target=_two_gpu_sequence_worker, args=(r, 2, port_no, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(150)
    codes = [p.exitcode for p in procs]
    for p in procs:          # a rank that died leaves its partner waiting for flags that never come: do not leave it behind
        if p.is_alive():
            p.kill()
            p.join(10)
    assert all(c == 0 for c in codes), codes
    assert ret.get(0) is True and ret.get(1) is True


@pytest.mark.gpu
def test_set_option_and_window_guard(golden, port, volumes, R):
    case = next(c for c in golden["cases"] if c["name"] == "synthA64_semi_ortho")
    sc = golden_scene(port, case, volumes(case["volume"]))
    scene, params = product_scene(R, sc)
    w, h = case["width"], case["height"]
    gpu = R.GPURenderer5(0)
    try:
        gpu.set_window_buffer(w, h)
        gpu.set_scene(scene)
        want = gpu.render_volume(params)
        for name, value in [("VR5_QUEUE_ORDER", "rings"), ("VR5_WARP_FETCH", "1"), ("VR5_QUEUE_ORDER", "cols"), ("VR5_QUEUE_ORDER", "rows")]:
            gpu.set_option(name, value)
            assert np.array_equal(gpu.render_volume(params), want), (name, value)
        with pytest.raises(R.Vr5Error):
            gpu.set_option("VR5_NO_SUCH_KNOB", "1")
        with pytest.raises(R.Vr5Error):
            gpu.set_option("VR5_QUEUE_ORDER", "diagonal")
        # a ring's slots are sized for the window it was created with
        gpu.peer_create(2)
        with pytest.raises(R.Vr5Error, match="ring"):
            gpu.set_window_buffer(w * 2, h)
        gpu.set_window_buffer(w, h)   # same size: allowed
        gpu.peer_close()
        gpu.set_window_buffer(w * 2, h)
    finally:
        gpu.close()
