"""Fused render + gather over peer memory (include/vr5.h "fused render + gather", csrc/vr5_peer.cu).

test_ring_protocol_single_gpu drives an owner and a writer renderer of ONE process on ONE device in an order in which no
wait kernel ever has to spin (the guide forbids kernels that wait on one another on one GPU): it checks the data path, the
slot arithmetic and the flag words. test_two_gpus_ipc is the real thing: two processes, two GPUs, CUDA IPC, concurrent."""
import os
import sys

import numpy as np
import pytest

from common import golden_scene, product_scene

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _scene(R, port, golden, volumes):
    case = next(c for c in golden["cases"] if c["name"] == "synthA64_default_persp")  # 317 x 203: ragged tiles
    sc = golden_scene(port, case, volumes(case["volume"]))
    scene, params = product_scene(R, sc)
    return case, scene, params


@pytest.mark.gpu
@pytest.mark.parametrize("lanes", [1, 2])
def test_ring_protocol_single_gpu(lanes, R, port, golden, volumes):
    case, scene, params = _scene(R, port, golden, volumes)
    w, h = case["width"], case["height"]
    owner, writer = R.GPURenderer5(0), R.GPURenderer5(0)
    try:
        for g, rank in ((owner, 0), (writer, 1)):
            g.set_window_buffer(w, h)
            g.set_scene(scene)
            g.set_partition(rank, 2, False)
            g.set_overlap(lanes)
        whole = R.GPURenderer5(0)
        whole.set_window_buffer(w, h)
        whole.set_scene(scene)
        want = whole.render_volume(params)
        whole.close()
        slots = 2
        owner.peer_create(slots)
        writer.peer_open_local(owner)
        ptrs = set()
        for seq in range(5):  # wraps the ring twice
            writer.peer_render(params, seq)   # its wait (seq >= slots) is already satisfied: the owner released seq - slots
            writer.sync()
            owner.peer_render(params, seq)
            ptr = owner.peer_wait(seq)        # satisfied: the writer arrived and was synchronised
            ptrs.add(ptr)
            got = np.zeros((h, w, 4), dtype=np.uint8)
            owner.peer_read(seq, got.ctypes.data)
            owner.peer_sync_frame(seq)
            assert np.array_equal(got, want), seq
            owner.peer_release(seq)
            owner.sync()
        assert len(ptrs) == slots
        owner.peer_status()
        writer.peer_status()
        # misuse is refused on the host, before anything can spin on the device
        with pytest.raises(R.Vr5Error):
            owner.peer_release(7)             # out of order
        with pytest.raises(R.Vr5Error):
            writer.peer_wait(5)               # only the owner waits
        with pytest.raises(R.Vr5Error):
            owner.peer_render(params, 5 + slots)  # slot still holds an unreleased frame
    finally:
        writer.peer_close()
        owner.close()
        writer.close()


def _two_gpu_worker(rank, world, port_no, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port_no), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    import vr5_loader
    vr5_loader.load()
    R = vr5_loader.renderer_module()
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n, w, h = 64, 317, 203
    gpu = R.GPURenderer5(rank)
    gpu.synth_volume(0, n, 1)
    bd, mm, bs = gpu.build_minmax_device()
    tf = R.premultiply_tf(R.default_base_tf())
    scene = R.Scene(None, tf, R.build_esl_bits(tf, mm), bd, bs, R.default_ray_step(n, n, n))
    gpu.set_transfer_fn(scene.tf, scene.esl_bits)
    gpu.set_window_buffer(w, h)
    frames = [scene.params(R.camera(w, h, i % 2, (20.0, 37.0 * i, 0.0), 2.0)) for i in range(9)]
    want = None
    if rank == 0:
        gpu.set_partition(0, 1, False)
        want = [gpu.render_volume(p) for p in frames]
    gpu.set_partition(rank, world, False)
    gpu.set_overlap(2)
    slots = 2
    box = [gpu.peer_create(slots) if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    if rank != 0:
        gpu.peer_open(box[0], slots)
    dist.barrier()
    ok = True
    for seq, p in enumerate(frames):      # no host synchronisation between ranks inside the loop: the flags do it
        gpu.peer_render(p, seq)
        if rank == 0:
            gpu.peer_wait(seq)
            got = np.zeros((h, w, 4), dtype=np.uint8)
            gpu.peer_read(seq, got.ctypes.data)
            gpu.peer_release(seq)
            gpu.peer_sync_frame(seq)
            ok = ok and np.array_equal(got, want[seq])
    gpu.peer_status()
    gpu.sync()
    dist.barrier()
    gpu.peer_close()
    gpu.close()
    ret[rank] = bool(ok)
    dist.destroy_process_group()


@pytest.mark.gpu
def test_two_gpus_ipc():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (one process per GPU)")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    procs = [ctx.Process(target=_two_gpu_worker, args=(r, 2, 29533, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=180)
    for p in procs:
        if p.is_alive():
            p.kill()
            pytest.fail("peer gather worker hung")
    assert dict(ret) == {0: True, 1: True}


@pytest.mark.gpu
@pytest.mark.parametrize("lanes", [2, 3])
def test_frame_overlap_equals_serial(lanes, R, port, golden, volumes):
    """vr5_set_overlap: frames in flight on several lanes (own queue counters, own internal framebuffers) give the same
    bytes as the serial renderer, for device and for host outputs."""
    case, scene, _ = _scene(R, port, golden, volumes)
    w, h = case["width"], case["height"]
    gpu = R.GPURenderer5(0)
    try:
        gpu.set_window_buffer(w, h)
        gpu.set_scene(scene)
        frames = [scene.params(R.camera(w, h, i % 2, (10.0 * i, 47.0 * i, 0.0), 2.0)) for i in range(7)]
        want = [gpu.render_volume(p) for p in frames]
        gpu.set_overlap(lanes)
        host = [np.zeros((h, w, 4), dtype=np.uint8) for _ in frames]
        for p, out in zip(frames, host):          # all in flight, distinct outputs
            gpu.render_async_host(p, out.ctypes.data)
        gpu.sync()
        for a, b in zip(host, want):
            assert np.array_equal(a, b)
        dev = [gpu.device_alloc(w * h * 4) for _ in frames]
        for p, d in zip(frames, dev):
            gpu.render_device(p, d, blocking=False)
        gpu.join()                                 # the renderer's stream now waits for every lane
        for d, b in zip(dev, want):
            got = np.zeros((h, w, 4), dtype=np.uint8)
            gpu.memcpy_d2h(got, d)                 # enqueued on the renderer's stream
            assert np.array_equal(got, b)
            gpu.device_free(d)
        assert gpu.count(frames[3]) == gpu.count(frames[3])
        gpu.set_overlap(1)
        assert np.array_equal(gpu.render_volume(frames[2]), want[2])
    finally:
        gpu.close()
