"""Host frame ring (include/vr5.h "host frame ring", csrc/vr5_hostring.cu): every rank copies its own strips into a ring of
frames in HOST memory shared by all ranks; the hand-shake is host-side counters, no device waits for another one.

test_hostring_single_process drives two ranks of ONE process on ONE device (legal here: no kernel ever waits on another);
test_hostring_two_gpus_shm is the deployment shape: one process per GPU, shared memory (/dev/shm), concurrent, no host
synchronisation between the ranks inside the loop."""
import os
import sys

import numpy as np
import pytest

from common import golden_scene, product_scene

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _frame(R, addr, h, w):
    import ctypes
    return np.frombuffer((ctypes.c_uint8 * (h * w * 4)).from_address(addr), dtype=np.uint8).reshape(h, w, 4)


@pytest.mark.gpu
@pytest.mark.parametrize("bands", [False, True], ids=["strips", "bands"])
@pytest.mark.parametrize("lanes,world", [(1, 2), (2, 3)])
def test_hostring_single_process(lanes, world, bands, R, port, golden, volumes):
    case = next(c for c in golden["cases"] if c["name"] == "synthA64_default_persp")  # 317 x 203: ragged tiles, short last strip
    sc = golden_scene(port, case, volumes(case["volume"]))
    scene, params = product_scene(R, sc)
    w, h = case["width"], case["height"]
    ranks = [R.GPURenderer5(0) for _ in range(world)]
    slots = 2 * lanes
    ring = bytearray(R.hostring_bytes(w, h, slots))
    try:
        whole = R.GPURenderer5(0)
        whole.set_window_buffer(w, h)
        whole.set_scene(scene)
        frames = [scene.params(R.camera(w, h, i % 2, (15.0 * i, 31.0 * i, 0.0), 2.0)) for i in range(2 * slots + 1)]
        want = [whole.render_volume(p) for p in frames]
        whole.close()
        for k, g in enumerate(ranks):
            g.set_window_buffer(w, h)
            g.set_scene(scene)
            if bands:
                g.set_partition_bands(k, world)   # this rank's cost-balanced band of every frame, copied as one block of rows
            else:
                g.set_partition(k, world, True)
            g.set_overlap(lanes)
            g.hostring_attach(ring, slots, init=(k == 0))
        consumer = ranks[0]
        for seq, p in enumerate(frames):             # wraps the ring twice
            for g in reversed(ranks):                # delivery order does not matter
                g.hostring_render(p, seq)
            if seq >= lanes - 1:                     # the consumer runs lanes - 1 frames behind, like bench.py
                f = seq - (lanes - 1)
                got = _frame(R, consumer.hostring_wait(f), h, w)
                assert np.array_equal(got, want[f]), f
                consumer.hostring_release(f)
        for f in range(len(frames) - (lanes - 1), len(frames)):
            assert np.array_equal(_frame(R, consumer.hostring_wait(f), h, w), want[f]), f
            consumer.hostring_release(f)
        # misuse is refused, waits are bounded
        with pytest.raises(R.Vr5Error):
            consumer.hostring_release(0)                     # released already
        unpacked = R.GPURenderer5(0)
        unpacked.set_window_buffer(w, h)
        unpacked.set_scene(scene)
        unpacked.set_partition(1, world, False)
        with pytest.raises(R.Vr5Error):
            unpacked.hostring_render(frames[0], 0)           # no ring attached
        unpacked.close()
    finally:
        for g in reversed(ranks):                    # the first attachment page-locked the block: detach it last
            g.hostring_detach()
            g.close()


def test_hostring_bytes(R):
    assert R.hostring_bytes(1920, 1080, 8) == 8192 + 8 * 1920 * 1080 * 4
    assert R.hostring_bytes(16, 16, 0) == 0 and R.hostring_bytes(16, 16, 17) == 0


def _shm_worker(rank, world, port_no, ret):
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
    tf = R.premultiply_tf(R.default_base_tf())
    scene = R.Scene(None, tf, R.build_esl_bits(tf, mm), bd, bs, R.default_ray_step(n, n, n))
    gpu.set_transfer_fn(scene.tf, scene.esl_bits)
    gpu.set_window_buffer(w, h)
    frames = [scene.params(R.camera(w, h, i % 2, (20.0, 37.0 * i, 0.0), 2.0)) for i in range(13)]
    want = None
    if rank == 0:
        gpu.set_partition(0, 1, False)
        want = [gpu.render_volume(p) for p in frames]
    lanes, slots = 2, 4
    gpu.set_partition(rank, world, True)
    gpu.set_overlap(lanes)
    nbytes = R.hostring_bytes(w, h, slots)
    name = "vr5_test_ring_%d" % port_no
    if rank == 0:
        R.ring_memory_unlink(name)
        mem = R.ring_memory(name, nbytes, create=True)
        gpu.hostring_attach(mem, slots, init=True)
    dist.barrier()
    if rank != 0:
        mem = R.ring_memory(name, nbytes, create=False)
        gpu.hostring_attach(mem, slots, init=False)
    dist.barrier()
    ok = True
    for seq, p in enumerate(frames):      # no host synchronisation between ranks inside the loop: the counters do it
        gpu.hostring_render(p, seq)
        if rank == 0 and seq >= lanes - 1:
            f = seq - (lanes - 1)
            ok = ok and np.array_equal(_frame(R, gpu.hostring_wait(f), h, w), want[f])
            gpu.hostring_release(f)
    if rank == 0:
        for f in range(len(frames) - (lanes - 1), len(frames)):
            ok = ok and np.array_equal(_frame(R, gpu.hostring_wait(f), h, w), want[f])
            gpu.hostring_release(f)
    gpu.sync()
    dist.barrier()
    gpu.hostring_detach()
    gpu.close()
    mem.close()
    dist.barrier()
    if rank == 0:
        R.ring_memory_unlink(name)
    ret[rank] = bool(ok)
    dist.destroy_process_group()


@pytest.mark.gpu
def test_hostring_two_gpus_shm():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (one process per GPU)")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    procs = [ctx.Process(target=_shm_worker, args=(r, 2, 29541, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=180)
    for p in procs:
        if p.is_alive():
            p.kill()
            pytest.fail("host ring worker hung")
    assert dict(ret) == {0: True, 1: True}
