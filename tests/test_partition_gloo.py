"""CPU test of the multi-GPU host logic with torch.distributed (gloo, world_size 2 and 3): strip ownership, packed
layout, gather to rank 0 and reassembly. The pixel source here is the CPU oracle (tests may use it); on the GPU box the
same plumbing carries the kernel's strips over NCCL (bench.py, tests/test_gpu_parity.py::test_partition_pack_unpack)."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port_no, w, h, ret):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port_no))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle
    import vr5_loader
    vr5_loader.load()
    import importlib
    part = importlib.import_module("volume_rendering_b200.partition")
    R = vr5_loader.renderer_module()
    port = oracle.Port()
    vol = port.synth_b(48, 2)
    view = oracle.View.from_buffer_copy(bytes(R.camera(w, h, 1, (20, 130, 0), 2.0)))
    sc = port.make_scene(vol, port.default_base_tf(), view)
    # this rank casts only the rows it owns (stand-in for cast_kernel with vr5_set_partition(rank, world, packed=1))
    mine = part.owned_rows(h, rank, world)
    full = np.zeros((h, w, 4), dtype=np.uint8)
    for r0 in sorted(set((mine // 4) * 4)):
        strip, _ = port.render(sc, rows=(int(r0), min(h, int(r0) + 4)), threads=1)
        full[r0:r0 + 4] = strip[r0:r0 + 4]
    padded = part.rows_padded(h, world)
    assert len(mine) == part.strip_rows(h, rank, world)
    packed = torch.from_numpy(part.pack(full, rank, world, padded))
    gathered = [torch.empty_like(packed) for _ in range(world)] if rank == 0 else None
    dist.gather(packed, gathered, dst=0)
    if rank == 0:
        frame = part.unpack(torch.stack(gathered).numpy(), h)
        want, _ = port.render(sc, threads=2)
        ret["equal"] = bool(np.array_equal(frame, want))
        ret["nonzero"] = int((want.reshape(-1, 4).max(1) > 0).sum())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,w,h", [(2, 96, 70), (3, 64, 33)])
def test_strips_gather_to_rank0(world, w, h):
    mgr = mp.Manager()
    ret = mgr.dict()
    port_no = 29500 + (os.getpid() % 2000) + world
    mp.spawn(_worker, args=(world, port_no, w, h, ret), nprocs=world, join=True)
    assert ret["equal"] and ret["nonzero"] > 100


def test_strip_rows_cover_the_frame(R):
    import importlib
    part = importlib.import_module("volume_rendering_b200.partition")
    for h in (1, 3, 4, 5, 33, 70, 1080, 2160):
        for world in (1, 2, 3, 4, 8):
            rows = [part.owned_rows(h, r, world) for r in range(world)]
            assert sorted(np.concatenate(rows).tolist()) == list(range(h))
            assert [len(x) for x in rows] == [part.strip_rows(h, r, world) for r in range(world)]


def _band_worker(rank, world, port_no, w, h, ret):
    """Band partition over gloo: every rank cuts the bands from the frame's parameters ON ITS OWN (no exchange), casts its band
    with the CPU oracle, and the bands reduced onto rank 0 (element-wise max over zero-initialised frames = disjoint union) give the
    whole frame - i.e. all ranks computed the same, gap-free, non-overlapping bounds."""
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port_no))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle
    import vr5_loader
    vr5_loader.load()
    R = vr5_loader.renderer_module()
    port = oracle.Port()
    vol = port.synth_a(40, 1)
    cam = R.camera(w, h, 1, (20, 130, 0), 2.0)
    sc = port.make_scene(vol, oracle.tf_semi(port.default_base_tf(), 40), oracle.View.from_buffer_copy(bytes(cam)))
    tf = R.premultiply_tf(R.tf_semi(40))
    params = R.Scene(None, tf, np.zeros(1024, np.uint32), sc.rc.esl_block_dims, sc.rc.esl_block_size.tup(), sc.rc.ray_step).params(cam)
    b = R.balance_bands(params, world).astype(int)
    bounds = [torch.zeros(world + 1, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(bounds, torch.from_numpy(b.astype(np.int64)))
    same = all(torch.equal(x, bounds[0]) for x in bounds)
    r0, r1 = int(b[rank]) * 4, min(h, int(b[rank + 1]) * 4)
    full = np.zeros((h, w, 4), dtype=np.uint8)
    if r1 > r0:
        part, _ = port.render(sc, rows=(r0, r1), threads=1)
        full[r0:r1] = part[r0:r1]
    t = torch.from_numpy(full)
    dist.reduce(t, dst=0, op=dist.ReduceOp.MAX)
    if rank == 0:
        want, _ = port.render(sc, threads=2)
        ret["equal"] = bool(np.array_equal(t.numpy(), want)) and same
        ret["bounds"] = b.tolist()
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,w,h", [(2, 96, 70), (3, 64, 57)])
def test_bands_from_every_rank_make_the_frame(world, w, h):
    mgr = mp.Manager()
    ret = mgr.dict()
    port_no = 31500 + (os.getpid() % 2000) + world
    mp.spawn(_band_worker, args=(world, port_no, w, h, ret), nprocs=world, join=True)
    assert ret["equal"], ret.get("bounds")
    assert ret["bounds"][0] == 0 and ret["bounds"][-1] == (h + 3) // 4
