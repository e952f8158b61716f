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
