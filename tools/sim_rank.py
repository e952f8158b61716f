#!/usr/bin/env python
"""One GPU playing ONE rank of an N-GPU image partition (its interleaved 4-row strips of every frame), without peers: how close
does the per-rank kernel time come to (whole frame on one GPU) / N? Separates the kernel's own tail / occupancy effects from
everything NVLink- or hand-shake-related. Prints one JSON line.   python tools/sim_rank.py --world 8 [--workload c3] [--steps 20]"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--world", type=int, default=8)
    ap.add_argument("--workload", default="c3")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--sampler", default="hybridw")
    ap.add_argument("--only", default=None, help="run one configuration only: whole | strips | band-centre | band-edge (for ncu captures)")
    args = ap.parse_args()
    import torch
    import bench, vr5_loader
    vr5_loader.load()
    R = vr5_loader.renderer_module()
    wl = bench.WORKLOADS[args.workload]
    n, w, h = wl["n"], wl["w"], wl["h"]
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    gpu = R.GPURenderer5(0, sampler=args.sampler)
    gpu.set_stream(stream.cuda_stream)
    gpu.synth_volume(wl["kind"], n, wl["seed"])
    bd, mm, bs = gpu.build_minmax_device()
    base = {"semi": lambda: R.tf_semi(n), "opaque": R.tf_opaque, "default": R.default_base_tf}[wl["tf"]]()
    tf = R.premultiply_tf(base)
    scene = R.Scene(None, tf, R.build_esl_bits(tf, mm), bd, bs, R.default_ray_step(n, n, n))
    gpu.set_transfer_fn(scene.tf, scene.esl_bits)
    gpu.set_window_buffer(w, h)
    frames = [scene.params(R.camera(w, h, 0, bench.orbit_angles(i, args.steps)[1], 2.0)) for i in range(args.steps)]
    bufs = [torch.empty((h, w, 4), dtype=torch.uint8, device=dev) for _ in range(min(8, args.steps))]
    ptrs = [bufs[i % len(bufs)].data_ptr() for i in range(args.steps)]

    def timed(fn):
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream); fn(); e1.record(stream); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / args.steps

    out = {"workload": wl["name"], "world": args.world, "steps": args.steps}
    strips = (h + 3) // 4
    per = (strips + args.world - 1) // args.world
    configs = [(0, 1), (0, args.world), (args.world // 2, args.world), ("band-centre", args.world), ("band-edge", args.world)]
    if args.only:
        configs = [c for c in configs if {"whole": c == (0, 1), "strips": c == (0, args.world), "band-centre": c[0] == "band-centre", "band-edge": c[0] == "band-edge"}[args.only]]
    for rank, world in configs:
        if rank == "band-centre":
            gpu.set_partition_band(per * (args.world // 2), per)
        elif rank == "band-edge":
            gpu.set_partition_band(0, per)
        else:
            gpu.set_partition(rank, world, False)
        samples = sum(gpu.count(p)["samples"] for p in frames) / len(frames)
        seq = timed(lambda: gpu.render_sequence(frames, ptrs))
        single = float(gpu.render_timed(frames, ptrs[0]).mean())
        gpu.set_overlap(4)
        def lanes():
            for i, p in enumerate(frames):
                gpu.render_device(p, ptrs[i], blocking=False)
            gpu.join()
        lan = timed(lanes)
        gpu.set_overlap(1)
        out["rank%s_of_%d" % (rank, world)] = {"sequence_ms_per_frame": seq, "single_frame_launch_ms": single, "four_lanes_ms_per_frame": lan,
                                               "samples_per_frame": samples, "gsamples_sequence": samples / (seq * 1e-3) / 1e9}
    if "rank0_of_1" not in out:
        print(json.dumps(out)); gpu.close(); return
    full = out["rank0_of_1"]["sequence_ms_per_frame"]
    for k, v in out.items():
        if isinstance(v, dict) and k != "rank0_of_1":
            v["ideal_ms"] = full / args.world
            v["efficiency_sequence"] = full / args.world / v["sequence_ms_per_frame"]
    print(json.dumps(out))
    gpu.close()


if __name__ == "__main__":
    main()
