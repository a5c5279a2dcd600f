"""Store layouts of the one-frame-over-N-ranks path (BASELINE config 4: README scene at 8K,
interleaved strips, every rank's kernel storing into rank 0's framebuffer over NVLink).

Run under torchrun on one node, one rank per GPU:
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
      --master-port 29511 tools/c4_exchange_sweep.py [--steps 40]

For every combination of patch shape (8x4 / 32x1 = RT_FLAG_WIDE_STORES) and 8-bit pixel size
(RGBA8 / RGB8 = RT_FLAG_RGB8) it times the frame like bench.py does (L2 flush, ranks started
together by a barrier, CUDA events on the launch stream, mean over steps, max over ranks) and
checks rank 0's frame against the first layout's.  GPU box only.
"""
import argparse
import ctypes as C
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from python_ray_tracer_b200 import scenes  # noqa: E402
from python_ray_tracer_b200.infrastructure.cuda import abi, affinity, sharding  # noqa: E402
from python_ray_tracer_b200.infrastructure.flatten import DefaultMaterial, flatten_scene  # noqa: E402


def main() -> None:
    parser = argparse.ArgumentParser()
    parser.add_argument("--steps", type=int, default=40)
    parser.add_argument("--warmup", type=int, default=4)
    parser.add_argument("--width", type=int, default=7680)
    parser.add_argument("--height", type=int, default=4320)
    parser.add_argument("--as-world", type=int, default=0,
                        help="cut the frame for this many ranks although fewer run (rank r still renders shard r): the "
                             "per-rank load of a larger job without its traffic into rank 0")
    parser.add_argument("--all-layouts", action="store_true", help="every combination, including 8x4 patches with 3-byte pixels (byte stores: slow over NVLink)")
    args = parser.parse_args()
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        affinity.bind_to_device(local)
        dist.init_process_group("nccl", device_id=dev)
    lib = abi.load()
    stream = torch.cuda.current_stream(local)
    w, h = args.width, args.height
    flat = flatten_scene(scenes.readme_scene(w, h), DefaultMaterial(**scenes.README_GAINS))
    cam = abi.make_camera(flat.eye, w, h)
    sph, lig = torch.from_numpy(flat.spheres).to(dev), torch.from_numpy(flat.point_lights).to(dev)
    cut = args.as_world or world
    shard = sharding.StripShard(h, cut, rank)
    flush = torch.empty(512 * 1024 * 1024, dtype=torch.uint8, device=dev)
    reference = None
    layouts = [(0, 4, 0), (1, 3, 0), (1, 3, 1)]
    if args.all_layouts:
        layouts = [(0, 4, 0), (0, 3, 0), (1, 4, 0), (1, 3, 0), (1, 4, 1), (1, 3, 1)]
    for wide, channels, even in layouts:
        flags = (abi.RT_FLAG_IMAGE_ORDER | (abi.RT_FLAG_WIDE_STORES if wide else 0) | (abi.RT_FLAG_RGB8 if channels == 3 else 0)
                 | (abi.RT_FLAG_EVEN_STORES if even else 0))
        env = abi.make_env(flat.eye, len(flat.point_lights), flat.dome_intensity, flat.dome_color, mode=abi.RT_MODE_FAST,
                           cutoff=1e-7, flags=flags)
        peer = sharding.PeerFrame(h, w, rank, world, local, dst=0, channels=channels)

        def launch() -> None:
            abi.check(lib.rt_render_strips_dev(local, sph.data_ptr(), flat.n_spheres, lig.data_ptr(), C.byref(env), C.byref(cam),
                                               shard.strip_rows, rank, cut, None, abi.RT_DTYPE_F32, 0, peer.ptr, None,
                                               int(stream.cuda_stream)), "rt_render_strips_dev")

        events = []
        for s in range(args.warmup + args.steps):
            flush.fill_(s & 0xFF)
            if world > 1:
                dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            launch()
            e1.record(stream)
            if s >= args.warmup:
                events.append((e0, e1))
        frame = peer.complete()
        ms = torch.tensor([float(np.mean([a.elapsed_time(b) for a, b in events]))], dtype=torch.float64, device=dev)
        per_rank = [ms.clone() for _ in range(world)]
        if world > 1:
            dist.all_gather(per_rank, ms)
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        if rank == 0:
            rgb = frame[:, :, :3].cpu().numpy()
            if reference is None:
                reference = rgb
            same = bool(np.array_equal(rgb, reference))
            gbs = (world - 1) / cut * h * w * channels / (ms.item() * 1e-3) / 1e9
            print(f"N={world} patch={'32x1' if wide else '8x4'} bytes/px={channels} order={'2 bottom : 1 top' if even else 'bottom-up'}: {ms.item():.4f} ms/frame  "
                  f"{w * h / (ms.item() * 1e-3) / 1e6:.0f} Mrays/s  into rank 0: {gbs:.0f} GB/s  same_frame={same}  "
                  f"per rank: {' '.join(f'{t.item():.3f}' for t in per_rank)}" + (f"  (frame cut for {cut} ranks)" if cut != world else ""),
                  flush=True)
        peer.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
