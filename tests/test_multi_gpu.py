"""Parity of the one-frame-over-several-GPUs path (BASELINE configs[3]) ACROSS devices.

One process per GPU (rank r on cuda:r, NCCL), as bench.py and torchrun run it.  The frame assembled
on rank 0 -- by kernels storing into its framebuffer over NVLink peer memory, in each of the three
store layouts, and by the NCCL gather of packed shards -- must be byte-identical to the frame one
GPU renders in a single launch (pixels are independent in the reference: base.py:102-119), and within
1 LSB of the reference's own pixels (tests/golden/c4_readme_8k_subset.npz).

Collected everywhere; skipped on boxes with fewer than two GPUs (`gpurun --gpus 2 -- python -m pytest
tests/test_multi_gpu.py -m gpu`).
"""

import os
import socket

import numpy as np
import pytest

from _cases import assert_png_parity, load_case

pytestmark = pytest.mark.gpu


def _device_count() -> int:
    try:
        import torch

        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:
        return 0


LAYOUTS = {
    # name: (channels, extra flags by name)
    "rgba8_8x4": (4, ()),
    "rgba8_32x1": (4, ("RT_FLAG_WIDE_STORES",)),
    "rgb8_32x1_both_ends": (3, ("RT_FLAG_WIDE_STORES", "RT_FLAG_RGB8", "RT_FLAG_EVEN_STORES")),
}


def _worker(rank, world, port, width, height, out):
    import torch
    import torch.distributed as dist

    from python_ray_tracer_b200 import Camera, scenes
    from python_ray_tracer_b200.infrastructure import CudaRenderer, NumpyVector3D
    from python_ray_tracer_b200.infrastructure.cuda import abi, sharding

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    report = {}
    try:
        scene = scenes.readme_scene(width, height)
        renderer = CudaRenderer(**scenes.README_GAINS, precision="fast", device=rank)
        flat = renderer.flatten(scene)
        cam = Camera(NumpyVector3D(*flat.eye), width, height)
        single = None
        if rank == 0:  # the frame one GPU renders in one launch
            single = renderer.render_rows(flat, cam, 0, height, want_rgba8=True, no_planes=True).rgba8.view(height, width, 4)
        shard = sharding.StripShard(height, world, rank)
        for name, (channels, flag_names) in LAYOUTS.items():
            flags = 0
            for f in flag_names:
                flags |= getattr(abi, f)
            peer = sharding.PeerFrame(height, width, rank, world, rank, dst=0, channels=channels)
            renderer.render_strips_into(flat, cam, shard, peer.ptr, flags=flags)
            frame = peer.complete()
            if rank == 0:
                report[name] = bool(torch.equal(frame[:, :, :3], single[:, :, :3]))
                if name == "rgba8_8x4":
                    report["alpha"] = bool((frame[:, :, 3] == 255).all())
            peer.close()
        # the NCCL alternative: packed shards gathered to rank 0 and permuted into image order
        gathered = renderer.render_frame_sharded(scene, rank, world)
        if rank == 0:
            report["nccl_gather"] = bool(torch.equal(gathered, single))
            report["frame"] = single[:, :, :3].cpu().numpy() if width * height <= 1 << 22 else None
            case = load_case("c4_readme_8k_subset") if (width, height) == (7680, 4320) else None
            if case is not None:  # the reference's own pixels of this frame
                report["subset8"] = single.view(-1, 4)[torch.from_numpy(case.pixels).to(single.device)][:, :3].cpu().numpy()
            out.put(report)
    finally:
        dist.destroy_process_group()


def _run(world, width, height):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    procs = [ctx.Process(target=_worker, args=(r, world, port, width, height, out)) for r in range(world)]
    for p in procs:
        p.start()
    report = out.get(timeout=600)
    for p in procs:
        p.join(timeout=600)
        assert p.exitcode == 0
    return report


@pytest.mark.parametrize("world", [2, 4, 8])
def test_frame_assembled_across_gpus_equals_the_single_gpu_frame(world, c_oracle):
    if _device_count() < world:
        pytest.skip(f"needs {world} GPUs, this box has {_device_count()}")
    # a frame whose height is not a multiple of world * strip_rows, then BASELINE configs[3] at size
    small = _run(world, 1000, 612)
    for name in list(LAYOUTS) + ["nccl_gather", "alpha"]:
        assert small[name], (world, name)
    big = _run(world, 7680, 4320)
    for name in list(LAYOUTS) + ["nccl_gather", "alpha"]:
        assert big[name], (world, name)
    case = load_case("c4_readme_8k_subset")
    assert_png_parity(big["subset8"], c_oracle.quantize(case.rgb), f"8K frame over {world} GPUs vs the reference's pixels")
