"""Host logic of the multi-GPU path: strip partition, frame sharding, gather + reassembly
(world_size 2 and 3 over gloo on CPU).  The local "render" in the multi-process tests is the
CPU oracle -- the checker standing in for the kernel, never the product path."""

import os
import socket

import numpy as np
import pytest

from _cases import load_case
from python_ray_tracer_b200.infrastructure.cuda import sharding


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("height", [1, 3, 16, 17, 90, 1080, 4320])
@pytest.mark.parametrize("world", [1, 2, 3, 8])
@pytest.mark.parametrize("strip_rows", [4, 16])
def test_strips_cover_every_row_once(height, world, strip_rows):
    assert sharding.covered_exactly_once(height, world, strip_rows)
    shards = sharding.strip_shards(height, world, strip_rows)
    sizes = [s.n_rows for s in shards]
    assert sum(sizes) == height and max(sizes) - min(sizes) <= strip_rows
    order = sharding.image_order(height, world, strip_rows)
    assert sorted(order.tolist()) == sorted(set(order.tolist())) and len(order) == height


def test_strip_rows_match_the_library():
    """The host partition and rt_strip_shard_rows() (what the kernel packs) agree."""
    from python_ray_tracer_b200.infrastructure.cuda import abi

    lib = abi.load()
    for height in (1, 5, 16, 33, 1080, 4321):
        for world in (1, 2, 3, 8):
            for strip_rows in (4, 16, 64):
                for s in sharding.strip_shards(height, world, strip_rows):
                    assert lib.rt_strip_shard_rows(height, strip_rows, s.rank, world) == s.n_rows
    assert lib.rt_strip_shard_rows(0, 4, 0, 1) == -1 and lib.rt_strip_shard_rows(8, 0, 0, 1) == -1


def test_band_rows_and_frame_sharding():
    for height in (1, 7, 90, 1080):
        for world in (1, 2, 4, 8):
            bands = [sharding.band_rows(height, world, r) for r in range(world)]
            assert bands[0][0] == 0 and bands[-1][1] == height
            assert all(a[1] == b[0] for a, b in zip(bands, bands[1:]))
    owned = [list(sharding.frames_of_rank(256, 8, r)) for r in range(8)]
    assert sorted(sum(owned, [])) == list(range(256)) and all(len(o) == 32 for o in owned)
    assert list(sharding.frames_of_rank(5, 8, 6)) == []
    with pytest.raises(ValueError):
        sharding.frames_of_rank(4, 2, 2)


def _worker(rank, world, port, name, strip_rows, out):
    import torch
    import torch.distributed as dist

    from oracle import c_oracle

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        case = load_case(name)
        w, h = case.width, case.height

        def render_local(shard):
            pixels = (shard.image_rows[:, None] * w + np.arange(w)[None, :]).reshape(-1)
            d = c_oracle.ray_directions(case.flat.eye, w, h, pixels)
            o = np.repeat(np.asarray(case.flat.eye)[:, None], pixels.size, axis=1)
            rgb, _ = c_oracle.trace_rays(case.flat, o, d, max_depth=case.max_depth)
            return torch.from_numpy(np.ascontiguousarray(rgb.T.reshape(shard.n_rows, w, 3)))

        frame = sharding.render_frame_sharded(render_local, h, rank, world, strip_rows=strip_rows)
        frames = sharding.render_frames_sharded(lambda k: k * k, 7, rank, world)
        assert sorted(frames) == list(range(rank, 7, world)) and all(v == k * k for k, v in frames.items())
        if rank == 0:
            out.put(frame.numpy())
        else:
            assert frame is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,strip_rows", [(2, 16), (3, 4)])
def test_gather_reassembles_the_frame_gloo(world, strip_rows):
    """world_size > 1 on CPU: every rank renders its strips (with the oracle), rank 0 gets
    the whole frame back in image order, equal to the reference's full-frame fixture."""
    import torch.multiprocessing as mp

    name = "readme_37x23"
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, name, strip_rows, out)) for r in range(world)]
    for p in procs:
        p.start()
    frame = out.get(timeout=180)
    for p in procs:
        p.join(timeout=180)
        assert p.exitcode == 0
    case = load_case(name)
    got = frame.reshape(-1, 3).T
    assert got.shape == case.rgb.shape
    assert np.abs(got - case.rgb).max() <= 1e-12
