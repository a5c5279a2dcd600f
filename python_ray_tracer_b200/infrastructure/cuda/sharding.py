"""Multi-GPU sharding of the hot path: one process per GPU, ``torch.distributed`` plumbing.

Pixels are independent in the reference (``base.py:102-119`` masks per pixel; a row band
or any pixel subset renders bit-identically), so the path
shards without a data-path collective.  Two partitions:

* **one frame over G ranks** (BASELINE config 4): the image is cut into strips of
  ``strip_rows`` rows and rank ``r`` renders strips ``r, r+G, r+2G, ...``
  (``rt_render_strips_dev``).  Every shard is a stratified sample of the frame, so the
  shards cost the same without a cost model (contiguous bands do not: the sky rows of the
  README scene are nearly free).  The one exchange step is a gather of the shards to
  rank 0 (``torch.distributed.gather``: NCCL over NVLink on GPUs, gloo in the CPU tests)
  followed by a row permutation into image order.
* **many frames over G ranks** (config 5, the turntable): frame ``k`` belongs to rank
  ``k mod G``; no exchange at all.

Nothing here computes pixels: the local render is a callable supplied by the caller
(``CudaRenderer.render_strips`` in the product; the CPU tests plug in a stand-in).
"""

from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Callable, List, Optional, Sequence

import numpy as np

DEFAULT_STRIP_ROWS = 16  # four 8x4 patches high: keeps warps of a block on adjacent rows


@dataclass(frozen=True)
class StripShard:
    """Rows of a ``height``-row frame owned by ``rank`` of ``world`` (interleaved strips)."""

    height: int
    world: int
    rank: int
    strip_rows: int = DEFAULT_STRIP_ROWS

    def __post_init__(self) -> None:
        if self.height <= 0 or self.world <= 0 or self.strip_rows <= 0:
            raise ValueError(f"bad shard {self}")
        if not 0 <= self.rank < self.world:
            raise ValueError(f"rank {self.rank} outside 0..{self.world - 1}")

    @property
    def image_rows(self) -> np.ndarray:
        """Image row of every local row, in local order (what ``rt_render_strips_dev`` packs)."""
        n_strips = -(-self.height // self.strip_rows)
        mine = np.arange(self.rank, n_strips, self.world, dtype=np.int64)
        rows = (mine[:, None] * self.strip_rows + np.arange(self.strip_rows, dtype=np.int64)[None, :]).reshape(-1)
        return rows[rows < self.height]

    @property
    def n_rows(self) -> int:
        return int(self.image_rows.size)


def strip_shards(height: int, world: int, strip_rows: int = DEFAULT_STRIP_ROWS) -> List[StripShard]:
    return [StripShard(height, world, r, strip_rows) for r in range(world)]


def band_rows(height: int, world: int, rank: int, align: int = 4) -> tuple:
    """Contiguous alternative: rows ``[row0, row1)`` of rank ``rank``, boundaries aligned to
    the patch height (for callers that need contiguous bands, e.g. tiled image writers)."""
    if not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside 0..{world - 1}")
    units = -(-height // align)
    lo = (units * rank // world) * align
    hi = (units * (rank + 1) // world) * align
    return min(lo, height), min(hi, height)


def frames_of_rank(n_frames: int, world: int, rank: int) -> range:
    """Frame-sharding of an animation batch: frame ``k`` is rendered by rank ``k mod world``."""
    if not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside 0..{world - 1}")
    return range(rank, n_frames, world)


def image_order(height: int, world: int, strip_rows: int = DEFAULT_STRIP_ROWS) -> np.ndarray:
    """Index ``idx`` such that ``gathered.reshape(world * max_rows, ...)[idx]`` is the frame in
    image row order, where ``gathered[r, :n_r]`` holds rank ``r``'s packed shard and every
    rank's block is padded to ``max_rows`` rows."""
    shards = strip_shards(height, world, strip_rows)
    max_rows = max(s.n_rows for s in shards)
    idx = np.empty(height, dtype=np.int64)
    for s in shards:
        idx[s.image_rows] = s.rank * max_rows + np.arange(s.n_rows, dtype=np.int64)
    return idx


_ORDER_CACHE: dict = {}


def gather_frame(local: Any, shard: StripShard, dst: int = 0, group: Any = None) -> Optional[Any]:
    """Gather the packed shards (``local``: tensor ``[shard.n_rows, ...]``) to rank ``dst``
    and return the frame in image row order there (``None`` on the other ranks).

    ``torch.distributed.gather`` needs equal shapes, so each rank pads its block to the
    largest shard (they differ by at most ``strip_rows`` rows).
    """
    import torch
    import torch.distributed as dist

    if local.shape[0] != shard.n_rows:
        raise ValueError(f"local block has {local.shape[0]} rows, shard owns {shard.n_rows}")
    world = shard.world
    if world == 1:
        return local
    max_rows = max(s.n_rows for s in strip_shards(shard.height, world, shard.strip_rows))
    block = local
    if local.shape[0] != max_rows:
        block = local.new_zeros((max_rows,) + tuple(local.shape[1:]))
        block[: shard.n_rows] = local
    block = block.contiguous()
    if shard.rank != dst:
        dist.gather(block, None, dst=dst, group=group)
        return None
    gathered = local.new_empty((world, max_rows) + tuple(local.shape[1:]))
    dist.gather(block, list(gathered.unbind(0)), dst=dst, group=group)
    key = (shard.height, world, shard.strip_rows, str(local.device))
    order = _ORDER_CACHE.get(key)
    if order is None:  # uploaded once per partition and device, not per frame
        order = _ORDER_CACHE[key] = torch.from_numpy(image_order(shard.height, world, shard.strip_rows)).to(local.device)
    return gathered.reshape((world * max_rows,) + tuple(local.shape[1:])).index_select(0, order)


def render_frame_sharded(render_local: Callable[[StripShard], Any], height: int, rank: int, world: int,
                         strip_rows: int = DEFAULT_STRIP_ROWS, dst: int = 0, group: Any = None) -> Optional[Any]:
    """One frame over ``world`` ranks: ``render_local(shard)`` returns this rank's packed rows
    (``[shard.n_rows, ...]``); the assembled frame comes back on rank ``dst``."""
    shard = StripShard(height, world, rank, strip_rows)
    return gather_frame(render_local(shard), shard, dst=dst, group=group)


def render_frames_sharded(render_frame: Callable[[int], Any], n_frames: int, rank: int, world: int) -> dict:
    """An animation batch over ``world`` ranks: calls ``render_frame(k)`` for this rank's
    frames and returns ``{k: result}``; nothing is exchanged."""
    return {k: render_frame(k) for k in frames_of_rank(n_frames, world, rank)}


def covered_exactly_once(height: int, world: int, strip_rows: int = DEFAULT_STRIP_ROWS) -> bool:
    """Partition check used by the tests and by ``bench.py`` before timing."""
    seen = np.zeros(height, dtype=np.int64)
    for s in strip_shards(height, world, strip_rows):
        seen[s.image_rows] += 1
    return bool((seen == 1).all())


class _DeviceBytes:
    """``__cuda_array_interface__`` view of raw device memory so torch can wrap it."""

    def __init__(self, ptr: int, shape: tuple, typestr: str) -> None:
        self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (ptr, False), "version": 3}


class PeerBuffer:
    """``nbytes`` of device memory on rank ``dst`` that every rank of the node can write.

    Rank ``dst`` owns the memory (``rt_peer_buffer_create``) and sends the 64-byte IPC handle
    to the other ranks (``broadcast_object_list``); they map it (``rt_peer_buffer_open``) and
    hand ``ptr`` to a kernel, whose stores then go straight into the owner's memory over NVLink.
    ``complete()`` is the only synchronisation: every rank drains its stream, then a barrier
    tells the owner that everything has landed.
    """

    def __init__(self, nbytes: int, rank: int, world: int, device: int, dst: int = 0, group: Any = None) -> None:
        import ctypes as C

        import torch
        import torch.distributed as dist

        from . import abi

        self.nbytes, self.rank, self.world, self.device, self.dst, self.group = int(nbytes), rank, world, device, dst, group
        self._lib = abi.load()
        self._owner = rank == dst
        ptr = C.c_void_p()
        handle = C.create_string_buffer(abi.RT_IPC_HANDLE_BYTES)
        if self._owner:
            abi.check(self._lib.rt_peer_buffer_create(device, self.nbytes, C.byref(ptr), handle), "rt_peer_buffer_create")
        box = [handle.raw if self._owner else None]
        if world > 1:
            dist.broadcast_object_list(box, src=dst, group=group)
        if not self._owner:
            abi.check(self._lib.rt_peer_buffer_open(device, box[0], C.byref(ptr)), "rt_peer_buffer_open")
        self.ptr = int(ptr.value)
        # the owner reads the buffer through torch; the writers only need the pointer
        self.bytes = None
        if self._owner:
            self.bytes = torch.as_tensor(_DeviceBytes(self.ptr, (self.nbytes,), "|u1"), device=torch.device("cuda", device))

    def complete(self) -> Optional[Any]:
        """Wait until every rank's kernels have finished writing; returns the owner's view (``None`` elsewhere)."""
        import torch
        import torch.distributed as dist

        torch.cuda.current_stream(self.device).synchronize()
        if self.world > 1:
            dist.barrier(group=self.group)
        return self.bytes

    def close(self) -> None:
        if self.ptr:
            if self.world > 1:
                import torch.distributed as dist

                dist.barrier(group=self.group)  # nobody unmaps or frees while a peer may still write
            self.bytes = None
            if self._owner:
                self._lib.rt_peer_buffer_destroy(self.device, self.ptr)
            else:
                self._lib.rt_peer_buffer_close(self.device, self.ptr)
            self.ptr = 0


class PeerFrame(PeerBuffer):
    """A whole-frame RGBA8 (or, ``channels=3``, RGB8) buffer on rank ``dst`` that every rank of
    the node can write: the ranks hand ``ptr`` to ``rt_render_strips_dev`` with
    ``RT_FLAG_IMAGE_ORDER`` and each rank's kernel stores its pixels at their image position while
    it renders, so the exchange step overlaps the compute and no gather or permutation follows."""

    def __init__(self, height: int, width: int, rank: int, world: int, device: int, dst: int = 0,
                 group: Any = None, channels: int = 4) -> None:
        if channels not in (3, 4):
            raise ValueError(f"a peer frame has 3 (RT_FLAG_RGB8) or 4 bytes per pixel, not {channels}")
        super().__init__(height * width * channels, rank, world, device, dst, group)
        self.height, self.width, self.channels = height, width, channels
        self.frame = None if self.bytes is None else self.bytes.view(height, width, channels)

    def complete(self) -> Optional[Any]:
        super().complete()
        return self.frame

    def close(self) -> None:
        self.frame = None
        super().close()


__all__: Sequence[str] = (
    "PeerBuffer", "PeerFrame",
    "DEFAULT_STRIP_ROWS", "StripShard", "strip_shards", "band_rows", "frames_of_rank", "image_order",
    "gather_frame", "render_frame_sharded", "render_frames_sharded", "covered_exactly_once",
)
