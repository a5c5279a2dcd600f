"""Is the device-to-host rate of N ranks copying at once ours or the box's?  (torchrun, one rank per GPU.)
Plain cudaMemcpyAsync D2H of one 1080p RGB8 image (6.2 MB) and of the f32 planes (24.9 MB) per rank, no kernel:
all ranks together vs one rank at a time; default pinned memory vs write-combined; after binding to the GPU's CPUs.
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/d2h_aggregate.py"""
import ctypes as C
import os
import sys
import time
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from python_ray_tracer_b200.infrastructure.cuda import affinity  # noqa: E402

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
bound = affinity.bind_to_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
cudart = C.CDLL("libcudart.so.12")


def host_alloc(nbytes: int, flags: int):
    p = C.c_void_p()
    assert cudart.cudaHostAlloc(C.byref(p), C.c_size_t(nbytes), C.c_uint(flags)) == 0
    return p


def run(nbytes: int, flags: int, together: bool, reps: int = 40) -> float:
    dev = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    host = host_alloc(nbytes, flags)
    stream = torch.cuda.current_stream()
    secs = 0.0
    for turn in range(1 if together else world):
        dist.barrier()
        torch.cuda.synchronize()
        if together or turn == rank:
            t0 = time.perf_counter()
            for _ in range(reps):
                assert cudart.cudaMemcpyAsync(host, C.c_void_p(dev.data_ptr()), C.c_size_t(nbytes), C.c_int(2), C.c_void_p(stream.cuda_stream)) == 0
            stream.synchronize()
            secs = (time.perf_counter() - t0) / reps
    dist.barrier()
    cudart.cudaFreeHost(host)
    t = torch.tensor([nbytes / secs / 1e9], device="cuda")
    out = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(out, t)
    return [float(v) for v in out]


if rank == 0:
    print(f"{world} ranks; rank 0 bound to {len(bound) if bound else 'no'} CPUs")
for label, nbytes in (("RGB8 1080p 6.2 MB", 1920 * 1080 * 3), ("f32 planes 24.9 MB", 1920 * 1080 * 12)):
    for fname, flags in (("pinned (portable)", 1), ("pinned write-combined", 1 | 4)):
        for together in (False, True):
            rates = run(nbytes, flags, together)
            if rank == 0:
                print(f"{label:20s} {fname:22s} {'all ranks at once' if together else 'one rank at a time'}: per rank "
                      f"{min(rates):5.1f}-{max(rates):5.1f} GB/s, sum {sum(rates):6.1f} GB/s")
dist.destroy_process_group()
