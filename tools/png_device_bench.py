"""Device PNG encoder alone (GPU box): time of rt_png_encode_bands_dev, rt_png_pack_dev and rt_png_encode_file_dev on a rendered frame.
    python tools/png_device_bench.py [width height] [reps]"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from python_ray_tracer_b200 import scenes  # noqa: E402
from python_ray_tracer_b200.infrastructure import CudaRenderer  # noqa: E402
from python_ray_tracer_b200.infrastructure.cuda import abi  # noqa: E402
from python_ray_tracer_b200.infrastructure.png import DevicePngEncoder  # noqa: E402

w, h = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1920, 1080)
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 20
lib = abi.load()
scene = scenes.readme_scene(w, h)
r = CudaRenderer(**scenes.README_GAINS)
img = r.render_rows(r.flatten(scene), scene.camera, 0, h, want_rgba8=True, flags=abi.RT_FLAG_RGB8, no_planes=True).rgba8
enc = DevicePngEncoder(w, h)
dev_file = torch.empty(enc.capacity, dtype=torch.uint8, device="cuda")
dev_size = torch.zeros(1, dtype=torch.int64, device="cuda")
stream = torch.cuda.current_stream()
sp = int(stream.cuda_stream)


def bands():
    abi.check(lib.rt_png_encode_bands_dev(0, img.data_ptr(), 3, w * 3, h, w, h, 0, h, 1, enc.scratch.data_ptr(), enc.scratch_bytes, sp), "bands")


def pack(host: bool):
    abi.check(lib.rt_png_pack_dev(0, w, h, enc.scratch.data_ptr(), enc.file.data_ptr() if host else dev_file.data_ptr(), enc.capacity,
                                  enc.size.data_ptr() if host else dev_size.data_ptr(), sp), "pack")


def whole(host: bool):
    abi.check(lib.rt_png_encode_file_dev(0, img.data_ptr(), 3, w * 3, w, h, enc.scratch.data_ptr(), enc.scratch_bytes,
                                         enc.file.data_ptr() if host else dev_file.data_ptr(), enc.capacity,
                                         enc.size.data_ptr() if host else dev_size.data_ptr(), sp), "file")


def timed(fn):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


print(f"{w}x{h}: rows per band {enc.rows_per_band}, bands {-(-h // enc.rows_per_band)}, raw {img.numel()} bytes")
print(f"  band kernel          {timed(bands):8.1f} us")
print(f"  pack -> device       {timed(lambda: pack(False)):8.1f} us   file {int(dev_size[0])} bytes")
print(f"  pack -> pinned host  {timed(lambda: pack(True)):8.1f} us")
print(f"  one launch -> device {timed(lambda: whole(False)):8.1f} us   file {int(dev_size[0])} bytes   (rt_png_encode_file_dev: memset + kernel)")
print(f"  one launch -> host   {timed(lambda: whole(True)):8.1f} us")
