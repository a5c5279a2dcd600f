"""Command line of the package.

    python -m python_ray_tracer_b200 render --scene scene.json [--settings settings.json] [--out frame.png]
    python -m python_ray_tracer_b200 turntable --frames 256 --out frames/
    python -m python_ray_tracer_b200 export-readme-scene scene.json settings.json

``render`` is ``render_image_pipeline`` (``application.py:43-52``) on a scene read from a file;
``turntable`` is BASELINE config 5 (see ``turntable.py``).
"""

from __future__ import annotations

import argparse
import sys
import time
from pathlib import Path


def _render(args: argparse.Namespace) -> int:
    from . import render_image_pipeline, scene_io
    from .infrastructure import CudaRenderer

    try:
        scene, settings = scene_io.load_scene(args.scene, args.settings)
    except (scene_io.SceneFormatError, OSError) as err:
        print(f"error: {err}", file=sys.stderr)
        return 2
    out = args.out or settings.output_path
    if out is None:
        print("error: no output path (give --out or output_path in the settings file)", file=sys.stderr)
        return 2
    kwargs = settings.renderer_kwargs()
    if args.precision:
        kwargs["precision"] = args.precision
    renderer = CudaRenderer(device=args.device, **kwargs)
    Path(out).parent.mkdir(parents=True, exist_ok=True)
    t0 = time.perf_counter()
    render_image_pipeline(scene, Path(out), renderer)
    print(f"{out}: {scene.camera.width}x{scene.camera.height}, {len(scene.shapes)} shapes, "
          f"{time.perf_counter() - t0:.3f} s (render + PNG)")
    return 0


def _export(args: argparse.Namespace) -> int:
    from . import scene_io, scenes

    scene = scenes.readme_scene(args.width, args.height)
    scene_io.save_scene(scene, args.scene, args.settings, default=scene_io.DefaultMaterial(**scenes.README_GAINS))
    return 0


def main(argv=None) -> int:
    parser = argparse.ArgumentParser(prog="python -m python_ray_tracer_b200", description=__doc__.split("\n")[0])
    sub = parser.add_subparsers(dest="command", required=True)
    render = sub.add_parser("render", help="render a scene file to a PNG")
    render.add_argument("--scene", type=Path, required=True)
    render.add_argument("--settings", type=Path, default=None)
    render.add_argument("--out", type=Path, default=None)
    render.add_argument("--device", type=int, default=0)
    render.add_argument("--precision", choices=["fast", "exact"], default=None)
    render.set_defaults(run=_render)
    export = sub.add_parser("export-readme-scene", help="write the README scene in the file format")
    export.add_argument("scene", type=Path)
    export.add_argument("settings", type=Path, nargs="?", default=None)
    export.add_argument("--width", type=int, default=1920)
    export.add_argument("--height", type=int, default=1080)
    export.set_defaults(run=_export)
    sub.add_parser("turntable", help="BASELINE config 5; see python -m python_ray_tracer_b200.turntable --help", add_help=False)
    args, rest = parser.parse_known_args(argv)
    if args.command == "turntable":
        from . import turntable

        sys.argv = [sys.argv[0] + " turntable"] + rest
        turntable.main()
        return 0
    if rest:
        parser.error(f"unrecognised arguments: {' '.join(rest)}")
    return args.run(args)


if __name__ == "__main__":
    sys.exit(main())
