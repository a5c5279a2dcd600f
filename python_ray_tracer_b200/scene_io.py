"""Scene and render-settings files.

The reference builds its scenes in Python (``main.py:11-51``, ``README.md:58-90``) and has no
loader; what it does have are two orphaned fixtures that show the intended shape of one:
``tests/testdata/input_scene.json`` (a list of ``{"type": "Sphere" | "Light", ...}`` entries
with ``centerXYZ`` / ``radius`` / ``colorRGB`` / ``reflection`` / ``roughness`` / ``texture``)
and ``tests/testdata/input_render_settings.json`` (``image_width``, ``image_height``,
``max_specular_depth``, ``output_path``, ...).  This module reads and writes that format,
extended with what the domain layer has and the fixtures lack (dome lights, the camera, the
other shading gains), so that a scene can be described in a file and rendered from a shell:

    python -m python_ray_tracer_b200 render --scene scene.json --settings settings.json

Entries (every key not listed is rejected, so typos fail loudly):

``Sphere``     ``centerXYZ``, ``radius``, ``colorRGB`` and optionally ``texture`` (``""`` or
               ``"checker"``), ``diffuse``, ``specular`` (alias ``reflection``, the fixture's
               name: the reference weights the bounce by the specular gain, ``shader.py:106``),
               ``roughness``, ``iridescence``.  Missing gains take the renderer's defaults.
``Light``      ``centerXYZ`` (``intensityRGB`` is accepted and ignored: ``PointLight`` has a
               position only, ``domain.py:26-30``).
``DomeLight``  ``intensity``, ``colorRGB``.
``Camera``     ``positionXYZ`` (at most one; default ``(0, 0.2, -2)``, ``README.md:88``).

Settings: ``image_width``, ``image_height``, ``max_specular_depth`` (ray generations traced,
0 or absent = unlimited like the reference), ``output_path``, ``precision`` (``"fast"`` /
``"exact"``).  ``max_samples_per_pixel`` must be 1 and ``denoise`` false if present (the
reference has neither at HEAD); ``background`` is accepted and ignored (the reference has no
image-based background: rays that miss are black, ``base.py:108``).
"""

from __future__ import annotations

import json
from dataclasses import dataclass, field
from pathlib import Path
from typing import Any, Dict, List, Optional, Sequence, Tuple, Union

from .domain import Camera, DomeLight, PointLight, Scene3D
from .infrastructure.flatten import DefaultMaterial
from .infrastructure.materials import NumpyShader, Texture, TextureChecker
from .infrastructure.shapes import NumpySphere
from .infrastructure.values import NumpyRGBColor, NumpyVector3D, as_float_triple

DEFAULT_EYE = (0.0, 0.2, -2.0)
PathLike = Union[str, Path]

_SPHERE_KEYS = {"type", "centerXYZ", "radius", "colorRGB", "texture", "diffuse", "specular", "reflection",
                "roughness", "iridescence"}
_LIGHT_KEYS = {"type", "centerXYZ", "intensityRGB"}
_DOME_KEYS = {"type", "intensity", "colorRGB"}
_CAMERA_KEYS = {"type", "positionXYZ"}
_SETTINGS_KEYS = {"image_width", "image_height", "max_samples_per_pixel", "max_specular_depth", "background",
                  "denoise", "output_path", "precision"}


class SceneFormatError(ValueError):
    """A scene or settings file that does not follow the format above."""


@dataclass
class RenderSettings:
    """``input_render_settings.json``: image size, depth cap, output path."""

    image_width: int = 1920
    image_height: int = 1080
    max_specular_depth: int = 0  # 0 = unlimited, like the reference
    output_path: Optional[str] = None
    precision: str = "fast"
    ignored: Dict[str, Any] = field(default_factory=dict)  # accepted keys with no effect

    def renderer_kwargs(self) -> Dict[str, Any]:
        return {"max_depth": self.max_specular_depth or None, "precision": self.precision}


def _triple(entry: Dict[str, Any], key: str, where: str) -> Tuple[float, float, float]:
    if key not in entry:
        raise SceneFormatError(f"{where}: missing {key!r}")
    value = entry[key]
    if not isinstance(value, (list, tuple)) or len(value) != 3:
        raise SceneFormatError(f"{where}: {key!r} must be a list of three numbers, got {value!r}")
    try:
        return tuple(float(v) for v in value)  # type: ignore[return-value]
    except (TypeError, ValueError):
        raise SceneFormatError(f"{where}: {key!r} must be a list of three numbers, got {value!r}") from None


def _number(entry: Dict[str, Any], key: str, where: str, default: Optional[float] = None) -> float:
    if key not in entry:
        if default is None:
            raise SceneFormatError(f"{where}: missing {key!r}")
        return default
    value = entry[key]
    if isinstance(value, bool) or not isinstance(value, (int, float)):
        raise SceneFormatError(f"{where}: {key!r} must be a number, got {value!r}")
    return float(value)


def _check_keys(entry: Dict[str, Any], allowed: set, where: str) -> None:
    unknown = sorted(set(entry) - allowed)
    if unknown:
        raise SceneFormatError(f"{where}: unknown key(s) {unknown}; allowed: {sorted(allowed)}")


def scene_from_entries(entries: Sequence[Dict[str, Any]], width: int, height: int,
                       default: Optional[DefaultMaterial] = None) -> Scene3D:
    """Build a ``Scene3D`` from the decoded entry list of a scene file."""
    default = default or DefaultMaterial()
    if not isinstance(entries, (list, tuple)):
        raise SceneFormatError("a scene file holds a JSON list of entries")
    shapes: List[Any] = []
    lights: List[Any] = []
    eye: Optional[Tuple[float, float, float]] = None
    for i, entry in enumerate(entries):
        if not isinstance(entry, dict) or "type" not in entry:
            raise SceneFormatError(f"entry {i}: expected an object with a 'type'")
        kind = entry["type"]
        where = f"entry {i} ({kind})"
        if kind == "Sphere":
            _check_keys(entry, _SPHERE_KEYS, where)
            if "specular" in entry and "reflection" in entry:
                raise SceneFormatError(f"{where}: 'reflection' is an alias of 'specular'; give one")
            radius = _number(entry, "radius", where)
            if not radius > 0.0:
                raise SceneFormatError(f"{where}: radius must be > 0, got {radius!r}")
            texture_name = entry.get("texture", "") or ""
            if texture_name not in ("", "checker"):
                raise SceneFormatError(f"{where}: texture must be '' or 'checker', got {texture_name!r}")
            colour = NumpyRGBColor(*_triple(entry, "colorRGB", where))
            texture = TextureChecker(colour) if texture_name == "checker" else Texture(colour)
            specular = _number(entry, "specular" if "specular" in entry else "reflection", where, default.specular_gain)
            shader = NumpyShader(
                default.reflection_gain,
                specular,
                _number(entry, "roughness", where, default.specular_roughness),
                _number(entry, "iridescence", where, default.iridescence_gain),
                _number(entry, "diffuse", where, default.diffuse_gain),
                texture,
            )
            shapes.append(NumpySphere(NumpyVector3D(*_triple(entry, "centerXYZ", where)), radius, shader))
        elif kind == "Light":
            _check_keys(entry, _LIGHT_KEYS, where)
            if "intensityRGB" in entry:
                _triple(entry, "intensityRGB", where)  # validated, then ignored (see module docstring)
            lights.append(PointLight(NumpyVector3D(*_triple(entry, "centerXYZ", where))))
        elif kind == "DomeLight":
            _check_keys(entry, _DOME_KEYS, where)
            lights.append(DomeLight(_number(entry, "intensity", where), NumpyRGBColor(*_triple(entry, "colorRGB", where))))
        elif kind == "Camera":
            _check_keys(entry, _CAMERA_KEYS, where)
            if eye is not None:
                raise SceneFormatError(f"{where}: a scene has one camera")
            eye = _triple(entry, "positionXYZ", where)
        else:
            raise SceneFormatError(f"entry {i}: unknown type {kind!r} (Sphere, Light, DomeLight, Camera)")
    if not shapes:
        raise SceneFormatError("a scene needs at least one Sphere")
    if not lights or not isinstance(lights[0], PointLight):
        # shader.py:75 reads lights[0].position: the first light of the file must be a Light
        raise SceneFormatError("the first light entry must be a Light (the reference shades with lights[0], shader.py:75)")
    return Scene3D(shapes, lights, Camera(NumpyVector3D(*(eye or DEFAULT_EYE)), int(width), int(height)))


def settings_from_dict(data: Dict[str, Any]) -> RenderSettings:
    if not isinstance(data, dict):
        raise SceneFormatError("a settings file holds a JSON object")
    _check_keys(data, _SETTINGS_KEYS, "settings")
    settings = RenderSettings()
    for key in ("image_width", "image_height", "max_specular_depth"):
        if key in data:
            value = data[key]
            if isinstance(value, bool) or not isinstance(value, int) or value < (0 if key == "max_specular_depth" else 1):
                raise SceneFormatError(f"settings: {key} must be a positive integer, got {value!r}")
            setattr(settings, key, value)
    if data.get("max_samples_per_pixel", 1) != 1:
        raise SceneFormatError("settings: max_samples_per_pixel must be 1 (the path has one sample per pixel, base.py:130-141)")
    if data.get("denoise", False):
        raise SceneFormatError("settings: denoise is not part of the path (denoiser.py is not imported by anything)")
    if "precision" in data:
        if data["precision"] not in ("fast", "exact"):
            raise SceneFormatError(f"settings: precision must be 'fast' or 'exact', got {data['precision']!r}")
        settings.precision = data["precision"]
    if data.get("output_path") is not None:
        settings.output_path = str(data["output_path"])
    for key in ("background", "max_samples_per_pixel", "denoise"):
        if key in data:
            settings.ignored[key] = data[key]
    return settings


def _read_json(path: PathLike) -> Any:
    try:
        return json.loads(Path(path).read_text())
    except json.JSONDecodeError as err:
        raise SceneFormatError(f"{path}: not valid JSON ({err})") from None


def load_settings(path: PathLike) -> RenderSettings:
    return settings_from_dict(_read_json(path))


def load_scene(scene_path: PathLike, settings: Union[None, PathLike, RenderSettings] = None,
               default: Optional[DefaultMaterial] = None) -> Tuple[Scene3D, RenderSettings]:
    """``(Scene3D, RenderSettings)`` from a scene file and an optional settings file."""
    if settings is None:
        settings = RenderSettings()
    elif not isinstance(settings, RenderSettings):
        settings = load_settings(settings)
    scene = scene_from_entries(_read_json(scene_path), settings.image_width, settings.image_height, default)
    return scene, settings


def scene_to_entries(scene: Any, default: Optional[DefaultMaterial] = None) -> List[Dict[str, Any]]:
    """The entry list that ``scene_from_entries`` turns back into an equal scene.  Reads the
    same attributes as the flattener, so scenes built with the reference's classes work."""
    from .infrastructure.flatten import _is_checker, _is_dome_light, _is_shader  # attribute-based tests

    default = default or DefaultMaterial()
    entries: List[Dict[str, Any]] = [{"type": "Camera", "positionXYZ": list(as_float_triple(scene.camera.position, "camera.position"))}]
    for shape in scene.shapes:
        centre = getattr(shape, "center", None)
        if centre is None:
            centre = shape.position
        entry: Dict[str, Any] = {"type": "Sphere", "centerXYZ": list(as_float_triple(centre, "center")), "radius": float(shape.radius)}
        shader = getattr(shape, "shader", None)
        if _is_shader(shader):
            texture = shader.diffuse_color
            entry.update(specular=float(shader.specular_gain), roughness=float(shader.specular_roughness),
                         iridescence=float(shader.iridescence_gain), diffuse=float(shader.diffuse_gain))
        else:
            texture = shape.material_texture() if hasattr(shape, "material_texture") else Texture(shader)
            entry.update(specular=default.specular_gain, roughness=default.specular_roughness,
                         iridescence=default.iridescence_gain, diffuse=default.diffuse_gain)
        colour = getattr(texture, "color", None)
        entry["colorRGB"] = list(as_float_triple(colour, "color")) if colour is not None else [1.0, 1.0, 1.0]
        entry["texture"] = "checker" if _is_checker(texture) else ""
        entries.append(entry)
    for light in scene.lights:
        if _is_dome_light(light):
            entries.append({"type": "DomeLight", "intensity": float(light.intensity), "colorRGB": list(as_float_triple(light.color, "color"))})
        else:
            entries.append({"type": "Light", "centerXYZ": list(as_float_triple(light.position, "light.position"))})
    return entries


def save_scene(scene: Any, scene_path: PathLike, settings_path: Optional[PathLike] = None,
               settings: Optional[RenderSettings] = None, default: Optional[DefaultMaterial] = None) -> None:
    """Write ``scene`` (and, if asked, its settings) in the format ``load_scene`` reads."""
    Path(scene_path).write_text(json.dumps(scene_to_entries(scene, default), indent=2) + "\n")
    if settings_path is not None:
        settings = settings or RenderSettings()
        data = {"image_width": int(scene.camera.width), "image_height": int(scene.camera.height),
                "max_specular_depth": int(settings.max_specular_depth), "precision": settings.precision}
        if settings.output_path is not None:
            data["output_path"] = settings.output_path
        Path(settings_path).write_text(json.dumps(data, indent=2) + "\n")
