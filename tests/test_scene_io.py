"""Scene / settings files: host logic, no GPU."""
import json

import numpy as np
import pytest

from python_ray_tracer_b200 import scene_io, scenes
from python_ray_tracer_b200.infrastructure.flatten import DefaultMaterial, flatten_scene

README = DefaultMaterial(**scenes.README_GAINS)

# the two orphaned fixtures of the reference (tests/testdata/*.json), restated
FIXTURE_SCENE = [
    {"type": "Sphere", "centerXYZ": [0.0, 0.0, 6], "radius": 4.0, "colorRGB": [0.5, 0.5, 0.5], "reflection": 1.0,
     "roughness": 0.5, "texture": ""},
    {"type": "Light", "centerXYZ": [0.0, 6.0, 6.689], "intensityRGB": [1.0, 1.0, 1.0]},
]
FIXTURE_SETTINGS = {"image_width": 30, "image_height": 30, "max_samples_per_pixel": 1, "max_specular_depth": 1,
                    "background": "sourceimages/hdri_1.jpg", "denoise": False, "output_path": "data/output_50.png"}


def _same(a, b):
    fa, fb = flatten_scene(a, README), flatten_scene(b, README)
    return fa.key() == fb.key() and (a.camera.width, a.camera.height) == (b.camera.width, b.camera.height)


@pytest.mark.parametrize("build", [lambda: scenes.readme_scene(64, 36), lambda: scenes.main_scene(48, 27),
                                   lambda: scenes.random_spheres_scene(32, 18, n_spheres=40, seed=3),
                                   lambda: scenes.turntable_scene(17, 256, 40, 30)])
def test_round_trip_gives_the_same_flat_scene(tmp_path, build):
    scene = build()
    scene_io.save_scene(scene, tmp_path / "s.json", tmp_path / "r.json", default=README)
    back, settings = scene_io.load_scene(tmp_path / "s.json", tmp_path / "r.json", default=README)
    assert _same(scene, back)
    assert (settings.image_width, settings.image_height) == (scene.camera.width, scene.camera.height)
    # and a second trip writes the same bytes
    scene_io.save_scene(back, tmp_path / "s2.json", default=README)
    assert (tmp_path / "s.json").read_text() == (tmp_path / "s2.json").read_text()


def test_reference_fixture_format_is_accepted(tmp_path):
    (tmp_path / "scene.json").write_text(json.dumps(FIXTURE_SCENE))
    (tmp_path / "settings.json").write_text(json.dumps(FIXTURE_SETTINGS))
    scene, settings = scene_io.load_scene(tmp_path / "scene.json", tmp_path / "settings.json")
    assert (scene.camera.width, scene.camera.height) == (30, 30) and settings.max_specular_depth == 1
    assert settings.output_path == "data/output_50.png" and settings.renderer_kwargs() == {"max_depth": 1, "precision": "fast"}
    assert settings.ignored == {"background": "sourceimages/hdri_1.jpg", "max_samples_per_pixel": 1, "denoise": False}
    flat = flatten_scene(scene)
    assert flat.n_spheres == 1 and flat.spheres[3, 0] == 4.0 and flat.spheres[5, 0] == 1.0  # 'reflection' -> specular gain
    assert flat.point_lights.tolist() == [[0.0, 6.0, 6.689]] and flat.eye == scene_io.DEFAULT_EYE
    # defaults when no settings file is given
    scene, settings = scene_io.load_scene(tmp_path / "scene.json")
    assert (scene.camera.width, scene.camera.height, settings.max_specular_depth) == (1920, 1080, 0)


@pytest.mark.parametrize("entries, message", [
    ({"type": "Sphere"}, "JSON list"),
    ([], "at least one Sphere"),
    ([{"centerXYZ": [0, 0, 0]}], "'type'"),
    ([{"type": "Cube"}], "unknown type"),
    ([{"type": "Sphere", "centerXYZ": [0, 0], "radius": 1, "colorRGB": [1, 1, 1]}], "three numbers"),
    ([{"type": "Sphere", "centerXYZ": [0, 0, 0], "radius": 0, "colorRGB": [1, 1, 1]}], "radius must be > 0"),
    ([{"type": "Sphere", "centerXYZ": [0, 0, 0], "radius": "1", "colorRGB": [1, 1, 1]}], "must be a number"),
    ([{"type": "Sphere", "centerXYZ": [0, 0, 0], "radius": 1, "colorRGB": [1, 1, 1], "colour": 1}], "unknown key"),
    ([{"type": "Sphere", "centerXYZ": [0, 0, 0], "radius": 1, "colorRGB": [1, 1, 1], "texture": "marble"}], "texture must be"),
    ([{"type": "Sphere", "centerXYZ": [0, 0, 0], "radius": 1, "colorRGB": [1, 1, 1], "specular": 1, "reflection": 1}], "alias"),
    ([{"type": "Sphere", "centerXYZ": [0, 0, 0], "radius": 1, "colorRGB": [1, 1, 1]}], "first light entry"),
    ([{"type": "Sphere", "centerXYZ": [0, 0, 0], "radius": 1, "colorRGB": [1, 1, 1]},
      {"type": "DomeLight", "intensity": 0.1, "colorRGB": [1, 1, 1]}, {"type": "Light", "centerXYZ": [0, 1, 0]}], "first light entry"),
    ([{"type": "Sphere", "centerXYZ": [0, 0, 0], "radius": 1, "colorRGB": [1, 1, 1]}, {"type": "Light", "centerXYZ": [0, 1, 0]},
      {"type": "Camera", "positionXYZ": [0, 0, -1]}, {"type": "Camera", "positionXYZ": [0, 0, -2]}], "one camera"),
])
def test_malformed_scenes_fail_loudly(entries, message):
    with pytest.raises(scene_io.SceneFormatError, match=message):
        scene_io.scene_from_entries(entries, 8, 8)


@pytest.mark.parametrize("data, message", [
    ([], "JSON object"), ({"image_width": 0}, "positive integer"), ({"image_height": 2.5}, "positive integer"),
    ({"max_samples_per_pixel": 4}, "must be 1"), ({"denoise": True}, "denoise"), ({"precision": "half"}, "precision"),
    ({"width": 3}, "unknown key"), ({"max_specular_depth": -1}, "positive integer"),
])
def test_malformed_settings_fail_loudly(data, message):
    with pytest.raises(scene_io.SceneFormatError, match=message):
        scene_io.settings_from_dict(data)


def test_cli_export_and_errors(tmp_path, capsys):
    from python_ray_tracer_b200.__main__ import main

    assert main(["export-readme-scene", str(tmp_path / "s.json"), str(tmp_path / "r.json"), "--width", "64", "--height", "36"]) == 0
    scene, settings = scene_io.load_scene(tmp_path / "s.json", tmp_path / "r.json", default=README)
    assert _same(scene, scenes.readme_scene(64, 36))
    (tmp_path / "bad.json").write_text("{not json")
    assert main(["render", "--scene", str(tmp_path / "bad.json"), "--out", str(tmp_path / "x.png")]) == 2
    assert "not valid JSON" in capsys.readouterr().err
    assert main(["render", "--scene", str(tmp_path / "s.json")]) == 2  # no output path anywhere
    assert "no output path" in capsys.readouterr().err
    assert np.isfinite(flatten_scene(scene, README).spheres).all()
