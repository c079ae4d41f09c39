"""The reference's JSON scene wire format: the `Into<JsonValue>` / `TryFrom<JsonValue>`
impls of the Rust tree, restated for the Python host side so that scenes can cross the C ABI as text.

    Vec3          {"x":..,"y":..,"z":..}                                   vec3.rs:22-50
    Material      {"metallicness":..,"opacity":..,"ir":..}                 materials.rs:21-54
                  (`emmited` is NOT serialised; parsing sets it to 0)
    ImageTexture  {"row":W,"col":H,"img":[Vec3,...]}                       texture.rs:28-59,268-276
                  (noise is NOT serialised; parsing sets noise=None, noise_scale=1)
    Sphere        {"origin","radius","col_mod","material","velocity","texture"}   sphere.rs:44-90
    Scene         {"spheres":[Sphere,...]}  -> Scene::new_sphere            viewport.rs:174-205
                  (quads and instances are NOT serialised)

Errors follow the reference: any missing or mistyped field is a ParseError (viewport.rs:2-26), never a
partial scene. Numbers are f32 on both sides (`as_f32`). Equality helpers restate the reference's
PartialEq impls (vec3.rs:17-21 tolerance 1e-7 per component; sphere.rs:22-30 ignores the texture;
viewport.rs:152-172 zips the lists, so a shorter list compares equal to its prefix).
"""
from __future__ import annotations

import json
from typing import Any, Dict, List

import numpy as np

from .scene import ImageTexture, Material, Scene, Sphere

F = np.float32


class ParseError(ValueError):
    """viewport.rs:2-26 `errors::ParseError { source: Option<json::Error> }`."""

    def __init__(self, source: Exception | None = None):
        super().__init__(f"ParseError: {source if source is not None else 'None'}")
        self.source = source


# ----------------------------------------------------------------------------- to JSON values
def _f(x) -> float:
    return float(F(x))


def vec3_to_json(v) -> Dict[str, float]:                      # vec3.rs:22-30
    a = np.asarray(v, dtype=F).reshape(3)
    return {"x": _f(a[0]), "y": _f(a[1]), "z": _f(a[2])}


def material_to_json(m: Material) -> Dict[str, float]:        # materials.rs:21-29
    return {"metallicness": _f(m.metallicness), "opacity": _f(m.opacity), "ir": _f(m.ir)}


def texture_to_json(t: ImageTexture) -> Dict[str, Any]:       # texture.rs:268-276
    return {"row": int(t.row), "col": int(t.col), "img": [vec3_to_json(p) for p in np.asarray(t.img, dtype=F).reshape(-1, 3)]}


def sphere_to_json(s: Sphere) -> Dict[str, Any]:              # sphere.rs:44-55
    return {"origin": vec3_to_json(s.origin), "radius": _f(s.radius), "col_mod": vec3_to_json(s.col_mod),
            "material": material_to_json(s.mat), "velocity": vec3_to_json(s.velocity),
            "texture": texture_to_json(s.texture)}


def scene_to_json(scene: Scene) -> Dict[str, Any]:            # viewport.rs:174-180
    return {"spheres": [sphere_to_json(s) for s in scene.spheres]}


# ----------------------------------------------------------------------------- from JSON values
def _as_f32(value: Any, key: str) -> float:
    """json::JsonValue::as_f32 on value[key]: a number or nothing. (bool is not a number there.)"""
    v = value.get(key) if isinstance(value, dict) else None
    if isinstance(v, bool) or not isinstance(v, (int, float)):
        raise ParseError()
    return _f(v)


def _as_usize(value: Any, key: str) -> int:
    v = value.get(key) if isinstance(value, dict) else None
    if isinstance(v, bool) or not isinstance(v, (int, float)) or v < 0 or v != int(v):
        raise ParseError()
    return int(v)


def vec3_from_json(value: Any) -> np.ndarray:                 # vec3.rs:31-50
    return np.array([_as_f32(value, "x"), _as_f32(value, "y"), _as_f32(value, "z")], dtype=F)


def material_from_json(value: Any) -> Material:               # materials.rs:30-54 (emmited = 0)
    return Material(_as_f32(value, "metallicness"), _as_f32(value, "opacity"), _as_f32(value, "ir"), (0.0, 0.0, 0.0))


def texture_from_json(value: Any) -> ImageTexture:            # texture.rs:28-59
    img = value.get("img") if isinstance(value, dict) else None
    if not isinstance(img, list):
        raise ParseError()
    texels = [vec3_from_json(p) for p in img]
    row, col = _as_usize(value, "row"), _as_usize(value, "col")
    try:    # the reference accepts row*col != len(img) and panics at the first texel fetch past the end;
        return ImageTexture(np.asarray(texels, dtype=F).reshape(-1, 3), row, col, None, 1.0)   # here it is a ParseError
    except ValueError as e:
        raise ParseError(e) from e


def sphere_from_json(value: Any) -> Sphere:                   # sphere.rs:56-90
    if not isinstance(value, dict):
        raise ParseError()
    return Sphere(vec3_from_json(value.get("origin")), _as_f32(value, "radius"), vec3_from_json(value.get("col_mod")),
                  material_from_json(value.get("material")), vec3_from_json(value.get("velocity")),
                  texture_from_json(value.get("texture")))


def scene_from_json(value: Any) -> Scene:                     # viewport.rs:181-205
    sp = value.get("spheres") if isinstance(value, dict) else None
    if not isinstance(sp, list):
        raise ParseError()
    return Scene.new_sphere([sphere_from_json(s) for s in sp])


# ----------------------------------------------------------------------------- text
def dumps(scene: Scene, pretty: bool = False) -> str:
    """`json::stringify(scene.into())` / `{:#}`."""
    return json.dumps(scene_to_json(scene), indent=4 if pretty else None, separators=None if pretty else (",", ":"))


def loads(text: str) -> Scene:
    """`Scene::try_from(json::parse(text)?)`."""
    try:
        value = json.loads(text)
    except json.JSONDecodeError as e:
        raise ParseError(e) from e
    return scene_from_json(value)


# ----------------------------------------------------------------------------- PartialEq restated
def vec3_eq(a, b) -> bool:                                    # vec3.rs:17-21 + close_to_zero :259-261
    d = np.asarray(a, dtype=F) - np.asarray(b, dtype=F)
    return bool((np.abs(d) < F(1e-7)).all())


def material_eq(a: Material, b: Material) -> bool:            # derive(PartialEq): exact, emmited via Vec3's eq
    return a.metallicness == b.metallicness and a.opacity == b.opacity and a.ir == b.ir and vec3_eq(a.emmited, b.emmited)


def sphere_eq(a: Sphere, b: Sphere) -> bool:                  # sphere.rs:22-30 (texture not compared)
    return (vec3_eq(a.origin, b.origin) and a.radius == b.radius and vec3_eq(a.col_mod, b.col_mod)
            and material_eq(a.mat, b.mat) and vec3_eq(a.velocity, b.velocity))


def scene_eq(a: Scene, b: Scene) -> bool:                     # viewport.rs:152-172 (zip: prefix equality; quads/instances
    return all(sphere_eq(x, y) for x, y in zip(a.spheres, b.spheres))   # only have derived/partial eq and are not serialised)
