"""JSON scene wire format: the reference's Into<JsonValue>/TryFrom<JsonValue> impls
(vec3.rs:22-50, materials.rs:21-54, texture.rs:28-59,268-276, sphere.rs:44-90, viewport.rs:174-205).
The reference's own asserting test is `deserialize_test` (viewport/json_tests.rs:49-92): scene -> JSON ->
scene, `assert_eq!(scene, json_s)`; it is restated here on the same three spheres, plus the error and
lossiness rules of the format and a render through the C ABI of a scene that came in as text."""
import json

import numpy as np
import pytest

import rtw_b200 as rtw
from rtw_b200 import _abi as A, scene_json as J


def _reference_test_scene():
    # json_tests.rs:51-86
    return rtw.Scene.new_sphere([
        rtw.Sphere.new(rtw.Vec3(-0.5, 0.0, -1.0), 0.5, rtw.Vec3(0.6, 0.6, 0.6), rtw.SCATTER_M),
        rtw.Sphere.new(rtw.Vec3(0.5, 0.0, -1.0), 0.5, rtw.Vec3(1.0, 1.0, 1.0), rtw.SCATTER_M),
        rtw.Sphere.new(rtw.Vec3(0.0, 0.0, -2.0), 1.0, rtw.Vec3(0.5, 1.0, 0.0), rtw.METALLIC_M),
    ])


def test_deserialize_test_of_the_reference():
    scene = _reference_test_scene()
    obj = J.scene_to_json(scene)
    back = J.scene_from_json(obj)
    assert J.scene_eq(scene, back)                       # json_tests.rs:92
    assert len(back.spheres) == 3 and not back.quads and not back.instances
    # and through text
    assert J.scene_eq(scene, J.loads(J.dumps(scene)))
    assert J.scene_eq(scene, J.loads(J.dumps(scene, pretty=True)))


def test_shape_of_the_document():
    obj = J.scene_to_json(_reference_test_scene())
    assert list(obj) == ["spheres"]
    s0 = obj["spheres"][0]
    assert list(s0) == ["origin", "radius", "col_mod", "material", "velocity", "texture"]     # sphere.rs:46-53
    assert s0["origin"] == {"x": -0.5, "y": 0.0, "z": -1.0}
    assert list(s0["material"]) == ["metallicness", "opacity", "ir"]                           # no `emmited`
    assert list(s0["texture"]) == ["row", "col", "img"] and s0["texture"]["row"] == 1          # Sphere::new: 1x1 texture of col_mod
    assert s0["texture"]["img"] == [{"x": float(np.float32(0.6))} | {"y": float(np.float32(0.6)), "z": float(np.float32(0.6))}]
    assert s0["radius"] == 0.5


def test_lossy_fields_follow_the_reference():
    glow = rtw.Material.new_emmiting(0.0, 0.0, 1.0, rtw.Vec3(4.0, 4.0, 4.0))
    tex = rtw.ImageTexture.from_color_noise(rtw.Vec3(1.0, 1.0, 1.0), 0.25)
    sc = rtw.Scene.new(
        [rtw.Sphere.new_moving_with_texture(rtw.Vec3(0, 1, 0), 2.0, None, glow, rtw.Vec3(1, 0, 0), tex)],
        [rtw.Quad.new(rtw.Vec3(0, 0, 0), rtw.Vec3(1, 0, 0), rtw.Vec3(0, 1, 0), rtw.SCATTER_M, rtw.Vec3(0, 0, 0),
                      rtw.ImageTexture.from_color(rtw.Vec3(1, 1, 1)))], [])
    back = J.loads(J.dumps(sc))
    s = back.spheres[0]
    assert not back.quads                                            # viewport.rs:174-180: spheres only
    assert np.array_equal(s.mat.emmited, np.zeros(3, np.float32))    # materials.rs:47-51
    assert s.texture.noise is None and s.texture.noise_scale == 1.0  # texture.rs:52-58
    assert np.array_equal(s.velocity, np.array([1, 0, 0], np.float32))
    assert not J.scene_eq(sc, back)                                  # the material differs (emmited)


@pytest.mark.parametrize("mutate", [
    lambda o: o.pop("spheres"),
    lambda o: o.__setitem__("spheres", {}),
    lambda o: o["spheres"][0].pop("radius"),
    lambda o: o["spheres"][0].__setitem__("radius", "0.5"),
    lambda o: o["spheres"][0]["origin"].pop("z"),
    lambda o: o["spheres"][1]["material"].pop("ir"),
    lambda o: o["spheres"][1].pop("velocity"),
    lambda o: o["spheres"][2].pop("texture"),
    lambda o: o["spheres"][2]["texture"].__setitem__("img", 3),
    lambda o: o["spheres"][2]["texture"].pop("row"),
    lambda o: o["spheres"][2]["texture"].__setitem__("col", 2),     # row*col != len(img)
])
def test_any_missing_or_mistyped_field_is_a_parse_error(mutate):
    obj = J.scene_to_json(_reference_test_scene())
    mutate(obj)
    with pytest.raises(J.ParseError):
        J.scene_from_json(obj)
    with pytest.raises(J.ParseError):
        J.loads("{not json")


def test_partial_eq_rules():
    a = _reference_test_scene()
    b = _reference_test_scene()
    b.spheres[0].origin = b.spheres[0].origin + np.float32(5e-8)     # inside Vec3's 1e-7 tolerance (vec3.rs:17-21)
    assert J.scene_eq(a, b)
    b.spheres[0].origin = b.spheres[0].origin + np.float32(1e-3)
    assert not J.scene_eq(a, b)
    c = _reference_test_scene()
    c.spheres.pop()                                                  # zip stops at the shorter list (viewport.rs:154)
    assert J.scene_eq(a, c)
    d = _reference_test_scene()
    d.spheres[1].texture = rtw.ImageTexture.from_color(rtw.Vec3(0, 0, 0))   # textures are not compared (sphere.rs:22-30)
    assert J.scene_eq(a, d)


def test_text_scene_renders_like_the_built_one(oracle):
    scene = _reference_test_scene()
    back = J.loads(J.dumps(scene))
    vp = rtw.Viewport.new_from_res(40, 30, 1, 10, 2.0, 90.0, rtw.Vec3(0, 0, 0), rtw.Vec3(0, 0, -1), rtw.Vec3(0, 1, 0), None, 0.0)
    a = vp.render_no_rand(rtw.ray_color_albedo, scene, oracle)
    b = vp.render_no_rand(rtw.ray_color_albedo, back, oracle)
    assert np.array_equal(a, b)


@pytest.mark.gpu
def test_gpu_renders_a_scene_that_came_in_as_text(oracle, cuda):
    from rtw_b200 import scenes
    job = scenes.CONFIGS["C2"]()                                     # ~488 spheres, metal / glass / diffuse
    text = J.dumps(job.scene)
    back = J.loads(text)
    assert len(back.spheres) == len(job.scene.spheres) and len(text) > 10000
    vp = rtw.Viewport.new_from_res(200, 112, 4, 20, 2.0, job.viewport.vfov, job.viewport.origin, job.viewport.direction,
                                   job.viewport.vup, None, 0.0)
    vp.flags = A.RTW_FLAG_STRICT_FP
    for ig in (rtw.ray_color_albedo, rtw.ray_color_normal):
        ref = vp.render_no_rand(ig, job.scene, oracle)
        tie = oracle.scene(job.scene).last_tie.astype(bool)
        got = vp.render_no_rand(ig, back, cuda)
        ok = ~tie[..., None] & np.isfinite(ref)
        err = np.abs(got - ref) / np.maximum(np.abs(ref), 1e-6)
        assert (err[np.broadcast_to(ok, err.shape)] > 1e-5).mean() <= 2e-4
    a = vp._run(job.ray_color, job.scene, A.RTW_MODE_RANDOM, cuda)
    b = vp._run(job.ray_color, back, A.RTW_MODE_RANDOM, cuda)
    assert np.array_equal(a, b, equal_nan=True) or np.allclose(a, b, rtol=1e-5, atol=1e-6, equal_nan=True)
