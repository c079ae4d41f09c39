"""Host-side mirror of the reference's scene-building API.

Same type names, constructor names, argument order and quirks as Rust/src of
Terence-23/RayTracing-in-a-weekend, so scenes written against the reference port line by line:

    Material, presets        Rust/src/objects/materials.rs:15-20, 56-87, 157-212
    PerlinNoise, ImageTexture Rust/src/texture.rs:63-69, 110-149, 196-257
    Sphere                   Rust/src/objects/sphere.rs:13-20, 151-247
    Quad                     Rust/src/objects/quad.rs:8-20, 85-106
    Instance                 Rust/src/objects/instance.rs:28-39, 83-242
    Scene                    Rust/src/viewport.rs:79-135

`Scene.flatten()` turns the object graph into the plain arrays of include/rtw.h (rtw_scene_desc).
Nothing here computes pixels.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Callable, List, Optional, Sequence

import numpy as np

from . import _abi as A
from .philox import PhiloxStream

F = np.float32


def Vec3(x: float, y: float, z: float) -> np.ndarray:
    """Vec3::new (vec3.rs:182-184): three f32."""
    return np.array([x, y, z], dtype=F)


def _v(v) -> np.ndarray:
    a = np.asarray(v, dtype=F).reshape(3)
    return a.copy()


# --------------------------------------------------------------------------- materials.rs
class Material:
    """materials.rs:15-20. `emmited` keeps the reference's spelling."""

    __slots__ = ("metallicness", "opacity", "ir", "emmited")

    def __init__(self, metallicness: float, opacity: float, ir: float, emmited=(0.0, 0.0, 0.0)):
        self.metallicness = float(F(metallicness))
        self.opacity = float(F(opacity))
        self.ir = float(F(ir))
        self.emmited = _v(emmited)

    @staticmethod
    def new(metallicness: float, opacity: float, ir: float) -> "Material":  # materials.rs:57-67
        return Material(metallicness, opacity, ir)

    @staticmethod
    def new_m(metallicness: float) -> "Material":  # materials.rs:68-79
        return Material(metallicness, 0.0, 1.0)

    @staticmethod
    def new_emmiting(metallicness: float, opacity: float, ir: float, emmited) -> "Material":  # :80-87
        return Material(metallicness, opacity, ir, emmited)

    def key(self):
        return (self.metallicness, self.opacity, self.ir, tuple(float(x) for x in self.emmited))


METALLIC_M = Material(1.0, 0.0, 1.0)        # materials.rs:157-166
SCATTER_M = Material(0.0, 0.0, 1.0)         # :168-177
FUZZY3_M = Material(0.7, 0.0, 1.0)          # :179-188
GLASS_M = Material(1.0, 1.0, 1.50)          # :190-199
GLASSR_M = Material(1.0, 1.0, float(F(1.0) / F(1.5)))   # :201-210  ir = 1.0 / GLASS_M.ir
EMPTY_M = SCATTER_M                         # :212


# --------------------------------------------------------------------------- texture.rs
class PerlinNoise:
    """texture.rs:63-69. `new` (texture.rs:110-135): ranvec = unit(U[-1,1]^3) with no ball rejection;
    the three permutations come from create_permute, whose shuffle loop runs over the empty range
    255..0, so they are the identity (texture.rs:137-149). The reference seeds from
    thread_rng; here the table stream is an explicit Philox key."""

    def __init__(self, ranvec: np.ndarray, perm_x, perm_y, perm_z):
        self.ranvec = np.ascontiguousarray(ranvec, dtype=F).reshape(256, 3)
        self.perm_x = np.asarray(perm_x, dtype=np.uint32).reshape(256)
        self.perm_y = np.asarray(perm_y, dtype=np.uint32).reshape(256)
        self.perm_z = np.asarray(perm_z, dtype=np.uint32).reshape(256)

    _auto_key = 7

    @staticmethod
    def new(key: Optional[int] = None) -> "PerlinNoise":
        if key is None:
            key = PerlinNoise._auto_key
            PerlinNoise._auto_key += 1
        rs = PhiloxStream(key)
        r = rs.uniform_f32(256 * 3).reshape(256, 3)
        v = r * F(2.0) + F(-1.0)                     # Vec3::random(-1,1): x*(max-min)+min
        length = np.sqrt(v[:, 0] * v[:, 0] + v[:, 1] * v[:, 1] + v[:, 2] * v[:, 2], dtype=F)
        ranvec = (v / length[:, None]).astype(F)     # .unit()
        ident = np.arange(256, dtype=np.uint32)
        return PerlinNoise(ranvec, ident, ident.copy(), ident.copy())


class ImageTexture:
    """texture.rs:21-27. `row` is the image width and `col` its height (texture.rs:196-204);
    texel (x, y) = img[y*row + x] (texture.rs:265)."""

    def __init__(self, img: np.ndarray, row: int, col: int, noise: Optional[PerlinNoise] = None,
                 noise_scale: float = 1.0):
        self.img = np.ascontiguousarray(img, dtype=F).reshape(-1, 3)
        if self.img.shape[0] != row * col:
            raise ValueError("ImageTexture: img must hold width*height texels")
        self.row, self.col = int(row), int(col)
        self.noise = noise
        self.noise_scale = float(F(noise_scale))

    @staticmethod
    def new(img, width: int, height: int) -> "ImageTexture":  # texture.rs:197-205
        return ImageTexture(img, width, height)

    @staticmethod
    def from_color(color) -> "ImageTexture":  # texture.rs:206-214
        return ImageTexture(_v(color).reshape(1, 3), 1, 1)

    @staticmethod
    def _decode(path: str):
        from PIL import Image  # decode only; into_rgb32f = u8 / 255 (texture.rs:216)
        im = Image.open(path).convert("RGB")
        a = (np.asarray(im, dtype=np.uint8).astype(F) / F(255.0)).astype(F)
        return a.reshape(-1, 3), im.width, im.height

    @staticmethod
    def from_path(path: str) -> "ImageTexture":  # texture.rs:215-226
        img, w, h = ImageTexture._decode(path)
        return ImageTexture(img, w, h)

    @staticmethod
    def new_with_noise(img, width: int, height: int, scale: float, noise: Optional[PerlinNoise] = None):
        return ImageTexture(img, width, height, noise or PerlinNoise.new(), scale)  # texture.rs:228-236

    @staticmethod
    def from_color_noise(color, scale: float, noise: Optional[PerlinNoise] = None):  # texture.rs:237-245
        return ImageTexture(_v(color).reshape(1, 3), 1, 1, noise or PerlinNoise.new(), scale)

    @staticmethod
    def from_path_noise(path: str, scale: float, noise: Optional[PerlinNoise] = None):  # texture.rs:246-257
        img, w, h = ImageTexture._decode(path)
        return ImageTexture(img, w, h, noise or PerlinNoise.new(), scale)


# --------------------------------------------------------------------------- sphere.rs
_ONE = (1.0, 1.0, 1.0)
_ZERO = (0.0, 0.0, 0.0)


class Sphere:
    """sphere.rs:13-20."""

    def __init__(self, origin, radius, col_mod, mat, velocity, texture):
        self.origin = _v(origin)
        self.radius = float(F(radius))
        self.col_mod = _v(col_mod)
        self.mat = mat
        self.velocity = _v(velocity)
        self.texture = texture

    @staticmethod
    def new(origin, r: float, col_mod=None, mat: Optional[Material] = None) -> "Sphere":
        """sphere.rs:151-172. NOTE: Some(c) sets BOTH col_mod = c and a 1x1 texture of
        colour c, so the effective albedo is c*c."""
        c = _ONE if col_mod is None else col_mod
        return Sphere(origin, r, c, mat or EMPTY_M, _ZERO, ImageTexture.from_color(c))

    @staticmethod
    def new_moving(origin, r: float, col_mod, mat, velocity) -> "Sphere":  # sphere.rs:173-196
        c = _ONE if col_mod is None else col_mod
        return Sphere(origin, r, c, mat or EMPTY_M, velocity, ImageTexture.from_color(c))

    @staticmethod
    def new_with_texture(origin, r: float, col_mod, mat, tex: ImageTexture) -> "Sphere":  # :197-221
        return Sphere(origin, r, _ONE if col_mod is None else col_mod, mat or EMPTY_M, _ZERO, tex)

    @staticmethod
    def new_moving_with_texture(origin, r, col_mod, mat, velocity, tex) -> "Sphere":  # :222-246
        return Sphere(origin, r, _ONE if col_mod is None else col_mod, mat or EMPTY_M, velocity, tex)


# --------------------------------------------------------------------------- quad.rs
class Quad:
    """quad.rs:8-20; normal/d/w (quad.rs:93-105) are derived by the backend from origin/u/v."""

    def __init__(self, origin, u, v, mat: Material, velocity, texture: ImageTexture):
        self.origin, self.u, self.v = _v(origin), _v(u), _v(v)
        self.mat = mat
        self.velocity = _v(velocity)   # stored, never read by the reference (quad.rs:13)
        self.texture = texture

    shape = 0   # RTW_SHAPE_PARALLELOGRAM

    @staticmethod
    def new(origin, u, v, mat, velocity, texture) -> "Quad":  # quad.rs:85-106
        return Quad(origin, u, v, mat, velocity, texture)


class Triangle(Quad):
    """Rust2/src/objects/triangle.rs:11-48 in the Rust/src object model: the half of
    the parallelogram (origin, u, v) next to `origin`. Same derived normal/d/w as a quad
    (triangle.rs:36-47), inside test `alfa >= 0, beta >= 0, alfa + beta <= 1` (:107-113), leaf box
    over the three corners padded to 0.005 (:50-89). It goes wherever a Quad goes: `Scene.new`'s and
    `Instance.new_quads`' quad lists. Triangle has no velocity field (triangle.rs:16)."""

    shape = 1   # RTW_SHAPE_TRIANGLE

    def __init__(self, origin, u, v, mat: Material, texture: ImageTexture):
        super().__init__(origin, u, v, mat, (0.0, 0.0, 0.0), texture)

    @staticmethod
    def new(origin, u, v, mat, texture) -> "Triangle":  # triangle.rs:27-48
        return Triangle(origin, u, v, mat, texture)


# --------------------------------------------------------------------------- instance.rs
def surface(_d: float) -> float:
    """instance.rs:21-23: the sentinel 'this is a surface, not a volume'."""
    return -1.0


def const_density(d: float) -> float:
    """instance.rs:24-26: -ln(xi)/d free-flight distance. Evaluated on the device (Philox volume
    sub-stream); this host object only tags the instance as RTW_DENSITY_CONST."""
    raise RuntimeError("const_density is sampled on the device; assign it to Instance.dist_fn")


class Instance:
    """instance.rs:28-39: owns quads + spheres, a translation and the reference's Euler 'rotation'."""

    def __init__(self, spheres: Sequence[Sphere], quads: Sequence[Quad]):
        self.spheres: List[Sphere] = list(spheres)
        self.quads: List[Quad] = list(quads)
        self.translation = Vec3(0.0, 0.0, 0.0)
        self.rotation = Vec3(0.0, 0.0, 0.0)
        self.dist_fn: Callable[[float], float] = surface
        self.density = 0.0

    @staticmethod
    def new_box(a, b, tex: ImageTexture, mat: Material) -> "Instance":
        """instance.rs:83-165: six quads, in the reference's order and winding."""
        a, b = _v(a), _v(b)
        mn, mx = np.minimum(a, b), np.maximum(a, b)
        dx = Vec3(mx[0] - mn[0], 0.0, 0.0)
        dy = Vec3(0.0, mx[1] - mn[1], 0.0)
        dz = Vec3(0.0, 0.0, mx[2] - mn[2])
        z = _ZERO
        return Instance.new_quads([
            Quad.new(Vec3(mn[0], mn[1], mx[2]), dx, dy, mat, z, tex),
            Quad.new(Vec3(mx[0], mn[1], mx[2]), -dz, dy, mat, z, tex),
            Quad.new(Vec3(mx[0], mn[1], mn[2]), -dx, dy, mat, z, tex),
            Quad.new(Vec3(mn[0], mn[1], mn[2]), dz, dy, mat, z, tex),
            Quad.new(Vec3(mn[0], mx[1], mx[2]), dx, -dz, mat, z, tex),
            Quad.new(Vec3(mn[0], mn[1], mn[2]), dx, dz, mat, z, tex),
        ])

    @staticmethod
    def new_sphere(spheres: Sequence[Sphere]) -> "Instance":  # instance.rs:166-186
        return Instance(spheres, [])

    @staticmethod
    def new_quads(quads: Sequence[Quad]) -> "Instance":  # instance.rs:187-207
        return Instance([], quads)

    @staticmethod
    def new(spheres: Sequence[Sphere], quads: Sequence[Quad]) -> "Instance":  # instance.rs:208-228
        return Instance(spheres, quads)

    def rotate(self, rot) -> None:  # instance.rs:230-232
        self.rotation = (self.rotation + _v(rot)).astype(F)

    def getr(self) -> np.ndarray:
        return self.rotation.copy()

    def translate(self, vec) -> None:  # instance.rs:237-239
        self.translation = (self.translation + _v(vec)).astype(F)

    def gett(self) -> np.ndarray:
        return self.translation.copy()


# --------------------------------------------------------------------------- viewport.rs Scene
class FlatScene:
    """rtw_scene_desc plus the numpy/ctypes buffers that keep its pointers alive."""

    def __init__(self):
        self.desc = A.rtw_scene_desc()
        self._keep = []


class Scene:
    """viewport.rs:79-88. BVH construction happens in the backend at rtw_scene_create."""

    def __init__(self, spheres, quads, instances):
        self.spheres: List[Sphere] = list(spheres)
        self.quads: List[Quad] = list(quads)
        self.instances: List[Instance] = list(instances)
        self.background_color = Vec3(0.0, 0.0, 0.0)
        # the `lights` argument of Rust2's light-biased integrators (Rust2/src/viewport/ray_color.rs:59,115): objects that
        # are also top-level spheres / quads of this scene (the same Python objects, as the same Arc in Rust2's tests)
        self.lights = []
        self._backend_cache = {}

    @staticmethod
    def new(spheres, quads, instances) -> "Scene":  # viewport.rs:120-135
        return Scene(spheres, quads, instances)

    @staticmethod
    def new_sphere(spheres) -> "Scene":  # viewport.rs:90-104
        return Scene(spheres, [], [])

    @staticmethod
    def new_quad(quads) -> "Scene":  # viewport.rs:105-119
        return Scene([], quads, [])

    def invalidate(self) -> None:
        """Drop cached device copies after mutating the scene."""
        for h in self._backend_cache.values():
            h.close()
        self._backend_cache = {}

    # ---- flatten to the C ABI
    def flatten(self) -> FlatScene:
        fs = FlatScene()
        mats, mat_ids = [], {}
        texs, tex_ids = [], {}
        noises, noise_ids = [], {}

        def mat_id(m: Material) -> int:
            k = m.key()
            if k not in mat_ids:
                mat_ids[k] = len(mats)
                mats.append(m)
            return mat_ids[k]

        def tex_id(t: ImageTexture) -> int:
            k = id(t)
            if k not in tex_ids:
                # 1x1 un-noised colours are deduplicated by value (every Sphere::new makes one)
                if t.row == 1 and t.col == 1 and t.noise is None:
                    vk = ("c",) + tuple(float(x) for x in t.img[0])
                    if vk in tex_ids:
                        tex_ids[k] = tex_ids[vk]
                        return tex_ids[k]
                    tex_ids[vk] = len(texs)
                tex_ids[k] = len(texs)
                texs.append(t)
            return tex_ids[k]

        def noise_id(n: Optional[PerlinNoise]) -> int:
            if n is None:
                return -1
            if id(n) not in noise_ids:
                noise_ids[id(n)] = len(noises)
                noises.append(n)
            return noise_ids[id(n)]

        all_spheres = list(self.spheres)
        all_quads = list(self.quads)
        inst_rec = []
        for ins in self.instances:
            inst_rec.append((len(all_spheres), len(ins.spheres), len(all_quads), len(ins.quads)))
            all_spheres.extend(ins.spheres)
            all_quads.extend(ins.quads)

        sp = (A.rtw_sphere * max(1, len(all_spheres)))()
        for i, s in enumerate(all_spheres):
            sp[i].origin[:] = s.origin.tolist()
            sp[i].radius = s.radius
            sp[i].col_mod[:] = s.col_mod.tolist()
            sp[i].velocity[:] = s.velocity.tolist()
            sp[i].mat_id = mat_id(s.mat)
            sp[i].tex_id = tex_id(s.texture)
        qd = (A.rtw_quad * max(1, len(all_quads)))()
        for i, q in enumerate(all_quads):
            qd[i].origin[:] = q.origin.tolist()
            qd[i].u[:] = q.u.tolist()
            qd[i].v[:] = q.v.tolist()
            qd[i].velocity[:] = q.velocity.tolist()
            qd[i].mat_id = mat_id(q.mat)
            qd[i].tex_id = tex_id(q.texture)
            qd[i].shape = q.shape
        ins_arr = (A.rtw_instance * max(1, len(self.instances)))()
        for i, (ins, rec) in enumerate(zip(self.instances, inst_rec)):
            ins_arr[i].first_sphere, ins_arr[i].n_spheres, ins_arr[i].first_quad, ins_arr[i].n_quads = rec
            ins_arr[i].translation[:] = ins.translation.tolist()
            ins_arr[i].rotation[:] = ins.rotation.tolist()
            if ins.dist_fn is const_density:
                ins_arr[i].density_kind = A.RTW_DENSITY_CONST
            elif ins.dist_fn is surface:
                ins_arr[i].density_kind = A.RTW_DENSITY_SURFACE
            else:
                raise ValueError("Instance.dist_fn must be `surface` or `const_density` "
                                 "(arbitrary closures cannot cross the ABI; SURVEY §7.4-5)")
            ins_arr[i].density = ins.density
        tx = (A.rtw_texture * max(1, len(texs)))()
        for i, t in enumerate(texs):
            tx[i].width, tx[i].height = t.row, t.col
            tx[i].rgb = t.img.ctypes.data_as(C.POINTER(C.c_float))
            tx[i].noise_id = noise_id(t.noise)
            tx[i].noise_scale = t.noise_scale
            fs._keep.append(t.img)
        ns = (A.rtw_noise * max(1, len(noises)))()
        for i, n in enumerate(noises):
            C.memmove(ns[i].ranvec, n.ranvec.ctypes.data, 256 * 3 * 4)
            C.memmove(ns[i].perm_x, n.perm_x.ctypes.data, 256 * 4)
            C.memmove(ns[i].perm_y, n.perm_y.ctypes.data, 256 * 4)
            C.memmove(ns[i].perm_z, n.perm_z.ctypes.data, 256 * 4)
        mt = (A.rtw_material * max(1, len(mats)))()
        for i, m in enumerate(mats):
            mt[i].metallicness, mt[i].opacity, mt[i].ir = m.metallicness, m.opacity, m.ir
            mt[i].emitted[:] = m.emmited.tolist()

        d = fs.desc
        d.abi_version = A.RTW_ABI_VERSION
        d.n_spheres, d.n_top_spheres = len(all_spheres), len(self.spheres)
        d.n_quads, d.n_top_quads = len(all_quads), len(self.quads)
        d.n_instances = len(self.instances)
        d.n_materials, d.n_textures, d.n_noises = len(mats), len(texs), len(noises)
        d.spheres, d.quads, d.instances = sp, qd, ins_arr
        d.materials, d.textures, d.noises = mt, tx, ns
        d.background[:] = _v(self.background_color).tolist()
        lt = (A.rtw_light * max(1, len(self.lights)))()
        for i, l in enumerate(self.lights):
            where = [k for k, s in enumerate(self.spheres) if s is l]
            kind = A.RTW_LIGHT_SPHERE
            if not where:
                where, kind = [k for k, q in enumerate(self.quads) if q is l], A.RTW_LIGHT_QUAD
            if not where:
                raise ValueError("Scene.lights: every light must be one of the scene's top-level spheres / quads")
            lt[i].kind, lt[i].index = kind, where[0]
        d.n_lights = len(self.lights)
        d.lights = lt
        fs._keep += [sp, qd, ins_arr, mt, tx, ns, lt]
        return fs
