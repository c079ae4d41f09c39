"""raytracing-in-a-weekend_b200 — B200-native path-tracing backend behind the scene/render API of
Terence-23/RayTracing-in-a-weekend's Rust tree (include/rtw.h).

The directory name carries a hyphen, so import it through the `rtw_b200` shim at the repo root:
    import rtw_b200 as rtw
"""
from ._abi import (RTW_MODE_NO_RAND, RTW_MODE_RANDOM, RTW_MODE_STRATIFIED, RTW_OUT_RESOLVED, RTW_OUT_SUM,
                   RtwError, load_library)
from .scene import (EMPTY_M, FUZZY3_M, GLASS_M, GLASSR_M, METALLIC_M, SCATTER_M, ImageTexture, Instance,
                    Material, PerlinNoise, Quad, Scene, Sphere, Triangle, Vec3, const_density, surface)
from .viewport import (Viewport, async_render, get_default_backend, light_biased_ray_cast, light_biased_ray_color,
                       ray_color_albedo, ray_color_bg_color,
                       ray_color_d, ray_color_glass_test, ray_color_gradient, ray_color_normal, render_multi,
                       set_default_backend)

from .write_img import quantize, write_img, write_img_f32
from .postprocessing import Proximity, ProximityType, bilateral_filter
from . import driver, scene_json
from .driver import dump_frames
from .scene_json import ParseError

__all__ = [n for n in dir() if not n.startswith("_")]
