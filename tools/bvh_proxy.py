"""BVH walk quality without a GPU: the kernels' own traversal headers compiled for the host (tests/emu, the
strict walk over the exact f32 nodes) count node visits and primitive tests per ray, for the host SAH tree
and for the LBVH rebuild. A proxy for builder experiments between GPU calls — the device numbers are the
RTW_FLAG_COUNT_WORK build's (bench.py `per_ray`; C5: 7.7 nodes, 2.3 sphere, 0.62 quad, 0.50 instance tests per ray
with the fast walk, which also drops the enclosing fog).
    python tools/bvh_proxy.py [C5] [width height spp]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import rtw_b200  # noqa: E402
from rtw_b200 import scenes  # noqa: E402
from emu_backend import EmuBackend  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "C5"
w, h, spp = (int(x) for x in sys.argv[2:5]) if len(sys.argv) > 4 else (96, 96, 2)
make = scenes.CONFIGS.get(name) or scenes.REFERENCE_TESTS[name]
for tag, be in (("SAH", EmuBackend()), ("LBVH", EmuBackend(lbvh=True))):
    job = make()
    vp = job.viewport
    nv = type(vp).new_from_res(w, h, spp, vp.depth, 1.0, vp.vfov, vp.origin, vp.direction, vp.vup, None, vp.lens_radius)
    nv.shutter_speed, nv.fps, nv.frame = vp.shutter_speed, vp.fps, vp.frame
    nv._run(job.ray_color, job.scene, job.mode, be)
    s, inf = be.scene(job.scene).stats(), be.scene(job.scene).info()
    r = float(s["rays"])
    print(f"{name} {tag:4s} rays/path {r / s['paths']:.2f}  per ray: nodes {s['nodes'] / r:.2f}  sphere {s['sphere_tests'] / r:.2f}  "
          f"quad {s['quad_tests'] / r:.2f}  instance {s['instance_tests'] / r:.2f}  hits {s['hits'] / r:.2f}   "
          f"(tlas nodes {inf.n_tlas_nodes}, blas nodes {inf.n_blas_nodes}, depth {inf.bvh_max_depth})")
