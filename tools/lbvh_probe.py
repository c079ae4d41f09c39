"""LBVH vs host SAH on one config (GPU box): build times, render throughput, node visits per ray.
    python tools/lbvh_probe.py [C5] [spp]"""
import ctypes as C
import sys
import time

sys.path.insert(0, ".")
import rtw_b200
from rtw_b200 import _abi as A, scenes

name = sys.argv[1] if len(sys.argv) > 1 else "C5"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 200
be = rtw_b200.get_default_backend()
job = scenes.CONFIGS[name](samples=spp)
vp = job.viewport
fs = job.scene.flatten()
t0 = time.perf_counter()
hd = C.c_void_p()
A.check(be.lib.rtw_scene_create(C.byref(fs.desc), be.device, C.byref(hd)))
create_ms = (time.perf_counter() - t0) * 1e3
be.lib.rtw_scene_destroy(hd)
h = be.scene(job.scene)


def run(tag):
    out = {}
    for flags, key in ((0, "fast"), (A.RTW_FLAG_COUNT_WORK, "count")):
        vp.flags = flags
        vp._run(job.ray_color, job.scene, job.mode, be)
        vp._run(job.ray_color, job.scene, job.mode, be)
        st = h.stats()
        out[key] = st
    f, c = out["fast"], out["count"]
    print(f"{tag:5s} {f.paths / f.kernel_ms / 1e3:8.1f} Mpaths/s  {f.rays / f.kernel_ms / 1e3:8.1f} Mrays/s  "
          f"nodes/ray {c.nodes / c.rays:5.2f}  sphere {c.sphere_tests / c.rays:4.2f}  quad {c.quad_tests / c.rays:4.2f}  inst {c.instance_tests / c.rays:4.2f}")


print(f"{name}: rtw_scene_create (flatten + host SAH build + upload) {create_ms:.1f} ms")
run("SAH")
ms = [h.rebuild_lbvh() for _ in range(5)]
print(f"rtw_scene_rebuild_lbvh device ms: first {ms[0]:.3f}, then {min(ms[1:]):.3f}")
t0 = time.perf_counter(); h.rebuild_lbvh(); print(f"rebuild wall {1e3 * (time.perf_counter() - t0):.3f} ms")
run("LBVH")
h.use_sah()
run("SAH")
