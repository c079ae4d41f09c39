"""Node visits per ray of the host SAH tree against the LBVH steps of rtw_lbvh.h, counted on the CPU by the test
emulator (tests/emu: the kernels' own traversal headers compiled for the host, instrumented build). No GPU needed.
    python tools/lbvh_quality.py [C5,C2] [width] [spp]"""
import sys

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import rtw_b200  # noqa: F401
from rtw_b200 import _abi as A, scenes
from emu_backend import EmuBackend

names = sys.argv[1].split(",") if len(sys.argv) > 1 else ["C5", "C2"]
width = int(sys.argv[2]) if len(sys.argv) > 2 else 96
spp = int(sys.argv[3]) if len(sys.argv) > 3 else 4
for name in names:
    for baked in (False, True):
        row = []
        for lbvh in (False, True):
            job = scenes.CONFIGS[name](samples=spp)
            vp = job.viewport
            nv = type(vp).new_from_res(width, max(1, width * vp.height // vp.width), spp, vp.depth, vp.gamma, vp.vfov, vp.origin,
                                       vp.direction, vp.vup, None, vp.lens_radius)
            nv.shutter_speed, nv.fps, nv.frame = vp.shutter_speed, vp.fps, vp.frame
            be = EmuBackend(lbvh=lbvh, baked=baked, inside=True)
            nv._run(job.ray_color, job.scene, job.mode, be)
            st = be.scene(job.scene).stats()
            row.append((st["nodes"] / st["rays"], st["sphere_tests"] / st["rays"], st["quad_tests"] / st["rays"], st["rays"]))
        (sn, ss, sq, sr), (ln, ls, lq, lr) = row
        print(f"{name} baked={int(baked)}  SAH nodes/ray {sn:5.2f} sphere {ss:4.2f} quad {sq:4.2f} | LBVH nodes/ray {ln:5.2f} sphere {ls:4.2f} "
              f"quad {lq:4.2f} | ratio {ln / sn:5.3f}  rays {sr} / {lr}", flush=True)
