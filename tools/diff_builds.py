"""First-hit and path renders of one config through two builds of the library (GPU box): are the images the same, and
if not, by how much. Each build in its own process (the library is loaded once per process).
    python tools/diff_builds.py render <lib.so> <out.npz>     |     python tools/diff_builds.py compare a.npz b.npz"""
import subprocess
import sys

import numpy as np

if sys.argv[1] == "render":
    import os
    os.environ["RTW_B200_LIB"] = sys.argv[2]
    sys.path.insert(0, ".")
    import rtw_b200
    from rtw_b200 import _abi as A, scenes
    be = rtw_b200.get_default_backend()
    out = {}
    for name in ("C3", "C2", "C5"):
        for rep in (0, 1):
            job = scenes.CONFIGS[name](samples=4)
            vp = job.viewport
            nv = type(vp).new_from_res(200, 150, 4, vp.depth, vp.gamma, vp.vfov, vp.origin, vp.direction, vp.vup, None, vp.lens_radius)
            nv.shutter_speed, nv.fps, nv.frame = vp.shutter_speed, vp.fps, vp.frame
            nv.flags = A.RTW_FLAG_WAVEFRONT
            out[f"{name}_normal_{rep}"] = nv._run(A.RTW_INT_NORMAL, job.scene, A.RTW_MODE_NO_RAND, be)
            out[f"{name}_albedo_{rep}"] = nv._run(A.RTW_INT_ALBEDO, job.scene, A.RTW_MODE_NO_RAND, be)
            out[f"{name}_path_{rep}"] = nv._run(job.ray_color, job.scene, job.mode, be)
    np.savez(sys.argv[3], **out)
else:
    a, b = np.load(sys.argv[2]), np.load(sys.argv[3])
    for k in sorted(a.files):
        if k.endswith("_1"):
            continue
        x0, x1, y0 = a[k], a[k[:-1] + "1"], b[k]
        def d(p, q):
            fin = np.isfinite(p) & np.isfinite(q)
            diff = np.abs(p - q)[fin]
            return f"differ {int((p[fin] != q[fin]).sum()):7d} of {fin.sum()}  max abs {diff.max():.3e}  > 1e-4: {int((diff > 1e-4).sum())}"
        print(f"{k[:-2]:12s} same build twice: {d(x0, x1)}   |   across builds: {d(x0, y0)}")
