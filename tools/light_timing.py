"""Device time of Rust2's light-biased integrators (megakernel, shadow rays) on Rust2's own test scene at 800x600.
    [RTW_B200_LIB=<variant .so>] python tools/light_timing.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rtw_b200
from rtw_b200 import scenes

be = rtw_b200.get_default_backend()
for cast, spp in ((True, 64), (False, 64)):
    job = scenes.light_biased_test(cast=cast, width=800, height=600, samples=spp)
    best = None
    for _ in range(4):
        img = job.viewport.render(job.ray_color, job.scene, backend=be)
        st = be.scene(job.scene).stats()
        best = st.kernel_ms if best is None else min(best, st.kernel_ms)
    print(f"{job.name:13s} 800x600 {spp} spp: {best:8.3f} ms  {st.paths / best / 1e3:8.1f} Mpaths/s  rays/path {st.rays / st.paths:.2f}  mean {float(img.mean()):.5f}")
