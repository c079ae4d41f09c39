import sys, time, os
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
mode = sys.argv[1]
if mode in ("torch", "cuda", "render"):
    import torch
if mode in ("cuda", "render"):
    torch.cuda.init(); torch.zeros(1, device="cuda")
import bench
from rtw_b200 import scenes
job = scenes.config_c5()
if mode == "render":
    import rtw_b200
    be = rtw_b200.get_default_backend()
    j2 = scenes.config_c5(samples=16)
    j2.viewport._run(j2.ray_color, j2.scene, j2.mode, be)
s = bench.cpu_sample(job, 6.0)
print(mode, round(s['paths'] / s['seconds'] / 1e6, 3), s['spp'], s['threads'], "affinity", len(os.sched_getaffinity(0)), "OMP", os.environ.get("OMP_NUM_THREADS"))
