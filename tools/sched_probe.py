"""Runs one COUNT_WORK | WAVEFRONT render of a config on cuda:0 and prints the warp-scheduling
counters of wf_extend (how many lanes each step kind ran with). Tuning aid; needs a GPU."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import rtw_b200  # noqa: E402
from rtw_b200 import _abi as A, scenes  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--config", default="C5")
ap.add_argument("--spp", type=int, default=32)
a = ap.parse_args()
job = scenes.CONFIGS[a.config](samples=a.spp)
vp = job.viewport
vp.flags = A.RTW_FLAG_WAVEFRONT | A.RTW_FLAG_COUNT_WORK
be = rtw_b200.get_default_backend()
img = vp._run(job.ray_color, job.scene, job.mode, be)
h = be.scene(job.scene)
st = h.stats()
d = h.sched_stats()
d["rays"] = int(st.rays)
d["per_ray"] = {"nodes": st.nodes / st.rays, "spheres": st.sphere_tests / st.rays, "quads": st.quad_tests / st.rays,
                "instances": st.instance_tests / st.rays, "iterations_x32": 32.0 * d["iterations"] / st.rays}
print(json.dumps(d))
