"""Megakernel against the wavefront pipeline as a function of the render's size (GPU box): every config at a
ladder of sample counts, each pipeline forced by its flag, device time of the whole render (rtw_stats.kernel_ms,
CUDA events on the render's stream; best of `reps` after one warm-up render). Feeds the pipeline rule of
rtw_api.cu (choose_pipeline).
    python tools/crossover.py [C1,C2,..] [spp,spp,..] [reps] [mega]      (mega: the megakernel column only, for A/Bs of it)"""
import sys

sys.path.insert(0, ".")
import rtw_b200
from rtw_b200 import _abi as A, scenes

names = sys.argv[1].split(",") if len(sys.argv) > 1 else ["C1", "C2", "C3", "C4", "C5"]
spps = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [1, 4, 16, 64, 256]
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
only_mega = len(sys.argv) > 4 and sys.argv[4] == "mega"
be = rtw_b200.get_default_backend()

print(f"{'config':6s} {'spp':>5s} {'Mpaths':>8s} | {'mega ms':>9s} {'Mpaths/s':>9s} | {'wavefront ms':>12s} {'Mpaths/s':>9s} | "
      f"{'default ms':>10s} {'pipeline':>9s} | wf/mega")
for name in names:
    for spp in spps:
        job = scenes.CONFIGS[name](samples=spp)
        vp = job.viewport
        h = be.scene(job.scene)
        res = {}
        for key, flags in ((("mega", A.RTW_FLAG_MEGAKERNEL),) if only_mega else (("mega", A.RTW_FLAG_MEGAKERNEL), ("wf", A.RTW_FLAG_WAVEFRONT), ("auto", 0))):
            vp.flags = flags
            best = None
            for r in range(reps + 1):
                vp._run(job.ray_color, job.scene, job.mode, be)
                st = h.stats()
                if r > 0 and (best is None or st.kernel_ms < best[0]):
                    best = (st.kernel_ms, st.paths, getattr(st, "iterations", 0))
            res[key] = best
        paths = res["mega"][1]
        if only_mega:
            print(f"{name:6s} {spp:5d} {paths / 1e6:8.2f} | {res['mega'][0]:9.3f} {paths / res['mega'][0] / 1e3:9.1f}", flush=True)
            h.close()
            continue
        mm, wm, am = res["mega"][0], res["wf"][0], res["auto"][0]
        pipe = "wavefront" if res["auto"][2] > 0 else "mega"
        print(f"{name:6s} {spp:5d} {paths / 1e6:8.2f} | {mm:9.3f} {paths / mm / 1e3:9.1f} | {wm:12.3f} {paths / wm / 1e3:9.1f} | "
              f"{am:10.3f} {pipe:>9s} | {mm / wm:6.2f}", flush=True)
        h.close()
