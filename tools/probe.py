"""Quick device probe: render each BASELINE config at reduced spp through the C ABI, print rates."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import rtw_b200
from rtw_b200 import scenes, _abi as A

be = rtw_b200.get_default_backend()
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 64
names = sys.argv[2].split(",") if len(sys.argv) > 2 else ["C1", "C2", "C3", "C4", "C5"]
flags = int(sys.argv[3]) if len(sys.argv) > 3 else 0
for name in names:
    job = scenes.CONFIGS[name](samples=spp)
    h = be.scene(job.scene)
    cam = job.viewport.camera(be)
    o = job.viewport.opts(job.ray_color, job.mode)
    o.flags = flags
    for rep in range(3):
        t = time.time(); img = h.render(cam, o); dt = time.time() - t
        st = h.stats()
    inf = h.info()
    fin = np.isfinite(img)
    print(f"{name} {cam.width}x{cam.height} spp={spp} flags={flags}: kernel {st.kernel_ms:.2f} ms total {st.total_ms:.2f} ms wall {dt*1e3:.2f} ms | "
          f"{st.paths/st.kernel_ms/1e3:.1f} Mpaths/s {st.rays/st.kernel_ms/1e3:.1f} Mrays/s rays/path {st.rays/st.paths:.2f} "
          f"nan_paths {st.nan_paths} mean {img[fin].mean():.4f} nonfinite_px {(~fin).any(-1).sum()} | tlas {inf.n_tlas_nodes} blas {inf.n_blas_nodes} depth {inf.bvh_max_depth}", flush=True)
