"""Generates tests/golden/*.npz. Run in the build container (needs /root/reference for the C++ part):

    python tests/golden/make_golden.py            # everything
    python tests/golden/make_golden.py triangle_mesh   # only the named oracle_*.npz fixtures

  ref_cpp_*.npz   outputs of the REFERENCE'S OWN C++ renderer (oracle/_ref/libref_cpp.so, compiled
                  from /root/reference/C++): Render_no_rand images of sphere scenes in four modes
                  (first-hit normal, first-hit t, the glass-test closure on the metal-control
                  scene of C++/src/tests.cpp:275-294 == Rust glass_tests.rs:57-98, mirror chains).
                  These pin the oracle's camera / sphere / normal / reflect arithmetic.
  rotation_kats.npz  the three assert_eq! of Rust/src/vec3.rs:364-404 as data.
  oracle_*.npz    outputs of the oracle itself (regression fixtures for scenes the C++ tree cannot
                  express: quads, instances, volumes, textures, Perlin, motion).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]

import rtw_b200  # noqa: E402
from rtw_b200 import _abi as A, scenes  # noqa: E402
from oracle_backend import OracleBackend  # noqa: E402
import ref_cpp_backend as ref  # noqa: E402

# (ox, oy, oz, r, cr, cg, cb, metallicness, opacity, ir)
SPHERE_SCENES = {
    # C++/src/tests.cpp:275-294 s_test == Rust viewport/glass_tests.rs:57-98 (metal control)
    "glass10_control": dict(spheres=[[0, 0, -1, 0.5, 1, 1, 1, 1.0, 0, 1], [0, -100.5, -1, 100, 0.8, 0.5, 1.0, 0.0, 0, 1]],
                            width=10, aspect=1.0, vfov=90.0, origin=(0, 0, 0), vup=(0, 1, 0), direction=(0, 0, -1)),
    # Rust viewport.rs:556-619 (four spheres, normal shading)
    "viewport_object": dict(spheres=[[-0.5, 0, -1, 0.5, 1, 1, 1, 0, 0, 1], [0.5, 0, -1, 0.5, 1, 1, 1, 0, 0, 1],
                                     [0, 0, -2, 1, 1, 1, 1, 0, 0, 1], [0, 0, 0, 1, 1, 1, 1, 0, 0, 1]],
                            width=80, aspect=80 / 60, vfov=90.0, origin=(0, 0, 0), vup=(0, 1, 0), direction=(0, 0, -1)),
    # all-mirror variant of camera_tests.rs:56-89 with a tilted unit view direction and vfov 60
    "mirrors": dict(spheres=[[-0.5, 0, -1, 0.5, 0.6, 0.6, 0.6, 1, 0, 1], [0.5, 0, -1, 0.5, 1, 1, 1, 1, 0, 1],
                             [0, 0, -2, 1, 0.5, 1, 0, 1, 0, 1], [0, -100.5, -1, 100, 0.8, 0.8, 0.0, 1, 0, 1]],
                    width=96, aspect=1.5, vfov=60.0, origin=(0, 0, 0), vup=(0, 1, 0), direction=(0.6, 0.0, -0.8)),
}
MODES = {"normal": (ref.MODE_NORMAL, 0.0, 1000.0), "t": (ref.MODE_T, 0.001, 1000.0),
         "glass_test": (ref.MODE_GLASS_TEST, 0.001, 1000.0), "mirror_sky": (ref.MODE_MIRROR_SKY, 0.001, 1000.0)}


def main():
    only = sys.argv[1:]
    if only:
        return oracle_fixtures(only)
    if not ref.available():
        raise SystemExit("oracle/_ref/libref_cpp.so missing: run `make -C oracle ref` where /root/reference exists")
    for name, sc in SPHERE_SCENES.items():
        out = {"spheres": np.asarray(sc["spheres"], dtype=np.float32),
               "camera": np.asarray([sc["width"], sc["aspect"], sc["vfov"], *sc["origin"], *sc["vup"], *sc["direction"]], dtype=np.float64)}
        for mname, (mode, mint, maxt) in MODES.items():
            out[mname] = ref.render_no_rand(sc["spheres"], sc["width"], sc["aspect"], sc["vfov"], sc["origin"], sc["vup"],
                                            sc["direction"], mode, mint, maxt)
        np.savez_compressed(os.path.join(HERE, f"ref_cpp_{name}.npz"), **out)
        print("wrote", f"ref_cpp_{name}.npz")
    # Rust/src/vec3.rs:364-404 rotation_tests: (vec, rot, expected) with expected spelled exactly as the asserts do
    PI = np.float32(np.pi)
    r1, r2 = np.array([PI / np.float32(6), 0, 0], np.float32), np.array([0, 0, PI / np.float32(6)], np.float32)
    kats = np.array([[[1, 0, 0], r1, [np.cos(r1[1]) * np.cos(r1[2]), 0.0, -0.0]],
                     [[1, 0, 0], r2, [np.cos(r2[2]), np.sin(r2[2]), 0.0]],
                     [[1, 0, 1], r2, [np.cos(r2[2]), np.sin(r2[2]), 1.0]]], dtype=np.float32)
    np.savez_compressed(os.path.join(HERE, "rotation_kats.npz"), kats=kats)
    oracle_fixtures(None)


def oracle_fixtures(only):
    # oracle regression fixtures
    orc = OracleBackend()
    reg = {"presentation": scenes.presentation_image, "volume": scenes.volume_test, "C3": scenes.config_c3,
           "C4": scenes.config_c4, "C5": scenes.config_c5, "first_frame": scenes.first_frame,
           "triangle_mesh": scenes.triangle_mesh_test}
    for name, f in reg.items():
        if only and name not in only:
            continue
        job = f()
        vp = job.viewport
        nv = type(vp).new_from_res(48, 36, 4, vp.depth, 1.0, vp.vfov, vp.origin, vp.direction, vp.vup, None, vp.lens_radius)
        nv.shutter_speed, nv.fps, nv.frame = vp.shutter_speed, vp.fps, vp.frame
        out = {}
        for ig_name, ig in (("normal", A.RTW_INT_NORMAL), ("albedo", A.RTW_INT_ALBEDO)):
            out[ig_name] = nv._run(ig, job.scene, A.RTW_MODE_NO_RAND, orc)
            out[ig_name + "_tie"] = orc.scene(job.scene).last_tie.copy()
        out["path4spp"] = nv._run(job.ray_color, job.scene, job.mode, orc)
        np.savez_compressed(os.path.join(HERE, f"oracle_{name}.npz"), **out)
        print("wrote", f"oracle_{name}.npz")


if __name__ == "__main__":
    main()
