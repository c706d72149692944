#!/usr/bin/env python
"""Small driver for ncu captures: one crt_render of a built-in scene through the C-ABI.
usage: prof_render.py [--scene S] [--mode wavefront|megakernel] [--width W] [--height H] [--spp N] [--bounces B] [--reps R]"""
import argparse, importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ap = argparse.ArgumentParser()
ap.add_argument("--scene", default="cornell_blob")
ap.add_argument("--param", type=int, default=0)
ap.add_argument("--mode", default="wavefront")
ap.add_argument("--width", type=int, default=2560)
ap.add_argument("--height", type=int, default=1440)
ap.add_argument("--spp", type=int, default=4)
ap.add_argument("--bounces", type=int, default=20)
ap.add_argument("--reps", type=int, default=1)
ap.add_argument("--counted", action="store_true")
ap.add_argument("--time-kernels", action="store_true")
ap.add_argument("--leaf-max", type=int, default=0)
a = ap.parse_args()
crt = importlib.import_module("raytracer-cuda_b200")
sc = crt.builtin_scene(a.scene, a.param)
r = crt.Renderer(a.width, a.height, mode=crt.MODE_WAVEFRONT if a.mode == "wavefront" else crt.MODE_MEGAKERNEL, bvh_leaf_max=a.leaf_max)
r.set_scene(sc)
r.update_camera(sc.camera)
r.initialize(crt.REFERENCE_SEED, 0)
for i in range(a.reps):
    st = r.render(a.spp, a.bounces, counted=a.counted, flags=crt.RENDER_TIME_KERNELS if a.time_kernels else 0)
    print("rep %d: %.3f ms, %d rays, %.1f Mrays/s, %d launches, nodes/ray %.2f prims/ray %.2f" % (
        i, st.render_ms, st.rays, st.rays / st.render_ms / 1e3, st.kernel_launches,
        st.nodes_fetched / max(st.rays, 1), st.prims_tested / max(st.rays, 1)))
    if a.time_kernels:
        print("   extend %.2f ms / %d launches, shade %.2f ms, tail %.2f ms" % (st.extend_ms, st.extend_launches, st.shade_ms, getattr(st, "tail_ms", 0.0)))
bs = r.stats()
print("bvh: %d nodes depth %d build %.2f ms; rng init %.2f ms" % (bs.bvh_nodes, bs.bvh_depth, bs.bvh_build_ms, bs.rng_init_ms))
r.close()
