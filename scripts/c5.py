#!/usr/bin/env python
"""BASELINE config C5: progressive 1920x1080, 1 spp per frame, 8 bounces, 200 frames after 20 warm-up; median / p99 frame ms."""
import importlib, json, os, statistics, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
crt = importlib.import_module("raytracer-cuda_b200")
sc = crt.builtin_scene("cornell_blob", 0)
out = {}
for mode, name in ((crt.MODE_WAVEFRONT, "wavefront"), (crt.MODE_MEGAKERNEL, "megakernel")):
    r = crt.Renderer(1920, 1080, mode=mode)
    r.set_scene(sc); r.update_camera(sc.camera); r.initialize(crt.REFERENCE_SEED, 0)
    ms, rays, st = [], 0, None
    for f in range(220):
        st = r.render(1, 8)
        if f >= 20: ms.append(st.render_ms); rays += int(st.rays)
    ms.sort()
    out[name] = {"median_ms": statistics.median(ms), "p99_ms": ms[int(0.99 * len(ms))], "min_ms": ms[0], "Mrays_s": rays / sum(ms) / 1e3,
                 "launches_per_frame": int(st.kernel_launches), "graph_launches": int(st.graph_launches), "host_syncs": int(st.host_syncs), "iterations": int(st.iterations)}
    r.close()
print(json.dumps({"config": "C5 1920x1080 1 spp 8 bounces, 200 frames", "env": {k: v for k, v in os.environ.items() if k.startswith("CRT_")}, **out}))
