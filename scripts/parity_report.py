#!/usr/bin/env python
"""Parity numbers for DESIGN.md / profiles: CUDA path (through the C-ABI) against the oracle on every BASELINE config,
at the config's own size on crops. Prints one JSON line per config: fraction of bit-exact pixels, of pixels within 1e-4,
primary-id equality, ray-count difference. TEST TOOLING (uses oracle/): run under gpurun, output kept in profiles/."""
import importlib, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
crt = importlib.import_module("raytracer-cuda_b200")
import oracle_lib as ol
SEED = crt.REFERENCE_SEED
CONFIGS = [("C1", "cornell", 0, 256, 256, 16, 8, [(0, 0, 256, 256)]),
           ("C2", "cornell_blob", 0, 2560, 1440, 4, 20, [(1100, 600, 256, 128), (0, 0, 128, 64), (2432, 1376, 128, 64), (700, 300, 128, 128)]),
           ("C3", "spheres", 0, 1200, 800, 8, 20, [(540, 360, 256, 128), (0, 0, 128, 64), (300, 420, 128, 64), (800, 300, 128, 64)]),
           ("C4", "terrain", 0, 2560, 1440, 4, 20, [(1408, 736, 128, 64), (128, 104, 128, 64), (1200, 200, 128, 64), (2432, 0, 128, 64)]),
           ("C5", "cornell_blob", 0, 1920, 1080, 1, 8, [(900, 500, 256, 128), (0, 0, 128, 64), (600, 250, 128, 128)])]
for cfg, scene, param, W, H, spp, bounces, crops in CONFIGS:
    sc = crt.builtin_scene(scene, param)
    osc = ol.OracleScene(sc.pointer())
    for mode, mname in ((crt.MODE_WAVEFRONT, "wavefront"), (crt.MODE_MEGAKERNEL, "megakernel")):
        r = crt.Renderer(W, H, mode=mode)
        r.set_scene(sc); r.update_camera(sc.camera); r.initialize(SEED, 0)
        ids, t = r.primary_ids(SEED, 0)
        r.render(spp, bounces)
        acc = r.accum()
        npx = exact = close = ids_eq = t_eq = 0
        worst = 0.0
        for (x0, y0, w, h) in crops:
            oid, ot = osc.primary(sc.camera, W, H, SEED, window=(x0, y0, w, h), threads=16)
            orad, _, _ = osc.render(sc.camera, W, H, spp, bounces, SEED, window=(x0, y0, w, h), threads=16)
            rad = acc[y0:y0 + h, x0:x0 + w, :3] * np.float32(1.0 / spp)
            rel = np.abs(rad - orad) / np.maximum(np.maximum(np.abs(rad), np.abs(orad)), 1e-3)
            npx += w * h
            exact += int((rad.view(np.uint32) == orad.view(np.uint32)).all(axis=-1).sum())
            close += int((rel <= 1e-4).all(axis=-1).sum())
            ids_eq += int((ids[y0:y0 + h, x0:x0 + w] == oid).sum())
            t_eq += int((t[y0:y0 + h, x0:x0 + w].view(np.uint32) == ot.view(np.uint32)).sum())
            worst = max(worst, float(rel.max()))
        print(json.dumps({"config": cfg, "scene": scene, "size": [W, H], "spp": spp, "bounces": bounces, "mode": mname, "pixels_compared": npx,
                          "primary_ids_equal": ids_eq / npx, "primary_t_bits_equal": t_eq / npx, "radiance_bit_exact": exact / npx,
                          "radiance_within_1e-4": close / npx, "worst_rel_err": worst,
                          "note": "pixels that are not bit-exact: a path met a Dielectric whose Schlick term (double pow) differs in the last ulp between CUDA and glibc"}), flush=True)
        r.close()
    osc.close(); sc.free()
