#!/usr/bin/env python
"""All five BASELINE.json configs, this library vs the reference kernel on the same GPU (run under gpurun).
Writes one JSON line per (config, implementation) to stdout. Times are device times (CUDA events)."""
import importlib, json, os, statistics, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
crt = importlib.import_module("raytracer-cuda_b200")
SEED = crt.REFERENCE_SEED
quick = "--quick" in sys.argv
ours_only = "--ours-only" in sys.argv   # the reference kernel's numbers of BASELINE.md section 5 are not re-measured

def ours(name, scene, param, W, H, spp, bounces, mode, frames=2, warm=1, aspect=None):
    sc = crt.builtin_scene(scene, param)
    r = crt.Renderer(W, H, mode=crt.MODE_WAVEFRONT if mode == "wavefront" else crt.MODE_MEGAKERNEL)
    t0 = time.perf_counter(); r.set_scene(sc); t_scene = time.perf_counter() - t0
    build1 = r.stats().bvh_build_ms
    r.set_scene(sc); build2 = r.stats().bvh_build_ms          # second build: kernels already loaded
    cam = sc.camera
    if aspect: cam.aspect = aspect
    r.update_camera(cam); r.initialize(SEED, 0)
    ms, rays = [], []
    for f in range(warm + frames):
        st = r.render(spp, bounces)
        if f >= warm: ms.append(st.render_ms); rays.append(int(st.rays))
    out = {"config": name, "impl": "b200-" + mode, "scene": scene, "W": W, "H": H, "spp": spp, "bounces": bounces,
           "frame_ms_median": statistics.median(ms), "frame_ms_p99": sorted(ms)[min(len(ms) - 1, int(0.99 * len(ms)))],
           "frame_ms_min": min(ms), "Mrays_s": sum(rays) / sum(ms) / 1e3, "rays_per_frame": rays[0], "frames": frames,
           "bvh_build_ms_first": build1, "bvh_build_ms": build2, "bvh_nodes": int(r.stats().bvh_nodes), "set_scene_wall_ms": t_scene * 1e3,
           "rng_init_ms": r.stats().rng_init_ms, "triangles": sc.num_triangles, "spheres": sc.num_spheres}
    print(json.dumps(out), flush=True)
    r.close(); sc.free()

def ref(name, scene, param, W, H, spp, bounces, frames=2, warm=1, timeout=900):
    if ours_only: return
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_render_b%d" % (8 if bounces == 8 else 20))
    cmd = [exe, "--scene", scene, "--param", str(param), "--width", str(W), "--height", str(H), "--spp", str(spp),
           "--frames", str(frames), "--warmup", str(warm), "--seed", str(SEED), "--twin"]
    try:
        p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=timeout)
        if p.returncode != 0: raise RuntimeError(p.stderr.strip()[-300:])
        j = json.loads(p.stdout.strip().split("\n")[-1])
        out = {"config": name, "impl": "reference-kernel", "scene": scene, "W": W, "H": H, "spp": spp, "bounces": bounces,
               "frame_ms_median": j["frame_ms_mean"], "frame_ms_min": j["frame_ms_min"], "frame_ms_max": j["frame_ms_max"],
               "Mrays_s": j["rays_frame0"] / j["frame_ms_mean"] / 1e3, "rays_per_frame": j["rays_frame0"], "frames": frames,
               "bvh_build_ms": j["mesh_bvh_build_ms"] + j["top_bvh_ms"], "rng_init_ms": j["rng_init_ms"], "twin_matches_render": j["twin_matches_render"]}
    except subprocess.TimeoutExpired:
        out = {"config": name, "impl": "reference-kernel", "scene": scene, "DNF": "timeout %d s (single-thread GPU BVH build)" % timeout}
    except Exception as e:
        out = {"config": name, "impl": "reference-kernel", "scene": scene, "error": str(e)[:300]}
    print(json.dumps(out), flush=True)

# C1
for m in ("wavefront", "megakernel"): ours("C1", "cornell", 0, 256, 256, 16, 8, m, frames=5, warm=2)
ref("C1", "cornell", 0, 256, 256, 16, 8, frames=5, warm=2)
# C3 (reference arm on a bounded spp)
for m in ("wavefront", "megakernel"): ours("C3", "spheres", 0, 1200, 800, 500 if not quick else 64, 20, m)
ref("C3", "spheres", 0, 1200, 800, 64, 20)
# C5: 1 spp per frame, 200 frames after 20 warm-up
for m in ("wavefront", "megakernel"): ours("C5", "cornell_blob", 0, 1920, 1080, 1, 8, m, frames=200, warm=20)
ref("C5", "cornell_blob", 0, 1920, 1080, 1, 8, frames=200, warm=20)
# C4: 5 M triangles
for m in ("wavefront", "megakernel"): ours("C4", "terrain", 0, 2560, 1440, 256 if not quick else 32, 20, m, frames=1, warm=1)
ref("C4", "terrain", 0, 2560, 1440, 16, 20, frames=1, warm=0, timeout=1500)
