#!/usr/bin/env python
"""bench.py — the reference's headline benchmark on B200: Mrays/s (and frame time) of the
Cornell-box + bunny-sized-mesh frame, 2560x1440, 2000 spp, 20 bounces (BASELINE.json configs[1]).

  python bench.py --gpus N --steps K --warmup W          (N>1: launched by torch.distributed.run)
  python bench.py --impl reference ...                   the UNMODIFIED reference CUDA kernel (oracle/_ref)

A step is one full frame. With N GPUs the spp budget is sharded over the ranks (rank g draws from cuRAND
subsequences g*W*H + pixel) and the float4 accumulation buffers are summed onto rank 0 with one NCCL
reduce, then tonemapped there (strong scaling: total work fixed).
One JSON line on stdout (rank 0).
"""
import argparse
import ctypes as C
import importlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
SEED = 1804289383


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--mode", default=os.environ.get("CRT_BENCH_MODE", "wavefront"), choices=["wavefront", "megakernel"])
    ap.add_argument("--scene", default="cornell_blob")
    ap.add_argument("--param", type=int, default=0)
    ap.add_argument("--width", type=int, default=2560)
    ap.add_argument("--height", type=int, default=1440)
    ap.add_argument("--spp", type=int, default=2000)
    ap.add_argument("--bounces", type=int, default=20)
    ap.add_argument("--ref-spp", type=int, default=128, help="bounded sample of the workload for the reference arm")
    ap.add_argument("--cpu-sample", default="512x288x256", help="oracle sample WxHxSPP (centre window of the frame)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag, self.proc = index, [], False, None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                if self.stop_flag:
                    break
                f = [x.strip() for x in line.split(",")]
                if len(f) >= 7:
                    self.samples.append(f)
        except Exception:
            pass

    def finish(self):
        self.stop_flag = True
        if self.proc:
            self.proc.terminate()
        sm = [float(s[0]) for s in self.samples if s[0].replace(".", "").isdigit()]
        mx = [float(s[1]) for s in self.samples if s[1].replace(".", "").isdigit()]
        reasons = set()
        for s in self.samples:
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], s[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(self.samples)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)), "measured"
        except Exception:
            pass
    return {"hbm_gbs": 6650.0}, "fallback"


def which_config(args):
    """The BASELINE.json config this run corresponds to (the default arguments are configs[1], the headline)."""
    key = (args.scene, args.width, args.height, args.spp, args.bounces)
    table = {("cornell", 256, 256, 16, 8): "BASELINE.json configs[0]", ("cornell_blob", 2560, 1440, 2000, 20): "BASELINE.json configs[1]",
             ("spheres", 1200, 800, 500, 20): "BASELINE.json configs[2]", ("terrain", 2560, 1440, 256, 20): "BASELINE.json configs[3]",
             ("cornell_blob", 1920, 1080, 1, 8): "BASELINE.json configs[4]"}
    return table.get(key, "not a BASELINE.json config: a reduced or custom run")


def cpu_baseline(crt, args):
    """The oracle (CPU port of the reference's algorithm) on the host cores, bounded sample."""
    import oracle_lib as ol
    w, h, spp = [int(x) for x in args.cpu_sample.split("x")]
    sc = crt.builtin_scene(args.scene, args.param)
    osc = ol.OracleScene(sc.pointer())
    W, H = args.width, args.height
    w, h = min(w, W), min(h, H)
    x0, y0 = (W - w) // 2, (H - h) // 2
    cores = os.cpu_count() or 1
    t0 = time.perf_counter()
    _, _, rays = osc.render(sc.camera, W, H, spp, args.bounces, SEED, window=(x0, y0, w, h), threads=cores)
    dt = time.perf_counter() - t0
    osc.close()
    sc.free()
    return {"value": rays / dt / 1e6, "unit": "Mrays/s", "cores": cores, "kind": "port",
            "sample": "%s %dx%d centre window of %dx%d, %d spp, %d bounces: %d rays in %.2f s" % (args.scene, w, h, W, H, spp, args.bounces, rays, dt)}


def run_reference(args):
    """Reference arm: the reference's own CUDA render kernel (no CPU renderer exists upstream), rebuilt
    unchanged for sm_100 by oracle/Makefile, same scene arrays / camera / seed, bounded spp."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_render_b%d" % (20 if args.bounces != 8 else 8))
    base = {"impl": "reference", "metric": "Mrays/s", "unit": "Mrays/s", "higher_is_better": True}
    if not os.path.exists(exe):
        print(json.dumps(dict(base, unavailable="oracle/_ref/ref_render_b20 not built (needs /root/reference at build time)")))
        return 0
    cmd = [exe, "--scene", args.scene, "--param", str(args.param), "--width", str(args.width), "--height", str(args.height),
           "--spp", str(args.ref_spp), "--frames", str(args.steps), "--warmup", str(args.warmup), "--seed", str(SEED), "--twin", "--rewind"]
    clk = ClockSampler(0)
    clk.start()
    t0 = time.perf_counter()
    out = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    wall = time.perf_counter() - t0
    clocks = clk.finish()
    if out.returncode != 0:
        print(json.dumps(dict(base, unavailable="reference kernel failed: " + out.stderr.strip()[-300:])))
        return 0
    r = json.loads(out.stdout.strip().split("\n")[-1])
    mrays = r["rays_frame0"] / (r["frame_ms_mean"] * 1e-3) / 1e6
    # frame-time metric scaled to the full spp budget (the kernel loops spp serially: time is linear in spp)
    line = dict(base, value=mrays, n_gpus=1, steps=args.steps, warmup=args.warmup, ms_per_step=r["frame_ms_mean"],
                scaling="strong", vs_baseline=None, dtype="f32", data="synthetic",
                config={"workload": "%s %dx%d %d bounces; reference arm renders a bounded sample of %d spp per step instead of %d "
                                    "(Mrays/s is a rate; working set > L2)" % (args.scene, args.width, args.height, args.bounces, args.ref_spp, args.spp),
                        "scene_triangles": r["triangles"], "stack_bytes": r["stack_bytes"]},
                frame_ms_2000spp_extrapolated=r["frame_ms_mean"] * args.spp / args.ref_spp,
                rays_per_step=r["rays_frame0"], twin_matches_render=r["twin_matches_render"],
                rays_counted_on="every timed frame: the RNG streams are rewound (untimed) before each frame, so each one is frame 0 bit for bit, "
                                "whose rays the untimed twin kernel counted" if r.get("rewind") else "frame 0 only",
                same_config=(args.ref_spp == args.spp),
                cpu_baseline={"value": mrays, "unit": "Mrays/s", "cores": 0, "kind": "reference",
                              "sample": "reference CUDA kernel on the GPU (upstream has no CPU renderer), %d spp" % args.ref_spp},
                e2e={"value": mrays, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                reference_build_ms={"mesh_bvh": r["mesh_bvh_build_ms"], "top_bvh": r["top_bvh_ms"], "rng_init": r["rng_init_ms"]},
                clocks=clocks, wall_s=wall, gpu_launches=args.steps)
    print(json.dumps(line))
    return 0


def crc32_of(arr):
    import zlib
    return zlib.crc32(np.ascontiguousarray(arr).tobytes()) & 0xffffffff


def kernel_rooflines(args, kst, st_counted, build, peaks, l2_gbs, traffic_json, world, W, H):
    """Per-kernel roofline rows (DESIGN.md section 6 states each bytes model). kst = crt_stats of the per-kernel timing pass."""
    hbm = peaks["hbm_gbs"]
    npx = W * H
    rays, samples = float(kst.rays), float(kst.samples)
    cache_resident = args.scene != "terrain"
    n_bar, t_bar = st_counted.nodes_fetched / st_counted.rays, st_counted.prims_tested / st_counted.rays
    b_ray = 80.0 * n_bar + 48.0 * t_bar + 44.0
    rows = []

    def row(name, bound, peak, nbytes, ms, launches, model, extra=None):
        ach = nbytes / (ms * 1e-3) / 1e9 if (nbytes and ms and ms > 0) else None
        r = {"kernel": name, "bound": bound, "bytes_model": model, "algorithmic_bytes": nbytes, "ms": ms, "launches": launches,
             "achieved_gbs": ach, "peak_gbs": peak, "frac": (ach / peak) if (ach and peak) else None}
        tj = traffic_json.get(name)
        if tj and launches:
            # DRAM bytes of ONE launch from the committed ncu --set full capture, scaled by this run's average launch time
            r["ncu_dram_bytes_per_launch"] = tj.get("dram_bytes_per_launch")
            r["ncu_dram_gbs"] = tj.get("dram_bytes_per_launch") / (tj.get("gpu_time_us") * 1e-6) / 1e9 if tj.get("gpu_time_us") else None
            r["ncu_dram_frac_of_hbm"] = (r["ncu_dram_gbs"] / hbm) if r.get("ncu_dram_gbs") else None
            r["ncu_source"] = tj.get("source")
            r["ncu_counters"] = tj.get("counters")   # L1 / L2 hit rates, pipe utilisation, lanes per instruction, occupancy of that capture
        if extra:
            r.update(extra)
        rows.append(r)
        return r

    ext = row("k_wf_extend", "l2" if cache_resident else "hbm", l2_gbs if cache_resident else hbm, b_ray * rays, kst.extend_ms, int(kst.extend_launches),
              "80 B x nodes/ray + 48 B x records/ray + 44 B (32 B ray read, 8 B hit write, 4 B queue word), counted by crt_render_counted",
              {"bytes_per_ray": b_ray, "nodes_per_ray": n_bar, "prims_per_ray": t_bar,
               "limiter": "SIMT divergence + dependent-load latency (profiles/: lanes per instruction, long-scoreboard); the scene is served from L1/L2"
                          if cache_resident else "dependent-load latency on a 306 MB scene (DRAM counters in ncu_dram_*)"})
    row("k_wf_shade", "hbm", hbm, rays * 140.0 + samples * 32.0, kst.shade_ms, int(kst.shade_launches),
        "per queue entry 48 B state + 24 B RNG read, 40 B state + 24 B RNG written, 4 B queue word; per finished path 16 B accumulator read + 16 B written")
    row("k_wf_begin", "hbm", hbm, npx * 116.0, kst.begin_ms, 1,
        "per pixel 24 B RNG read + 24 B written, 48 B state, 16 B accumulator, 4 B queue word written")
    row("k_compact_count+scan+scatter", "hbm (launch-latency at this size)", hbm, None, kst.compact_ms, None,
        "12 B per queue entry per compaction (count reads 4 B, scatter reads 4 B and writes <= 4 B); three launches of a few microseconds each, "
        "every 16 iterations", {"algorithmic_bytes": None})
    row("k_wf_tail", "latency (the frame's serial tail: few pixels, each a chain of dependent path segments)", None, None, kst.tail_ms, None,
        "no byte model: the last <= 150 k pixels per part, one thread per pixel")
    row("k_rng_init_coop", "alu + shuffles (block-cooperative XORWOW skip-ahead: one generic 2^67 x subsequence jump per CTA, 255 warp-level "
        "matrix-vector products for its other pixels)", hbm, npx * 24.0, build.rng_init_ms, 1,
        "24 B written per pixel; outside the frame (once per crt_init_rng); the first call of a process also pays the module load")
    row("bvh build (k_for_n<...> + cub sort/scan)", "launch latency (one short launch per PLOC step)", None, None, build.bvh_build_ms, None,
        "outside the frame (once per crt_set_scene); %d wide nodes" % int(build.bvh_nodes))
    if world > 1:
        rows.append({"kernel": "ncclReduce + k_tonemap", "bound": "nvlink / hbm", "bytes_model": "16 B per pixel per rank over NVLink, then 16 B read + 4 B written per pixel on the root",
                     "algorithmic_bytes": npx * 16.0 * (world - 1) + npx * 20.0})
    tot_ms = kst.extend_ms + kst.shade_ms + kst.begin_ms + kst.compact_ms + kst.tail_ms
    for r in rows:
        if r.get("ms") is not None and r["kernel"] in ("k_wf_extend", "k_wf_shade", "k_wf_begin", "k_compact_count+scan+scatter", "k_wf_tail"):
            r["share_of_frame_kernel_time"] = r["ms"] / tot_ms if tot_ms > 0 else None
    return ext, rows


def issue_slot_roofline(counters, rays, extend_ms, sm_mhz):
    if not (counters and counters.get("warp_instructions") and rays and extend_ms and sm_mhz):
        return None
    per_ray = counters["warp_instructions"] / (2560 * 1440 / 2)
    ach = per_ray * float(rays) / (extend_ms * 1e-3) / 1e9
    peak = 148 * 4 * sm_mhz * 1e6 / 1e9
    return {"warp_instructions_per_ray": per_ray, "achieved_ginst_s": ach, "peak_ginst_s": peak, "frac": ach / peak,
            "lanes_per_instruction": counters.get("lanes_per_instruction")}


def main():
    args = parse()
    # a run that stalls says where: Python stacks of every thread go to stderr every CRT_BENCH_WATCHDOG_S seconds (default 600)
    import faulthandler
    faulthandler.dump_traceback_later(float(os.environ.get("CRT_BENCH_WATCHDOG_S", "600")), repeat=True, file=sys.stderr)
    if args.impl == "reference":
        return run_reference(args)
    # stdout carries exactly ONE line (the JSON): everything libraries print there on their own (NCCL's version banner
    # for one) goes to stderr; the JSON line is written to the saved descriptor at the end
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    import torch
    crt = importlib.import_module("raytracer-cuda_b200")
    mg = importlib.import_module("raytracer-cuda_b200.multigpu")
    crt.load_library()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: libcrt_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    # a stream of our own: the library launches on it, torch's events are recorded on it, and the library's NCCL reduce
    # runs on it too (handle 0, the legacy default stream, cannot be handed to crt_set_stream)
    stream = torch.cuda.Stream(device=local)
    torch.cuda.set_stream(stream)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    W, H = args.width, args.height
    mode = crt.MODE_WAVEFRONT if args.mode == "wavefront" else crt.MODE_MEGAKERNEL
    sc = crt.builtin_scene(args.scene, args.param)
    cam = sc.camera
    r = crt.Renderer(W, H, device=local, mode=mode)
    r.set_scene(sc)
    r.update_camera(cam)
    r.initialize(SEED, mg.subsequence_base(rank, W, H))     # disjoint cuRAND subsequences per rank
    build = r.stats()
    my_spp = mg.shard_spp(args.spp, world, rank)
    flags = crt.RENDER_NO_TONEMAP if world > 1 else 0
    r.set_stream(stream.cuda_stream)
    if world > 1:
        # the library's own communicator (crt_comm_init = ncclCommInitRank): rank 0's id travels over torch.distributed,
        # which is plumbing here (rendezvous, barrier, the max over ranks of the timings); the frame's ONE collective,
        # the sum-reduce of the float4 buffer, is crt_reduce_accum inside the library, on the render's stream
        idt = torch.zeros(crt.COMM_ID_BYTES, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt.copy_(torch.frombuffer(bytearray(crt.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(idt, src=0)
        r.comm_init(bytes(idt.cpu().numpy().tobytes()), world, rank)
    host_img = np.empty((H, W, 4), np.uint8)

    def sync_all():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    def step(time_kernels=False, e2e=False, spp=None):
        if e2e:
            r.update_camera(cam)              # per-frame host->device input, like CUDARenderer::updateCamera
        st = r.render(my_spp if spp is None else spp, args.bounces, flags=flags | (crt.RENDER_TIME_KERNELS if time_kernels else 0))
        if world > 1:
            r.reduce_accum(0)                 # ncclReduce(sum) onto rank 0 + tonemap there, inside the library
        if e2e and rank == 0:
            r.image_rgba8(host_img)           # device->host frame, like WindowManager::drawFrame
        return st

    def timed(n, **kw):
        """n steps bracketed by barrier + synchronize, timed with CUDA events on the launching stream."""
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        acc = {"rays": 0, "launches": 0, "host_syncs": 0, "iterations": 0, "graphs": 0, "last": None}
        sync_all()
        ev0.record(stream)
        for _ in range(n):
            st = step(**kw)
            acc["rays"] += int(st.rays)
            acc["launches"] += int(st.kernel_launches) + (1 if world > 1 and rank == 0 else 0)
            acc["host_syncs"] += int(st.host_syncs)
            acc["iterations"] += int(st.iterations)
            acc["graphs"] += int(st.graph_launches)
            acc["last"] = st
        ev1.record(stream)
        sync_all()
        acc["ms"] = ev0.elapsed_time(ev1)
        return acc

    for _ in range(args.warmup):
        step()
    clk = ClockSampler(local)
    if rank == 0:
        clk.start()
    main_run = timed(args.steps)
    clocks = clk.finish() if rank == 0 else None
    # check: the frame the timed region just produced (rank 0 holds the reduced buffer)
    check = None
    if rank == 0:
        accw = r.accum()[..., 3]
        check = {"accum_w_min": float(accw.min()), "accum_w_max": float(accw.max()), "expected_w": float(args.spp),
                 "rgba8_crc32": crc32_of(r.image_rgba8()), "frame": "last frame of the timed region, rank 0 after the reduce"}
    # per-kernel-class times (events between the launches perturb the pipeline a little: separate, untimed-for-value pass)
    k_run = timed(1, time_kernels=True)
    # e2e: the same frames through the public API with host buffers inside the timed region
    e_run = timed(args.steps, e2e=True)

    tot = torch.tensor([float(main_run["rays"]), float(e_run["rays"]), float(main_run["launches"]), float(main_run["host_syncs"])],
                       dtype=torch.float64, device="cuda")
    mx = torch.tensor([main_run["ms"], e_run["ms"]], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    rays_all, e_rays_all, launches_all, syncs_all = tot[0].item(), tot[1].item(), tot[2].item(), tot[3].item()
    dt, dt_e2e = mx[0].item() * 1e-3, mx[1].item() * 1e-3

    # N > 1: does every rank's shard land in rank 0's buffer, and is a shard what a single GPU renders from that base?
    if world > 1:
        vs = 4                                            # samples per rank of the verification frame
        x0, y0, cw, ch = W // 2 - 64, H // 2 - 32, 128, 64
        r.initialize(SEED, mg.subsequence_base(rank, W, H))
        r.render(vs, args.bounces, flags=crt.RENDER_NO_TONEMAP)
        mine = torch.from_numpy(r.accum()[y0:y0 + ch, x0:x0 + cw].copy()).cuda()
        shards = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(shards, mine)
        r.reduce_accum(0)
        if rank == 0:
            red = r.accum()
            want = np.zeros((ch, cw, 4), np.float32)
            for t in shards:
                want += t.cpu().numpy()
            got = red[y0:y0 + ch, x0:x0 + cw]
            g = world - 1
            r.initialize(SEED, mg.subsequence_base(g, W, H))
            r.render(vs, args.bounces, flags=crt.RENDER_NO_TONEMAP)
            solo = r.accum()[y0:y0 + ch, x0:x0 + cw]
            check["multi_gpu"] = {
                "verification_frame": "%d spp per rank, crop %dx%d at (%d,%d)" % (vs, cw, ch, x0, y0),
                "reduced_w_min": float(red[..., 3].min()), "reduced_w_max": float(red[..., 3].max()), "expected_w": float(vs * world),
                "reduced_crop_equals_sum_of_shards": bool(np.allclose(got, want, rtol=1e-6, atol=0.0)),
                "reduced_crop_max_rel_err": float((np.abs(got - want) / np.maximum(np.abs(want), 1e-6)).max()),
                "shard_checked": g, "shard_equals_single_gpu_render_from_its_base_bitwise": bool(np.array_equal(solo.view(np.uint32), shards[g].cpu().numpy().view(np.uint32))),
                "comm": dict(zip(("nranks", "rank", "nccl_version"), r.comm_info()))}
        dist.barrier()

    if rank == 0:
        # algorithmic bytes per ray of the dominant kernel: counted by an instrumented (untimed) bounded render
        r.initialize(SEED, 0)
        c = r.render(2, args.bounces, flags=flags, counted=True)
        peaks, how = measured_peaks()
        l2 = None   # L2-resident streaming-read bandwidth measured on this pool's B200 by scripts/l2_bw.cu
        try:
            l2j = json.load(open(os.path.join(ROOT, "profiles", "l2_peak.json")))
            l2 = [pt["gbs"] for pt in l2j["points"] if pt["working_set_mb"] == 4][0]
        except Exception:
            pass
        traffic_json = {}
        try:
            traffic_json = json.load(open(os.path.join(ROOT, "profiles", "traffic_terrain.json" if args.scene == "terrain" else "traffic.json")))
        except Exception:
            pass
        kst = k_run["last"]
        if args.mode == "wavefront":
            ext, rows = kernel_rooflines(args, kst, c, build, peaks, l2, traffic_json, world, W, H)
            cache_resident = args.scene != "terrain"
            roof = {"bound": "l2" if cache_resident else "hbm", "kernel": "k_wf_extend", "achieved": ext["achieved_gbs"],
                    "peak": ext["peak_gbs"], "unit": "GB/s", "frac": ext["frac"],
                    "traffic": ext.get("ncu_dram_bytes_per_launch"), "traffic_source": ext.get("ncu_source"),
                    "peak_source": ("profiles/l2_peak.json (scripts/l2_bw.cu: 4 MB working set streamed past L1 by all SMs, measured on this pool's B200); "
                                    "the scene (<= 4.3 MB of nodes + records) is cache resident, SURVEY.md 8(d) names L2 as the bound" if cache_resident
                                    else how + " (MEASURED_PEAKS.json hbm_gbs, copy bandwidth): the 306 MB scene does not fit the 126 MB L2"),
                    "bytes_per_ray": ext["bytes_per_ray"], "nodes_per_ray": ext["nodes_per_ray"], "prims_per_ray": ext["prims_per_ray"],
                    "algorithmic_bytes_per_launch": ext["algorithmic_bytes"] / max(ext["launches"], 1),
                    "kernel_ms_per_launch": ext["ms"] / max(ext["launches"], 1), "kernel_launches_per_step": ext["launches"],
                    "kernel_share_of_step": ext["ms"] / k_run["ms"], "hbm_peak_gbs": peaks["hbm_gbs"], "l2_peak_gbs": l2,
                    "frac_of_hbm_copy_peak": ext["achieved_gbs"] / peaks["hbm_gbs"] if ext["achieved_gbs"] else None,
                    "achieved_dram_gbs_ncu": ext.get("ncu_dram_gbs"), "achieved_dram_frac_of_hbm_ncu": ext.get("ncu_dram_frac_of_hbm"),
                    # the bound that explains the kernel: warp instructions per ray from the committed ncu capture (one
                    # steady-state launch = one part = half the pixels of a 2560x1440 frame) x this run's rays / extend time, against
                    # 148 SMs x 4 schedulers x the SM clock sampled during the timed region
                    "issue_slots": issue_slot_roofline(ext.get("ncu_counters"), kst.rays, ext["ms"], (clocks or {}).get("sm_mhz")),
                    "kernels": rows,
                    "note": "achieved = algorithmic bytes of k_wf_extend / its summed launch time (CUDA events around every launch, separate pass); "
                            "kernels[] carries the same for every kernel of the frame; see DESIGN.md section 6"}
        else:
            b_ray = 80.0 * c.nodes_fetched / c.rays + 48.0 * c.prims_tested / c.rays
            ach = b_ray * k_run["rays"] / (kst.extend_ms * 1e-3) / 1e9 if kst.extend_ms > 0 else None
            roof = {"bound": "l2", "kernel": "k_render_mega", "achieved": ach, "peak": l2, "unit": "GB/s", "frac": (ach / l2) if (ach and l2) else None,
                    "traffic": None, "bytes_per_ray": b_ray}
        cpu = None if args.no_cpu_baseline or world > 1 else cpu_baseline(crt, args)
        value = rays_all / dt / 1e6
        line = {"metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic",
                "config": {"workload": "%s %dx%d %d spp %d bounces (%s); one step = one frame; spp sharded over "
                                       "ranks, one NCCL reduce of the float4 buffer per frame (crt_reduce_accum, inside the library); per-step state %d MB > L2 "
                                       "(126 MB), so no L2 flush between steps" % (args.scene, W, H, args.spp, args.bounces, which_config(args), (24 + 16 + 48 + 8) * W * H // 2 ** 20),
                           "mode": args.mode, "triangles": sc.num_triangles, "spheres": sc.num_spheres, "seed": SEED,
                           "bvh_nodes": int(build.bvh_nodes), "bvh_build_ms": build.bvh_build_ms, "rng_init_ms": build.rng_init_ms},
                "frame_ms": dt / args.steps * 1e3, "rays_per_step": rays_all / args.steps,
                "e2e": {"value": e_rays_all / dt_e2e / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": C.sizeof(crt.Camera),
                        "d2h_bytes_per_step": W * H * 4, "ms_per_step": dt_e2e / args.steps * 1e3},
                "gpu_launches": int(launches_all), "launches_per_frame": launches_all / args.steps / world,
                "host_syncs_per_frame": syncs_all / args.steps / world, "iterations_per_frame": main_run["iterations"] / args.steps,
                "graph_launches_per_frame": main_run["graphs"] / args.steps,
                "check": check, "roofline": roof, "cpu_baseline": cpu, "clocks": clocks}
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    r.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
