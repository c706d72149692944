#!/usr/bin/env python
"""Turn the ncu captures of scripts/profile_round.sh (gpurun_out/<tag>_*) into the tracked summaries under profiles/:
  <out>_launch_list.txt      per-kernel totals and shares of the bench command's launch list
  <out>_ncu_kernels.txt      key metrics of one ncu --set full capture per kernel, local-memory counters included
  <out>_ncu_kernels_c4.txt   the same on the 5 M-triangle terrain
  <out>_extend_by_line.txt   instruction issue and stall samples of k_wf_extend by source line (and of k_wf_shade)
  <out>_extend_node_visit.sass  SASS of the node-visit loop body of k_wf_extend (8 quantised child boxes)
  traffic.json / traffic_terrain.json   DRAM bytes and duration per launch, read by bench.py's roofline object
usage: python scripts/make_profiles.py <tag> <out-prefix>     e.g.  r2 r2"""
import collections, csv, json, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag, out = sys.argv[1], sys.argv[2]
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
os.makedirs(P, exist_ok=True)

# 1. launch list -> per-kernel totals and shares
lp = os.path.join(G, tag + "_launches.csv")
if os.path.exists(lp):
    rows = [r for r in csv.reader(open(lp)) if len(r) > 10]
    hdr = rows[0]; ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    d = collections.defaultdict(lambda: [0, 0.0])
    for r in rows[1:]:
        try: v = float(r[vi].replace(",", ""))
        except ValueError: continue
        v = v / 1e3 if r[ui] == "ns" else v * 1e3 if r[ui] == "ms" else v
        k = r[ki].split("(")[0][:70]; d[k][0] += 1; d[k][1] += v
    tot = sum(v[1] for v in d.values())
    with open(os.path.join(P, out + "_launch_list.txt"), "w") as f:
        f.write("# CRT_WF_GRAPH=0 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv  python bench.py --spp 32 --steps 2 --warmup 3 --no-cpu-baseline\n")
        f.write("# (graph replay off under the profiler so that every kernel is a launch of its own; the plain run below uses the default)\n")
        f.write("# per-launch times are cold-cache and serialised: compare SHARES with bench.py's roofline.kernel_share_of_step, not absolutes\n")
        f.write("# plain (un-profiled) run of the same command: " + open(os.path.join(G, tag + "_bench_spp32.json")).read().strip()[:400] + " ...\n")
        for k, v in sorted(d.items(), key=lambda x: -x[1][1]):
            f.write("%-72s n=%5d total_us=%11.1f share=%.4f avg_us=%.1f\n" % (k, v[0], v[1], v[1] / tot, v[1] / v[0]))

# 2. full captures -> key metrics
want = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__shared_mem_per_block_static",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed.avg.per_cycle_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "l1tex__throughput.avg.pct_of_peak_sustained_active", "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "lts__t_sectors.sum",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum",
        "smsp__inst_executed_op_local_ld.sum", "smsp__inst_executed_op_local_st.sum", "smsp__inst_executed_op_shared_ld.sum", "smsp__inst_executed_op_shared_st.sum",
        "smsp__inst_executed_op_global_ld.sum", "smsp__inst_executed_op_global_st.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio", "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio"]
KERN = {"extend": "k_wf_extend", "shade": "k_wf_shade", "tail": "k_wf_tail", "begin": "k_wf_begin", "compact_scatter": "k_compact_scatter",
        "rng_init": "k_rng_init", "mega": "k_render_mega", "shade_bulk": "k_wf_shade_bulk"}


def summarise(prefix, names, out_file, header, traffic_file):
    traffic = {}
    wrote = False
    with open(out_file, "w") as f:
        f.write(header)
        for name in names:
            rep = os.path.join(G, "%s_%s.ncu-rep" % (prefix, name))
            if not os.path.exists(rep): continue
            txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
            rr = list(csv.reader(txt.splitlines()))
            if len(rr) < 3: continue
            h, u = rr[0], rr[1]
            for r in rr[2:]:
                wrote = True
                f.write("== %s: %s\n" % (name, r[h.index("Kernel Name")][:90]))
                for w in want:
                    if w in h: f.write("  %-86s %s %s\n" % (w, r[h.index(w)], u[h.index(w)]))
                def val(m):
                    x, un = float(r[h.index(m)].replace(",", "")), u[h.index(m)]
                    return x * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "us": 1, "ms": 1e3, "ns": 1e-3}.get(un, 1)
                def num(m):
                    try: return float(r[h.index(m)].replace(",", ""))
                    except (ValueError, IndexError): return None
                traffic[KERN.get(name, name)] = {"dram_bytes_per_launch": val("dram__bytes_read.sum") + val("dram__bytes_write.sum"),
                                                 "gpu_time_us": val("gpu__time_duration.sum"),
                                                 # the counters north_star names, for bench.py's roofline.kernels[].ncu_counters
                                                 "counters": {"l1_hit_pct": num("l1tex__t_sector_hit_rate.pct"), "l2_hit_pct": num("lts__t_sector_hit_rate.pct"),
                                                              "l2_throughput_pct_of_peak": num("lts__throughput.avg.pct_of_peak_sustained_elapsed"),
                                                              "dram_throughput_pct_of_peak": num("dram__throughput.avg.pct_of_peak_sustained_elapsed"),
                                                              "fma_pipe_pct": num("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"),
                                                              "alu_pipe_pct": num("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
                                                              "issue_slots_busy_pct": num("smsp__issue_active.avg.pct_of_peak_sustained_active"),
                                                              "lanes_per_instruction": num("smsp__thread_inst_executed_per_inst_executed.ratio"),
                                                              "warps_active_pct": num("sm__warps_active.avg.pct_of_peak_sustained_active"),
                                                              "warp_instructions": num("smsp__inst_executed.sum"),
                                                              "registers_per_thread": num("launch__registers_per_thread"),
                                                              "stall_long_scoreboard_per_issue": num("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio")},
                                                 "source": "profiles/%s (%s, one ncu --set full capture)" % (os.path.basename(out_file), name)}
    if wrote: json.dump(traffic, open(traffic_file, "w"), indent=1)
    else: os.remove(out_file)
    return traffic


t1 = summarise(tag, ["extend", "shade", "shade_bulk", "tail", "begin", "compact_scatter", "rng_init", "mega"], os.path.join(P, out + "_ncu_kernels.txt"),
               "# ncu --set full --clock-control none --import-source on, one launch each (steady state: launch #100 of a 64-spp frame of the headline scene,\n"
               "# one part = half of the 3.69 M pixels per launch); l1tex__t_sectors_pipe_lsu_mem_local_* and smsp__inst_executed_op_local_* are the traversal stack's\n"
               "# local-memory traffic (the first 4 stack entries live in shared memory: smsp__inst_executed_op_shared_*)\n", os.path.join(P, "traffic.json"))
t2 = summarise(tag + "_terrain", ["extend", "shade"], os.path.join(P, out + "_ncu_kernels_c4.txt"),
               "# the same captures on the 5 002 296-triangle terrain (BASELINE configs[3]): launch #60 of a 16-spp frame\n", os.path.join(P, "traffic_terrain.json"))

# 3. source-line view + SASS of the node-visit loop
lib = os.path.join(ROOT, "raytracer-cuda_b200", "libcrt_b200.so")
tmp = "/tmp/crt_profiles_cubin"
os.makedirs(tmp, exist_ok=True)
subprocess.run(["cuobjdump", "-xelf", "all", lib], cwd=tmp, capture_output=True)
cubin = os.path.join(tmp, "crt_wavefront.sm_100a.cubin")
if os.path.exists(cubin):
    with open(os.path.join(P, out + "_by_source_line.txt"), "w") as f:
        f.write("# scripts/ncu_by_source_line.py over the source page of the captures above: share of stall samples, of executed warp instructions, and lanes\n"
                "# per instruction, by source line (file:line of raytracer-cuda_b200/csrc at the commit of the capture)\n")
        for name, sym in (("extend", "k_wf_extendILb0ELb0"), ("shade", "k_wf_shadeE"), ("shade_bulk", "k_wf_shade_bulk")):
            rep = os.path.join(G, "%s_%s.ncu-rep" % (tag, name))
            if not os.path.exists(rep): continue
            src = os.path.join(tmp, name + ".csv")
            open(src, "w").write(subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout)
            f.write("== %s\n" % name)
            f.write(subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "ncu_by_source_line.py"), cubin, sym, src, "40"], capture_output=True, text=True).stdout)
    # SASS of k_wf_extend<false,false>: the span between the five node loads (LDG.E.128.CONSTANT) and the hit-mask assembly
    sass = subprocess.run(["cuobjdump", "-sass", "-fun", "_ZN3crt11k_wf_extendILb0ELb0EEEvNS_9SceneViewENS_6WfViewEPjS3_iPyiiP5uint2PKS5_ii", cubin],
                          capture_output=True, text=True).stdout.splitlines()
    ins = [l for l in sass if re.match(r"\s+/\*[0-9a-f]{4}\*/", l)]
    prmt = [i for i, l in enumerate(ins) if " PRMT " in l]
    if prmt:   # the window of 330 instructions that holds the most byte decodes = one node visit
        best = max(prmt, key=lambda i0: sum(1 for i in prmt if i0 <= i < i0 + 330))
        prmt = [i for i in prmt if best <= i < best + 330]
    if prmt:
        lo, hi = max(prmt[0] - 45, 0), min(prmt[-1] + 30, len(ins))
        with open(os.path.join(P, out + "_extend_node_visit.sass"), "w") as f:
            f.write("// cuobjdump -sass of k_wf_extend<false,false> (sm_100a), the node visit: five 16-byte node loads, per-axis set-up, 8 children x\n"
                    "// (6 PRMT byte->float, 6 FFMA, FMNMX3+FMNMX per bound, 1 compare, hit-mask insert). %d instructions between the first and the last PRMT.\n"
                    % (prmt[-1] - prmt[0] + 1))
            ops = collections.Counter(re.sub(r"\..*", "", l.split("*/")[1].split()[0] if not l.split("*/")[1].split()[0].startswith("@") else l.split("*/")[1].split()[1]) for l in ins[lo:hi])
            f.write("// opcode histogram of the excerpt: " + ", ".join("%s %d" % kv for kv in ops.most_common(14)) + "\n")
            f.write("\n".join(ins[lo:hi]) + "\n")
print(json.dumps({"traffic": t1, "traffic_terrain": t2})[:600])
