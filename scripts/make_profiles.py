#!/usr/bin/env python
"""Turn the ncu captures of scripts/profile_round.sh (gpurun_out/<tag>_*) into the tracked summaries under profiles/.
usage: python scripts/make_profiles.py <tag> <out-prefix>     e.g.  r1g r1"""
import collections, csv, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag, out = sys.argv[1], sys.argv[2]
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
os.makedirs(P, exist_ok=True)

# 1. launch list -> per-kernel totals and shares
rows = [r for r in csv.reader(open(os.path.join(G, tag + "_launches.csv"))) if len(r) > 10]
hdr = rows[0]; ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
d = collections.defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    try: v = float(r[vi].replace(",", ""))
    except ValueError: continue
    v = v / 1e3 if r[ui] == "ns" else v * 1e3 if r[ui] == "ms" else v
    k = r[ki].split("(")[0][:70]; d[k][0] += 1; d[k][1] += v
tot = sum(v[1] for v in d.values())
with open(os.path.join(P, out + "_launch_list.txt"), "w") as f:
    f.write("# ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv  python bench.py --spp 32 --steps 2 --warmup 3 --no-cpu-baseline\n")
    f.write("# per-launch times are cold-cache and serialised: compare SHARES with bench.py's roofline.kernel_share_of_step, not absolutes\n")
    f.write("# plain (un-profiled) run of the same command: " + open(os.path.join(G, tag + "_bench_spp32.json")).read().strip()[:400] + " ...\n")
    for k, v in sorted(d.items(), key=lambda x: -x[1][1]):
        f.write("%-72s n=%5d total_us=%11.1f share=%.4f avg_us=%.1f\n" % (k, v[0], v[1], v[1] / tot, v[1] / v[0]))

# 2. full captures -> key metrics
want = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed.avg.per_cycle_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "l1tex__throughput.avg.pct_of_peak_sustained_active", "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio", "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio"]
traffic = {}
with open(os.path.join(P, out + "_ncu_kernels.txt"), "w") as f:
    f.write("# ncu --set full --clock-control none --import-source on, one launch each (steady state: launch #100 of a 64-spp frame, one part = half of the 3.69 M pixels per launch)\n")
    for name, kern in (("extend", "k_wf_extend"), ("shade", "k_wf_shade"), ("tail", "k_wf_tail"), ("mega", "k_render_mega")):
        rep = os.path.join(G, "%s_%s.ncu-rep" % (tag, name))
        if not os.path.exists(rep): continue
        txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rr = list(csv.reader(txt.splitlines())); h, u = rr[0], rr[1]
        for r in rr[2:]:
            f.write("== %s: %s\n" % (name, r[h.index("Kernel Name")][:90]))
            for w in want:
                if w in h: f.write("  %-86s %s %s\n" % (w, r[h.index(w)], u[h.index(w)]))
            def val(m):
                x, un = float(r[h.index(m)].replace(",", "")), u[h.index(m)]
                return x * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(un, 1)
            traffic[kern] = {"dram_bytes_per_launch": val("dram__bytes_read.sum") + val("dram__bytes_write.sum"),
                             "source": "profiles/%s_ncu_kernels.txt (%s, one ncu --set full capture)" % (out, name)}
json.dump(traffic, open(os.path.join(P, "traffic.json"), "w"), indent=1)
print(open(os.path.join(P, out + "_launch_list.txt")).read()[:1500])
print(json.dumps(traffic))
