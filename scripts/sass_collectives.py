#!/usr/bin/env python
"""Lists the warp collectives (VOTE / SHFL / REDUX / MATCH) of the library's own kernels that have neither a BRA.DIV
divergence check, a WARPSYNC nor another collective within the 14 instructions in front of them. Straight-line code at
the top of a kernel is fine; one at a loop's back edge after a divergent region is what stalled k_wf_tail.
usage: python scripts/sass_collectives.py [object files]   (default: raytracer-cuda_b200/build/*.o)"""
import glob, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
objs = sys.argv[1:] or sorted(glob.glob(os.path.join(ROOT, "raytracer-cuda_b200", "build", "crt_*.o")))
for o in objs:
    out = subprocess.run(["cuobjdump", "-sass", o], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    fn, buf = None, []
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            fn, buf = m.group(1), []
            continue
        m = re.search(r"/\*([0-9a-f]{4,5})\*/\s+(.*?);", line)
        if not m or fn is None or "cub" in fn:
            continue
        ins = m.group(2).strip()
        buf.append(ins)
        if re.match(r"(@!?U?P\d+ )?(VOTE\.|SHFL|REDUX|MATCH)", ins):
            prev = buf[-15:-1]
            if not any("BRA.DIV" in p or "WARPSYNC" in p or re.match(r"(VOTE\.|SHFL|REDUX)", p) for p in prev):
                print("%-70s %s  %s" % (fn[:70], m.group(1), ins[:60]))
