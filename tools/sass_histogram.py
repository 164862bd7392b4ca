#!/usr/bin/env python
"""Static view of the hot kernels in prophet_b200/libprophet_b200.so: per kernel the resource usage (`cuobjdump -res-usage`:
registers, stack, shared memory) and a histogram of the SASS mnemonics that matter for this path (`cuobjdump -sass`).
Usage: python tools/sass_histogram.py [kernel-name-substring ...] > profiles/rNN_sass_summary.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "prophet_b200", "libprophet_b200.so")
HOT = ["mlp_fused2_kernel", "mlpw_forward_kernel", "mlpw_backward_kernel", "mlp32_fused_kernel", "mlp32_umma_kernel", "mlp_kernel", "md_desc_kernel", "md_force_kernel",
       "md_fused_kernel", "desc_periodic_kernel", "reduce_partials_kernel", "reduce_push_kernel", "wait_sum_kernel", "struct_errors_kernel"]
WATCH = ["DMMA", "HMMA", "UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTCATOMSWS", "DFMA", "DMUL", "DADD", "FFMA", "MUFU", "UTMALDG", "UBLKCP", "SYNCS", "LDS", "STS", "LDG", "STG", "LDL", "STL", "RED", "ATOM",
         "BAR", "SHFL", "LDSM", "LDC", "NOP", "BRA"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def main():
    want = sys.argv[1:] or HOT
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
    usage = {}
    cur = None
    for line in res.splitlines():
        m = re.match(r"\s*Function (\S+):", line)
        if m:
            cur = m.group(1)
            continue
        if cur and "REG:" in line:
            usage[cur] = line.strip()
            cur = None
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    hist = {}
    cur = None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            hist[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m:
            op = m.group(1)
            hist[cur]["_total"] += 1
            hist[cur][op.split(".")[0]] += 1
            if op.startswith("LDS.128") or op.startswith("LDS.64") or op.startswith("LDS.U.128"):
                hist[cur]["LDS(wide)"] += 1
    names = demangle(sorted(hist))
    print("# %s: resource usage and SASS mnemonic counts per kernel (static instruction counts, not executed counts)" % os.path.relpath(LIB, ROOT))
    for mangled in sorted(hist, key=lambda k: names[k]):
        pretty = names[mangled]
        if not any(w in pretty for w in want):
            continue
        h = hist[mangled]
        print("\n== %s" % pretty)
        print("   %s" % usage.get(mangled, "(no res-usage line)"))
        print("   instructions %d: " % h["_total"] + ", ".join("%s %d" % (k, h[k]) for k in WATCH + ["LDS(wide)"] if h[k]))


if __name__ == "__main__":
    main()
