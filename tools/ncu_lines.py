#!/usr/bin/env python
"""Attribute ncu per-instruction counters to CUDA source lines.

ncu's `--page source --csv` export lists SASS instructions with their counters but no line numbers; nvdisasm -g
lists the same instruction sequence annotated with `//## File "...", line N`.  This tool joins the two by
instruction order.

    ncu -i prof.ncu-rep --page source --csv --kernel-name regex:NAME --launch-count 1 > sass.csv
    cuobjdump -xelf all lib.so ; nvdisasm -g -c X.cubin > x.sass
    python tools/ncu_lines.py sass.csv x.sass MANGLED_FUNCTION_NAME [top]
"""
import collections
import csv
import re
import sys


def load_ncu(path):
    rows = list(csv.reader(open(path)))
    hdr = None
    out = []
    for r in rows:
        if r and r[0] == "Address":
            hdr = r
            continue
        if hdr and len(r) == len(hdr):
            out.append(dict(zip(hdr, r)))
    return out


def load_sass(path, func):
    lines = open(path).read().split("\n")
    start = None
    for i, l in enumerate(lines):
        if l.startswith(".text." + func + ":"):
            start = i
            break
    if start is None:
        raise SystemExit("function %s not found" % func)
    cur = ("?", 0)
    seq = []
    for l in lines[start + 1:]:
        if l.startswith(".text.") or l.startswith("//--------------------- .") and ".text." in l and seq:
            break
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
        if m:
            seq.append((cur, m.group(2).strip()))
    return seq


def main():
    ncu = load_ncu(sys.argv[1])
    seq = load_sass(sys.argv[2], sys.argv[3])
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    n = min(len(ncu), len(seq))
    if abs(len(ncu) - len(seq)) > 8:
        print("warning: %d ncu instructions vs %d disassembled" % (len(ncu), len(seq)))
    agg = collections.defaultdict(lambda: [0.0, 0.0, collections.Counter()])
    ti = ts = 0.0
    for k in range(n):
        d = ncu[k]
        inst = float(d.get("Instructions Executed", "0") or 0)
        samp = float(d.get("# Samples", "0") or 0)
        key = seq[k][0]
        agg[key][0] += inst
        agg[key][1] += samp
        agg[key][2][seq[k][1].split()[0]] += inst
        ti += inst
        ts += samp
    print("instructions %.3g  samples %.0f" % (ti, ts))
    srcs = {}
    for key, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
        f, ln = key
        txt = ""
        try:
            if f not in srcs:
                import glob
                hits = glob.glob("/root/repo/**/" + f, recursive=True)
                srcs[f] = open(hits[0]).read().split("\n") if hits else []
            txt = srcs[f][ln - 1].strip()[:70] if srcs[f] else ""
        except Exception:
            pass
        ops = ",".join("%s" % o for o, _ in v[2].most_common(3))
        print("%-12s L%-4d inst %5.1f%%  stall-samples %5.1f%%  [%s]  %s" % (f, ln, 100 * v[0] / ti, 100 * v[1] / max(ts, 1), ops, txt))


if __name__ == "__main__":
    main()
