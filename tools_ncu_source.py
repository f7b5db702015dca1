#!/usr/bin/env python
"""Aggregate the ncu source page (per-SASS-instruction stall samples) by opcode and list hot spots.
usage: python tools_ncu_source.py <rep> [kernel-id like :::1] [topN]"""
import csv, subprocess, sys, collections, re
rep = sys.argv[1]; kid = sys.argv[2] if len(sys.argv) > 2 else ":::1"; top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-id", kid], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]; data = rows[hi + 1:]
ci = {n: i for i, n in enumerate(h)}
stalls = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
tot = collections.Counter(); byop = collections.defaultdict(collections.Counter); nexec = collections.Counter()
recs = []
for k, r in enumerate(data):
    if len(r) < len(h) or r[0] == "Address" or not r[0].startswith("0x"): continue
    src = r[ci["Source"]].strip()
    m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_.]+)", src)
    op = m.group(2).split(".")[0] if m else "?"
    ns = int(r[ci["# Samples"]] or 0)
    ex = int(r[ci["Instructions Executed"]] or 0)
    nexec[op] += ex
    for s in stalls:
        v = int(r[ci[s]] or 0)
        tot[s] += v; byop[op][s] += v
    recs.append((ns, k, src, {s: int(r[ci[s]] or 0) for s in stalls if int(r[ci[s]] or 0) > 0}, ex))
T = sum(tot.values())
print("total samples", T)
for s, v in tot.most_common(): print("  %-28s %8d %5.1f%%" % (s, v, 100 * v / T))
print("\nby opcode (samples, %% of all, executed warp-instr):")
for op, c in sorted(byop.items(), key=lambda x: -sum(x[1].values()))[:18]:
    sv = sum(c.values())
    print("  %-10s %8d %5.1f%%  exec %10d  top: %s" % (op, sv, 100 * sv / T, nexec[op], ", ".join("%s %d" % (a[6:], b) for a, b in c.most_common(3))))
print("\nhot instructions:")
for ns, k, src, st, ex in sorted(recs, reverse=True)[:top]:
    print("  %6d  #%4d  %-60s %s" % (ns, k, src[:60], {a[6:]: b for a, b in st.items()}))
