"""Extracts the time-series rows the reference's own CPU binary printed for samples/gputest
(/root/reference/samples/gputest/cpu-test.out:273-292: 32^3, hydro+density+entropy+magnetic, eos_idealgas, no forcing,
inituu='ABC' k=2, lnrho=1.23, nu=eta=1e-3, no heat conduction, fixed dt=1e-3, 20 steps, it1=1) into
tests/golden/gputest_cpu_rows.txt.  Run in the build container (the reference tree does not exist on the GPU box):
    python tests/golden/make_gputest_golden.py
Columns kept (print.in): it, dt(e9.2), urms(f10.5), umax(f10.5), epsK(e10.3), rhom(f9.4), rhorms(f9.4)."""
import os
import re

SRC = "/root/reference/samples/gputest/cpu-test.out"
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "gputest_cpu_rows.txt")

rows = []
with open(SRC) as fh:
    lines = fh.read().splitlines()
start = next(i for i, l in enumerate(lines) if l.lstrip().startswith("----it----"))
for l in lines[start + 1:]:
    if not re.match(r"^\s+\d+\s", l):
        break
    t = l.split()
    rows.append((int(t[0]), t[2], t[3], t[4], t[5], t[6], t[7]))
assert len(rows) == 20 and rows[0][0] == 0 and rows[-1][0] == 19, rows
with open(DST, "w") as out:
    out.write("# it dt urms umax epsK rhom rhorms  (samples/gputest/cpu-test.out:%d-%d, formats of print.in)\n" % (start + 2, start + 21))
    for r in rows:
        out.write(" ".join(str(x) for x in r) + "\n")
print("wrote", DST, len(rows), "rows")
