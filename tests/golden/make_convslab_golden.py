"""Extracts the double-precision time series the reference ships for samples/conv-slab
(/root/reference/samples/conv-slab/reference.out.double: 32^3 convection in a vertically stratified slab, hydro + density +
entropy + gravity_simple, 41 steps with Courant dt, rows every 10 steps, 13-14 significant digits) into
tests/golden/convslab_rows.txt.  Run in the build container (the reference tree does not exist on the GPU box):
    python tests/golden/make_convslab_golden.py
Columns (samples/conv-slab/print.in.double): it t dt urms umax rhom ssm dtc dtu dtnu dtchi."""
import os

SRC = "/root/reference/samples/conv-slab/reference.out.double"
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "convslab_rows.txt")

rows = [l.split() for l in open(SRC) if l.strip() and not l.startswith("#")]
assert [int(r[0]) for r in rows] == [0, 10, 20, 30, 40] and all(len(r) == 11 for r in rows), rows
# samples/conv-slab_cp_2 (cp = 2, hcond0 doubled) holds the same rows digit for digit, its entropy column being ssbycpm = ssm/cp
rows2 = [l.split() for l in open("/root/reference/samples/conv-slab_cp_2/reference.out.double") if l.strip() and not l.startswith("#")]
assert rows2 == rows
with open(DST, "w") as out:
    out.write("# it t dt urms umax rhom ssm dtc dtu dtnu dtchi  (samples/conv-slab/reference.out.double, formats of print.in.double)\n")
    for r in rows:
        out.write(" ".join(r) + "\n")
print("wrote", DST, len(rows), "rows")
