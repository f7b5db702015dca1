"""Extracts the time series the reference holds for its own CPU run of samples/decay-reconnection (32^3 periodic isothermal
MHD decaying from a random non-helical field, 6th-order upwinding of u, lnrho and A, nu = 2e-5, eta = 4e-6, Courant time step,
4 processors 1 x 2 x 2, default random_gen='min_std'; in the reference's GPU auto-test list, bin/pc_auto-test:275) into
tests/golden/decay_reconnection_rows.txt.  Run in the build container (the reference tree does not exist on the GPU box):
    python tests/golden/make_decay_golden.py
"""
import os

SRC = "/root/reference/samples/decay-reconnection/reference.out"
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "decay_reconnection_rows.txt")

lines = open(SRC).read().splitlines()
names = [n for n in lines[0].replace("#", "").split("-") if n]
rows = [l.split() for l in lines[1:] if l.strip()]
assert [r[0] for r in rows] == ["0", "10", "20", "30", "40", "50"] and all(len(r) == len(names) for r in rows)
with open(DST, "w") as out:
    out.write("# " + " ".join(names) + "   (samples/decay-reconnection/reference.out, formats of print.in: e10.3 / e11.3)\n")
    for r in rows:
        out.write(" ".join(r) + "\n")
print("wrote", DST, names)
