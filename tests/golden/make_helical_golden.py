"""Collects everything the reference holds about its own CPU run of samples/helical-MHDturb (32^3 isothermal MHD with
helical forcing, nu = eta = 5e-3, fixed dt = 4e-2, random_gen='nr_f90': config[1]'s physics) into
tests/golden/helical_mhdturb.json:

* rows     -- /root/reference/samples/helical-MHDturb/reference.out (it = 0, 2, 4, 6, 8; formats of print.in),
* kdat     -- samples/helical-MHDturb/k.dat (the run's forcing wavevectors: an INPUT file of the run, 98 vectors),
* aa       -- samples/helical-MHDturb/tests/read_data.ref: A_y(x=5, y=5, z=0..4) of var.dat after nt = 10 steps, %g,
* bymz     -- samples/helical-MHDturb/tests/read_xyaver.ref: <B_y>_xy(z=0..3) of the 4th record (t = 0.24), %g,
* ux2mxz, bymxy -- tests/read_yaver.ref, tests/read_zaver.ref: <u_x^2>_y(x=10, z=4..8) and <B_y>_z(x=10, y=4..8) of the records
              at t = 0 and t = 0.12, %g.

Run in the build container (the reference tree does not exist on the GPU box):
    python tests/golden/make_helical_golden.py
"""
import json
import os

S = "/root/reference/samples/helical-MHDturb"
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "helical_mhdturb.json")

lines = open(os.path.join(S, "reference.out")).read().splitlines()
header = lines[0]
names = [n for n in header.replace("#", "").split("-") if n]
rows = [l.split() for l in lines[1:] if l.strip()]
assert [r[0] for r in rows] == ["0", "2", "4", "6", "8"] and all(len(r) == len(names) for r in rows), (names, rows)


def ref_values(path, key):
    for l in open(path):
        if l.startswith(key):
            return l.split(":")[-1].split()
    raise KeyError(key)


out = {
    "source": "samples/helical-MHDturb/{reference.out,k.dat,tests/read_data.ref,tests/read_xyaver.ref,tests/read_yaver.ref,tests/read_zaver.ref,tests/read_slices.ref}",
    "columns": names,
    "rows": [" ".join(r) for r in rows],
    "kdat": open(os.path.join(S, "k.dat")).read(),
    "aa_5_5_0to4_y": ref_values(os.path.join(S, "tests", "read_data.ref"), "aa(5,5,0:4,1)"),
    "bymz_rec3_z0to3": ref_values(os.path.join(S, "tests", "read_xyaver.ref"), "av.xy.bymz[3,:4]"),
    # y- and z-averages, records 0 and 1 (t = 0 and t = 0.12, i.e. after 0 and 3 steps); index order of the readers: [rec, x, z|y]
    "ux2mxz_rec1_x10_z4to8": ref_values(os.path.join(S, "tests", "read_yaver.ref"), "av.y.ux2mxz[1,10,4:9]"),
    # video slice of B_z in the lowest xy plane, second record (dvid = 0.1: t = 0.12), [y=4..8, x=13]
    "bb3_xy_rec1_y4to8_x13": ref_values(os.path.join(S, "tests", "read_slices.ref"), "sl.xy.bb3[1,4:9,13]"),
    "bymxy_rec0_x10_y4to8": ref_values(os.path.join(S, "tests", "read_zaver.ref"), "av.z.bymxy[0,10,4:9]"),
    "bymxy_rec1_x10_y4to8": ref_values(os.path.join(S, "tests", "read_zaver.ref"), "av.z.bymxy[1,10,4:9]"),
}
with open(DST, "w") as fh:
    json.dump(out, fh, indent=1)
print("wrote", DST, names)
