"""The two time series the reference holds for config[1] at its own size: samples/turbulence/helical-MHDturb-nopart and
samples/turbulence/helical-MHDturb32-4procs (128^3 isothermal MHD, Gaussian-noise A of amplitude 1e-4, helical forcing
force=0.05 relhel=1 with k.dat = the one of samples/helical-MHDturb, nu = eta = 2e-3, Courant time step, default generator
'min_std', 10 steps, it1 = 5) -> tests/golden/helical128_rows.txt.  The two runs differ only in the processor layout that drew
the initial noise; the tree does not record it (both cparam.local say 1 x 4 x 8) -- tests/test_oracle_helical128.py finds that
the first was made on 1 x 4 x 8 and the second on 1 x 2 x 2 processors.
    python tests/golden/make_helical128_golden.py
"""
import os

S = "/root/reference/samples/turbulence"
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "helical128_rows.txt")
with open(DST, "w") as out:
    for name in ("helical-MHDturb-nopart", "helical-MHDturb32-4procs"):
        lines = open(os.path.join(S, name, "reference.out")).read().splitlines()
        cols = [n for n in lines[0].replace("#", "").split("-") if n]
        out.write("# %s: %s\n" % (name, " ".join(cols)))
        for l in lines[1:]:
            if l.strip():
                assert len(l.split()) == len(cols)
                out.write(name + " " + " ".join(l.split()) + "\n")
assert open(os.path.join(S, "helical-MHDturb-nopart", "k.dat")).read() == open("/root/reference/samples/helical-MHDturb/k.dat").read()
print(open(DST).read())
