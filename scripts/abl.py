#!/usr/bin/env python
"""Lab harness (no torch): per-substep time of the fused kernel for one library build.
usage: PCB200_LIB=build/abl/libpcb200_<m>.so python scripts/abl.py [n=512] [steps=3] [variant=0]"""
import os, sys, types
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ctypes as C
from pencil_code_b200 import gpu as G
if os.environ.get("PCB200_LIB"):
    G.LIB_PATH = os.path.abspath(os.environ["PCB200_LIB"])
import bench
n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
kv = int(sys.argv[3]) if len(sys.argv) > 3 else 0
wl = sys.argv[4] if len(sys.argv) > 4 else "mhd_entropy"
args = types.SimpleNamespace(workload=wl, n=n)
cfg, wname = bench.workload_config(args, n, 1, 0, 1, 0)
cfg["device"] = 0; cfg["kernel_variant"] = kv
from pencil_code_b200 import config as pcfg
dt = pcfg.fixed_dt(cfg)
g = G.GPU(cfg)
g.register_GPU()
z = C.c_int(0); ct = C.c_double(0.0)
g.L.initialize_gpu_c(None, C.byref(z), C.byref(ct), C.byref(z), C.byref(z), C.byref(z), C.byref(z))
g.init_synthetic(12345)
forcing = G.HelicalForcing(pcfg.kvectors())
t = 0.0
for _ in range(2):
    _, t = g.time_step(None, dt, False, t, forcing=forcing.draw(dt))
g.synchronize()
g.enable_kernel_timing(True); g.kernel_timing(reset=True)
tot = [0.0, 0.0, 0.0]
for _ in range(steps):
    g.set_dt(dt, False)
    for it in (1, 2, 3):
        if it == 3: g.set_forcing_hel(forcing.draw(dt))
        g.before_boundary_gpu(None, False, it, t)
        g.rhs_GPU(None, it, t)
        g.synchronize()
        ms, nl = g.kernel_timing(reset=True)
        tot[it - 1] += ms
cells = n ** 3
per = [x / steps for x in tot]
print("%s n=%d variant=%d  substep ms %.3f %.3f %.3f  sum %.3f  -> %.2f G cell-substeps/s (kernel only)" %
      (os.path.basename(G.LIB_PATH), n, kv, per[0], per[1], per[2], sum(per), cells * 3 / (sum(per) * 1e-3) / 1e9), flush=True)
g.finalize_GPU()
