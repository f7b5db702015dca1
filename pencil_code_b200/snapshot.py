"""var.dat snapshots of one rank in the reference's distributed-I/O layout.

Host-side mirror of `output_snap` / `input_snap` (src/io_dist.f90:108-199, 657-788; called from `wsnap` / `rsnap`,
src/snapshot.f90:70-383) over `pcb200_write_snapshot` / `pcb200_read_snapshot` of libpcb200.so.  The f-array is the
host array of `module GPU` (shape (nvar, mz, my, mx), C order = Fortran f(mx,my,mz,nvar)); together with
`GPU.copy_farray_from_GPU_async` / `copy_farray_wait` the device keeps stepping while the snapshot is written.
"""
import os

import numpy as np

from . import gpu as _gpu

NGHOST = 3


def mesh(cfg):
    """x(1:mx), y(1:my), z(1:mz) of this rank: equidistant, cfg x0/y0/z0 = x(l1), y(m1), z(n1) (grid.f90:123-149)."""
    out = []
    for n, o, d in ((cfg["nx"], cfg["x0"], cfg["dx"]), (cfg["ny"], cfg["y0"], cfg["dy"]), (cfg["nz"], cfg["z0"], cfg["dz"])):
        out.append(o + d * (np.arange(n + 2 * NGHOST, dtype=np.float64) - NGHOST))
    return out


def snapshot_path(datadir, iproc, name="var.dat"):
    """data/proc<iproc>/var.dat (directory_names_std, src/io_dist.f90:71-106)."""
    return os.path.join(datadir, "proc%d" % iproc, name)


def _check(rc, what, path):
    if rc == 0:
        return
    why = {1: "I/O error (cannot open, short file or write failure)", 2: "not a var.dat of the expected sizes", 3: "bad argument"}
    raise IOError("%s %s: %s" % (what, path, why.get(rc, "status %d" % rc)))


def output_snap(path, f, t, cfg):
    """write f (nvar, mz, my, mx) and {t, x, y, z, dx, dy, dz} as the two records of var.dat"""
    L = _gpu.load_library()
    nvar, mz, my, mx = f.shape
    if (mx, my, mz) != (cfg["nx"] + 6, cfg["ny"] + 6, cfg["nz"] + 6):
        raise ValueError("f-array shape %s does not match the configuration" % (f.shape,))
    x, y, z = mesh(cfg)
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    rc = L.pcb200_write_snapshot(path.encode(), _gpu._dp(f), mx, my, mz, nvar, float(t), _gpu._dp(x), _gpu._dp(y),
                                 _gpu._dp(z), cfg["dx"], cfg["dy"], cfg["dz"])
    _check(rc, "output_snap", path)


def input_snap(path, cfg, nvar=None):
    """read var.dat; returns f (nvar, mz, my, mx), t, (x, y, z), (dx, dy, dz)"""
    L = _gpu.load_library()
    nvar = cfg["nvar"] if nvar is None else nvar
    mx, my, mz = cfg["nx"] + 6, cfg["ny"] + 6, cfg["nz"] + 6
    f = np.empty((nvar, mz, my, mx), dtype=np.float64)
    x, y, z = np.empty(mx), np.empty(my), np.empty(mz)
    t, d3 = np.zeros(1), np.zeros(3)
    rc = L.pcb200_read_snapshot(path.encode(), _gpu._dp(f), mx, my, mz, nvar, _gpu._dp(t), _gpu._dp(x), _gpu._dp(y),
                                _gpu._dp(z), _gpu._dp(d3))
    _check(rc, "input_snap", path)
    return f, float(t[0]), (x, y, z), tuple(d3)
