"""Domain decomposition of the hot path across GPUs (host-side logic, no device code).

Mirrors the reference's processor layout and halo slabs for nprocx=1:
  * rank = ipy + nprocy*ipz                               (mpicomm.f90: iproc = ipx + nprocx*(ipy + nprocy*ipz))
  * y faces  f(:, m1:m1i, n1:n2, :) / f(:, m2i:m2, n1:n2, :)  -- mx wide, interior z   (mpicomm.f90:1604-1731)
  * z faces  f(:, :,     n1:n1i, :) / f(:, :,     n2i:n2, :)  -- mx*my wide: sent after the y exchange so that
    they carry the freshly filled y ghosts (edges/corners without separate corner messages)
libpcb200.so implements exactly these slabs (exchange_y / exchange_z in csrc/pcb200.cu); this module is what
bench.py uses to place ranks and what the gloo CPU tests exercise.
"""
import numpy as np

NG = 3


def choose_decomposition(world, kind="slab"):
    """(nprocy, nprocz) for `world` ranks: z-slabs 1 x G, or the squarest yz-pencil grid with nprocy <= nprocz."""
    if kind == "slab" or world == 1:
        return 1, world
    best = (1, world)
    for py in range(1, int(world ** 0.5) + 1):
        if world % py == 0:
            best = (py, world // py)
    return best


def rank_position(rank, nprocy, nprocz):
    assert 0 <= rank < nprocy * nprocz
    return rank % nprocy, rank // nprocy  # ipy, ipz


def rank_of(ipy, ipz, nprocy, nprocz):
    return (ipy % nprocy) + nprocy * (ipz % nprocz)


def neighbours(rank, nprocy, nprocz):
    """periodic neighbours {'ylo','yhi','zlo','zhi'} -> rank"""
    ipy, ipz = rank_position(rank, nprocy, nprocz)
    return dict(ylo=rank_of(ipy - 1, ipz, nprocy, nprocz), yhi=rank_of(ipy + 1, ipz, nprocy, nprocz),
                zlo=rank_of(ipy, ipz - 1, nprocy, nprocz), zhi=rank_of(ipy, ipz + 1, nprocy, nprocz))


def halo_message_doubles(nx, ny, nz, nvar):
    """doubles per message: (y face, z face) for local interior sizes nx, ny, nz"""
    mx, my = nx + 2 * NG, ny + 2 * NG
    return NG * mx * nz * nvar, NG * mx * my * nvar


def wrap_x(f):
    """periodic x ghosts, local (boundcond.f90:2387-2409, nprocx=1); f[var, z, y, x]"""
    nx = f.shape[-1] - 2 * NG
    f[..., :NG] = f[..., nx:nx + NG]
    f[..., nx + NG:] = f[..., NG:2 * NG]


def y_faces(f):
    """(send_to_yhi, send_to_ylo): top / bottom 3 interior rows, all x, interior z"""
    ny = f.shape[-2] - 2 * NG
    return (np.ascontiguousarray(f[:, NG:-NG, ny:ny + NG, :]), np.ascontiguousarray(f[:, NG:-NG, NG:2 * NG, :]))


def put_y_ghosts(f, from_ylo, from_yhi):
    ny = f.shape[-2] - 2 * NG
    f[:, NG:-NG, :NG, :] = from_ylo
    f[:, NG:-NG, ny + NG:, :] = from_yhi


def z_faces(f):
    """(send_to_zhi, send_to_zlo): top / bottom 3 interior planes, all x and y (ghosts included)"""
    nz = f.shape[-3] - 2 * NG
    return (np.ascontiguousarray(f[:, nz:nz + NG]), np.ascontiguousarray(f[:, NG:2 * NG]))


def put_z_ghosts(f, from_zlo, from_zhi):
    nz = f.shape[-3] - 2 * NG
    f[:, :NG] = from_zlo
    f[:, nz + NG:] = from_zhi


def exchange_halos(f, rank, nprocy, nprocz, sendrecv):
    """Complete ghost update of a local f-array in the library's order: x wrap, y faces, z faces.
    sendrecv(send_array, dest_rank, src_rank, tag) -> received array (same shape)."""
    nb = neighbours(rank, nprocy, nprocz)
    wrap_x(f)
    up, dn = y_faces(f)
    lo = sendrecv(up, nb["yhi"], nb["ylo"], 0)  # my top rows go up; my low ghosts come from below
    hi = sendrecv(dn, nb["ylo"], nb["yhi"], 1)
    put_y_ghosts(f, lo, hi)
    up, dn = z_faces(f)
    lo = sendrecv(up, nb["zhi"], nb["zlo"], 2)
    hi = sendrecv(dn, nb["zlo"], nb["zhi"], 3)
    put_z_ghosts(f, lo, hi)


def scatter(fglobal, rank, nprocy, nprocz):
    """local f-array (with ghost frame, ghosts unset) of `rank` cut from a global array with ghost frame"""
    ipy, ipz = rank_position(rank, nprocy, nprocz)
    nvar, mzg, myg, mx = fglobal.shape
    ny, nz = (myg - 2 * NG) // nprocy, (mzg - 2 * NG) // nprocz
    loc = np.zeros((nvar, nz + 2 * NG, ny + 2 * NG, mx))
    loc[:, NG:-NG, NG:-NG, :] = fglobal[:, NG + ipz * nz:NG + (ipz + 1) * nz, NG + ipy * ny:NG + (ipy + 1) * ny, :]
    return loc


def gather_into(fglobal, loc, rank, nprocy, nprocz):
    ipy, ipz = rank_position(rank, nprocy, nprocz)
    ny, nz = loc.shape[-2] - 2 * NG, loc.shape[-3] - 2 * NG
    fglobal[:, NG + ipz * nz:NG + (ipz + 1) * nz, NG + ipy * ny:NG + (ipy + 1) * ny, NG:-NG] = loc[:, NG:-NG, NG:-NG, NG:-NG]
