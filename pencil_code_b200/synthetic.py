"""Synthetic initial condition of SURVEY.md 8(d), numpy edition (host arrays for parity tests).

lnrho=0; u = 0.1*ABC(k=1); A = 0.01*ABC(k=2) (Beltrami); ss = 0.01 sin x cos y; plus hashed
noise 1e-3*(U-1/2) from a splitmix64 counter over the GLOBAL cell index, so every domain
decomposition reproduces the same field.  The device generator in csrc/pcb200.cu
(init_synthetic_kernel) uses the same formulae.
"""
import numpy as np

NG = 3
_M = np.uint64(0xFFFFFFFFFFFFFFFF)


def _splitmix(z):
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def hash_noise(gx, gy, gz, k, seed):
    with np.errstate(over="ignore"):
        cnt = np.uint64(1) + gx + np.uint64(4096) * (gy + np.uint64(4096) * (gz + np.uint64(4096) * np.uint64(k)))
        z = np.uint64(seed) + np.uint64(0x9E3779B97F4A7C15) * cnt
        z = _splitmix(z)
    return (z >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def farray_shape(cfg):
    return (cfg["nvar"], cfg["nz"] + 2 * NG, cfg["ny"] + 2 * NG, cfg["nx"] + 2 * NG)


def initial_condition(cfg, seed=12345, noise=1e-3):
    """Returns f (nvar, mz, my, mx) float64 with interior set and ghosts zero."""
    nx, ny, nz = cfg["nx"], cfg["ny"], cfg["nz"]
    f = np.zeros(farray_shape(cfg))
    i = np.arange(nx); j = np.arange(ny); l = np.arange(nz)
    x = (cfg["x0"] + i * cfg["dx"])[None, None, :]
    y = (cfg["y0"] + j * cfg["dy"])[None, :, None]
    z = (cfg["z0"] + l * cfg["dz"])[:, None, None]
    ent, mag = cfg["iss"] >= 0, cfg["iax"] >= 0
    vals = [0.1 * (np.sin(z) + np.cos(y)) + 0 * x,
            0.1 * (np.sin(x) + np.cos(z)) + 0 * y,
            0.1 * (np.sin(y) + np.cos(x)) + 0 * z,
            np.zeros((nz, ny, nx))]
    if ent:
        vals.append(0.01 * np.sin(x) * np.cos(y) + 0 * z)
    if mag:
        vals += [0.01 * (np.sin(2 * z) + np.cos(2 * y)) + 0 * x,
                 0.01 * (np.sin(2 * x) + np.cos(2 * z)) + 0 * y,
                 0.01 * (np.sin(2 * y) + np.cos(2 * x)) + 0 * z]
    gx = (np.uint64(cfg.get("ipx", 0) * nx) + i.astype(np.uint64))[None, None, :]
    gy = (np.uint64(cfg.get("ipy", 0) * ny) + j.astype(np.uint64))[None, :, None]
    gz = (np.uint64(cfg.get("ipz", 0) * nz) + l.astype(np.uint64))[:, None, None]
    for k, v in enumerate(vals):
        n = hash_noise(gx, gy, gz, k, seed) if noise else 0.0
        f[k, NG:NG + nz, NG:NG + ny, NG:NG + nx] = v + noise * (n - 0.5)
    return f
