"""Generates tests/golden/deriv_golden.npz by IMPORTING the reference's own numpy stencils
(/root/reference/python/pencil/math/derivatives/der_6th_order_w_ghosts.py, loaded standalone
because the full `pencil` package needs h5py).  Run in the build container only; the .npz is
committed because /root/reference does not exist on the GPU box.

    python tests/golden/make_deriv_golden.py
"""
import importlib.util
import os

import numpy as np

REF = "/root/reference/python/pencil/math/derivatives/der_6th_order_w_ghosts.py"
spec = importlib.util.spec_from_file_location("der6", REF)
der6 = importlib.util.module_from_spec(spec)
spec.loader.exec_module(der6)

rng = np.random.default_rng(20241104)
nx, ny, nz = 12, 10, 8
dx, dy, dz = 0.37, 0.23, 0.51
f = rng.standard_normal((nz + 6, ny + 6, nx + 6))
# periodic ghosts so that the interior stencils see consistent data
f[:, :, :3] = f[:, :, nx:nx + 3]; f[:, :, nx + 3:] = f[:, :, 3:6]
f[:, :3, :] = f[:, ny:ny + 3, :]; f[:, ny + 3:, :] = f[:, 3:6, :]
f[:3, :, :] = f[nz:nz + 3, :, :]; f[nz + 3:, :, :] = f[3:6, :, :]

out = {"f": f, "dxyz": np.array([dx, dy, dz])}
I = (slice(3, nz + 3), slice(3, ny + 3), slice(3, nx + 3))
for name, fn, h in [("der_x", der6.xder_6th, dx), ("der_y", der6.yder_6th, dy), ("der_z", der6.zder_6th, dz),
                    ("der2_x", der6.xder2_6th, dx), ("der2_y", der6.yder2_6th, dy), ("der2_z", der6.zder2_6th, dz),
                    ("der5_x", der6.xder5_6th, dx), ("der5_y", der6.yder5_6th, dy), ("der5_z", der6.zder5_6th, dz),
                    ("der6_x", der6.xder6_6th, dx), ("der6_y", der6.yder6_6th, dy), ("der6_z", der6.zder6_6th, dz)]:
    out[name] = np.ascontiguousarray(fn(f, h)[I])
np.savez_compressed(os.path.join(os.path.dirname(__file__), "deriv_golden.npz"), **out)
print("wrote deriv_golden.npz", {k: v.shape for k, v in out.items()})
