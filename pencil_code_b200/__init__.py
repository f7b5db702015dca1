"""pencil_code_b200 -- B200-native (sm_100a) GPU module for Pencil Code's per-substep RHS + RK3 path.

Only the hot path lives here: csrc/ (CUDA kernels + C ABI, built into libpcb200.so), the host-side
mirror of the reference's `module GPU` (gpu.py), run configurations (config.py), the synthetic
initial condition (synthetic.py) and the Fortran drop-in module (fortran/gpu_b200.f90).
"""
from . import config, synthetic  # noqa: F401
from .gpu import GPU, load_library, nccl_unique_id  # noqa: F401
