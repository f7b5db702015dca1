"""Run under torchrun (one rank per GPU): the NCCL halo exchange of libpcb200.so (z-slabs, then yz-pencils when
the world allows) must reproduce the single-domain oracle.  Prints MULTI_GPU_CHECK PASS on rank 0.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tests/multi_gpu_check.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from pencil_code_b200 import config as pcfg, decomp, synthetic  # noqa: E402
from pencil_code_b200.gpu import GPU, nccl_unique_id  # noqa: E402


def run_case(rank, world, npy, npz, variant, ldt, p2p, nsteps=3):
    ipy, ipz = decomp.rank_position(rank, npy, npz)
    kw = dict(ny=32, nz=48, nu=5e-3, eta=4e-3, chi=3e-3, cdt=0.4, cdtv=0.3)
    gcfg = pcfg.make_config(64, True, True, **kw)
    cfg = pcfg.make_config(64, True, True, nprocy=npy, ipy=ipy, nprocz=npz, ipz=ipz, **kw)
    cfg.update(device=int(os.environ.get("LOCAL_RANK", "0")), kernel_variant=variant)
    fglob = synthetic.initial_condition(gcfg)
    f = decomp.scatter(fglob, rank, npy, npz)
    ids = [nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ids, src=0)
    g = GPU(cfg, nccl=(ids[0], rank, world))
    g.register_GPU()
    g.initialize_GPU(f)
    if p2p:  # z halos (and y faces of pencils) as copy-engine puts into the neighbours' ghost zones (CUDA IPC) instead of NCCL
        blobs = [None] * world
        dist.all_gather_object(blobs, g.p2p_export())
        oks = [None] * world
        dist.all_gather_object(oks, g.p2p_connect(blobs))
        assert all(oks), "peer mapping failed on some rank: %s" % oks
        g.p2p_enable()
        # the Fortran order (connect_p2p_halo and p2p_enable come before load_farray_to_GPU): upload again with the host's ghost
        # zones poisoned -- a neighbour's put that lands before the tail of my upload must not be overwritten by stale host ghosts
        f[:, :3] = np.nan; f[:, -3:] = np.nan
        f[:, :, :3] = np.nan; f[:, :, -3:] = np.nan
        f[:, :, :, :3] = np.nan; f[:, :, :, -3:] = np.nan
        g.load_farray_to_GPU(f)
    dt, t = (0.0 if ldt else pcfg.fixed_dt(gcfg)), 0.0
    series = []
    for s in range(nsteps):
        dt, t = g.time_step(f, dt, ldt, t, diagnostics=True)
        series.append(g.get_gpu_reduced_vars())
    g.copy_farray_from_GPU(f)
    dist.barrier()  # nobody frees ghost planes a neighbour may still be writing to
    g.finalize_GPU()
    parts = [None] * world
    dist.all_gather_object(parts, f)
    ok = True
    if rank == 0:
        import pc_oracle
        o = pc_oracle.Oracle(gcfg)
        fc, dfc = fglob.copy(), np.zeros_like(fglob)
        dtc, tc = (0.0 if ldt else pcfg.fixed_dt(gcfg)), 0.0
        sc = []
        for s in range(nsteps):
            dtc, tc, d = o.time_step(fc, dfc, ldt, dtc, tc, diag=True)
            sc.append(d)
        got = np.zeros_like(fglob)
        for r in range(world):
            decomp.gather_into(got, parts[r], r, npy, npz)
        err = 0.0
        for k in range(got.shape[0]):
            a, b = got[k, 3:-3, 3:-3, 3:-3], fc[k, 3:-3, 3:-3, 3:-3]
            err = max(err, np.abs(a - b).max() / np.abs(b).max())
        derr = max(abs(a[k] - b[k]) / max(abs(b[k]), 1e-300) for a, b in zip(series, sc) for k in ("urms", "brms", "jrms", "rhom"))
        ok = err <= 1e-11 and derr <= 1e-11 and abs(dt - dtc) <= 1e-12 * dtc
        print("case %dx%d variant %d ldt %d %s: rel Linf %.2e  diag %.2e  dt %.3e/%.3e  %s" %
              (npy, npz, variant, ldt, "p2p" if p2p else "nccl", err, derr, dt, dtc, "ok" if ok else "FAIL"), flush=True)
    flag = [ok]
    dist.broadcast_object_list(flag, src=0)
    return flag[0]


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("gloo")
    cases = [(1, world, 10, False, True), (1, world, 5, True, True), (1, world, 10, True, False), (1, world, 4, True, True)]
    if world % 2 == 0:
        cases += [(2, world // 2, 10, True, True), (2, world // 2, 4, False, False), (world, 1, 1, False, False), (world, 1, 10, True, False),
                  (world, 1, 10, False, True), (2, world // 2, 5, False, True)]
    ok = True
    for npy, npz, variant, ldt, p2p in cases:
        ok = run_case(rank, world, npy, npz, variant, ldt, p2p) and ok
    if rank == 0:
        print("MULTI_GPU_CHECK " + ("PASS" if ok else "FAIL"), flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
