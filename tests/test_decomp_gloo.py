"""CPU tests of the multi-GPU host logic (world_size > 1 over gloo): rank placement, neighbour ranks and the
halo slabs of pencil_code_b200.decomp -- the same slabs libpcb200.so exchanges with NCCL (exchange_y /
exchange_z) -- checked by decomposition invariance: a domain advanced rank-by-rank with exchanged halos must
equal the single-domain result of the oracle bit for bit (SURVEY.md 8e "correctness check")."""
import os
import socket
import sys

import numpy as np
import pytest

from pencil_code_b200 import config as pcfg
from pencil_code_b200 import decomp, synthetic

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ALPHA = (0.0, -5.0 / 9.0, -153.0 / 128.0)  # timestep.f90:39-44
BETA = (1.0 / 3.0, 15.0 / 16.0, 8.0 / 15.0)


def test_rank_layout_and_neighbours():
    assert decomp.choose_decomposition(8, "slab") == (1, 8)
    assert decomp.choose_decomposition(8, "pencil") == (2, 4)
    assert decomp.choose_decomposition(4, "pencil") == (2, 2)
    assert decomp.choose_decomposition(1, "pencil") == (1, 1)
    for npy, npz in ((1, 2), (2, 4), (1, 8), (2, 2)):
        seen = set()
        for r in range(npy * npz):
            ipy, ipz = decomp.rank_position(r, npy, npz)
            assert decomp.rank_of(ipy, ipz, npy, npz) == r
            seen.add((ipy, ipz))
            nb = decomp.neighbours(r, npy, npz)
            # my upper neighbour's lower neighbour is me
            assert decomp.neighbours(nb["zhi"], npy, npz)["zlo"] == r
            assert decomp.neighbours(nb["yhi"], npy, npz)["ylo"] == r
        assert len(seen) == npy * npz
    # 512^3 / 8 z-slabs, 8 variables: 51.5 MB per z face
    ydbl, zdbl = decomp.halo_message_doubles(512, 512, 64, 8)
    assert zdbl * 8 == 3 * 518 * 518 * 8 * 8 and ydbl == 3 * 518 * 64 * 8


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, npy, npz, port, outdir, ldt):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pc_oracle
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ipy, ipz = decomp.rank_position(rank, npy, npz)
    n = 12
    gcfg = pcfg.make_config(n, True, True, ny=16, nz=24, nu=5e-3, eta=4e-3, chi=3e-3, cdt=0.4, cdtv=0.3)
    cfg = pcfg.make_config(n, True, True, ny=16, nz=24, nu=5e-3, eta=4e-3, chi=3e-3, cdt=0.4, cdtv=0.3,
                           nprocy=npy, ipy=ipy, nprocz=npz, ipz=ipz)
    fglob = synthetic.initial_condition(gcfg)
    f = decomp.scatter(fglob, rank, npy, npz)
    # the per-rank generator (hashed noise over GLOBAL indices) gives the same field up to coordinate rounding
    floc = synthetic.initial_condition(cfg)
    assert np.abs(floc - f)[:, 3:-3, 3:-3, 3:-3].max() < 1e-15
    o = pc_oracle.Oracle(cfg, nthreads=1)
    df = np.zeros_like(f)

    def sendrecv(arr, dest, src, tag):
        if dest == rank and src == rank:
            return arr.copy()
        out = torch.empty(arr.shape, dtype=torch.float64)
        reqs = [dist.isend(torch.from_numpy(arr), dest, tag=tag), dist.irecv(out, src, tag=tag)]
        for r in reqs:
            r.wait()
        return out.numpy()

    dt = 0.0 if ldt else pcfg.fixed_dt(gcfg)
    for step in range(2):
        for it in range(3):
            decomp.exchange_halos(f, rank, npy, npz, sendrecv)
            df *= ALPHA[it]
            dt1, _ = o.rhs(f, df, lcourant=(ldt and it == 0))
            if ldt and it == 0:  # sub.f90:4241-4254: dt = 1/allreduce_max(dt1)
                t1 = torch.tensor([dt1], dtype=torch.float64)
                dist.all_reduce(t1, op=dist.ReduceOp.MAX)
                dt = 1.0 / float(t1.item())
            f[:, 3:-3, 3:-3, 3:-3] += dt * BETA[it] * df[:, 3:-3, 3:-3, 3:-3]
    np.save(os.path.join(outdir, "f_%d.npy" % rank), f)
    if rank == 0:
        np.save(os.path.join(outdir, "dt.npy"), np.array([dt]))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("npy,npz,ldt", [(1, 2, False), (1, 2, True), (2, 2, False)])
def test_decomposition_invariance_over_gloo(oracle_mod, tmp_path, npy, npz, ldt):
    import torch.multiprocessing as mp
    world = npy * npz
    mp.spawn(_worker, args=(world, npy, npz, _free_port(), str(tmp_path), ldt), nprocs=world, join=True)
    gcfg = pcfg.make_config(12, True, True, ny=16, nz=24, nu=5e-3, eta=4e-3, chi=3e-3, cdt=0.4, cdtv=0.3)
    fg = synthetic.initial_condition(gcfg)
    o = oracle_mod.Oracle(gcfg, nthreads=1)
    df = np.zeros_like(fg)
    dt, t = (0.0 if ldt else pcfg.fixed_dt(gcfg)), 0.0
    for _ in range(2):
        dt, t, _ = o.time_step(fg, df, ldt, dt, t)
    got = np.zeros_like(fg)
    for r in range(world):
        decomp.gather_into(got, np.load(tmp_path / ("f_%d.npy" % r)), r, npy, npz)
    assert np.array_equal(got[:, 3:-3, 3:-3, 3:-3], fg[:, 3:-3, 3:-3, 3:-3])
    if ldt:
        assert float(np.load(tmp_path / "dt.npy")[0]) == dt
