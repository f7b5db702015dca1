#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric for the per-substep RHS path.

  metric : MHD cell-updates/s (RK3 substep), i.e. nx*ny*nz*3 cell-substep-updates per full step
  config : periodic MHD with entropy 512^3 (8 variables, FP64); z-slab decomposed over N GPUs
           (strong scaling: the global grid is fixed)
  step   : one full 2N-RK3 time step (3 fused rhs+update substeps, fixed dt) on synthetic data

  python bench.py [--gpus N] [--steps K] [--warmup W] [--n 512] [--impl b200|reference]

One JSON line on stdout (rank 0).  `value` is device-resident throughput, `e2e` goes through the
reference-facing module-GPU calls with HOST buffers (load_farray_to_GPU + 3 x rhs_GPU +
copy_farray_from_GPU every step), `roofline` is measured with CUDA events around the fused kernel,
`cpu_baseline` times the CPU oracle (a restatement, not the Fortran binary: no gfortran here).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "MHD cell-updates/s (RK3 substep)"
UNIT = "cell-substep-updates/s"


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """samples nvidia-smi during the timed region (B200_PROFILING.md clocks line)"""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); pw.append(float(r[2]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                continue
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def host_threads():
    """threads this process may actually use (torchrun exports OMP_NUM_THREADS=1: the affinity mask is what counts)"""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


CPU_WINDOW_NZ = 64  # z-planes of the bounded CPU sample


def cpu_oracle_throughput(args, steps, warmup=1):
    """cell-substep-updates/s of the CPU restatement (-O3 -march=native, OpenMP over (m,n) pencils, all host threads)
    on a bounded sample of the bench workload: the global grid's full x and y extent (same pencil length and
    plane size as the real run) and a CPU_WINDOW_NZ-plane periodic window in z; one step = one RK3 step with the
    workload's forcing.  Returns (value, seconds per step, threads, description)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import numpy as np
    import pc_oracle
    from pencil_code_b200 import config as pcfg
    from pencil_code_b200 import synthetic
    n = args.n
    nzw = min(CPU_WINDOW_NZ, n)
    cfg, wname = workload_config(args, n, 1, 0, 1, 0)
    cfg = dict(cfg, nz=nzw)
    nthreads = host_threads()
    os.environ["OMP_NUM_THREADS"] = str(nthreads)
    outdir = tempfile.mkdtemp(prefix="pc_oracle_fast_")  # -march=native must be built on the box that runs it
    o = pc_oracle.Oracle(cfg, nthreads=nthreads, fast=True, outdir=outdir)
    f = synthetic.initial_condition(cfg)
    df = np.zeros_like(f)
    dt, t = pcfg.fixed_dt(cfg), 0.0
    for _ in range(max(warmup, 1)):  # page faults, thread pool
        dt, t, _ = o.time_step(f, df, False, dt, t)
    t0 = time.perf_counter()
    for _ in range(steps):
        dt, t, _ = o.time_step(f, df, False, dt, t)
    el = time.perf_counter() - t0
    desc = ("%d x %d x %d z-window (periodic) of the %d^3 grid, %d variables, same physics, %d RK3 steps in %.1f s on %d "
            "threads; CPU restatement of the reference (oracle, g++ -O3 -march=native -fopenmp) -- the Fortran code "
            "cannot be built here (no gfortran/MPI)" % (n, n, nzw, n, cfg["nvar"], steps, el, nthreads))
    return n * n * nzw * 3 * steps / el, el / max(steps, 1), nthreads, desc


def workload_config(args, n, npy, ipy, npz, ipz):
    from pencil_code_b200 import config as pcfg
    if args.workload == "iso_mhd":
        cfg = pcfg.make_config(n, entropy=False, magnetic=True, nu=5e-3, eta=5e-3, cdt=0.4, cdtv=0.3,
                               nprocy=npy, ipy=ipy, nprocz=npz, ipz=ipz)
        wname = "isothermal forced MHD turbulence %d^3 (lnrho, uu, aa: 7 vars, FP64)" % n
    elif args.workload == "hydro_hyper3":
        cfg = pcfg.config_hydro_hyper3(n, nprocy=npy, ipy=ipy, nprocz=npz, ipz=ipz)
        wname = "hydro-only forced turbulence %d^3 with 6th-order hyperdiffusion (4 vars, FP64)" % n
    else:
        cfg = pcfg.config_mhd_entropy(n, nprocy=npy, ipy=ipy, nprocz=npz, ipz=ipz)
        wname = "periodic MHD with entropy %d^3 (8 vars, FP64)" % n
    return cfg, wname


def config_dict(args, world, wname, nvar, nbytes_f):
    """the `config` object of the JSON line -- identical for the b200 arm and the reference arm"""
    from pencil_code_b200 import decomp
    npy, npz = decomp.choose_decomposition(world, args.decomp)
    return {"workload": wname + ", helical forcing, fixed dt, 2N-RK3",
            "decomposition": "1x%dx%d %s" % (npy, npz, "z-slabs" if npy == 1 else "yz-pencils"), "nvar": nvar,
            "l2": "inputs larger than L2 (%.1f GB per f buffer per GPU)" % (nbytes_f / 1e9)}


def run_reference(args, rank):
    """--impl reference: the reference's CPU implementation of the path on the box's host cores.  The Fortran code
    cannot be compiled in this image (no gfortran/MPI), so this times the CPU oracle (restatement, kind "port") with
    all host threads; every step is one RK3 step of a bounded sample of the workload (cpu_oracle_throughput)."""
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    val, s_per_step, cores, desc = cpu_oracle_throughput(args, max(args.steps, 1), max(args.warmup, 1))
    n = args.n
    cfg, wname = workload_config(args, n, 1, 0, 1, 0)
    import numpy as np
    from pencil_code_b200 import decomp
    npy, npz = decomp.choose_decomposition(world, args.decomp)
    nbytes_f = (n + 6) * (n // npy + 6) * (n // npz + 6) * cfg["nvar"] * 8
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": s_per_step * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_dict(args, world, wname, cfg["nvar"], nbytes_f),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--size", "--n", dest="n", type=int, default=512, help="global grid n^3 (use --size under torchrun: --n is ambiguous there)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--e2e-it1", type=int, default=10, help="steps per asynchronous snapshot copy of the second e2e figure (0 = skip)")
    ap.add_argument("--kernel-variant", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="mhd_entropy", choices=["mhd_entropy", "iso_mhd", "hydro_hyper3"],
                    help="mhd_entropy: BASELINE configs[2]/[3] (headline, default 512^3); iso_mhd: configs[1] (use --size 128); "
                         "hydro_hyper3: configs[4] (use --size 256)")
    ap.add_argument("--halo", default="p2p", choices=["p2p", "nccl"],
                    help="z-halo transport: copy-engine puts into the neighbours' ghost planes (CUDA IPC) or NCCL send/recv")
    ap.add_argument("--decomp", default="slab", choices=["slab", "pencil"],
                    help="slab: 1x1xN z-slabs (config 3); pencil: squarest nprocy x nprocz yz-pencils (config 4)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3  # timing rule: W >= 3

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            sys.exit("bench.py --gpus %d must be launched with torchrun (one rank per GPU)" % args.gpus)

    import numpy as np
    import torch
    import torch.distributed as dist
    from pencil_code_b200 import config as pcfg
    from pencil_code_b200 import decomp, synthetic
    from pencil_code_b200.gpu import GPU, nccl_unique_id

    if not torch.cuda.is_available():
        sys.exit("bench.py: no CUDA device; the gpu_b200 module has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    nccl = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group(backend="nccl", device_id=dev)
        ids = [nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        nccl = (ids[0], rank, world)

    n = args.n
    npy, npz = decomp.choose_decomposition(world, args.decomp)
    ipy, ipz = decomp.rank_position(rank, npy, npz)
    cfg, wname = workload_config(args, n, npy, ipy, npz, ipz)
    cfg["device"] = local_rank
    cfg["kernel_variant"] = args.kernel_variant
    nvar = cfg["nvar"]
    dt = pcfg.fixed_dt(cfg)
    cells_global = n ** 3
    shape = synthetic.farray_shape(cfg)
    nbytes_f = int(np.prod(shape)) * 8

    # host f-array in pinned memory (what the Fortran driver owns)
    fh_t = torch.empty(shape, dtype=torch.float64, pin_memory=True)
    fh = fh_t.numpy()
    fh[...] = 0.0

    g = GPU(cfg, nccl=nccl)
    stream = torch.cuda.Stream(device=dev)  # a real (non-default) stream so that torch events see the library's work
    torch.cuda.set_stream(stream)
    g.L.pcb200_set_stream(stream.cuda_stream)
    g.register_GPU()
    g.initialize_GPU(fh, nt=0)
    if world > 1 and args.halo == "p2p":
        blobs = [None] * world  # CUDA IPC handles of every rank's f buffers: z halos become copy-engine puts over NVLink
        dist.all_gather_object(blobs, g.p2p_export())
        ok = torch.tensor([1 if g.p2p_connect(blobs) else 0], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok.item()) == 1:
            g.p2p_enable()
        else:
            args.halo = "nccl"  # some rank cannot map its neighbour (GPU not visible to the process): all ranks stay on NCCL
    g.init_synthetic(12345)  # synthetic forced-turbulence-like state generated on the device
    g.synchronize()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    from pencil_code_b200.gpu import HelicalForcing
    forcing = HelicalForcing(pcfg.kvectors())  # iforce='helical', force=0.07, relhel=1; same stream on every rank

    def step(t):
        return g.time_step(fh, dt, False, t, forcing=forcing.draw(dt))

    t = 0.0
    for _ in range(args.warmup):
        _, t = step(t)
    # ---- device-resident timing
    g.enable_kernel_timing(True)
    g.kernel_timing(reset=True)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    l0 = g.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        _, t = step(t)
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    launches = g.launch_count() - l0
    clocks = sampler.stop() if rank == 0 else None
    kms, klaunch = g.kernel_timing(reset=True)
    g.enable_kernel_timing(False)
    tmax = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms = float(tmax.item())
    value = cells_global * 3 * args.steps / (ms * 1e-3)

    # ---- end-to-end through the module-GPU calls with host buffers (H2D + 3 substeps + D2H per step)
    g.copy_farray_from_GPU(fh)
    barrier()
    t_e2e0 = time.perf_counter()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record(stream)
    for _ in range(args.e2e_steps):
        g.load_farray_to_GPU(fh)
        _, t = step(t)
        g.copy_farray_from_GPU(fh)
    e3.record(stream)
    barrier()
    ms_e2e = max(e2.elapsed_time(e3), (time.perf_counter() - t_e2e0) * 1e3)
    tm2 = torch.tensor([ms_e2e], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tm2, op=dist.ReduceOp.MAX)
    ms_e2e = float(tm2.item())
    e2e_value = cells_global * 3 * args.e2e_steps / (ms_e2e * 1e-3)
    finite = bool(np.isfinite(fh[:, 3:-3, 3:-3, 3:-3]).all())

    # ---- the same at the reference's real cadence: the f-array stays on the device and comes back every it1-th step
    # only (run.f90: lout steps; samples/*/run.in it1=10), as an asynchronous snapshot copy that overlaps the next steps
    # (pcb200_copy_farray_async: device snapshot + D2H on a copy stream, gpu_astaroth.f90:390-412 nowait_ protocol)
    it1 = args.e2e_it1
    e2e_cadence = None
    if it1 > 0:
        g.synchronize()
        barrier()
        t_c0 = time.perf_counter()
        g.copy_farray_from_GPU_async(fh)
        for _ in range(it1):
            _, t = step(t)
        g.copy_farray_wait()
        g.synchronize()
        ms_c = (time.perf_counter() - t_c0) * 1e3
        tm3 = torch.tensor([ms_c], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tm3, op=dist.ReduceOp.MAX)
        ms_c = float(tm3.item())
        e2e_cadence = {"value": cells_global * 3 * it1 / (ms_c * 1e-3), "unit": UNIT, "it1": it1, "ms_per_step": ms_c / it1,
                       "h2d_bytes_per_step": 0, "d2h_bytes_per_step": nbytes_f * world / it1,
                       "path": "f resident on the device; copy_farray_from_GPU(nowait) once per it1 steps, overlapped with the steps"}

    # ---- roofline of the dominant kernel (fused rhs+rk), measured live with CUDA events
    peak, peak_src = hbm_peak()
    cells_local = cfg["nx"] * cfg["ny"] * cfg["nz"]
    # 80*nvar B per cell per full step.  With N>1 a substep is 3 launches of the fused kernel
    # (interior + 2 boundary slabs), so bytes and time are summed over the timed region rather than per launch.
    substeps = 3 * args.steps
    bytes_total = cells_local * 80.0 * nvar / 3.0 * substeps
    bytes_per_launch = bytes_total / max(klaunch, 1)
    avg_kernel_ms = kms / max(klaunch, 1)
    achieved = bytes_total / (kms * 1e-3) / 1e9 if kms > 0 else 0.0
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": None, "kernel": "rhs_update (fused RHS + 2N-RK3)", "avg_kernel_ms": avg_kernel_ms,
                "kernel_launches": klaunch, "kernel_share_of_step": kms / ms if ms > 0 else None,
                "algorithmic_bytes_per_launch": bytes_per_launch, "peak_source": peak_src,
                "frac_of_nominal_8TBs": achieved / 8000.0}
    traffic_file = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(traffic_file) and args.workload == "mhd_entropy" and n == 512 and world == 1:
        try:
            roofline["traffic"] = json.load(open(traffic_file)).get("dram_bytes_per_launch_512")
        except Exception:
            pass

    # ---- decomposition-invariance checksum: global diagnostics of the state after warmup + steps + e2e steps,
    # reduced on the device over all ranks (sums: ncclAllReduce) -- must agree between N = 1, 2, 4, 8 to ~1e-12
    g.time_step(fh, dt, False, t, forcing=forcing.draw(dt), diagnostics=True)
    red = g.get_gpu_reduced_vars()
    checksum = {k: red[k] for k in ("urms", "umax", "rhom", "brms", "jrms", "abm", "ssm") if k in red}
    checksum["after_steps"] = args.warmup + args.steps + args.e2e_steps + max(args.e2e_it1, 0) + 1

    if world > 1:
        barrier()  # nobody frees ghost planes a neighbour may still be writing to
    g.finalize_GPU()

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        v, s_per_step, cores, desc = cpu_oracle_throughput(args, 16)
        cpu_baseline = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc}
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": config_dict(args, world, wname, nvar, nbytes_f),
            "z_halo": ("none (1 GPU)" if world == 1 else ("peer puts over NVLink (copy engines)" + (", y faces as strided peer puts" if npy > 1 else "") if args.halo == "p2p" else "NCCL send/recv")),
            "kernel_variant": args.kernel_variant, "finite": finite, "checksum": checksum,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": nbytes_f * world,
                    "d2h_bytes_per_step": nbytes_f * world, "steps": args.e2e_steps,
                    "ms_per_step": ms_e2e / args.e2e_steps,
                    "path": "load_farray_to_GPU + 3 x rhs_GPU + copy_farray_from_GPU per step, pinned host f-array"},
            "e2e_at_output_cadence": e2e_cadence,
            "gpu_launches": launches,
            "clocks": clocks,
            "roofline": roofline,
            "cpu_baseline": cpu_baseline,
            "us_per_point_per_step": ms * 1e3 / args.steps / cells_global,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
