"""Host-side mirror of Pencil Code's Fortran `module GPU` (src/gpu.h:1-12) over libpcb200.so.

Same public names, argument meaning and error behaviour as src/gpu_astaroth.f90:121-518:
initialize_GPU(f) retains the host f-array, rhs_GPU(f, itsub) performs one complete substep on
the device, copy_farray_from_GPU / load_farray_to_GPU move the f-array, finalize_GPU frees the
device.  Errors abort the process (the reference's fatal_error convention) -- there is no CPU
fallback: if libpcb200.so is missing this module raises at import of the library.

All calls go through the C ABI declared in include/pcb200.h with HOST numpy buffers, exactly
what the Fortran module would pass.
"""
import ctypes as C
import os

import numpy as np

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "libpcb200.so")

N_GPU_REDUCED_VARS = 20  # gpu.h:12
REDUCED_NAMES = ("urms", "umax", "rhom", "brms", "bmax", "jrms", "jmax", "abm", "jbm", "oum", "epsK", "ssm",
                 "dtu", "dtc", "dtb", "dtnu", "dt1_max", "dt", "dtchi", "spare1")
BC_CODES = {"p": 0, "s": 1, "a": 2, "a2": 3, "c1": 4, "cT": 5}  # include/pcb200.h PCB200_BC_*


class Config(C.Structure):
    """struct pcb200_config (include/pcb200.h)."""
    _fields_ = [
        ("nx", C.c_int), ("ny", C.c_int), ("nz", C.c_int),
        ("nprocx", C.c_int), ("nprocy", C.c_int), ("nprocz", C.c_int),
        ("ipx", C.c_int), ("ipy", C.c_int), ("ipz", C.c_int),
        ("nvar", C.c_int),
        ("iux", C.c_int), ("ilnrho", C.c_int), ("iss", C.c_int), ("iax", C.c_int),
        ("dx", C.c_double), ("dy", C.c_double), ("dz", C.c_double),
        ("x0", C.c_double), ("y0", C.c_double), ("z0", C.c_double),
        ("cs20", C.c_double), ("gamma", C.c_double), ("cp", C.c_double), ("lnrho0", C.c_double),
        ("mu0", C.c_double),
        ("nu", C.c_double), ("nu_hyper3", C.c_double),
        ("lvisc_nu_const", C.c_int), ("lvisc_hyper3_simplified", C.c_int), ("lvisc_hyper3_nu_const", C.c_int),
        ("diffrho_hyper3", C.c_double), ("ldiff_hyper3lnrho", C.c_int),
        ("chi", C.c_double), ("lheatc_chiconst", C.c_int), ("lviscosity_heat", C.c_int),
        ("eta", C.c_double), ("lresi_eta_const", C.c_int), ("lohmic_heat", C.c_int), ("llorentzforce", C.c_int),
        ("lupw_uu", C.c_int), ("lupw_lnrho", C.c_int), ("lupw_ss", C.c_int), ("lupw_aa", C.c_int),
        ("cdt", C.c_double), ("cdtv", C.c_double), ("cdtv2", C.c_double), ("cdtv3", C.c_double),
        ("dtmax", C.c_double),
        ("lperi", C.c_int * 3),
        ("device", C.c_int), ("kernel_variant", C.c_int),
        ("bcz", (C.c_int * 8) * 2),
        ("gravz", C.c_double), ("Omega", C.c_double),
        ("lheatc_Kprof", C.c_int),
        ("hcond0", C.c_double), ("hcond1", C.c_double), ("hcond2", C.c_double),
        ("widthss", C.c_double), ("z1", C.c_double), ("z2", C.c_double),
        ("cool", C.c_double), ("wcool", C.c_double), ("cs2cool", C.c_double), ("ztop", C.c_double),
        ("FbotKbot", C.c_double), ("FtopKtop", C.c_double), ("cs2bot", C.c_double), ("cs2top", C.c_double),
    ]


class DriverState(C.Structure):
    """struct pcb200_driver_state (include/pcb200.h)."""
    _fields_ = [("t", C.c_double), ("dt", C.c_double), ("it", C.c_int), ("ldt", C.c_int),
                ("lforcing", C.c_int), ("force", C.c_double), ("relhel", C.c_double), ("kav", C.c_double),
                ("nk", C.c_int), ("kkx", C.POINTER(C.c_double)), ("kky", C.POINTER(C.c_double)),
                ("kkz", C.POINTER(C.c_double)), ("rng_state", C.c_ulonglong), ("it1", C.c_int)]


# every symbol include/pcb200.h declares (checked by tests/test_abi.py on CPU)
ABI_SYMBOLS = (
    "pcb200_configure", "pcb200_reconfigure", "pcb200_sizeof_config", "pcb200_set_dt", "pcb200_get_dt1_max", "pcb200_get_dt",
    "pcb200_set_forcing_hel", "pcb200_forcing_coefs_hel", "pcb200_request_diagnostics", "pcb200_nccl_get_unique_id", "pcb200_nccl_init",
    "pcb200_set_stream", "pcb200_synchronize", "pcb200_init_synthetic", "pcb200_launch_count", "pcb200_host_sync_count",
    "pcb200_kernel_timing", "pcb200_enable_kernel_timing",
    "pcb200_p2p_blob_bytes", "pcb200_p2p_export", "pcb200_p2p_connect", "pcb200_p2p_enable",
    "pcb200_copy_farray_async", "pcb200_copy_farray_wait", "pcb200_write_snapshot", "pcb200_read_snapshot",
    "initialize_gpu_c", "register_gpu_c", "finalize_gpu_c", "rhs_gpu_c", "before_boundary_gpu_c",
    "update_after_substep_gpu_c", "load_farray_c", "copy_farray_c", "gpu_prepare_for_first_substep_c",
    "get_farray_ptr_gpu_c", "reload_gpu_config_c", "update_on_gpu_scal_by_name_c", "update_on_gpu_vec_by_name_c",
    "update_on_gpu_arr_by_name_c", "update_on_gpu_scal_by_ind_c", "update_on_gpu_vec_by_ind_c",
    "update_on_gpu_arr_by_ind_c", "pos_real_ptr_c", "split_update_gpu_c", "radtransfer_gpu_c", "test_bcs_c",
    "get_gpu_reduced_vars_c", "pcb200_driver_run", "pcb200_test_bcs", "pcb200_dt1_max_noalloc",
)
# the reference binds these with a trailing underscore (FTNIZE, src/headers_c.h:31-37)
FTN_SYMBOLS = tuple(s + "_" for s in ABI_SYMBOLS if s.endswith("_c"))

_lib = None


def load_library():
    """dlopen libpcb200.so; fails loudly when the CUDA extension has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "libpcb200.so not found at %s -- build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(the gpu_b200 module has no CPU fallback)" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int)
    L.pcb200_configure.argtypes = [C.POINTER(Config)]
    L.pcb200_sizeof_config.restype = C.c_int
    L.pcb200_set_dt.argtypes = [dp, ip]
    L.pcb200_get_dt1_max.argtypes = [dp]
    L.pcb200_get_dt.argtypes = [dp]
    L.pcb200_set_forcing_hel.argtypes = [ip, dp, dp, dp, dp, dp, dp, dp]
    L.pcb200_nccl_get_unique_id.argtypes = [C.c_char_p]
    L.pcb200_nccl_init.argtypes = [C.c_char_p, ip, ip]
    L.pcb200_set_stream.argtypes = [C.c_void_p]
    L.pcb200_p2p_blob_bytes.restype = C.c_int
    L.pcb200_p2p_export.argtypes = [C.c_char_p]
    L.pcb200_p2p_connect.argtypes = [C.c_char_p, ip]
    L.pcb200_p2p_connect.restype = C.c_int
    L.pcb200_write_snapshot.argtypes = [C.c_char_p, dp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, dp, dp, dp,
                                        C.c_double, C.c_double, C.c_double]
    L.pcb200_write_snapshot.restype = C.c_int
    L.pcb200_read_snapshot.argtypes = [C.c_char_p, dp, C.c_int, C.c_int, C.c_int, C.c_int, dp, dp, dp, dp, dp]
    L.pcb200_read_snapshot.restype = C.c_int
    L.pcb200_init_synthetic.argtypes = [ip]
    L.pcb200_launch_count.restype = C.c_longlong
    L.pcb200_host_sync_count.restype = C.c_longlong
    L.pcb200_kernel_timing.argtypes = [dp, C.POINTER(C.c_longlong), ip]
    L.pcb200_enable_kernel_timing.argtypes = [ip]
    L.initialize_gpu_c.argtypes = [dp, ip, dp, ip, ip, ip, ip]
    L.rhs_gpu_c.argtypes = [ip, dp]
    L.before_boundary_gpu_c.argtypes = [ip, ip, dp, ip]
    L.copy_farray_c.argtypes = [dp]
    L.pcb200_copy_farray_async.argtypes = [dp]
    L.gpu_prepare_for_first_substep_c.argtypes = [dp]
    L.get_farray_ptr_gpu_c.argtypes = [C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
    L.get_gpu_reduced_vars_c.argtypes = [dp]
    L.update_on_gpu_scal_by_name_c.argtypes = [C.c_char_p, dp]
    L.update_on_gpu_scal_by_name_c.restype = C.c_int
    L.split_update_gpu_c.argtypes = [dp]
    L.pcb200_driver_run.argtypes = [C.POINTER(DriverState), C.c_int, dp, C.c_int]
    L.pcb200_driver_run.restype = C.c_int
    L.pcb200_test_bcs.restype = C.c_longlong
    L.pcb200_dt1_max_noalloc.argtypes = [C.c_void_p, C.c_void_p, ip]
    L.pcb200_dt1_max_noalloc.restype = C.c_int
    L.pcb200_reconfigure.argtypes = [C.POINTER(Config)]
    L.pos_real_ptr_c.argtypes = [C.POINTER(C.c_void_p), ip]
    L.pos_real_ptr_c.restype = C.c_void_p
    if L.pcb200_sizeof_config() != C.sizeof(Config):
        raise RuntimeError("pcb200_config layout mismatch between gpu.py and libpcb200.so")
    _lib = L
    return L


def _dp(a):
    if a.dtype != np.float64 or not a.flags["C_CONTIGUOUS"]:
        raise ValueError("f-array must be C-contiguous float64 of shape (mfarray, mz, my, mx)")
    return a.ctypes.data_as(C.POINTER(C.c_double))


def make_struct(cfg):
    c = Config()
    for name, _ in Config._fields_:
        if name in ("lperi", "bcz"):
            continue
        if name in cfg:
            setattr(c, name, cfg[name])
    closed_z = False
    for j, spec in enumerate(cfg.get("bcz", ())):  # 'bot:top' per variable as in run.in (boundcond.f90:175-260)
        bot, _, top = spec.partition(":")
        c.bcz[0][j] = BC_CODES[bot]
        c.bcz[1][j] = BC_CODES[top or bot]
        closed_z = closed_z or c.bcz[0][j] != 0 or c.bcz[1][j] != 0
    lp = cfg.get("lperi", (1, 1, 0 if closed_z else 1))
    c.lperi[0], c.lperi[1], c.lperi[2] = [int(v) for v in lp]
    return c


class GPU:
    """`module GPU` for one rank.  Public method names are those of src/gpu.h."""

    lgpu = True  # ! CPARAM logical, parameter :: lgpu = .true.
    n_gpu_reduced_vars = N_GPU_REDUCED_VARS

    def __init__(self, cfg, nccl=None):
        """cfg: dict from pencil_code_b200.config; nccl=(id_bytes, rank, nranks) for multi-GPU."""
        self.L = load_library()
        self.cfg = dict(cfg)
        self._c = make_struct(cfg)
        self.L.pcb200_configure(C.byref(self._c))
        if nccl is not None:
            idb, rank, nranks = nccl
            self.L.pcb200_nccl_init(idb, C.byref(C.c_int(rank)), C.byref(C.c_int(nranks)))
        self._f = None
        self.ltest_bcs = False
        self.ltest_rhs = False
        self.it_test_rhs = 1

    # --- gpu.h public interface -------------------------------------------------
    def register_GPU(self):
        self.L.register_gpu_c()

    def initialize_GPU(self, f, t=0.0, nt=1):
        """gpu_astaroth.f90:192-245: initialize_gpu_c retains f; `if (nt>0) load_farray_to_GPU(f)`."""
        self._f = f
        ct, cnt, z = C.c_double(t), C.c_int(nt), C.c_int(0)
        comm = C.c_int(0)
        self.L.initialize_gpu_c(_dp(f), C.byref(comm), C.byref(ct), C.byref(cnt), C.byref(z), C.byref(z), C.byref(z))
        if nt > 0:
            self.load_farray_to_GPU(f)

    def copy_farray_from_GPU_async(self, f):
        """start a snapshot + asynchronous device-to-host copy of the current f; pair with copy_farray_wait()"""
        self.L.pcb200_copy_farray_async(_dp(f))

    def copy_farray_wait(self):
        self.L.pcb200_copy_farray_wait()

    # --- z-halo over NVLink peer memory (include/pcb200.h: pcb200_p2p_*) ---------
    def p2p_export(self):
        """IPC blob of this rank (call after initialize_GPU); all-gather the blobs of all ranks in rank order."""
        buf = C.create_string_buffer(self.L.pcb200_p2p_blob_bytes())
        self.L.pcb200_p2p_export(buf)
        return bytes(buf.raw)

    def p2p_connect(self, blobs):
        """blobs: list of every rank's p2p_export() in rank order.  Returns True when this rank mapped both z
        neighbours; the caller reduces the result over all ranks and then calls p2p_enable() on all of them."""
        table = b"".join(blobs)
        return self.L.pcb200_p2p_connect(table, C.byref(C.c_int(len(blobs)))) == 0

    def p2p_enable(self):
        """switch the z exchange from NCCL to peer puts (all ranks, after all of them connected)"""
        self.L.pcb200_p2p_enable()

    def finalize_GPU(self):
        self.L.finalize_gpu_c()
        self._f = None

    def rhs_GPU(self, f, isubstep, t=0.0):
        ci, ct = C.c_int(isubstep), C.c_double(t)
        self.L.rhs_gpu_c(C.byref(ci), C.byref(ct))

    def before_boundary_gpu(self, f, lrmv, isubstep, t):
        a, b, c, d = C.c_int(int(lrmv)), C.c_int(isubstep), C.c_double(t), C.c_int(0)
        self.L.before_boundary_gpu_c(C.byref(a), C.byref(b), C.byref(c), C.byref(d))

    def update_after_substep_gpu(self):
        self.L.update_after_substep_gpu_c()

    def gpu_prepare_for_first_substep(self, t=0.0):
        self.L.gpu_prepare_for_first_substep_c(C.byref(C.c_double(t)))

    def copy_farray_from_GPU(self, f, nowait_=False):
        self.L.copy_farray_c(_dp(f))

    def load_farray_to_GPU(self, f):
        if self._f is None or f.ctypes.data != self._f.ctypes.data:
            raise ValueError("load_farray_to_GPU uses the f-array retained by initialize_GPU (gpu_astaroth.f90:419)")
        self.L.load_farray_c()

    def reload_GPU_config(self):
        self.L.reload_gpu_config_c()

    def update_on_gpu(self, index, varname, value):
        v = C.c_double(value)
        idx = self.L.update_on_gpu_scal_by_name_c(varname.encode() + b"\0", C.byref(v))
        if idx < 0:
            raise KeyError('update_on_gpu: variable "%s" not found' % varname)
        if varname in self.cfg:
            self.cfg[varname] = value  # a later reconfigure() sends the complete struct
        return idx

    def get_gpu_reduced_vars(self):
        dst = np.zeros(N_GPU_REDUCED_VARS)
        self.L.get_gpu_reduced_vars_c(_dp(dst))
        return dict(zip(REDUCED_NAMES, dst))

    def split_update_gpu(self, f):
        self.L.split_update_gpu_c(_dp(f))

    def test_gpu_bcs(self):
        """ltest_bcs (gpu_astaroth.f90 test_gpu_bcs): exits the process with EXIT_FAILURE when the check fails"""
        self.L.test_bcs_c()

    def test_bcs_count(self):
        """number of ghost values that differ from the CPU boundary conditions (0 = passed); never exits"""
        return int(self.L.pcb200_test_bcs())

    def reconfigure(self, **changes):
        """RELOAD path of gpu_b200.f90: new physics parameters while initialized (pcb200_reconfigure + reload)"""
        self.cfg.update(changes)
        self._c = make_struct(self.cfg)
        self.L.pcb200_reconfigure(C.byref(self._c))
        self.L.reload_gpu_config_c()

    def pos_real_ptr(self, base, ind):
        b = C.c_void_p(base)
        return self.L.pos_real_ptr_c(C.byref(b), C.byref(C.c_int(ind)))

    def driver_run(self, nt, dt, ldt=False, t=0.0, it=0, forcing=None, it1=0, seed=2024):
        """pcb200_driver_run: the C++ mirror of run.f90/timestep.f90's loop (forcing: HelicalForcing or None).
        Returns (t, dt, it, rows) with rows = [(it, t, {reduced vars})] of the it1-th steps."""
        st = DriverState()
        st.t, st.dt, st.it, st.ldt = t, dt, it, int(ldt)
        st.lforcing = 0 if forcing is None else 1
        st.it1 = it1
        keep = None
        if forcing is not None:
            n = len(forcing.kvecs)
            arr = [(C.c_double * n)(*[k[i] for k in forcing.kvecs]) for i in range(3)]
            keep = arr
            st.force, st.relhel, st.kav, st.nk = forcing.force, forcing.relhel, forcing.kav, n
            st.kkx, st.kky, st.kkz = [C.cast(a, C.POINTER(C.c_double)) for a in arr]
            st.rng_state = seed & 0xFFFFFFFFFFFFFFFF
        rowlen = N_GPU_REDUCED_VARS + 2
        max_rows = nt + 1
        buf = (C.c_double * (rowlen * max_rows))()
        nrows = self.L.pcb200_driver_run(C.byref(st), nt, buf, max_rows)
        rows = []
        for r in range(nrows):
            v = buf[r * rowlen:(r + 1) * rowlen]
            rows.append((int(v[0]), v[1], dict(zip(REDUCED_NAMES, v[2:]))))
        del keep
        return st.t, st.dt, st.it, rows

    def get_farray_ptr_gpu(self):
        a, b = C.c_void_p(), C.c_void_p()
        self.L.get_farray_ptr_gpu_c(C.byref(a), C.byref(b))
        return a.value, b.value

    # --- timestep exchange / forcing / diagnostics (include/pcb200.h part 2) ----
    def set_dt(self, dt, ldt):
        self.L.pcb200_set_dt(C.byref(C.c_double(dt)), C.byref(C.c_int(int(ldt))))

    def get_dt(self):
        d = C.c_double()
        self.L.pcb200_get_dt(C.byref(d))
        return d.value

    def get_dt1_max(self):
        d = C.c_double()
        self.L.pcb200_get_dt1_max(C.byref(d))
        return d.value

    def set_forcing_hel(self, frc):
        """frc: object with coef1, coef2, fda, kk (3-sequences), phase, fact, k1_ff; None disables."""
        a3 = C.c_double * 3
        if frc is None:
            z = a3(0, 0, 0)
            d = C.c_double(0)
            self.L.pcb200_set_forcing_hel(C.byref(C.c_int(0)), z, z, z, z, C.byref(d), C.byref(d), C.byref(d))
            return
        self.L.pcb200_set_forcing_hel(C.byref(C.c_int(1)), a3(*frc.coef1), a3(*frc.coef2), a3(*frc.fda), a3(*frc.kk),
                                      C.byref(C.c_double(frc.phase)), C.byref(C.c_double(frc.fact)),
                                      C.byref(C.c_double(frc.k1_ff)))

    def request_diagnostics(self):
        self.L.pcb200_request_diagnostics()

    def synchronize(self):
        self.L.pcb200_synchronize()

    def init_synthetic(self, seed=12345):
        self.L.pcb200_init_synthetic(C.byref(C.c_int(seed)))

    def launch_count(self):
        return int(self.L.pcb200_launch_count())

    def host_sync_count(self):
        return int(self.L.pcb200_host_sync_count())

    def enable_kernel_timing(self, on=True):
        self.L.pcb200_enable_kernel_timing(C.byref(C.c_int(int(on))))

    def kernel_timing(self, reset=True):
        ms, n = C.c_double(), C.c_longlong()
        self.L.pcb200_kernel_timing(C.byref(ms), C.byref(n), C.byref(C.c_int(int(reset))))
        return ms.value, n.value

    # --- the reference driver's call order for one full step (timestep.f90:206-270) ----
    def time_step(self, f, dt, ldt, t, forcing=None, diagnostics=False):
        """advance_substeps with lgpu=T: for itsub=1..3 { before_boundary_gpu; rhs_gpu; t+=dtsub }.
        Returns (dt, t)."""
        alpha = (0.0, -5.0 / 9.0, -153.0 / 128.0)
        beta = (1.0 / 3.0, 15.0 / 16.0, 8.0 / 15.0)
        if diagnostics:
            self.request_diagnostics()
        self.set_dt(dt, ldt)
        ds = 0.0
        for itsub in (1, 2, 3):
            ds = 0.0 if itsub == 1 else alpha[itsub - 1] * ds
            if itsub == 3:
                self.set_forcing_hel(forcing)
            self.before_boundary_gpu(f, False, itsub, t)
            self.rhs_GPU(f, itsub, t)
            if itsub == 1 and ldt:
                dt = self.get_dt()
            ds += 1.0
            t += ds * (dt * beta[itsub - 1])
            self.update_after_substep_gpu()
        return dt, t


class HelicalForcing:
    """Per-step helical forcing coefficients (forcing.f90:1773-2001 fconst_coefs_hel, draw order :1804-1811)
    from the library's own host routine pcb200_forcing_coefs_hel and a splitmix64 stream (seed 2024,
    force=0.07, relhel=1, all integer k with 4.5 <= |k| < 5.5)."""

    class Coefs:
        pass

    def __init__(self, kvecs, force=0.07, relhel=1.0, seed=2024):
        self.L = load_library()
        dp = C.POINTER(C.c_double)
        self.L.pcb200_forcing_coefs_hel.argtypes = [C.c_double] * 5 + [dp, C.c_double, C.c_double, dp, dp, dp, dp, dp]
        self.kvecs, self.force, self.relhel = list(kvecs), force, relhel
        self.kav = sum((k[0] ** 2 + k[1] ** 2 + k[2] ** 2) ** 0.5 for k in self.kvecs) / len(self.kvecs)
        self.state = seed & 0xFFFFFFFFFFFFFFFF

    def _uniform(self):
        m = 0xFFFFFFFFFFFFFFFF
        self.state = (self.state + 0x9E3779B97F4A7C15) & m
        z = self.state
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & m
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & m
        z ^= z >> 31
        return (z >> 11) * (1.0 / 9007199254740992.0)

    def draw(self, dt):
        import math
        f1, f2, f3 = self._uniform(), self._uniform(), self._uniform()
        phase = math.pi * (2 * f1 - 1.0)
        ik = int(len(self.kvecs) * (0.9999 * f2) + 1) - 1
        phi = f3 * 2 * math.pi
        a3 = C.c_double * 3
        c1, c2, fda, kk, fact = a3(), a3(), a3(), a3(), C.c_double()
        self.L.pcb200_forcing_coefs_hel(self.force, self.relhel, self.kav, 1.0, dt, a3(*self.kvecs[ik]), phase, phi,
                                        c1, c2, fda, kk, C.byref(fact))
        o = HelicalForcing.Coefs()
        o.coef1, o.coef2, o.fda, o.kk = list(c1), list(c2), list(fda), list(kk)
        o.phase, o.fact, o.k1_ff = phase, fact.value, 1.0
        return o


def nccl_unique_id():
    L = load_library()
    buf = C.create_string_buffer(128)
    L.pcb200_nccl_get_unique_id(buf)
    return buf.raw
