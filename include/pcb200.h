/*
 * pcb200.h -- C ABI of libpcb200.so, the B200-native GPU module for Pencil Code
 * (selected with GPU=gpu_b200 in src/Makefile.local).
 *
 * Part 1 are the entry points the reference's Fortran `module GPU`
 * (src/gpu.h:1-12; src/gpu_astaroth.f90:21-53 `external *_c`) binds; names,
 * argument order and by-reference convention are those of
 * src/gpu_astaroth_ansi.c / src/astaroth/gpu_astaroth.cc.  Every symbol is
 * exported twice: plain (for ISO_C_BINDING `bind(C,name=...)`) and with a
 * trailing underscore (the reference's FTNIZE() convention,
 * src/headers_c.h:31-37).
 *
 * Part 2 replaces what the reference obtains at compile time from
 * cparam_c.h (scripts/cparam2c) and at run time from the pushpars2c address
 * tables (src/global_pushpars.f90:26-101 and each module's pushpars2c): one
 * plain struct handed over before initialize_gpu_c.
 *
 * Error convention (as gpu_astaroth.cc:1595-1597,2074-2078): no status codes;
 * an unsupported configuration or a CUDA/NCCL error prints to stderr and
 * aborts the process.  There is NO CPU fallback.
 */
#ifndef PCB200_H
#define PCB200_H

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ */
/* Part 2: configuration                                               */
/* ------------------------------------------------------------------ */
typedef struct pcb200_config {
  /* local interior sizes per rank and decomposition (cparam.f90:54-75; cparam.local) */
  int nx, ny, nz;
  int nprocx, nprocy, nprocz; /* nprocx must be 1 */
  int ipx, ipy, ipz;          /* this rank's position */
  int nvar;                   /* mvar */
  int iux, ilnrho, iss, iax;  /* 0-based slots (Fortran index-1); -1 = module absent */
  double dx, dy, dz;          /* equidistant Cartesian (grid.f90:123-149) */
  double x0, y0, z0;          /* x(l1), y(m1), z(n1) of THIS rank */
  /* eos_idealgas pushpars (eos_idealgas.f90:4589-4606) */
  double cs20, gamma, cp, lnrho0;
  double mu0; /* cdata (global_pushpars.f90: mu0) */
  /* viscosity.f90:3250-3255 */
  double nu, nu_hyper3;
  int lvisc_nu_const, lvisc_hyper3_simplified, lvisc_hyper3_nu_const;
  /* density.f90:4591-4599 */
  double diffrho_hyper3;
  int ldiff_hyper3lnrho;
  /* entropy.f90:8675-8816 */
  double chi;
  int lheatc_chiconst, lviscosity_heat;
  /* magnetic.f90:12063-12069 */
  double eta;
  int lresi_eta_const, lohmic_heat, llorentzforce;
  /* upwinding: hydro.f90:9367, density.f90:4591-4599, entropy.f90, magnetic.f90 */
  int lupw_uu, lupw_lnrho, lupw_ss, lupw_aa;
  /* cdata: cdt, cdtv, cdtv2, cdtv3 (global_pushpars.f90:27-32), dtmax */
  double cdt, cdtv, cdtv2, cdtv3;
  double dtmax;
  /* boundary conditions: x and y must be periodic; z is periodic (lperi[2]=1) or closed (lperi[2]=0, see bcz below) */
  int lperi[3];
  /* device selection: -1 = local rank modulo device count */
  int device;
  /* kernel selection: 0 = auto (10 on isotropic grids with even mx, 4 on anisotropic ones, 1 when mx is odd or the run is
     vertically stratified), 1 = global-load kernel (also the stratified one), 4 = TMA/shared-memory z-marching kernel with
     lead/trail z-mixed derivatives and per-direction weights, 5 = 4 specialised for dx=dy=dz (round-1 default, kept as an
     independent cross-check), 10 = all global traffic through the TMA engine, z histories and all-trail accumulators in
     tensor memory.  Other numbers (earlier lab variants 2, 3, 6-9) are refused. */
  int kernel_variant;
  /* ---- vertically stratified set-ups (samples/conv-slab); all zero = triply periodic box --------------------
     z boundary condition per [bottom/top][variable slot] when lperi[2]=0 (boundcond.f90:1421-1787, bcz12):
     PCB200_BC_S 's', PCB200_BC_A 'a', PCB200_BC_A2 'a2', PCB200_BC_C1 'c1' (entropy: constant heat flux,
     entropy_bcs.f90:93-235), PCB200_BC_CT 'cT' (entropy: constant temperature, eos_idealgas.f90:3112-3260).
     Requires nprocy = nprocz = 1 and selects kernel variant 1. */
  int bcz[2][8];
  double gravz;     /* gravity_simple.f90:521-523 gravz_profile='const'; addgravity :1083-1162 */
  double Omega;     /* hydro.f90:6339-6366 coriolis_cartesian (theta=0) */
  int lheatc_Kprof; /* iheatcond='K-profile' (entropy.f90:1176-1178, 5913-6151) */
  double hcond0, hcond1, hcond2, widthss, z1, z2; /* get_gravz_heatcond entropy.f90:7741-7769 */
  double cool, wcool, cs2cool, ztop;              /* cooling layer, entropy.f90:1323,1367,6761-6828 ('gaussian') */
  double FbotKbot, FtopKtop;                      /* entropy.f90:854-887 (for 'c1') */
  double cs2bot, cs2top;                          /* for 'cT' (density.f90:1605-1610, entropy.f90:818-838) */
} pcb200_config;

enum { PCB200_BC_P = 0, PCB200_BC_S = 1, PCB200_BC_A = 2, PCB200_BC_A2 = 3, PCB200_BC_C1 = 4, PCB200_BC_CT = 5 };

/* must be called before initialize_gpu_c; copies *cfg */
void pcb200_configure(const pcb200_config *cfg);
/* RELOAD: new physics parameters while initialized (grid, decomposition and variable slots must be unchanged;
   device and kernel_variant are kept); follow with reload_gpu_config_c.  Before initialize_gpu_c it equals
   pcb200_configure. */
void pcb200_reconfigure(const pcb200_config *cfg);
int pcb200_sizeof_config(void);

/* Timestep exchange with the host driver.  The Fortran module calls
   pcb200_set_dt before rhs_gpu_c when ldt=F, and pcb200_get_dt1_max after
   substep 1 when ldt=T to feed Sub::set_dt (sub.f90:4223-4254). */
void pcb200_set_dt(const double *dt, const int *ldt);
void pcb200_get_dt1_max(double *dt1_max);
void pcb200_get_dt(double *dt);

/* helical forcing coefficients for the step in progress (forcing.f90:1731-1746
   forcing_pars_hel; consumed after substep 3, astaroth/forcing.h:1-36).
   lforce=0 disables. */
void pcb200_set_forcing_hel(const int *lforce, const double coef1[3], const double coef2[3],
                            const double fda[3], const double kk[3], const double *phase,
                            const double *fact, const double *k1_ff);

/* host-side fconst_coefs_hel (forcing.f90:1773-2001, default branches: no shear, lscale_kvector_*=F,
   old_forcing_evector=F, cs0eff=cs0) for drivers that do not have the Fortran forcing module
   (pcb200_driver_run, bench.py); by value, not a Fortran binding */
void pcb200_forcing_coefs_hel(double force, double relhel, double kav, double cs0eff, double dt, const double kvec[3],
                              double phase, double phi, double coef1[3], double coef2[3], double fda[3], double kk[3],
                              double *fact);

/* ask for time-series diagnostics to be reduced during the next substep 1
   (ldiagnos=lfirst.and.lout, equ.f90:120) */
void pcb200_request_diagnostics(void);

/* multi-GPU: NCCL bootstrap without MPI.  Rank 0 obtains an id (128 bytes),
   the host broadcasts it (MPI_Bcast in Fortran, torch.distributed in bench.py),
   every rank calls pcb200_nccl_init before initialize_gpu_c. */
void pcb200_nccl_get_unique_id(char id[128]);
void pcb200_nccl_init(const char id[128], const int *rank, const int *nranks);

/* use an externally owned CUDA stream (cudaStream_t as void*) for all work; NULL = own stream */
void pcb200_set_stream(void *cuda_stream);
/* z-halo exchange as copy-engine "puts" into the neighbours' ghost planes over NVLink peer memory (CUDA IPC),
   replacing the NCCL send/recv pair (and with it mpicomm.f90:1544-2271 for the z direction) without using SMs:
   every rank exports one blob after initialize_gpu_c, the host all-gathers them (MPI_Allgather in gpu_b200.f90)
   and every rank connects with the table of all ranks' blobs (rank order).  With yz-pencils (nprocy > 1) the y faces
   travel the same way: one strided 2-D copy per side into the y neighbour's ghost rows (no pack/unpack kernels;
   load_farray_c keeps the two-sided exchange).  Optional: without these calls all exchanges run on NCCL.  Ranks must not call finalize_gpu_c before all ranks have finished stepping. */
int pcb200_p2p_blob_bytes(void);
void pcb200_p2p_export(unsigned char *blob);
/* returns 0 when both z neighbours' buffers are mapped, 1 when they cannot be (peer GPU not visible, no P2P) */
int pcb200_p2p_connect(const unsigned char *blobs, const int *nranks);
/* call on all ranks after every rank's connect returned 0 (the host reduces the status: MPI_ALLREDUCE / all_reduce) */
void pcb200_p2p_enable(void);
/* asynchronous copy_farray (extension): device-side snapshot in stream order, device-to-host transfer on a
   copy stream overlapping the following substeps; _wait blocks until the host array f is complete */
void pcb200_copy_farray_async(double *f);
void pcb200_copy_farray_wait(void);
/* var.dat snapshot of this rank in the reference's distributed-I/O layout (src/io_dist.f90:108-199
   output_snap, :697-788 read_snap): two gfortran sequential unformatted records, {f(mx,my,mz,nvar)} and
   {t, x(mx), y(my), z(mz), dx, dy, dz}, records above 2 GiB as subrecords.  Host-only (no device work): f is the host
   f-array, e.g. after pcb200_copy_farray_wait.  By value, not Fortran bindings (the Fortran build keeps its own Io
   module).  Unlike the device entry points these RETURN a status: 0 ok, 1 I/O error (errno set), 2 file does not have
   the expected record structure/sizes, 3 bad argument.  dxyz receives dx, dy, dz. */
int pcb200_write_snapshot(const char *path, const double *f, int mx, int my, int mz, int nvar, double t, const double *x,
                          const double *y, const double *z, double dx, double dy, double dz);
int pcb200_read_snapshot(const char *path, double *f, int mx, int my, int mz, int nvar, double *t, double *x, double *y,
                         double *z, double *dxyz);
/* block until all queued device work has finished */
void pcb200_synchronize(void);
/* device-side synthetic initial condition written straight into the device f-array */
void pcb200_init_synthetic(const int *seed);
/* number of kernels this library has launched so far (bench.py: gpu_launches) */
long long pcb200_launch_count(void);
/* number of blocking host synchronisations (cudaStream/Event/DeviceSynchronize) the library has issued so far: a
   fixed-dt step without diagnostics must not add any (tests/test_parity_gpu.py) */
long long pcb200_host_sync_count(void);
/* device time in ms spent in the fused rhs+rk kernel since the last reset, and its launch count */
void pcb200_kernel_timing(double *ms, long long *launches, const int *reset);
void pcb200_enable_kernel_timing(const int *on);

/* ------------------------------------------------------------------ */
/* Part 1: the reference's GPU-module C interface                       */
/* ------------------------------------------------------------------ */
/* gpu_astaroth.cc:2009-2122 initializeGPU; RETAINS f for load/copy */
void initialize_gpu_c(double *f, int *comm_fint, double *t, int *nt, int *lread_all_vars_from_device,
                      int *lcpu_timestep_on_gpu, int *lac_sparse_autotuning);
void register_gpu_c(void);
void finalize_gpu_c(void);
/* gpu_astaroth.cc:1672-1717 substepGPU: ghost exchange + RHS + 2N-RK update (+forcing after 3) */
void rhs_gpu_c(int *isubstep, double *t);
void before_boundary_gpu_c(int *lrmv, int *isubstep, double *t, int *lsubstepping);
void update_after_substep_gpu_c(void);
/* gpu_astaroth.cc:1947-1971 loadFarray (uses the retained pointer) / :1719-1814 copyFarray */
void load_farray_c(void);
void copy_farray_c(double *f);
void gpu_prepare_for_first_substep_c(double *t);
/* gpu_astaroth.cc:1833-1849 copyVBApointers.  The device layout is the library's own (as the reference's VBA is
   Astaroth's): x fastest, row pitch mx (+2 with one pad cell at either end for kernel variant 10), and with kernel
   variant 10 and the entropy module the variables are stored as u, s, lnrho, A (lnrho and s exchanged);
   load_farray_c / copy_farray_c translate to and from the Fortran f-array. */
void get_farray_ptr_gpu_c(double **p_in, double **p_out);
void reload_gpu_config_c(void);
/* gpu_astaroth.cc:2179-2241 updateInConfig* ; return index>=0 or -1 if unknown */
int update_on_gpu_scal_by_name_c(const char *name, double *value);
int update_on_gpu_vec_by_name_c(const char *name, double *value);
int update_on_gpu_arr_by_name_c(const char *name);
void update_on_gpu_scal_by_ind_c(int *index, double *value);
void update_on_gpu_vec_by_ind_c(int *index, double *value);
void update_on_gpu_arr_by_ind_c(int *index);
void *pos_real_ptr_c(void **ptr, int *ind);
void split_update_gpu_c(double *f);
void radtransfer_gpu_c(void);
void test_bcs_c(void);
/* the check behind test_bcs_c (astaroth/gpu_astaroth.cc:2283-2454 testBCs): number of ghost values that differ from the
   CPU boundary conditions on a randomised mesh by more than 1e-12 relative (0 = passed); does not exit */
long long pcb200_test_bcs(void);

/* No counterpart in the reference: added for callers that cannot call
   cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
   Courant reduction of a device f-array WITHOUT any allocation: dt1_max_dev[0] = max over all interior cells of this rank
   of the inverse time step (equ.f90:1858-1972, the quantity pcb200_get_dt1_max returns after substep 1 of a Courant
   step, before the reduction over ranks), computed by the same kernels; the per-block partial maxima live in a
   file-scope static __device__ array of PCB200_NOALLOC_MAX_BLOCKS doubles (32 KB, held by every GPU that loads the
   library until exit) instead of cudaMalloc'ed scratch.
     f_dev        device f-array in the library's layout with valid ghost zones (get_farray_ptr_gpu_c, first pointer)
     dt1_max_dev  device pointer to one double (output)
     nblocks      size of the reduction grid = number of scratch slots used, 1..PCB200_NOALLOC_MAX_BLOCKS
   Work is queued on the library's stream (pcb200_set_stream); no host synchronisation, no cudaMalloc/cudaFree.
   Returns 0, or prints the reason to stderr and returns 1 (nblocks beyond the maximum or < 1), 2 (not initialized),
   3 (null argument), 4 (scratch address unavailable), 5 (launch failed).  All calls share the scratch array, which
   keeps the last call's contents (nothing in it needs clearing); calls that overlap on the GPU race on it. */
#define PCB200_NOALLOC_MAX_BLOCKS 4096
int pcb200_dt1_max_noalloc(const double *f_dev, double *dt1_max_dev, const int *nblocks);

/* gpu.h:12 n_gpu_reduced_vars=20.  Slot layout of dst[]: */
enum {
  PCB200_RED_URMS = 0, PCB200_RED_UMAX, PCB200_RED_RHOM, PCB200_RED_BRMS, PCB200_RED_BMAX,
  PCB200_RED_JRMS, PCB200_RED_JMAX, PCB200_RED_ABM, PCB200_RED_JBM, PCB200_RED_OUM,
  PCB200_RED_EPSK, PCB200_RED_SSM, PCB200_RED_DTU, PCB200_RED_DTC, PCB200_RED_DTB,
  PCB200_RED_DTNU, PCB200_RED_DT1MAX, PCB200_RED_DT, PCB200_RED_DTCHI, PCB200_RED_SPARE1,
  PCB200_N_REDUCED = 20
};
void get_gpu_reduced_vars_c(double *dst);

/* ------------------------------------------------------------------ */
/* Part 3: host-side mirror of the reference driver for this path       */
/* (time_step/advance_substeps with lgpu=T: timestep.f90:206-270,       */
/*  77-204; pde: equ.f90:106-120,197-201,307-332; run.f90:575)          */
/* ------------------------------------------------------------------ */
typedef struct pcb200_driver_state {
  double t, dt;
  int it, ldt;     /* ldt = (dt==0) at start: timestep.f90:60-75 */
  int lforcing;    /* iforce='helical' */
  double force, relhel, kav; /* forcing_run_pars */
  int nk;          /* number of wave vectors in k-list */
  const double *kkx, *kky, *kkz;
  unsigned long long rng_state; /* host RNG for fconst_coefs_hel */
  int it1;         /* diagnostics cadence (run_pars it1) */
} pcb200_driver_state;

/* advance nt steps; if ts_rows!=NULL, writes one row of PCB200_N_REDUCED+2 doubles
   (it, t, reduced vars) for every step with mod(it-1,it1)==0; returns rows written */
int pcb200_driver_run(pcb200_driver_state *st, int nt, double *ts_rows, int max_rows);

#ifdef __cplusplus
}
#endif
#endif
