/*
 * pc_oracle.h -- C interface of the CPU ORACLE (test infrastructure only).
 *
 * The oracle is a plain C++ restatement of Pencil Code's per-substep RHS
 * evaluation (pde -> rhs_cpu -> calc_all_pencils/calc_all_rhs -> set_dt1_max)
 * and the 2N-RK3 update of timestep.f90, in the reference's pencil order.
 * It is NOT part of the product: only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load it.
 *
 * Parity status: the Fortran reference cannot be compiled in this image
 * (no gfortran, no MPI), so tight (1e-12) parity is pinned by this
 * restatement only; against the reference's own artefacts it is pinned by
 *   - python/pencil/math/derivatives/der_6th_order_w_ghosts.py (imported from
 *     /root/reference when generating tests/golden/deriv_golden.npz),
 *   - the closed-form it=0 row of samples/gputest/cpu-test.out:273
 *     (urms, umax, rhom, epsK) -- see tests/test_oracle_golden.py.
 * => "parity coarsely pinned" (SURVEY.md section 8c).
 *
 * Layout: f[ivar][iz][iy][ix], x fastest, mx=nx+6 etc. (src/cdata.f90:90-92).
 */
#ifndef PC_ORACLE_H
#define PC_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct pc_oracle_pars {
  /* local interior grid and variables (cparam.f90:54-75) */
  int nx, ny, nz;
  int nvar;
  int iux, ilnrho, iss, iax; /* 0-based slot of first component, -1 = module absent */
  double dx, dy, dz;
  double x0, y0, z0; /* coordinate of first interior point: x(l1), y(m1), z(n1) */
  /* eos_idealgas.f90:46-47,154-270 */
  double cs20, gamma, cp, lnrho0;
  /* cdata.f90:64 */
  double mu0;
  /* viscosity.f90:1633-1664,1982-1989,2250-2257 */
  double nu, nu_hyper3;
  int lvisc_nu_const, lvisc_hyper3_simplified, lvisc_hyper3_nu_const;
  /* density.f90:3197-3288 */
  double diffrho_hyper3;
  int ldiff_hyper3lnrho;
  /* entropy.f90:4852-4947, viscosity.f90:2743-2789 */
  double chi;
  int lheatc_chiconst, lviscosity_heat;
  /* magnetic.f90:5584-5590,6138-6144 */
  double eta;
  int lresi_eta_const, lohmic_heat, llorentzforce;
  /* upwinding switches: hydro.f90:286, density.f90:122, entropy.f90:139, magnetic.f90:359 */
  int lupw_uu, lupw_lnrho, lupw_ss, lupw_aa;
  /* cdata.f90:235-239 */
  double cdt, cdtv, cdtv2, cdtv3;
  double dtmax; /* equ.f90:270-280: dt1_max initialised to 1/dtmax */
  int nthreads; /* OpenMP threads over (m,n) pencils; <=0 -> default */
  /* ---- vertically stratified set-ups (samples/conv-slab); all zero = triply periodic box ---- */
  /* z boundary condition per [bottom/top][variable slot] (boundcond.f90:1421-1787): PC_BC_P 'p', PC_BC_S 's',
     PC_BC_A 'a', PC_BC_A2 'a2', PC_BC_C1 'c1' (entropy: constant heat flux, entropy_bcs.f90:93-235),
     PC_BC_CT 'cT' (entropy: constant temperature, eos_idealgas.f90:3112-3260) */
  int bcz[2][8];
  double gravz;     /* gravity_simple.f90:521-523 (gravz_profile='const'), addgravity :1083-1162 */
  double Omega;     /* hydro.f90:6339-6366 coriolis_cartesian, theta=0 */
  int lheatc_Kprof; /* entropy.f90:1176-1178, calc_heatcond_arrays :5913-6119 */
  double hcond0, hcond1, hcond2, widthss, z1, z2; /* get_gravz_heatcond entropy.f90:7741-7769 */
  double cool, wcool, cs2cool, ztop;              /* entropy.f90:1323,1367; get_heat_cool_gravz :6761-6828 */
  double FbotKbot, FtopKtop;                      /* entropy.f90:854-887 */
  double cs2bot, cs2top;                          /* density.f90:1605-1610, entropy.f90:818-838 */
} pc_oracle_pars;

enum { PC_BC_P = 0, PC_BC_S = 1, PC_BC_A = 2, PC_BC_A2 = 3, PC_BC_C1 = 4, PC_BC_CT = 5 };

/* helical forcing coefficients for one step (forcing.f90:1773-2001) */
typedef struct pc_oracle_forcing {
  double coef1[3], coef2[3], fda[3], kk[3];
  double phase, fact;
  double k1_ff;
} pc_oracle_forcing;

/* diagnostics accumulated during substep 1 */
typedef struct pc_oracle_diag {
  double urms, umax, rhom, brms, bmax, jrms, jmax, abm, jbm, oum, epsK, ssm;
  double dtu, dtc, dtb, dtnu; /* max(advec_uu)/cdt etc. (not yet multiplied by dt) */
  double dt1_max;
  double dtchi; /* max(diffus_chi)/cdtv (entropy.f90:4184-4220) */
} pc_oracle_diag;

/* ghost fill in x, then y (periodic, boundcond.f90:2387-2433), then z (periodic, or bcz bottom then top in
   variable order, boundcond.f90:1421-1787; 'a' and 'cT' also set the boundary plane itself) */
void pc_oracle_update_ghosts(const pc_oracle_pars *p, double *f);

/* one derivative of variable k along j (1..3) for the whole interior; out is nx*ny*nz.
   kind: 1=der, 2=der2, 5=der5, 6=der6, 7=der6 upwind; for kind 12/23/31 mixed derij. */
void pc_oracle_deriv(const pc_oracle_pars *p, const double *f, int k, int j, int kind, double *out);

/* df(interior) += RHS(f); ghosts of f must be valid.  Returns max dt1 (0 if !lcourant).
   diag may be NULL. */
double pc_oracle_rhs(const pc_oracle_pars *p, const double *f, double *df, int lcourant,
                     pc_oracle_diag *diag);

/* one substep: ghosts, df=alpha*df (or 0), rhs, optional dt from CFL (itsub==1 && ldt),
   f += dt*beta*df.  *dt is in/out. */
void pc_oracle_substep(const pc_oracle_pars *p, double *f, double *df, int itsub, int ldt,
                       double *dt, double *t, pc_oracle_diag *diag);

/* full RK3 step (timestep.f90:206-270) + optional forcing (run.f90:575) */
void pc_oracle_time_step(const pc_oracle_pars *p, double *f, double *df, int ldt, double *dt,
                         double *t, const pc_oracle_forcing *frc, pc_oracle_diag *diag);

/* u += Re((coef1 + i coef2) fx fy fz) (forcing.f90:2170-2345) */
void pc_oracle_forcing_hel(const pc_oracle_pars *p, double *f, const pc_oracle_forcing *frc);

/* host-side coefficient generation (forcing.f90:1773-2001) given k-vector, phase and phi */
void pc_oracle_forcing_coefs(double force, double relhel, double kav, double cs0eff, double dt,
                             const double kvec[3], double phase, double phi,
                             pc_oracle_forcing *out);

int pc_oracle_sizeof_pars(void);

#ifdef __cplusplus
}
#endif
#endif
