/*                             gpu_b200_ansi.c
                              ----------------
   C shim compiled inside Pencil Code's src/ (GPU_INTERFACE object for GPU=gpu_b200; the role
   gpu_astaroth_ansi.c plays for Astaroth).  It fetches the physics parameters that are private
   to their Fortran modules through the existing pushpars2c address tables and forwards the
   helical-forcing coefficients, then everything goes to libpcb200.so (include/pcb200.h).

   Table indices (1-based in Fortran, so -1 here) are those of the reference:
     viscosity.f90:3250-3255,3311   nu(1) nu_hyper3(3) lvisc_nu_const(5) lvisc_hyper3_nu_const(6)
                                    lvisc_hyper3_simplified(62)
     magnetic.f90:12063-12070,12166 eta(1) lresi_eta_const(4) lupw_aa(7) llorentzforce(8) lohmic_heat(105)
     eos_idealgas.f90:4589-4594     cs20(1) gamma(2) cp(4) lnrho0(5)
     density.f90:4591-4593          ldiff_hyper3lnrho(3) diffrho_hyper3(4) lupw_lnrho(5)
     hydro.f90:9367                 lupw_uu(2)
     entropy.f90:8675-8816          chi(1) lheatc_chiconst(18) lupw_ss(29) lviscosity_heat(141)
     eos_idealgas.f90:4598-4599     cs2bot(10) cs2top(11)
     gravity_simple.f90:1571-1609   gravz(1) gravx_xpencil(11) gravy_ypencil(13) enum_gravz_profile(20) z1(23) z2(24)
     entropy.f90:8686-8864          FbotKbot(14) FtopKtop(15) widthss(31) luminosity(34) cool(36) wcool(43) cs2cool(53)
                                    hcond0(72) hcond1(73) hcond2(76) lheatc_kprof(116) lcooling_general(132)
                                    lcooling_to_cs2cool(133) lmultilayer(139) lhcond0_density_dep(156) lcalc_heat_cool(167)
   Module-procedure symbol names follow the compiler's mangling; the Makefile passes
   MODULE_PREFIX/INFIX/SUFFIX exactly as it does for the Astaroth glue (gfortran: __ , _MOD_ , "").
*/
#include <stddef.h>
#include <math.h>
#include "headers_c.h"
#include "pcb200.h"

#ifndef MODULE_PREFIX
#define MODULE_PREFIX __
#define MODULE_INFIX _MOD_
#define MODULE_SUFFIX
#endif
#define PC_CAT5_(a, b, c, d, e) a##b##c##d##e
#define PC_CAT5(a, b, c, d, e) PC_CAT5_(a, b, c, d, e)
#define MODPROC(mod, proc) PC_CAT5(MODULE_PREFIX, mod, MODULE_INFIX, proc, MODULE_SUFFIX)

#define NPAR_MAX 2048
extern void MODPROC(viscosity, pushpars2c)(FINT8 *);
extern void MODPROC(magnetic, pushpars2c)(FINT8 *);
extern void MODPROC(equationofstate, pushpars2c)(FINT8 *);
extern void MODPROC(density, pushpars2c)(FINT8 *);
extern void MODPROC(hydro, pushpars2c)(FINT8 *);
extern void MODPROC(energy, pushpars2c)(FINT8 *);
extern void MODPROC(gravity, pushpars2c)(FINT8 *);
extern void MODPROC(forcing, forcing_pars_hel)(REAL *coef1, REAL *coef2, REAL *coef3, REAL *fda, REAL *kk,
                                               REAL *phase, REAL *fact);
extern void MODPROC(forcing, pushpars2c)(FINT8 *);
extern void MODPROC(sub, set_dt)(REAL *dt1_); /* sub.f90:4223-4254; the name scripts/sub2c builds for the Astaroth glue */

#define PD(tab, i) (*(double *)(tab)[(i)-1])
#define PI_(tab, i) (*(int *)(tab)[(i)-1])

void pcb200_fill_physics_pars_c(pcb200_config *c)
{
  static FINT8 p[NPAR_MAX];
  int i;

  for (i = 0; i < NPAR_MAX; i++) p[i] = 0;
  MODPROC(equationofstate, pushpars2c)(p);
  c->cs20 = PD(p, 1); c->gamma = PD(p, 2); c->cp = PD(p, 4); c->lnrho0 = PD(p, 5);

  for (i = 0; i < NPAR_MAX; i++) p[i] = 0;
  MODPROC(viscosity, pushpars2c)(p);
  c->nu = PD(p, 1); c->nu_hyper3 = PD(p, 3);
  c->lvisc_nu_const = PI_(p, 5); c->lvisc_hyper3_nu_const = PI_(p, 6); c->lvisc_hyper3_simplified = PI_(p, 62);

  for (i = 0; i < NPAR_MAX; i++) p[i] = 0;
  MODPROC(density, pushpars2c)(p);
  c->ldiff_hyper3lnrho = PI_(p, 3); c->diffrho_hyper3 = PD(p, 4); c->lupw_lnrho = PI_(p, 5);

  for (i = 0; i < NPAR_MAX; i++) p[i] = 0;
  MODPROC(hydro, pushpars2c)(p);
  c->lupw_uu = PI_(p, 2);

  c->chi = 0.; c->lheatc_chiconst = 0; c->lupw_ss = 0; c->lviscosity_heat = 1;
  if (c->iss >= 0) {
    for (i = 0; i < NPAR_MAX; i++) p[i] = 0;
    MODPROC(energy, pushpars2c)(p);
    c->chi = PD(p, 1); c->lheatc_chiconst = PI_(p, 18); c->lupw_ss = PI_(p, 29); c->lviscosity_heat = PI_(p, 141);
  }
  c->eta = 0.; c->lresi_eta_const = 0; c->lupw_aa = 0; c->llorentzforce = 1; c->lohmic_heat = 1;
  if (c->iax >= 0) {
    for (i = 0; i < NPAR_MAX; i++) p[i] = 0;
    MODPROC(magnetic, pushpars2c)(p);
    c->eta = PD(p, 1); c->lresi_eta_const = PI_(p, 4); c->lupw_aa = PI_(p, 7); c->llorentzforce = PI_(p, 8);
    c->lohmic_heat = PI_(p, 105);
  }
}

/* Vertically stratified set-ups (lgrav): constant gravz, K-profile conduction, 'gaussian' cooling layer, the constants
   of the 'c1' / 'cT' entropy boundary conditions.  Returns 0 when the configuration uses something else.
   The profile selectors are string enums on the Fortran side (enum_gravz_profile, enum_cooling_profile are hashes of
   the strings), so the supported case is recognised by its effect instead: gravx/gravy pencils all zero and
   gravz_zpencil == gravz everywhere is what gravz_profile='const' produces (gravity_simple.f90:521-523). */
int pcb200_fill_strat_pars_c(pcb200_config *c)
{
  static FINT8 p[NPAR_MAX];
  int i;
  const int mx = c->nx + 6, my = c->ny + 6, mz = c->nz + 6;

  for (i = 0; i < NPAR_MAX; i++) p[i] = 0;
  MODPROC(gravity, pushpars2c)(p);
  c->gravz = PD(p, 1);
  {
    const REAL *gz = (const REAL *)p[1], *gx = (const REAL *)p[10], *gy = (const REAL *)p[12];
    for (i = 0; i < mz; i++) if (gz[i] != (REAL)c->gravz) return 0;
    for (i = 0; i < mx; i++) if (gx[i] != (REAL)0.) return 0;
    for (i = 0; i < my; i++) if (gy[i] != (REAL)0.) return 0;
  }
  c->z1 = PD(p, 23); c->z2 = PD(p, 24);
  if (PI_(p, 8) || PI_(p, 9)) return 0; /* lxyzdependence, lboussinesq_grav */

  for (i = 0; i < NPAR_MAX; i++) p[i] = 0;
  MODPROC(equationofstate, pushpars2c)(p);
  c->cs2bot = PD(p, 10); c->cs2top = PD(p, 11);

  c->lheatc_Kprof = 0; c->cool = 0.;
  if (c->iss >= 0) {
    for (i = 0; i < NPAR_MAX; i++) p[i] = 0;
    MODPROC(energy, pushpars2c)(p);
    c->lheatc_Kprof = PI_(p, 116);
    c->hcond0 = PD(p, 72); c->hcond1 = PD(p, 73); c->hcond2 = PD(p, 76); c->widthss = PD(p, 31);
    if (c->lheatc_Kprof && (!PI_(p, 139) || PI_(p, 156))) return 0; /* lmultilayer=T, lhcond0_density_dep=F */
    if (c->lheatc_Kprof && c->hcond0 == 0.) c->lheatc_Kprof = 0;      /* entropy.f90:6139: nothing is added */
    c->FbotKbot = PD(p, 14); c->FtopKtop = PD(p, 15);
    if (PI_(p, 167)) {                                                /* lcalc_heat_cool */
      if (PD(p, 34) != 0. || PI_(p, 132) || !PI_(p, 133)) return 0;   /* luminosity, lcooling_general, lcooling_to_cs2cool */
      c->cool = PD(p, 36); c->wcool = PD(p, 43); c->cs2cool = PD(p, 53);
      { /* cooling_profile='gaussian': profz_cool(n) = cool*exp(-0.5*((ztop-z)/wcool)^2) (entropy.f90:1323,1367) */
        const REAL *pc = (const REAL *)p[20];
        const double q = (c->ztop - c->z0) / c->wcool, want = c->cool * exp(-0.5 * q * q);
        if (fabs((double)pc[0] - want) > 1e-6 * fabs(c->cool)) return 0;
      }
    }
  }
  return 1;
}

/* called by rhs_GPU before the last substep: draw the step's coefficients with the host RNG
   (forcing.f90:1731-1746 forcing_pars_hel) and hand them to the device */
void pcb200_push_forcing_c(void)
{
  static FINT8 p[NPAR_MAX];
  REAL coef1[3], coef2[3], coef3[3], fda[3], kk[3], phase, fact, k1_ff;
  int on = 1;
  MODPROC(forcing, forcing_pars_hel)(coef1, coef2, coef3, fda, kk, &phase, &fact);
  MODPROC(forcing, pushpars2c)(p);
  k1_ff = PD(p, 1); /* forcing.f90:6620 */
  pcb200_set_forcing_hel(&on, coef1, coef2, fda, kk, &phase, &fact, &k1_ff);
}

/* Courant step: the device has reduced dt1 = max over all cells and GPUs inside substep 1.  Pencil's own dt goes
   through Sub::set_dt exactly as in astaroth/gpu_astaroth.cc:1617-1634 (lgpu: no MPI reduction, dt=1./dt1). */
void pcb200_hand_dt1_to_set_dt_c(void)
{
  double dt1;
  REAL dt1_;
  pcb200_get_dt1_max(&dt1);
  dt1_ = (REAL)dt1;
  MODPROC(sub, set_dt)(&dt1_);
}

/* 1 if the forcing configured in forcing_run_pars is the one this module implements on the device: iforce='helical'
   (lforce_helical(1), forcing.f90:6609-6612), no second forcing function, no continuous forcing, no momentum
   forcing, no z symmetrisation.  Table indices: forcing.f90:6618-6706 (iforcing_zsym 3, n_forcing_cont 10,
   lmomentum_ff 21, lforce_helical 82, lsecond_force 83).  iforce_profile is not pushed as a switch, only the profile
   arrays are (4-9): iforce_profile='nothing' leaves all of them 1 (forcing.f90:373-379), which is what is checked. */
int pcb200_forcing_supported_c(const int *nx, const int *my, const int *mz)
{
  static FINT8 p[NPAR_MAX];
  int i, k;
  const int len[6] = {*nx, *my, *mz, *nx, *my, *mz};
  for (i = 0; i < NPAR_MAX; i++) p[i] = 0;
  MODPROC(forcing, pushpars2c)(p);
  if (!PI_(p, 82)) return 0;      /* lforce_helical(1) */
  if (PI_(p, 83)) return 0;       /* lsecond_force */
  if (PI_(p, 10) != 0) return 0;  /* n_forcing_cont */
  if (PI_(p, 21)) return 0;       /* lmomentum_ff */
  if (PI_(p, 3) != 0) return 0;   /* iforcing_zsym */
  for (k = 0; k < 6; k++) {       /* profx_ampl, profy_ampl, profz_ampl, profx_hel, profy_hel, profz_hel */
    const REAL *prof = (const REAL *)p[3 + k];
    for (i = 0; i < len[k]; i++)
      if (prof[i] != (REAL)1.) return 0;
  }
  return 1;
}
