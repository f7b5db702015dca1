// pcb200_driver.cpp -- host-side mirror (C++) of the reference driver for the GPU path.
//
// No Fortran compiler exists in this image, so the call sequence the Fortran driver performs
// with lgpu=.true. is restated here above the C ABI, in the reference's order:
//   run.f90:381      lout=(mod(it-1,it1)==0)
//   timestep.f90:206-270  time_step  -> advance_substeps (:77-204), itorder=3
//   equ.f90:106-120,197-201,307-332  pde: before_boundary_gpu, rhs_gpu(f,itsub)
//   timestep.f90:146-152,189,201     ds, dtsub, update_after_substep_gpu, t=t+dtsub
//   run.f90:575 / astaroth/forcing.h:1-36  forcing coefficients drawn on the host
//                (forcing.f90:1773-2001 fconst_coefs_hel) and consumed by the device at substep 3
// Everything numerical happens behind the C ABI (libpcb200.so kernels).
#include "../../include/pcb200.h"

#include <cmath>
#include <cstdio>

namespace {

// host RNG standing in for random_number_wrapper (general.f90:644-1032): splitmix64 -> U[0,1)
inline double rng_uniform(unsigned long long &s) {
  s += 0x9E3779B97F4A7C15ULL;
  unsigned long long z = s;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  z = z ^ (z >> 31);
  return (double)(z >> 11) * (1.0 / 9007199254740992.0);
}

inline void cross3(const double *a, const double *b, double *c) {
  c[0] = a[1] * b[2] - a[2] * b[1];
  c[1] = a[2] * b[0] - a[0] * b[2];
  c[2] = a[0] * b[1] - a[1] * b[0];
}

} // namespace

// fconst_coefs_hel (forcing.f90:1773-2001), default branches (no shear, lscale_kvector_*=F,
// old_forcing_evector=F, force_strength=0, slope_ff=0, cs0eff=cs0)
extern "C" void pcb200_forcing_coefs_hel(double force, double relhel, double kav, double cs0eff, double dt,
                                         const double kvec[3], double phase, double phi, double coef1[3], double coef2[3],
                                         double fda[3], double kk[3], double *fact) {
  const double pi = 3.14159265358979323846;
  (void)pi;
  const double kx = kvec[0], ky = kvec[1], kz = kvec[2];
  const double k2 = kx * kx + ky * ky + kz * kz, k = std::sqrt(k2);
  double ee[3] = {1., 0., 0.};
  if (ky == 0 && kz == 0) { ee[0] = 0.; ee[1] = 1.; }
  kk[0] = kx; kk[1] = ky; kk[2] = kz;
  double e1[3], e2[3];
  cross3(kk, ee, e1);
  double n = std::sqrt(e1[0] * e1[0] + e1[1] * e1[1] + e1[2] * e1[2]);
  for (double &v : e1) v /= n;
  cross3(kk, e1, e2);
  n = std::sqrt(e2[0] * e2[0] + e2[1] * e2[1] + e2[2] * e2[2]);
  for (double &v : e2) v /= n;
  for (int q = 0; q < 3; ++q) ee[q] = std::cos(phi) * e1[q] + std::sin(phi) * e2[q];
  const double kde = kk[0] * ee[0] + kk[1] * ee[1] + kk[2] * ee[2];
  double ke[3], kke[3];
  cross3(kk, ee, ke);
  cross3(kk, ke, kke);
  const double ffnorm = std::sqrt(1. + relhel * relhel) * k * std::sqrt(k2 - kde * kde) / std::sqrt(kav * cs0eff * cs0eff * cs0eff);
  *fact = force / ffnorm * std::sqrt(dt);
  for (int q = 0; q < 3; ++q) {
    coef1[q] = k * ke[q];
    coef2[q] = relhel * kke[q];
    fda[q] = 1.;
  }
  (void)phase;
}

extern "C" int pcb200_driver_run(pcb200_driver_state *st, int nt, double *ts_rows, int max_rows) {
  static const double alpha_ts[3] = {0.0, -5.0 / 9.0, -153.0 / 128.0}; // timestep.f90:39-44
  static const double beta_ts[3] = {1.0 / 3.0, 15.0 / 16.0, 8.0 / 15.0};
  const double pi = 3.14159265358979323846;
  const int rowlen = PCB200_N_REDUCED + 2;
  int rows = 0;
  for (int step = 0; step < nt; ++step) {
    st->it += 1;
    const bool lout = st->it1 > 0 && ((st->it - 1) % st->it1 == 0); // run.f90:381
    if (lout) pcb200_request_diagnostics();                          // ldiagnos=lfirst.and.lout
    pcb200_set_dt(&st->dt, &st->ldt);                                // timestep.f90:240
    const double t_diag = st->t; // diagnostics refer to the state before the step (tdiagnos)
    double ds = 0.0;
    for (int itsub = 1; itsub <= 3; ++itsub) {
      ds = (itsub == 1) ? 0.0 : alpha_ts[itsub - 1] * ds; // :109-116
      if (itsub == 3) {
        int lf = st->lforcing;
        double coef1[3] = {0, 0, 0}, coef2[3] = {0, 0, 0}, fda[3] = {1, 1, 1}, kk[3] = {0, 0, 0}, phase = 0, fact = 0, k1 = 1;
        if (lf) {
          // forcing.f90:1804-1811: phase=pi*(2*fran(1)-1); ik=int(nk*(.9999*fran(2))+1); then phi=fran*2pi
          const double f1 = rng_uniform(st->rng_state), f2 = rng_uniform(st->rng_state);
          phase = pi * (2 * f1 - 1.);
          const int ik = (int)(st->nk * (.9999 * f2) + 1) - 1;
          const double phi = rng_uniform(st->rng_state) * 2 * pi;
          const double kv[3] = {st->kkx[ik], st->kky[ik], st->kkz[ik]};
          pcb200_forcing_coefs_hel(st->force, st->relhel, st->kav, 1.0, st->dt, kv, phase, phi, coef1, coef2, fda, kk, &fact);
        }
        pcb200_set_forcing_hel(&lf, coef1, coef2, fda, kk, &phase, &fact, &k1);
      }
      int lrmv = 0, lsub = 0;
      before_boundary_gpu_c(&lrmv, &itsub, &st->t, &lsub); // equ.f90:198
      rhs_gpu_c(&itsub, &st->t);                           // equ.f90:319
      if (itsub == 1 && st->ldt) pcb200_get_dt(&st->dt);   // set_dt: equ.f90:408 -> sub.f90:4223
      ds = ds + 1.0;                                       // timestep.f90:146
      const double dtsub = ds * (st->dt * beta_ts[itsub - 1]); // :150-152
      update_after_substep_gpu_c();                        // :189
      st->t += dtsub;                                      // :201
    }
    if (lout && ts_rows && rows < max_rows) {
      double *r = ts_rows + (size_t)rows * rowlen;
      r[0] = (double)st->it;
      r[1] = t_diag;
      get_gpu_reduced_vars_c(r + 2);
      rows++;
    }
  }
  return rows;
}
