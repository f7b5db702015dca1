/* test_rhs_replay.c -- the reference's ltest_rhs test mode (src/equ.f90:1974-2143, subroutine test_rhs_gpu) replayed
 * in C against libpcb200.so, because no Fortran compiler exists in this image.  What Fortran would do is done here with
 * the same symbols: every `*_c` entry point is resolved by its FTNIZE'd name (trailing underscore, src/headers_c.h:31-37,
 * as `external` procedures of gpu_astaroth.f90:21-53 are) and called with by-reference arguments.  The CPU side of the
 * comparison (`cpu_version`, i.e. rhs_cpu) is the oracle -- TEST INFRASTRUCTURE, this program lives under tests/.
 *
 *   test_rhs_replay <libpcb200.so> <libpc_oracle.so> <pcb200_config blob> <pc_oracle_pars blob> <f0 raw doubles> <dt>
 *
 * Protocol (line numbers of equ.f90): copy_farray_from_GPU(f); f_copy=f; df_copy=0 (2026-2029); per substep: df_copy=
 * alpha*df_copy (2037), before_boundary_gpu, rhs_gpu (2038-2039), update_ghosts(f_copy), cpu_version(f_copy,df_copy)
 * (2044-2054), f_copy += dt*beta*df_copy on the interior (2059), copy_farray_from_GPU (2065); afterwards
 * update_after_substep_gpu, copy_farray_from_GPU, f_diff = |(f_copy-f)/(f_copy+tini)| and the "Max comp diff" report
 * (2068-2137), then test_gpu_bcs (2141).  The reference's measure is pointwise relative (it is large wherever a field
 * crosses zero), so the north-star's measure -- max|f_copy-f| / max|f_copy| per variable -- is printed as well and
 * decides the exit status: 0 when every variable agrees to 1e-12. */
#include <dlfcn.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "pcb200.h"
#include "pc_oracle.h"

static void *must(void *h, const char *name) {
  void *p = dlsym(h, name);
  if (!p) { fprintf(stderr, "symbol %s missing: %s\n", name, dlerror()); exit(2); }
  return p;
}
static void *slurp(const char *path, size_t *n) {
  FILE *f = fopen(path, "rb");
  if (!f) { perror(path); exit(2); }
  fseek(f, 0, SEEK_END);
  *n = (size_t)ftell(f);
  fseek(f, 0, SEEK_SET);
  void *p = malloc(*n);
  if (fread(p, 1, *n, f) != *n) { perror("fread"); exit(2); }
  fclose(f);
  return p;
}

int main(int argc, char **argv) {
  if (argc != 7) { fprintf(stderr, "usage: %s libpcb200.so libpc_oracle.so config.bin pars.bin f0.bin dt\n", argv[0]); return 2; }
  void *L = dlopen(argv[1], RTLD_NOW | RTLD_GLOBAL), *O = dlopen(argv[2], RTLD_NOW);
  if (!L || !O) { fprintf(stderr, "dlopen: %s\n", dlerror()); return 2; }
  /* the Fortran-visible names (trailing underscore) */
  void (*register_gpu)(void) = must(L, "register_gpu_c_");
  void (*initialize_gpu)(double *, int *, double *, int *, int *, int *, int *) = must(L, "initialize_gpu_c_");
  void (*load_farray)(void) = must(L, "load_farray_c_");
  void (*copy_farray)(double *) = must(L, "copy_farray_c_");
  void (*before_boundary_gpu)(int *, int *, double *, int *) = must(L, "before_boundary_gpu_c_");
  void (*rhs_gpu)(int *, double *) = must(L, "rhs_gpu_c_");
  void (*update_after_substep_gpu)(void) = must(L, "update_after_substep_gpu_c_");
  void (*test_bcs)(void) = must(L, "test_bcs_c_");
  void (*finalize_gpu)(void) = must(L, "finalize_gpu_c_");
  /* what gpu_b200.f90 binds by name (bind(C)): sizes/parameters and the time step */
  void (*configure)(const pcb200_config *) = must(L, "pcb200_configure");
  int (*sizeof_config)(void) = must(L, "pcb200_sizeof_config");
  void (*set_dt)(double *, int *) = must(L, "pcb200_set_dt");
  /* oracle */
  void (*o_ghosts)(const pc_oracle_pars *, double *) = must(O, "pc_oracle_update_ghosts");
  double (*o_rhs)(const pc_oracle_pars *, const double *, double *, int, pc_oracle_diag *) = must(O, "pc_oracle_rhs");
  int (*o_sizeof)(void) = must(O, "pc_oracle_sizeof_pars");

  size_t nc, np, nf;
  pcb200_config *cfg = slurp(argv[3], &nc);
  pc_oracle_pars *par = slurp(argv[4], &np);
  double *f = slurp(argv[5], &nf);
  const double dt0 = atof(argv[6]);
  if ((int)nc != sizeof_config() || nc != sizeof(pcb200_config)) { fprintf(stderr, "pcb200_config size mismatch\n"); return 2; }
  if ((int)np != o_sizeof() || np != sizeof(pc_oracle_pars)) { fprintf(stderr, "pc_oracle_pars size mismatch\n"); return 2; }
  const int mx = cfg->nx + 6, my = cfg->ny + 6, mz = cfg->nz + 6, mvar = cfg->nvar;
  const size_t nreal = (size_t)mx * my * mz * mvar;
  if (nf != nreal * sizeof(double)) { fprintf(stderr, "f0 has %zu bytes, expected %zu\n", nf, nreal * sizeof(double)); return 2; }

  /* initialize_GPU (gpu_astaroth.f90:192-245): sizes, then the C side retains f; nt>0 -> load_farray_to_GPU */
  configure(cfg);
  register_gpu();
  int comm = 0, nt = 1, zero = 0;
  double t = 0.0;
  initialize_gpu(f, &comm, &t, &nt, &zero, &zero, &zero);
  load_farray();

  /* ---- test_rhs_gpu ---- */
  const double beta_ts[3] = {1 / 3.0, 15 / 16.0, 8 / 15.0}, alpha_ts[3] = {0.0, -5 / 9.0, -153 / 128.0};
  double *f_copy = malloc(nreal * sizeof(double)), *df_copy = calloc(nreal, sizeof(double));
  copy_farray(f);
  memcpy(f_copy, f, nreal * sizeof(double));
  double dt = dt0;
  int ldt = 0, lrmv = 0, lsub = 0;
  set_dt(&dt, &ldt); /* gpu_b200.f90 rhs_GPU: before the first substep */
  for (int itsub = 1; itsub <= 3; ++itsub) {
    if (itsub > 1)
      for (size_t i = 0; i < nreal; ++i) df_copy[i] *= alpha_ts[itsub - 1];
    before_boundary_gpu(&lrmv, &itsub, &t, &lsub);
    rhs_gpu(&itsub, &t);
    o_ghosts(par, f_copy);
    o_rhs(par, f_copy, df_copy, 0, NULL); /* cpu_version: df_copy += rhs */
    for (int k = 0; k < mvar; ++k)
      for (int n = 3; n < mz - 3; ++n)
        for (int m = 3; m < my - 3; ++m)
          for (int l = 3; l < mx - 3; ++l) {
            const size_t i = (((size_t)k * mz + n) * my + m) * mx + l;
            f_copy[i] += dt * beta_ts[itsub - 1] * df_copy[i];
          }
    copy_farray(f);
  }
  update_after_substep_gpu();
  copy_farray(f);

  const double tini = 1e-300;
  const char *names[8] = {"ux", "uy", "uz", "lnrho", "ss", "ax", "ay", "az"};
  double worst = 0.0, worst_abs = 0.0, worst_linf = 0.0;
  for (int k = 0; k < mvar; ++k) {
    double mxd = 0.0, sum = 0.0, mxa = 0.0, mxf = 0.0;
    int loc[3] = {0, 0, 0};
    for (int n = 3; n < mz - 3; ++n)
      for (int m = 3; m < my - 3; ++m)
        for (int l = 3; l < mx - 3; ++l) {
          const size_t i = (((size_t)k * mz + n) * my + m) * mx + l;
          const double ad = fabs(f_copy[i] - f[i]);
          double d = fabs((f_copy[i] - f[i]) / (f_copy[i] + tini));
          if (f_copy[i] == 0.0 && ad < 1e-20) d = 0.0;
          sum += d;
          if (d > mxd) { mxd = d; loc[0] = l + 1; loc[1] = m + 1; loc[2] = n + 1; }
          if (ad > worst_abs) worst_abs = ad;
          if (ad > mxa) mxa = ad;
          if (fabs(f_copy[i]) > mxf) mxf = fabs(f_copy[i]);
        }
    const char *nm = (cfg->iss < 0 && k >= 4) ? names[k + 1] : names[k];
    printf(" Max comp diff for %s %d : %.6e\n", nm, k + 1, mxd);
    printf(" Avg comp diff for %s %d : %.6e\n", nm, k + 1, sum / ((double)cfg->nx * cfg->ny * cfg->nz));
    printf(" Max comp loc  for %s %d : %d %d %d\n", nm, k + 1, loc[0], loc[1], loc[2]);
    printf(" rel Linf      for %s %d : %.6e\n", nm, k + 1, mxa / mxf);
    if (mxd > worst) worst = mxd;
    if (mxa / mxf > worst_linf) worst_linf = mxa / mxf;
  }
  printf(" Max comp diff: %.6e\n", worst);
  printf(" Max comp loc abs diff: %.6e\n", worst_abs);
  fflush(stdout);
  test_bcs(); /* call test_gpu_bcs: exits with EXIT_FAILURE when the ghost zones differ */
  finalize_gpu();
  return worst_linf <= 1e-12 ? 0 : 1;
}
