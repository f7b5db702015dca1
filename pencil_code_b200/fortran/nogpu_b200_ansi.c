/*                            nogpu_b200_ansi.c
                              -----------------
   Dummy of libpcb200.so's C ABI for linking start.x (role of src/nogpu_astaroth_ansi.c:1-60,
   which src/Makefile.src:3139 links into start.x): start.x only needs the symbols that
   gpu_b200.f90 binds to resolve; it never calls the device.  Signatures: include/pcb200.h.
*/
#include <stdio.h>
#include <stdlib.h>
#include "pcb200.h"

static void not_in_start(const char *who)
{
  fprintf(stderr, "gpu_b200: %s called from start.x -- the GPU module is only active in run.x\n", who);
  exit(EXIT_FAILURE);
}
void pcb200_configure(const pcb200_config *cfg) { (void)cfg; }
void pcb200_reconfigure(const pcb200_config *cfg) { (void)cfg; }
void pcb200_hand_dt1_to_set_dt_c(void) { not_in_start("pcb200_hand_dt1_to_set_dt_c"); }
int pcb200_forcing_supported_c(const int *nx, const int *my, const int *mz) { (void)nx; (void)my; (void)mz; return 1; }
int pcb200_sizeof_config(void) { return (int)sizeof(pcb200_config); }
void pcb200_fill_physics_pars_c(pcb200_config *cfg) { (void)cfg; }
int pcb200_fill_strat_pars_c(pcb200_config *cfg) { (void)cfg; return 1; }
void pcb200_push_forcing_c(void) {}
void pcb200_set_dt(const double *dt, const int *ldt) { (void)dt; (void)ldt; }
void pcb200_get_dt(double *dt) { (void)dt; not_in_start("pcb200_get_dt"); }
void pcb200_get_dt1_max(double *d) { (void)d; not_in_start("pcb200_get_dt1_max"); }
void pcb200_request_diagnostics(void) {}
void pcb200_nccl_get_unique_id(char id[128]) { (void)id; }
void pcb200_nccl_init(const char id[128], const int *rank, const int *nranks) { (void)id; (void)rank; (void)nranks; }
int pcb200_p2p_blob_bytes(void) { return 192; }
void pcb200_p2p_export(unsigned char *blob) { (void)blob; }
int pcb200_p2p_connect(const unsigned char *blobs, const int *nranks) { (void)blobs; (void)nranks; return 1; }
void pcb200_p2p_enable(void) {}
void initialize_gpu_c(double *f, int *comm, double *t, int *nt, int *a, int *b, int *c)
{ (void)f; (void)comm; (void)t; (void)nt; (void)a; (void)b; (void)c; }
void register_gpu_c(void) {}
void finalize_gpu_c(void) {}
void rhs_gpu_c(int *isubstep, double *t) { (void)isubstep; (void)t; not_in_start("rhs_gpu_c"); }
void before_boundary_gpu_c(int *lrmv, int *isubstep, double *t, int *lsub) { (void)lrmv; (void)isubstep; (void)t; (void)lsub; }
void update_after_substep_gpu_c(void) {}
void load_farray_c(void) {}
void copy_farray_c(double *f) { (void)f; }
void gpu_prepare_for_first_substep_c(double *t) { (void)t; }
void get_farray_ptr_gpu_c(double **p_in, double **p_out) { *p_in = 0; *p_out = 0; }
void reload_gpu_config_c(void) {}
int update_on_gpu_scal_by_name_c(const char *name, double *value) { (void)name; (void)value; return -1; }
void update_on_gpu_scal_by_ind_c(int *index, double *value) { (void)index; (void)value; }
void *pos_real_ptr_c(void **ptr, int *ind) { (void)ptr; (void)ind; return 0; }
void get_gpu_reduced_vars_c(double *dst) { (void)dst; }
void split_update_gpu_c(double *f) { (void)f; }
void test_bcs_c(void) {}
