// pcb200_v9.cuh -- kernel variant 9: z-marching with "all-trail" z-mixed derivatives and delayed finalisation.
//
// What changed against variant 5 (profiles/r01_v5_ncu_full_256.txt: 228 LDS.64 + 655 FP64 + 1270 issue slots per
// cell, shared-memory pipe 74 % busy at 8 warps per SM):
//   * all variables are staged alike: ONE TMA box (38 x 14 x 1 x nvar) per xy-plane into a 6-slot shared-memory
//     ring (planes z..z+3 in use, two in flight); only the centre plane is read for in-plane taps, the plane z+3
//     only for the thread's own column (the newest entry of its z history);
//   * the own-column z taps f(z-3..z+2) of every variable live in tensor memory (TMEM, tcgen05.ld/st: per-thread
//     columns), 6 doubles per variable, shifted by one entry per plane with 2-column stores straight from the loaded
//     registers (wider stores would need re-packed, aligned register blocks) -- no register pipelines, no ring reads;
//   * the z-mixed bidiagonal derivatives (deriv.f90:1400-1452)
//        d_zx f(z) = sum_k w_k [Dx_k f(z+k) - Dx_k f(z-k)],   Dx_k f(z') = f(x+k,y,z') - f(x-k,y,z')
//     are never gathered: when plane z' is the centre plane the differences Dx_k, Dy_k that its first derivatives
//     form anyway are scattered into the accumulators of the six output planes z'-3..z'+3 ("trail" for the later
//     planes, as in variant 4/5, and now also for the EARLIER planes, which replaces the 48 "lead" loads and 48 FP64
//     operations per cell).  The price: the terms that contain them -- grad div u in the viscous force, the current
//     density j = curl curl A / mu0 in the Lorentz force and the Ohmic heating -- are complete three planes late, so
//     u (and s when there is a magnetic field) of plane z-3 are finalised and written while plane z is evaluated;
//     what they need besides the accumulators (b, 1/rho, eta mu0/(rho T), the partial ds/dt) waits in TMEM.
//     The accumulators themselves are a circular TMEM ring indexed by output plane (8 slots of 16 columns, updated by
//     read-modify-write with a dynamic column address): no registers are held and nothing is shifted.
//     Every z chunk therefore runs 3 warm-up and 3 cool-down stages that only form the in-plane differences.
// Isotropic grids only (dx == dy == dz); arithmetic and reference citations: pcb200_device.cuh / pcb200_fast.cuh.
#pragma once
#include "pcb200_fast.cuh"
#ifndef PCB200_ABL
#define PCB200_ABL 0
#endif
#ifndef V9_DFPF
#define V9_DFPF 0 // per-thread L2 prefetch of df this many stages ahead (0: off)
#endif
#ifndef V9_RING
#define V9_RING 1 // 1: shifting accumulator slots with static addresses, 0: rotating ring (dynamic addresses)
#endif

namespace pcb200 {

constexpr int V9_NSLOT = 6; // planes z..z+3 in use + 2 in flight

template <int NV> struct V9Layout {
  static constexpr int SLOT = ((NV * V2_PLANE + 15) / 16) * 16; // doubles per slot (TMA destination: 128-byte aligned)
  static constexpr unsigned TX_BYTES = NV * V2_PLANE * 8;
  static constexpr size_t SMEM_BYTES = (size_t)V9_NSLOT * SLOT * 8 + V9_NSLOT * 8 + 64 + 64 * 8 + (size_t)NV * V2_THREADS * 8; // + exp table + df staging
};

// ---- tensor memory as per-thread scratch (shape 32x32b: lane = thread, 32-bit columns).  No "memory" clobber: TMEM
// is invisible to every other memory operation of the kernel, so shared/global loads may be scheduled across these
// blocks; `volatile` keeps the tcgen05 operations in program order among themselves and each load block contains
// its own wait, so no result can be used early.
__device__ __forceinline__ void t9_ld6(uint32_t taddr, double h[6]) { // columns 0..11
  asm volatile(
      "{\n\t.reg .b32 t<12>;\n\t"
      "tcgen05.ld.sync.aligned.32x32b.x8.b32 {t0, t1, t2, t3, t4, t5, t6, t7}, [%6];\n\t"
      "tcgen05.ld.sync.aligned.32x32b.x4.b32 {t8, t9, t10, t11}, [%7];\n\t"
      "tcgen05.wait::ld.sync.aligned;\n\t"
      "mov.b64 %0, {t0, t1};\n\tmov.b64 %1, {t2, t3};\n\tmov.b64 %2, {t4, t5};\n\t"
      "mov.b64 %3, {t6, t7};\n\tmov.b64 %4, {t8, t9};\n\tmov.b64 %5, {t10, t11};\n\t}"
      : "=d"(h[0]), "=d"(h[1]), "=d"(h[2]), "=d"(h[3]), "=d"(h[4]), "=d"(h[5])
      : "r"(taddr), "r"(taddr + 8));
}
__device__ __forceinline__ void t9_ld8(uint32_t taddr, double h[8]) { // columns 0..15
  asm volatile(
      "{\n\t.reg .b32 t<16>;\n\t"
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {t0, t1, t2, t3, t4, t5, t6, t7, t8, t9, t10, t11, t12, t13, t14, t15}, [%8];\n\t"
      "tcgen05.wait::ld.sync.aligned;\n\t"
      "mov.b64 %0, {t0, t1};\n\tmov.b64 %1, {t2, t3};\n\tmov.b64 %2, {t4, t5};\n\tmov.b64 %3, {t6, t7};\n\t"
      "mov.b64 %4, {t8, t9};\n\tmov.b64 %5, {t10, t11};\n\tmov.b64 %6, {t12, t13};\n\tmov.b64 %7, {t14, t15};\n\t}"
      : "=d"(h[0]), "=d"(h[1]), "=d"(h[2]), "=d"(h[3]), "=d"(h[4]), "=d"(h[5]), "=d"(h[6]), "=d"(h[7])
      : "r"(taddr));
}
__device__ __forceinline__ double t9_ld1(uint32_t taddr) { // columns 0..1
  double v;
  asm volatile(
      "{\n\t.reg .b32 t<2>;\n\t"
      "tcgen05.ld.sync.aligned.32x32b.x2.b32 {t0, t1}, [%1];\n\t"
      "tcgen05.wait::ld.sync.aligned;\n\t"
      "mov.b64 %0, {t0, t1};\n\t}"
      : "=d"(v)
      : "r"(taddr));
  return v;
}
__device__ __forceinline__ void t9_st4(uint32_t taddr, double a, double b, double c, double d) { // columns 0..7
  asm volatile(
      "{\n\t.reg .b32 t<8>;\n\tmov.b64 {t0, t1}, %1;\n\tmov.b64 {t2, t3}, %2;\n\tmov.b64 {t4, t5}, %3;\n\tmov.b64 {t6, t7}, %4;\n\t"
      "tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {t0, t1, t2, t3, t4, t5, t6, t7};\n\t}" ::"r"(taddr),
      "d"(a), "d"(b), "d"(c), "d"(d));
}
__device__ __forceinline__ void t9_st2(uint32_t taddr, double a, double b) { // columns 0..3
  asm volatile(
      "{\n\t.reg .b32 t<4>;\n\tmov.b64 {t0, t1}, %1;\n\tmov.b64 {t2, t3}, %2;\n\t"
      "tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {t0, t1, t2, t3};\n\t}" ::"r"(taddr),
      "d"(a), "d"(b));
}
__device__ __forceinline__ void t9_st1(uint32_t taddr, double a) { // columns 0..1
  asm volatile("{\n\t.reg .b32 t<2>;\n\tmov.b64 {t0, t1}, %1;\n\ttcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {t0, t1};\n\t}" ::"r"(taddr), "d"(a));
}
__device__ __forceinline__ void t9_st8(uint32_t taddr, const double h[8]) { // columns 0..15
  asm volatile(
      "{\n\t.reg .b32 t<16>;\n\tmov.b64 {t0, t1}, %1;\n\tmov.b64 {t2, t3}, %2;\n\tmov.b64 {t4, t5}, %3;\n\tmov.b64 {t6, t7}, %4;\n\t"
      "mov.b64 {t8, t9}, %5;\n\tmov.b64 {t10, t11}, %6;\n\tmov.b64 {t12, t13}, %7;\n\tmov.b64 {t14, t15}, %8;\n\t"
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {t0, t1, t2, t3, t4, t5, t6, t7, t8, t9, t10, t11, t12, t13, t14, t15};\n\t}" ::"r"(taddr),
      "d"(h[0]), "d"(h[1]), "d"(h[2]), "d"(h[3]), "d"(h[4]), "d"(h[5]), "d"(h[6]), "d"(h[7]));
}
// two 16-column loads behind one wait
__device__ __forceinline__ void t9_ld8x2(uint32_t ta, uint32_t tb, double h[8], double g[8]) {
  asm volatile(
      "{\n\t.reg .b32 t<32>;\n\t"
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {t0, t1, t2, t3, t4, t5, t6, t7, t8, t9, t10, t11, t12, t13, t14, t15}, [%16];\n\t"
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {t16, t17, t18, t19, t20, t21, t22, t23, t24, t25, t26, t27, t28, t29, t30, t31}, [%17];\n\t"
      "tcgen05.wait::ld.sync.aligned;\n\t"
      "mov.b64 %0, {t0, t1};\n\tmov.b64 %1, {t2, t3};\n\tmov.b64 %2, {t4, t5};\n\tmov.b64 %3, {t6, t7};\n\t"
      "mov.b64 %4, {t8, t9};\n\tmov.b64 %5, {t10, t11};\n\tmov.b64 %6, {t12, t13};\n\tmov.b64 %7, {t14, t15};\n\t"
      "mov.b64 %8, {t16, t17};\n\tmov.b64 %9, {t18, t19};\n\tmov.b64 %10, {t20, t21};\n\tmov.b64 %11, {t22, t23};\n\t"
      "mov.b64 %12, {t24, t25};\n\tmov.b64 %13, {t26, t27};\n\tmov.b64 %14, {t28, t29};\n\tmov.b64 %15, {t30, t31};\n\t}"
      : "=d"(h[0]), "=d"(h[1]), "=d"(h[2]), "=d"(h[3]), "=d"(h[4]), "=d"(h[5]), "=d"(h[6]), "=d"(h[7]), "=d"(g[0]), "=d"(g[1]),
        "=d"(g[2]), "=d"(g[3]), "=d"(g[4]), "=d"(g[5]), "=d"(g[6]), "=d"(g[7])
      : "r"(ta), "r"(tb));
}
// two 8-column loads behind one wait
__device__ __forceinline__ void t9_ld4x2(uint32_t ta, uint32_t tb, double h[4], double g[4]) {
  asm volatile(
      "{\n\t.reg .b32 t<16>;\n\t"
      "tcgen05.ld.sync.aligned.32x32b.x8.b32 {t0, t1, t2, t3, t4, t5, t6, t7}, [%8];\n\t"
      "tcgen05.ld.sync.aligned.32x32b.x8.b32 {t8, t9, t10, t11, t12, t13, t14, t15}, [%9];\n\t"
      "tcgen05.wait::ld.sync.aligned;\n\t"
      "mov.b64 %0, {t0, t1};\n\tmov.b64 %1, {t2, t3};\n\tmov.b64 %2, {t4, t5};\n\tmov.b64 %3, {t6, t7};\n\t"
      "mov.b64 %4, {t8, t9};\n\tmov.b64 %5, {t10, t11};\n\tmov.b64 %6, {t12, t13};\n\tmov.b64 %7, {t14, t15};\n\t}"
      : "=d"(h[0]), "=d"(h[1]), "=d"(h[2]), "=d"(h[3]), "=d"(g[0]), "=d"(g[1]), "=d"(g[2]), "=d"(g[3])
      : "r"(ta), "r"(tb));
}
__device__ __forceinline__ void t9_ld4(uint32_t ta, double h[4]) {
  asm volatile(
      "{\n\t.reg .b32 t<8>;\n\t"
      "tcgen05.ld.sync.aligned.32x32b.x8.b32 {t0, t1, t2, t3, t4, t5, t6, t7}, [%4];\n\t"
      "tcgen05.wait::ld.sync.aligned;\n\t"
      "mov.b64 %0, {t0, t1};\n\tmov.b64 %1, {t2, t3};\n\tmov.b64 %2, {t4, t5};\n\tmov.b64 %3, {t6, t7};\n\t}"
      : "=d"(h[0]), "=d"(h[1]), "=d"(h[2]), "=d"(h[3])
      : "r"(ta));
}

// ---- accumulator slots of the z-mixed sums ("shift ring", V9_RING=1): position q = 0..5 of a vector field holds the
// partial sums {c0, c1, c2, info} of the output plane centre + q - 3; every stage adds this plane's contribution and
// moves each slot one position down, so all addresses are static.  Each block is ONE asm statement whose tcgen05.ld
// and tcgen05.st name the same registers: the sums are updated in place and the untouched `info` entries ride along
// (separate load/store statements made ptxas copy every slot through a staging block: 120 moves per cell).
#define T9_R8(o) "{t" #o "0, t" #o "1, t" #o "2, t" #o "3, t" #o "4, t" #o "5, t" #o "6, t" #o "7}"
#define T9_FMA(lo, hi, w, e) "mov.b64 d, {" lo ", " hi "};\n\tfma.rn.f64 d, " w ", " e ", d;\n\tmov.b64 {" lo ", " hi "}, d;\n\t"
// final = old[0] + w3 E3 -> scratch slot (+48); new[0] = old[1] + w2 E2, new[1] = old[2] + w1 E1
__device__ __forceinline__ void t9_ring_a(uint32_t b, double w1, double w2, double w3, const double (&E)[3][3]) {
  asm volatile(
      "{\n\t.reg .b32 ta<8>, tb<8>, tc<8>;\n\t.reg .f64 d;\n\t"
      "tcgen05.ld.sync.aligned.32x32b.x8.b32 {ta0, ta1, ta2, ta3, ta4, ta5, ta6, ta7}, [%0];\n\t"
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {tb0, tb1, tb2, tb3, tb4, tb5, tb6, tb7, tc0, tc1, tc2, tc3, tc4, tc5, tc6, tc7}, [%1];\n\t"
      "tcgen05.wait::ld.sync.aligned;\n\t"
      T9_FMA("ta0", "ta1", "%5", "%8") T9_FMA("ta2", "ta3", "%5", "%11") T9_FMA("ta4", "ta5", "%5", "%14")
      T9_FMA("tb0", "tb1", "%4", "%7") T9_FMA("tb2", "tb3", "%4", "%10") T9_FMA("tb4", "tb5", "%4", "%13")
      T9_FMA("tc0", "tc1", "%3", "%6") T9_FMA("tc2", "tc3", "%3", "%9") T9_FMA("tc4", "tc5", "%3", "%12")
      "tcgen05.st.sync.aligned.32x32b.x8.b32 [%2], {ta0, ta1, ta2, ta3, ta4, ta5, ta6, ta7};\n\t"
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {tb0, tb1, tb2, tb3, tb4, tb5, tb6, tb7, tc0, tc1, tc2, tc3, tc4, tc5, tc6, tc7};\n\t}"
      ::"r"(b), "r"(b + 8u), "r"(b + 48u), "d"(w1), "d"(w2), "d"(w3),
      "d"(E[0][0]), "d"(E[0][1]), "d"(E[0][2]), "d"(E[1][0]), "d"(E[1][1]), "d"(E[1][2]), "d"(E[2][0]), "d"(E[2][1]), "d"(E[2][2]));
}
// new[2] = old[3] + part, info; new[3] = old[4] - w1 E1
__device__ __forceinline__ void t9_ring_b(uint32_t b, double mw1, const double (&E)[3][3], const double part[3], double info) {
  asm volatile(
      "{\n\t.reg .b32 tb<8>, tc<8>;\n\t.reg .f64 d;\n\t"
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {tb0, tb1, tb2, tb3, tb4, tb5, tb6, tb7, tc0, tc1, tc2, tc3, tc4, tc5, tc6, tc7}, [%0];\n\t"
      "tcgen05.wait::ld.sync.aligned;\n\t"
      "mov.b64 d, {tb0, tb1};\n\tadd.rn.f64 d, d, %6;\n\tmov.b64 {tb0, tb1}, d;\n\t"
      "mov.b64 d, {tb2, tb3};\n\tadd.rn.f64 d, d, %7;\n\tmov.b64 {tb2, tb3}, d;\n\t"
      "mov.b64 d, {tb4, tb5};\n\tadd.rn.f64 d, d, %8;\n\tmov.b64 {tb4, tb5}, d;\n\t"
      "mov.b64 {tb6, tb7}, %9;\n\t"
      T9_FMA("tc0", "tc1", "%2", "%3") T9_FMA("tc2", "tc3", "%2", "%4") T9_FMA("tc4", "tc5", "%2", "%5")
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%1], {tb0, tb1, tb2, tb3, tb4, tb5, tb6, tb7, tc0, tc1, tc2, tc3, tc4, tc5, tc6, tc7};\n\t}"
      ::"r"(b + 24u), "r"(b + 16u), "d"(mw1), "d"(E[0][0]), "d"(E[1][0]), "d"(E[2][0]), "d"(part[0]), "d"(part[1]), "d"(part[2]), "d"(info));
}
// new[4] = old[5] - w2 E2; new[5] = -w3 E3
__device__ __forceinline__ void t9_ring_c(uint32_t b, double mw2, double mw3, const double (&E)[3][3]) {
  asm volatile(
      "{\n\t.reg .b32 tb<8>, tc<8>;\n\t.reg .f64 d;\n\t"
      "tcgen05.ld.sync.aligned.32x32b.x8.b32 {tb0, tb1, tb2, tb3, tb4, tb5, tb6, tb7}, [%0];\n\t"
      "mul.rn.f64 d, %3, %7;\n\tmov.b64 {tc0, tc1}, d;\n\t"
      "mul.rn.f64 d, %3, %8;\n\tmov.b64 {tc2, tc3}, d;\n\t"
      "mul.rn.f64 d, %3, %9;\n\tmov.b64 {tc4, tc5}, d;\n\t"
      "mov.b32 tc6, 0;\n\tmov.b32 tc7, 0;\n\t"
      "tcgen05.wait::ld.sync.aligned;\n\t"
      T9_FMA("tb0", "tb1", "%2", "%4") T9_FMA("tb2", "tb3", "%2", "%5") T9_FMA("tb4", "tb5", "%2", "%6")
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%1], {tb0, tb1, tb2, tb3, tb4, tb5, tb6, tb7, tc0, tc1, tc2, tc3, tc4, tc5, tc6, tc7};\n\t}"
      ::"r"(b + 40u), "r"(b + 32u), "d"(mw2), "d"(mw3), "d"(E[0][1]), "d"(E[1][1]), "d"(E[2][1]), "d"(E[0][2]), "d"(E[1][2]), "d"(E[2][2]));
}
// slot += a (the pressure force, known only after the entropy section, joins the centre slot at its new position)
__device__ __forceinline__ void t9_ring_add(uint32_t ta, const double a[3]) {
  asm volatile(
      "{\n\t.reg .b32 tb<8>;\n\t.reg .f64 d;\n\t"
      "tcgen05.ld.sync.aligned.32x32b.x8.b32 {tb0, tb1, tb2, tb3, tb4, tb5, tb6, tb7}, [%0];\n\t"
      "tcgen05.wait::ld.sync.aligned;\n\t"
      "mov.b64 d, {tb0, tb1};\n\tadd.rn.f64 d, d, %1;\n\tmov.b64 {tb0, tb1}, d;\n\t"
      "mov.b64 d, {tb2, tb3};\n\tadd.rn.f64 d, d, %2;\n\tmov.b64 {tb2, tb3}, d;\n\t"
      "mov.b64 d, {tb4, tb5};\n\tadd.rn.f64 d, d, %3;\n\tmov.b64 {tb4, tb5}, d;\n\t"
      "tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {tb0, tb1, tb2, tb3, tb4, tb5, tb6, tb7};\n\t}"
      ::"r"(ta), "d"(a[0]), "d"(a[1]), "d"(a[2]));
}
__device__ __forceinline__ void t9_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;"); }

// predicated global stores (the compiler turns `if (c) *p = v;` into a branch around the store, which costs a
// reconvergence point before the next warp-collective tcgen05 operation)
__device__ __forceinline__ void st_pred(double *p, double v, bool c) {
  asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %2, 0;\n\t@q st.global.f64 [%0], %1;\n\t}" ::"l"(p), "d"(v), "r"((int)c) : "memory");
}
__device__ __forceinline__ void st_pred_cs(double *p, double v, bool c) {
  asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %2, 0;\n\t@q st.global.cs.f64 [%0], %1;\n\t}" ::"l"(p), "d"(v), "r"((int)c) : "memory");
}

// df staging: global -> shared without a register destination (LDGSTS), issued at the top of a stage and read back
// right before the RK update.  A register load does not work here: with the register file full ptxas sinks the
// load down to its first use (measured: LDG 14 instructions before the DFMA, long_scoreboard 0.3 -> 1.3 per issue).
// bytes = 0 writes zeros and reads nothing.
__device__ __forceinline__ void cp_async8(uint32_t dst, const double *src, bool on) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(dst), "l"(src), "r"(on ? 8 : 0));
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
__device__ __forceinline__ double lds_volatile(uint32_t addr) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
  return v;
}

// exp by table: x = (32 m + j) ln2/32 + r, |r| <= ln2/64; exp(x) = 2^m 2^(j/32) e^r with a degree-6 polynomial
// (truncation 3.5e-18 relative) and the 32-entry table of 2^(j/32) (and its inverse at +32) in shared memory.
// ~2 ulp; 11 FP64 operations instead of 18 (exp_fast), and exp(-x) comes with 4 more instead of a reciprocal.
struct ExpRed {
  double r;
  int j, m;
};
__device__ __forceinline__ ExpRed exp_reduce(double x) {
  const double shifter = 6755399441055744.0; // 1.5 * 2^52: the low word of t is rint(x * 32/ln2)
  const double t = fma(x, cP.expk[0], shifter);
  const int k = __double2loint(t);
  const double kd = t - shifter;
  ExpRed e;
  e.r = fma(kd, cP.expk[2], fma(kd, cP.expk[1], x)); // ln2/32 in two parts (high part: 32 bits)
  e.j = k & 31;
  e.m = k >> 5;
  return e;
}
__device__ __forceinline__ double exp_scale(double v, int m) { // v * 2^m (v normal, result normal)
  return __hiloint2double(__double2hiint(v) + (m << 20), __double2loint(v));
}
// (all coefficients are constant-memory operands of the FMAs: cP.expc[n] = 1/n!)
__device__ __forceinline__ double exp_tab(double x, const double *tab) {
  const ExpRed e = exp_reduce(x);
  double p = cP.expc[6];
#pragma unroll
  for (int n = 5; n >= 0; --n) p = fma(p, e.r, cP.expc[n]);
  return exp_scale(p * tab[e.j], e.m);
}
// exp(x) and exp(-x): even and odd part of the polynomial
__device__ __forceinline__ void exp_tab_pm(double x, const double *tab, double &ep, double &em) {
  const ExpRed e = exp_reduce(x);
  const double r2 = e.r * e.r;
  const double pe = fma(r2, fma(r2, fma(r2, cP.expc[6], cP.expc[4]), cP.expc[2]), cP.expc[0]);
  const double po = e.r * fma(r2, fma(r2, cP.expc[5], cP.expc[3]), cP.expc[1]);
  ep = exp_scale((pe + po) * tab[e.j], e.m);
  em = exp_scale((pe - po) * tab[32 + e.j], -e.m);
}
// stencil weights applied as one FMA chain (one operation less than the balanced form of d2_line / SymSum::weighted)
__device__ __forceinline__ double d2_chain(double c, const double m[3], const double p[3]) {
  return fma(cP.d2c[0][0], c, fma(cP.d2c[0][1], p[0] + m[0], fma(cP.d2c[0][2], p[1] + m[1], cP.d2c[0][3] * (p[2] + m[2]))));
}
__device__ __forceinline__ double sym_chain(const SymSum &s, double c, double c0n) {
  return fma(c0n, c, fma(cP.d2c[0][1], s.s0, fma(cP.d2c[0][2], s.s1, cP.d2c[0][3] * s.s2)));
}

// taps of one line: m[i] / p[i] at -(i+1) / +(i+1) along x, y (shared memory, centre plane) or z (TMEM history)
__device__ __forceinline__ void v9_xline(const double *q, double m[3], double p[3]) {
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    m[i] = q[-(i + 1)];
    p[i] = q[i + 1];
  }
}
__device__ __forceinline__ void v9_yline(const double *q, double m[3], double p[3]) {
  if (PCB200_ABL & 16) { for (int i = 0; i < 3; ++i) { m[i] = q[-(i + 1)] * 1.5; p[i] = q[i + 1] * 0.5; } return; }
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    m[i] = q[-(i + 1) * V2_HX];
    p[i] = q[(i + 1) * V2_HX];
  }
}
__device__ __forceinline__ void v9_zline(const double h[7], double m[3], double p[3]) { // h = f(z-3..z+3)
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    m[i] = h[2 - i];
    p[i] = h[4 + i];
  }
}
// bidiagonal xy derivative in the centre plane (deriv.f90:1400-1452)
__device__ __forceinline__ double v9_dxy(const double *q) {
  if (PCB200_ABL & 4) return q[0];
  double g[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const int a = (i + 1), b = (i + 1) * V2_HX;
    g[i] = (q[a + b] - q[-a + b]) + (q[-a - b] - q[a - b]);
  }
  return cP.dmc[0][0] * g[0] + (cP.dmc[0][1] * g[1] + cP.dmc[0][2] * g[2]);
}

// Kernel arguments.  Global addresses are formed as (uniform per-variable base from the parameter bank) + (one
// per-thread element offset that advances by a plane per stage): the bases live in uniform registers / constant
// memory, so no per-variable 64-bit offsets are held in vector registers.  The bases of the variables that are
// finalised three planes late (u, and s with a magnetic field) are shifted by -3 planes.
struct V9Args {
  const double *fin;
  const double *dt_dev;
  double alpha, beta;
  int read_df, write_df;
  int z_lo, z_hi, lz;
  Forcing frc;
  double *df_k[MAXVAR];   // df + k*sv (- 3*sz if delayed): read and written
  double *fo_k[MAXVAR];   // f_out + k*sv (- 3*sz if delayed)
};

extern __shared__ __align__(1024) double v9_smem[];
template <int V> struct cuda_std_ic { static constexpr int value = V; }; // tag for the stage specialisations

template <bool ENT, bool MAG, bool HYP, bool UPW>
__global__ void __launch_bounds__(V2_THREADS, 1) rhs_update_v9(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ V9Args a) {
  typedef Slots<ENT> S;
  constexpr int NV = 4 + (ENT ? 1 : 0) + (MAG ? 3 : 0);
  constexpr int NGRP = MAG ? 2 : 1;   // vector fields with a grad div: u (and A)
  constexpr bool SDEL = ENT && MAG;   // ds/dt contains the Ohmic heating -> finalised three planes late as well
  typedef V9Layout<NV> L;
  const Pars &P = cP;

  double *planes = v9_smem;
  uint64_t *bars = reinterpret_cast<uint64_t *>(planes + V9_NSLOT * L::SLOT);
  uint32_t *tm_slot = reinterpret_cast<uint32_t *>(bars + V9_NSLOT);
  double *exptab = reinterpret_cast<double *>(bars + V9_NSLOT + 2); // 64 doubles: 2^(j/32), 2^(-j/32)
  double *dfs = exptab + 64;                                         // [NV][256]: df of this thread's cells

  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * V2_TX + tx;
  const int bx0 = blockIdx.x * V2_TX, by0 = blockIdx.y * V2_TY; // origin of the halo'd tile (incl. ghosts)
  const int zc0 = a.z_lo + blockIdx.z * a.lz;
  const int zc1 = min(zc0 + a.lz, a.z_hi);
  if (zc0 >= zc1) return;
  const int nst = (zc1 - zc0) + 2 * NG; // stages: centre planes zc0-3 .. zc1+2
  const int ix = bx0 + tx + NG, iy = by0 + ty + NG;
  const bool active = ix < P.nx + NG && iy < P.ny + NG;
  const int ixc = min(ix, P.nx + NG - 1), iyc = min(iy, P.ny + NG - 1);

  if (tid == 0) {
    if (smem_u32(planes) & 127u) __trap();
    for (int s = 0; s < V9_NSLOT; ++s) mbar_init(&bars[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (tid >= 64 && tid < 128) exptab[tid - 64] = P.exptab[tid - 64];
  if (tid < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tm_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

  auto issue = [&](int j) { // stage j <-> plane zc0-3+j, slot j % 6
    if (j >= nst) return;
    const int s = j % V9_NSLOT;
    mbar_expect_tx(&bars[s], L::TX_BYTES);
    tma_load_4d(planes + s * L::SLOT, &tmap, &bars[s], bx0, by0, zc0 - NG + j, 0);
  };
  if (tid == 0) {
#pragma unroll 1
    for (int j = 0; j < V9_NSLOT; ++j) issue(j);
  }

  // TMEM columns of this thread (two warps share each lane quarter -> 256 columns each):
  //   [0,96)    z histories: variable k at 12k, entries f(z-3..z+2)
  //   [96,160)  accumulator ring of u: the output plane of stage index i owns the 8 columns at 96 + 8 (i & 7):
  //             {U0, U1, U2, 1/rho}: sums for du/dt (z-mixed part scaled by nu/3); 1/rho from the centre stage on
  //   [160,224) accumulator ring of A at 160 + 8 (i & 7): {A0, A1, A2, eta mu0/(rho T)}: sums for mu0 j
  //   [224,256) {partial ds/dt, b_x, b_y, b_z} of the planes z-3..z: stage index i at 224 + 8 (i & 3)
  // warp-uniform by construction; the shuffle tells the compiler so (TMEM addresses are uniform-register operands:
  // without it every ring address is computed in vector registers and moved over with R2UR, 31 per stage)
  const uint32_t tmb = __shfl_sync(0xffffffffu, *tm_slot + ((uint32_t)((tid >> 5) & 3) << 21) + (uint32_t)(tid >> 7) * 256u, 0);
  #if V9_RING
  //   (shift ring: [96,152) u: six slots of 8 columns + the finished slot at +48, [152,208) A likewise, [208,240) pending)
  const uint32_t tmh = tmb, tmau = tmb + 96u, tmaa = tmb + 152u, tmp = tmb + 208u;
#else
  const uint32_t tmh = tmb, tmau = tmb + 96u, tmaa = tmb + 160u, tmp = tmb + 224u;
#endif

  const int toff = (ty + NG) * V2_HX + (tx + NG);
  const long long sv = P.sv, sz = P.sz;
  const long long col = ixc + P.sy * iyc;
  { // z histories of the first evaluated plane zc0: own column, planes zc0-3..zc0+2 (coalesced global loads)
    const double *g = a.fin + col + sz * (zc0 - NG);
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      double h[6];
#pragma unroll
      for (int d = 0; d < 6; ++d) h[d] = __ldg(g + k * sv + sz * d);
      t9_st4(tmh + 12u * k, h[0], h[1], h[2], h[3]);
      t9_st2(tmh + 12u * k + 8u, h[4], h[5]);
    }
  }

  if (a.frc.on) { // forcing.f90:2170-2345: Re, Im of fx(x) fy(y) of this thread's column -> TMEM columns 240..243
    const double2 fx = __ldg(a.frc.fx + ixc), fy = __ldg(a.frc.fy + iyc);
    t9_st2(tmb + 240u, fx.x * fy.x - fx.y * fy.y, fx.x * fy.y + fx.y * fy.x);
  }
  const uint32_t dfs_u32 = smem_u32(dfs) + (uint32_t)tid * 8u;
  auto dfv = [&](int k) { if (PCB200_ABL & 32) return a.alpha; return a.read_df ? lds_volatile(dfs_u32 + (uint32_t)k * (V2_THREADS * 8u)) : 0.0; };
  const double dtb = (*a.dt_dev) * a.beta;
  const bool rdf = a.read_df != 0 && active;
  // running element offset (advances by one plane per stage): cell (ix, iy) of the centre plane of the stage
  long long o = ix + P.sy * iy + sz * (zc0 - NG);
  auto dfn = [&](int) {};
  // 2N-RK3 update of one variable (timestep.f90:113,160-161); off = offset from the running pointers
  // (arithmetic unconditional, only the two stores are predicated: no divergent branch, no reconvergence)
  auto rk = [&](bool on, double r, double dfo, double fcv, int k, auto extra) {
    const double d = a.alpha * dfo + r;
    double fn = fcv + dtb * d;
    if (sizeof(extra) == sizeof(double)) fn += extra; // forcing (velocity only); callers without one pass an int
    if (PCB200_ABL & 8) { if (d == 1.2345 && fn == 6.789) st_pred(a.fo_k[k] + o, fn, on); return; }
    st_pred_cs(a.df_k[k] + o, d, on && a.write_df);
    st_pred(a.fo_k[k] + o, fn, on);
  };

  // Accumulator ring of one vector field (base: its TMEM columns): this plane is z+k of the output planes centre-k and
  // z-k of the output planes centre+k:  sum_k w_k [E_k(z+k) - E_k(z-k)]  ->  slot(centre-k) += w_k E_k,
  // slot(centre+k) -= w_k E_k.  Slot centre-3 is complete afterwards (it is read back and finalised at the end of the stage), the slot of
  // the centre plane is updated by ring_centre, the slot of plane centre+3 is started here.
  auto ring_scatter = [&](uint32_t base, int j, const double (&E)[3][3], const double w[3]) {
    if (PCB200_ABL & 1) return;
    auto at = [&](int d) { return base + 8u * (uint32_t)((j + d) & 7); };
    double s0[4], s1[4];
    t9_ld4x2(at(-3), at(-2), s0, s1);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      s0[c] = fma(w[2], E[c][2], s0[c]);
      s1[c] = fma(w[1], E[c][1], s1[c]);
    }
    t9_st4(at(-3), s0[0], s0[1], s0[2], s0[3]);
    t9_st4(at(-2), s1[0], s1[1], s1[2], s1[3]);
    t9_ld4x2(at(-1), at(1), s0, s1);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      s0[c] = fma(w[0], E[c][0], s0[c]);
      s1[c] = fma(-w[0], E[c][0], s1[c]);
    }
    t9_st4(at(-1), s0[0], s0[1], s0[2], s0[3]);
    t9_st4(at(1), s1[0], s1[1], s1[2], s1[3]);
    t9_ld4(at(2), s0);
#pragma unroll
    for (int c = 0; c < 3; ++c) s0[c] = fma(-w[1], E[c][1], s0[c]);
    t9_st4(at(2), s0[0], s0[1], s0[2], s0[3]);
    t9_st4(at(3), -w[2] * E[0][2], -w[2] * E[1][2], -w[2] * E[2][2], 0.0);
  };
  // the slot of the centre plane receives the plane's own part and its `info` entry
  auto ring_centre = [&](uint32_t base, int j, const double part[3], double info) {
    if (PCB200_ABL & 1) return;
    const uint32_t ta = base + 8u * (uint32_t)(j & 7);
    double s0[4];
    t9_ld4(ta, s0);
    t9_st4(ta, s0[0] + part[0], s0[1] + part[1], s0[2] + part[2], info);
  };
  const double wu[3] = {P.dmcu[0], P.dmcu[1], P.dmcu[2]}, wa[3] = {P.dmc[0][0], P.dmc[0][1], P.dmc[0][2]};

  int slot = 0, par = 0;
  // One stage = one centre plane.  MODE 0: warm-up (planes zc0-3..zc0-1: in-plane differences only), 1: evaluated
  // plane, 2: cool-down (planes zc1..zc1+2: differences + finalisation).  Three specialised copies, so the main
  // loop is straight-line code.
  auto stage = [&](auto mode_tag, const int j) {
    constexpr int MODE = decltype(mode_tag)::value;
    constexpr bool full = MODE == 1;
    const bool fin = MODE != 0 && j >= 2 * NG; // plane (centre - 3) >= zc0: finalise it (gates memory operations only)
    // slot release + TMA issue: the first warp waits until every warp has finished stage j-1, then refills its slot.
    // Eight rotating named barriers: a warp can be at most 5 stages ahead of the slowest one.
    if (tid < 32 && j > 0) {
      named_bar_sync(1 + ((j - 1) & 7), V2_THREADS);
      if (tid == 0) issue(j - 1 + V9_NSLOT);
      __syncwarp();
    }
    // early global loads: df of the planes written in this stage (L2-prefetched two stages earlier).  All at the
    // top of the stage: issued later they are still in flight when the shared-memory and TMEM loads that share their
    // scoreboard slots are consumed (measured: long_scoreboard 0.3 -> 1.5 stall cycles per issue).
    if (a.read_df && !(PCB200_ABL & 32)) {
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        const bool del = (k < 3) || (SDEL && k == S::ISS);
        const bool on = rdf && (del ? fin : full);
        cp_async8(dfs_u32 + (uint32_t)k * (V2_THREADS * 8u), a.df_k[k] + o, on);
#if V9_DFPF
        if (rdf && j + V9_DFPF < nst) asm volatile("prefetch.global.L2 [%0];" ::"l"(a.df_k[k] + (o + V9_DFPF * sz)));
#endif
      }
    }
    double2 fpq[3];
    if (MODE != 0 && a.frc.on && fin) { // Forcing::pq of the plane that is finalised in this stage
      const double2 *q2 = reinterpret_cast<const double2 *>(a.frc.pq) + 3 * (zc0 - 2 * NG + j);
#pragma unroll
      for (int i = 0; i < 3; ++i) fpq[i] = __ldg(q2 + i);
    }

    // the slot of this stage's centre plane was already waited for three stages ago (as the newest plane of an
    // evaluated stage), except in the first six stages of the chunk
    if (j < 2 * NG) mbar_wait(&bars[slot], par);
    const double *cp = planes + slot * L::SLOT + toff;
    t9_wait_st(); // TMEM stores of the previous stage are complete

    // in-plane differences of this plane: Eg[g][c][k-1], c = 0: Dx_k f_z, 1: Dy_k f_z, 2: Dx_k f_x + Dy_k f_y
    double Eg[NGRP][3][3];
    double part[NGRP][3];            // centre plane: everything else of du/dt (g = 0) and of mu0 j (g = 1)
    double rho1 = 0.0, hT = 0.0;     // 1/rho and eta mu0/(rho T) of the centre plane
    double fq[4] = {0., 0., 0., 0.}; // f(centre - 3) of the variables that are finalised late
#pragma unroll
    for (int g = 0; g < NGRP; ++g)
#pragma unroll
      for (int c = 0; c < 3; ++c) part[g][c] = 0.0;

    if (full) {
      const int slot3 = slot + 3 >= V9_NSLOT ? slot + 3 - V9_NSLOT : slot + 3;
      mbar_wait(&bars[slot3], slot + 3 >= V9_NSLOT ? par ^ 1 : par);
      const double *cp3 = planes + slot3 * L::SLOT + toff;
      // z history of variable k: f(z-3..z+2) from TMEM, f(z+3) from the newest plane; stored back shifted by one
      auto hist = [&](int k, double h[7]) {
        const uint32_t ta = tmh + 12u * k;
        t9_ld6(ta, h);
        h[6] = cp3[k * V2_PLANE];
#pragma unroll
        for (int e = 0; e < ((PCB200_ABL & 2) ? 1 : 6); ++e) t9_st1(ta + 2u * e, h[e + 1]);
      };
      double h[7], m[3], pp[3], dxx[3];
      double au[3] = {0., 0., 0.};
      if (UPW) {
#pragma unroll
        for (int i = 0; i < 3; ++i) au[i] = fabs(cp[(S::IUX + i) * V2_PLANE]);
      }
      const bool hyp_visc = HYP && (P.lvisc_hyper3_simplified || P.lvisc_hyper3_nu_const);

      // ---------------- lnrho (density.f90:2727-2811)
      hist(S::ILN, h);
      const double lnrho = h[3];
      double gl[3], del2lnrho = 0.0, del6lnrho = 0.0, upw_l = 0.0;
      {
        SymSum sl;
#pragma unroll
        for (int jd = 0; jd < 3; ++jd) {
          if (jd == 0) v9_xline(cp + S::ILN * V2_PLANE, m, pp);
          else if (jd == 1) v9_yline(cp + S::ILN * V2_PLANE, m, pp);
          else v9_zline(h, m, pp);
          gl[jd] = d1_line(m, pp, 0);
          if (ENT) sl.add(m, pp);
          if (HYP || UPW) {
            const double r6 = d6raw_line(lnrho, m, pp);
            if (HYP) del6lnrho += P.d6fac[0] * r6;
            if (UPW) upw_l += au[jd] * (P.d6upfac[0] * r6);
          }
        }
        if (ENT) del2lnrho = sym_chain(sl, lnrho, P.d2c0x3);
      }

      // ---------------- velocity (hydro.f90:3925-4063, sub.f90:2610-2749)
      double u[3], uij[3][3], del2u[3], gdu[3], hyp[3] = {0., 0., 0.}, upw_u[3] = {0., 0., 0.};
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        hist(S::IUX + i, h);
        u[i] = h[3];
        fq[i] = h[0];
        const double *q = cp + (S::IUX + i) * V2_PLANE;
        double d2s = 0.0;
        SymSum su;
#pragma unroll
        for (int jd = 0; jd < 3; ++jd) {
          if (jd == 0) v9_xline(q, m, pp);
          else if (jd == 1) v9_yline(q, m, pp);
          else v9_zline(h, m, pp);
          uij[i][jd] = d1_line(m, pp, 0);
          // in-plane differences of this plane feed the z-mixed sums of the planes z-3..z+3
          if (i == 2 && jd == 0) {
#pragma unroll
            for (int k = 0; k < 3; ++k) Eg[0][0][k] = pp[k] - m[k];
          }
          if (i == 2 && jd == 1) {
#pragma unroll
            for (int k = 0; k < 3; ++k) Eg[0][1][k] = pp[k] - m[k];
          }
          if (i == 0 && jd == 0) {
#pragma unroll
            for (int k = 0; k < 3; ++k) dxx[k] = pp[k] - m[k];
          }
          if (i == 1 && jd == 1) {
#pragma unroll
            for (int k = 0; k < 3; ++k) Eg[0][2][k] = dxx[k] + (pp[k] - m[k]);
          }
          if (jd != i) {
            su.add(m, pp); // the two off-diagonal second derivatives only enter through their sum
          } else {
            const double d2 = d2_chain(u[i], m, pp);
            d2s += d2;
            gdu[i] = d2; // fjji(i,i)=der2 (sub.f90:2651-2653)
          }
          if (HYP || UPW) {
            const double r6 = d6raw_line(u[i], m, pp);
            if (HYP) {
              hyp[i] += P.d6fac[0] * r6;                                       // del6v
              if (P.lvisc_hyper3_nu_const) hyp[i] += d5_line(m, pp, 0) * gl[jd]; // uij5glnrho
            }
            if (UPW) upw_u[i] += au[jd] * (P.d6upfac[0] * r6);
          }
        }
        d2s += sym_chain(su, u[i], P.d2c0x2);
        del2u[i] = d2s;
      }
#if !V9_RING
      ring_scatter(tmau, j, Eg[0], wu);
#endif
      // xy parts of grad div u (the z-mixed parts arrive through the accumulator ring)
      gdu[0] += v9_dxy(cp + (S::IUX + 1) * V2_PLANE);
      gdu[1] += v9_dxy(cp + (S::IUX + 0) * V2_PLANE);

      const double divu = uij[0][0] + uij[1][1] + uij[2][2];
      double sij[3][3];
#pragma unroll
      for (int jd = 0; jd < 3; ++jd) {
        sij[jd][jd] = uij[jd][jd] - (1. / 3.) * divu;
#pragma unroll
        for (int i = jd + 1; i < 3; ++i) {
          sij[i][jd] = .5 * (uij[i][jd] + uij[jd][i]);
          sij[jd][i] = sij[i][jd];
        }
      }
      double ru[3];
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        double ugu = u[0] * uij[i][0] + u[1] * uij[i][1] + u[2] * uij[i][2];
        if (UPW && P.lupw_uu) ugu -= upw_u[i];
        ru[i] = -ugu;
      }
      double uglnrho = u[0] * gl[0] + u[1] * gl[1] + u[2] * gl[2];
      if (UPW && P.lupw_lnrho) uglnrho -= upw_l;
      // viscosity (viscosity.f90:1633-1664, :1982-1989, :2250-2257)
      double sij2 = 0.0;
      if (P.lvisc_nu_const) {
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          const double sgl = sij[i][0] * gl[0] + sij[i][1] * gl[1] + sij[i][2] * gl[2];
          ru[i] += P.nu * (del2u[i] + 2. * sgl + (1. / 3.) * gdu[i]);
        }
        if (ENT) {
          sij2 = sij[0][0] * sij[0][0] + sij[1][1] * sij[1][1] + sij[2][2] * sij[2][2] +
                 2 * (sij[1][0] * sij[1][0] + sij[2][0] * sij[2][0] + sij[2][1] * sij[2][1]);
        }
      }
      if (hyp_visc) {
#pragma unroll
        for (int i = 0; i < 3; ++i) ru[i] += P.nu_hyper3 * hyp[i];
      }
#if V9_RING
      // everything of du/dt that is known now (advection, viscosity; isothermal pressure) -> accumulator slots of u;
      // with entropy the pressure force follows below (t9_ring_add)
      if (!ENT) {
#pragma unroll
        for (int jd = 0; jd < 3; ++jd) ru[jd] += -P.cs20 * gl[jd];
      }
      if (MAG) rho1 = exp_tab(-lnrho, exptab); // density.f90:2750
      if (!(PCB200_ABL & 1)) {
        t9_ring_a(tmau, wu[0], wu[1], wu[2], Eg[0]);
        t9_ring_b(tmau, -wu[0], Eg[0], ru, rho1);
        t9_ring_c(tmau, -wu[1], -wu[2], Eg[0]);
      }
#endif
      // continuity (density.f90:2887-3060, 3197-3288)
      double rl = (-divu) - uglnrho;
      if (HYP && P.ldiff_hyper3lnrho) rl += P.diffrho_hyper3 * del6lnrho;
      cp_async_wait_all(); // df of this stage (issued at its top)
      rk(active, rl, dfv(S::ILN), lnrho, S::ILN, 0);

      // eos + energy (eos_idealgas.f90:851-888, entropy.f90:3554-3876, 4852-4947)
      double rs = 0.0, TT1 = 0.0;
      if (ENT) {
        hist(S::ISS, h);
        const double ss = h[3];
        fq[3] = h[0];
        double gss[3], del2ss = 0.0, upw_s = 0.0;
        SymSum sss;
#pragma unroll
        for (int jd = 0; jd < 3; ++jd) {
          if (jd == 0) v9_xline(cp + S::ISS * V2_PLANE, m, pp);
          else if (jd == 1) v9_yline(cp + S::ISS * V2_PLANE, m, pp);
          else v9_zline(h, m, pp);
          gss[jd] = d1_line(m, pp, 0);
          sss.add(m, pp);
          if (UPW) upw_s += au[jd] * (P.d6upfac[0] * d6raw_line(ss, m, pp));
        }
        del2ss = sym_chain(sss, ss, P.d2c0x3);
        const double e = P.cv1 * ss + P.gamma_m1 * (lnrho - P.lnrho0);
        double ex, exm;
        exp_tab_pm(e, exptab, ex, exm);
        const double mcs2g = P.mcs20g1 * ex; // -cs2/gamma
        TT1 = P.TT0_1 * exm;                            // exp(-(lnTT0+e)) (eos_idealgas.f90:851-888)
        double g2 = 0.0;
#pragma unroll
        for (int jd = 0; jd < 3; ++jd) {
          const double glnTT = P.gamma_m1 * gl[jd] + P.cv1 * gss[jd];
          const double gsum = gl[jd] + glnTT;
#if V9_RING
          ru[jd] = mcs2g * gsum; // fpres = -cs2*(glnrho+glnTT)/gamma (entropy.f90:3596-3600)
#else
          ru[jd] = fma(mcs2g, gsum, ru[jd]);
#endif
          g2 = fma(gsum, glnTT, g2);
        }
        double ugss = u[0] * gss[0] + u[1] * gss[1] + u[2] * gss[2];
        if (UPW && P.lupw_ss) ugss -= upw_s;
        rs = -ugss;
        if (P.lviscosity_heat) rs += TT1 * (P.nu2 * sij2);
        if (P.lheatc_chiconst) rs += P.chicp * ((P.gamma_m1 * del2lnrho + P.cv1 * del2ss) + g2);
        if (!SDEL) rk(active, rs, dfv(S::ISS), ss, S::ISS, 0);
      } else {
#if !V9_RING
#pragma unroll
        for (int jd = 0; jd < 3; ++jd) ru[jd] += -P.cs20 * gl[jd];
#endif
      }
#if V9_RING
      if (ENT && !(PCB200_ABL & 1)) t9_ring_add(tmau + 16u, ru);
      hT = (MAG && ENT && P.lohmic_heat) ? P.etamu0 * (rho1 * TT1) : 0.0;
#else
      // everything of du/dt except the z-mixed part of grad div u and the Lorentz force -> accumulator ring of u
      if (MAG) rho1 = exp_tab(-lnrho, exptab); // density.f90:2750
      hT = (MAG && ENT && P.lohmic_heat) ? P.etamu0 * (rho1 * TT1) : 0.0;
      ring_centre(tmau, j, ru, rho1);
#endif

      // ---------------- magnetic (magnetic.f90:4279-5225, 5392-6640)
      if (MAG) {
        const bool upa = UPW && P.lupw_aa;
        double aij[3][3], dia[3], rest[3], upw_a[3] = {0., 0., 0.}, av[3];
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          hist(S::IAX + i, h);
          av[i] = h[3];
          const double *q = cp + (S::IAX + i) * V2_PLANE;
          SymSum sa;
#pragma unroll
          for (int jd = 0; jd < 3; ++jd) {
            if (jd == 0) v9_xline(q, m, pp);
            else if (jd == 1) v9_yline(q, m, pp);
            else v9_zline(h, m, pp);
            if (i == 2 && jd == 0) {
#pragma unroll
              for (int k = 0; k < 3; ++k) Eg[NGRP - 1][0][k] = pp[k] - m[k];
            }
            if (i == 2 && jd == 1) {
#pragma unroll
              for (int k = 0; k < 3; ++k) Eg[NGRP - 1][1][k] = pp[k] - m[k];
            }
            if (i == 0 && jd == 0) {
#pragma unroll
              for (int k = 0; k < 3; ++k) dxx[k] = pp[k] - m[k];
            }
            if (i == 1 && jd == 1) {
#pragma unroll
              for (int k = 0; k < 3; ++k) Eg[NGRP - 1][2][k] = dxx[k] + (pp[k] - m[k]);
            }
            aij[i][jd] = (i != jd || upa) ? d1_line(m, pp, 0) : 0.0;
            if (jd == i) dia[i] = d2_chain(av[i], m, pp);
            else sa.add(m, pp);
            if (UPW && jd != i) upw_a[i] += au[jd] * (P.d6upfac[0] * d6raw_line(av[i], m, pp)); // mask=i (magnetic.f90:6301)
          }
          rest[i] = sym_chain(sa, av[i], P.d2c0x2);
        }
        const double bb[3] = {aij[2][1] - aij[1][2], aij[0][2] - aij[2][0], aij[1][0] - aij[0][1]};
        // mu0 j = grad div A - del2 A: everything but the z-mixed parts
        part[NGRP - 1][0] = v9_dxy(cp + (S::IAX + 1) * V2_PLANE) - rest[0];
        part[NGRP - 1][1] = v9_dxy(cp + (S::IAX + 0) * V2_PLANE) - rest[1];
        part[NGRP - 1][2] = -rest[2];
        double fres[3] = {0., 0., 0.};
        if (P.lresi_eta_const) {
#pragma unroll
          for (int i = 0; i < 3; ++i) fres[i] = P.eta * (dia[i] + rest[i]);
        }
        double ra[3];
        if (!upa) { // dAdt = uxb + fres (magnetic.f90:6247)
          ra[0] = (u[1] * bb[2] - u[2] * bb[1]) + fres[0];
          ra[1] = (u[2] * bb[0] - u[0] * bb[2]) + fres[1];
          ra[2] = (u[0] * bb[1] - u[1] * bb[0]) + fres[2];
        } else { // upwinded induction (magnetic.f90:6257-6313)
#pragma unroll
          for (int jd = 0; jd < 3; ++jd) {
            double x = 0.0;
#pragma unroll
            for (int k = 0; k < 3; ++k)
              if (k != jd) x += u[k] * (aij[k][jd] - aij[jd][k]);
            ra[jd] = (x + upw_a[jd]) + fres[jd];
          }
        }
#pragma unroll
        for (int i = 0; i < 3; ++i) rk(active, ra[i], dfv(S::IAX + i), av[i], S::IAX + i, 0);
        // what the Lorentz force and the Ohmic heating of this plane need three stages from now
        t9_st4(tmp + 8u * (uint32_t)(j & 3), rs, bb[0], bb[1], bb[2]);
#if V9_RING
        if (!(PCB200_ABL & 1)) {
          t9_ring_a(tmaa, wa[0], wa[1], wa[2], Eg[NGRP - 1]);
          t9_ring_b(tmaa, -wa[0], Eg[NGRP - 1], part[NGRP - 1], hT);
          t9_ring_c(tmaa, -wa[1], -wa[2], Eg[NGRP - 1]);
        }
#else
        ring_scatter(tmaa, j, Eg[NGRP - 1], wa);
        ring_centre(tmaa, j, part[NGRP - 1], hT);
#endif
      }
    } else {
      // warm-up / cool-down stage: only the in-plane differences of this plane
      double m[3], pp[3], d[3];
#pragma unroll
      for (int g = 0; g < NGRP; ++g) {
        const double *q = cp + (g == 0 ? S::IUX : S::IAX) * V2_PLANE;
        v9_xline(q + 2 * V2_PLANE, m, pp);
#pragma unroll
        for (int k = 0; k < 3; ++k) Eg[g][0][k] = pp[k] - m[k];
        v9_yline(q + 2 * V2_PLANE, m, pp);
#pragma unroll
        for (int k = 0; k < 3; ++k) Eg[g][1][k] = pp[k] - m[k];
        v9_xline(q, m, pp);
#pragma unroll
        for (int k = 0; k < 3; ++k) d[k] = pp[k] - m[k];
        v9_yline(q + V2_PLANE, m, pp);
#pragma unroll
        for (int k = 0; k < 3; ++k) Eg[g][2][k] = d[k] + (pp[k] - m[k]);
      }
#if V9_RING
      {
        const double zero3[3] = {0., 0., 0.};
        t9_ring_a(tmau, wu[0], wu[1], wu[2], Eg[0]);
        t9_ring_b(tmau, -wu[0], Eg[0], zero3, 0.0);
        t9_ring_c(tmau, -wu[1], -wu[2], Eg[0]);
        if (MAG) {
          t9_ring_a(tmaa, wa[0], wa[1], wa[2], Eg[NGRP - 1]);
          t9_ring_b(tmaa, -wa[0], Eg[NGRP - 1], zero3, 0.0);
          t9_ring_c(tmaa, -wa[1], -wa[2], Eg[NGRP - 1]);
        }
      }
#else
      ring_scatter(tmau, j, Eg[0], wu);
      if (MAG) ring_scatter(tmaa, j, Eg[NGRP - 1], wa);
#endif
      if (MODE == 2) { // cool-down: the histories stopped at planes zc1-3..zc1+2
        const uint32_t e = 2u * (uint32_t)(j - (nst - NG));
#pragma unroll
        for (int i = 0; i < 3; ++i) fq[i] = t9_ld1(tmh + 12u * (S::IUX + i) + e);
        if (SDEL) fq[3] = t9_ld1(tmh + 12u * S::ISS + e);
      }
    }

    // the slots of plane centre-3 are complete: all six neighbouring planes have contributed
    double sA[4], sB[4];
    t9_wait_st();
    if (PCB200_ABL & 1) { for (int q = 0; q < 4; ++q) { sA[q] = 1.0; sB[q] = 1.0; } }
#if V9_RING
    else t9_ld4x2(tmau + 48u, tmaa + 48u, sA, sB);
#else
    else t9_ld4x2(tmau + 8u * (uint32_t)((j - 3) & 7), tmaa + 8u * (uint32_t)((j - 3) & 7), sA, sB);
#endif
    double rr[3] = {sA[0], sA[1], sA[2]};
    const double jq[3] = {P.mu01 * sB[0], P.mu01 * sB[1], P.mu01 * sB[2]};
    const double rho1q = sA[3], hTq = sB[3];

    // ---------------- finalise plane (centre - 3): all six neighbouring planes have contributed
    if (MODE != 0) {
      double rsq = 0.0;
      if (MAG) {
        double pd[4];
        { // {partial ds/dt, b} written in stage j-3
          t9_ld4(tmp + 8u * (uint32_t)((j + 1) & 3), pd);
        }
        const double *bq = pd + 1;
        if (P.llorentzforce) { // jxbr = rho1*(jj x bb) (magnetic.f90:4870-4888, 5531)
          rr[0] += rho1q * (jq[1] * bq[2] - jq[2] * bq[1]);
          rr[1] += rho1q * (jq[2] * bq[0] - jq[0] * bq[2]);
          rr[2] += rho1q * (jq[0] * bq[1] - jq[1] * bq[0]);
        }
        rsq = pd[0] + hTq * (jq[0] * jq[0] + jq[1] * jq[1] + jq[2] * jq[2]); // Ohmic heating (magnetic.f90:6138-6144)
      }
      double force[3] = {0., 0., 0.};
      if (a.frc.on && fin) { // force_i = Re(fx fy) P_i(z) + Im(fx fy) Q_i(z) (substep 3 only)
        double fxy[2];
        {
          double t4[4];
          t9_ld4(tmb + 240u, t4);
          fxy[0] = t4[0]; fxy[1] = t4[1];
        }
        force[0] = fxy[0] * fpq[0].x + fxy[1] * fpq[1].y;
        force[1] = fxy[0] * fpq[0].y + fxy[1] * fpq[2].x;
        force[2] = fxy[0] * fpq[1].x + fxy[1] * fpq[2].y;
      }
      if (!full) cp_async_wait_all();
#pragma unroll
      for (int i = 0; i < 3; ++i) rk(active && fin, rr[i], dfv(i), fq[i], S::IUX + i, force[i]);
      if (SDEL) rk(active && fin, rsq, dfv(S::ISS), fq[3], S::ISS, 0);
    }

    o += sz;
    if (++slot == V9_NSLOT) { slot = 0; par ^= 1; }
    if (j + 1 < nst && tid >= 32) named_bar_arrive(1 + (j & 7), V2_THREADS);
  };
  typedef cuda_std_ic<0> M0;
  typedef cuda_std_ic<1> M1;
  typedef cuda_std_ic<2> M2;
#pragma unroll 1
  for (int j = 0; j < NG; ++j) stage(M0(), j);
#pragma unroll 1
  for (int j = NG; j < nst - NG; ++j) stage(M1(), j);
#pragma unroll 1
  for (int j = nst - NG; j < nst; ++j) stage(M2(), j);
  t9_wait_st();
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (tid < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(*tm_slot) : "memory");
}

} // namespace pcb200
