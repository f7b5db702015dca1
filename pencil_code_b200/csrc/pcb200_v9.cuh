// pcb200_v9.cuh -- the building blocks of the "all-trail" z-marching scheme (TMEM scratch accessors, accumulator slots,
// table-based exp, line helpers) that kernel variant 10 (pcb200_v10.cuh) is made of.  The kernel they were developed in,
// variant 9 (per-thread LDGSTS/STG for df and the outputs, 15.06 G/s at 512^3), was removed once variant 10 had replaced it;
// its description stays because it is the description of the scheme:
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
template <int V> struct cuda_std_ic { static constexpr int value = V; }; // tag for the stage specialisations

} // namespace pcb200
