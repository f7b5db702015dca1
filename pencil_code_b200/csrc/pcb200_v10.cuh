// pcb200_v10.cuh -- kernel variant 10: variant 9 (z-marching, all-trail z-mixed derivatives, TMEM histories and
// accumulator slots; see pcb200_v9.cuh) with ALL global traffic of the loop moved to the TMA engine.
//
// Measured on variant 9 at 512^3 (scripts/abl.py, lab builds): without the 16 predicated STG and the 8 LDGSTS per cell
// and their 64-bit address arithmetic the kernel is 20 % faster, although DRAM is only 55 % busy -- the LSU path and
// the ordering constraints of the volatile store statements cost more than the bytes.  Here
//   * df is loaded into a shared-memory staging buffer by TMA (one 32 x 8 box per variable), updated IN PLACE by the
//     thread that owns the cell (LDS, FMA, STS to the same word) and written back by a TMA store from the same buffer;
//     the new f goes to a second staging buffer and leaves by TMA store as well.  Both buffers are double-buffered by
//     stage parity.  Tensor maps of the outputs end at the last interior cell, so partial tiles need no predicates:
//     the TMA engine clips them.
//   * shared memory for the staging buffers comes from the plane ring: variant 9 kept the planes z+1..z+3 resident
//     only to read the thread's own column of plane z+3 (the newest entry of its z history).  Now two rings: halo'd
//     (38 x 14) boxes for the centre plane (in use + 2 in flight) and interior-only (32 x 8) boxes of plane z+3
//     (in use + 2 in flight): 150 KB instead of 204 KB.  A plane is therefore fetched twice, three stages apart; the
//     second fetch hits L2 (reuse distance: 3 stages x ~15 MB over all SMs).
//   * a ninth warp is the producer: it waits (named barrier) until the eight compute warps have finished a stage,
//     issues that stage's stores, refills the ring slots the stage released and requests df one stage ahead.  A TMA
//     instruction costs its issuing warp 30-60 ns (measured: 26 of them per stage by a compute warp made the kernel
//     60 % slower), so they must not sit in a compute warp; 196 registers per thread leave room for the extra warp.
//     Variables that share a plane are moved as ONE box per array: on the device this variant stores the variables as
//     u, s, lnrho, A (Slots<ENT, true>), so the ones finished three planes late (u; s with a magnetic field) and the ones
//     finished at once (lnrho, A) are neighbours: 2 instead of 4 TMA operations per array and stage (measured: substep 2
//     9.6 -> 8.75 ms at 512^3; a TMA operation on a 2 KB box costs the engine as much as one on a 16 KB box,
//     profiles/r02_tma_rate.txt).
#pragma once
#include "pcb200_v9.cuh"

namespace pcb200 {

// Device layout of this variant (pcb200.cu: State::xpad): rows are padded by one double at either end, so the tensor
// maps see x_physical = x_logical + 1 and the first interior cell of a row (logical 3) is 16-byte aligned -- a TMA box
// of doubles must start at an even x (an odd start raises "illegal instruction", scripts/ubench/tma_align.cu).  The
// halo'd box therefore starts one cell early (physical bx0 = logical bx0 - 1) and is 40 wide instead of 38.
constexpr int V10_HX = V2_TX + 2 * NG + 2, V10_PLANE = V10_HX * V2_HY;
constexpr int V10_NH = 3; // halo'd ring: centre plane + 2 in flight
// interior ring: the box of plane centre+3 gives every thread the newest entries of its z histories; once read, the
// thread overwrites its own words with the new f of the stage, and the slot is the source of the f_out stores
// (same [variable][256] layout): in use, being stored, 2 in flight
constexpr int V10_NI = 4;
constexpr int V10_NS = 3; // df staging buffers (in, updated in place, out): stage j uses buffer j % 3
#ifndef V10_PWARP
#define V10_PWARP 1 // 1: producer warp in a third warpgroup, 0: first lane of the first compute warp (3x slower: lab only)
#endif
// With the producer warp the CTA has three warpgroups (384 threads, 168 registers each at launch).  The third one
// gives its registers back (setmaxnreg.dec 24) and the eight compute warps grow to 240 (setmaxnreg.inc): per SM
// sub-partition 2 x 32 x 240 + 32 x 24 = 16128 of 16384 registers.  Only its first warp works (the producer).
constexpr int V10_CTA = V10_PWARP ? 384 : V2_THREADS;      // threads per CTA
constexpr int V10_THREADS = V2_THREADS + 32 * V10_PWARP;   // participants of the stage barriers

template <int NV> struct V10Layout {
  static constexpr int HSLOT = ((NV * V10_PLANE + 15) / 16) * 16; // doubles per halo'd slot (128-byte aligned)
  static constexpr int ISLOT = NV * V2_THREADS;                  // doubles per interior slot
  static constexpr int STG = NV * V2_THREADS;                    // doubles per staging buffer
  static constexpr unsigned HTX = NV * V10_PLANE * 8, ITX = NV * V2_THREADS * 8;
  static constexpr size_t SMEM_BYTES = (size_t)(V10_NH * HSLOT + V10_NI * ISLOT + V10_NS * STG) * 8 + (V10_NH + V10_NI + V10_NS) * 8 + 64 + 64 * 8;
};

struct V10Args {
  const double *fin;
  const double *dt_dev;
  double alpha, beta;
  int read_df, write_df;
  int z_lo, z_hi, lz;
  int nch1, z_lo2, z_hi2; // a second z range served by the chunks blockIdx.z >= nch1 (the two boundary slabs of a z-decomposed run in one launch)
  // z-decomposed runs with the peer-put halo: ONE launch over all planes; the chunks run top-down (blockIdx.z = 0 is the
  // uppermost), and a CTA waits for the arrival counter of a neighbour (pcb200.cu: put_z) only when it is about to read
  // that neighbour's ghost planes -- the lowest chunks start when most of the kernel is done, the halo landed long before
  const unsigned long long *halo_flags; // [0]: from the lower neighbour (planes < NG), [1]: from the upper one; nullptr: no waits
  unsigned long long halo_seq;
  int nch;
  Forcing frc;
};

__device__ __forceinline__ void sts_volatile(uint32_t addr, double v) { asm volatile("st.shared.f64 [%0], %1;" ::"r"(addr), "d"(v) : "memory"); }
__device__ __forceinline__ void tma_store_4d(const CUtensorMap *map, const void *src, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(map), "r"(smem_u32(src)),
               "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}

__device__ __forceinline__ void v10_yline(const double *q, double m[3], double p[3]) {
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    m[i] = q[-(i + 1) * V10_HX];
    p[i] = q[(i + 1) * V10_HX];
  }
}
// bidiagonal xy derivative in the centre plane (deriv.f90:1400-1452)
__device__ __forceinline__ double v10_dxy(const double *q) {
  double g[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const int a = (i + 1), b = (i + 1) * V10_HX;
    g[i] = (q[a + b] - q[-a + b]) + (q[-a - b] - q[a - b]);
  }
  return cP.dmc[0][0] * g[0] + (cP.dmc[0][1] * g[1] + cP.dmc[0][2] * g[2]);
}

extern __shared__ __align__(1024) double v10_smem[];

// (x coordinates are physical) tm_h: f_in, box 40 x 14 x 1 x NV (halo'd planes); tm_i: f_in, box 32 x 8 x 1 x NV (interior boxes);
// dl/dn: df, boxes 32 x 8 x 1 x {variables finished late, variables finished at once}; ol/on: f_out likewise -- these four end
// at the last interior cell in x and y
struct V10Maps {
  CUtensorMap h, i, dl, dn, ol, on;
};
// device slots of this variant: u, [s,] lnrho, [A] (Slots<ENT, true>); the first v10_nlate() of them are finished three planes late
__host__ __device__ constexpr int v10_nlate(bool ent, bool mag) { return 3 + ((ent && mag) ? 1 : 0); }
template <bool ENT, bool MAG, bool HYP, bool UPW>
__global__ void __launch_bounds__(V10_CTA, 1) rhs_update_v10(const __grid_constant__ V10Maps tm, const __grid_constant__ V10Args a) {
  typedef Slots<ENT, true> S;
  constexpr int NV = 4 + (ENT ? 1 : 0) + (MAG ? 3 : 0);
  constexpr int NGRP = MAG ? 2 : 1;   // vector fields with a grad div: u (and A)
  constexpr bool SDEL = ENT && MAG;   // ds/dt contains the Ohmic heating -> finalised three planes late as well
  typedef V10Layout<NV> L;
  const Pars &P = cP;

  double *hring = v10_smem;
  double *iring = hring + V10_NH * L::HSLOT;
  double *dstg = iring + V10_NI * L::ISLOT; // [3][NV][256]: df in and out
  uint64_t *hbar = reinterpret_cast<uint64_t *>(dstg + V10_NS * L::STG);
  uint64_t *ibar = hbar + V10_NH;
  uint64_t *dbar = ibar + V10_NI;
  uint32_t *tm_slot = reinterpret_cast<uint32_t *>(dbar + V10_NS);
  double *exptab = reinterpret_cast<double *>(dbar + V10_NS + 2); // 64 doubles: 2^(j/32), 2^(-j/32)

  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * V2_TX + tx;
  const int bx0 = blockIdx.x * V2_TX, by0 = blockIdx.y * V2_TY; // origin of the halo'd tile (incl. ghosts)
  const bool second = (int)blockIdx.z >= a.nch1;
  const int chunk = a.halo_flags ? a.nch - 1 - (int)blockIdx.z : (int)blockIdx.z; // top-down when the halo is awaited in the kernel
  const int zc0 = second ? a.z_lo2 + ((int)blockIdx.z - a.nch1) * a.lz : a.z_lo + chunk * a.lz;
  const int zc1 = min(zc0 + a.lz, second ? a.z_hi2 : a.z_hi);
  if (zc0 >= zc1) return;
  const int nst = (zc1 - zc0) + 2 * NG; // stages: centre planes zc0-3 .. zc1+2
  const int ix = bx0 + tx + NG, iy = by0 + ty + NG;
  const int ixc = min(ix, P.nx + NG - 1), iyc = min(iy, P.ny + NG - 1);

  if (tid == 0) {
    if (smem_u32(hring) & 127u) __trap();
    for (int s = 0; s < V10_NH + V10_NI + V10_NS; ++s) mbar_init(&hbar[s], 1);
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

  // ---- producer side (first lane of the ninth warp) -----------------------------------------------------------
  auto halo_wait = [&](int plane) { // ghost planes of a z neighbour: wait until its put of this exchange has landed
    if (!a.halo_flags) return;
    const int d = plane < NG ? 0 : (plane >= NG + P.nz ? 1 : -1);
    if (d < 0) return;
    unsigned long long v;
    do {
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(a.halo_flags + d) : "memory");
    } while (v < a.halo_seq);
    asm volatile("fence.proxy.async;" ::: "memory"); // the planes were written by a copy engine: order the TMA reads behind the flag
  };
  auto load_h = [&](int j) { // halo'd box of the centre plane of stage j
    if (j >= nst) return;
    halo_wait(zc0 - NG + j);
    const int sh = j % V10_NH;
    mbar_expect_tx(&hbar[sh], L::HTX);
    tma_load_4d(hring + sh * L::HSLOT, &tm.h, &hbar[sh], bx0, by0, zc0 - NG + j, 0);
  };
  auto load_i = [&](int j) { // interior box of plane centre + 3 of stage j
    if (j >= nst) return;
    halo_wait(zc0 + j);
    const int si = j % V10_NI;
    mbar_expect_tx(&ibar[si], L::ITX);
    tma_load_4d(iring + si * L::ISLOT, &tm.i, &ibar[si], bx0 + NG + 1, by0 + NG, zc0 + j, 0);
  };
  // variable groups that share a plane: g = 0: the NLATE leading slots (u, and s with a magnetic field), finished three
  // planes late; g = 1: the others (lnrho, A; s without a magnetic field), finished at once.  One box per group and array.
  constexpr int NGR = 2, NLATE = v10_nlate(ENT, MAG);
  auto group = [&](int j, int g, int &k0, int &nv, int &plane) { // -> does stage j update this group?
    const bool del = g == 0;
    k0 = del ? 0 : NLATE;
    nv = del ? NLATE : NV - NLATE;
    plane = zc0 - NG + j - (del ? NG : 0);
    return del ? (j >= 2 * NG) : (j >= NG && j < nst - NG);
  };
  auto store_stage = [&](int j) { // results of stage j: staging buffers -> global (one bulk group)
    const int b = j % V10_NS;
#pragma unroll 1
    for (int g = 0; g < NGR; ++g) {
      int k0, nv, pl;
      if (!group(j, g, k0, nv, pl) || (PCB200_ABL & 8)) continue;
      tma_store_4d(g == 0 ? &tm.ol : &tm.on, iring + (j % V10_NI) * L::ISLOT + k0 * V2_THREADS, bx0 + NG + 1, by0 + NG, pl, k0);
      if (a.write_df) tma_store_4d(g == 0 ? &tm.dl : &tm.dn, dstg + b * L::STG + k0 * V2_THREADS, bx0 + NG + 1, by0 + NG, pl, k0);
    }
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  };
  auto request_df = [&](int j) { // df of the groups stage j updates -> dstg[j % 3]; always one arrival on dbar[j % 3]
    if (j >= nst) return;
    const int b = j % V10_NS;
    unsigned bytes = 0;
    int k0, nv, pl;
    if (a.read_df && !(PCB200_ABL & 32)) {
#pragma unroll 1
      for (int g = 0; g < NGR; ++g) bytes += group(j, g, k0, nv, pl) ? nv * V2_THREADS * 8u : 0u;
    }
    mbar_expect_tx(&dbar[b], bytes);
    if (bytes) {
#pragma unroll 1
      for (int g = 0; g < NGR; ++g)
        if (group(j, g, k0, nv, pl))
          tma_load_4d(dstg + b * L::STG + k0 * V2_THREADS, g == 0 ? &tm.dl : &tm.dn, &dbar[b], bx0 + NG + 1, by0 + NG, pl, k0);
    }
  };
  auto produce = [&](int j) { // top of stage j >= 1 (every compute warp has finished stage j - 1), by one lane
    store_stage(j - 1);
    // the buffers of stage j - 2 (interior slot, df staging) are free: all bulk groups but the newest have been read
    asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
    load_h(j - 1 + V10_NH);
    load_i(j - 2 + V10_NI);
    request_df(j + 1);
  };
  if (ty == V2_TY || (!V10_PWARP && tid == 0)) { // prologue requests
    if (tx == 0) {
#pragma unroll 1
      for (int j = 0; j < V10_NH; ++j) load_h(j);
#pragma unroll 1
      for (int j = 0; j < V10_NI - 1; ++j) load_i(j);
      request_df(0);
      request_df(1);
    }
  }
  if (V10_PWARP) {
    if (ty >= V2_TY) asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");
    else asm volatile("setmaxnreg.inc.sync.aligned.u32 240;");
  }
  if (V10_PWARP && ty > V2_TY) { // idle warps of the producer warpgroup
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    return;
  }
  if (V10_PWARP && ty == V2_TY) { // ---- the producer warp
#pragma unroll 1
    for (int j = 1; j <= nst; ++j) { // "top of stage j": every compute warp has finished stage j - 1
      named_bar_sync(1 + ((j - 1) & 7), V10_THREADS);
      if (tx == 0) produce(j);
      __syncwarp();
    }
    if (tx == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); // shared memory stays valid until read
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    return;
  }
  const int wid = tid >> 5;

  // TMEM columns of this thread (two warps share each lane quarter -> 256 columns each):
  //   [0,96)    z histories: variable k at 12k, entries f(z-3..z+2)
  //   [96,152)  accumulator slots of u by lag: position q = 0..5 (8 columns each: {U0, U1, U2, 1/rho}) holds the sums
  //             of the output plane centre + q - 3, the finished slot is parked at +48
  //   [152,208) accumulator slots of A likewise: {A0, A1, A2, eta mu0/(rho T)}: sums for mu0 j
  //   [208,240) {partial ds/dt, b_x, b_y, b_z} of the planes z-3..z: stage index i at 208 + 8 (i & 3)
  //   [240,244) Re, Im of fx(x) fy(y) (helical forcing)
  const uint32_t tmb = __shfl_sync(0xffffffffu, *tm_slot + ((uint32_t)((tid >> 5) & 3) << 21) + (uint32_t)(tid >> 7) * 256u, 0);
  const uint32_t tmh = tmb, tmau = tmb + 96u, tmaa = tmb + 152u, tmp = tmb + 208u;

  const int toff = (ty + NG) * V10_HX + (tx + NG + 1);
  const long long sv = P.sv, sz = P.sz;
  const long long col = ixc + P.sy * iyc;
  if (a.halo_flags && zc0 - NG < NG) { // the history loads below read the lower neighbour's planes
    unsigned long long v;
    do {
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(a.halo_flags) : "memory");
    } while (v < a.halo_seq);
  }
  { // z histories of the first evaluated plane zc0: own column, planes zc0-3..zc0+2 (coalesced global loads)
    const double *g = a.fin + col + sz * (zc0 - NG);
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      double h[6];
#pragma unroll
      for (int d = 0; d < 6; ++d) h[d] = a.halo_flags ? __ldcg(g + k * sv + sz * d) : __ldg(g + k * sv + sz * d);
      t9_st4(tmh + 12u * k, h[0], h[1], h[2], h[3]);
      t9_st2(tmh + 12u * k + 8u, h[4], h[5]);
    }
  }
  if (a.frc.on) { // forcing.f90:2170-2345: Re, Im of fx(x) fy(y) of this thread's column -> TMEM columns 240..243
    const double2 fx = __ldg(a.frc.fx + ixc), fy = __ldg(a.frc.fy + iyc);
    t9_st2(tmb + 240u, fx.x * fy.x - fx.y * fy.y, fx.x * fy.y + fx.y * fy.x);
  }
  const uint32_t dst_u32 = smem_u32(dstg) + (uint32_t)tid * 8u, fst_u32 = smem_u32(iring) + (uint32_t)tid * 8u;
  const double dtb = (*a.dt_dev) * a.beta;
  const double wu[3] = {P.dmcu[0], P.dmcu[1], P.dmcu[2]}, wa[3] = {P.dmc[0][0], P.dmc[0][1], P.dmc[0][2]};

  int hs = 0, is = 0, ds = 0;
  unsigned hpar = 0, ipar = 0, dpar = 0;
  // One stage = one centre plane.  MODE 0: warm-up (planes zc0-3..zc0-1: in-plane differences only), 1: evaluated
  // plane, 2: cool-down (planes zc1..zc1+2: differences + finalisation).  Three specialised copies, so the main
  // loop is straight-line code.
  auto stage = [&](auto mode_tag, const int j) {
    constexpr int MODE = decltype(mode_tag)::value;
    constexpr bool full = MODE == 1;
    const bool fin = MODE != 0 && j >= 2 * NG; // plane (centre - 3) >= zc0: finalise it (gates the forcing loads only)
    if (!V10_PWARP && j > 0 && wid == 0) { // without a producer warp: first warp (bulk groups are per thread)
      named_bar_sync(1 + ((j - 1) & 7), V10_THREADS);
      if (tid == 0) produce(j);
      __syncwarp();
    }
    double2 fpq[3];
    if (MODE != 0 && a.frc.on && fin) { // Forcing::pq of the plane that is finalised in this stage
      const double2 *q2 = reinterpret_cast<const double2 *>(a.frc.pq) + 3 * (zc0 - 2 * NG + j);
#pragma unroll
      for (int i = 0; i < 3; ++i) fpq[i] = __ldg(q2 + i);
    }
    // staging buffers of this stage: ds = j % 3, phase dpar of dbar[ds] (running counters)
    const uint32_t dbuf = dst_u32 + (uint32_t)ds * (L::STG * 8u), fbuf = fst_u32 + (uint32_t)is * (L::ISLOT * 8u);
    bool dwaited = false; // (compile-time after inlining: the first rk of a stage waits for the staging buffers)
    auto dready = [&]() {
      if (!dwaited) { // the df box has landed; the interior box too (its slot takes the new f)
        mbar_wait(&dbar[ds], dpar);
        if (!full) mbar_wait(&ibar[is], ipar);
        dwaited = true;
      }
    };
    // 2N-RK3 update of one variable (timestep.f90:113,160-161) in the staging buffers: df is read from and written
    // back to the same shared-memory word, the new f goes to the output buffer; both leave through TMA stores
    auto rk = [&](double r, double fcv, int k, auto extra) {
      dready();
      const uint32_t wd = dbuf + (uint32_t)k * (V2_THREADS * 8u);
      const double dfo = a.read_df ? lds_volatile(wd) : 0.0;
      const double d = a.alpha * dfo + r;
      double fn = fcv + dtb * d;
      if (sizeof(extra) == sizeof(double)) fn += extra; // forcing (velocity only); callers without one pass an int
      if (a.write_df) sts_volatile(wd, d);
      sts_volatile(fbuf + (uint32_t)k * (V2_THREADS * 8u), fn);
    };

    mbar_wait(&hbar[hs], hpar); // centre plane (with halo)
    const double *cp = hring + hs * L::HSLOT + toff;
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
      mbar_wait(&ibar[is], ipar); // interior box of plane centre + 3: this thread's newest history entries
      const double *cp3 = iring + is * L::ISLOT + tid;
      // z history of variable k: f(z-3..z+2) from TMEM, f(z+3) from the newest plane; stored back shifted by one
      auto hist = [&](int k, double h[7]) {
        const uint32_t ta = tmh + 12u * k;
        t9_ld6(ta, h);
        h[6] = cp3[k * V2_THREADS];
#pragma unroll
        for (int e = 0; e < ((PCB200_ABL & 2) ? 1 : 6); ++e) t9_st1(ta + 2u * e, h[e + 1]);
      };
      double h[7], m[3], pp[3], dxx[3];
      double au[3] = {0., 0., 0.};
      if (UPW) {
#pragma unroll
        for (int i = 0; i < 3; ++i) au[i] = fabs(cp[(S::IUX + i) * V10_PLANE]);
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
          if (jd == 0) v9_xline(cp + S::ILN * V10_PLANE, m, pp);
          else if (jd == 1) v10_yline(cp + S::ILN * V10_PLANE, m, pp);
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
        const double *q = cp + (S::IUX + i) * V10_PLANE;
        double d2s = 0.0;
        SymSum su;
#pragma unroll
        for (int jd = 0; jd < 3; ++jd) {
          if (jd == 0) v9_xline(q, m, pp);
          else if (jd == 1) v10_yline(q, m, pp);
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
      // xy parts of grad div u (the z-mixed parts arrive through the accumulator ring)
      gdu[0] += v10_dxy(cp + (S::IUX + 1) * V10_PLANE);
      gdu[1] += v10_dxy(cp + (S::IUX + 0) * V10_PLANE);

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
      // continuity (density.f90:2887-3060, 3197-3288)
      double rl = (-divu) - uglnrho;
      if (HYP && P.ldiff_hyper3lnrho) rl += P.diffrho_hyper3 * del6lnrho;
      rk(rl, lnrho, S::ILN, 0);

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
          if (jd == 0) v9_xline(cp + S::ISS * V10_PLANE, m, pp);
          else if (jd == 1) v10_yline(cp + S::ISS * V10_PLANE, m, pp);
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
          ru[jd] = mcs2g * gsum; // fpres = -cs2*(glnrho+glnTT)/gamma (entropy.f90:3596-3600)
          g2 = fma(gsum, glnTT, g2);
        }
        double ugss = u[0] * gss[0] + u[1] * gss[1] + u[2] * gss[2];
        if (UPW && P.lupw_ss) ugss -= upw_s;
        rs = -ugss;
        if (P.lviscosity_heat) rs += TT1 * (P.nu2 * sij2);
        if (P.lheatc_chiconst) rs += P.chicp * ((P.gamma_m1 * del2lnrho + P.cv1 * del2ss) + g2);
        if (!SDEL) rk(rs, ss, S::ISS, 0);
      } else {
      }
      if (ENT && !(PCB200_ABL & 1)) t9_ring_add(tmau + 16u, ru);
      hT = (MAG && ENT && P.lohmic_heat) ? P.etamu0 * (rho1 * TT1) : 0.0;

      // ---------------- magnetic (magnetic.f90:4279-5225, 5392-6640)
      if (MAG) {
        const bool upa = UPW && P.lupw_aa;
        double aij[3][3], dia[3], rest[3], upw_a[3] = {0., 0., 0.}, av[3];
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          hist(S::IAX + i, h);
          av[i] = h[3];
          const double *q = cp + (S::IAX + i) * V10_PLANE;
          SymSum sa;
#pragma unroll
          for (int jd = 0; jd < 3; ++jd) {
            if (jd == 0) v9_xline(q, m, pp);
            else if (jd == 1) v10_yline(q, m, pp);
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
        part[NGRP - 1][0] = v10_dxy(cp + (S::IAX + 1) * V10_PLANE) - rest[0];
        part[NGRP - 1][1] = v10_dxy(cp + (S::IAX + 0) * V10_PLANE) - rest[1];
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
        for (int i = 0; i < 3; ++i) rk(ra[i], av[i], S::IAX + i, 0);
        // what the Lorentz force and the Ohmic heating of this plane need three stages from now
        t9_st4(tmp + 8u * (uint32_t)(j & 3), rs, bb[0], bb[1], bb[2]);
        if (!(PCB200_ABL & 1)) {
          t9_ring_a(tmaa, wa[0], wa[1], wa[2], Eg[NGRP - 1]);
          t9_ring_b(tmaa, -wa[0], Eg[NGRP - 1], part[NGRP - 1], hT);
          t9_ring_c(tmaa, -wa[1], -wa[2], Eg[NGRP - 1]);
        }
      }
    } else {
      // warm-up / cool-down stage: only the in-plane differences of this plane
      double m[3], pp[3], d[3];
#pragma unroll
      for (int g = 0; g < NGRP; ++g) {
        const double *q = cp + (g == 0 ? S::IUX : S::IAX) * V10_PLANE;
        v9_xline(q + 2 * V10_PLANE, m, pp);
#pragma unroll
        for (int k = 0; k < 3; ++k) Eg[g][0][k] = pp[k] - m[k];
        v10_yline(q + 2 * V10_PLANE, m, pp);
#pragma unroll
        for (int k = 0; k < 3; ++k) Eg[g][1][k] = pp[k] - m[k];
        v9_xline(q, m, pp);
#pragma unroll
        for (int k = 0; k < 3; ++k) d[k] = pp[k] - m[k];
        v10_yline(q + V10_PLANE, m, pp);
#pragma unroll
        for (int k = 0; k < 3; ++k) Eg[g][2][k] = d[k] + (pp[k] - m[k]);
      }
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
    else t9_ld4x2(tmau + 48u, tmaa + 48u, sA, sB);
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
#pragma unroll
      for (int i = 0; i < 3; ++i) rk(rr[i], fq[i], S::IUX + i, force[i]);
      if (SDEL) rk(rsq, fq[3], S::ISS, 0);
    }

    if (++hs == V10_NH) { hs = 0; hpar ^= 1; }
    if (++is == V10_NI) { is = 0; ipar ^= 1; }
    if (++ds == V10_NS) { ds = 0; dpar ^= 1; }
    // results of this stage -> visible to the TMA stores the producer issues once every warp has arrived
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (V10_PWARP || (wid != 0 && j + 1 < nst)) named_bar_arrive(1 + (j & 7), V10_THREADS);
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
  __syncthreads(); // (the producer warp arrives here after its last stores have been read)
  if (!V10_PWARP && tid == 0) { // results of the last stage; shared memory must stay valid until the engine has read it
    store_stage(nst - 1);
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  }
  if (tid < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(*tm_slot) : "memory");
}

} // namespace pcb200
