// pcb200_v5.cuh -- kernel variant 5: the z-marching TMA ring with lead/trail z-mixed derivatives
// (variant 4), restructured after the ncu source page of variant 4 (profiles/r01_v5_ncu_full_256.txt):
//   * isotropic grids only (dx == dy == dz bit for bit): one set of stencil weights serves all
//     directions, 10 instead of 30 FP64 constants; anisotropic grids keep variant 4;
//   * running global pointers (one 64-bit add per variable instead of an index multiply + scale per
//     access); no L2 prefetch of df (with TMA two planes ahead it no longer pays) and plain read-only df
//     loads: 1468 -> ~1270 SASS instructions per cell and substep;
//   * the own-column value at z-3 of every ring field is carried in a register (it was loaded as the
//     z-2 tap one plane earlier), so a ring slot is free one plane sooner: with the same 8 slots and
//     3 scalar slots TMA now runs two planes ahead instead of one (7.5 % of the stall samples of the
//     first version sat in mbarrier waits), and 6 LDS per cell disappear;
//   * one TMA box per 3-vector (38 x 14 x 1 x 3) and one for {lnrho, ss}: the tighter padding pays for
//     the third scalar slot (230.7 KB of the 227 KiB = 232448 B an sm_100a CTA may use).
// Arithmetic and reference citations: pcb200_device.cuh / pcb200_fast.cuh.
#pragma once
#include "pcb200_fast.cuh"

namespace pcb200 {

constexpr int V5_GROUP = 1600; // 3 planes of 532 doubles = 12768 B -> 12800 B (TMA needs 128-byte aligned destinations)
constexpr int V5_NSLOT = 8;    // ring: planes z-2..z+3 in use + 2 in flight
constexpr int V5_NSCAL = 3;    // scalar planes: z in use + 2 in flight

template <bool ENT, bool MAG> struct V5Layout {
  static constexpr int NGROUP = MAG ? 2 : 1;
  static constexpr int NSCAL = ENT ? 2 : 1;
  static constexpr int SLOT = NGROUP * V5_GROUP;     // doubles per ring slot
  static constexpr int RING_DOUBLES = V5_NSLOT * SLOT;
  static constexpr int SCAL_SLOT = ENT ? 1072 : 544; // 8512 B -> 8576 B ; 4256 B -> 4352 B
  static constexpr int SCAL_DOUBLES = V5_NSCAL * SCAL_SLOT;
  static constexpr unsigned RING_TX_BYTES = NGROUP * 3 * V2_PLANE * 8;
  static constexpr unsigned SCAL_TX_BYTES = NSCAL * V2_PLANE * 8;
  static constexpr size_t SMEM_BYTES = (size_t)(RING_DOUBLES + SCAL_DOUBLES) * 8 + (V5_NSLOT + V5_NSCAL) * 8 + 64;
};

// ---- tensor memory (TMEM) as per-thread scratch: tcgen05.ld/st, shape 32x32b (lane = thread, 32-bit columns)
// The 32-bit halves are packed / unpacked inside the asm block (mov.b64 with register pairs costs nothing once ptxas
// has allocated aligned pairs), and the wait sits in the same block, so no use can be scheduled before it.
__device__ __forceinline__ void tmem_ld6d(uint32_t taddr, double h[6]) { // 6 doubles = columns 0..11 (x8 + x4)
  asm volatile(
      "{\n\t.reg .b32 t<12>;\n\t"
      "tcgen05.ld.sync.aligned.32x32b.x8.b32 {t0, t1, t2, t3, t4, t5, t6, t7}, [%6];\n\t"
      "tcgen05.ld.sync.aligned.32x32b.x4.b32 {t8, t9, t10, t11}, [%7];\n\t"
      "tcgen05.wait::ld.sync.aligned;\n\t"
      "mov.b64 %0, {t0, t1};\n\tmov.b64 %1, {t2, t3};\n\tmov.b64 %2, {t4, t5};\n\t"
      "mov.b64 %3, {t6, t7};\n\tmov.b64 %4, {t8, t9};\n\tmov.b64 %5, {t10, t11};\n\t}"
      : "=d"(h[0]), "=d"(h[1]), "=d"(h[2]), "=d"(h[3]), "=d"(h[4]), "=d"(h[5])
      : "r"(taddr), "r"(taddr + 8)
      : "memory");
}
__device__ __forceinline__ void tmem_ld7d(uint32_t taddr, double h[7]) { // 7 doubles (+1 spare) = columns 0..15
  double spare;
  asm volatile(
      "{\n\t.reg .b32 t<16>;\n\t"
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {t0, t1, t2, t3, t4, t5, t6, t7, t8, t9, t10, t11, t12, t13, t14, t15}, [%8];\n\t"
      "tcgen05.wait::ld.sync.aligned;\n\t"
      "mov.b64 %0, {t0, t1};\n\tmov.b64 %1, {t2, t3};\n\tmov.b64 %2, {t4, t5};\n\tmov.b64 %3, {t6, t7};\n\t"
      "mov.b64 %4, {t8, t9};\n\tmov.b64 %5, {t10, t11};\n\tmov.b64 %6, {t12, t13};\n\tmov.b64 %7, {t14, t15};\n\t}"
      : "=d"(h[0]), "=d"(h[1]), "=d"(h[2]), "=d"(h[3]), "=d"(h[4]), "=d"(h[5]), "=d"(h[6]), "=d"(spare)
      : "r"(taddr)
      : "memory");
  (void)spare;
}
__device__ __forceinline__ void tmem_st2d(uint32_t taddr, double a, double b) { // 2 doubles = 4 columns
  asm volatile(
      "{\n\t.reg .b32 t<4>;\n\tmov.b64 {t0, t1}, %1;\n\tmov.b64 {t2, t3}, %2;\n\t"
      "tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {t0, t1, t2, t3};\n\t}" ::"r"(taddr), "d"(a), "d"(b)
      : "memory");
}
__device__ __forceinline__ void tmem_st4d(uint32_t taddr, double a, double b, double c, double d) { // 4 doubles = 8 columns
  asm volatile(
      "{\n\t.reg .b32 t<8>;\n\tmov.b64 {t0, t1}, %1;\n\tmov.b64 {t2, t3}, %2;\n\tmov.b64 {t4, t5}, %3;\n\tmov.b64 {t6, t7}, %4;\n\t"
      "tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {t0, t1, t2, t3, t4, t5, t6, t7};\n\t}" ::"r"(taddr), "d"(a), "d"(b), "d"(c), "d"(d)
      : "memory");
}
__device__ __forceinline__ void tmem_st1d(uint32_t taddr, double a) { // 1 double = 2 columns
  asm volatile("{\n\t.reg .b32 t<2>;\n\tmov.b64 {t0, t1}, %1;\n\ttcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {t0, t1};\n\t}" ::"r"(taddr), "d"(a)
               : "memory");
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// tap policy (interface of SmemLines, pcb200_fast.cuh).  TM >= 1: the z histories of lnrho / ss live in TMEM (8 doubles
// = 16 columns per variable and thread, entries f(z-3..z+3) + 1 spare) instead of two 7-deep register pipelines.
// TM == 2: the own-column z taps f(z-3..z+2) of the six ring fields live there as well; only the newest tap f(z+3) is
// read from the ring: 4 LDS per field and plane less, no z-3 registers.  All eight histories are circular (32 columns
// per variable and thread = the whole 512-column TMEM for the two warps of a lane quarter).
template <bool ENT, bool MAG, int TM = 0> struct RingLines5 {
  typedef Slots<ENT> S;
  uint32_t tm; // TMEM address (lane base of the warp, first column) of this thread's lnrho history; ss at +16
  // TM == 2: circular histories, 8 entries per variable stored twice (entry e at 2e and 2e+16 of the variable's 32
  // columns) so that every window of consecutive planes is contiguous and nothing is ever shifted:
  uint32_t tw;    // column offset of the window start (plane z-3) in this iteration: 2*(j0 mod 8)
  uint32_t tnew;  // column offset of the entry that receives this iteration's new plane
  static constexpr bool TRAIL = true, ISO = true;
  static constexpr int NR = MAG ? 6 : 3;
  static __device__ __forceinline__ int dr(int) { return 0; }
  static __device__ __forceinline__ int pr(int) { return 0; }
  const double *pl[6]; // planes z-2..z+3 of ring field u_x at this thread's cell (centre: pl[2])
  const double *scal;  // current {lnrho, ss} planes at this thread's cell
  double lz[7], sz[7]; // register pipelines of lnrho / ss along z
  double zm3[NR];      // own-column values at z-3 of the ring fields
  double T[MAG ? 2 : 1][3][3];

  static __device__ __forceinline__ bool is_ring(int k) { return k < 3 || (MAG && k >= S::IAX); }
  static __device__ __forceinline__ int ridx(int k) { return k < 3 ? k : 3 + (k - S::IAX); }
  static __device__ __forceinline__ int foff(int k) { return k < 3 ? k * V2_PLANE : V5_GROUP + (k - S::IAX) * V2_PLANE; }
  static __device__ __forceinline__ int goff(int g) { return g == 0 ? 0 : V5_GROUP; }

  __device__ __forceinline__ void trail_zero() {
#pragma unroll
    for (int g = 0; g < (MAG ? 2 : 1); ++g)
#pragma unroll
      for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int j = 0; j < 3; ++j) T[g][c][j] = 0.0;
  }
  __device__ __forceinline__ void trail_pop(int g, double out[3]) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      out[c] = T[g][c][0];
      T[g][c][0] = T[g][c][1];
      T[g][c][1] = T[g][c][2];
      T[g][c][2] = 0.0;
    }
  }
  __device__ __forceinline__ void trail_add(int g, int c, int, const double m[3], const double p[3]) {
#pragma unroll
    for (int j = 0; j < 3; ++j) T[g][c][j] -= cP.dmc[0][j] * (p[j] - m[j]);
  }
  __device__ __forceinline__ void lead(int g, double out[3]) const {
    const int fx = goff(g), fy = fx + V2_PLANE, fz = fy + V2_PLANE;
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const double *q = pl[3 + j];
      const int a = j + 1, b = (j + 1) * V2_HX;
      out[0] += cP.dmc[0][j] * (q[fz + a] - q[fz - a]);
      out[1] += cP.dmc[0][j] * (q[fz + b] - q[fz - b]);
      out[2] += cP.dmc[0][j] * ((q[fx + a] - q[fx - a]) + (q[fy + b] - q[fy - b]));
    }
  }
  // first plane of a z chunk: contributions of the planes z-3..z-1 (q3[0..2]) to the outputs z..z+2
  __device__ __forceinline__ void trail_prologue(int g, const double *const q3[3]) {
    const int fx = goff(g), fy = fx + V2_PLANE, fz = fy + V2_PLANE;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      const double *q = q3[d];
#pragma unroll
      for (int k = 3 - d; k <= 3; ++k) {
        const int j = d - 3 + k, a = k, b = k * V2_HX;
        T[g][0][j] -= cP.dmc[0][k - 1] * (q[fz + a] - q[fz - a]);
        T[g][1][j] -= cP.dmc[0][k - 1] * (q[fz + b] - q[fz - b]);
        T[g][2][j] -= cP.dmc[0][k - 1] * ((q[fx + a] - q[fx - a]) + (q[fy + b] - q[fy - b]));
      }
    }
  }

  __device__ __forceinline__ double c(int k) const {
    if (is_ring(k)) return pl[2][foff(k)];
    if (TM) return scal[k == S::ILN ? 0 : V2_PLANE];
    return k == S::ILN ? lz[3] : sz[3];
  }
  // TMEM history of a scalar: taps of this plane, and the history shifted by one plane written back
  // (entry 6 = f(z+4) follows at the end of the iteration, when the global load has landed)
  __device__ __forceinline__ void tm_zline(int k, double m[3], double p[3]) {
    double h[7];
    if (TM == 2) { // circular: nothing to write back
      tmem_ld7d(tm + (k == S::ILN ? 0 : 32) + tw, h);
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        m[i] = h[2 - i];
        p[i] = h[4 + i];
      }
      return;
    }
    const uint32_t a = tm + (k == S::ILN ? 0 : 16);
    tmem_ld7d(a, h);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      m[i] = h[2 - i];
      p[i] = h[4 + i];
    }
    tmem_st4d(a + 0, h[1], h[2], h[3], h[4]);
    tmem_st2d(a + 8, h[5], h[6]);
  }
  __device__ __forceinline__ void line(int k, int dir, double m[3], double p[3]) {
    if (is_ring(k)) {
      const int f = foff(k);
      if (dir == 2 && TM == 2) {
        const uint32_t a = tm + 64 + 32 * ridx(k);
        double h[6];
        tmem_ld6d(a + tw, h);          // f(z-3..z+2)
        const double fresh = pl[5][f]; // f(z+3): the new entry of the circular history (both copies)
        m[0] = h[2]; m[1] = h[1]; m[2] = h[0];
        p[0] = h[4]; p[1] = h[5]; p[2] = fresh;
        tmem_st1d(a + tnew, fresh);
        tmem_st1d(a + tnew + 16, fresh);
      } else if (dir == 2) {
        m[0] = pl[1][f];
        m[1] = pl[0][f];
        m[2] = zm3[ridx(k)];
        zm3[ridx(k)] = m[1]; // f(z-2) is the z-3 tap of the next plane
#pragma unroll
        for (int i = 0; i < 3; ++i) p[i] = pl[3 + i][f];
      } else {
        const double *q = pl[2] + f;
        const int s = dir == 0 ? 1 : V2_HX;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          m[i] = q[-(i + 1) * s];
          p[i] = q[(i + 1) * s];
        }
      }
    } else {
      if (dir == 2) {
        if (TM) {
          tm_zline(k, m, p);
        } else {
          const double *r = (k == S::ILN) ? lz : sz;
#pragma unroll
          for (int i = 0; i < 3; ++i) {
            m[i] = r[2 - i];
            p[i] = r[4 + i];
          }
        }
      } else {
        const double *q = scal + (k == S::ILN ? 0 : V2_PLANE);
        const int s = dir == 0 ? 1 : V2_HX;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          m[i] = q[-(i + 1) * s];
          p[i] = q[(i + 1) * s];
        }
      }
    }
  }
  // bidiagonal xy derivative in the centre plane (the z-mixed ones are lead + trail)
  __device__ __forceinline__ double dij(int k, int) const {
    const double *q = pl[2] + foff(k);
    double g[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const int a = (i + 1), b = (i + 1) * V2_HX;
      g[i] = (q[a + b] - q[-a + b]) + (q[-a - b] - q[a - b]);
    }
    return cP.dmc[0][0] * g[0] + (cP.dmc[0][1] * g[1] + cP.dmc[0][2] * g[2]);
  }
};

extern __shared__ __align__(1024) double v5_smem[];

template <bool ENT, bool MAG, bool HYP, bool UPW, int TM = 0>
__global__ void __launch_bounds__(V2_THREADS, 1)
    rhs_update_v5(const __grid_constant__ CUtensorMap tmap_ring, const __grid_constant__ CUtensorMap tmap_scal, const V2Args a) {
  typedef V5Layout<ENT, MAG> L;
  typedef Slots<ENT> S;
  constexpr int NV = 4 + (ENT ? 1 : 0) + (MAG ? 3 : 0);
  double *ring = v5_smem;
  double *scal = ring + L::RING_DOUBLES;
  uint64_t *bar_ring = reinterpret_cast<uint64_t *>(scal + L::SCAL_DOUBLES);
  uint64_t *bar_scal = bar_ring + V5_NSLOT;

  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * V2_TX + tx;
  const int bx0 = blockIdx.x * V2_TX, by0 = blockIdx.y * V2_TY; // origin of the halo'd tile (incl. ghosts)
  const int zc0 = a.z_lo + blockIdx.z * a.lz;
  const int zc1 = min(zc0 + a.lz, a.z_hi);
  if (zc0 >= zc1) return;
  const int ix = bx0 + tx + NG, iy = by0 + ty + NG;
  const bool active = ix < cP.nx + NG && iy < cP.ny + NG;
  const int ixc = min(ix, cP.nx + NG - 1), iyc = min(iy, cP.ny + NG - 1); // clamped (register pipelines)
  const int last_plane = zc1 - 1 + NG;

  if (tid == 0) {
    if (smem_u32(ring) & 127u) __trap();
    for (int s = 0; s < V5_NSLOT + V5_NSCAL; ++s) mbar_init(&bar_ring[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  uint32_t *tm_slot = reinterpret_cast<uint32_t *>(bar_scal + V5_NSCAL);
  constexpr int TM_COLS = TM == 2 ? 512 : 64; // 16 (TM=1) or 32 (TM=2) columns per variable and thread, two warps share each lane quarter
  if (TM) {
    if (tid < 32) {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tm_slot)), "n"(TM_COLS) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  }
  __syncthreads();
  if (TM) asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

  auto issue_ring = [&](int j) { // ring plane j <-> z = zc0-3+j, slot j % 8
    const int p = zc0 - NG + j;
    if (p > last_plane) return;
    const int slot = j % V5_NSLOT;
    mbar_expect_tx(&bar_ring[slot], L::RING_TX_BYTES);
    tma_load_4d(ring + slot * L::SLOT, &tmap_ring, &bar_ring[slot], bx0, by0, p, S::IUX);
    if (MAG) tma_load_4d(ring + slot * L::SLOT + V5_GROUP, &tmap_ring, &bar_ring[slot], bx0, by0, p, S::IAX);
  };
  auto issue_scal = [&](int j, int slot) { // scalar plane j <-> z = zc0+j, slot j % 3
    const int p = zc0 + j;
    if (p >= zc1) return;
    mbar_expect_tx(&bar_scal[slot], L::SCAL_TX_BYTES);
    tma_load_4d(scal + slot * L::SCAL_SLOT, &tmap_scal, &bar_scal[slot], bx0, by0, p, S::ILN);
  };

  if (tid == 0) {
#pragma unroll 1
    for (int j = 0; j < V5_NSLOT; ++j) issue_ring(j);
#pragma unroll 1
    for (int j = 0; j < V5_NSCAL; ++j) issue_scal(j, j);
  }

  RingLines5<ENT, MAG, TM> tp;
  tp.trail_zero();
  if (TM) tp.tm = *tm_slot + ((uint32_t)((tid >> 5) & 3) << 21) + (uint32_t)(tid >> 7) * (TM_COLS / 2); // lane (warp%4)*32, column half
  const int toff = (ty + NG) * V2_HX + (tx + NG);
  const long long sv = cP.sv, sz = cP.sz;
  const long long col = ixc + cP.sy * iyc;
  { // register pipelines of the scalars (own column, straight from global; coalesced)
    const double *gl = a.fin + col + S::ILN * sv + sz * (zc0 - NG);
    if (TM) {
#pragma unroll
      for (int v = 0; v < (ENT ? 2 : 1); ++v) {
        double h[8];
#pragma unroll
        for (int d = 0; d < 7; ++d) h[d] = __ldg(gl + v * sv + sz * d);
        h[7] = 0.0;
#pragma unroll
        for (int d = 0; d < 2; ++d) {
          if (TM == 2) { // planes 0..6 at entries 0..6, both copies
            tmem_st4d(tp.tm + 32 * v + 8 * d, h[4 * d], h[4 * d + 1], h[4 * d + 2], h[4 * d + 3]);
            tmem_st4d(tp.tm + 32 * v + 16 + 8 * d, h[4 * d], h[4 * d + 1], h[4 * d + 2], h[4 * d + 3]);
          } else {
            tmem_st4d(tp.tm + 16 * v + 8 * d, h[4 * d], h[4 * d + 1], h[4 * d + 2], h[4 * d + 3]);
          }
        }
      }
    } else {
#pragma unroll
      for (int d = 0; d < 7; ++d) tp.lz[d] = __ldg(gl + sz * d);
      if (ENT) {
#pragma unroll
        for (int d = 0; d < 7; ++d) tp.sz[d] = __ldg(gl + sv + sz * d);
      }
    }
  }
  const double dtb = (*a.dt_dev) * a.beta;
  const bool rdf = a.read_df != 0 && active;

  // running pointers of this thread (all advance by one plane per iteration)
  const long long o0 = ix + cP.sy * iy + sz * zc0;
  double *pdf = a.df + o0;                                        // df(ix,iy,z,0)
  double *pfo = a.fout + o0;                                      // f_out(ix,iy,z,0)
  const double *pnx = a.fin + col + S::ILN * sv + sz * (zc0 + 4); // lnrho(ixc,iyc,z+4): next pipeline entry
  int sc_slot = 0, sc_par = 0; // slot and phase parity of the scalar plane of this iteration
  // helical forcing (substep 3 only): fx(x)*fy(y) is fixed for this thread, fz(z) is fetched at the top of each plane
  double fxy_r = 0.0, fxy_i = 0.0;
  if (a.frc.on) {
    const double2 fx = a.frc.fx[ixc], fy = a.frc.fy[iyc];
    fxy_r = fx.x * fy.x - fx.y * fy.y;
    fxy_i = fx.x * fy.y + fx.y * fy.x;
  }

#pragma unroll 1
  for (int z = zc0; z < zc1; ++z) {
    const int j0 = z - zc0;
    // slot release + TMA issue: the first warp waits until every warp has finished plane z-1, then refills the
    // slots that plane freed.  Named barriers 1..4 rotate: with TMA two planes ahead a warp can arrive for
    // planes z-1, z and z+1 before the slowest warp has arrived for z-1 (plane z+2 needs data that is only
    // requested after the barrier of z-1 completed), so three generations can be open at once.
    if (tid < 32 && j0 > 0) {
      named_bar_sync(1 + ((j0 - 1) & 3), V2_THREADS);
      if (tid == 0) {
        if (j0 == 1) issue_ring(V5_NSLOT); // slot 0 held plane z-3 of the first iteration (trail prologue)
        issue_ring(j0 + V5_NSLOT);
        issue_scal(j0 + 2, sc_slot == 0 ? 2 : sc_slot - 1);
      }
      __syncwarp();
    }
    double dfv[NV];
    {
      const double *q = pdf;
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        dfv[k] = rdf ? __ldg(q) : 0.0; // not ld.global.cs: evict-first lines were refetched (+6.7 GB DRAM reads in substep 3 at 512^3)
        q += sv;
      }
    }
    double2 fzv = make_double2(0.0, 0.0);
    if (a.frc.on) fzv = __ldg(a.frc.fz + z);
    double nl = 0.0, ns = 0.0;
    if (z + 4 < cP.mz) {
      nl = __ldg(pnx);
      if (ENT) ns = __ldg(pnx + sv);
    }
    if (j0 == 0) {
#pragma unroll 1
      for (int d = 0; d < 6; ++d) mbar_wait(&bar_ring[d], 0);
    }
    mbar_wait(&bar_ring[(j0 + 6) % V5_NSLOT], ((j0 + 6) / V5_NSLOT) & 1);
    mbar_wait(&bar_scal[sc_slot], sc_par);
#pragma unroll
    for (int d = 0; d < 6; ++d) tp.pl[d] = ring + ((j0 + 1 + d) % V5_NSLOT) * L::SLOT + toff;
    tp.scal = scal + sc_slot * L::SCAL_SLOT + toff;
    if (j0 == 0) {
      const double *const q3[3] = {ring + toff, tp.pl[0], tp.pl[1]}; // planes z-3 (slot 0), z-2, z-1
      tp.trail_prologue(0, q3);
      if (MAG) tp.trail_prologue(1, q3);
#pragma unroll
      for (int r = 0; r < RingLines5<ENT, MAG>::NR; ++r) {
        const int fo = r < 3 ? r * V2_PLANE : V5_GROUP + (r - 3) * V2_PLANE;
        if (TM == 2) { // own-column history f(z-3..z+2) = ring planes 0..5 at entries 0..5, both copies
          double h[6];
#pragma unroll
          for (int d = 0; d < 6; ++d) h[d] = ring[d * L::SLOT + toff + fo];
#pragma unroll
          for (int d = 0; d < 3; ++d) {
            tmem_st2d(tp.tm + 64 + 32 * r + 4 * d, h[2 * d], h[2 * d + 1]);
            tmem_st2d(tp.tm + 64 + 32 * r + 16 + 4 * d, h[2 * d], h[2 * d + 1]);
          }
        } else {
          tp.zm3[r] = q3[0][fo];
        }
      }
      if (TM == 2) tmem_wait_st();
    }

    if (TM) tmem_wait_st(); // the history stores of the previous plane (or of the prologue) are complete
    if (TM == 2) {
      tp.tw = 2u * (uint32_t)(j0 & 7);
      tp.tnew = 2u * (uint32_t)((j0 + 6) & 7); // ring fields: plane z+3
    }
    double rhs[NV], fc[NV];
    // TM: every thread evaluates the cell (tcgen05.ld/st are warp-collective; threads outside the grid work on
    // zero-filled TMA data and store nothing)
    if (TM) rhs_cell_lines<ENT, MAG, HYP, UPW>(tp, rhs, fc);
    if (active) {
      if (!TM) rhs_cell_lines<ENT, MAG, HYP, UPW>(tp, rhs, fc);
      double force[3] = {0., 0., 0.};
      if (a.frc.on) { // forcing.f90:2170-2345
        const double re = fxy_r * fzv.x - fxy_i * fzv.y, im = fxy_r * fzv.y + fxy_i * fzv.x;
#pragma unroll
        for (int j = 0; j < 3; ++j) force[j] = a.frc.fda[j] * (a.frc.coef1[j] * re - a.frc.coef2[j] * im);
      }
      double *qd = pdf, *qf = pfo;
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        const double d = a.alpha * dfv[k] + rhs[k]; // timestep.f90:113 (alpha=0 in substep 1)
        if (a.write_df) __stcs(qd, d);
        double fn = fc[k] + dtb * d; // timestep.f90:160-161
        if (k < 3) fn += force[k];
        *qf = fn;
        qd += sv;
        qf += sv;
      }
    }
    // advance the scalar pipelines and the running pointers
    if (TM == 2) { // scalars: plane z+4 = entry (j0+7) mod 8, both copies
      const uint32_t e = 2u * (uint32_t)((j0 + 7) & 7);
      tmem_st1d(tp.tm + e, nl);
      tmem_st1d(tp.tm + e + 16, nl);
      if (ENT) {
        tmem_st1d(tp.tm + 32 + e, ns);
        tmem_st1d(tp.tm + 32 + e + 16, ns);
      }
    } else if (TM) {
      tmem_st1d(tp.tm + 12, nl);
      if (ENT) tmem_st1d(tp.tm + 16 + 12, ns);
    } else {
#pragma unroll
      for (int d = 0; d < 6; ++d) {
        tp.lz[d] = tp.lz[d + 1];
        if (ENT) tp.sz[d] = tp.sz[d + 1];
      }
      tp.lz[6] = nl;
      if (ENT) tp.sz[6] = ns;
    }
    pdf += sz;
    pfo += sz;
    pnx += sz;
    if (++sc_slot == V5_NSCAL) { sc_slot = 0; sc_par ^= 1; }
    // this warp is done with plane z-2's ring slot and with this scalar slot
    if (z + 1 < zc1 && tid >= 32) named_bar_arrive(1 + (j0 & 3), V2_THREADS);
  }
  if (TM) {
    tmem_wait_st();
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (tid < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(*tm_slot), "n"(TM_COLS) : "memory");
  }
}

} // namespace pcb200
