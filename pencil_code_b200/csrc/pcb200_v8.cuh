// pcb200_v8.cuh -- kernel variant 8: variant 7 (all z histories in tensor memory) on a (32 x 12)-cell tile with
// 384 threads = 12 warps per SM instead of 8.  What makes it fit:
//   * registers: with the z histories in TMEM the cell evaluation compiles to 167 registers without spills
//     (384 x 168 = 64512 of the 65536 registers of an SM);
//   * shared memory: the ring only has to hold the centre plane and z+1..z+3 (in-plane taps, xy and leading z-mixed
//     taps); 6 slots (2 in flight) x {u box, A box} of 38 x 18 x 3 doubles + 3 {lnrho, ss} slots = 231.3 KB;
//   * TMEM: three warps share a lane quarter, 170 of the 512 columns each: the six ring fields keep a period-6
//     circular history stored 1.83 times (entries 0..10, so that the window z-3..z+2 is always contiguous: 22 columns),
//     lnrho / ss keep a shifted 7-entry history (16 columns): 164 columns per thread.
// Arithmetic and reference citations: pcb200_device.cuh / pcb200_fast.cuh; TMA/mbarrier/TMEM helpers: pcb200_v5.cuh.
#pragma once
#include "pcb200_v5.cuh"

namespace pcb200 {

constexpr int V8_TX = 32, V8_TY = 12;
constexpr int V8_HX = V8_TX + 2 * NG, V8_HY = V8_TY + 2 * NG; // 38 x 18
constexpr int V8_PLANE = V8_HX * V8_HY;                       // 684 doubles
constexpr int V8_GROUP = 2064;                                // 3 planes = 16416 B -> 16512 B (128-byte aligned)
constexpr int V8_NSLOT = 6, V8_NSCAL = 3;
constexpr int V8_THREADS = V8_TX * V8_TY;                     // 384
constexpr int V8_TMCOLS = 170;                                // TMEM columns per thread (3 warps per lane quarter)

template <bool ENT, bool MAG> struct V8Layout {
  static constexpr int NGROUP = MAG ? 2 : 1;
  static constexpr int NSCAL = ENT ? 2 : 1;
  static constexpr int SLOT = NGROUP * V8_GROUP;
  static constexpr int RING_DOUBLES = V8_NSLOT * SLOT;
  static constexpr int SCAL_SLOT = ENT ? 1376 : 688; // 10944 B -> 11008 B ; 5472 B -> 5504 B
  static constexpr int SCAL_DOUBLES = V8_NSCAL * SCAL_SLOT;
  static constexpr unsigned RING_TX_BYTES = NGROUP * 3 * V8_PLANE * 8;
  static constexpr unsigned SCAL_TX_BYTES = NSCAL * V8_PLANE * 8;
  static constexpr size_t SMEM_BYTES = (size_t)(RING_DOUBLES + SCAL_DOUBLES) * 8 + (V8_NSLOT + V8_NSCAL) * 8 + 64;
};

template <bool ENT, bool MAG> struct RingLines8 {
  typedef Slots<ENT> S;
  static constexpr bool TRAIL = true, ISO = true;
  static constexpr int NR = MAG ? 6 : 3;
  static __device__ __forceinline__ int dr(int) { return 0; }
  static __device__ __forceinline__ int pr(int) { return 0; }
  const double *pl[4]; // planes z, z+1, z+2, z+3 of ring field u_x at this thread's cell
  const double *scal;  // current {lnrho, ss} planes at this thread's cell
  uint32_t tm;         // TMEM address of this thread's columns: ring field r at 22 r, lnrho at 132, ss at 148
  uint32_t tw;         // 2 * (j0 mod 6): column offset of the window start (plane z-3) = slot of the new entry
  double T[MAG ? 2 : 1][3][3];

  static __device__ __forceinline__ bool is_ring(int k) { return k < 3 || (MAG && k >= S::IAX); }
  static __device__ __forceinline__ int ridx(int k) { return k < 3 ? k : 3 + (k - S::IAX); }
  static __device__ __forceinline__ int foff(int k) { return k < 3 ? k * V8_PLANE : V8_GROUP + (k - S::IAX) * V8_PLANE; }
  static __device__ __forceinline__ int goff(int g) { return g == 0 ? 0 : V8_GROUP; }

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
    const int fx = goff(g), fy = fx + V8_PLANE, fz = fy + V8_PLANE;
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const double *q = pl[1 + j];
      const int a = j + 1, b = (j + 1) * V8_HX;
      out[0] += cP.dmc[0][j] * (q[fz + a] - q[fz - a]);
      out[1] += cP.dmc[0][j] * (q[fz + b] - q[fz - b]);
      out[2] += cP.dmc[0][j] * ((q[fx + a] - q[fx - a]) + (q[fy + b] - q[fy - b]));
    }
  }
  __device__ __forceinline__ void trail_prologue(int g, const double *const q3[3]) {
    const int fx = goff(g), fy = fx + V8_PLANE, fz = fy + V8_PLANE;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      const double *q = q3[d];
#pragma unroll
      for (int k = 3 - d; k <= 3; ++k) {
        const int j = d - 3 + k, a = k, b = k * V8_HX;
        T[g][0][j] -= cP.dmc[0][k - 1] * (q[fz + a] - q[fz - a]);
        T[g][1][j] -= cP.dmc[0][k - 1] * (q[fz + b] - q[fz - b]);
        T[g][2][j] -= cP.dmc[0][k - 1] * ((q[fx + a] - q[fx - a]) + (q[fy + b] - q[fy - b]));
      }
    }
  }

  __device__ __forceinline__ double c(int k) const {
    if (is_ring(k)) return pl[0][foff(k)];
    return scal[k == S::ILN ? 0 : V8_PLANE];
  }
  __device__ __forceinline__ void line(int k, int dir, double m[3], double p[3]) {
    if (is_ring(k)) {
      const int f = foff(k);
      if (dir == 2) {
        const uint32_t a = tm + 22 * ridx(k) + tw;
        double h[6];
        tmem_ld6d(a, h);               // f(z-3..z+2)
        const double fresh = pl[3][f]; // f(z+3) replaces f(z-3) in the circular history
        m[0] = h[2]; m[1] = h[1]; m[2] = h[0];
        p[0] = h[4]; p[1] = h[5]; p[2] = fresh;
        tmem_st1d(a, fresh);
        if (tw < 10) tmem_st1d(a + 12, fresh); // second copy of entries 0..4 (tw is warp-uniform)
      } else {
        const double *q = pl[0] + f;
        const int s = dir == 0 ? 1 : V8_HX;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          m[i] = q[-(i + 1) * s];
          p[i] = q[(i + 1) * s];
        }
      }
    } else {
      if (dir == 2) { // shifted 7-entry history; entry 6 = f(z+4) follows at the end of the iteration
        const uint32_t a = tm + 132 + (k == S::ILN ? 0 : 16);
        double h[7];
        tmem_ld7d(a, h);
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          m[i] = h[2 - i];
          p[i] = h[4 + i];
        }
        tmem_st4d(a + 0, h[1], h[2], h[3], h[4]);
        tmem_st2d(a + 8, h[5], h[6]);
      } else {
        const double *q = scal + (k == S::ILN ? 0 : V8_PLANE);
        const int s = dir == 0 ? 1 : V8_HX;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          m[i] = q[-(i + 1) * s];
          p[i] = q[(i + 1) * s];
        }
      }
    }
  }
  __device__ __forceinline__ double dij(int k, int) const {
    const double *q = pl[0] + foff(k);
    double g[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const int a = (i + 1), b = (i + 1) * V8_HX;
      g[i] = (q[a + b] - q[-a + b]) + (q[-a - b] - q[a - b]);
    }
    return cP.dmc[0][0] * g[0] + (cP.dmc[0][1] * g[1] + cP.dmc[0][2] * g[2]);
  }
};

extern __shared__ __align__(1024) double v8_smem[];

template <bool ENT, bool MAG, bool HYP, bool UPW>
__global__ void __launch_bounds__(V8_THREADS, 1)
    rhs_update_v8(const __grid_constant__ CUtensorMap tmap_ring, const __grid_constant__ CUtensorMap tmap_scal, const V2Args a) {
  typedef V8Layout<ENT, MAG> L;
  typedef Slots<ENT> S;
  typedef RingLines8<ENT, MAG> TP;
  constexpr int NV = 4 + (ENT ? 1 : 0) + (MAG ? 3 : 0);
  double *ring = v8_smem;
  double *scal = ring + L::RING_DOUBLES;
  uint64_t *bar_ring = reinterpret_cast<uint64_t *>(scal + L::SCAL_DOUBLES);
  uint64_t *bar_scal = bar_ring + V8_NSLOT;
  uint32_t *tm_slot = reinterpret_cast<uint32_t *>(bar_scal + V8_NSCAL);

  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * V8_TX + tx;
  const int bx0 = blockIdx.x * V8_TX, by0 = blockIdx.y * V8_TY;
  const int zc0 = a.z_lo + blockIdx.z * a.lz;
  const int zc1 = min(zc0 + a.lz, a.z_hi);
  if (zc0 >= zc1) return;
  const int ix = bx0 + tx + NG, iy = by0 + ty + NG;
  const bool active = ix < cP.nx + NG && iy < cP.ny + NG;
  const int ixc = min(ix, cP.nx + NG - 1), iyc = min(iy, cP.ny + NG - 1);
  const int last_plane = zc1 - 1 + NG;

  if (tid == 0) {
    if (smem_u32(ring) & 127u) __trap();
    for (int s = 0; s < V8_NSLOT + V8_NSCAL; ++s) mbar_init(&bar_ring[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (tid < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tm_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

  auto issue_ring = [&](int j) { // ring plane j <-> z = zc0-3+j, slot j % 6
    const int p = zc0 - NG + j;
    if (p > last_plane) return;
    const int slot = j % V8_NSLOT;
    mbar_expect_tx(&bar_ring[slot], L::RING_TX_BYTES);
    tma_load_4d(ring + slot * L::SLOT, &tmap_ring, &bar_ring[slot], bx0, by0, p, S::IUX);
    if (MAG) tma_load_4d(ring + slot * L::SLOT + V8_GROUP, &tmap_ring, &bar_ring[slot], bx0, by0, p, S::IAX);
  };
  auto issue_scal = [&](int j, int slot) { // scalar plane j <-> z = zc0+j, slot j % 3
    const int p = zc0 + j;
    if (p >= zc1) return;
    mbar_expect_tx(&bar_scal[slot], L::SCAL_TX_BYTES);
    tma_load_4d(scal + slot * L::SCAL_SLOT, &tmap_scal, &bar_scal[slot], bx0, by0, p, S::ILN);
  };
  if (tid == 0) {
#pragma unroll 1
    for (int j = 0; j < V8_NSLOT; ++j) issue_ring(j);
#pragma unroll 1
    for (int j = 0; j < V8_NSCAL; ++j) issue_scal(j, j);
  }

  TP tp;
  tp.trail_zero();
  tp.tm = *tm_slot + ((uint32_t)((tid >> 5) & 3) << 21) + (uint32_t)(tid >> 7) * V8_TMCOLS;
  const int toff = (ty + NG) * V8_HX + (tx + NG);
  const long long sv = cP.sv, sz = cP.sz;
  const long long col = ixc + cP.sy * iyc;
  { // scalar histories f(z-3..z+3) of the first plane, straight from global (coalesced)
    const double *gl = a.fin + col + S::ILN * sv + sz * (zc0 - NG);
#pragma unroll
    for (int v = 0; v < (ENT ? 2 : 1); ++v) {
      double h[8];
#pragma unroll
      for (int d = 0; d < 7; ++d) h[d] = __ldg(gl + v * sv + sz * d);
      h[7] = 0.0;
#pragma unroll
      for (int d = 0; d < 2; ++d) tmem_st4d(tp.tm + 132 + 16 * v + 8 * d, h[4 * d], h[4 * d + 1], h[4 * d + 2], h[4 * d + 3]);
    }
  }
  // ---- chunk prologue on ring planes 0..5 (z-3..z+2): trailing halves of the z-mixed sums, ring histories
  {
#pragma unroll 1
    for (int d = 0; d < 6; ++d) mbar_wait(&bar_ring[d], 0);
    const double *const q3[3] = {ring + toff, ring + L::SLOT + toff, ring + 2 * L::SLOT + toff};
    tp.trail_prologue(0, q3);
    if (MAG) tp.trail_prologue(1, q3);
#pragma unroll
    for (int r = 0; r < TP::NR; ++r) {
      const int fo = r < 3 ? r * V8_PLANE : V8_GROUP + (r - 3) * V8_PLANE;
      double h[6];
#pragma unroll
      for (int d = 0; d < 6; ++d) h[d] = ring[d * L::SLOT + toff + fo];
      const uint32_t b = tp.tm + 22 * r; // entries 0..5 and their copies 6..10
      tmem_st4d(b + 0, h[0], h[1], h[2], h[3]);
      tmem_st2d(b + 8, h[4], h[5]);
      tmem_st4d(b + 12, h[0], h[1], h[2], h[3]);
      tmem_st1d(b + 20, h[4]);
    }
    tmem_wait_st();
    __syncthreads(); // planes 0..2 are dead now: their slots take planes 6..8
    if (tid == 0) {
      issue_ring(6);
      issue_ring(7);
      issue_ring(8);
    }
  }
  const double dtb = (*a.dt_dev) * a.beta;
  const bool rdf = a.read_df != 0 && active;

  const long long o0 = ix + cP.sy * iy + sz * zc0;
  double *pdf = a.df + o0;
  double *pfo = a.fout + o0;
  const double *pnx = a.fin + col + S::ILN * sv + sz * (zc0 + 4);
  int sc_slot = 0, sc_par = 0;
  int cs = 3;  // ring slot of the centre plane (plane j0+3)
  int w6 = 0;  // j0 mod 6
  double fxy_r = 0.0, fxy_i = 0.0;
  if (a.frc.on) {
    const double2 fx = a.frc.fx[ixc], fy = a.frc.fy[iyc];
    fxy_r = fx.x * fy.x - fx.y * fy.y;
    fxy_i = fx.x * fy.y + fx.y * fy.x;
  }

#pragma unroll 1
  for (int z = zc0; z < zc1; ++z) {
    const int j0 = z - zc0;
    // slot release + TMA issue by the first warp once every warp has finished plane z-1 (4 rotating named barriers)
    if (tid < 32 && j0 > 0) {
      named_bar_sync(1 + ((j0 - 1) & 3), V8_THREADS);
      if (tid == 0) {
        issue_ring(j0 + 8); // into the slot of plane j0+2, the centre plane of the previous iteration
        issue_scal(j0 + 2, sc_slot == 0 ? 2 : sc_slot - 1);
      }
      __syncwarp();
    }
    double dfv[NV];
    {
      const double *q = pdf;
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        dfv[k] = rdf ? __ldg(q) : 0.0;
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
    {
      const int jn = j0 + 6; // newest plane needed: z+3
      mbar_wait(&bar_ring[jn % V8_NSLOT], (jn / V8_NSLOT) & 1);
    }
    mbar_wait(&bar_scal[sc_slot], sc_par);
#pragma unroll
    for (int d = 0; d < 4; ++d) {
      int s = cs + d;
      if (s >= V8_NSLOT) s -= V8_NSLOT;
      tp.pl[d] = ring + s * L::SLOT + toff;
    }
    tp.scal = scal + sc_slot * L::SCAL_SLOT + toff;
    tp.tw = 2u * (uint32_t)w6;
    tmem_wait_st(); // history stores of the previous plane (or of the prologue) are complete

    double rhs[NV], fc[NV];
    // every thread evaluates the cell (tcgen05.ld/st are warp-collective); threads outside the grid work on zero-filled
    // TMA data and store nothing
    rhs_cell_lines<ENT, MAG, HYP, UPW>(tp, rhs, fc);
    if (active) {
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
    // newest scalar history entries, running pointers and counters
    tmem_st1d(tp.tm + 132 + 12, nl);
    if (ENT) tmem_st1d(tp.tm + 148 + 12, ns);
    pdf += sz;
    pfo += sz;
    pnx += sz;
    if (++sc_slot == V8_NSCAL) { sc_slot = 0; sc_par ^= 1; }
    if (++cs == V8_NSLOT) cs = 0;
    if (++w6 == 6) w6 = 0;
    if (z + 1 < zc1 && tid >= 32) named_bar_arrive(1 + (j0 & 3), V8_THREADS);
  }
  tmem_wait_st();
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (tid < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(*tm_slot) : "memory");
}

} // namespace pcb200
