// pcb200_v3.cuh -- kernel variant 3: the z-marching TMA ring of variant 2 with the physics split
// over two warp roles that share one (32x8)-cell tile:
//   role 0 (warps 0-7)  "hydro": u, lnrho, ss   -> d(u, lnrho, ss)/dt and their 2N-RK3 update
//   role 1 (warps 8-15) "mag":   A (+u centre)  -> dA/dt and its update; j x b and j^2 are handed to
//                                the hydro warps through a 8 KB shared-memory mailbox
// 512 threads x <=128 registers per SM (16 resident warps) instead of 256 x 255 (8 warps): variant 2
// was latency bound (ncu: 12.5 % warps active, shared-memory pipe 53 %, FP64 pipe 37 %).
//
// Shared memory (MHD+entropy): ring of 8 plane slots x {u group, A group}; one TMA box per group
// (38 x 14 x 1 x 3 variables = 12768 B, padded to 12800 B), 2 slots of the {lnrho, ss} group,
// the mailbox and 10 mbarriers: 230.2 KB of the 227 KiB (232448 B) an sm_100a CTA may use.
//
// Arithmetic and reference citations: pcb200_device.cuh / pcb200_fast.cuh (hydro_part, mag_part).
#pragma once
#include "pcb200_fast.cuh"

namespace pcb200 {

constexpr int V3_TX = 32, V3_TY = 8;
constexpr int V3_HX = V3_TX + 2 * NG, V3_HY = V3_TY + 2 * NG; // 38 x 14
constexpr int V3_PLANE = V3_HX * V3_HY;                       // 532 doubles, field stride inside a group
constexpr int V3_GROUP = 1600;                                // 3 planes = 12768 B -> 12800 B (128 B aligned)
constexpr int V3_NSLOT = 8;
constexpr int V3_CELLS = V3_TX * V3_TY;

template <bool ENT, bool MAG> struct V3Layout {
  static constexpr int NGROUP = MAG ? 2 : 1;
  static constexpr int NSCAL = ENT ? 2 : 1;
  static constexpr int SLOT = NGROUP * V3_GROUP;
  static constexpr int RING_DOUBLES = V3_NSLOT * SLOT;
  static constexpr int SCAL_SLOT = ENT ? 1072 : 544; // 8512 B -> 8576 B ; 4256 B -> 4352 B
  static constexpr int SCAL_DOUBLES = 2 * SCAL_SLOT;
  static constexpr int XCH_DOUBLES = MAG ? 4 * V3_CELLS : 0;
  static constexpr unsigned RING_TX_BYTES = NGROUP * 3 * V3_PLANE * 8;
  static constexpr unsigned SCAL_TX_BYTES = NSCAL * V3_PLANE * 8;
  static constexpr size_t SMEM_BYTES = (size_t)(RING_DOUBLES + SCAL_DOUBLES + XCH_DOUBLES) * 8 + (V3_NSLOT + 2) * 8 + 64;
  static constexpr int THREADS = V3_CELLS * (MAG ? 2 : 1);
};

// tap policy on the v3 layout (same interface as SmemLines)
template <bool ENT, bool MAG, bool ISO_ = false> struct V3Lines {
  typedef Slots<ENT> S;
  static constexpr bool TRAIL = false;
  static constexpr bool ISO = ISO_; // dx == dy == dz: one set of stencil weights
  static __device__ __forceinline__ int dr(int j) { return ISO ? 0 : j; }
  static __device__ __forceinline__ int pr(int p) { return ISO ? 0 : p; }
  __device__ __forceinline__ void trail_pop(int, double *) {}
  __device__ __forceinline__ void trail_add(int, int, int, const double *, const double *) {}
  __device__ __forceinline__ void lead(int, double *) const {}
  const double *pl[7]; // planes z-3..z+3, field u_x, this thread's cell
  const double *scal;  // current {lnrho, ss} planes, this thread's cell
  double lz[7], sz[7]; // z register pipelines of lnrho / ss (hydro role)

  static __device__ __forceinline__ bool is_ring(int k) { return k < 3 || (MAG && k >= S::IAX); }
  static __device__ __forceinline__ int roff(int k) { return k < 3 ? k * V3_PLANE : V3_GROUP + (k - S::IAX) * V3_PLANE; }
  static __device__ __forceinline__ int soff(int k) { return k == S::ILN ? 0 : V3_PLANE; }

  __device__ __forceinline__ double c(int k) const {
    if (is_ring(k)) return pl[3][roff(k)];
    return k == S::ILN ? lz[3] : sz[3];
  }
  __device__ __forceinline__ void line(int k, int dir, double m[3], double p[3]) const {
    if (is_ring(k)) {
      const int f = roff(k);
      if (dir == 2) {
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          m[i] = pl[2 - i][f];
          p[i] = pl[4 + i][f];
        }
      } else {
        const double *q = pl[3] + f;
        const int s = dir == 0 ? 1 : V3_HX;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          m[i] = q[-(i + 1) * s];
          p[i] = q[(i + 1) * s];
        }
      }
    } else {
      if (dir == 2) {
        const double *r = (k == S::ILN) ? lz : sz;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          m[i] = r[2 - i];
          p[i] = r[4 + i];
        }
      } else {
        const double *q = scal + soff(k);
        const int s = dir == 0 ? 1 : V3_HX;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          m[i] = q[-(i + 1) * s];
          p[i] = q[(i + 1) * s];
        }
      }
    }
  }
  __device__ __forceinline__ double dij(int k, int pair) const {
    const int f = roff(k);
    double g[3];
    if (pair == 0) {
      const double *q = pl[3] + f;
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        const int a = (i + 1), b = (i + 1) * V3_HX;
        g[i] = (q[a + b] - q[-a + b]) + (q[-a - b] - q[a - b]);
      }
    } else {
      const int a1 = pair == 1 ? V3_HX : 1;
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        const double *qp = pl[4 + i] + f, *qm = pl[2 - i] + f;
        const int a = (i + 1) * a1;
        g[i] = (qp[a] - qp[-a]) + (qm[-a] - qm[a]);
      }
    }
    return cP.dmc[pr(pair)][0] * g[0] + (cP.dmc[pr(pair)][1] * g[1] + cP.dmc[pr(pair)][2] * g[2]);
  }
};


extern __shared__ __align__(1024) double v3_smem[];

// named barriers (0 is __syncthreads): mailbox full / mailbox free / plane done (alternating ids,
// because a warp may run one plane ahead of the slowest one)
enum { BAR_MBOX_FULL = 1, BAR_MBOX_FREE = 2, BAR_PLANE_EVEN = 3, BAR_PLANE_ODD = 4 };

template <bool ENT, bool MAG, bool HYP, bool UPW, bool ISO = false>
__global__ void __launch_bounds__((V3Layout<ENT, MAG>::THREADS), 1)
    rhs_update_v3(const __grid_constant__ CUtensorMap tmap_ring, const __grid_constant__ CUtensorMap tmap_scal, const V2Args a) {
  typedef V3Layout<ENT, MAG> L;
  typedef Slots<ENT> S;
  double *ring = v3_smem;
  double *scal = ring + L::RING_DOUBLES;
  double *xch = scal + L::SCAL_DOUBLES;
  uint64_t *bar_ring = reinterpret_cast<uint64_t *>(xch + L::XCH_DOUBLES);
  uint64_t *bar_scal = bar_ring + V3_NSLOT;

  const int tx = threadIdx.x, ty = threadIdx.y, role = threadIdx.z;
  const int cell = ty * V3_TX + tx;
  const int tid = role * V3_CELLS + cell;
  const bool producer_warp = tid < 32;
  const int bx0 = blockIdx.x * V3_TX, by0 = blockIdx.y * V3_TY;
  const int zc0 = a.z_lo + blockIdx.z * a.lz;
  const int zc1 = min(zc0 + a.lz, a.z_hi);
  if (zc0 >= zc1) return;
  const int ix = bx0 + tx + NG, iy = by0 + ty + NG;
  const bool active = ix < cP.nx + NG && iy < cP.ny + NG;
  const int ixc = min(ix, cP.nx + NG - 1), iyc = min(iy, cP.ny + NG - 1);
  const int last_plane = zc1 - 1 + NG;

  if (tid == 0) {
    if (smem_u32(ring) & 127u) __trap();
    for (int s = 0; s < V3_NSLOT; ++s) mbar_init(&bar_ring[s], 1);
    mbar_init(&bar_scal[0], 1);
    mbar_init(&bar_scal[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();

  auto issue_ring = [&](int j) { // ring plane j <-> z = zc0-3+j
    const int p = zc0 - NG + j;
    if (p > last_plane) return;
    const int slot = j % V3_NSLOT;
    mbar_expect_tx(&bar_ring[slot], L::RING_TX_BYTES);
    tma_load_4d(ring + slot * L::SLOT, &tmap_ring, &bar_ring[slot], bx0, by0, p, S::IUX);
    if (MAG) tma_load_4d(ring + slot * L::SLOT + V3_GROUP, &tmap_ring, &bar_ring[slot], bx0, by0, p, S::IAX);
  };
  auto issue_scal = [&](int j) { // scalar plane j <-> z = zc0+j
    const int p = zc0 + j;
    if (p >= zc1) return;
    const int slot = j & 1;
    mbar_expect_tx(&bar_scal[slot], L::SCAL_TX_BYTES);
    tma_load_4d(scal + slot * L::SCAL_SLOT, &tmap_scal, &bar_scal[slot], bx0, by0, p, S::ILN);
  };
  if (tid == 0) {
#pragma unroll 1
    for (int j = 0; j < 7; ++j) issue_ring(j);
    issue_scal(0);
  }

  V3Lines<ENT, MAG, ISO> tp;
  const int toff = (ty + NG) * V3_HX + (tx + NG);
  const long long col = ixc + cP.sy * iyc;
  if (role == 0) {
    const double *gl = a.fin + col + S::ILN * cP.sv;
#pragma unroll
    for (int d = 0; d < 7; ++d) tp.lz[d] = __ldg(gl + cP.sz * (zc0 - NG + d));
    if (ENT) {
      const double *gs = a.fin + col + S::ISS * cP.sv;
#pragma unroll
      for (int d = 0; d < 7; ++d) tp.sz[d] = __ldg(gs + cP.sz * (zc0 - NG + d));
    }
  }
  const double dtb = (*a.dt_dev) * a.beta;
  const bool rdf = a.read_df != 0 && active;
  constexpr int NVH = 4 + (ENT ? 1 : 0); // variables updated by the hydro role (slots 0..NVH-1)

#pragma unroll 1
  for (int z = zc0; z < zc1; ++z) {
    const int j0 = z - zc0;
    // ---- slot release + TMA issue: only the producer warp waits for the whole CTA
    if (producer_warp) {
      if (j0 > 0) named_bar_sync(BAR_PLANE_EVEN + ((j0 - 1) & 1), L::THREADS);
      if (tid == 0) {
        issue_ring(j0 + 7);
        issue_scal(j0 + 1);
      }
      __syncwarp();
    }
    const long long o = ix + cP.sy * iy + cP.sz * z;
    if (j0 == 0) {
#pragma unroll 1
      for (int d = 0; d < 6; ++d) mbar_wait(&bar_ring[d], 0);
    }
#pragma unroll
    for (int d = 0; d < 7; ++d) tp.pl[d] = ring + ((j0 + d) % V3_NSLOT) * L::SLOT + toff;
    tp.scal = scal + (j0 & 1) * L::SCAL_SLOT + toff;

    if (role == 0) {
      // ------------------------------------------------------------------ hydro role
      double dfv[NVH];
#pragma unroll
      for (int k = 0; k < NVH; ++k) dfv[k] = rdf ? __ldcs(a.df + o + k * cP.sv) : 0.0;
      double nl = 0.0, ns = 0.0;
      if (z + 4 < cP.mz) {
        nl = __ldg(a.fin + col + S::ILN * cP.sv + cP.sz * (z + 4));
        if (ENT) ns = __ldg(a.fin + col + S::ISS * cP.sv + cP.sz * (z + 4));
      }
      mbar_wait(&bar_ring[(j0 + 6) % V3_NSLOT], ((j0 + 6) / V3_NSLOT) & 1);
      mbar_wait(&bar_scal[j0 & 1], (j0 >> 1) & 1);
      double rhs[NVH], fc[NVH];
      HydroOut ho;
      if (active) hydro_part<ENT, HYP, UPW, V3Lines<ENT, MAG, ISO>, !MAG>(tp, rhs, fc, ho);
      if (MAG) {
        named_bar_sync(BAR_MBOX_FULL, L::THREADS); // mailbox of this plane is full
        if (active) { // Lorentz force + (nu/3) mixed part of grad div u; Ohmic heating / TT1
#pragma unroll
          for (int i = 0; i < 3; ++i) rhs[S::IUX + i] += xch[i * V3_CELLS + cell];
          if (ENT) rhs[S::ISS] += xch[3 * V3_CELLS + cell] * ho.TT1;
        }
        if (z + 1 < zc1) named_bar_arrive(BAR_MBOX_FREE, L::THREADS);
      }
      if (active) {
        double force[3] = {0., 0., 0.};
        if (a.frc.on) {
          const double2 fx = a.frc.fx[ix], fy = a.frc.fy[iy], fz = a.frc.fz[z];
          const double xr = fx.x * fy.x - fx.y * fy.y, xi = fx.x * fy.y + fx.y * fy.x;
          const double re = xr * fz.x - xi * fz.y, im = xr * fz.y + xi * fz.x;
#pragma unroll
          for (int j = 0; j < 3; ++j) force[j] = a.frc.fda[j] * (a.frc.coef1[j] * re - a.frc.coef2[j] * im);
        }
#pragma unroll
        for (int k = 0; k < NVH; ++k) {
          const long long q = o + k * cP.sv;
          const double d = a.alpha * dfv[k] + rhs[k];
          if (a.write_df) __stcs(a.df + q, d);
          double fn = fc[k] + dtb * d;
          if (k < 3) fn += force[k];
          a.fout[q] = fn;
        }
      }
#pragma unroll
      for (int d = 0; d < 6; ++d) {
        tp.lz[d] = tp.lz[d + 1];
        if (ENT) tp.sz[d] = tp.sz[d + 1];
      }
      tp.lz[6] = nl;
      if (ENT) tp.sz[6] = ns;
    } else if (MAG) {
      // ------------------------------------------------------------------ magnetic role
      const Pars &P = cP;
      double dfv[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) dfv[k] = rdf ? __ldcs(a.df + o + (S::IAX + k) * cP.sv) : 0.0;
      mbar_wait(&bar_ring[(j0 + 6) % V3_NSLOT], ((j0 + 6) / V3_NSLOT) & 1);
      mbar_wait(&bar_scal[j0 & 1], (j0 >> 1) & 1);
      double rhs[8], fc[8], y[4] = {0., 0., 0., 0.};
      if (active) {
        const double u[3] = {tp.pl[3][0], tp.pl[3][V3_PLANE], tp.pl[3][2 * V3_PLANE]};
        MagOut mo;
        mag_part<ENT, UPW>(tp, u, rhs, fc, mo);
        const double rho1 = exp(-tp.scal[0]); // density.f90:2750
        if (P.llorentzforce) {
#pragma unroll
          for (int i = 0; i < 3; ++i) y[i] = rho1 * mo.jxb[i];
        }
        if (ENT && P.lohmic_heat) y[3] = P.eta * P.mu0 * mo.j2 * rho1;
        if (P.lvisc_nu_const) { // mixed part of grad div u (sub.f90:2610-2749), weighted nu/3
          const double u1_xy = tp.dij(S::IUX + 0, 0), u1_zx = tp.dij(S::IUX + 0, 2);
          const double u2_xy = tp.dij(S::IUX + 1, 0), u2_yz = tp.dij(S::IUX + 1, 1);
          const double u3_yz = tp.dij(S::IUX + 2, 1), u3_zx = tp.dij(S::IUX + 2, 2);
          const double w = P.nu * (1. / 3.);
          y[0] += w * (u2_xy + u3_zx);
          y[1] += w * (u1_xy + u3_yz);
          y[2] += w * (u1_zx + u2_yz);
        }
      }
      if (j0 > 0) named_bar_sync(BAR_MBOX_FREE, L::THREADS); // hydro warps have read the previous plane's mail
      if (active) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
          if (i < 3 || ENT) xch[i * V3_CELLS + cell] = y[i];
      }
      named_bar_arrive(BAR_MBOX_FULL, L::THREADS);
      if (active) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          const long long q = o + (S::IAX + k) * cP.sv;
          const double d = a.alpha * dfv[k] + rhs[S::IAX + k];
          if (a.write_df) __stcs(a.df + q, d);
          a.fout[q] = fc[S::IAX + k] + dtb * d;
        }
      }
    }
    // ---- this thread is done with plane z-3's ring slot and with this scalar slot
    if (z + 1 < zc1) {
      if (!producer_warp) named_bar_arrive(BAR_PLANE_EVEN + (j0 & 1), L::THREADS);
    }
  }
}

} // namespace pcb200
