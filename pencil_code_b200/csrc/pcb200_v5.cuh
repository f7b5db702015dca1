// pcb200_v5.cuh -- kernel variant 5: variant 4 (z-marching TMA ring, lead/trail z-mixed derivatives)
// with the non-arithmetic instruction overhead removed.  Per cell and substep variant 4 executes
// 1468 SASS instructions of which 693 are FP64 and 234 LDS.64; of the other 541, ~180 were 64-bit
// address arithmetic for the 8+8+8+8 global accesses of df / f_out, 63 constant loads and ~50 register
// moves.  Variant 5
//   * walks the global arrays with per-thread running pointers (one 64-bit add per variable instead
//     of an index multiply + scale per access),
//   * issues the L2 prefetch of the df tile two planes ahead once per warp (lane -> variable, cache
//     line) instead of once per cell and variable,
//   * is instantiated for isotropic grids only (dx == dy == dz bit for bit, template flag ISO of the
//     tap policy): one set of stencil weights serves all directions, 10 instead of 30 FP64 constants
//     live in registers.  Anisotropic grids keep variant 4.
// Arithmetic and reference citations: pcb200_device.cuh / pcb200_fast.cuh.
#pragma once
#include "pcb200_fast.cuh"

namespace pcb200 {

template <bool ENT, bool MAG, bool HYP, bool UPW>
__global__ void __launch_bounds__(V2_THREADS, 1)
    rhs_update_v5(const __grid_constant__ CUtensorMap tmap, const V2Args a) {
  typedef V2Layout<ENT, MAG> L;
  typedef Slots<ENT> S;
  constexpr int NV = 4 + (ENT ? 1 : 0) + (MAG ? 3 : 0);
  double *ring = v2_smem;
  double *scal = ring + L::RING_DOUBLES;
  uint64_t *bar_ring = reinterpret_cast<uint64_t *>(scal + L::SCAL_DOUBLES);
  uint64_t *bar_scal = bar_ring + V2_NSLOT;

  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * V2_TX + tx;
  const int bx0 = blockIdx.x * V2_TX, by0 = blockIdx.y * V2_TY;
  const int zc0 = a.z_lo + blockIdx.z * a.lz;
  const int zc1 = min(zc0 + a.lz, a.z_hi);
  if (zc0 >= zc1) return;
  const int ix = bx0 + tx + NG, iy = by0 + ty + NG;
  const bool active = ix < cP.nx + NG && iy < cP.ny + NG;
  const int ixc = min(ix, cP.nx + NG - 1), iyc = min(iy, cP.ny + NG - 1);
  const int last_plane = zc1 - 1 + NG;

  if (tid == 0) {
    if (smem_u32(ring) & 127u) __trap();
    for (int s = 0; s < V2_NSLOT; ++s) mbar_init(&bar_ring[s], 1);
    mbar_init(&bar_scal[0], 1);
    mbar_init(&bar_scal[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();

  auto issue_ring = [&](int j) { // ring plane j <-> z = zc0-3+j
    const int p = zc0 - NG + j;
    if (p > last_plane) return;
    const int slot = j % V2_NSLOT;
    mbar_expect_tx(&bar_ring[slot], L::NRING * V2_BOX_BYTES);
#pragma unroll
    for (int f = 0; f < L::NRING; ++f) {
      const int var = f < 3 ? f : S::IAX + (f - 3);
      tma_load_4d(ring + (slot * L::NRING + f) * V2_PLANE_PAD, &tmap, &bar_ring[slot], bx0, by0, p, var);
    }
  };
  auto issue_scal = [&](int j) { // scalar plane j <-> z = zc0+j
    const int p = zc0 + j;
    if (p >= zc1) return;
    const int slot = j & 1;
    mbar_expect_tx(&bar_scal[slot], L::NSCAL * V2_BOX_BYTES);
    tma_load_4d(scal + (slot * L::NSCAL + 0) * V2_PLANE_PAD, &tmap, &bar_scal[slot], bx0, by0, p, S::ILN);
    if (ENT) tma_load_4d(scal + (slot * L::NSCAL + 1) * V2_PLANE_PAD, &tmap, &bar_scal[slot], bx0, by0, p, S::ISS);
  };

  if (tid == 0) {
#pragma unroll 1
    for (int j = 0; j < 7; ++j) issue_ring(j);
    issue_scal(0);
  }

  SmemLines<ENT, MAG, true, true> tp;
  tp.trail_zero();
  const int toff = (ty + NG) * V2_HX + (tx + NG);
  const long long sv = cP.sv, sz = cP.sz;
  const long long col = ixc + cP.sy * iyc;
  {
    const double *gl = a.fin + col + S::ILN * sv + sz * (zc0 - NG);
#pragma unroll
    for (int d = 0; d < 7; ++d) tp.lz[d] = __ldg(gl + sz * d);
    if (ENT) {
#pragma unroll
      for (int d = 0; d < 7; ++d) tp.sz[d] = __ldg(gl + sv + sz * d);
    }
  }
  const double dtb = (*a.dt_dev) * a.beta;
  const bool rdf = a.read_df != 0 && active;

  // running pointers of this thread (all advance by one plane per iteration)
  const long long o0 = ix + cP.sy * iy + sz * zc0;
  double *pdf = a.df + o0;                                         // df(ix,iy,z,0)
  double *pfo = a.fout + o0;                                       // f_out(ix,iy,z,0)
  const double *pnx = a.fin + col + S::ILN * sv + sz * (zc0 + 4);  // lnrho(ixc,iyc,z+4): next pipeline entry
  // warp-wide L2 prefetch of the df tile row two planes ahead: lane -> (variable, 128-byte line of the row)
  const int pf_k = tx / 3, pf_l = tx - 3 * pf_k;
  const bool pf_on = a.read_df != 0 && a.df_prefetch != 0 && pf_k < NV && iy < cP.ny + NG;
  const double *ppf = a.df + (bx0 + NG + min(pf_l * 16, V2_TX - 1)) + cP.sy * iyc + sz * (zc0 + 2) + pf_k * sv;

#pragma unroll 1
  for (int z = zc0; z < zc1; ++z) {
    const int j0 = z - zc0;
    if (tid < 32) {
      if (j0 > 0) named_bar_sync(1 + ((j0 - 1) & 1), V2_THREADS);
      if (tid == 0) {
        issue_ring(j0 + 7);
        issue_scal(j0 + 1);
      }
      __syncwarp();
    }
    double dfv[NV];
    {
      const double *q = pdf;
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        dfv[k] = rdf ? __ldcs(q) : 0.0;
        q += sv;
      }
    }
    if (pf_on && z + 2 < zc1) asm volatile("prefetch.global.L2 [%0];" ::"l"(ppf));
    double nl = 0.0, ns = 0.0;
    if (z + 4 < cP.mz) {
      nl = __ldg(pnx);
      if (ENT) ns = __ldg(pnx + sv);
    }
    if (j0 == 0) {
#pragma unroll 1
      for (int d = 0; d < 6; ++d) mbar_wait(&bar_ring[d], 0);
    }
    mbar_wait(&bar_ring[(j0 + 6) % V2_NSLOT], ((j0 + 6) / V2_NSLOT) & 1);
    mbar_wait(&bar_scal[j0 & 1], (j0 >> 1) & 1);
#pragma unroll
    for (int d = 0; d < 7; ++d) tp.pl[d] = ring + ((j0 + d) % V2_NSLOT) * (L::NRING * V2_PLANE_PAD) + toff;
    tp.scal = scal + (j0 & 1) * (L::NSCAL * V2_PLANE_PAD) + toff;
    if (j0 == 0) {
      tp.trail_prologue(0);
      if (MAG) tp.trail_prologue(1);
    }

    if (active) {
      double rhs[NV], fc[NV];
      rhs_cell_lines<ENT, MAG, HYP, UPW>(tp, rhs, fc);
      double force[3] = {0., 0., 0.};
      if (a.frc.on) {
        const double2 fx = a.frc.fx[ix], fy = a.frc.fy[iy], fz = a.frc.fz[z];
        const double xr = fx.x * fy.x - fx.y * fy.y, xi = fx.x * fy.y + fx.y * fy.x;
        const double re = xr * fz.x - xi * fz.y, im = xr * fz.y + xi * fz.x;
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
#pragma unroll
    for (int d = 0; d < 6; ++d) {
      tp.lz[d] = tp.lz[d + 1];
      if (ENT) tp.sz[d] = tp.sz[d + 1];
    }
    tp.lz[6] = nl;
    if (ENT) tp.sz[6] = ns;
    pdf += sz;
    pfo += sz;
    pnx += sz;
    ppf += sz;
    if (z + 1 < zc1 && tid >= 32) named_bar_arrive(1 + (j0 & 1), V2_THREADS);
  }
}

} // namespace pcb200
