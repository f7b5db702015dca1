// Microbenchmarks that calibrate the per-SM rates the fused kernel's budget is built from:
// FP64 pipe, shared-memory loads (64/128 bit), tensor-memory loads/stores used as per-thread scratch.
// One CTA per SM, NW warps; prints cycles per warp-instruction per SM sub-partition and bytes/clk/SM.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench ubench.cu && ./ubench
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__device__ __forceinline__ long long clk() { long long c; asm volatile("mov.u64 %0, %%clock64;" : "=l"(c)); return c; }

template <int ILP> __global__ void k_dfma(double *out, long long *cyc, int iters, double a, double b) {
  double v[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) v[i] = threadIdx.x + i;
  __syncthreads();
  long long t0 = clk();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
      for (int i = 0; i < ILP; ++i) v[i] = fma(v[i], a, b);
  }
  __syncthreads();
  long long t1 = clk();
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += v[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

// DADD/DFMA mix interleaved with independent integer work (issue-slot sharing)
template <int ILP, int NI> __global__ void k_mix(double *out, long long *cyc, int iters, double a, double b, int q) {
  double v[ILP];
  int w[4] = {1, 2, 3, 4};
#pragma unroll
  for (int i = 0; i < ILP; ++i) v[i] = threadIdx.x + i;
  __syncthreads();
  long long t0 = clk();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 8; ++r) {
#pragma unroll
      for (int i = 0; i < ILP; ++i) {
        v[i] = fma(v[i], a, b);
#pragma unroll
        for (int n = 0; n < NI; ++n) w[(i + n) & 3] = (w[(i + n) & 3] ^ q) + n;
      }
    }
  }
  __syncthreads();
  long long t1 = clk();
  double s = w[0] + w[1] + w[2] + w[3];
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += v[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

extern __shared__ __align__(16) double sm[];
template <int W> __global__ void k_lds(double *out, long long *cyc, int iters, int stride) {
  for (int i = threadIdx.x; i < 16384; i += blockDim.x) sm[i] = i;
  __syncthreads();
  const uint32_t base = (uint32_t)__cvta_generic_to_shared(sm) + threadIdx.x * (W * 8);
  double s = 0;
  long long t0 = clk();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      const uint32_t ad = base + ((it + r * stride) & 15) * 4096;
      if (W == 1) { double x; asm volatile("ld.shared.f64 %0, [%1];" : "=d"(x) : "r"(ad)); s += x; }
      else { double x, y; asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(x), "=d"(y) : "r"(ad)); s += x + y; }
    }
  }
  __syncthreads();
  long long t1 = clk();
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

// tensor memory as per-thread scratch: X 32-bit columns per instruction
template <int X, bool ST> __global__ void k_tmem(double *out, long long *cyc, int iters) {
  __shared__ uint32_t slot;
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(&slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const int tid = threadIdx.x;
  const uint32_t tb = slot + ((uint32_t)((tid >> 5) & 3) << 21) + (uint32_t)(tid >> 7) * 256u;
  uint32_t r[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) r[i] = tid + i;
  // initialise the columns
  for (int c = 0; c < 256; c += 16)
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(tb + c),
                 "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]));
  asm volatile("tcgen05.wait::st.sync.aligned;");
  __syncthreads();
  uint32_t acc = 0;
  long long t0 = clk();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const uint32_t ta = tb + ((it + q) & 7) * 16;
      if (ST) {
        if (X == 16)
          asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(ta),
                       "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]));
        else if (X == 4)
          asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1,%2,%3,%4};" ::"r"(ta), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]));
        else
          asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1,%2};" ::"r"(ta), "r"(r[0]), "r"(r[1]));
      } else {
        if (X == 16) {
          asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                       : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                       : "r"(ta));
        } else if (X == 8) {
          asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                       : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(ta));
        } else {
          asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0,%1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(ta));
        }
      }
    }
    if (!ST) { asm volatile("tcgen05.wait::ld.sync.aligned;"); acc += r[0] + r[X - 1]; }
  }
  if (ST) asm volatile("tcgen05.wait::st.sync.aligned;");
  __syncthreads();
  long long t1 = clk();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(slot) : "memory");
}

static double *out; static long long *cyc; static long long hc[148];
static double med() { cudaMemcpy(hc, cyc, sizeof(hc), cudaMemcpyDeviceToHost); long long s = 0; for (int i = 0; i < 148; ++i) s += hc[i]; return (double)s / 148; }

int main() {
  CK(cudaMalloc(&out, 148 * 1024 * 8)); CK(cudaMalloc(&cyc, 148 * 8));
  const int it = 2000;
#define RUN(name, kern, nw, smem, instr_per_warp, bytes_per_warp, ...) do { \
    kern<<<148, (nw) * 32, smem>>>(out, cyc, __VA_ARGS__); CK(cudaDeviceSynchronize()); \
    kern<<<148, (nw) * 32, smem>>>(out, cyc, __VA_ARGS__); CK(cudaDeviceSynchronize()); \
    double c = med(); double wi = (double)(instr_per_warp) * (nw) / 4.0; \
    printf("%-34s warps %2d  cycles %9.0f  clk per warp-instr per SMSP %6.3f  bytes/clk/SM %7.1f\n", name, nw, c, c / wi, (double)(bytes_per_warp) * (nw) / c); } while (0)
  for (int nw : {4, 8, 12, 16}) {
    RUN("DFMA ilp1", k_dfma<1>, nw, 0, (double)it * 8 * 1, 0, it, 1.0000001, 1e-9);
    RUN("DFMA ilp2", k_dfma<2>, nw, 0, (double)it * 8 * 2, 0, it, 1.0000001, 1e-9);
    RUN("DFMA ilp4", k_dfma<4>, nw, 0, (double)it * 8 * 4, 0, it, 1.0000001, 1e-9);
    RUN("DFMA ilp8", k_dfma<8>, nw, 0, (double)it * 8 * 8, 0, it, 1.0000001, 1e-9);
  }
  for (int nw : {4, 8}) {
    RUN("DFMA ilp4 + 1 int each", (k_mix<4, 1>), nw, 0, (double)it * 8 * 4, 0, it, 1.0000001, 1e-9, 3);
    RUN("DFMA ilp4 + 2 int each", (k_mix<4, 2>), nw, 0, (double)it * 8 * 4, 0, it, 1.0000001, 1e-9, 3);
    RUN("DFMA ilp4 + 3 int each", (k_mix<4, 3>), nw, 0, (double)it * 8 * 4, 0, it, 1.0000001, 1e-9, 3);
  }
  CK(cudaFuncSetAttribute(k_lds<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072));
  CK(cudaFuncSetAttribute(k_lds<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072));
  for (int nw : {4, 8, 16}) {
    RUN("LDS.64", k_lds<1>, nw, 131072, (double)it * 16, (double)it * 16 * 256, it, 1);
    RUN("LDS.128", k_lds<2>, nw, 131072, (double)it * 16, (double)it * 16 * 512, it, 1);
  }
  for (int nw : {4, 8}) {
    RUN("LDTM.x2", (k_tmem<2, false>), nw, 0, (double)it * 8, (double)it * 8 * 32 * 8, it);
    RUN("LDTM.x8", (k_tmem<8, false>), nw, 0, (double)it * 8, (double)it * 8 * 32 * 32, it);
    RUN("LDTM.x16", (k_tmem<16, false>), nw, 0, (double)it * 8, (double)it * 8 * 32 * 64, it);
    RUN("STTM.x2", (k_tmem<2, true>), nw, 0, (double)it * 8, (double)it * 8 * 32 * 8, it);
    RUN("STTM.x4", (k_tmem<4, true>), nw, 0, (double)it * 8, (double)it * 8 * 32 * 16, it);
    RUN("STTM.x16", (k_tmem<16, true>), nw, 0, (double)it * 8, (double)it * 8 * 32 * 64, it);
  }
  return 0;
}
