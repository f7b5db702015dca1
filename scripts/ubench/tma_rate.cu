// Per-SM throughput of the TMA engine for FP64 boxes of the shapes the fused kernel moves (loads into a shared-memory
// ring, stores from it), as a function of the box shape: is the engine limited by bytes or by box rows?
// One CTA per SM, one issuing thread, DEPTH boxes in flight; the tensor is small enough to stay in L2 (rates of the
// engine, not of HBM) or large (HBM-backed), selectable.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tma_rate tma_rate.cu -lcuda
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_wait(uint64_t *bar, unsigned parity) {
  asm volatile("{\n.reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}" ::"r"(s32(bar)), "r"(parity) : "memory");
}
constexpr int DEPTH = 4;
extern __shared__ __align__(1024) double smem[];

// mode 0: loads, 1: stores, 2: loads + stores alternating (a load and a store of the same shape per iteration)
__global__ void __launch_bounds__(128, 1) k(const __grid_constant__ CUtensorMap tin, const __grid_constant__ CUtensorMap tout, int box_doubles,
                                            int bx, int by, int ntx, int nty, int nz, int iters, int mode, long long *clk) {
  uint64_t *bar = reinterpret_cast<uint64_t *>(smem + DEPTH * box_doubles);
  if (threadIdx.x == 0) {
    for (int i = 0; i < DEPTH; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar[i])));
    asm volatile("fence.mbarrier_init.release.cluster;");
    asm volatile("fence.proxy.async.shared::cta;");
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  const unsigned bytes = box_doubles * 8u;
  // tiles of this CTA: a column of tiles marched along z (like the fused kernel)
  const int tile = blockIdx.x % (ntx * nty);
  const int x0 = (tile % ntx) * bx, y0 = (tile / ntx) * by;
  long long t0 = clock64();
  unsigned par[DEPTH] = {0, 0, 0, 0};
  for (int it = 0; it < iters + DEPTH; ++it) {
    const int s = it % DEPTH;
    if (it >= DEPTH && mode != 1) { mbar_wait(&bar[s], par[s]); par[s] ^= 1; }
    if (it < iters) {
      const int z = it % nz;
      if (mode != 1) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&bar[s])), "r"(bytes));
        asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(s32(smem + s * box_doubles)),
                     "l"(&tin), "r"(s32(&bar[s])), "r"(x0), "r"(y0), "r"(z), "r"(0)
                     : "memory");
      }
      if (mode != 0) {
        asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(&tout), "r"(s32(smem + s * box_doubles)), "r"(x0),
                     "r"(y0), "r"(z), "r"(0)
                     : "memory");
        asm volatile("cp.async.bulk.commit_group;");
        asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(DEPTH - 1));
      }
    }
  }
  if (mode != 0) asm volatile("cp.async.bulk.wait_group 0;");
  clk[blockIdx.x] = clock64() - t0;
}

int main(int argc, char **argv) {
  const int big = argc > 1 ? atoi(argv[1]) : 0; // 1: HBM-backed tensor
  cuInit(0);
  int nsm = 148;
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, 0);
  const int nv = 8;
  struct Shape { int bx, by, bv; const char *what; };
  const Shape shapes[] = {{40, 14, 8, "halo'd plane box 40x14x8"}, {32, 8, 8, "interior box 32x8x8"}, {32, 8, 3, "32x8x3"}, {32, 8, 1, "32x8x1"},
                          {72, 10, 8, "halo'd 72x10x8 (64x4 tile)"}, {64, 4, 8, "interior 64x4x8"}, {64, 4, 3, "64x4x3"}, {136, 8, 4, "halo'd 136x8x4 (128x2 tile)"},
                          {128, 2, 8, "interior 128x2x8"}, {256, 1, 8, "256x1x8"}};
  long long *clk;
  cudaMalloc(&clk, sizeof(long long) * nsm);
  std::vector<long long> h(nsm);
  printf("%s tensor, %d SMs, %d boxes in flight per SM\n", big ? "HBM-backed" : "L2-resident", nsm, DEPTH);
  printf("%-34s %5s %9s | %28s | %28s | %28s\n", "box", "rows", "bytes", "loads  B/clk/SM  clk/row", "stores B/clk/SM  clk/row", "ld+st  B/clk/SM  clk/row");
  for (const Shape &sh : shapes) {
    // tensor: ntx x nty tiles in the plane, nz planes
    const int ntx = big ? 8 : 4, nty = big ? 19 : 4 * 148 / 16 / 4 + 1; // enough distinct tiles for all CTAs in the big case
    const int mx = ((ntx * sh.bx + 1) / 2) * 2, my = nty * sh.by, nz = big ? 64 : 6;
    const size_t n = (size_t)mx * my * nz * nv;
    double *a, *b;
    if (cudaMalloc(&a, n * 8) != cudaSuccess || cudaMalloc(&b, n * 8) != cudaSuccess) { printf("alloc failed\n"); return 1; }
    cudaMemset(a, 0, n * 8); cudaMemset(b, 0, n * 8);
    CUtensorMap tin, tout;
    cuuint64_t dims[4] = {(cuuint64_t)mx, (cuuint64_t)my, (cuuint64_t)nz, (cuuint64_t)nv};
    cuuint64_t str[3] = {(cuuint64_t)mx * 8, (cuuint64_t)mx * my * 8, (cuuint64_t)mx * my * nz * 8};
    cuuint32_t box[4] = {(cuuint32_t)sh.bx, (cuuint32_t)sh.by, 1, (cuuint32_t)sh.bv}, es[4] = {1, 1, 1, 1};
    CUresult r1 = cuTensorMapEncodeTiled(&tin, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 4, a, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                         CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    CUresult r2 = cuTensorMapEncodeTiled(&tout, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 4, b, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                         CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r1 || r2) { printf("encode failed %d %d for %s\n", (int)r1, (int)r2, sh.what); return 1; }
    const int box_doubles = ((sh.bx * sh.by * sh.bv + 15) / 16) * 16;
    const size_t smem_bytes = (size_t)DEPTH * box_doubles * 8 + 64;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
    const int iters = 2000;
    const int rows = sh.by * sh.bv;
    const double bytes = (double)sh.bx * sh.by * sh.bv * 8;
    printf("%-34s %5d %9.0f |", sh.what, rows, bytes);
    for (int mode = 0; mode < 3; ++mode) {
      k<<<nsm, 128, smem_bytes>>>(tin, tout, box_doubles, sh.bx, sh.by, ntx, nty, nz, iters, mode, clk); // warm-up
      k<<<nsm, 128, smem_bytes>>>(tin, tout, box_doubles, sh.bx, sh.by, ntx, nty, nz, iters, mode, clk);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf(" %s\n", cudaGetErrorString(e)); return 1; }
      cudaMemcpy(h.data(), clk, sizeof(long long) * nsm, cudaMemcpyDeviceToHost);
      long long mx_clk = 0;
      for (long long v : h) mx_clk = v > mx_clk ? v : mx_clk;
      const double per = (double)mx_clk / iters, mult = mode == 2 ? 2.0 : 1.0;
      printf(" %12.1f B/clk %8.2f clk/row |", mult * bytes / per, per / (mult * rows));
    }
    printf("\n");
    cudaFree(a); cudaFree(b);
  }
  return 0;
}
