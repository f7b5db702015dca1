// Does a TMA tensor copy of FP64 boxes need a 16-byte aligned box start (even x coordinate)?  Loads and stores of a
// (32 x 8) box at x0 = 2 and x0 = 3.   nvcc -gencode arch=compute_100a,code=sm_100a -o tma_align tma_align.cu -lcuda
#include <cstdio>
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void k(const __grid_constant__ CUtensorMap in, const __grid_constant__ CUtensorMap out, int x0, int do_store, double *chk) {
  __shared__ __align__(128) double buf[256];
  __shared__ uint64_t bar;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
    asm volatile("fence.proxy.async.shared::cta;");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], 2048;" ::"r"(s32(&bar)));
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(s32(buf)), "l"(&in), "r"(s32(&bar)), "r"(x0), "r"(1) : "memory");
  }
  asm volatile("{\n.reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n@p bra D;\nbra W;\nD:\n}" ::"r"(s32(&bar)) : "memory");
  chk[threadIdx.x] = buf[threadIdx.x];
  buf[threadIdx.x] += 1000.0;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0 && do_store) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(&out), "r"(s32(buf)), "r"(x0), "r"(1) : "memory");
    asm volatile("cp.async.bulk.commit_group;");
    asm volatile("cp.async.bulk.wait_group 0;");
  }
}
int main() {
  const int mx = 70, my = 20;
  double *a, *b, *chk, h[70 * 20], hc[256];
  cudaMalloc(&a, sizeof(h)); cudaMalloc(&b, sizeof(h)); cudaMalloc(&chk, sizeof(hc));
  for (int i = 0; i < mx * my; ++i) h[i] = i;
  cudaMemcpy(a, h, sizeof(h), cudaMemcpyHostToDevice);
  CUtensorMap tin, tout;
  cuuint64_t dims[2] = {mx, my}, str[1] = {mx * 8};
  cuuint32_t box[2] = {32, 8}, es[2] = {1, 1};
  cuInit(0);
  CUresult r1 = cuTensorMapEncodeTiled(&tin, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, a, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  CUresult r2 = cuTensorMapEncodeTiled(&tout, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, b, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("encode %d %d\n", (int)r1, (int)r2);
  for (int st = 0; st < 2; ++st)
    for (int x0 : {2, 3}) {
      cudaMemset(b, 0, sizeof(h));
      k<<<1, 256>>>(tin, tout, x0, st, chk);
      cudaError_t e = cudaDeviceSynchronize();
      printf("store=%d x0=%d: %s", st, x0, cudaGetErrorString(e));
      if (e == cudaSuccess) {
        cudaMemcpy(hc, chk, sizeof(hc), cudaMemcpyDeviceToHost);
        printf("  loaded[0]=%g (expect %d) loaded[33]=%g (expect %d)", hc[0], mx + x0, hc[33], 2 * mx + x0 + 1);
        if (st) { cudaMemcpy(h, b, sizeof(h), cudaMemcpyDeviceToHost); printf("  stored(%d,1)=%g stored(%d,1)=%g", x0, h[mx + x0], x0 - 1, h[mx + x0 - 1]); }
      }
      printf("\n");
      if (e != cudaSuccess) return 0;
    }
  return 0;
}
