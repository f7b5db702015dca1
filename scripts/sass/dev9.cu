// Static check of variant 9 (see dev8.cu):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -I../../pencil_code_b200/csrc -I../../include \
//       -Xptxas -v -cubin -o /tmp/dev9.cubin dev9.cu && cuobjdump -sass /tmp/dev9.cubin > /tmp/dev9.sass
#include "pcb200_device.cuh"
#include "pcb200_fast.cuh"
#include "pcb200_v9.cuh"
namespace pcb200 {
template __global__ void rhs_update_v9<true, true, false, false>(const __grid_constant__ CUtensorMap, const __grid_constant__ V9Args);
} // namespace pcb200
