// Static check of one kernel instantiation without rebuilding the library (seconds instead of minutes):
//   cd scripts/sass && nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 \
//       -I../../pencil_code_b200/csrc -I../../include -Xptxas -v -cubin -o /tmp/dev8.cubin dev8.cu
//   cuobjdump -sass /tmp/dev8.cubin > /tmp/dev8.sass && python loop_stat.py /tmp/dev8.sass rhs_update_v8
// prints registers/spills (ptxas) and the instruction histogram of the z loop (FP64, LDS.64, the rest).
#include "pcb200_device.cuh"
#include "pcb200_fast.cuh"
#include "pcb200_v8.cuh"
namespace pcb200 {
template __global__ void rhs_update_v8<true, true, false, false>(const __grid_constant__ CUtensorMap, const __grid_constant__ CUtensorMap, const V2Args);
template __global__ void rhs_update_v5<true, true, false, false, 0>(const __grid_constant__ CUtensorMap, const __grid_constant__ CUtensorMap, const V2Args);
} // namespace pcb200
