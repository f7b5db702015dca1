// Static check of variant 10 (see dev9.cu)
#include "pcb200_device.cuh"
#include "pcb200_fast.cuh"
#include "pcb200_v10.cuh"
namespace pcb200 {
template __global__ void rhs_update_v10<true, true, false, false>(const __grid_constant__ V10Maps, const __grid_constant__ V10Args);
} // namespace pcb200
