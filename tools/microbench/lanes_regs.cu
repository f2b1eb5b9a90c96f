// lanes_regs.cu - instantiates only the float32 section-pipelined kernels, to read their register use / SASS in seconds:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xptxas -v -Idelayed_speech_synthesis_b200/csrc -cubin -o /tmp/lanes_regs.cubin tools/microbench/lanes_regs.cu
#include <cuda.h>
#include <cuda_runtime.h>

#include "hga_kernels_lane.cuh"
namespace hga {
template __global__ void k_hga_lanes<LaneF32, 10, 60, 2, 2, 2, 3, 5, false, false, false>(const __grid_constant__ LaneConst, const __grid_constant__ LaneArgs,
                                                                                         const __grid_constant__ CUtensorMap);
template __global__ void k_hga_lanes<LaneF32, 10, 60, 2, 2, 2, 3, 5, false, false, true>(const __grid_constant__ LaneConst, const __grid_constant__ LaneArgs,
                                                                                        const __grid_constant__ CUtensorMap);
}  // namespace hga
