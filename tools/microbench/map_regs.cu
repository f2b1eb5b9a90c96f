// map_regs.cu — instantiates only the fused channel-map variant of the lane kernel, to read its register use / SASS quickly:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xptxas -v -Idelayed_speech_synthesis_b200/csrc -cubin -o /tmp/map_regs.cubin tools/microbench/map_regs.cu
#include <cuda.h>
#include <cuda_runtime.h>

#include "hga_kernels_lane.cuh"
namespace hga {
template __global__ void k_hga_lanes<LaneF32, 10, 20, 2, 2, 2, 3, 5, false, true>(const __grid_constant__ LaneConst, const __grid_constant__ LaneArgs,
                                                                                  const __grid_constant__ CUtensorMap);
}  // namespace hga
