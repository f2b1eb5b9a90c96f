// fix_regs.cu — instantiates only the head fix-up kernels, to read their register use / SASS without the whole library:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xptxas -v -Idelayed_speech_synthesis_b200/csrc -cubin -o /tmp/fix_regs.cubin tools/microbench/fix_regs.cu
#include <cuda.h>
#include <cuda_runtime.h>

#include "hga_kernels_lane.cuh"
namespace hga {
template __global__ void k_lane_fix<float, 10>(const __grid_constant__ LaneConst, const LaneFix*, const LaneSub*, int, const double*, int, int, const void*, const double*, double*,
                                               const double*, const double*, double*, int64_t, double*);
template __global__ void k_lane_fix<double, 10>(const __grid_constant__ LaneConst, const LaneFix*, const LaneSub*, int, const double*, int, int, const void*, const double*, double*,
                                                const double*, const double*, double*, int64_t, double*);
}  // namespace hga
