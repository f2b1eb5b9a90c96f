// hga_kernels_pipe.cuh — warp-pipelined variant of K-hga (f64); see DESIGN.md §4.
#pragma once
#include "hga_kernels_fused.cuh"

namespace hga {
static inline bool pipe_supported(const Cascade&, int, int) { return false; }
static inline int pipe_preferred(int, int, int, const int64_t*, int) { return 0; }
static inline int64_t pipe_item_slots(int n_sm, int) { return n_sm; }
static inline int launch_pipe(const Cascade&, const FusedArgs&, int, cudaStream_t) { return HGA_ERR_UNSUPPORTED; }
}  // namespace hga
