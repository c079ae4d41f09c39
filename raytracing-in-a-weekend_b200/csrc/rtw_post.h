// rtw_post.h — device driver of the 8-bit post-processing stage (SURVEY 8f-4 tail): the reference's
// bilateral filter, Rust2/src/postprocessing.rs:11-135. Kernels in rtw_post.cu.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace rtw {

enum : int { POST_PROX_SQUARE = 0, POST_PROX_EDGES = 1 };     // ProximityType (postprocessing.rs:11-14)

// d_in / d_out: height rows of width RGB8 pixels on the device. d_params (4 floats on the device, may be
// null) receives {spatial, avg_gradient, inv_range, inv_spatial}. Everything is enqueued on `st`;
// scratch comes from the stream-ordered allocator and is released on the stream.
cudaError_t bilateral_filter_enqueue(const uint8_t* d_in, uint32_t width, uint32_t height, uint32_t size, int kind,
                                     uint8_t* d_out, float* d_params, cudaStream_t st, int* n_launches);

}  // namespace rtw
