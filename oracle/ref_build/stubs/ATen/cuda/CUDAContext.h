// Build-time stand-in for <ATen/cuda/CUDAContext.h>: pointnet2/_ext_src launches on
// at::cuda::getCurrentCUDAStream(); the oracle build pins that to a settable stream (default 0).
#pragma once
#include <cuda_runtime_api.h>
#include <cstdio>
#include <cstdlib>
namespace at { class Tensor; namespace cuda {
inline cudaStream_t& geot_ref_stream_slot() { static cudaStream_t s = 0; return s; }
inline cudaStream_t getCurrentCUDAStream() { return geot_ref_stream_slot(); }
} }
