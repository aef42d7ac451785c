// Build-time stand-in for <torch/serialize/tensor.h>.
// The reference's *_gpu.h / *_kernel.h headers only *declare* host wrappers that take at::Tensor
// by value; the .cu files we compile never touch a Tensor.  A forward declaration is enough, which
// keeps the reference-kernel oracle (oracle/_ref) free of any libtorch dependency and makes it build
// in seconds.  Test infrastructure only.
#pragma once
namespace at { class Tensor; }
namespace torch { using Tensor = at::Tensor; }
