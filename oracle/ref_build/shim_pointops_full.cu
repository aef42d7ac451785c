// extern "C" doorway onto openpoints/cpp/pointops/src launchers (the full Point-Transformer op set,
// offset format).  Test infrastructure only.
#include "knnquery/knnquery_cuda_kernel.h"
#include "sampling/sampling_cuda_kernel.h"
#include "ballquery/ballquery_cuda_kernel.h"
#include "grouping/grouping_cuda_kernel.h"
#include "interpolation/interpolation_cuda_kernel.h"
#include "subtraction/subtraction_cuda_kernel.h"
#include "aggregation/aggregation_cuda_kernel.h"
#include <cuda_runtime.h>
extern "C" {
int ref_sync() { return (int)cudaDeviceSynchronize(); }
void ref_knnquery(int m, int nsample, const float* xyz, const float* new_xyz, const int* offset, const int* new_offset, int* idx, float* dist2) {
  knnquery_cuda_launcher(m, nsample, xyz, new_xyz, offset, new_offset, idx, dist2);
}
void ref_furthestsampling(int b, int n, const float* xyz, const int* offset, const int* new_offset, float* tmp, int* idx) {
  furthestsampling_cuda_launcher(b, n, xyz, offset, new_offset, tmp, idx);
}
void ref_ballquery(int m, float radius, int nsample, const float* xyz, const float* new_xyz, const int* offset, const int* new_offset, int* idx) {
  ballquery_launcher(m, radius, nsample, xyz, new_xyz, offset, new_offset, idx);
}
void ref_grouping_forward(int m, int nsample, int c, const float* input, const int* idx, float* output) {
  grouping_forward_cuda_launcher(m, nsample, c, input, idx, output);
}
void ref_grouping_backward(int m, int nsample, int c, const float* grad_output, const int* idx, float* grad_input) {
  grouping_backward_cuda_launcher(m, nsample, c, grad_output, idx, grad_input);
}
void ref_interpolation_forward(int n, int c, int k, const float* input, const int* idx, const float* weight, float* output) {
  interpolation_forward_cuda_launcher(n, c, k, input, idx, weight, output);
}
void ref_interpolation_backward(int n, int c, int k, const float* grad_output, const int* idx, const float* weight, float* grad_input) {
  interpolation_backward_cuda_launcher(n, c, k, grad_output, idx, weight, grad_input);
}
void ref_subtraction_forward(int n, int nsample, int c, const float* input1, const float* input2, const int* idx, float* output) {
  subtraction_forward_cuda_launcher(n, nsample, c, input1, input2, idx, output);
}
void ref_subtraction_backward(int n, int nsample, int c, const int* idx, const float* grad_output, float* grad_input1, float* grad_input2) {
  subtraction_backward_cuda_launcher(n, nsample, c, idx, grad_output, grad_input1, grad_input2);
}
void ref_aggregation_forward(int n, int nsample, int c, int w_c, const float* input, const float* position, const float* weight, const int* idx, float* output) {
  aggregation_forward_cuda_launcher(n, nsample, c, w_c, input, position, weight, idx, output);
}
void ref_aggregation_backward(int n, int nsample, int c, int w_c, const float* input, const float* position, const float* weight, const int* idx, const float* grad_output, float* grad_input, float* grad_position, float* grad_weight) {
  aggregation_backward_cuda_launcher(n, nsample, c, w_c, input, position, weight, idx, grad_output, grad_input, grad_position, grad_weight);
}
}
