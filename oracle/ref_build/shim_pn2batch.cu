// extern "C" doorway onto the reference's own launchers in openpoints/cpp/pointnet2_batch/src/*.cu
// (compiled unchanged from /root/reference by oracle/ref_build/Makefile into oracle/_ref/).
// Test infrastructure: the GPU parity oracle and "kernel to beat"; never linked by the product.
#include "ball_query_gpu.h"
#include "group_points_gpu.h"
#include "interpolate_gpu.h"
#include "sampling_gpu.h"
#include <cuda_runtime.h>

extern "C" {
int ref_sync() { return (int)cudaDeviceSynchronize(); }
// ball_query_gpu.cu:54 defines (new_xyz, xyz) — the header's names are swapped, types agree.
void ref_ball_query(int b, int n, int m, float radius, int nsample, const float* new_xyz, const float* xyz, int* idx) {
  ball_query_kernel_launcher_fast(b, n, m, radius, nsample, new_xyz, xyz, idx);
}
void ref_group_points(int b, int c, int n, int npoints, int nsample, const float* points, const int* idx, float* out) {
  group_points_kernel_launcher_fast(b, c, n, npoints, nsample, points, idx, out);
}
void ref_group_points_grad(int b, int c, int n, int npoints, int nsample, const float* grad_out, const int* idx, float* grad_points) {
  group_points_grad_kernel_launcher_fast(b, c, n, npoints, nsample, grad_out, idx, grad_points);
}
void ref_three_nn(int b, int n, int m, const float* unknown, const float* known, float* dist2, int* idx) {
  three_nn_kernel_launcher_fast(b, n, m, unknown, known, dist2, idx);
}
void ref_three_interpolate(int b, int c, int m, int n, const float* points, const int* idx, const float* weight, float* out) {
  three_interpolate_kernel_launcher_fast(b, c, m, n, points, idx, weight, out);
}
void ref_three_interpolate_grad(int b, int c, int n, int m, const float* grad_out, const int* idx, const float* weight, float* grad_points) {
  three_interpolate_grad_kernel_launcher_fast(b, c, n, m, grad_out, idx, weight, grad_points);
}
void ref_gather_points(int b, int c, int n, int npoints, const float* points, const int* idx, float* out) {
  gather_points_kernel_launcher_fast(b, c, n, npoints, points, idx, out);
}
void ref_gather_points_grad(int b, int c, int n, int npoints, const float* grad_out, const int* idx, float* grad_points) {
  gather_points_grad_kernel_launcher_fast(b, c, n, npoints, grad_out, idx, grad_points);
}
void ref_furthest_point_sampling(int b, int n, int m, const float* dataset, float* temp, int* idxs) {
  furthest_point_sampling_kernel_launcher(b, n, m, dataset, temp, idxs);
}
}
