// extern "C" doorway onto pointnet2/_ext_src/src/*.cu launchers (variant #4: 512-thread FPS with
// the |p|^2 <= 1e-3 skip).  Test infrastructure only.
#include "cuda_utils.h"
void query_ball_point_kernel_wrapper(int b, int n, int m, float radius, int nsample, const float* new_xyz, const float* xyz, int* idx);
void group_points_kernel_wrapper(int b, int c, int n, int npoints, int nsample, const float* points, const int* idx, float* out);
void group_points_grad_kernel_wrapper(int b, int c, int n, int npoints, int nsample, const float* grad_out, const int* idx, float* grad_points);
void three_nn_kernel_wrapper(int b, int n, int m, const float* unknown, const float* known, float* dist2, int* idx);
void three_interpolate_kernel_wrapper(int b, int c, int m, int n, const float* points, const int* idx, const float* weight, float* out);
void three_interpolate_grad_kernel_wrapper(int b, int c, int n, int m, const float* grad_out, const int* idx, const float* weight, float* grad_points);
void gather_points_kernel_wrapper(int b, int c, int n, int npoints, const float* points, const int* idx, float* out);
void gather_points_grad_kernel_wrapper(int b, int c, int n, int npoints, const float* grad_out, const int* idx, float* grad_points);
void furthest_point_sampling_kernel_wrapper(int b, int n, int m, const float* dataset, float* temp, int* idxs);

extern "C" {
int ref_sync() { return (int)cudaDeviceSynchronize(); }
void ref_ball_query(int b, int n, int m, float radius, int nsample, const float* new_xyz, const float* xyz, int* idx) {
  query_ball_point_kernel_wrapper(b, n, m, radius, nsample, new_xyz, xyz, idx);
}
void ref_group_points(int b, int c, int n, int npoints, int nsample, const float* points, const int* idx, float* out) {
  group_points_kernel_wrapper(b, c, n, npoints, nsample, points, idx, out);
}
void ref_group_points_grad(int b, int c, int n, int npoints, int nsample, const float* grad_out, const int* idx, float* grad_points) {
  group_points_grad_kernel_wrapper(b, c, n, npoints, nsample, grad_out, idx, grad_points);
}
void ref_three_nn(int b, int n, int m, const float* unknown, const float* known, float* dist2, int* idx) {
  three_nn_kernel_wrapper(b, n, m, unknown, known, dist2, idx);
}
void ref_three_interpolate(int b, int c, int m, int n, const float* points, const int* idx, const float* weight, float* out) {
  three_interpolate_kernel_wrapper(b, c, m, n, points, idx, weight, out);
}
void ref_three_interpolate_grad(int b, int c, int n, int m, const float* grad_out, const int* idx, const float* weight, float* grad_points) {
  three_interpolate_grad_kernel_wrapper(b, c, n, m, grad_out, idx, weight, grad_points);
}
void ref_gather_points(int b, int c, int n, int npoints, const float* points, const int* idx, float* out) {
  gather_points_kernel_wrapper(b, c, n, npoints, points, idx, out);
}
void ref_gather_points_grad(int b, int c, int n, int npoints, const float* grad_out, const int* idx, float* grad_points) {
  gather_points_grad_kernel_wrapper(b, c, n, npoints, grad_out, idx, grad_points);
}
void ref_furthest_point_sampling(int b, int n, int m, const float* dataset, float* temp, int* idxs) {
  furthest_point_sampling_kernel_wrapper(b, n, m, dataset, temp, idxs);
}
}
