// extern "C" doorway onto the top-level pointops/src launchers (the variant GeoT actually imports:
// knnquery with 1000-slot heaps, furthestsampling, weighted furthestsampling).  Test infrastructure only.
#include "knnquery/knnquery_cuda_kernel.h"
#include "sampling/sampling_cuda_kernel.h"
#include <cuda_runtime.h>
extern "C" {
int ref_sync() { return (int)cudaDeviceSynchronize(); }
void ref_knnquery(int m, int nsample, const float* xyz, const float* new_xyz, const int* offset, const int* new_offset, int* idx, float* dist2) {
  knnquery_cuda_launcher(m, nsample, xyz, new_xyz, offset, new_offset, idx, dist2);
}
void ref_furthestsampling(int b, int n, const float* xyz, const int* offset, const int* new_offset, float* tmp, int* idx) {
  furthestsampling_cuda_launcher(b, n, xyz, offset, new_offset, tmp, idx);
}
void ref_furthestsampling_weights(int b, int n, const float* xyz, const int* offset, const int* new_offset, const float* weights, float* tmp, int* idx) {
  furthestsampling_weights_cuda_launcher(b, n, xyz, offset, new_offset, weights, tmp, idx);
}
}
