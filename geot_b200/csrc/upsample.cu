// geot_b200/csrc/upsample.cu — the step on the inference side of the path, sm_100a.
//
// Replaces examples/segmentation/train.py:781-800 (get_pred_whole: three_nn -> 1/(d+1e-8) -> sum -> divide ->
// three_interpolate -> argmax, per cloud; 5 torch kernels and a (1,C,n_whole) tensor between the two ops) and
// train.py:803-832 (get_seg_metrics: a Python loop over classes on CPU tensors, 4 reductions per class), plus the
// weight expression of upsampling.py:98-100 / pointnet2_modules.py:619-623.
//   interp_weights_kernel : squared distances -> inverse-distance weights (and sqrt distances), one pass
//   interp_argmax_kernel  : weighted 3-neighbour blend of every class row + first-max arg-max; the upsampled
//                           (C, n) score tensor is never written (HBM: 24 n in, 8 n out; the class rows live in L2)
//   seg_counts_kernel     : per-class |pred = c and label = c|, |pred = c|, |label = c| and the number of correct
//                           vertices (integer shared-memory histograms, integer global atomics: order-free, exact)
#include <algorithm>
#include "common.cuh"

namespace geot {

void count_launch(int n);

constexpr int kUpThreads = 256;

// torch: dist = sqrt(d2); r = 1.0 / (dist + 1e-8); norm = r.sum(2, keepdim=True); w = r / norm   (all fp32, IEEE)
__global__ void __launch_bounds__(kUpThreads)
interp_weights_kernel(int64_t total, const float* __restrict__ dist2, float* __restrict__ dist, float* __restrict__ weight) {
  for (int64_t j = (int64_t)blockIdx.x * kUpThreads + threadIdx.x; j < total; j += (int64_t)gridDim.x * kUpThreads) {
    const float d0 = __fsqrt_rn(__ldg(dist2 + 3 * j)), d1 = __fsqrt_rn(__ldg(dist2 + 3 * j + 1)), d2 = __fsqrt_rn(__ldg(dist2 + 3 * j + 2));
    const float r0 = __fdiv_rn(1.0f, __fadd_rn(d0, 1e-8f)), r1 = __fdiv_rn(1.0f, __fadd_rn(d1, 1e-8f)),
                r2 = __fdiv_rn(1.0f, __fadd_rn(d2, 1e-8f));
    const float norm = __fadd_rn(__fadd_rn(r0, r1), r2);
    if (dist != nullptr) { dist[3 * j] = d0; dist[3 * j + 1] = d1; dist[3 * j + 2] = d2; }
    weight[3 * j] = __fdiv_rn(r0, norm);
    weight[3 * j + 1] = __fdiv_rn(r1, norm);
    weight[3 * j + 2] = __fdiv_rn(r2, norm);
  }
}

// thread per target vertex; the blend is the expression of three_interpolate (interpolate_gpu.cu:103 as nvcc contracts it),
// so the scores are the bits three_interpolate would have written; arg-max keeps the FIRST maximum (torch.argmax)
template <int UNROLL>
__global__ void __launch_bounds__(kUpThreads)
interp_argmax_kernel(int C, int m, int n, const float* __restrict__ points, const int* __restrict__ idx,
                     const float* __restrict__ weight, int64_t* __restrict__ pred, float* __restrict__ score) {
  const int b = blockIdx.y;
  const int j = blockIdx.x * kUpThreads + threadIdx.x;
  if (j >= n) return;
  const int64_t t = ((int64_t)b * n + j) * 3;
  const int i0 = __ldg(idx + t), i1 = __ldg(idx + t + 1), i2 = __ldg(idx + t + 2);
  const float w0 = __ldg(weight + t), w1 = __ldg(weight + t + 1), w2 = __ldg(weight + t + 2);
  const float* __restrict__ base = points + (int64_t)b * C * m;
  float best = 0.f;
  int arg = -1;
  int c = 0;
  for (; c + UNROLL <= C; c += UNROLL) {
    float a[UNROLL], bb[UNROLL], cc[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; u++) {  // all 3*UNROLL gathers requested before the first use
      const float* __restrict__ row = base + (int64_t)(c + u) * m;
      a[u] = __ldg(row + i0); bb[u] = __ldg(row + i1); cc[u] = __ldg(row + i2);
    }
#pragma unroll
    for (int u = 0; u < UNROLL; u++) {
      const float v = __fmaf_rn(w2, cc[u], __fmaf_rn(w0, a[u], __fmul_rn(w1, bb[u])));
      if (arg < 0 || v > best || (v != v && best == best)) { best = v; arg = c + u; }  // NaN wins once, like torch
    }
  }
  for (; c < C; c++) {
    const float* __restrict__ row = base + (int64_t)c * m;
    const float v = __fmaf_rn(w2, __ldg(row + i2), __fmaf_rn(w0, __ldg(row + i0), __fmul_rn(w1, __ldg(row + i1))));
    if (arg < 0 || v > best || (v != v && best == best)) { best = v; arg = c; }
  }
  pred[(int64_t)b * n + j] = arg;
  if (score != nullptr) score[(int64_t)b * n + j] = best;
}

// counts layout: inter[C] | pred_cnt[C] | label_cnt[C] | correct | out_of_range  (uint64)
__global__ void __launch_bounds__(kUpThreads)
seg_counts_kernel(int64_t n, int C, const int64_t* __restrict__ pred, const int64_t* __restrict__ label,
                  unsigned long long* __restrict__ counts) {
  extern __shared__ unsigned s_cnt[];  // 3C + 2; a CTA sees < 2^32 vertices
  for (int i = threadIdx.x; i < 3 * C + 2; i += kUpThreads) s_cnt[i] = 0u;
  __syncthreads();
  for (int64_t g = (int64_t)blockIdx.x * kUpThreads + threadIdx.x; g < n; g += (int64_t)gridDim.x * kUpThreads) {
    const int64_t p = pred[g], l = label[g];
    const bool pin = p >= 0 && p < C, lin = l >= 0 && l < C;
    if (pin) atomicAdd(&s_cnt[C + (int)p], 1u);
    if (lin) atomicAdd(&s_cnt[2 * C + (int)l], 1u);
    if (p == l) {
      atomicAdd(&s_cnt[3 * C], 1u);
      if (pin) atomicAdd(&s_cnt[(int)p], 1u);
    }
    if (!pin || !lin) atomicAdd(&s_cnt[3 * C + 1], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 3 * C + 2; i += kUpThreads)
    if (s_cnt[i] != 0u) atomicAdd(&counts[i], (unsigned long long)s_cnt[i]);
}

}  // namespace geot

using namespace geot;

GEOT_API int geot_interp_weights(int64_t total, const float* dist2, float* dist, float* weight, geot_stream_t stream) {
  GEOT_REQUIRE(total >= 0, "interp_weights: bad size %lld", (long long)total);
  if (total == 0) return GEOT_OK;
  GEOT_REQUIRE(dist2 && weight, "interp_weights: NULL pointer");
  const int grid = (int)std::min<int64_t>((total + kUpThreads - 1) / kUpThreads, (int64_t)num_sms() * 16);
  interp_weights_kernel<<<grid, kUpThreads, 0, as_stream(stream)>>>(total, dist2, dist, weight);
  count_launch(1);
  return check_launch("interp_weights_kernel");
}

GEOT_API int geot_interp_argmax(int B, int C, int m, int n, const float* points, const int* idx, const float* weight,
                                int64_t* pred, float* score, geot_stream_t stream) {
  GEOT_REQUIRE(B >= 0 && C >= 1 && m >= 1 && n >= 0 && B <= 65535, "interp_argmax: bad sizes B=%d C=%d m=%d n=%d", B, C, m, n);
  if (B == 0 || n == 0) return GEOT_OK;
  GEOT_REQUIRE(points && idx && weight && pred, "interp_argmax: NULL pointer");
  dim3 grid((n + kUpThreads - 1) / kUpThreads, B);
  interp_argmax_kernel<4><<<grid, kUpThreads, 0, as_stream(stream)>>>(C, m, n, points, idx, weight, pred, score);
  count_launch(1);
  return check_launch("interp_argmax_kernel");
}

GEOT_API int geot_seg_counts(int64_t n, int C, const int64_t* pred, const int64_t* label, int64_t* counts,
                             geot_stream_t stream) {
  GEOT_REQUIRE(n >= 0 && C >= 1 && C <= 4096, "seg_counts: bad sizes n=%lld C=%d", (long long)n, C);
  GEOT_REQUIRE(counts && (n == 0 || (pred && label)), "seg_counts: NULL pointer");
  cudaStream_t s = as_stream(stream);
  cudaError_t e = cudaMemsetAsync(counts, 0, (size_t)(3 * C + 2) * sizeof(int64_t), s);
  if (e != cudaSuccess) { set_error("seg_counts: memset: %s", cudaGetErrorString(e)); return GEOT_ERR_CUDA; }
  if (n == 0) return GEOT_OK;
  const int grid = (int)std::min<int64_t>((n + kUpThreads * 8 - 1) / (kUpThreads * 8), (int64_t)num_sms() * 4);
  seg_counts_kernel<<<std::max(grid, 1), kUpThreads, (size_t)(3 * C + 2) * sizeof(unsigned), s>>>(
      n, C, pred, label, reinterpret_cast<unsigned long long*>(counts));
  count_launch(1);
  return check_launch("seg_counts_kernel");
}
