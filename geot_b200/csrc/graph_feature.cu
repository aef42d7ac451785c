// geot_b200/csrc/graph_feature.cu — DGCNN edge features of the Point-Transformer decoder, sm_100a.
//
// Replaces the body of DGCNN_Propagation.get_graph_feature after the kNN (openpoints/models/backbone/transformer.py:
// 343-364): x_k.transpose().contiguous() -> [idx, :] row gather -> view/permute/contiguous -> expand -> subtract -> cat,
// i.e. three (B,C,Nq,k)-sized intermediates and a (B,Nk,C) transpose before the (B,2C,Nq,k) result exists.  Here one
// pass: a CTA stages one feature row x_k[b,c,:] in shared memory (Nk floats), reads each query's k neighbour indices
// once and writes both halves of the concatenation,
//     out[b, c, q, j] = x_k[b, c, idx[b,q,j]] - x_q[b, c, q]        out[b, C+c, q, j] = x_q[b, c, q],
// as 16-byte streaming stores when k = 4 (GeoT's value, transformer.py:313).  One fp32 subtraction per element: the
// result is bit-identical to the reference expression.  HBM: 4 C Nk + 4 C Nq + 8 C Nq k (write) + 4 Nq k (idx, L2).
#include <algorithm>
#include "common.cuh"

namespace geot {

void count_launch(int n);

constexpr int kGfThreads = 256;

// grid (ceil(Nq / q_tile), C, B).  STAGED: the row lives in shared memory, else gathers go to global memory (L2).
template <int K, bool STAGED>
__global__ void __launch_bounds__(kGfThreads)
graph_feature_kernel(int C, int Nk, int Nq, int k, int q_tile, const float* __restrict__ x_k,
                     const float* __restrict__ x_q, const int* __restrict__ idx, float* __restrict__ out) {
  extern __shared__ __align__(16) float gf_row[];
  const int t = threadIdx.x, c = blockIdx.y, b = blockIdx.z;
  const float* __restrict__ src = x_k + ((int64_t)b * C + c) * Nk;
  if (STAGED) {
    if ((Nk & 3) == 0 && ((uintptr_t)src & 15) == 0) {
      const float4* s4 = reinterpret_cast<const float4*>(src);
      float4* d4 = reinterpret_cast<float4*>(gf_row);
      for (int i = t; i < Nk / 4; i += kGfThreads) d4[i] = __ldg(s4 + i);
    } else {
      for (int i = t; i < Nk; i += kGfThreads) gf_row[i] = __ldg(src + i);
    }
    __syncthreads();
  }
  const float* __restrict__ row = STAGED ? gf_row : src;
  const int q0 = blockIdx.x * q_tile, q1 = min(Nq, q0 + q_tile);
  const float* __restrict__ xq_row = x_q + ((int64_t)b * C + c) * Nq;
  const int* __restrict__ id_b = idx + (int64_t)b * Nq * k;
  float* __restrict__ o_diff = out + (((int64_t)b * 2 * C + c) * Nq) * k;
  float* __restrict__ o_self = out + (((int64_t)b * 2 * C + C + c) * Nq) * k;
  for (int q = q0 + t; q < q1; q += kGfThreads) {
    const float xq = __ldg(xq_row + q);
    if (K == 4) {
      const int4 i4 = __ldg(reinterpret_cast<const int4*>(id_b) + q);
      float4 d;
      d.x = __fsub_rn(STAGED ? row[i4.x] : __ldg(row + i4.x), xq);
      d.y = __fsub_rn(STAGED ? row[i4.y] : __ldg(row + i4.y), xq);
      d.z = __fsub_rn(STAGED ? row[i4.z] : __ldg(row + i4.z), xq);
      d.w = __fsub_rn(STAGED ? row[i4.w] : __ldg(row + i4.w), xq);
      __stcs(reinterpret_cast<float4*>(o_diff) + q, d);
      __stcs(reinterpret_cast<float4*>(o_self) + q, make_float4(xq, xq, xq, xq));
    } else {
      const int* __restrict__ id = id_b + (int64_t)q * k;
      for (int j = 0; j < k; j++) {
        const int i = __ldg(id + j);
        __stcs(o_diff + (int64_t)q * k + j, __fsub_rn(STAGED ? row[i] : __ldg(row + i), xq));
        __stcs(o_self + (int64_t)q * k + j, xq);
      }
    }
  }
}

// grad_x_q[b,c,q] = sum_j (g[b,C+c,q,j] - g[b,c,q,j]) in ascending j (fixed order); thread per (b,c,q), coalesced along q*k
__global__ void __launch_bounds__(kGfThreads)
graph_feature_grad_q_kernel(int64_t total, int C, int Nq, int k, const float* __restrict__ grad_out, float* __restrict__ grad_x_q) {
  for (int64_t g = (int64_t)blockIdx.x * kGfThreads + threadIdx.x; g < total; g += (int64_t)gridDim.x * kGfThreads) {
    const int64_t bc = g / Nq, q = g - bc * Nq;
    const int64_t b = bc / C, c = bc - b * C;
    const float* __restrict__ gd = grad_out + (((b * 2 * C + c) * Nq) + q) * k;
    const float* __restrict__ gs = grad_out + (((b * 2 * C + C + c) * Nq) + q) * k;
    float acc = 0.f;
    for (int j = 0; j < k; j++) acc += __ldcs(gs + j) - __ldg(gd + j);  // the diff half is read again by the scatter
    grad_x_q[g] = acc;
  }
}

}  // namespace geot

using namespace geot;

GEOT_API int geot_graph_feature(int B, int C, int Nk, int Nq, int k, const float* x_k, const float* x_q, const int* idx,
                                float* out, geot_stream_t stream) {
  GEOT_REQUIRE(B >= 0 && C >= 0 && Nk >= 1 && Nq >= 0 && k >= 1, "graph_feature: bad sizes B=%d C=%d Nk=%d Nq=%d k=%d", B, C, Nk, Nq, k);
  if (B == 0 || C == 0 || Nq == 0) return GEOT_OK;
  GEOT_REQUIRE(x_k && x_q && idx && out, "graph_feature: NULL pointer");
  GEOT_REQUIRE(B <= 65535 && C <= 65535, "graph_feature: B or C exceeds grid limits");
  cudaStream_t s = as_stream(stream);
  const bool k4 = k == 4 && (((uintptr_t)idx | (uintptr_t)out) & 15) == 0;
  // stage the row when it is re-used enough: Nq*k gathers against Nk staged floats, and it fits
  const bool staged = (size_t)Nk * 4 <= 96 * 1024 && (int64_t)Nq * k * 4 >= Nk;
  int q_tile = Nq;
  if (staged) {  // several CTAs per row only when each still amortises its own copy of the row
    q_tile = std::max(2 * Nk / std::max(k, 1), 4 * kGfThreads);
    q_tile = std::min(q_tile, Nq);
  } else {
    q_tile = std::min(Nq, 8 * kGfThreads);
  }
  dim3 grid((Nq + q_tile - 1) / q_tile, C, B);
  const size_t smem = staged ? (size_t)Nk * 4 : 0;
  static DynSmem ds4, ds0;
  if (int rc = ds4.need(graph_feature_kernel<4, true>, 96 * 1024, "graph_feature")) return rc;
  if (int rc = ds0.need(graph_feature_kernel<0, true>, 96 * 1024, "graph_feature")) return rc;
  if (k4 && staged) graph_feature_kernel<4, true><<<grid, kGfThreads, smem, s>>>(C, Nk, Nq, k, q_tile, x_k, x_q, idx, out);
  else if (k4) graph_feature_kernel<4, false><<<grid, kGfThreads, 0, s>>>(C, Nk, Nq, k, q_tile, x_k, x_q, idx, out);
  else if (staged) graph_feature_kernel<0, true><<<grid, kGfThreads, smem, s>>>(C, Nk, Nq, k, q_tile, x_k, x_q, idx, out);
  else graph_feature_kernel<0, false><<<grid, kGfThreads, 0, s>>>(C, Nk, Nq, k, q_tile, x_k, x_q, idx, out);
  count_launch(1);
  return check_launch("graph_feature_kernel");
}

GEOT_API int geot_graph_feature_grad_q(int B, int C, int Nq, int k, const float* grad_out, float* grad_x_q,
                                       geot_stream_t stream) {
  GEOT_REQUIRE(B >= 0 && C >= 0 && Nq >= 0 && k >= 1, "graph_feature_grad_q: bad sizes");
  const int64_t total = (int64_t)B * C * Nq;
  if (total == 0) return GEOT_OK;
  GEOT_REQUIRE(grad_out && grad_x_q, "graph_feature_grad_q: NULL pointer");
  const int grid = (int)std::min<int64_t>((total + kGfThreads - 1) / kGfThreads, (int64_t)num_sms() * 32);
  graph_feature_grad_q_kernel<<<grid, kGfThreads, 0, as_stream(stream)>>>(total, C, Nq, k, grad_out, grad_x_q);
  count_launch(1);
  return check_launch("graph_feature_grad_q_kernel");
}
