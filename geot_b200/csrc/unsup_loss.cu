// geot_b200/csrc/unsup_loss.cu — unsupervised poly-1 focal loss on the transition-corrected logits, sm_100a.
//
// Replaces openpoints/loss/build.py:830-892 (Poly1FocalLoss_U_corr.forward, the criterion_u of
// cfgs/tooth_semi/transformer_finetune_fixmatch_ntm.yaml:54) and its autograd: sigmoid, one_hot + transpose,
// binary_cross_entropy_with_logits, the focal modulation, the poly-1 term, the confidence mask
// (.ge(thresh).unsqueeze(1).repeat(1,C,1)) and the masked mean are 8-10 passes over (B,C,N) tensors plus their
// saved activations.  Here: one pass forward (thread per point, loop over the C classes, coalesced along N;
// per-CTA double partials, fixed-order final sum - deterministic), one pass backward that recomputes the
// element from the logit.  exp / log1p / pow are the CUDA libm functions torch's own kernels call.
#include <algorithm>
#include "common.cuh"

namespace geot {

void count_launch(int n);

// torch.pow(tensor, scalar) takes x*x, x*x*x, sqrt(x) and x for the exponents 2, 3, 0.5 and 1 (PowKernel.cu); gamma = 2
// in every shipped config, so the default path carries no transcendental pow at all
__device__ __forceinline__ float pow_scalar(float x, float e) {
  if (e == 2.0f) return x * x;
  if (e == 3.0f) return (x * x) * x;
  if (e == 1.0f) return x;
  if (e == 0.5f) return sqrtf(x);
  if (e == 0.0f) return 1.0f;
  return powf(x, e);
}

struct Poly1 {
  float value, dvalue;  // poly1 term of one (point, class) element and its derivative w.r.t. the logit
};
__device__ __forceinline__ Poly1 poly1_element(float x, bool y, float eps, float alpha, float gamma, bool want_grad) {
  // torch.sigmoid; binary_cross_entropy_with_logits: (1 - y) * x + max(-x, 0) + log1p(exp(-|x|))
  const float p = 1.0f / (1.0f + expf(-x));
  const float ce = (y ? 0.f : x) + fmaxf(-x, 0.f) + log1pf(expf(-fabsf(x)));
  const float pt = y ? p : 1.0f - p;          // labels * p + (1 - labels) * (1 - p)
  const float om = 1.0f - pt;
  const float mod = pow_scalar(om, gamma);     // (1 - pt) ** gamma
  const float at = alpha >= 0.f ? (y ? alpha : 1.0f - alpha) : 1.0f;
  Poly1 r;
  r.value = at * (ce * mod) + eps * pow_scalar(om, gamma + 1.0f);
  r.dvalue = 0.f;
  if (want_grad) {
    const float dp = p * (1.0f - p);
    const float dpt = y ? dp : -dp;
    const float dce = p - (y ? 1.0f : 0.f);
    // d/dx (1-pt)^g = -g (1-pt)^(g-1) dpt ; powf(0, g-1) with g < 1 would be inf: (1-pt) > 0 for finite logits
    const float dmod = -gamma * pow_scalar(om, gamma - 1.0f) * dpt;
    r.dvalue = at * (dce * mod + ce * dmod) - eps * (gamma + 1.0f) * mod * dpt;
  }
  return r;
}
__device__ __forceinline__ float keep_of(const float* conf, const float* mask, float thresh, int64_t g) {
  return mask != nullptr ? mask[g] : (conf[g] >= thresh ? 1.0f : 0.f);
}

// Poly1FocalLoss_U_T (build.py:661-670): beta = logits_pred / pred_u_t[b, label, n]
__device__ __forceinline__ float beta_of(const float* conf, const float* pred_t, int64_t b, int C, int N, int lab, int64_t n) {
  return conf[b * N + n] / __ldg(pred_t + (b * C + lab) * N + n);
}

constexpr int kLossThreads = 256;
// grid-stride over points; partial[blockIdx] = {sum of kept poly1 terms, sum of keep}
__global__ void __launch_bounds__(kLossThreads)
poly1_fwd_kernel(int B, int C, int N, const float* __restrict__ logits, const int64_t* __restrict__ labels,
                 const float* __restrict__ conf, const float* __restrict__ mask, float thresh, float eps, float alpha,
                 float gamma, const float* __restrict__ pred_t, double* __restrict__ partial) {
  __shared__ double s_a[kLossThreads / 32], s_b[kLossThreads / 32];
  const int64_t BN = (int64_t)B * N;
  double acc = 0.0, cnt = 0.0;
  for (int64_t g = (int64_t)blockIdx.x * kLossThreads + threadIdx.x; g < BN; g += (int64_t)gridDim.x * kLossThreads) {
    const float k = keep_of(conf, mask, thresh, g);
    cnt += (double)k;
    if (k != 0.f) {
      const int64_t b = g / N, n = g - b * N;
      const int lab = (int)labels[g];
      const float* __restrict__ x = logits + b * C * N + n;
      float s = 0.f;
      for (int c = 0; c < C; c++) s += poly1_element(__ldg(x + (int64_t)c * N), c == lab, eps, alpha, gamma, false).value;
      if (pred_t != nullptr) s *= beta_of(conf, pred_t, b, C, N, lab, n);  // Poly1FocalLoss_U_T: poly1 * beta
      acc += (double)(s * k);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    acc += __shfl_xor_sync(0xffffffffu, acc, o);
    cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  }
  if ((threadIdx.x & 31) == 0) { s_a[threadIdx.x >> 5] = acc; s_b[threadIdx.x >> 5] = cnt; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0.0, c = 0.0;
    for (int w = 0; w < kLossThreads / 32; w++) { a += s_a[w]; c += s_b[w]; }
    partial[2 * blockIdx.x] = a;
    partial[2 * blockIdx.x + 1] = c;
  }
}
// loss = sum / (C * kept + 0.001); stats = {loss, denominator}
__global__ void poly1_final_kernel(int nparts, int C, const double* __restrict__ partial, float* __restrict__ loss,
                                   float* __restrict__ denom) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  double a = 0.0, c = 0.0;
  for (int i = 0; i < nparts; i++) { a += partial[2 * i]; c += partial[2 * i + 1]; }
  const float d = (float)(c * (double)C) + 0.001f;  // thresh_mask.sum() + 0.001 in fp32
  loss[0] = (float)a / d;
  denom[0] = d;
}
__global__ void __launch_bounds__(kLossThreads)
poly1_bwd_kernel(int B, int C, int N, const float* __restrict__ logits, const int64_t* __restrict__ labels,
                 const float* __restrict__ conf, const float* __restrict__ mask, float thresh, float eps, float alpha,
                 float gamma, const float* __restrict__ denom, const float* __restrict__ grad_loss,
                 float* __restrict__ grad_logits, const float* __restrict__ pred_t, float* __restrict__ grad_pred_t) {
  const int64_t BN = (int64_t)B * N;
  const float scale = grad_loss[0] / denom[0];
  for (int64_t g = (int64_t)blockIdx.x * kLossThreads + threadIdx.x; g < BN; g += (int64_t)gridDim.x * kLossThreads) {
    const float k = keep_of(conf, mask, thresh, g) * scale;
    const int64_t b = g / N, n = g - b * N;
    const int lab = (int)labels[g];
    const float* __restrict__ x = logits + b * C * N + n;
    float* __restrict__ o = grad_logits + b * C * N + n;
    if (pred_t == nullptr) {
      for (int c = 0; c < C; c++)
        o[(int64_t)c * N] = k != 0.f ? k * poly1_element(__ldg(x + (int64_t)c * N), c == lab, eps, alpha, gamma, true).dvalue : 0.f;
    } else {  // loss element = beta * sum_c poly1, beta = conf / after: d/d logit = beta * dpoly1, d/d after = -beta / after * sum
      float* __restrict__ ot = grad_pred_t + b * C * N + n;
      const float after = __ldg(pred_t + (b * C + lab) * N + n);
      const float beta = conf[g] / after;
      float sum = 0.f;
      for (int c = 0; c < C; c++) {
        float d = 0.f;
        if (k != 0.f) {
          const Poly1 e = poly1_element(__ldg(x + (int64_t)c * N), c == lab, eps, alpha, gamma, true);
          d = k * beta * e.dvalue;
          sum += e.value;
        }
        o[(int64_t)c * N] = d;
        ot[(int64_t)c * N] = 0.f;
      }
      if (k != 0.f) ot[(int64_t)lab * N] = -k * sum * (beta / after);
    }
  }
}

static int loss_grid(int64_t BN) {
  const int64_t want = (BN + kLossThreads - 1) / kLossThreads;
  return (int)std::max<int64_t>(1, std::min<int64_t>(want, (int64_t)num_sms() * 8));
}

}  // namespace geot

using namespace geot;

GEOT_API size_t geot_poly1_focal_workspace_bytes(int64_t BN) { return BN > 0 ? (size_t)loss_grid(BN) * 2 * sizeof(double) + 256 : 0; }

static int poly1_forward_impl(const char* who, int B, int C, int N, const float* logits, const int64_t* labels, const float* conf,
                              const float* mask, const float* pred_t, float thresh, float epsilon, float alpha, float gamma,
                              float* loss, float* denom, void* workspace, size_t workspace_bytes, geot_stream_t stream) {
  GEOT_REQUIRE(B >= 1 && C >= 1 && N >= 1, "%s: bad sizes B=%d C=%d N=%d", who, B, C, N);
  GEOT_REQUIRE(logits && labels && (conf || mask) && loss && denom, "%s: NULL pointer", who);
  GEOT_REQUIRE(pred_t == nullptr || conf != nullptr, "%s: the beta factor needs logits_pred (conf)", who);
  const int64_t BN = (int64_t)B * N;
  if (workspace == nullptr || workspace_bytes < geot_poly1_focal_workspace_bytes(BN)) {
    set_error("%s: workspace %zu < %zu bytes", who, workspace_bytes, geot_poly1_focal_workspace_bytes(BN));
    return GEOT_ERR_WORKSPACE;
  }
  cudaStream_t s = as_stream(stream);
  const int grid = loss_grid(BN);
  double* partial = reinterpret_cast<double*>(workspace);
  // with a mask the keep weights come from the mask; conf is then only the numerator of beta
  poly1_fwd_kernel<<<grid, kLossThreads, 0, s>>>(B, C, N, logits, labels, conf, mask, thresh, epsilon, alpha, gamma, pred_t, partial);
  count_launch(1);
  int rc = check_launch("poly1_fwd_kernel");
  if (rc) return rc;
  poly1_final_kernel<<<1, 32, 0, s>>>(grid, C, partial, loss, denom);
  count_launch(1);
  return check_launch("poly1_final_kernel");
}

static int poly1_backward_impl(const char* who, int B, int C, int N, const float* logits, const int64_t* labels, const float* conf,
                               const float* mask, const float* pred_t, float thresh, float epsilon, float alpha, float gamma,
                               const float* denom, const float* grad_loss, float* grad_logits, float* grad_pred_t,
                               geot_stream_t stream) {
  GEOT_REQUIRE(B >= 1 && C >= 1 && N >= 1, "%s: bad sizes B=%d C=%d N=%d", who, B, C, N);
  GEOT_REQUIRE(logits && labels && (conf || mask) && denom && grad_loss && grad_logits, "%s: NULL pointer", who);
  GEOT_REQUIRE(pred_t == nullptr || (conf != nullptr && grad_pred_t != nullptr), "%s: NULL pointer (beta path)", who);
  poly1_bwd_kernel<<<loss_grid((int64_t)B * N), kLossThreads, 0, as_stream(stream)>>>(
      B, C, N, logits, labels, conf, mask, thresh, epsilon, alpha, gamma, denom, grad_loss, grad_logits, pred_t, grad_pred_t);
  count_launch(1);
  return check_launch("poly1_bwd_kernel");
}

GEOT_API int geot_poly1_focal_forward(int B, int C, int N, const float* logits, const int64_t* labels, const float* conf,
                                      const float* mask, float thresh, float epsilon, float alpha, float gamma,
                                      float* loss, float* denom, void* workspace, size_t workspace_bytes,
                                      geot_stream_t stream) {
  return poly1_forward_impl("poly1_focal_forward", B, C, N, logits, labels, conf, mask, nullptr, thresh, epsilon, alpha, gamma, loss,
                            denom, workspace, workspace_bytes, stream);
}

GEOT_API int geot_poly1_focal_backward(int B, int C, int N, const float* logits, const int64_t* labels, const float* conf,
                                       const float* mask, float thresh, float epsilon, float alpha, float gamma,
                                       const float* denom, const float* grad_loss, float* grad_logits,
                                       geot_stream_t stream) {
  return poly1_backward_impl("poly1_focal_backward", B, C, N, logits, labels, conf, mask, nullptr, thresh, epsilon, alpha, gamma,
                             denom, grad_loss, grad_logits, nullptr, stream);
}

GEOT_API int geot_poly1_focal_t_forward(int B, int C, int N, const float* logits, const int64_t* labels, const float* conf,
                                        const float* mask, const float* pred_t, float thresh, float epsilon, float alpha,
                                        float gamma, float* loss, float* denom, void* workspace, size_t workspace_bytes,
                                        geot_stream_t stream) {
  GEOT_REQUIRE(pred_t != nullptr && conf != nullptr, "poly1_focal_t_forward: NULL pointer");
  return poly1_forward_impl("poly1_focal_t_forward", B, C, N, logits, labels, conf, mask, pred_t, thresh, epsilon, alpha, gamma, loss,
                            denom, workspace, workspace_bytes, stream);
}

GEOT_API int geot_poly1_focal_t_backward(int B, int C, int N, const float* logits, const int64_t* labels, const float* conf,
                                         const float* mask, const float* pred_t, float thresh, float epsilon, float alpha,
                                         float gamma, const float* denom, const float* grad_loss, float* grad_logits,
                                         float* grad_pred_t, geot_stream_t stream) {
  GEOT_REQUIRE(pred_t != nullptr && conf != nullptr, "poly1_focal_t_backward: NULL pointer");
  return poly1_backward_impl("poly1_focal_t_backward", B, C, N, logits, labels, conf, mask, pred_t, thresh, epsilon, alpha, gamma,
                             denom, grad_loss, grad_logits, grad_pred_t, stream);
}
