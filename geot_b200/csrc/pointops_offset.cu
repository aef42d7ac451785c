// geot_b200/csrc/pointops_offset.cu — channel-last ("offset format") ops of pointops_cuda for sm_100a.
//
// Replaces openpoints/cpp/pointops/src/
//   grouping/grouping_cuda_kernel.cu:5-25          grouping fwd / bwd (atomicAdd :24)
//   interpolation/interpolation_cuda_kernel.cu:5-33   interpolation fwd / bwd (atomicAdd :31)
//   subtraction/subtraction_cuda_kernel.cu:5-30    subtraction fwd / bwd (2 atomicAdds)
//   aggregation/aggregation_cuda_kernel.cu:5-39    aggregation fwd / bwd (2 atomicAdds + direct write)
//
// Rows are channel-last, so a gathered row is contiguous: one thread moves VEC = 4 channels with
// 16-byte loads/stores when c % 4 == 0 (the reference moves 4 bytes per thread and spends three integer
// div/mods per element).  All backward scatter-adds go through the deterministic scatter plan of
// group.cu (B = 1): each target row sums its sources in ascending source order.
#include "common.cuh"

namespace geot {

void count_launch(int n);

template <int VEC> struct Vec;
template <> struct Vec<4> { using T = float4; };
template <> struct Vec<1> { using T = float; };

__device__ __forceinline__ float4 ld4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ void st4_stream(float* p, float4 v) { __stcs(reinterpret_cast<float4*>(p), v); }

// ---- grouping fwd: output (m*ns, c) = input[idx] ------------------------------------------------
template <int VEC>
__global__ void __launch_bounds__(256)
grouping_fwd_kernel(int64_t rows, int c, const float* __restrict__ input, const int* __restrict__ idx,
                    float* __restrict__ output) {
  const int cv = c / VEC;
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= rows * cv) return;
  const int64_t r = g / cv;
  const int ci = (int)(g - r * cv) * VEC;
  const int64_t src = (int64_t)__ldg(idx + r) * c + ci;
  if (VEC == 4) st4_stream(output + r * c + ci, ld4(input + src));
  else __stcs(output + r * c + ci, __ldg(input + src));
}

// ---- generic channel-last segmented reduction ----------------------------------------------------
//   dst[t, ci] = sum_{p in seg(t)} sign * g[(perm[p] / DIV), ci] * (WEIGHTED ? w[perm[p]] : 1)
template <int VEC, bool WEIGHTED>
__global__ void __launch_bounds__(256)
seg_reduce_clast_kernel(int n_targets, int c, int div, float sign, const float* __restrict__ g,
                        const float* __restrict__ w, const int* __restrict__ row_ptr,
                        const int* __restrict__ perm, float* __restrict__ dst) {
  const int cv = c / VEC;
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= (int64_t)n_targets * cv) return;
  const int t = (int)(gid / cv);
  const int ci = (int)(gid - (int64_t)t * cv) * VEC;
  const int s = row_ptr[t], e = row_ptr[t + 1];
  float acc[VEC];
#pragma unroll
  for (int v = 0; v < VEC; v++) acc[v] = 0.f;
  for (int p = s; p < e; p++) {
    const int src = __ldg(perm + p);
    const float wt = WEIGHTED ? __fmul_rn(sign, __ldg(w + src)) : sign;
    const float* gp = g + (int64_t)(src / div) * c + ci;
    if (VEC == 4) {
      const float4 v = ld4(gp);
      acc[0] = __fadd_rn(acc[0], __fmul_rn(v.x, wt));
      acc[1 % VEC] = __fadd_rn(acc[1 % VEC], __fmul_rn(v.y, wt));
      acc[2 % VEC] = __fadd_rn(acc[2 % VEC], __fmul_rn(v.z, wt));
      acc[3 % VEC] = __fadd_rn(acc[3 % VEC], __fmul_rn(v.w, wt));
    } else {
      acc[0] = __fadd_rn(acc[0], __fmul_rn(__ldg(gp), wt));
    }
  }
#pragma unroll
  for (int v = 0; v < VEC; v++) dst[(int64_t)t * c + ci + v] = acc[v];
}

struct PlanPtrs { const int* row_ptr; const int* perm; };
static inline PlanPtrs plan_ptrs(const void* plan, int N) {
  size_t rb = ((size_t)(N + 1) * sizeof(int) + 255) & ~(size_t)255;
  return PlanPtrs{reinterpret_cast<const int*>(plan),
                  reinterpret_cast<const int*>(reinterpret_cast<const char*>(plan) + rb)};
}

static inline bool vec_ok(int c, const void* a, const void* b) {
  return (c % 4 == 0) && (((uintptr_t)a & 15) == 0) && (((uintptr_t)b & 15) == 0);
}

template <bool WEIGHTED>
static int launch_clast_reduce(int n_targets, int c, int div, float sign, const float* g, const float* w,
                               const void* plan, float* dst, cudaStream_t s, const char* what) {
  if (n_targets == 0 || c == 0) return GEOT_OK;
  GEOT_REQUIRE(g && plan && dst, "%s: NULL pointer", what);
  const PlanPtrs pp = plan_ptrs(plan, n_targets);
  if (vec_ok(c, g, dst)) {
    const int64_t tot = (int64_t)n_targets * (c / 4);
    seg_reduce_clast_kernel<4, WEIGHTED><<<(unsigned)((tot + 255) / 256), 256, 0, s>>>(n_targets, c, div, sign, g, w, pp.row_ptr, pp.perm, dst);
  } else {
    const int64_t tot = (int64_t)n_targets * c;
    seg_reduce_clast_kernel<1, WEIGHTED><<<(unsigned)((tot + 255) / 256), 256, 0, s>>>(n_targets, c, div, sign, g, w, pp.row_ptr, pp.perm, dst);
  }
  count_launch(1);
  return check_launch(what);
}

// ---- interpolation fwd: out[p,ci] = fma(in[idx[p,i],ci], w[p,i], out) for i = 0..k-1 ----------------
template <int VEC>
__global__ void __launch_bounds__(256)
interpolation_fwd_kernel(int n, int c, int k, const float* __restrict__ input, const int* __restrict__ idx,
                         const float* __restrict__ weight, float* __restrict__ output) {
  const int cv = c / VEC;
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= (int64_t)n * cv) return;
  const int p = (int)(g / cv);
  const int ci = (int)(g - (int64_t)p * cv) * VEC;
  float acc[VEC];
#pragma unroll
  for (int v = 0; v < VEC; v++) acc[v] = 0.f;
  for (int i = 0; i < k; i++) {
    const float w = __ldg(weight + (int64_t)p * k + i);
    const float* src = input + (int64_t)__ldg(idx + (int64_t)p * k + i) * c + ci;
    if (VEC == 4) {
      const float4 v = ld4(src);
      acc[0] = __fmaf_rn(v.x, w, acc[0]); acc[1 % VEC] = __fmaf_rn(v.y, w, acc[1 % VEC]);
      acc[2 % VEC] = __fmaf_rn(v.z, w, acc[2 % VEC]); acc[3 % VEC] = __fmaf_rn(v.w, w, acc[3 % VEC]);
    } else {
      acc[0] = __fmaf_rn(__ldg(src), w, acc[0]);
    }
  }
#pragma unroll
  for (int v = 0; v < VEC; v++) output[(int64_t)p * c + ci + v] = acc[v];
}

// ---- subtraction fwd: out[(p,s),ci] = in1[p,ci] - in2[idx[p,s],ci] --------------------------------
template <int VEC>
__global__ void __launch_bounds__(256)
subtraction_fwd_kernel(int n, int ns, int c, const float* __restrict__ in1, const float* __restrict__ in2,
                       const int* __restrict__ idx, float* __restrict__ output) {
  const int cv = c / VEC;
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= (int64_t)n * ns * cv) return;
  const int64_t r = g / cv;  // (p, s)
  const int ci = (int)(g - r * cv) * VEC;
  const int p = (int)(r / ns);
  const int64_t s2 = (int64_t)__ldg(idx + r) * c + ci;
  if (VEC == 4) {
    const float4 a = ld4(in1 + (int64_t)p * c + ci), b = ld4(in2 + s2);
    st4_stream(output + r * c + ci, make_float4(__fsub_rn(a.x, b.x), __fsub_rn(a.y, b.y), __fsub_rn(a.z, b.z), __fsub_rn(a.w, b.w)));
  } else {
    __stcs(output + r * c + ci, __fsub_rn(__ldg(in1 + (int64_t)p * c + ci), __ldg(in2 + s2)));
  }
}
// grad_input1[p,ci] = sum_s g[p,s,ci]  (fixed s order)
__global__ void __launch_bounds__(256)
subtraction_bwd1_kernel(int n, int ns, int c, const float* __restrict__ g, float* __restrict__ gi1) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)n * c) return;
  const int p = (int)(t / c), ci = (int)(t - (int64_t)p * c);
  float acc = 0.f;
  for (int s = 0; s < ns; s++) acc = __fadd_rn(acc, __ldg(g + ((int64_t)p * ns + s) * c + ci));
  gi1[t] = acc;
}

__global__ void __launch_bounds__(256)
subtraction_bwd1_vec_kernel(int n, int ns, int c, const float* __restrict__ g, float* __restrict__ gi1) {
  const int c4 = c >> 2;
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)n * c4) return;
  const int p = (int)(t / c4), ci = (int)(t - (int64_t)p * c4) * 4;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 8
  for (int s = 0; s < ns; s++) {
    const float4 v = ld4(g + ((int64_t)p * ns + s) * c + ci);
    acc.x = __fadd_rn(acc.x, v.x); acc.y = __fadd_rn(acc.y, v.y); acc.z = __fadd_rn(acc.z, v.z); acc.w = __fadd_rn(acc.w, v.w);
  }
  *reinterpret_cast<float4*>(gi1 + (int64_t)p * c + ci) = acc;
}

// ---- aggregation fwd: out[p,ci] = sum_s (in[idx[p,s],ci] + pos[p,s,ci]) * w[p,s,ci % w_c] ----------
__global__ void __launch_bounds__(256)
aggregation_fwd_kernel(int n, int ns, int c, int w_c, const float* __restrict__ input,
                       const float* __restrict__ position, const float* __restrict__ weight,
                       const int* __restrict__ idx, float* __restrict__ output) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)n * c) return;
  const int p = (int)(t / c), ci = (int)(t - (int64_t)p * c);
  const int wi = ci % w_c;
  float acc = 0.f;
  for (int s = 0; s < ns; s++) {
    const int64_t r = (int64_t)p * ns + s;
    const float a = __fadd_rn(__ldg(input + (int64_t)__ldg(idx + r) * c + ci), __ldg(position + r * c + ci));
    acc = __fmaf_rn(a, __ldg(weight + r * w_c + wi), acc);
  }
  output[t] = acc;
}
// grad_position[p,s,ci] = g[p,ci] * w[p,s,ci%w_c]
__global__ void __launch_bounds__(256)
aggregation_bwd_pos_kernel(int n, int ns, int c, int w_c, const float* __restrict__ weight,
                           const float* __restrict__ g, float* __restrict__ gpos) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)n * ns * c) return;
  const int64_t r = t / c;
  const int ci = (int)(t - r * c);
  const int p = (int)(r / ns);
  gpos[t] = __fmul_rn(__ldg(g + (int64_t)p * c + ci), __ldg(weight + r * w_c + ci % w_c));
}
// grad_weight[p,s,wi] = sum_{ci = wi, wi+w_c, ...} g[p,ci] * (in[idx[p,s],ci] + pos[p,s,ci])  (ascending ci)
__global__ void __launch_bounds__(256)
aggregation_bwd_w_kernel(int n, int ns, int c, int w_c, const float* __restrict__ input,
                         const float* __restrict__ position, const int* __restrict__ idx,
                         const float* __restrict__ g, float* __restrict__ gw) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)n * ns * w_c) return;
  const int64_t r = t / w_c;
  const int wi = (int)(t - r * w_c);
  const int p = (int)(r / ns);
  const int64_t src = (int64_t)__ldg(idx + r) * c;
  float acc = 0.f;
  for (int ci = wi; ci < c; ci += w_c)
    acc = __fadd_rn(acc, __fmul_rn(__ldg(g + (int64_t)p * c + ci), __fadd_rn(__ldg(input + src + ci), __ldg(position + r * c + ci))));
  gw[t] = acc;
}
// grad_input[t,ci] = sum_{p in seg(t)} g[perm/ns, ci] * w[perm, ci % w_c]
__global__ void __launch_bounds__(256)
aggregation_bwd_in_kernel(int n, int ns, int c, int w_c, const float* __restrict__ weight,
                          const float* __restrict__ g, const int* __restrict__ row_ptr,
                          const int* __restrict__ perm, float* __restrict__ gin) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)n * c) return;
  const int tgt = (int)(t / c), ci = (int)(t - (int64_t)tgt * c);
  const int wi = ci % w_c;
  float acc = 0.f;
  for (int p = row_ptr[tgt]; p < row_ptr[tgt + 1]; p++) {
    const int src = __ldg(perm + p);
    acc = __fadd_rn(acc, __fmul_rn(__ldg(g + (int64_t)(src / ns) * c + ci), __ldg(weight + (int64_t)src * w_c + wi)));
  }
  gin[t] = acc;
}

// ---- float4 variants (c % 4 == 0, w_c % 4 == 0, 16-byte aligned): the scalar kernels above mirror the reference's
// thread-per-(point, channel) mapping and run at its speed; here a thread owns 4 consecutive channels, whose weights
// w[.., ci % w_c] are 4 consecutive floats as well, so every access is a 16-byte vector.
__global__ void __launch_bounds__(256)
aggregation_fwd_vec_kernel(int n, int ns, int c, int w_c, const float* __restrict__ input,
                           const float* __restrict__ position, const float* __restrict__ weight,
                           const int* __restrict__ idx, float* __restrict__ output) {
  const int c4 = c >> 2;
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)n * c4) return;
  const int p = (int)(t / c4), ci = (int)(t - (int64_t)p * c4) * 4;
  const int wi = ci % w_c;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
  for (int s = 0; s < ns; s++) {
    const int64_t r = (int64_t)p * ns + s;
    const float4 a = ld4(input + (int64_t)__ldg(idx + r) * c + ci);
    const float4 q = ld4(position + r * c + ci);
    const float4 w = ld4(weight + r * w_c + wi);
    acc.x = __fmaf_rn(__fadd_rn(a.x, q.x), w.x, acc.x);
    acc.y = __fmaf_rn(__fadd_rn(a.y, q.y), w.y, acc.y);
    acc.z = __fmaf_rn(__fadd_rn(a.z, q.z), w.z, acc.z);
    acc.w = __fmaf_rn(__fadd_rn(a.w, q.w), w.w, acc.w);
  }
  *reinterpret_cast<float4*>(output + (int64_t)p * c + ci) = acc;
}

// grad_position and grad_weight in one pass over (p, s): the T = c/4 threads of a row are T consecutive lanes;
// lanes whose channel groups share a weight slot (q % G equal, G = w_c/4) are summed with a fixed xor tree.
__global__ void __launch_bounds__(256)
aggregation_bwd_pw_vec_kernel(int n, int ns, int c, int w_c, const float* __restrict__ input,
                              const float* __restrict__ position, const float* __restrict__ weight,
                              const int* __restrict__ idx, const float* __restrict__ g,
                              float* __restrict__ gpos, float* __restrict__ gw) {
  const int T = c >> 2, G = w_c >> 2;
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t rows = (int64_t)n * ns;
  const bool live = t < rows * T;
  const int64_t r = live ? t / T : 0;
  const int q = live ? (int)(t - r * T) : 0;
  const int ci = q * 4, wi = ci % w_c;
  const int p = (int)(r / ns);
  float4 part = make_float4(0.f, 0.f, 0.f, 0.f);
  if (live) {
    const float4 go = ld4(g + (int64_t)p * c + ci);
    const float4 w = ld4(weight + r * w_c + wi);
    __stcs(reinterpret_cast<float4*>(gpos + r * c + ci), make_float4(__fmul_rn(go.x, w.x), __fmul_rn(go.y, w.y), __fmul_rn(go.z, w.z), __fmul_rn(go.w, w.w)));
    const float4 a = ld4(input + (int64_t)__ldg(idx + r) * c + ci);
    const float4 ps = ld4(position + r * c + ci);
    part = make_float4(__fmul_rn(go.x, __fadd_rn(a.x, ps.x)), __fmul_rn(go.y, __fadd_rn(a.y, ps.y)),
                       __fmul_rn(go.z, __fadd_rn(a.z, ps.z)), __fmul_rn(go.w, __fadd_rn(a.w, ps.w)));
  }
  for (int o = T >> 1; o >= G; o >>= 1) {  // T and G are powers of two, T <= 32: rows never straddle a warp
    part.x = __fadd_rn(part.x, __shfl_xor_sync(0xffffffffu, part.x, o));
    part.y = __fadd_rn(part.y, __shfl_xor_sync(0xffffffffu, part.y, o));
    part.z = __fadd_rn(part.z, __shfl_xor_sync(0xffffffffu, part.z, o));
    part.w = __fadd_rn(part.w, __shfl_xor_sync(0xffffffffu, part.w, o));
  }
  if (live && q < G) *reinterpret_cast<float4*>(gw + r * w_c + q * 4) = part;
}

__global__ void __launch_bounds__(256)
aggregation_bwd_in_vec_kernel(int n, int ns, int c, int w_c, const float* __restrict__ weight,
                              const float* __restrict__ g, const int* __restrict__ row_ptr,
                              const int* __restrict__ perm, float* __restrict__ gin) {
  const int c4 = c >> 2;
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)n * c4) return;
  const int tgt = (int)(t / c4), ci = (int)(t - (int64_t)tgt * c4) * 4;
  const int wi = ci % w_c;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  const int p0 = row_ptr[tgt], p1 = row_ptr[tgt + 1];
#pragma unroll 4
  for (int p = p0; p < p1; p++) {
    const int src = __ldg(perm + p);
    const float4 go = ld4(g + (int64_t)(src / ns) * c + ci);
    const float4 w = ld4(weight + (int64_t)src * w_c + wi);
    acc.x = __fadd_rn(acc.x, __fmul_rn(go.x, w.x)); acc.y = __fadd_rn(acc.y, __fmul_rn(go.y, w.y));
    acc.z = __fadd_rn(acc.z, __fmul_rn(go.z, w.z)); acc.w = __fadd_rn(acc.w, __fmul_rn(go.w, w.w));
  }
  *reinterpret_cast<float4*>(gin + (int64_t)tgt * c + ci) = acc;
}

static inline bool pow2(int v) { return v > 0 && (v & (v - 1)) == 0; }
static inline bool aligned16(const void* a, const void* b, const void* c2, const void* d) {
  return ((((uintptr_t)a) | ((uintptr_t)b) | ((uintptr_t)c2) | ((uintptr_t)d)) & 15) == 0;
}

static inline unsigned blocks_for(int64_t tot) { return (unsigned)((tot + 255) / 256); }

}  // namespace geot

using namespace geot;

GEOT_API int geot_grouping_forward(int m, int nsample, int c, const float* input, const int* idx,
                                   float* output, geot_stream_t stream) {
  GEOT_REQUIRE(m >= 0 && nsample >= 0 && c >= 0, "grouping_forward: bad sizes");
  const int64_t rows = (int64_t)m * nsample;
  if (rows == 0 || c == 0) return GEOT_OK;
  GEOT_REQUIRE(input && idx && output, "grouping_forward: NULL pointer");
  if (vec_ok(c, input, output)) grouping_fwd_kernel<4><<<blocks_for(rows * (c / 4)), 256, 0, as_stream(stream)>>>(rows, c, input, idx, output);
  else grouping_fwd_kernel<1><<<blocks_for(rows * c), 256, 0, as_stream(stream)>>>(rows, c, input, idx, output);
  count_launch(1);
  return check_launch("grouping_fwd_kernel");
}

GEOT_API int geot_grouping_backward(int m, int nsample, int c, int n, const float* grad_output,
                                    const void* plan, float* grad_input, geot_stream_t stream) {
  GEOT_REQUIRE(m >= 0 && nsample >= 0 && c >= 0 && n >= 0, "grouping_backward: bad sizes");
  return launch_clast_reduce<false>(n, c, 1, 1.f, grad_output, nullptr, plan, grad_input, as_stream(stream), "grouping_backward");
}

GEOT_API int geot_interpolation_forward(int n, int c, int k, const float* input, const int* idx,
                                        const float* weight, float* output, geot_stream_t stream) {
  GEOT_REQUIRE(n >= 0 && c >= 0 && k >= 0, "interpolation_forward: bad sizes");
  if (n == 0 || c == 0) return GEOT_OK;
  GEOT_REQUIRE(output && (k == 0 || (input && idx && weight)), "interpolation_forward: NULL pointer");
  if (vec_ok(c, input, output)) interpolation_fwd_kernel<4><<<blocks_for((int64_t)n * (c / 4)), 256, 0, as_stream(stream)>>>(n, c, k, input, idx, weight, output);
  else interpolation_fwd_kernel<1><<<blocks_for((int64_t)n * c), 256, 0, as_stream(stream)>>>(n, c, k, input, idx, weight, output);
  count_launch(1);
  return check_launch("interpolation_fwd_kernel");
}

GEOT_API int geot_interpolation_backward(int n, int c, int k, int m, const float* grad_output,
                                         const float* weight, const void* plan, float* grad_input,
                                         geot_stream_t stream) {
  GEOT_REQUIRE(n >= 0 && c >= 0 && k >= 1 && m >= 0, "interpolation_backward: bad sizes");
  return launch_clast_reduce<true>(m, c, k, 1.f, grad_output, weight, plan, grad_input, as_stream(stream), "interpolation_backward");
}

GEOT_API int geot_subtraction_forward(int n, int nsample, int c, const float* input1, const float* input2,
                                      const int* idx, float* output, geot_stream_t stream) {
  GEOT_REQUIRE(n >= 0 && nsample >= 0 && c >= 0, "subtraction_forward: bad sizes");
  const int64_t rows = (int64_t)n * nsample;
  if (rows == 0 || c == 0) return GEOT_OK;
  GEOT_REQUIRE(input1 && input2 && idx && output, "subtraction_forward: NULL pointer");
  if (vec_ok(c, input1, output) && (((uintptr_t)input2 & 15) == 0)) subtraction_fwd_kernel<4><<<blocks_for(rows * (c / 4)), 256, 0, as_stream(stream)>>>(n, nsample, c, input1, input2, idx, output);
  else subtraction_fwd_kernel<1><<<blocks_for(rows * c), 256, 0, as_stream(stream)>>>(n, nsample, c, input1, input2, idx, output);
  count_launch(1);
  return check_launch("subtraction_fwd_kernel");
}

GEOT_API int geot_subtraction_backward(int n, int nsample, int c, const float* grad_output,
                                       const void* plan, float* grad_input1, float* grad_input2,
                                       geot_stream_t stream) {
  GEOT_REQUIRE(n >= 0 && nsample >= 1 && c >= 0, "subtraction_backward: bad sizes");
  if (n == 0 || c == 0) return GEOT_OK;
  GEOT_REQUIRE(grad_output && grad_input1 && grad_input2, "subtraction_backward: NULL pointer");
  if (c % 4 == 0 && ((((uintptr_t)grad_output) | ((uintptr_t)grad_input1)) & 15) == 0)
    subtraction_bwd1_vec_kernel<<<blocks_for((int64_t)n * (c / 4)), 256, 0, as_stream(stream)>>>(n, nsample, c, grad_output, grad_input1);
  else
    subtraction_bwd1_kernel<<<blocks_for((int64_t)n * c), 256, 0, as_stream(stream)>>>(n, nsample, c, grad_output, grad_input1);
  count_launch(1);
  int rc = check_launch("subtraction_bwd1_kernel");
  if (rc) return rc;
  return launch_clast_reduce<false>(n, c, 1, -1.f, grad_output, nullptr, plan, grad_input2, as_stream(stream), "subtraction_backward");
}

GEOT_API int geot_aggregation_forward(int n, int nsample, int c, int w_c, const float* input,
                                      const float* position, const float* weight, const int* idx,
                                      float* output, geot_stream_t stream) {
  GEOT_REQUIRE(n >= 0 && nsample >= 0 && c >= 0 && w_c >= 1, "aggregation_forward: bad sizes");
  if (n == 0 || c == 0) return GEOT_OK;
  GEOT_REQUIRE(output && (nsample == 0 || (input && position && weight && idx)), "aggregation_forward: NULL pointer");
  if (c % 4 == 0 && w_c % 4 == 0 && aligned16(input, position, weight, output))
    aggregation_fwd_vec_kernel<<<blocks_for((int64_t)n * (c / 4)), 256, 0, as_stream(stream)>>>(n, nsample, c, w_c, input, position, weight, idx, output);
  else
    aggregation_fwd_kernel<<<blocks_for((int64_t)n * c), 256, 0, as_stream(stream)>>>(n, nsample, c, w_c, input, position, weight, idx, output);
  count_launch(1);
  return check_launch("aggregation_fwd_kernel");
}

GEOT_API int geot_aggregation_backward(int n, int nsample, int c, int w_c, const float* input,
                                       const float* position, const float* weight, const int* idx,
                                       const float* grad_output, const void* plan, float* grad_input,
                                       float* grad_position, float* grad_weight, geot_stream_t stream) {
  GEOT_REQUIRE(n >= 0 && nsample >= 1 && c >= 0 && w_c >= 1, "aggregation_backward: bad sizes");
  if (n == 0 || c == 0) return GEOT_OK;
  GEOT_REQUIRE(input && position && weight && idx && grad_output && plan && grad_input && grad_position && grad_weight, "aggregation_backward: NULL pointer");
  cudaStream_t s = as_stream(stream);
  const PlanPtrs pp = plan_ptrs(plan, n);
  const bool vec = c % 4 == 0 && w_c % 4 == 0 && aligned16(input, position, weight, grad_output) && aligned16(grad_input, grad_position, grad_weight, nullptr);
  if (vec) aggregation_bwd_in_vec_kernel<<<blocks_for((int64_t)n * (c / 4)), 256, 0, s>>>(n, nsample, c, w_c, weight, grad_output, pp.row_ptr, pp.perm, grad_input);
  else aggregation_bwd_in_kernel<<<blocks_for((int64_t)n * c), 256, 0, s>>>(n, nsample, c, w_c, weight, grad_output, pp.row_ptr, pp.perm, grad_input);
  if (vec && pow2(c / 4) && pow2(w_c / 4) && c / 4 <= 32 && c % w_c == 0) {
    aggregation_bwd_pw_vec_kernel<<<blocks_for((int64_t)n * nsample * (c / 4)), 256, 0, s>>>(n, nsample, c, w_c, input, position, weight, idx, grad_output, grad_position, grad_weight);
  } else {
    aggregation_bwd_pos_kernel<<<blocks_for((int64_t)n * nsample * c), 256, 0, s>>>(n, nsample, c, w_c, weight, grad_output, grad_position);
    aggregation_bwd_w_kernel<<<blocks_for((int64_t)n * nsample * w_c), 256, 0, s>>>(n, nsample, c, w_c, input, position, idx, grad_output, grad_weight);
  }
  count_launch(3);
  return check_launch("aggregation_backward");
}
