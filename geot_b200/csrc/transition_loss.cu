// geot_b200/csrc/transition_loss.cu — manifold regularisers of the transition matrices and the class-level
// statistics, for sm_100a.
//
// Replaces utils/insT_loss.py:22-58 (feature_space_loss) and :72-110 (threeD_space_loss) after their kNN:
// the reference runs k rounds of index_select + torch.cat that grow a (BN,k,C,C) tensor twice (2 x 4.5 GB at
// C=33, k=32, BN=32000), then .repeat()s the (BN,C,C) tensor k times.  Here a warp owns a point, streams the
// k neighbour matrices straight from L2/HBM (4*C*C contiguous bytes each, coalesced) against its own matrix
// and never materialises anything larger than the (BN,k) coefficient map the backward needs.
// The backward is a deterministic gather: target a sums its own k terms and then the terms of the points
// that list a as a neighbour, in the fixed order of the scatter plan (group.cu) — no float atomics.
//
// Also: the per-class anchor pick (train.py:509-526), the Gaussian-prior / EMA update of the class-level
// transition matrix (train.py:528-545,556-557) and the running class-mean update (train.py:884-895), which
// the reference runs as C*C + C Python-level tiny launches.
#include "common.cuh"

namespace geot {

void count_launch(int n);

// ---------------------------------------------------------------------------------------------
// Processing order of the manifold kernels.  The (BN, C, C) tensor (139 MB at C=33, BN=32000) does not fit the 126 MB
// L2, and in dataset order (the clouds are shuffled) the k neighbour matrices of the ~1200 points in flight are
// scattered over all of it: round 1 measured 715 MB (forward) and 2087 MB (backward) of DRAM reads per launch against
// an ideal 139 / 278 MB.  Visiting the points of a cloud in Morton-cell order keeps a neighbourhood's 4.3 KB matrices
// L2-resident between the warps that share them.  The order only decides WHEN a point is processed: every point
// writes its own outputs and the final sums run in index order, so no result bit depends on it (the counting sort
// below is free to use integer atomics).  One CTA per cloud: bounding cube -> 16^3 cells -> histogram, scan, scatter.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned spread4(unsigned v) { return (v & 1u) | ((v & 2u) << 2) | ((v & 4u) << 4) | ((v & 8u) << 6); }
__global__ void __launch_bounds__(1024)
spatial_order_kernel(int N, int D, const float* __restrict__ points, int* __restrict__ order) {
  __shared__ int hist[4096];
  __shared__ float s_red[6][32];
  __shared__ int s_scan[32];
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const float* __restrict__ p = points + (int64_t)blockIdx.x * N * D;
  int* __restrict__ ord = order + (int64_t)blockIdx.x * N;
  const float kInf = __int_as_float(0x7f800000);
  const int d1 = D > 1 ? 1 : 0, d2 = D > 2 ? 2 : 0;
  float lo[3] = {kInf, kInf, kInf}, hi[3] = {-kInf, -kInf, -kInf};
  for (int i = t; i < N; i += 1024) {
    const float c[3] = {p[(int64_t)i * D], p[(int64_t)i * D + d1], p[(int64_t)i * D + d2]};
#pragma unroll
    for (int a = 0; a < 3; a++) { lo[a] = fminf(lo[a], c[a]); hi[a] = fmaxf(hi[a], c[a]); }
  }
#pragma unroll
  for (int a = 0; a < 3; a++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
      hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
    }
    if (lane == 0) { s_red[a][warp] = lo[a]; s_red[3 + a][warp] = hi[a]; }
  }
  for (int i = t; i < 4096; i += 1024) hist[i] = 0;
  __syncthreads();
  float ext = 0.f;
#pragma unroll
  for (int a = 0; a < 3; a++) {
    float l = s_red[a][lane], h = s_red[3 + a][lane];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      l = fminf(l, __shfl_xor_sync(0xffffffffu, l, o));
      h = fmaxf(h, __shfl_xor_sync(0xffffffffu, h, o));
    }
    lo[a] = l;
    ext = fmaxf(ext, h - l);
  }
  const float scale = (ext > 0.f && ext < kInf) ? 16.0f / ext : 0.f;
  auto cell = [&](int i) -> unsigned {  // NaN -> cell 0: scheduling only
    const int qx = min(max(__float2int_rd((p[(int64_t)i * D] - lo[0]) * scale), 0), 15);
    const int qy = min(max(__float2int_rd((p[(int64_t)i * D + d1] - lo[1]) * scale), 0), 15);
    const int qz = min(max(__float2int_rd((p[(int64_t)i * D + d2] - lo[2]) * scale), 0), 15);
    return spread4((unsigned)qx) | (spread4((unsigned)qy) << 1) | (spread4((unsigned)qz) << 2);
  };
  for (int i = t; i < N; i += 1024) atomicAdd(&hist[cell(i)], 1);
  __syncthreads();
  {
    int v[4], sum = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) { v[i] = hist[t * 4 + i]; sum += v[i]; }
    int inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += y;
    }
    if (lane == 31) s_scan[warp] = inc;
    __syncthreads();
    int wtot = s_scan[lane], winc = wtot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o) winc += y;
    }
    int run = __shfl_sync(0xffffffffu, winc - wtot, warp) + inc - sum;
#pragma unroll
    for (int i = 0; i < 4; i++) { hist[t * 4 + i] = run; run += v[i]; }
  }
  __syncthreads();
  for (int i = t; i < N; i += 1024) ord[atomicAdd(&hist[cell(i)], 1)] = i;
}

// ---------------------------------------------------------------------------------------------
// Manifold loss forward.  grid-stride over points (in `order` when given), one warp per point.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
manifold_fwd_kernel(int B, int N, int D, int C2, int k, const float* __restrict__ points,
                    const int64_t* __restrict__ labels, const float* __restrict__ insT,
                    const int* __restrict__ idx, const int* __restrict__ order, float inv_two_sigma2, int mode,
                    float* __restrict__ coef, float* __restrict__ per_point) {
  extern __shared__ float mf_smem[];  // 8 warps x C2 floats
  const int lane = threadIdx.x & 31;
  float* ti = mf_smem + (size_t)(threadIdx.x >> 5) * C2;
  const int64_t BN = (int64_t)B * N;
  const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t slot = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; slot < BN; slot += warps) {
    const int64_t base = (slot / N) * N;  // first point of this cloud
    const int64_t g = order ? base + __ldg(order + slot) : slot;
    const float* __restrict__ Ti = insT + g * C2;
    const int64_t lab_i = labels[g];
    // T_i goes to shared memory once: re-reading it per neighbour from global doubled the L2 traffic (L1 keeps only
    // ~25 % of it), and 35 registers per lane for it cost more occupancy than it saved
    __syncwarp();
    for (int f = lane; f < C2; f += 32) ti[f] = __ldg(Ti + f);
    __syncwarp();
    float num = 0.f, den = 0.f;  // lane-partial sums over this lane's neighbours
    for (int s0 = 0; s0 < k; s0 += 32) {
      const int s = s0 + lane;
      const int cnt = min(32, k - s0);
      const int64_t nb = s < k ? base + __ldg(idx + g * k + s) : g;
      // T distances: neighbour q of this chunk is processed by the whole warp, lane q keeps the result
      float td = 0.f;
      for (int q = 0; q < cnt; q++) {
        const int64_t o = __shfl_sync(0xffffffffu, nb, q);
        const float* __restrict__ Tj = insT + o * C2;
        float a = 0.f;
#pragma unroll 4
        for (int f = lane; f < C2; f += 32) {
          const float d = ti[f] - __ldg(Tj + f);
          a = __fmaf_rn(d, d, a);
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) a += __shfl_xor_sync(0xffffffffu, a, off);
        if (lane == q) td = a;
      }
      float w = 0.f;
      if (s < k) {
        float d2 = 0.f;
        for (int c = 0; c < D; c++) {
          const float d = __ldg(points + g * D + c) - __ldg(points + nb * D + c);
          d2 = __fmaf_rn(d, d, d2);
        }
        const float e = expf(-d2 * inv_two_sigma2);
        const bool same = labels[nb] == lab_i;
        w = (mode == GEOT_MANIFOLD_3D) ? (same ? e : 0.f) : (same ? e : -e);
      }
      num = __fmaf_rn(w, td, num);
      den += w;
      if (s < k) coef[g * k + s] = w;  // finalised below for the 3-D mode
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      num += __shfl_xor_sync(0xffffffffu, num, off);
      den += __shfl_xor_sync(0xffffffffu, den, off);
    }
    if (mode == GEOT_MANIFOLD_3D) {
      const float inv = 1.0f / (den + 0.001f);  // insT_loss.py:107
      for (int s = lane; s < k; s += 32) coef[g * k + s] *= inv;  // same lane wrote it: no hazard
      if (lane == 0) per_point[g] = num * inv;
    } else {
      const float inv = 1.0f / (float)k;  // torch.mean over (BN, k): the 1/k lives in coef, the 1/BN in the final sum
      for (int s = lane; s < k; s += 32) coef[g * k + s] *= inv;
      if (lane == 0) per_point[g] = num * inv;
    }
  }
}

// loss = scale * sum(per_point): one CTA, fixed order, double accumulation.
__global__ void __launch_bounds__(1024)
sum_scale_kernel(int64_t n, const float* __restrict__ v, double scale, float* __restrict__ out) {
  __shared__ double s_part[32];
  double acc = 0.0;
  for (int64_t i = threadIdx.x; i < n; i += 1024) acc += (double)v[i];
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    double a = s_part[threadIdx.x];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) a += __shfl_xor_sync(0xffffffffu, a, off);
    if (threadIdx.x == 0) out[0] = (float)(a * scale);
  }
}

// ---------------------------------------------------------------------------------------------
// Manifold loss backward w.r.t. insT.  One warp per target point a:
//   grad[a] = 2*g*scale * ( sum_{s<k} coef[a,s] (T_a - T_{nbr(a,s)})  +  sum_{(i,s): nbr(i,s)=a} coef[i,s] (T_a - T_i) )
// PARTS warps share a point, each owning a contiguous range of 32*NT matrix entries (NT register slots per lane):
// at C=33 one warp per point needs 184 registers (7 warps per SM, too few loads in flight to fill the L2 pipe).
// ---------------------------------------------------------------------------------------------
template <int NT, int PARTS>
__global__ void __launch_bounds__(256)
manifold_bwd_kernel(int B, int N, int C2, int k, const float* __restrict__ insT, const int* __restrict__ idx,
                    const float* __restrict__ coef, const int* __restrict__ row_ptr,
                    const int* __restrict__ perm, const int* __restrict__ order, const float* __restrict__ grad_loss,
                    float scale, float* __restrict__ grad_insT) {
  const int lane = threadIdx.x & 31;
  const int64_t BN = (int64_t)B * N;
  const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const float gs = 2.0f * scale * __ldg(grad_loss);
  const int64_t E = (int64_t)N * k;
  for (int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < BN * PARTS; w += warps) {
    const int64_t slot = w / PARTS;
    const int f0 = (int)(w - slot * PARTS) * (32 * NT);  // first entry of this warp's part
    const int64_t b = slot / N, base = b * N;
    const int64_t a = order ? (int64_t)__ldg(order + slot) : slot - base, g = base + a;
    const float* __restrict__ Ta = insT + g * C2;
    float ta[NT], acc[NT];
#pragma unroll
    for (int t = 0; t < NT; t++) {
      const int f = f0 + lane + 32 * t;
      ta[t] = f < C2 ? __ldg(Ta + f) : 0.f;
      acc[t] = 0.f;
    }
    auto add_term = [&](int64_t other, float c) {
      const float* __restrict__ To = insT + other * C2;
#pragma unroll
      for (int t = 0; t < NT; t++) {
        const int f = f0 + lane + 32 * t;
        if (f < C2) acc[t] = __fmaf_rn(c, ta[t] - __ldg(To + f), acc[t]);
      }
    };
    for (int s = 0; s < k; s++) add_term(base + __ldg(idx + g * k + s), __ldg(coef + g * k + s));
    const int* __restrict__ rp = row_ptr + b * (N + 1);
    const int* __restrict__ pm = perm + b * E;
    const int p0 = rp[a], p1 = rp[a + 1];
    for (int p = p0; p < p1; p++) {
      const int e = __ldg(pm + p);  // e = i*k + s  (local source point i, slot s)
      const int i = e / k;
      add_term(base + i, __ldg(coef + base * k + e));
    }
    float* __restrict__ out = grad_insT + g * C2;
#pragma unroll
    for (int t = 0; t < NT; t++) {
      const int f = f0 + lane + 32 * t;
      if (f < C2) out[f] = gs * acc[t];
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Class-level statistics.
// ---------------------------------------------------------------------------------------------
struct Pick {
  float v;
  int64_t i;
};
__device__ __forceinline__ Pick pick_better(Pick a, Pick b) {  // larger value, then smaller index (torch.argmax: first max)
  return (b.v > a.v || (b.v == a.v && b.i < a.i)) ? b : a;
}
__device__ __forceinline__ Pick pick_warp(Pick p) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    Pick o;
    o.v = __shfl_xor_sync(0xffffffffu, p.v, off);
    o.i = __shfl_xor_sync(0xffffffffu, p.i, off);
    p = pick_better(p, o);
  }
  return p;
}

constexpr int kPickParts = 64;

// grid (kPickParts, C): partial arg-max of pred_u[:, cc, :] over a slice of the flat (b, n) index.
__global__ void __launch_bounds__(256)
class_pick_partial_kernel(int B, int C, int N, const float* __restrict__ pred_u, Pick* __restrict__ parts) {
  __shared__ Pick s_p[8];
  const int cc = blockIdx.y;
  const int64_t BN = (int64_t)B * N;
  const int64_t per = (BN + kPickParts - 1) / kPickParts;
  const int64_t lo = (int64_t)blockIdx.x * per, hi = min(BN, lo + per);
  Pick best{-INFINITY, INT64_MAX};
  for (int64_t g = lo + threadIdx.x; g < hi; g += 256) {
    const int64_t b = g / N, n = g - b * N;
    best = pick_better(best, Pick{__ldg(pred_u + (b * C + cc) * N + n), g});
  }
  best = pick_warp(best);
  if ((threadIdx.x & 31) == 0) s_p[threadIdx.x >> 5] = best;
  __syncthreads();
  if (threadIdx.x < 32) {
    Pick p = threadIdx.x < 8 ? s_p[threadIdx.x] : Pick{-INFINITY, INT64_MAX};
    p = pick_warp(p);
    if (threadIdx.x == 0) parts[cc * kPickParts + blockIdx.x] = p;
  }
}

// grid C, 64 threads: final pick + row copy class_T[cc] = pred_u[b, :, n].
__global__ void __launch_bounds__(64)
class_pick_final_kernel(int B, int C, int N, const float* __restrict__ pred_u, const Pick* __restrict__ parts,
                        float* __restrict__ class_T, float* __restrict__ best_val, int64_t* __restrict__ best_idx) {
  __shared__ Pick s_best;
  const int cc = blockIdx.x;
  if (threadIdx.x < 32) {
    Pick p = pick_better(parts[cc * kPickParts + threadIdx.x], parts[cc * kPickParts + 32 + threadIdx.x]);
    p = pick_warp(p);
    if (threadIdx.x == 0) s_best = p;
  }
  __syncthreads();
  const Pick p = s_best;
  if (threadIdx.x == 0) { best_val[cc] = p.v; best_idx[cc] = p.i; }
  const int64_t b = p.i / N, n = p.i - b * N;
  for (int c = threadIdx.x; c < C; c += 64) class_T[cc * C + c] = pred_u[(b * C + c) * N + n];
}

// One CTA.  Every "X / torch.sum(X, 1)" of the reference lacks keepdim, i.e. element [r][c] is divided by the
// ROW-sum of row c; reproduced verbatim.
__global__ void __launch_bounds__(1024)
class_T_step_kernel(int C, const float* __restrict__ class_T, const float* __restrict__ ema_t,
                    const float* __restrict__ sigma, const int* __restrict__ label_proj, float geo, float decay,
                    float* __restrict__ ema_t_corr, float* __restrict__ ema_t_out) {
  __shared__ float m[64 * 64], m2[64 * 64], rs[64], rs2[64];
  const int C2 = C * C;
  auto rowsums = [&](const float* src, float* dst) {
    __syncthreads();
    for (int r = threadIdx.x; r < C; r += blockDim.x) {
      float s = 0.f;
      for (int c = 0; c < C; c++) s += src[r * C + c];
      dst[r] = s;
    }
    __syncthreads();
  };
  // prior_T (train.py:528-537)
  for (int e = threadIdx.x; e < C2; e += blockDim.x) {
    const int cc = e / C, ck = e - cc * C;
    float v = 0.f;
    if (cc >= 1 && ck >= 1) {
      const float s = sigma[cc];
      const int dx = label_proj[ck] - label_proj[cc];
      v = (1.0f / (s * sqrtf(6.2831855f))) * expf(-(float)(dx * dx) / (2.0f * (s * s)));
    }
    if (cc == 0 && ck == 0) v = 1.0f;
    m[e] = v;
  }
  rowsums(m, rs);
  // new_T = geo*class_T + (1-geo)*prior/rowsum ; new_T[0] = class_T[0]   (:538-540)
  for (int e = threadIdx.x; e < C2; e += blockDim.x) {
    const int cc = e / C, ck = e - cc * C;
    const float ct = class_T[e];
    m2[e] = cc == 0 ? ct : geo * ct + (1.0f - geo) * (m[e] / rs[ck]);
  }
  rowsums(m2, rs);
  // ema_t_corr = normalise(ema*decay + (new_T/rowsum)*(1-decay))           (:541-545)
  for (int e = threadIdx.x; e < C2; e += blockDim.x) {
    const int ck = e % C;
    m[e] = ema_t[e] * decay + (m2[e] / rs[ck]) * (1.0f - decay);
  }
  rowsums(m, rs);
  // ema_t = normalise(ema*decay + class_T*(1-decay))                        (:556-557)
  for (int e = threadIdx.x; e < C2; e += blockDim.x) m2[e] = ema_t[e] * decay + class_T[e] * (1.0f - decay);
  rowsums(m2, rs2);
  for (int e = threadIdx.x; e < C2; e += blockDim.x) {
    const int ck = e % C;
    ema_t_corr[e] = m[e] / rs[ck];
    ema_t_out[e] = m2[e] / rs2[ck];
  }
}

__global__ void __launch_bounds__(256)
label_hist_kernel(int64_t BN, int C, const int64_t* __restrict__ target, unsigned long long* __restrict__ hist) {
  __shared__ unsigned int s_h[64];
  if (threadIdx.x < 64) s_h[threadIdx.x] = 0;
  __syncthreads();
  for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < BN; g += (int64_t)gridDim.x * blockDim.x) {
    const int64_t t = target[g];
    if (t >= 0 && t < C) atomicAdd(&s_h[(int)t], 1u);  // integer: order-free
  }
  __syncthreads();
  if (threadIdx.x < C && s_h[threadIdx.x]) atomicAdd(hist + threadIdx.x, (unsigned long long)s_h[threadIdx.x]);
}

// train.py:887-895 verbatim: `cur_feats = logits[target]` gathers ROWS (points) by class id, so its mean is
// sum_c hist[c] * probs[c, :] / BN — the same vector for every class kk that is present.
__global__ void __launch_bounds__(64)
mean_feature_update_kernel(int64_t BN, int C, const float* __restrict__ probs,
                           const unsigned long long* __restrict__ hist, float* __restrict__ cm,
                           float* __restrict__ c_num) {
  __shared__ float mean[64];
  const int j = threadIdx.x;
  if (j < C) {
    double s = 0.0;
    for (int c = 0; c < C; c++) s += (double)hist[c] * (double)probs[(int64_t)c * C + j];
    mean[j] = (float)(s / (double)BN);
  }
  __syncthreads();
  for (int kk = 0; kk < C; kk++) {
    const float cur = (float)hist[kk];
    if (cur == 0.f) continue;
    const float cn = c_num[kk];
    if (j < C) cm[kk * C + j] = (cm[kk * C + j] * cn + mean[j] * cur) / (cn + cur);
    __syncthreads();
    if (j == 0) c_num[kk] = cn + cur;
    __syncthreads();
  }
}

static inline size_t align256(size_t b) { return (b + 255) & ~(size_t)255; }

}  // namespace geot

using namespace geot;

GEOT_API size_t geot_manifold_workspace_bytes(int64_t BN) { return BN > 0 ? align256((size_t)BN * sizeof(float)) : 0; }

GEOT_API int geot_spatial_order(int B, int N, int D, const float* points, int* order, geot_stream_t stream) {
  GEOT_REQUIRE(B >= 0 && N >= 1 && D >= 1, "spatial_order: bad sizes B=%d N=%d D=%d", B, N, D);
  if (B == 0) return GEOT_OK;
  GEOT_REQUIRE(points && order, "spatial_order: NULL pointer");
  spatial_order_kernel<<<B, 1024, 0, as_stream(stream)>>>(N, D, points, order);
  count_launch(1);
  return check_launch("spatial_order_kernel");
}

GEOT_API int geot_manifold_loss_forward(int B, int N, int D, int C, int k, const float* points,
                                        const int64_t* labels, const float* insT, const int* idx, const int* order,
                                        float sigma, int mode, float* coef, float* loss, void* workspace,
                                        size_t workspace_bytes, geot_stream_t stream) {
  GEOT_REQUIRE(B >= 1 && N >= 1 && D >= 1 && C >= 1 && C <= 64 && k >= 1, "manifold_loss_forward: bad sizes (1 <= C <= 64) B=%d N=%d D=%d C=%d k=%d", B, N, D, C, k);
  GEOT_REQUIRE(mode == GEOT_MANIFOLD_3D || mode == GEOT_MANIFOLD_FEATURE, "manifold_loss_forward: unknown mode %d", mode);
  GEOT_REQUIRE(points && labels && insT && idx && coef && loss, "manifold_loss_forward: NULL pointer");
  GEOT_REQUIRE(sigma > 0.f, "manifold_loss_forward: sigma must be positive");
  const int64_t BN = (int64_t)B * N;
  if (workspace == nullptr || workspace_bytes < geot_manifold_workspace_bytes(BN)) {
    set_error("manifold_loss_forward: workspace %zu < %zu bytes", workspace_bytes, geot_manifold_workspace_bytes(BN));
    return GEOT_ERR_WORKSPACE;
  }
  cudaStream_t s = as_stream(stream);
  float* per_point = reinterpret_cast<float*>(workspace);
  const int64_t want = (BN + 7) / 8;
  const int grid = (int)(want < (int64_t)kNumSMs * 8 ? want : (int64_t)kNumSMs * 8);
  const size_t smem = (size_t)8 * C * C * sizeof(float);
  if (smem > 48 * 1024) {
    static DynSmem ds;
    if (int rc = ds.need(manifold_fwd_kernel, smem, "manifold_loss_forward")) return rc;
  }
  manifold_fwd_kernel<<<grid, 256, smem, s>>>(B, N, D, C * C, k, points, labels, insT, idx, order,
                                              1.0f / (2.0f * (sigma * sigma)), mode, coef, per_point);
  count_launch(1);
  int rc = check_launch("manifold_fwd_kernel");
  if (rc) return rc;
  const double scale = 1.0 / (double)BN;
  sum_scale_kernel<<<1, 1024, 0, s>>>(BN, per_point, scale, loss);
  count_launch(1);
  return check_launch("sum_scale_kernel");
}

GEOT_API int geot_manifold_loss_backward(int B, int N, int C, int k, const float* insT, const int* idx,
                                         const float* coef, const void* plan, const int* order, const float* grad_loss,
                                         float* grad_insT, geot_stream_t stream) {
  GEOT_REQUIRE(B >= 1 && N >= 1 && C >= 1 && C <= 64 && k >= 1, "manifold_loss_backward: bad sizes");
  GEOT_REQUIRE(insT && idx && coef && plan && grad_loss && grad_insT, "manifold_loss_backward: NULL pointer");
  const int64_t BN = (int64_t)B * N;
  // plan layout of group.cu: row_ptr (B, N+1) | perm (B, N*k), row_ptr block padded to 256 bytes
  const int* row_ptr = reinterpret_cast<const int*>(plan);
  const size_t rp_bytes = (((size_t)B * (N + 1) * sizeof(int)) + 255) & ~(size_t)255;
  const int* perm = reinterpret_cast<const int*>(reinterpret_cast<const char*>(plan) + rp_bytes);
  const int64_t want = (BN + 7) / 8;
  const int grid = (int)(want < (int64_t)kNumSMs * 8 ? want : (int64_t)kNumSMs * 8);
  cudaStream_t s = as_stream(stream);
  // coef holds d(per-point loss)/d(Tdist) for either mode; the mean over points (1/BN) is applied here.
  const int C2 = C * C;
  const int nt = (C2 + 31) / 32;
  const float scale = 1.0f / (float)BN;
  auto grid_of = [&](int parts) {
    const int64_t want_w = (BN * parts + 7) / 8;
    return (int)(want_w < (int64_t)kNumSMs * 16 ? want_w : (int64_t)kNumSMs * 16);
  };
  if (nt <= 10) manifold_bwd_kernel<10, 1><<<grid_of(1), 256, 0, s>>>(B, N, C2, k, insT, idx, coef, row_ptr, perm, order, grad_loss, scale, grad_insT);
  else if (nt <= 36) manifold_bwd_kernel<18, 2><<<grid_of(2), 256, 0, s>>>(B, N, C2, k, insT, idx, coef, row_ptr, perm, order, grad_loss, scale, grad_insT);
  else if (nt <= 64) manifold_bwd_kernel<16, 4><<<grid_of(4), 256, 0, s>>>(B, N, C2, k, insT, idx, coef, row_ptr, perm, order, grad_loss, scale, grad_insT);
  else manifold_bwd_kernel<16, 8><<<grid_of(8), 256, 0, s>>>(B, N, C2, k, insT, idx, coef, row_ptr, perm, order, grad_loss, scale, grad_insT);
  count_launch(1);
  return check_launch("manifold_bwd_kernel");
}

GEOT_API size_t geot_class_pick_workspace_bytes(int C) {
  return C > 0 ? align256((size_t)C * kPickParts * sizeof(Pick)) : 0;
}

GEOT_API int geot_class_pick(int B, int C, int N, const float* pred_u, float* class_T, float* best_val,
                             int64_t* best_idx, void* workspace, size_t workspace_bytes, geot_stream_t stream) {
  GEOT_REQUIRE(B >= 1 && N >= 1 && C >= 1 && C <= 64, "class_pick: bad sizes B=%d C=%d N=%d", B, C, N);
  GEOT_REQUIRE(pred_u && class_T && best_val && best_idx, "class_pick: NULL pointer");
  if (workspace == nullptr || workspace_bytes < geot_class_pick_workspace_bytes(C)) {
    set_error("class_pick: workspace %zu < %zu bytes", workspace_bytes, geot_class_pick_workspace_bytes(C));
    return GEOT_ERR_WORKSPACE;
  }
  cudaStream_t s = as_stream(stream);
  Pick* parts = reinterpret_cast<Pick*>(workspace);
  class_pick_partial_kernel<<<dim3(kPickParts, C), 256, 0, s>>>(B, C, N, pred_u, parts);
  count_launch(1);
  int rc = check_launch("class_pick_partial_kernel");
  if (rc) return rc;
  class_pick_final_kernel<<<C, 64, 0, s>>>(B, C, N, pred_u, parts, class_T, best_val, best_idx);
  count_launch(1);
  return check_launch("class_pick_final_kernel");
}

GEOT_API int geot_class_t_step(int C, const float* class_T, const float* ema_t, const float* sigma,
                               const int* label_proj, float geo_lambda, float ema_decay, float* ema_t_corr,
                               float* ema_t_out, geot_stream_t stream) {
  GEOT_REQUIRE(C >= 1 && C <= 64, "class_T_step: C=%d outside [1,64]", C);
  GEOT_REQUIRE(class_T && ema_t && sigma && label_proj && ema_t_corr && ema_t_out, "class_T_step: NULL pointer");
  class_T_step_kernel<<<1, 1024, 0, as_stream(stream)>>>(C, class_T, ema_t, sigma, label_proj, geo_lambda, ema_decay, ema_t_corr, ema_t_out);
  count_launch(1);
  return check_launch("class_T_step_kernel");
}

GEOT_API int geot_mean_feature_update(int64_t BN, int C, const float* probs, const int64_t* target, float* cm,
                                      float* c_num, void* workspace, size_t workspace_bytes, geot_stream_t stream) {
  GEOT_REQUIRE(BN >= 1 && C >= 1 && C <= 64, "mean_feature_update: bad sizes");
  GEOT_REQUIRE(BN >= C, "mean_feature_update: the reference's probs[target] row gather needs BN >= C rows");
  GEOT_REQUIRE(probs && target && cm && c_num, "mean_feature_update: NULL pointer");
  if (workspace == nullptr || workspace_bytes < 64 * sizeof(unsigned long long)) {
    set_error("mean_feature_update: workspace %zu < 512 bytes", workspace_bytes);
    return GEOT_ERR_WORKSPACE;
  }
  cudaStream_t s = as_stream(stream);
  unsigned long long* hist = reinterpret_cast<unsigned long long*>(workspace);
  cudaError_t e = cudaMemsetAsync(hist, 0, 64 * sizeof(unsigned long long), s);
  if (e != cudaSuccess) { set_error("mean_feature_update: memset: %s", cudaGetErrorString(e)); return GEOT_ERR_CUDA; }
  const int64_t want = (BN + 255) / 256;
  label_hist_kernel<<<(int)(want < 1024 ? want : 1024), 256, 0, s>>>(BN, C, target, hist);
  count_launch(1);
  int rc = check_launch("label_hist_kernel");
  if (rc) return rc;
  mean_feature_update_kernel<<<1, 64, 0, s>>>(BN, C, probs, hist, cm, c_num);
  count_launch(1);
  return check_launch("mean_feature_update_kernel");
}
