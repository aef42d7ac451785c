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

// Fine order for clouds of up to 16384 points: 18-bit Morton keys (64^3 cells) sorted in shared memory by a bitonic
// network, key << 14 | index as one 32-bit word.  The tile kernels below want consecutive points of the order to be
// spatial neighbours: with the 16^3 cells above a tile of 32 points touches ~245 distinct matrices, with 64^3 cells
// ~153, in full 30-bit Morton order ~150 (numpy model of the cfg4 clouds).  A thread owns EPT consecutive keys: the
// steps of the network that stay inside them run in registers.  Same contract: a permutation, scheduling only.
__device__ __forceinline__ unsigned spread6(unsigned v) {  // 6 bits -> every third bit
  v &= 0x3fu;
  v = (v | (v << 8)) & 0x0000300fu;
  v = (v | (v << 4)) & 0x000030c3u;
  v = (v | (v << 2)) & 0x00009249u;
  return v;
}
// R consecutive steps (j, j/2, .. j >> (R-1)) of merge size k in one pass: a thread takes the 2^R keys that differ in
// those R index bits through registers - a third of the shared-memory round trips and barriers of one step per pass.
template <int R, int N2>
__device__ __forceinline__ void so_pass(unsigned* keys, int t, int k, int j) {
  const int jl = j >> (R - 1);  // the lowest step of this pass
  for (int gi = t; gi < (N2 >> R); gi += 1024) {
    const int base = ((gi & ~(jl - 1)) << R) | (gi & (jl - 1));
    const bool up = (base & k) == 0;
    unsigned w[1 << R];
#pragma unroll
    for (int c = 0; c < (1 << R); c++) w[c] = keys[base + c * jl];
#pragma unroll
    for (int st = 1 << (R - 1); st > 0; st >>= 1) {
#pragma unroll
      for (int c = 0; c < (1 << R); c++) {
        if ((c & st) == 0) {
          const unsigned x = w[c], y = w[c | st];
          if ((x > y) == up) { w[c] = y; w[c | st] = x; }
        }
      }
    }
#pragma unroll
    for (int c = 0; c < (1 << R); c++) keys[base + c * jl] = w[c];
  }
}
template <int EPT>
__global__ void __launch_bounds__(1024)
spatial_order_fine_kernel(int N, int D, const float* __restrict__ points, int* __restrict__ order) {
  constexpr int n2 = 1024 * EPT;
  extern __shared__ __align__(16) unsigned so_keys[];  // n2
  __shared__ float s_red[6][32];
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
  const float scale = (ext > 0.f && ext < kInf) ? 64.0f / ext : 0.f;
  unsigned v[EPT];  // keys 16t .. 16t + EPT - 1
#pragma unroll
  for (int e = 0; e < EPT; e++) {
    const int i = t * EPT + e;
    unsigned key = ~0u;  // padding sorts last
    if (i < N) {  // NaN -> cell 0: scheduling only
      const int qx = min(max(__float2int_rd((p[(int64_t)i * D] - lo[0]) * scale), 0), 63);
      const int qy = min(max(__float2int_rd((p[(int64_t)i * D + d1] - lo[1]) * scale), 0), 63);
      const int qz = min(max(__float2int_rd((p[(int64_t)i * D + d2] - lo[2]) * scale), 0), 63);
      key = ((spread6((unsigned)qx) | (spread6((unsigned)qy) << 1) | (spread6((unsigned)qz) << 2)) << 14) | (unsigned)i;
    }
    v[e] = key;
  }
  auto local_steps = [&](int k, int jmax) {  // steps j = jmax, jmax/2, .. 1 of merge size k on this thread's keys
#pragma unroll
    for (int j = EPT / 2; j > 0; j >>= 1) {
      if (j <= jmax) {
#pragma unroll
        for (int e = 0; e < EPT; e++) {
          if ((e & j) == 0) {
            const bool up = ((t * EPT + e) & k) == 0;
            const unsigned x = v[e], y = v[e | j];
            if ((x > y) == up) { v[e] = y; v[e | j] = x; }
          }
        }
      }
    }
  };
#pragma unroll
  for (int k = 2; k <= EPT; k <<= 1) local_steps(k, k >> 1);
#pragma unroll
  for (int e = 0; e < EPT; e++) so_keys[t * EPT + e] = v[e];
  __syncthreads();
  constexpr int RMAX = EPT >= 8 ? 3 : (EPT >= 4 ? 2 : 1);  // network steps per pass over shared memory
  for (int k = 2 * EPT; k <= n2; k <<= 1) {
    int j = k >> 1;
    while (j >= EPT) {
      int avail = 1;
      while ((j >> avail) >= EPT) avail++;
      if (RMAX >= 3 && avail >= 3) { so_pass<3, n2>(so_keys, t, k, j); j >>= 3; }
      else if (RMAX >= 2 && avail >= 2) { so_pass<2, n2>(so_keys, t, k, j); j >>= 2; }
      else { so_pass<1, n2>(so_keys, t, k, j); j >>= 1; }
      __syncthreads();
    }
    if (EPT > 1) {
#pragma unroll
      for (int e = 0; e < EPT; e++) v[e] = so_keys[t * EPT + e];
      local_steps(k, EPT / 2);
#pragma unroll
      for (int e = 0; e < EPT; e++) so_keys[t * EPT + e] = v[e];
      __syncthreads();
    }
  }
  for (int i = t; i < N; i += 1024) ord[i] = (int)(so_keys[i] & 0x3fffu);
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
// Mutual neighbours.  When a lists i and i lists a, the pair contributes c[a,s] (T_a - T_i) and c[i,s'] (T_a - T_i) to
// grad[a]: the two coefficients are added first and the difference is taken once (83 % of the pairs of the 32-NN graph
// of a cfg4 cloud are mutual: 64 -> ~37 terms per point).  Definition for k <= 32, shared by both backward kernels:
//   grad[a] = 2*g*scale * ( sum_s (c[a,s] + the coefficients of the in-terms from nbr(a,s), in plan order) (T_a - T_nbr(a,s))
//                           + the in-terms whose source is none of a's neighbours, in plan order )
// A repeated id among a's neighbours takes the matching in-terms at its first slot.
// oid = this lane's neighbour id (-1 for lanes >= k), oc = its coefficient; returns the merged coefficient.
// ---------------------------------------------------------------------------------------------
// The in-terms are handled 32 at a time: lane l loads term pb + l (source i_l, coefficient c_l, -1 / 0 past the end).
__device__ __forceinline__ void manifold_in_batch(int lane, int k, int pb, int p1, const int* __restrict__ pm,
                                                  const float* __restrict__ coef_cloud, int& i_l, float& c_l) {
  i_l = -1;
  c_l = 0.f;
  if (pb + lane < p1) {
    const int e = __ldg(pm + pb + lane);  // e = i*k + s  (local source point i, slot s)
    i_l = e / k;
    c_l = __ldg(coef_cloud + e);
  }
}
// bit q of the result: term q of the batch comes from one of the out ids; with ACC its coefficient is added to that
// lane's oc (terms in order, the first lane with the id)
template <bool ACC>
__device__ __forceinline__ unsigned manifold_match_batch(int lane, int oid, float& oc, int i_l, float c_l, int cnt) {
  unsigned matched = 0u;
  for (int q = 0; q < cnt; q++) {
    const int i_q = __shfl_sync(0xffffffffu, i_l, q);
    const float c_q = __shfl_sync(0xffffffffu, c_l, q);
    const unsigned m = __ballot_sync(0xffffffffu, oid == i_q);
    if (m != 0u) {
      matched |= 1u << q;
      if (ACC && lane == __ffs(m) - 1) oc = __fadd_rn(oc, c_q);
    }
  }
  return matched;
}
__device__ __forceinline__ float manifold_merge_coef(int lane, int k, int oid, float oc, int p0, int p1, const int* __restrict__ pm,
                                                     const float* __restrict__ coef_cloud) {
  for (int pb = p0; pb < p1; pb += 32) {
    int i_l;
    float c_l;
    manifold_in_batch(lane, k, pb, p1, pm, coef_cloud, i_l, c_l);
    manifold_match_batch<true>(lane, oid, oc, i_l, c_l, min(32, p1 - pb));
  }
  return oc;
}

// ---------------------------------------------------------------------------------------------
// Manifold loss backward w.r.t. insT.  One warp per target point a:
//   grad[a] = 2*g*scale * ( sum_{s<k} coef[a,s] (T_a - T_{nbr(a,s)})  +  sum_{(i,s): nbr(i,s)=a} coef[i,s] (T_a - T_i) )
// (for k <= 32 with the coefficients of mutual neighbours added first, see manifold_merge_coef)
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
    const int* __restrict__ rp = row_ptr + b * (N + 1);
    const int* __restrict__ pm = perm + b * E;
    const int p0 = rp[a], p1 = rp[a + 1];
    if (k <= 32) {  // mutual neighbours merged (see manifold_merge_coef)
      const int oid = lane < k ? __ldg(idx + g * k + lane) : -1;
      const float oc = manifold_merge_coef(lane, k, oid, lane < k ? __ldg(coef + g * k + lane) : 0.f, p0, p1, pm, coef + base * k);
      for (int s = 0; s < k; s++) add_term(base + __shfl_sync(0xffffffffu, oid, s), __shfl_sync(0xffffffffu, oc, s));
      for (int pb = p0; pb < p1; pb += 32) {
        int i_l;
        float c_l, unused = 0.f;
        manifold_in_batch(lane, k, pb, p1, pm, coef + base * k, i_l, c_l);
        const int cnt = min(32, p1 - pb);
        const unsigned matched = manifold_match_batch<false>(lane, oid, unused, i_l, c_l, cnt);
        for (int q = 0; q < cnt; q++) {
          const int i_q = __shfl_sync(0xffffffffu, i_l, q);
          const float c_q = __shfl_sync(0xffffffffu, c_l, q);
          if (((matched >> q) & 1u) == 0u) add_term(base + i_q, c_q);
        }
      }
    } else {
      for (int s = 0; s < k; s++) add_term(base + __ldg(idx + g * k + s), __ldg(coef + g * k + s));
      for (int p = p0; p < p1; p++) {
        const int e = __ldg(pm + p);  // e = i*k + s  (local source point i, slot s)
        const int i = e / k;
        add_term(base + i, __ldg(coef + base * k + e));
      }
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
// Tile form of the two manifold kernels (default for k <= 32).  The warp-per-point kernels above pull every neighbour
// matrix from L2 once per PAIR: 32000 points x 32 neighbours x 4.3 KB = 4.4 GB of L2 -> SM traffic in the forward and
// twice that in the backward, and sit on its latency.  Neighbourhoods of spatially adjacent points overlap: a tile of
// 32 consecutive points of the Morton order touches 150-250 distinct points, not 1024.  Here a CTA owns such a tile:
// the distinct points it touches are numbered through a shared-memory hash table, their matrices are staged in shared
// memory ONE SLICE of the C*C entries at a time (as wide as the number of rows allows: 256 / 128 / 64 / 32 entries),
// one bulk async copy per row and slice, and every pair is evaluated from shared memory by the warp that owns the
// point.  Per entry and per pair the arithmetic - and its order - is exactly that of the warp-per-point kernels: same
// bits, whatever the tiles or the slice width.  A row of C*C floats is not 16-byte aligned when C is odd: the copy
// starts at the aligned address below the slice and the row's handle carries the 0..3 floats of shift.
// ---------------------------------------------------------------------------------------------
constexpr int kMtThreads = 1024;                   // 32 warps, one point of the tile each
constexpr int kMtTile = 32;
constexpr int kMtHash = 4096;                      // slots of the id -> row table
constexpr int kMtRowFloats = 44 * 1024;            // 176 KB of staged matrix slices
constexpr int kMtRowsMax = kMtRowFloats / 36;      // 1251 rows at the narrowest slice (32 entries + 4 of shift room)
constexpr int kMtTerms = 64;                       // backward: merged terms of a target kept in shared memory
struct MtTable {
  int* keys;       // kMtHash, -1 = empty
  short* rowof;    // kMtHash
  int* rowid;      // kMtRowsMax: row -> local point id
  int* nrows;
  uint64_t* bar;   // completion of a slice's bulk copies
  int2* terms;     // backward: kMtTile x kMtTerms {row handle (or ~id when not staged), coefficient bits}
};
__device__ __forceinline__ unsigned mt_hash(int id) { return ((unsigned)id * 2654435761u) >> 20; }  // 12 bits
__device__ __forceinline__ void mt_insert(const MtTable& tb, int id) {
  unsigned h = mt_hash(id);
  for (;;) {
    const int cur = *reinterpret_cast<volatile int*>(tb.keys + h);
    if (cur == id) return;
    if (cur == -1) {
      if (*reinterpret_cast<volatile int*>(tb.nrows) >= kMtRowsMax) return;  // table closed: the id stays a global read
      const int old = atomicCAS(tb.keys + h, -1, id);
      if (old == -1) {
        const int r = atomicAdd(tb.nrows, 1);
        if (r < kMtRowsMax) { tb.rowof[h] = (short)r; tb.rowid[r] = id; } else tb.rowof[h] = (short)-1;
        return;
      }
      if (old == id) return;
    }
    h = (h + 1) & (kMtHash - 1);
  }
}
__device__ __forceinline__ int mt_row(const MtTable& tb, int id) {  // after the inserts (and a barrier); -1 = not staged
  unsigned h = mt_hash(id);
  for (;;) {
    const int cur = tb.keys[h];
    if (cur == id) return tb.rowof[h];
    if (cur == -1) return -1;
    h = (h + 1) & (kMtHash - 1);
  }
}
// offset (in floats) of a point's slice in the row buffer: row * (W + 4) + the shift of its matrix against 16 bytes
__device__ __forceinline__ int mt_handle(const MtTable& tb, int id, int64_t base, int C2, int Wp) {
  const int r = mt_row(tb, id);
  return r < 0 ? -1 : r * Wp + (int)(((base + id) * C2) & 3);
}
__device__ __forceinline__ MtTable mt_table(unsigned char* smem) {
  MtTable tb;
  tb.keys = reinterpret_cast<int*>(smem + (size_t)kMtRowFloats * 4);
  tb.rowid = tb.keys + kMtHash;
  tb.nrows = tb.rowid + kMtRowsMax + 1;
  tb.bar = reinterpret_cast<uint64_t*>(tb.nrows + 2);  // 8-byte aligned: kMtRowFloats*4 + (4096 + 1252 + 2) * 4
  tb.rowof = reinterpret_cast<short*>(tb.bar + 1);
  tb.terms = reinterpret_cast<int2*>(tb.rowof + kMtHash);
  return tb;
}
constexpr size_t kMtSmemBytes = (size_t)kMtRowFloats * 4 + (size_t)(kMtHash + kMtRowsMax + 1 + 2) * 4 + 8 + (size_t)kMtHash * 2 +
                                (size_t)kMtTile * kMtTerms * 8;
static_assert(((size_t)kMtRowFloats * 4 + (size_t)(kMtHash + kMtRowsMax + 1 + 2) * 4) % 8 == 0, "mbarrier alignment");

__device__ __forceinline__ void mt_mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mt_mbar_wait(uint64_t* bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "MTWAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@!p bra MTWAIT_%=;\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}

// Stage slice [s0, s0 + 32*NJ) of the first `nr` rows: thread t copies rows t, t + 1024, ... with one bulk copy each and
// arrives on the barrier with their byte count (threads without a row arrive empty); everyone then waits for the phase.
// Entries at or beyond C2 are zeroed afterwards (a zero difference adds nothing; the copy may have brought the next
// matrix's first entries there).  total = number of floats in insT (the copy of the very last matrix is clipped to it).
template <int NJ>
__device__ __forceinline__ void mt_fill(float* rows, const MtTable& tb, int nr, const float* __restrict__ insT, int64_t base, int C2,
                                        int s0, int64_t total, unsigned& parity) {
  constexpr int W = 32 * NJ, Wp = W + 4;
  const int t = threadIdx.x;
  unsigned bytes = 0;
  if (t < nr) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // earlier generic reads of the rows
  for (int r = t; r < nr; r += kMtThreads) {
    const int64_t e0 = (base + tb.rowid[r]) * C2 + s0, ea = e0 & ~(int64_t)3;
    const int want = min(Wp, (int)(((e0 - ea) + min(W, C2 - s0) + 3) & ~3));  // floats, a multiple of 4
    const int have = (int)min((int64_t)want, (total - ea) & ~(int64_t)3);
    float* dst = rows + r * Wp;
    if (have > 0) {
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(smem_u32(dst)), "l"(insT + ea), "r"((unsigned)have * 4u), "r"(smem_u32(tb.bar)) : "memory");
      bytes += (unsigned)have * 4u;
    }
    for (int i = have; i < want && ea + i < total; i++) dst[i] = __ldg(insT + ea + i);  // the tensor's last 1..3 floats
  }
  if (bytes) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(tb.bar)), "r"(bytes) : "memory");
  else asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(tb.bar)) : "memory");
  mt_mbar_wait(tb.bar, parity);
  parity ^= 1u;
  if (s0 + W > C2) {  // last slice: zero from the end of the matrix to the end of the row
    const int lane = t & 31, warp = t >> 5;
    for (int r = warp; r < nr; r += kMtThreads / 32) {
      const int first = (int)(((base + tb.rowid[r]) * C2) & 3) + (C2 - s0);
      for (int i = first + lane; i < Wp; i += 32) rows[r * Wp + i] = 0.f;
    }
  }
}

// the sum of part[q] over the 32 lanes lands in lane q: the xor butterfly (16, 8, 4, 2, 1) of the warp-per-point
// kernel, each lane keeping only the half of the values it will still need - same additions, 31 shuffles instead of 160
__device__ __forceinline__ float mt_reduce32(float (&v)[32], int lane) {
#pragma unroll
  for (int i = 0; i < 16; i++) {
    const bool up = lane & 16;
    const float send = up ? v[i] : v[i + 16], keep = up ? v[i + 16] : v[i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
  }
#pragma unroll
  for (int i = 0; i < 8; i++) {
    const bool up = lane & 8;
    const float send = up ? v[i] : v[i + 8], keep = up ? v[i + 8] : v[i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
  }
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const bool up = lane & 4;
    const float send = up ? v[i] : v[i + 4], keep = up ? v[i + 4] : v[i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
  }
#pragma unroll
  for (int i = 0; i < 2; i++) {
    const bool up = lane & 2;
    const float send = up ? v[i] : v[i + 2], keep = up ? v[i + 2] : v[i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
  }
  {
    const bool up = lane & 1;
    const float send = up ? v[0] : v[1], keep = up ? v[1] : v[0];
    v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 1);
  }
  return v[0];
}

// one slice of the forward: part[q] += sum over this lane's entries of (T_i - T_nbr(q))^2.  Every row is staged: a
// forward tile touches at most 32 + 32*32 = 1056 <= kMtRowsMax distinct points.
template <int NJ>
__device__ __forceinline__ void mt_fwd_slice(const float* rows, int h_i, int h_q, float (&part)[32]) {
  const int lane = threadIdx.x & 31;
  float ti[NJ];
#pragma unroll
  for (int j = 0; j < NJ; j++) ti[j] = rows[h_i + lane + 32 * j];
#pragma unroll
  for (int q = 0; q < 32; q++) {
    const float* __restrict__ tj = rows + __shfl_sync(0xffffffffu, h_q, q) + lane;
    float a = part[q];
#pragma unroll
    for (int j = 0; j < NJ; j++) {
      const float d = ti[j] - tj[32 * j];
      a = __fmaf_rn(d, d, a);
    }
    part[q] = a;  // lanes q >= k point at the tile's own row: finite, never read back
  }
}

template <int NJ>
__device__ __forceinline__ void mt_fwd_slices(float* rows, const MtTable& tb, int nr, const float* __restrict__ insT, int64_t base, int C2,
                                              int64_t total, unsigned& parity, int a, int nbl, bool live, float (&part)[32]) {
  constexpr int Wp = 32 * NJ + 4;
  const int h_i = live ? mt_handle(tb, a, base, C2, Wp) : 0, h_q = live ? mt_handle(tb, nbl, base, C2, Wp) : 0;
  for (int s0 = 0; s0 < C2; s0 += 32 * NJ) {
    __syncthreads();  // the previous slice is no longer read
    mt_fill<NJ>(rows, tb, nr, insT, base, C2, s0, total, parity);
    __syncthreads();
    if (live) mt_fwd_slice<NJ>(rows, h_i, h_q, part);
  }
}

// grid: persistent CTAs over the tiles (B * ceil(N / 32)), kMtThreads threads, kMtSmemBytes dynamic shared memory
__global__ void __launch_bounds__(kMtThreads, 1)
manifold_fwd_tile_kernel(int B, int N, int D, int C2, int k, const float* __restrict__ points,
                         const int64_t* __restrict__ labels, const float* __restrict__ insT,
                         const int* __restrict__ idx, const int* __restrict__ order, float inv_two_sigma2, int mode,
                         float* __restrict__ coef, float* __restrict__ per_point) {
  extern __shared__ __align__(16) unsigned char mt_smem[];
  float* rows = reinterpret_cast<float*>(mt_smem);
  const MtTable tb = mt_table(mt_smem);
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int tpc = (N + kMtTile - 1) / kMtTile;
  const int64_t ntiles = (int64_t)B * tpc, total = (int64_t)B * N * C2;
  if (t == 0) mt_mbar_init(tb.bar, kMtThreads);
  unsigned parity = 0;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t b = tile / tpc, base = b * N;
    const int slot0 = (int)(tile - b * tpc) * kMtTile;
    const int npts = min(kMtTile, N - slot0);
    __syncthreads();  // the previous tile's table and rows are no longer read
    for (int i = t; i < kMtHash; i += kMtThreads) tb.keys[i] = -1;
    if (t == 0) *tb.nrows = 0;
    __syncthreads();
    // the tile's own points first (their rows are read by every pair), then the neighbours
    if (t < npts) mt_insert(tb, order ? __ldg(order + base + slot0 + t) : slot0 + t);
    __syncthreads();
    for (int e = t; e < npts * k; e += kMtThreads) {
      const int p = e / k, s = e - p * k;
      const int a = order ? __ldg(order + base + slot0 + p) : slot0 + p;
      mt_insert(tb, __ldg(idx + (base + a) * k + s));
    }
    __syncthreads();
    const int nr = *tb.nrows;  // <= 1056
    const bool live = warp < npts;
    const int a = live ? (order ? __ldg(order + base + slot0 + warp) : slot0 + warp) : 0;
    const int64_t g = base + a;
    const int nbl = (live && lane < k) ? __ldg(idx + g * k + lane) : a;
    float part[32];
#pragma unroll
    for (int q = 0; q < 32; q++) part[q] = 0.f;
    if (nr * 260 <= kMtRowFloats) mt_fwd_slices<8>(rows, tb, nr, insT, base, C2, total, parity, a, nbl, live, part);
    else if (nr * 132 <= kMtRowFloats) mt_fwd_slices<4>(rows, tb, nr, insT, base, C2, total, parity, a, nbl, live, part);
    else if (nr * 68 <= kMtRowFloats) mt_fwd_slices<2>(rows, tb, nr, insT, base, C2, total, parity, a, nbl, live, part);
    else mt_fwd_slices<1>(rows, tb, nr, insT, base, C2, total, parity, a, nbl, live, part);
    if (live) {  // warp-uniform
      const float td = mt_reduce32(part, lane);  // lane q: squared distance to neighbour q
      const int64_t nb = base + nbl;
      float w = 0.f;
      if (lane < k) {
        float d2 = 0.f;
        for (int c = 0; c < D; c++) {
          const float d = __ldg(points + g * D + c) - __ldg(points + nb * D + c);
          d2 = __fmaf_rn(d, d, d2);
        }
        const float e = expf(-d2 * inv_two_sigma2);
        const bool same = labels[nb] == labels[g];
        w = (mode == GEOT_MANIFOLD_3D) ? (same ? e : 0.f) : (same ? e : -e);
      }
      float num = lane < k ? __fmaf_rn(w, td, 0.f) : 0.f, den = 0.f + w;
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) {
        num += __shfl_xor_sync(0xffffffffu, num, off);
        den += __shfl_xor_sync(0xffffffffu, den, off);
      }
      const float inv = (mode == GEOT_MANIFOLD_3D) ? 1.0f / (den + 0.001f) : 1.0f / (float)k;
      if (lane < k) coef[g * k + lane] = w * inv;
      if (lane == 0) per_point[g] = num * inv;
    }
  }
}

// The merged term list of this warp's target (manifold_merge_coef) goes to shared memory once per tile: the k own
// neighbours with their merged coefficients, then the in-terms that matched none of them (plan order).  Returns the
// number of terms; only the first kMtTerms are in the table, a longer list (a point listed by many that are not its
// neighbours) is finished from global memory in every slice.  Not staged (OVF tiles): the entry holds ~id.
template <int NJ>
__device__ __forceinline__ int mt_bwd_terms(const MtTable& tb, int a, int64_t base, int oid, int C2, int k, const float* __restrict__ coef,
                                            int p0, int p1, const int* __restrict__ pm) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int Wp = 32 * NJ + 4;
  int2* tt = tb.terms + warp * kMtTerms;
  const float oc = manifold_merge_coef(lane, k, oid, lane < k ? __ldg(coef + (base + a) * k + lane) : 0.f, p0, p1, pm, coef + base * k);
  if (lane < k) {
    const int h = mt_handle(tb, oid, base, C2, Wp);
    tt[lane] = make_int2(h >= 0 ? h : ~oid, __float_as_int(oc));
  }
  int n = k;
  for (int pb = p0; pb < p1; pb += 32) {
    int i_l;
    float c_l, unused = 0.f;
    manifold_in_batch(lane, k, pb, p1, pm, coef + base * k, i_l, c_l);
    const int cnt = min(32, p1 - pb);
    const unsigned unmatched = ~manifold_match_batch<false>(lane, oid, unused, i_l, c_l, cnt) & (cnt == 32 ? 0xffffffffu : ((1u << cnt) - 1u));
    if ((unmatched >> lane) & 1u) {
      const int pos = n + __popc(unmatched & ((1u << lane) - 1u));
      if (pos < kMtTerms) {
        const int h = mt_handle(tb, i_l, base, C2, Wp);
        tt[pos] = make_int2(h >= 0 ? h : ~i_l, __float_as_int(c_l));
      }
    }
    n += __popc(unmatched);
  }
  __syncwarp();
  return n;
}

// one slice of the backward for one target: acc[j] += c * (T_a - T_other) over the target's term list.  OVF: the tile
// touches more distinct points than the table holds (a hub listed by thousands): rows that are not staged come from global.
template <int NJ, bool OVF>
__device__ __forceinline__ void mt_bwd_slice(const float* rows, const MtTable& tb, int a, int64_t base, int oid, int nterms,
                                             const float* __restrict__ insT, const float* __restrict__ coef, int p0, int p1,
                                             const int* __restrict__ pm, int C2, int s0, int k, float gs, float* __restrict__ grad_insT) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int Wp = 32 * NJ + 4;
  const int64_t g = base + a;
  const int h_a = mt_handle(tb, a, base, C2, Wp);  // the tile's own points are inserted first: always staged
  float ta[NJ], acc[NJ];
#pragma unroll
  for (int j = 0; j < NJ; j++) {
    ta[j] = rows[h_a + lane + 32 * j];
    acc[j] = 0.f;
  }
  auto add_global = [&](int other, float c) {
    const float* __restrict__ to = insT + (base + other) * C2 + s0 + lane;
#pragma unroll
    for (int j = 0; j < NJ; j++)
      if (s0 + lane + 32 * j < C2) acc[j] = __fmaf_rn(c, ta[j] - __ldg(to + 32 * j), acc[j]);
  };
  const int2* __restrict__ tt = tb.terms + warp * kMtTerms;
  const int nt = min(nterms, kMtTerms);
#pragma unroll 4
  for (int q = 0; q < nt; q++) {
    const int2 tc = tt[q];
    const float c = __int_as_float(tc.y);
    if (!OVF || tc.x >= 0) {
      const float* __restrict__ to = rows + tc.x + lane;
#pragma unroll
      for (int j = 0; j < NJ; j++) acc[j] = __fmaf_rn(c, ta[j] - to[32 * j], acc[j]);
    } else {
      add_global(~tc.x, c);
    }
  }
  if (nterms > kMtTerms) {  // the tail of a very long list: the same walk as mt_bwd_terms, from global memory
    int n = k;
    for (int pb = p0; pb < p1; pb += 32) {
      int i_l;
      float c_l, unused = 0.f;
      manifold_in_batch(lane, k, pb, p1, pm, coef + base * k, i_l, c_l);
      const int cnt = min(32, p1 - pb);
      const unsigned matched = manifold_match_batch<false>(lane, oid, unused, i_l, c_l, cnt);
      for (int q = 0; q < cnt; q++) {
        const int i_q = __shfl_sync(0xffffffffu, i_l, q);
        const float c_q = __shfl_sync(0xffffffffu, c_l, q);
        if ((matched >> q) & 1u) continue;
        if (n >= kMtTerms) add_global(i_q, c_q);
        n++;
      }
    }
  }
  float* __restrict__ out = grad_insT + g * C2 + s0;
#pragma unroll
  for (int j = 0; j < NJ; j++)
    if (s0 + lane + 32 * j < C2) out[lane + 32 * j] = gs * acc[j];
}

template <int NJ, bool OVF>
__device__ __forceinline__ void mt_bwd_slices(float* rows, const MtTable& tb, int nr, bool live, int a, int64_t base, int64_t total,
                                              unsigned& parity, const float* __restrict__ insT, const int* __restrict__ idx,
                                              const float* __restrict__ coef, const int* __restrict__ rp, const int* __restrict__ pm,
                                              int C2, int k, float gs, float* __restrict__ grad_insT) {
  const int lane = threadIdx.x & 31;
  int oid = -1, nterms = 0, p0 = 0, p1 = 0;
  if (live) {
    oid = lane < k ? __ldg(idx + (base + a) * k + lane) : -1;
    p0 = __ldg(rp + a); p1 = __ldg(rp + a + 1);
    nterms = mt_bwd_terms<NJ>(tb, a, base, oid, C2, k, coef, p0, p1, pm);
  }
  for (int s0 = 0; s0 < C2; s0 += 32 * NJ) {
    __syncthreads();  // the previous slice is no longer read
    mt_fill<NJ>(rows, tb, nr, insT, base, C2, s0, total, parity);
    __syncthreads();
    if (live) mt_bwd_slice<NJ, OVF>(rows, tb, a, base, oid, nterms, insT, coef, p0, p1, pm, C2, s0, k, gs, grad_insT);
  }
}

__global__ void __launch_bounds__(kMtThreads, 1)
manifold_bwd_tile_kernel(int B, int N, int C2, int k, const float* __restrict__ insT, const int* __restrict__ idx,
                         const float* __restrict__ coef, const int* __restrict__ row_ptr, const int* __restrict__ perm,
                         const int* __restrict__ order, const float* __restrict__ grad_loss, float scale,
                         float* __restrict__ grad_insT) {
  extern __shared__ __align__(16) unsigned char mt_smem[];
  float* rows = reinterpret_cast<float*>(mt_smem);
  const MtTable tb = mt_table(mt_smem);
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int tpc = (N + kMtTile - 1) / kMtTile;
  const int64_t ntiles = (int64_t)B * tpc, total = (int64_t)B * N * C2;
  const float gs = 2.0f * scale * __ldg(grad_loss);
  const int64_t E = (int64_t)N * k;
  if (t == 0) mt_mbar_init(tb.bar, kMtThreads);
  unsigned parity = 0;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t b = tile / tpc, base = b * N;
    const int slot0 = (int)(tile - b * tpc) * kMtTile;
    const int npts = min(kMtTile, N - slot0);
    const int* __restrict__ rp = row_ptr + b * (N + 1);
    const int* __restrict__ pm = perm + b * E;
    __syncthreads();  // the previous tile's table and rows are no longer read
    for (int i = t; i < kMtHash; i += kMtThreads) tb.keys[i] = -1;
    if (t == 0) *tb.nrows = 0;
    __syncthreads();
    if (t < npts) mt_insert(tb, order ? __ldg(order + base + slot0 + t) : slot0 + t);
    __syncthreads();
    for (int e = t; e < npts * k; e += kMtThreads) {
      const int p = e / k, s = e - p * k;
      const int a = order ? __ldg(order + base + slot0 + p) : slot0 + p;
      mt_insert(tb, __ldg(idx + (base + a) * k + s));
    }
    const bool live = warp < npts;
    const int a = live ? (order ? __ldg(order + base + slot0 + warp) : slot0 + warp) : 0;
    if (live) {  // the points that list this warp's target
      const int p0 = __ldg(rp + a), p1 = __ldg(rp + a + 1);
      for (int pp = p0 + lane; pp < p1; pp += 32) mt_insert(tb, __ldg(pm + pp) / k);
    }
    __syncthreads();
    const int nins = *tb.nrows;
    const int nr = min(nins, kMtRowsMax);
    if (nins >= kMtRowsMax) mt_bwd_slices<1, true>(rows, tb, nr, live, a, base, total, parity, insT, idx, coef, rp, pm, C2, k, gs, grad_insT);
    else if (nr * 260 <= kMtRowFloats) mt_bwd_slices<8, false>(rows, tb, nr, live, a, base, total, parity, insT, idx, coef, rp, pm, C2, k, gs, grad_insT);
    else if (nr * 132 <= kMtRowFloats) mt_bwd_slices<4, false>(rows, tb, nr, live, a, base, total, parity, insT, idx, coef, rp, pm, C2, k, gs, grad_insT);
    else if (nr * 68 <= kMtRowFloats) mt_bwd_slices<2, false>(rows, tb, nr, live, a, base, total, parity, insT, idx, coef, rp, pm, C2, k, gs, grad_insT);
    else mt_bwd_slices<1, false>(rows, tb, nr, live, a, base, total, parity, insT, idx, coef, rp, pm, C2, k, gs, grad_insT);
  }
}

bool manifold_tiles_disabled();  // geot_group_points_config(6) (group.cu): A/B against the warp-per-point kernels
bool manifold_tiles_forced();    // geot_group_points_config(7): the tile kernels whatever C (tests)

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
  if (N <= 16384) {
    cudaStream_t s = as_stream(stream);
    static DynSmem ds16;
    if (N > 8192) {
      if (int rc = ds16.need(spatial_order_fine_kernel<16>, 16384 * 4, "spatial_order")) return rc;
      spatial_order_fine_kernel<16><<<B, 1024, 16384 * 4, s>>>(N, D, points, order);
    } else if (N > 4096) spatial_order_fine_kernel<8><<<B, 1024, 8192 * 4, s>>>(N, D, points, order);
    else if (N > 2048) spatial_order_fine_kernel<4><<<B, 1024, 4096 * 4, s>>>(N, D, points, order);
    else if (N > 1024) spatial_order_fine_kernel<2><<<B, 1024, 2048 * 4, s>>>(N, D, points, order);
    else spatial_order_fine_kernel<1><<<B, 1024, 1024 * 4, s>>>(N, D, points, order);
    count_launch(1);
    return check_launch("spatial_order_fine_kernel");
  }
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
  int rc = GEOT_OK;
  // same bits from either pair of kernels; the tiles pay a fixed ~0.1 ms of table building per launch (measured at
  // 2 x 16000, k = 32: C = 33 / 17 / 5 forward 0.33 / 0.15 / 0.104 ms against 0.59 / 0.21 / 0.097 for the warp kernel)
  if (k <= 32 && (((uintptr_t)insT) & 15) == 0 && !manifold_tiles_disabled() && (C * C >= 64 || manifold_tiles_forced())) {
    static DynSmem dst;
    if (int rcs = dst.need(manifold_fwd_tile_kernel, kMtSmemBytes, "manifold_loss_forward")) return rcs;
    const int64_t ntiles = (int64_t)B * ((N + kMtTile - 1) / kMtTile);
    const int grid = (int)(ntiles < num_sms() ? ntiles : num_sms());
    manifold_fwd_tile_kernel<<<grid, kMtThreads, kMtSmemBytes, s>>>(B, N, D, C * C, k, points, labels, insT, idx, order,
                                                                     1.0f / (2.0f * (sigma * sigma)), mode, coef, per_point);
    count_launch(1);
    rc = check_launch("manifold_fwd_tile_kernel");
  } else {
    const int64_t want = (BN + 7) / 8;
    const int grid = (int)(want < (int64_t)num_sms() * 8 ? want : (int64_t)num_sms() * 8);
    const size_t smem = (size_t)8 * C * C * sizeof(float);
    if (smem > 48 * 1024) {
      static DynSmem ds;
      if (int rcs = ds.need(manifold_fwd_kernel, smem, "manifold_loss_forward")) return rcs;
    }
    manifold_fwd_kernel<<<grid, 256, smem, s>>>(B, N, D, C * C, k, points, labels, insT, idx, order,
                                                1.0f / (2.0f * (sigma * sigma)), mode, coef, per_point);
    count_launch(1);
    rc = check_launch("manifold_fwd_kernel");
  }
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
  const int grid = (int)(want < (int64_t)num_sms() * 8 ? want : (int64_t)num_sms() * 8);
  cudaStream_t s = as_stream(stream);
  // coef holds d(per-point loss)/d(Tdist) for either mode; the mean over points (1/BN) is applied here.
  const int C2 = C * C;
  const int nt = (C2 + 31) / 32;
  const float scale = 1.0f / (float)BN;
  // backward: C = 33 / 17 / 5 tiles 0.46 / 0.23 / 0.16 ms, warp kernel 0.65 / 0.17 / 0.14 ms
  if (k <= 32 && (((uintptr_t)insT) & 15) == 0 && !manifold_tiles_disabled() && (C2 >= 512 || manifold_tiles_forced())) {
    static DynSmem dst;
    if (int rcs = dst.need(manifold_bwd_tile_kernel, kMtSmemBytes, "manifold_loss_backward")) return rcs;
    const int64_t ntiles = (int64_t)B * ((N + kMtTile - 1) / kMtTile);
    const int tgrid = (int)(ntiles < num_sms() ? ntiles : num_sms());
    manifold_bwd_tile_kernel<<<tgrid, kMtThreads, kMtSmemBytes, s>>>(B, N, C2, k, insT, idx, coef, row_ptr, perm, order, grad_loss,
                                                                      scale, grad_insT);
    count_launch(1);
    return check_launch("manifold_bwd_tile_kernel");
  }
  auto grid_of = [&](int parts) {
    const int64_t want_w = (BN * parts + 7) / 8;
    return (int)(want_w < (int64_t)num_sms() * 16 ? want_w : (int64_t)num_sms() * 16);
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
