// geot_b200/csrc/transition_apply.cu — per-point application of the blended transition matrix, sm_100a.
//
// Replaces examples/segmentation/train.py:549-552 and its autograd:
//     newT = F.normalize(lam * ema_t_corr + (1 - lam) * insT, p=1, dim=2)          (BN,C,C) temp #1, #2
//     out  = bmm(logits.permute(0,2,1).view(-1,C).unsqueeze(1), newT).squeeze(1)    batched 1xC . CxC
//     out  = out.view(B,N,C).permute(0,2,1).contiguous()
// The reference streams the (BN,C,C) tensor through HBM four times forward (blend, norm, divide, bmm) and
// keeps two full-size temporaries for autograd.  Here one warp owns a point: T_n is read ONCE (coalesced,
// 4*C*C contiguous bytes), blended into a per-warp shared tile, row L1 norms and the 1xC . CxC product come
// from that tile, and the (B,C,N) <-> point-major permutes are folded into a 32-point CTA tile so that
// logits / grad_out loads and out / grad_logits stores are 128-byte rows.  HBM-bound: 4*C*C bytes per point
// forward, 8*C*C backward (read insT, write grad_insT); nothing else of that size is materialised.
#include "common.cuh"

namespace geot {

void count_launch(int n);

constexpr int kApplyWarps = 8;
constexpr int kApplyThreads = kApplyWarps * 32;
constexpr int kApplyTile = 32;  // points per CTA tile (4 per warp)

struct ApplySmem {
  // floats: E[C*C] | lt[C][33] | gt[C][33] | ot[C][33] | per warp { sv[nbuf][C*Cp] | ca[64] | cb[64] }
  int C, Cp, nbuf;  // nbuf = 2: the next point's matrix is prefetched while this one is used (needs 2 tiles per warp)
  __host__ __device__ ApplySmem(int c, int nb) : C(c), Cp(c | 1), nbuf(nb) {}
  __host__ __device__ size_t e_off() const { return 0; }
  __host__ __device__ size_t lt_off() const { return ((size_t)C * C + 3) & ~(size_t)3; }
  __host__ __device__ size_t gt_off() const { return lt_off() + (size_t)C * 33; }
  __host__ __device__ size_t ot_off() const { return gt_off() + (size_t)C * 33; }
  __host__ __device__ size_t warp_off(int w) const { return ot_off() + (size_t)C * 33 + (size_t)w * ((size_t)nbuf * C * Cp + 128); }
  __host__ __device__ size_t bytes() const { return warp_off(kApplyWarps) * sizeof(float); }
};

// BWD == false: computes out.  BWD == true: computes grad_logits and/or grad_insT from grad_out.
template <bool BWD>
__global__ void __launch_bounds__(kApplyThreads)
transition_apply_kernel(int B, int C, int N, int nbuf, const float* __restrict__ logits, const float* __restrict__ ema_t,
                        const float* __restrict__ insT, float lam, const float* __restrict__ grad_out,
                        float* __restrict__ out_or_gl, float* __restrict__ grad_insT) {
  extern __shared__ __align__(16) float apply_smem[];
  const ApplySmem L(C, nbuf);
  const int Cp = L.Cp, C2 = C * C;
  float* E_s = apply_smem + L.e_off();    // lam * ema_t
  float* lt = apply_smem + L.lt_off();    // logits tile      [kk][pt]
  float* gt = apply_smem + L.gt_off();    // grad_out tile    [j][pt]            (BWD)
  float* ot = apply_smem + L.ot_off();    // out / grad_logits tile [c][pt]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* svbuf = apply_smem + L.warp_off(warp);  // two T_n buffers: the next point's matrix lands while this one is used
  float* ca = svbuf + (size_t)nbuf * C * Cp;
  float* cb = ca + 64;
  // T_n (4*C*C contiguous bytes, only 4-byte aligned for odd C) -> padded [kk][Cp] layout, asynchronously
  auto prefetch = [&](int64_t g, float* buf) {
    const float* __restrict__ T = insT + g * C2;
    int kk = 0, j = lane;
    while (j >= C) { j -= C; kk++; }
    for (int f = lane; f < C2; f += 32) {
      const unsigned dst = (unsigned)__cvta_generic_to_shared(buf + kk * Cp + j);
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(T + f) : "memory");
      j += 32;
      while (j >= C) { j -= C; kk++; }
    }
  };
  const int64_t BN = (int64_t)B * N;
  const float oml = 1.0f - lam;

  for (int e = threadIdx.x; e < C2; e += kApplyThreads) E_s[e] = lam * ema_t[e];

  const int64_t ntiles = (BN + kApplyTile - 1) / kApplyTile;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t g0 = tile * kApplyTile;
    if (nbuf == 2 && g0 + warp < BN) prefetch(g0 + warp, svbuf);  // flies under the tile loads below
    asm volatile("cp.async.commit_group;" ::: "memory");
    __syncthreads();  // previous tile's ot fully stored; E_s visible on the first pass
    for (int e = threadIdx.x; e < C * kApplyTile; e += kApplyThreads) {
      const int c = e >> 5, pt = e & 31;
      const int64_t g = g0 + pt;
      float lv = 0.f, gv = 0.f;
      if (g < BN) {
        const int64_t b = g / N, n = g - b * N;
        lv = __ldg(logits + (b * C + c) * N + n);
        if (BWD) gv = __ldg(grad_out + (b * C + c) * N + n);
      }
      lt[c * 33 + pt] = lv;
      if (BWD) gt[c * 33 + pt] = gv;
    }
    __syncthreads();

    for (int pt = warp, it = 0; pt < kApplyTile; pt += kApplyWarps, it++) {
      const int64_t g = g0 + pt;
      if (g >= BN) break;  // warp-uniform
      float* sv = svbuf + (nbuf == 2 ? (size_t)(it & 1) * C * Cp : 0);
      if (nbuf == 2) {
        if (pt + kApplyWarps < kApplyTile && g + kApplyWarps < BN) prefetch(g + kApplyWarps, svbuf + (size_t)((it + 1) & 1) * C * Cp);
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 1;" ::: "memory");  // this point's matrix has landed (own copies; same lane blends them)
      } else {  // large C: one tile per warp, no overlap
        prefetch(g, sv);
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 0;" ::: "memory");
      }
      {  // blend in place: sv[kk][j] = lam*E + (1-lam)*T
        int kk = 0, j = lane;
        while (j >= C) { j -= C; kk++; }
        for (int f = lane; f < C2; f += 32) {
          sv[kk * Cp + j] = __fmaf_rn(oml, sv[kk * Cp + j], E_s[f]);
          j += 32;
          while (j >= C) { j -= C; kk++; }
        }
      }
      __syncwarp();
      for (int kk = lane; kk < C; kk += 32) {  // row L1 norms (F.normalize: x / max(||x||_1, 1e-12))
        const float* row = sv + kk * Cp;
        float s = 0.f, d = 0.f;
        for (int j = 0; j < C; j++) {
          const float v = row[j];
          s += fabsf(v);
          if (BWD) d = __fmaf_rn(gt[j * 33 + pt], v, d);
        }
        const float inv = 1.0f / fmaxf(s, 1e-12f);
        const float l = lt[kk * 33 + pt];
        if (BWD) {
          ot[kk * 33 + pt] = d * inv;      // grad_logits[kk] = sum_j go[j] * newT[kk][j]
          ca[kk] = oml * l * inv;          // d newT -> d insT scale
          cb[kk] = d * inv;                // sum_m go[m] * newT[kk][m]
        } else {
          ca[kk] = l * inv;
        }
      }
      __syncwarp();
      if (!BWD) {
        for (int j = lane; j < C; j += 32) {
          float o = 0.f;
          for (int kk = 0; kk < C; kk++) o = __fmaf_rn(ca[kk], sv[kk * Cp + j], o);
          ot[j * 33 + pt] = o;
        }
      } else if (grad_insT != nullptr) {
        float* __restrict__ gT = grad_insT + g * C2;
        int kk = 0, j = lane;
        while (j >= C) { j -= C; kk++; }
        for (int f = lane; f < C2; f += 32) {
          const float v = sv[kk * Cp + j];
          const float sgn = v > 0.f ? 1.0f : (v < 0.f ? -1.0f : 0.f);
          gT[f] = ca[kk] * (gt[j * 33 + pt] - sgn * cb[kk]);
          j += 32;
          while (j >= C) { j -= C; kk++; }
        }
      }
      __syncwarp();
    }
    __syncthreads();
    if (out_or_gl != nullptr) {
      for (int e = threadIdx.x; e < C * kApplyTile; e += kApplyThreads) {
        const int c = e >> 5, pt = e & 31;
        const int64_t g = g0 + pt;
        if (g < BN) {
          const int64_t b = g / N, n = g - b * N;
          out_or_gl[(b * C + c) * N + n] = ot[c * 33 + pt];
        }
      }
    }
  }
}

template <bool BWD>
static int launch_apply(int B, int C, int N, const float* logits, const float* ema_t, const float* insT,
                        float lam, const float* grad_out, float* out_or_gl, float* grad_insT, cudaStream_t s) {
  int nbuf = 2;
  if (ApplySmem(C, 2).bytes() > (size_t)112 * 1024) nbuf = 1;  // keep two CTAs per SM
  const ApplySmem L(C, nbuf);
  const size_t smem = L.bytes();
  static size_t configured = 0;
  if (smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(transition_apply_kernel<BWD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) { set_error("transition_apply: cudaFuncSetAttribute(%zu): %s", smem, cudaGetErrorString(e)); return GEOT_ERR_CUDA; }
    configured = smem;
  }
  const int64_t tiles = ((int64_t)B * N + kApplyTile - 1) / kApplyTile;
  const int per_sm = smem > 110 * 1024 ? 1 : (smem > 72 * 1024 ? 2 : (smem > 54 * 1024 ? 3 : 4));
  const int64_t cap = (int64_t)kNumSMs * per_sm;
  const int grid = (int)(tiles < cap ? tiles : cap);
  transition_apply_kernel<BWD><<<grid, kApplyThreads, smem, s>>>(B, C, N, nbuf, logits, ema_t, insT, lam, grad_out, out_or_gl, grad_insT);
  count_launch(1);
  return check_launch(BWD ? "transition_apply_kernel<bwd>" : "transition_apply_kernel<fwd>");
}

}  // namespace geot

using namespace geot;

GEOT_API int geot_transition_apply_forward(int B, int C, int N, const float* logits, const float* ema_t,
                                           const float* insT, float lam, float* out, geot_stream_t stream) {
  GEOT_REQUIRE(B >= 0 && N >= 0 && C >= 2 && C <= 64, "transition_apply_forward: bad sizes B=%d C=%d N=%d (2 <= C <= 64)", B, C, N);
  if (B == 0 || N == 0) return GEOT_OK;
  GEOT_REQUIRE(logits && ema_t && insT && out, "transition_apply_forward: NULL pointer");
  return launch_apply<false>(B, C, N, logits, ema_t, insT, lam, nullptr, out, nullptr, as_stream(stream));
}

GEOT_API int geot_transition_apply_backward(int B, int C, int N, const float* logits, const float* ema_t,
                                            const float* insT, float lam, const float* grad_out,
                                            float* grad_logits, float* grad_insT, geot_stream_t stream) {
  GEOT_REQUIRE(B >= 0 && N >= 0 && C >= 2 && C <= 64, "transition_apply_backward: bad sizes B=%d C=%d N=%d (2 <= C <= 64)", B, C, N);
  if (B == 0 || N == 0 || (grad_logits == nullptr && grad_insT == nullptr)) return GEOT_OK;
  GEOT_REQUIRE(logits && ema_t && insT && grad_out, "transition_apply_backward: NULL pointer");
  return launch_apply<true>(B, C, N, logits, ema_t, insT, lam, grad_out, grad_logits, grad_insT, as_stream(stream));
}
