// geot_b200/csrc/transition.cu — instance-dependent transition matrices: prediction (sig_t_mean / sig_t)
// and its backward, for sm_100a.
//
// Replaces openpoints/models/backbone/transformer.py:1120-1131 (sig_t_mean.forward: a Python loop of C
// Linear(2C->C) calls, each materialising a (BN,2C) concat and a strided slice write, then clamp and
// F.normalize) and :1089-1097 (sig_t.forward), plus their autograd.
//
// Formulation.  row kk of T_n is  W_kk . [p_n ; cm_kk] = A_kk p_n + b_kk  with  A_kk = W_kk[:, :C] and the
// point-independent  b_kk = W_kk[:, C:] cm_kk  (computed once by a C*C-thread pre-kernel).  So the whole
// predictor is one (BN x C) . (C x C*C) product with a row-wise epilogue (clamp, L1 normalise over j).
// FP32 FFMA on CUDA cores by design (north_star: no tensor cores; fp32 parity 1e-5): the weights A
// (C*C*C floats, 144 KB at C=33) stay resident in shared memory for the whole persistent CTA, a lane owns
// two points and one row kk at a time (2*C accumulators in registers), every weight fetch is a broadcast
// LDS.128 feeding 8 FFMAs, and finished rows leave through a per-warp staging tile so that the (BN,C,C)
// store is coalesced.
#include "common.cuh"

namespace geot {

void count_launch(int n);

constexpr int kSigtThreads = 384;  // 12 warps: 33 rows in 3 rounds; 8 warps left the FMA pipe idle half the time
constexpr int kSigtWarps = kSigtThreads / 32;
constexpr int kSigtTile = 64;  // points per CTA tile: lane l owns points l and l+32

template <int C>
struct SigtSmem {
  static constexpr int CP = (C + 3) & ~3;  // padded row of A_s so that float4 reads stay aligned
  static constexpr size_t a_floats = (size_t)C * C * CP;
  static constexpr size_t bias_floats = ((size_t)C * C + 3) & ~(size_t)3;
  static constexpr size_t p_floats = (size_t)C * kSigtTile;
  static constexpr size_t stage_floats = (size_t)kSigtWarps * 32 * C;
  static constexpr size_t bytes = (a_floats + bias_floats + p_floats + stage_floats) * sizeof(float);
};

// bias[kk*C + j] = sum_i W[(kk*C+j)*stride + C + i] * cm[kk*C + i]    (zero when cm == nullptr)
__global__ void sigt_bias_kernel(int C, const float* __restrict__ W, int stride, const float* __restrict__ cm,
                                 float* __restrict__ bias) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= C * C) return;
  float acc = 0.f;
  if (cm != nullptr) {
    const int kk = r / C;
    for (int i = 0; i < C; i++) acc = __fmaf_rn(W[(int64_t)r * stride + C + i], cm[kk * C + i], acc);
  }
  bias[r] = acc;
}

// Shared skeleton of the forward and of backward stage 1 (which recomputes the products).
// BWD == false: insT_or_graw = insT (out).  BWD == true: grad_in = grad_insT, insT_or_graw = g_raw (out).
template <int C, bool BWD>
__global__ void __launch_bounds__(kSigtThreads, 1)
sigt_kernel(int B, int N, const float* __restrict__ x, const float* __restrict__ W, int stride,
            const float* __restrict__ bias, const float* __restrict__ grad_in,
            float* __restrict__ insT_or_graw) {
  using S = SigtSmem<C>;
  constexpr int CP = S::CP;
  extern __shared__ __align__(16) float sigt_smem[];
  float* A_s = sigt_smem;                      // [kk][i][CP]  = W[(kk*C + j)*stride + i]
  float* bias_s = A_s + S::a_floats;           // [kk*C + j]
  float* p_s = bias_s + S::bias_floats;        // [i][64]
  float* stage_all = p_s + S::p_floats;        // [warp][32*C]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* stage = stage_all + (size_t)warp * 32 * C;
  const int64_t BN = (int64_t)B * N;

  for (int e = threadIdx.x; e < C * C * C; e += kSigtThreads) {
    const int r = e / C, i = e - r * C;        // coalesced read of W rows
    const int kk = r / C, j = r - kk * C;
    A_s[((size_t)kk * C + i) * CP + j] = W[(int64_t)r * stride + i];
  }
  for (int e = threadIdx.x; e < C * C; e += kSigtThreads) bias_s[e] = bias[e];

  const int64_t ntiles = (BN + kSigtTile - 1) / kSigtTile;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t g0 = tile * kSigtTile;
    __syncthreads();  // previous tile's p_s fully consumed (and A_s / bias_s visible on the first pass)
    for (int e = threadIdx.x; e < C * kSigtTile; e += kSigtThreads) {
      const int i = e / kSigtTile, pt = e - i * kSigtTile;
      const int64_t g = g0 + pt;
      float v = 0.f;
      if (g < BN) {
        const int64_t b = g / N, n = g - b * N;
        v = __ldg(x + (b * C + i) * N + n);
      }
      p_s[e] = v;
    }
    __syncthreads();

    for (int kk = warp; kk < C; kk += kSigtWarps) {
      float acc0[C], acc1[C];
#pragma unroll
      for (int j = 0; j < C; j++) { acc0[j] = 0.f; acc1[j] = 0.f; }
      const float* __restrict__ Ak = A_s + (size_t)kk * C * CP;
#pragma unroll 3
      for (int i = 0; i < C; i++) {
        const float p0 = p_s[i * kSigtTile + lane], p1 = p_s[i * kSigtTile + 32 + lane];
        const float4* __restrict__ a4 = reinterpret_cast<const float4*>(Ak + i * CP);
#pragma unroll
        for (int q = 0; q < CP / 4; q++) {
          const float4 a = a4[q];
          const float av[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
          for (int e = 0; e < 4; e++) {
            if (4 * q + e < C) {
              acc0[4 * q + e] = __fmaf_rn(av[e], p0, acc0[4 * q + e]);
              acc1[4 * q + e] = __fmaf_rn(av[e], p1, acc1[4 * q + e]);
            }
          }
        }
      }
#pragma unroll
      for (int h = 0; h < 2; h++) {
        float* acc = h ? acc1 : acc0;
        const int64_t gbase = g0 + h * 32;  // first point of this half
        if (BWD) {  // stage-in the upstream gradient rows of these 32 points, coalesced
          const float* __restrict__ gin = grad_in + (gbase * C + kk) * C;  // row kk of point gbase; next point C*C floats on
          const int nvalid = (int)min((int64_t)32, BN - gbase);
          for (int f = lane; f < 32 * C; f += 32) {
            const int pt = f / C, j = f - pt * C;
            stage[f] = (pt < nvalid) ? __ldg(gin + pt * (C * C) + j) : 0.f;
          }
          __syncwarp();
        }
        float s = 0.f, dot = 0.f;
#pragma unroll
        for (int j = 0; j < C; j++) {
          const float raw = acc[j] + bias_s[kk * C + j];
          const float y = fminf(fmaxf(raw, 1e-5f), 1.0f - 1e-5f);  // torch.clamp(min=1e-5, max=1-1e-5)
          s += y;
          if (BWD) {
            dot = __fmaf_rn(stage[lane * C + j], y, dot);
            acc[j] = raw;
          } else {
            acc[j] = y;
          }
        }
        const float inv = 1.0f / fmaxf(s, 1e-12f);  // F.normalize(p=1): x / max(||x||_1, eps); all y > 0
        if (BWD) {
          // d/dy of y/s:  (g_j - sum_l g_l y_l / s) / s ; clamp passes the gradient inside [min, max]
          const float corr = dot * inv;
#pragma unroll
          for (int j = 0; j < C; j++) {
            const bool pass = (acc[j] >= 1e-5f) && (acc[j] <= 1.0f - 1e-5f);
            const float gj = stage[lane * C + j];
            stage[lane * C + j] = pass ? (gj - corr) * inv : 0.f;
          }
        } else {
#pragma unroll
          for (int j = 0; j < C; j++) stage[lane * C + j] = acc[j] * inv;
        }
        __syncwarp();
        {
          float* __restrict__ orow = insT_or_graw + (gbase * C + kk) * C;
          const int nvalid = (int)min((int64_t)32, BN - gbase);
#pragma unroll 4
          for (int f = lane; f < 32 * C; f += 32) {
            const int pt = f / C, j = f - pt * C;
            if (pt < nvalid) orow[pt * (C * C) + j] = stage[f];
          }
        }
        __syncwarp();
      }
    }
  }
}

// Any C (2..64): one thread per (point, kk), weights through L1.  Correct, not fast; C in {17, 33} (the
// reference's and BASELINE's class counts) take the shared-memory kernel above.
template <bool BWD>
__global__ void __launch_bounds__(128)
sigt_generic_kernel(int B, int C, int N, const float* __restrict__ x, const float* __restrict__ W, int stride,
                    const float* __restrict__ bias, const float* __restrict__ grad_in,
                    float* __restrict__ insT_or_graw) {
  const int64_t BN = (int64_t)B * N;
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= BN * C) return;
  const int64_t g = t / C;
  const int kk = (int)(t - g * C);
  const int64_t b = g / N, n = g - b * N;
  const float* __restrict__ xp = x + b * C * N + n;
  auto raw_of = [&](int j) {
    const float* __restrict__ w = W + (int64_t)(kk * C + j) * stride;
    float a = 0.f;
    for (int i = 0; i < C; i++) a = __fmaf_rn(__ldg(w + i), __ldg(xp + (int64_t)i * N), a);
    return a + bias[kk * C + j];
  };
  float s = 0.f, dot = 0.f;
  for (int j = 0; j < C; j++) {
    const float y = fminf(fmaxf(raw_of(j), 1e-5f), 1.0f - 1e-5f);
    s += y;
    if (BWD) dot = __fmaf_rn(grad_in[(g * C + kk) * C + j], y, dot);
  }
  const float inv = 1.0f / fmaxf(s, 1e-12f);
  for (int j = 0; j < C; j++) {
    const float raw = raw_of(j);
    if (BWD) {
      const bool pass = (raw >= 1e-5f) && (raw <= 1.0f - 1e-5f);
      insT_or_graw[(g * C + kk) * C + j] = pass ? (grad_in[(g * C + kk) * C + j] - dot * inv) * inv : 0.f;
    } else {
      insT_or_graw[(g * C + kk) * C + j] = fminf(fmaxf(raw, 1e-5f), 1.0f - 1e-5f) * inv;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Backward stage 2:  G[r][i] = sum_n g_raw[n][r] * p[n][i]  (i < C)  and  G[r][C] = sum_n g_raw[n][r].
// Each CTA reduces a contiguous slice of points into a private (C*C) x (C+1) partial (registers), the
// finalize kernel adds the partials in CTA order: bit-reproducible, no float atomics.
// ---------------------------------------------------------------------------------------------
constexpr int kGwThreads = 384;
constexpr int kGwTile = 8;  // points per shared tile

// The g_raw tile of the NEXT 8 points travels (16-byte cp.async: pts_per_cta is a multiple of 8, so a tile starts on a
// 32-byte boundary) while the current one is multiplied: with a single buffer the kernel alternated between waiting
// for 35 KB and 1.4 us of FMAs (ncu: 29 % issue-active, 176 us at C=33, BN=32000).
__device__ __forceinline__ void gw_cp_async16(void* smem, const void* gmem) {
  const unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gmem) : "memory");
}
template <int C>
__global__ void __launch_bounds__(kGwThreads, 1)
sigt_gw_partial_kernel(int B, int N, int64_t pts_per_cta, const float* __restrict__ x,
                       const float* __restrict__ g_raw, float* __restrict__ partial) {
  constexpr int R = C * C;
  constexpr int RPT = (R + kGwThreads - 1) / kGwThreads;
  constexpr int CI = C + 1;
  constexpr int CIP = (CI + 3) & ~3;
  constexpr int GT = (kGwTile * R + 3) & ~3;  // floats per g buffer
  extern __shared__ __align__(16) float gw_smem[];
  auto g_of = [&](int b2) { return gw_smem + b2 * GT; };
  auto p_of = [&](int b2) { return gw_smem + 2 * GT + b2 * (kGwTile * CIP); };
  const int64_t BN = (int64_t)B * N;
  const int64_t first = (int64_t)blockIdx.x * pts_per_cta;
  const int64_t last = min(BN, first + pts_per_cta);
  float acc[RPT][CI];
#pragma unroll
  for (int q = 0; q < RPT; q++)
#pragma unroll
    for (int i = 0; i < CI; i++) acc[q][i] = 0.f;

  auto issue_g = [&](int buf, int64_t g0, int cnt) {  // g_raw[g0 .. g0 + cnt) -> g_buf[buf]
    const float* __restrict__ src = g_raw + g0 * R;
    const int nfl = cnt * R;
    if (((reinterpret_cast<uintptr_t>(src)) & 15) == 0) {
      const int n4 = nfl >> 2;
      for (int e = threadIdx.x; e < n4; e += kGwThreads) gw_cp_async16(g_of(buf) + 4 * e, src + 4 * e);
      for (int e = (n4 << 2) + threadIdx.x; e < nfl; e += kGwThreads) g_of(buf)[e] = __ldg(src + e);
    } else {
      for (int e = threadIdx.x; e < nfl; e += kGwThreads) g_of(buf)[e] = __ldg(src + e);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  auto load_p = [&](int64_t g0, int cnt) -> float {  // this thread's entry of the (8 x CIP) input tile: [p | 1 | 0..]
    const int e = threadIdx.x;
    if (e >= kGwTile * CIP) return 0.f;
    const int pt = e / CIP, i = e - pt * CIP;
    if (pt >= cnt) return 0.f;
    if (i < C) {
      const int64_t g = g0 + pt, b = g / N, n = g - b * N;
      return __ldg(x + (b * C + i) * N + n);
    }
    return i == C ? 1.0f : 0.f;
  };
  static_assert(kGwTile * CIP <= kGwThreads, "one input-tile entry per thread");

  int buf = 0;
  if (first < last) {
    const int cnt0 = (int)min((int64_t)kGwTile, last - first);
    issue_g(0, first, cnt0);
    const float pv0 = load_p(first, cnt0);
    if (threadIdx.x < kGwTile * CIP) p_of(0)[threadIdx.x] = pv0;
    asm volatile("cp.async.wait_group 0;" ::: "memory");
  }
  __syncthreads();
  for (int64_t g0 = first; g0 < last; g0 += kGwTile) {
    const int cnt = (int)min((int64_t)kGwTile, last - g0);
    const int64_t gn = g0 + kGwTile;
    const bool more = gn < last;
    const int cntn = more ? (int)min((int64_t)kGwTile, last - gn) : 0;
    float pvn = 0.f;
    if (more) {
      issue_g(buf ^ 1, gn, cntn);
      pvn = load_p(gn, cntn);
    }
    const float* __restrict__ g_s = g_of(buf);
    const float* __restrict__ p_s = p_of(buf);
    for (int pt = 0; pt < cnt; pt++) {
      float gv[RPT];
#pragma unroll
      for (int q = 0; q < RPT; q++) {
        const int r = threadIdx.x + q * kGwThreads;
        gv[q] = r < R ? g_s[pt * R + r] : 0.f;
      }
      const float4* __restrict__ p4 = reinterpret_cast<const float4*>(p_s + pt * CIP);
#pragma unroll
      for (int w = 0; w < CIP / 4; w++) {
        const float4 p = p4[w];
        const float pv[4] = {p.x, p.y, p.z, p.w};
#pragma unroll
        for (int e = 0; e < 4; e++) {
          if (4 * w + e < CI) {
#pragma unroll
            for (int q = 0; q < RPT; q++) acc[q][4 * w + e] = __fmaf_rn(gv[q], pv[e], acc[q][4 * w + e]);
          }
        }
      }
    }
    if (more) {
      if (threadIdx.x < kGwTile * CIP) p_of(buf ^ 1)[threadIdx.x] = pvn;
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();  // the next tile is complete and everyone is done with this one
    buf ^= 1;
  }
  float* __restrict__ dst = partial + (size_t)blockIdx.x * R * CI;
#pragma unroll
  for (int q = 0; q < RPT; q++) {
    const int r = threadIdx.x + q * kGwThreads;
    if (r < R) {
#pragma unroll
      for (int i = 0; i < CI; i++) dst[(size_t)r * CI + i] = acc[q][i];
    }
  }
}
template <int C>
constexpr size_t gw_smem_bytes() { return (size_t)(2 * (((kGwTile * C * C) + 3) & ~3) + 2 * kGwTile * (((C + 1) + 3) & ~3)) * sizeof(float); }

// Generic C: one thread per (r, i) column walks its CTA's slice of points.
__global__ void __launch_bounds__(256)
sigt_gw_partial_generic_kernel(int B, int C, int N, int64_t pts_per_cta, const float* __restrict__ x,
                               const float* __restrict__ g_raw, float* __restrict__ partial) {
  const int R = C * C, CI = C + 1;
  const int64_t BN = (int64_t)B * N;
  const int64_t first = (int64_t)blockIdx.y * pts_per_cta, last = min(BN, first + pts_per_cta);
  const int o = blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= R * CI) return;
  const int r = o / CI, i = o - r * CI;
  float acc = 0.f;
  for (int64_t g = first; g < last; g++) {
    const int64_t b = g / N, n = g - b * N;
    const float p = i < C ? __ldg(x + (b * C + i) * N + n) : 1.0f;
    acc = __fmaf_rn(__ldg(g_raw + g * R + r), p, acc);
  }
  partial[(size_t)blockIdx.y * R * CI + o] = acc;
}

// grad_W[r*stride + i] = sum_cta partial[cta][r][i]            (i < C)
// grad_W[r*stride + C + i] = (sum_cta partial[cta][r][C]) * cm[kk*C + i]   (sig_t_mean only)
__global__ void __launch_bounds__(256)
sigt_gw_final_kernel(int C, int stride, int nparts, const float* __restrict__ partial,
                     const float* __restrict__ cm, float* __restrict__ grad_W) {
  const int R = C * C, CI = C + 1;
  const int o = blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= R * stride) return;
  const int r = o / stride, i = o - r * stride;
  const int col = i < C ? i : C;
  float acc = 0.f;
  for (int c = 0; c < nparts; c++) acc += partial[(size_t)c * R * CI + (size_t)r * CI + col];
  if (i >= C) acc *= cm[(r / C) * C + (i - C)];
  grad_W[o] = acc;
}

static int sigt_grid(int64_t BN) {
  const int64_t tiles = (BN + kSigtTile - 1) / kSigtTile;
  return (int)(tiles < num_sms() ? tiles : num_sms());
}

template <int C, bool BWD>
static int launch_sigt(int B, int N, const float* x, const float* W, int stride, const float* bias,
                       const float* grad_in, float* out, cudaStream_t s) {
  static DynSmem ds;  // per instantiation
  if (int rc = ds.need(sigt_kernel<C, BWD>, SigtSmem<C>::bytes, "sigt")) return rc;
  sigt_kernel<C, BWD><<<sigt_grid((int64_t)B * N), kSigtThreads, SigtSmem<C>::bytes, s>>>(B, N, x, W, stride, bias, grad_in, out);
  count_launch(1);
  return check_launch(BWD ? "sigt_kernel<bwd>" : "sigt_kernel<fwd>");
}

template <bool BWD>
static int dispatch_sigt(int B, int C, int N, const float* x, const float* W, int stride, const float* bias,
                         const float* grad_in, float* out, cudaStream_t s) {
  if (C == 17) return launch_sigt<17, BWD>(B, N, x, W, stride, bias, grad_in, out, s);
  if (C == 33) return launch_sigt<33, BWD>(B, N, x, W, stride, bias, grad_in, out, s);
  const int64_t total = (int64_t)B * N * C;
  sigt_generic_kernel<BWD><<<(unsigned)((total + 127) / 128), 128, 0, s>>>(B, C, N, x, W, stride, bias, grad_in, out);
  count_launch(1);
  return check_launch("sigt_generic_kernel");
}

static inline size_t align256(size_t b) { return (b + 255) & ~(size_t)255; }
static inline int gw_parts(int64_t BN) {
  const int64_t p = (BN + 63) / 64;
  return (int)(p < num_sms() ? (p < 1 ? 1 : p) : num_sms());
}

}  // namespace geot

using namespace geot;

GEOT_API size_t geot_sigt_workspace_bytes(int C) { return C > 0 ? align256((size_t)C * C * sizeof(float)) : 0; }

static int sigt_check(const char* what, int B, int C, int N, const float* x, const float* W, int stride, const float* cm) {
  GEOT_REQUIRE(B >= 0 && N >= 0 && C >= 2 && C <= 64, "%s: bad sizes B=%d C=%d N=%d (2 <= C <= 64)", what, B, C, N);
  GEOT_REQUIRE(x && W, "%s: NULL pointer", what);
  GEOT_REQUIRE(stride == (cm ? 2 * C : C), "%s: w_row_stride=%d must be 2C with cm (sig_t_mean) or C without (sig_t)", what, stride);
  return GEOT_OK;
}

GEOT_API int geot_sigt_forward(int B, int C, int N, const float* x, const float* W, int w_row_stride,
                               const float* cm, float* insT, void* workspace, size_t workspace_bytes,
                               geot_stream_t stream) {
  if (B == 0 || N == 0) return GEOT_OK;
  int rc = sigt_check("sigt_forward", B, C, N, x, W, w_row_stride, cm);
  if (rc) return rc;
  GEOT_REQUIRE(insT, "sigt_forward: NULL insT");
  if (workspace == nullptr || workspace_bytes < geot_sigt_workspace_bytes(C)) {
    set_error("sigt_forward: workspace %zu < %zu bytes", workspace_bytes, geot_sigt_workspace_bytes(C));
    return GEOT_ERR_WORKSPACE;
  }
  cudaStream_t s = as_stream(stream);
  float* bias = reinterpret_cast<float*>(workspace);
  sigt_bias_kernel<<<(C * C + 127) / 128, 128, 0, s>>>(C, W, w_row_stride, cm, bias);
  count_launch(1);
  rc = check_launch("sigt_bias_kernel");
  if (rc) return rc;
  return dispatch_sigt<false>(B, C, N, x, W, w_row_stride, bias, nullptr, insT, s);
}

GEOT_API size_t geot_sigt_backward_workspace_bytes(int B, int C, int N) {
  if (B <= 0 || C <= 0 || N <= 0) return 0;
  return geot_sigt_workspace_bytes(C) + align256((size_t)gw_parts((int64_t)B * N) * C * C * (C + 1) * sizeof(float));
}

GEOT_API int geot_sigt_backward(int B, int C, int N, const float* x, const float* W, int w_row_stride,
                                const float* cm, const float* grad_insT, float* g_raw, float* grad_W,
                                void* workspace, size_t workspace_bytes, geot_stream_t stream) {
  int rc = sigt_check("sigt_backward", B, C, N, x, W, w_row_stride, cm);
  if (rc) return rc;
  GEOT_REQUIRE(grad_W && (B == 0 || N == 0 || (grad_insT && g_raw)), "sigt_backward: NULL pointer");
  cudaStream_t s = as_stream(stream);
  if (B == 0 || N == 0) {
    cudaError_t e = cudaMemsetAsync(grad_W, 0, (size_t)C * C * w_row_stride * sizeof(float), s);
    if (e != cudaSuccess) { set_error("sigt_backward: memset: %s", cudaGetErrorString(e)); return GEOT_ERR_CUDA; }
    return GEOT_OK;
  }
  if (workspace == nullptr || workspace_bytes < geot_sigt_backward_workspace_bytes(B, C, N)) {
    set_error("sigt_backward: workspace %zu < %zu bytes", workspace_bytes, geot_sigt_backward_workspace_bytes(B, C, N));
    return GEOT_ERR_WORKSPACE;
  }
  float* bias = reinterpret_cast<float*>(workspace);
  float* partial = reinterpret_cast<float*>(reinterpret_cast<char*>(workspace) + geot_sigt_workspace_bytes(C));
  sigt_bias_kernel<<<(C * C + 127) / 128, 128, 0, s>>>(C, W, w_row_stride, cm, bias);
  count_launch(1);
  rc = check_launch("sigt_bias_kernel");
  if (rc) return rc;
  rc = dispatch_sigt<true>(B, C, N, x, W, w_row_stride, bias, grad_insT, g_raw, s);
  if (rc) return rc;
  const int64_t BN = (int64_t)B * N;
  const int parts = gw_parts(BN);
  const int64_t per = (((BN + parts - 1) / parts) + 7) & ~(int64_t)7;  // whole 8-point tiles per CTA (16-byte aligned tile starts)
  if (C == 17) {
    static DynSmem ds17;
    if (int rcs = ds17.need(sigt_gw_partial_kernel<17>, gw_smem_bytes<17>(), "sigt_backward")) return rcs;
    sigt_gw_partial_kernel<17><<<parts, kGwThreads, gw_smem_bytes<17>(), s>>>(B, N, per, x, g_raw, partial);
  } else if (C == 33) {
    static DynSmem ds33;
    if (int rcs = ds33.need(sigt_gw_partial_kernel<33>, gw_smem_bytes<33>(), "sigt_backward")) return rcs;
    sigt_gw_partial_kernel<33><<<parts, kGwThreads, gw_smem_bytes<33>(), s>>>(B, N, per, x, g_raw, partial);
  } else {
    dim3 grid((C * C * (C + 1) + 255) / 256, parts);
    sigt_gw_partial_generic_kernel<<<grid, 256, 0, s>>>(B, C, N, per, x, g_raw, partial);
  }
  count_launch(1);
  rc = check_launch("sigt_gw_partial_kernel");
  if (rc) return rc;
  sigt_gw_final_kernel<<<(C * C * w_row_stride + 255) / 256, 256, 0, s>>>(C, w_row_stride, parts, partial, cm, grad_W);
  count_launch(1);
  return check_launch("sigt_gw_final_kernel");
}
