// geot_b200/csrc/fps.cu — furthest point sampling for sm_100a.
//
// Replaces (bit-exact indices) the three FPS kernels of the reference:
//   openpoints/cpp/pointnet2_batch/src/sampling_gpu.cu:101-216   batched, first index 0, 1024-thread tree
//   pointops/src/sampling/sampling_cuda_kernel.cu:15-129, :175-307   offset format (+ weighted)
//   pointnet2/_ext_src/src/sampling_gpu.cu:73-176                 512-thread tree, |p|^2 <= 1e-3 skip
//
// Design (not a port): one CTA per cloud.  Coordinates are staged once into shared memory as SoA
// float4 planes, the running min-distance of every point lives in REGISTERS for the whole kernel
// (the reference re-reads and re-writes a global temp array every round), and the per-round arg-max
// is two warp REDUX reductions + one 32-slot shared exchange instead of a 10-level __syncthreads tree.
//
// Tie rule.  The reference's winner among equal distances is decided by its strided thread loop and
// shared-memory tree (sampling_gpu.cu:93-98,131-210): thread tid keeps the FIRST maximum of
// k = tid, tid+BS, ...; the tree keeps the lower slot on ties at strides BS/2 ... 1, so the winning
// thread minimises bit-reversed tid.  Hence  rank(k) = bitrev(k mod BS) * ceil(n/BS) + k div BS  and
// the reference picks the max distance, then the min rank.  We reproduce exactly that order without
// reproducing its thread layout.
#include "common.cuh"

namespace geot {

void count_launch(int n);

struct TieRule {
  int log2bs;  // BS = 1 << log2bs, the reference's block size for this n
};

__device__ __forceinline__ unsigned tie_rank(int k, int log2bs, int q_count) {
  const unsigned tid = (unsigned)k & ((1u << log2bs) - 1u);
  const unsigned q = (unsigned)k >> log2bs;
  const unsigned br = log2bs ? (__brev(tid) >> (32 - log2bs)) : 0u;
  return br * (unsigned)q_count + q;
}
__device__ __forceinline__ int tie_unrank(unsigned rank, int log2bs, int q_count) {
  const unsigned br = rank / (unsigned)q_count, q = rank % (unsigned)q_count;
  const unsigned tid = log2bs ? (__brev(br) >> (32 - log2bs)) : 0u;
  return (int)((q << log2bs) | tid);
}

// Segment of block b.  Batched: offset == nullptr, indices local.  Offset format: global indices.
__device__ __forceinline__ void fps_segment(const int* __restrict__ offset, const int* __restrict__ new_offset,
                                            int N, int M, int b, int64_t& start_n, int& n,
                                            int64_t& start_m, int& m, int& idx_base) {
  if (offset == nullptr) {
    start_n = (int64_t)b * N; n = N; start_m = (int64_t)b * M; m = M; idx_base = 0;
  } else {
    const int s_n = b ? offset[b - 1] : 0, s_m = b ? new_offset[b - 1] : 0;
    start_n = s_n; n = offset[b] - s_n; start_m = s_m; m = new_offset[b] - s_m; idx_base = s_n;
  }
}

// ---------------------------------------------------------------------------------------------
// Fast path: n <= CH*4096 points, 1024 threads, CH*4 points per thread.
// Thread t owns points c*4096 + 4t + {0,1,2,3}, c < CH  (one LDS.128 per plane per chunk).
// ---------------------------------------------------------------------------------------------
template <int CH>
__global__ void __launch_bounds__(1024, 1)
fps_smem_kernel(const float* __restrict__ xyz, const int* __restrict__ offset,
                const int* __restrict__ new_offset, int N, int M, int* __restrict__ idx, int log2bs,
                int mag_skip) {
  extern __shared__ float4 fps_smem[];
  float4* sx4 = fps_smem;
  float4* sy4 = sx4 + CH * 1024;
  float4* sz4 = sy4 + CH * 1024;
  float* sx = reinterpret_cast<float*>(sx4);
  float* sy = reinterpret_cast<float*>(sy4);
  float* sz = reinterpret_cast<float*>(sz4);
  __shared__ int s_val[32];
  __shared__ unsigned s_rank[32];

  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  int64_t start_n, start_m;
  int n, m, idx_base;
  fps_segment(offset, new_offset, N, M, blockIdx.x, start_n, n, start_m, m, idx_base);
  if (m <= 0) return;
  if (n > CH * 4096) return;  // caller understated n_max (the Python layer computes it from offset)
  const int q_count = (n + (1 << log2bs) - 1) >> log2bs;

  const float* __restrict__ src = xyz + start_n * 3;
  for (int k = t; k < CH * 4096; k += 1024) {
    float x = 0.f, y = 0.f, z = 0.f;
    if (k < n) { x = src[(int64_t)k * 3 + 0]; y = src[(int64_t)k * 3 + 1]; z = src[(int64_t)k * 3 + 2]; }
    sx[k] = x; sy[k] = y; sz[k] = z;
  }
  if (t == 0) idx[start_m] = idx_base;  // first sample = segment start (sampling_gpu.cu:120-122)
  __syncthreads();

  // Running min-distance, registers only.  -1 marks "never a candidate": padding beyond n and, for the
  // pointnet2/_ext variant, points the reference skips (|p|^2 <= 1e-3, _ext_src sampling_gpu.cu:103-104).
  float mind[CH * 4];
#pragma unroll
  for (int c = 0; c < CH; c++) {
    const float4 X = sx4[c * 1024 + t], Y = sy4[c * 1024 + t], Z = sz4[c * 1024 + t];
    const float xs[4] = {X.x, X.y, X.z, X.w}, ys[4] = {Y.x, Y.y, Y.z, Y.w}, zs[4] = {Z.x, Z.y, Z.z, Z.w};
#pragma unroll
    for (int e = 0; e < 4; e++) {
      const int k = c * 4096 + 4 * t + e;
      bool live = k < n;
      if (mag_skip) {
        const float mag = __fmaf_rn(zs[e], zs[e], __fmaf_rn(xs[e], xs[e], __fmul_rn(ys[e], ys[e])));
        live = live && !((double)mag <= 1e-3);
      }
      mind[c * 4 + e] = live ? 1e10f : -1.0f;
    }
  }

  int old = 0;
  for (int j = 1; j < m; j++) {
    const float x1 = sx[old], y1 = sy[old], z1 = sz[old];
    float best = -1.0f;
#pragma unroll
    for (int c = 0; c < CH; c++) {
      const float4 X = sx4[c * 1024 + t], Y = sy4[c * 1024 + t], Z = sz4[c * 1024 + t];
      const float xs[4] = {X.x, X.y, X.z, X.w}, ys[4] = {Y.x, Y.y, Y.z, Y.w}, zs[4] = {Z.x, Z.y, Z.z, Z.w};
#pragma unroll
      for (int e = 0; e < 4; e++) {
        const float d = dist2_ref(xs[e], ys[e], zs[e], x1, y1, z1);
        const float v = fminf(d, mind[c * 4 + e]);
        mind[c * 4 + e] = v;
        best = fmaxf(best, v);
      }
    }
    // Phase 1: block-wide max of the value.  Candidates are >= 0 or exactly -1, so the int view of the
    // float bits is order preserving.
    const int bits = __float_as_int(best);
    const int wmax = __reduce_max_sync(0xffffffffu, bits);
    if (lane == 0) s_val[warp] = wmax;
    __syncthreads();
    const int gmax = __reduce_max_sync(0xffffffffu, s_val[lane]);
    // Phase 2: among the points that hold the max, the reference's tie order (min rank).
    unsigned r = 0xffffffffu;
    if (bits == gmax) {
#pragma unroll
      for (int i = 0; i < CH * 4; i++) {
        if (__float_as_int(mind[i]) == gmax) {
          const int k = (i >> 2) * 4096 + 4 * t + (i & 3);
          r = min(r, tie_rank(k, log2bs, q_count));
        }
      }
    }
    const unsigned wr = __reduce_min_sync(0xffffffffu, r);
    if (lane == 0) s_rank[warp] = wr;
    __syncthreads();
    const unsigned grank = __reduce_min_sync(0xffffffffu, s_rank[lane]);
    old = tie_unrank(grank, log2bs, q_count);
    if (t == 0) idx[start_m + j] = idx_base + old;
  }
}

// ---------------------------------------------------------------------------------------------
// Fallback: any n, optional weights; min-distances in the caller's temp array (global), like the
// reference.  64-bit (value, ~rank) keys, shuffle reduction.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long fps_key(float v, unsigned rank) {
  const unsigned b = __float_as_uint(v);
  const unsigned u = (b & 0x80000000u) ? ~b : (b | 0x80000000u);  // order-preserving float -> uint
  return ((unsigned long long)u << 32) | (unsigned long long)(0xffffffffu - rank);
}

__global__ void __launch_bounds__(1024, 1)
fps_generic_kernel(const float* __restrict__ xyz, const int* __restrict__ offset,
                   const int* __restrict__ new_offset, int N, int M, const float* __restrict__ weights,
                   float* __restrict__ temp, int* __restrict__ idx, int log2bs, int mag_skip) {
  __shared__ unsigned long long s_key[32];
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  int64_t start_n, start_m;
  int n, m, idx_base;
  fps_segment(offset, new_offset, N, M, blockIdx.x, start_n, n, start_m, m, idx_base);
  if (m <= 0) return;
  const int q_count = (n + (1 << log2bs) - 1) >> log2bs;
  const float* __restrict__ src = xyz + start_n * 3;
  float* __restrict__ tmp = temp + start_n;
  const float* __restrict__ w = weights ? weights + start_n : nullptr;
  if (t == 0) idx[start_m] = idx_base;

  int old = 0;
  for (int j = 1; j < m; j++) {
    const float x1 = src[(int64_t)old * 3 + 0], y1 = src[(int64_t)old * 3 + 1], z1 = src[(int64_t)old * 3 + 2];
    unsigned long long best = fps_key(-1.0f, 0u);
    for (int k = t; k < n; k += 1024) {
      const float x2 = src[(int64_t)k * 3 + 0], y2 = src[(int64_t)k * 3 + 1], z2 = src[(int64_t)k * 3 + 2];
      if (mag_skip) {
        const float mag = __fmaf_rn(z2, z2, __fmaf_rn(x2, x2, __fmul_rn(y2, y2)));
        if ((double)mag <= 1e-3) continue;
      }
      float d = dist2_ref(x2, y2, z2, x1, y1, z1);
      // sampling_cuda_kernel.cu:222 — max(float, 1e-12) is a double, so the product is formed in double.
      if (w) d = (float)((double)d * fmax((double)w[k], 1e-12));
      const float d2 = fminf(d, tmp[k]);
      tmp[k] = d2;
      const unsigned long long key = fps_key(d2, tie_rank(k, log2bs, q_count));
      best = key > best ? key : best;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
      best = other > best ? other : best;
    }
    if (lane == 0) s_key[warp] = best;
    __syncthreads();
    unsigned long long g = s_key[lane];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const unsigned long long other = __shfl_xor_sync(0xffffffffu, g, o);
      g = other > g ? other : g;
    }
    __syncthreads();  // s_key is rewritten next round
    old = tie_unrank(0xffffffffu - (unsigned)(g & 0xffffffffull), log2bs, q_count);
    if (t == 0) idx[start_m + j] = idx_base + old;
  }
}

static int ilog2(int v) {
  int l = 0;
  while ((1 << (l + 1)) <= v) l++;
  return l;
}

template <int CH>
static int launch_smem(int blocks, const float* xyz, const int* offset, const int* new_offset, int N,
                       int M, int* idx, int log2bs, int mag_skip, cudaStream_t s) {
  const size_t smem = (size_t)3 * CH * 4096 * sizeof(float);
  static bool configured = false;  // per instantiation
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(fps_smem_kernel<CH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) { set_error("fps: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return GEOT_ERR_CUDA; }
    configured = true;
  }
  fps_smem_kernel<CH><<<blocks, 1024, smem, s>>>(xyz, offset, new_offset, N, M, idx, log2bs, mag_skip);
  count_launch(1);
  return check_launch("fps_smem_kernel");
}

static int fps_dispatch(int blocks, int n_max, const float* xyz, const int* offset, const int* new_offset,
                        int N, int M, const float* weights, float* temp, int* idx, int flags,
                        cudaStream_t s) {
  const int cap = (flags & GEOT_FPS_CAP512) ? 512 : 1024;
  const int log2bs = ilog2(ref_opt_n_threads(n_max, cap));
  const int mag_skip = (flags & GEOT_FPS_MAGSKIP) ? 1 : 0;
  const bool generic = (flags & GEOT_FPS_FORCE_GENERIC) || weights != nullptr || n_max > 4 * 4096;
  if (!generic) {
    const int ch = (n_max + 4095) / 4096;
    switch (ch) {
      case 1: return launch_smem<1>(blocks, xyz, offset, new_offset, N, M, idx, log2bs, mag_skip, s);
      case 2: return launch_smem<2>(blocks, xyz, offset, new_offset, N, M, idx, log2bs, mag_skip, s);
      case 3: return launch_smem<3>(blocks, xyz, offset, new_offset, N, M, idx, log2bs, mag_skip, s);
      default: return launch_smem<4>(blocks, xyz, offset, new_offset, N, M, idx, log2bs, mag_skip, s);
    }
  }
  if (temp == nullptr) {
    set_error("fps: this shape/variant needs the temp scratch (n_max=%d > 16384, weights or forced fallback)", n_max);
    return GEOT_ERR_WORKSPACE;
  }
  fps_generic_kernel<<<blocks, 1024, 0, s>>>(xyz, offset, new_offset, N, M, weights, temp, idx, log2bs, mag_skip);
  count_launch(1);
  return check_launch("fps_generic_kernel");
}

}  // namespace geot

using namespace geot;

GEOT_API int geot_furthest_point_sampling(int B, int N, int M, const float* xyz, float* temp, int* idx,
                                          int flags, geot_stream_t stream) {
  GEOT_REQUIRE(B >= 0 && N >= 1 && M >= 0, "furthest_point_sampling: bad sizes B=%d N=%d M=%d", B, N, M);
  if (B == 0 || M == 0) return GEOT_OK;
  GEOT_REQUIRE(xyz && idx, "furthest_point_sampling: NULL xyz/idx");
  return fps_dispatch(B, N, xyz, nullptr, nullptr, N, M, nullptr, temp, idx, flags, as_stream(stream));
}

GEOT_API int geot_furthestsampling(int b, int n_max, const float* xyz, const int* offset,
                                   const int* new_offset, const float* weights, float* tmp, int* idx,
                                   int flags, geot_stream_t stream) {
  GEOT_REQUIRE(b >= 0 && n_max >= 1, "furthestsampling: bad sizes b=%d n_max=%d", b, n_max);
  if (b == 0) return GEOT_OK;
  GEOT_REQUIRE(xyz && idx && offset && new_offset, "furthestsampling: NULL pointer");
  return fps_dispatch(b, n_max, xyz, offset, new_offset, 0, 0, weights, tmp, idx, flags, as_stream(stream));
}
