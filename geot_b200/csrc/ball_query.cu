// geot_b200/csrc/ball_query.cu — ball query for sm_100a.
//
// Replaces (bit-exact indices)
//   openpoints/cpp/pointnet2_batch/src/ball_query_gpu.cu:15-51     (batched)
//   pointnet2/_ext_src/src/ball_query_gpu.cu:12-47                 (batched, same result)
//   openpoints/cpp/pointops/src/ballquery/ballquery_cuda_kernel.cu:26-78   (offset format)
// Semantics: the first `nsample` indices k, ascending, with d2 < radius^2; the remaining slots repeat
// the first hit; a row without any hit is all zeros (the reference relies on the caller's zero fill,
// group.py:194 — here the kernel writes every slot).
//
// Design (not a port).  The reference runs one THREAD per query: every thread streams the whole cloud
// from L2, the 32 queries of a warp diverge on their early exits and each hit is a scattered 4-byte
// store.  Here one WARP owns a query: the 32 lanes test 32 consecutive points per step, a ballot gives
// the hits already in ascending index order, and the warp stops exactly when its query is full.  The
// 32 warps of a CTA work on 32 queries of the same cloud and share ONE stream of the cloud through
// shared memory (cp.async double-buffered tiles, raw AoS: the stride-3 word reads are conflict free),
// so a cloud tile is fetched from L2 once per 32 queries instead of once per query; the CTA leaves as
// soon as all of its queries are full.
#include "common.cuh"

namespace geot {

void count_launch(int n);

constexpr int kBqTile = 1024;  // points per shared-memory tile (12 KB), two buffers

__device__ __forceinline__ void cp_async4(void* smem, const void* gmem) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// Stage nfloats contiguous floats global -> shared with the whole CTA.
__device__ __forceinline__ void stage_floats(float* dst, const float* __restrict__ src, int nfloats,
                                             bool aligned16) {
  const int t = threadIdx.x, nt = blockDim.x;
  if (aligned16) {
    const int n4 = nfloats >> 2;
    for (int i = t; i < n4; i += nt) cp_async16(dst + 4 * i, src + 4 * i);
    for (int i = (n4 << 2) + t; i < nfloats; i += nt) cp_async4(dst + i, src + i);
  } else {
    for (int i = t; i < nfloats; i += nt) cp_async4(dst + i, src + i);
  }
  cp_async_commit();
}

// One warp scans `count` staged points (tile-local) for its query.  Returns true when the row is full.
__device__ __forceinline__ bool ball_scan_tile(const float* __restrict__ s, int count, int base_k,
                                               float qx, float qy, float qz, float r2, int nsample,
                                               int* __restrict__ row, int& cnt, int& first) {
  const int lane = threadIdx.x & 31;
  for (int i = 0; i < count; i += 32) {
    const int p = i + lane;
    const float x = s[3 * p + 0], y = s[3 * p + 1], z = s[3 * p + 2];
    const float d2 = dist2_ref(qx, qy, qz, x, y, z);
    const bool hit = (p < count) && (d2 < r2);
    const unsigned mask = __ballot_sync(0xffffffffu, hit);
    if (mask) {
      if (cnt == 0) first = base_k + i + (__ffs(mask) - 1);
      const int slot = cnt + __popc(mask & lanemask_lt());
      if (hit && slot < nsample) row[slot] = base_k + p;
      cnt += __popc(mask);
      if (cnt >= nsample) return true;
    }
  }
  return false;
}

// Batched: grid (ceil(M/32), B), 1024 threads; warp w of CTA x handles query 32x + w of cloud y.
__global__ void __launch_bounds__(1024, 2)
ball_query_shared_kernel(int N, int M, float radius, int nsample, const float* __restrict__ new_xyz,
                         const float* __restrict__ xyz, int* __restrict__ idx, int aligned16) {
  __shared__ __align__(16) float tile[2][kBqTile * 3];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = blockIdx.y;
  const int q = blockIdx.x * 32 + warp;
  const float r2 = __fmul_rn(radius, radius);
  const float* __restrict__ cloud = xyz + (int64_t)b * N * 3;

  bool done = q >= M;
  float qx = 0.f, qy = 0.f, qz = 0.f;
  int* row = nullptr;
  if (!done) {
    const float* qp = new_xyz + ((int64_t)b * M + q) * 3;
    qx = qp[0]; qy = qp[1]; qz = qp[2];
    row = idx + ((int64_t)b * M + q) * nsample;
  }
  int cnt = 0, first = 0;

  const int ntiles = (N + kBqTile - 1) / kBqTile;
  stage_floats(tile[0], cloud, min(kBqTile, N) * 3, aligned16);
  for (int tI = 0; tI < ntiles; tI++) {
    const int buf = tI & 1;
    cp_async_wait_all();
    __syncthreads();
    if (tI + 1 < ntiles) {
      const int nb = (tI + 1) * kBqTile;
      stage_floats(tile[buf ^ 1], cloud + (int64_t)nb * 3, min(kBqTile, N - nb) * 3, aligned16);
    }
    if (!done) {
      const int base = tI * kBqTile;
      done = ball_scan_tile(tile[buf], min(kBqTile, N - base), base, qx, qy, qz, r2, nsample, row, cnt, first);
    }
    if (__syncthreads_and(done)) break;
  }
  cp_async_wait_all();
  if (row != nullptr) {
    const int fill = cnt > 0 ? first : 0;
    for (int l = min(cnt, nsample) + lane; l < nsample; l += 32) row[l] = fill;
  }
}

// Offset format (ragged segments): one warp per query, points straight from global/L2.
__global__ void __launch_bounds__(256)
ballquery_offset_kernel(int b, int m, float radius, int nsample, const float* __restrict__ xyz,
                        const float* __restrict__ new_xyz, const int* __restrict__ offset,
                        const int* __restrict__ new_offset, int* __restrict__ idx) {
  const int lane = threadIdx.x & 31;
  const int q = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (q >= m) return;
  int seg = 0;
  while (seg < b - 1 && q >= new_offset[seg]) seg++;  // get_bt_idx, knnquery_cuda_kernel.cu:51-62
  const int start = seg ? offset[seg - 1] : 0, end = offset[seg];
  const float r2 = __fmul_rn(radius, radius);
  const float qx = new_xyz[(int64_t)q * 3 + 0], qy = new_xyz[(int64_t)q * 3 + 1], qz = new_xyz[(int64_t)q * 3 + 2];
  int* row = idx + (int64_t)q * nsample;
  int cnt = 0, first = 0;
  for (int k0 = start; k0 < end; k0 += 32) {
    const int k = k0 + lane;
    float x = 0.f, y = 0.f, z = 0.f;
    if (k < end) { x = xyz[(int64_t)k * 3 + 0]; y = xyz[(int64_t)k * 3 + 1]; z = xyz[(int64_t)k * 3 + 2]; }
    const float d2 = dist2_ref(qx, qy, qz, x, y, z);
    const bool hit = (k < end) && (d2 < r2);
    const unsigned mask = __ballot_sync(0xffffffffu, hit);
    if (mask) {
      if (cnt == 0) first = k0 + (__ffs(mask) - 1);
      const int slot = cnt + __popc(mask & lanemask_lt());
      if (hit && slot < nsample) row[slot] = k;
      cnt += __popc(mask);
      if (cnt >= nsample) break;
    }
  }
  const int fill = cnt > 0 ? first : 0;
  for (int l = min(cnt, nsample) + lane; l < nsample; l += 32) row[l] = fill;
}

}  // namespace geot

using namespace geot;

GEOT_API int geot_ball_query(int B, int N, int M, float radius, int nsample, const float* new_xyz,
                             const float* xyz, int* idx, geot_stream_t stream) {
  GEOT_REQUIRE(B >= 0 && N >= 0 && M >= 0 && nsample >= 0, "ball_query: bad sizes");
  if (B == 0 || M == 0 || nsample == 0) return GEOT_OK;
  GEOT_REQUIRE(new_xyz && xyz && idx, "ball_query: NULL pointer");
  GEOT_REQUIRE(B <= 65535, "ball_query: B=%d exceeds grid.y", B);
  const int aligned16 = (((uintptr_t)xyz & 15) == 0) && ((N * 3) % 4 == 0);
  dim3 grid((M + 31) / 32, B);
  ball_query_shared_kernel<<<grid, 1024, 0, as_stream(stream)>>>(N, M, radius, nsample, new_xyz, xyz, idx, aligned16);
  count_launch(1);
  return check_launch("ball_query_shared_kernel");
}

GEOT_API int geot_ballquery_offset(int b, int m, float radius, int nsample, const float* xyz,
                                   const float* new_xyz, const int* offset, const int* new_offset,
                                   int* idx, geot_stream_t stream) {
  GEOT_REQUIRE(b >= 0 && m >= 0 && nsample >= 0, "ballquery_offset: bad sizes");
  if (b == 0 || m == 0 || nsample == 0) return GEOT_OK;
  GEOT_REQUIRE(new_xyz && xyz && idx && offset && new_offset, "ballquery_offset: NULL pointer");
  ballquery_offset_kernel<<<(m + 7) / 8, 256, 0, as_stream(stream)>>>(b, m, radius, nsample, xyz, new_xyz, offset, new_offset, idx);
  count_launch(1);
  return check_launch("ballquery_offset_kernel");
}
