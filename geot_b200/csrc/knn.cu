// geot_b200/csrc/knn.cu — k-nearest-neighbour query for sm_100a.
//
// Replaces pointops/src/knnquery/knnquery_cuda_kernel.cu:21-108 (and its openpoints/cpp/pointops twin):
// for every query the nsample nearest points of the SAME segment, global int32 indices, ascending d2,
// unfilled slots (segment shorter than nsample) = (1e10, segment start).
//
// Design (not a port).  The reference runs one thread per query with an 8 KB local-memory max-heap
// (best_dist[1000], best_idx[1000]) and streams the cloud from L2 once per thread.  Here a WARP owns a
// query: the 32 warps of a CTA share one cp.async stream of the segment through shared memory, the 32
// lanes evaluate 32 points per step, and the running top-(k+1) is a sorted list held in REGISTERS
// across the lanes (slot-major, 1-4 slots per lane); a candidate that beats the current threshold is
// inserted with a ballot + shuffle shift.  Candidates are taken in index order and inserted after
// equal distances, i.e. the list is the stable ascending order.
//
// Exactness.  With pairwise-distinct distances among the k+1 best, the reference's heap + heap_sort
// yields exactly that ascending list.  Under exact ties its membership/order depend on the heap
// layout (strict '<' insert :97, reheap swaps on equality :29), so the fast path keeps ONE extra
// element, flags any query whose k+1 best contain an equal pair, and a second kernel redoes the
// flagged queries with the literal heap algorithm (also used for nsample >= 128).
#include "common.cuh"

namespace geot {

void count_launch(int n);
__device__ __forceinline__ void knn_cp_async4(void* smem, const void* gmem) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void knn_cp_async16(void* smem, const void* gmem) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem));
}

constexpr int kKnnTile = 1024;

// Segment bounds for block-row `seg`.
__device__ __forceinline__ void knn_segment(const int* __restrict__ offset, const int* __restrict__ new_offset,
                                            int n_per, int m_per, int seg, int& start, int& end,
                                            int& qstart, int& qend) {
  if (offset == nullptr) {
    start = seg * n_per; end = start + n_per; qstart = seg * m_per; qend = qstart + m_per;
  } else {
    start = seg ? offset[seg - 1] : 0; end = offset[seg];
    qstart = seg ? new_offset[seg - 1] : 0; qend = new_offset[seg];
  }
}

// S slots per lane: list element j = s*32 + lane, ascending.  K1 = nsample + 1 <= 32*S.
template <int S>
__global__ void __launch_bounds__(1024, 1)
knn_warp_kernel(int nsample, const float* __restrict__ xyz, const float* __restrict__ new_xyz,
                const int* __restrict__ offset, const int* __restrict__ new_offset, int n_per,
                int m_per, int* __restrict__ idx, float* __restrict__ dist2,
                unsigned char* __restrict__ flags) {
  __shared__ __align__(16) float tile[2][kKnnTile * 3];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int start, end, qstart, qend;
  knn_segment(offset, new_offset, n_per, m_per, blockIdx.y, start, end, qstart, qend);
  if (qstart + (int)blockIdx.x * 32 >= qend) return;  // whole CTA beyond this segment's queries
  const int q = qstart + blockIdx.x * 32 + warp;
  const bool active = q < qend;
  const int K1 = nsample + 1;

  float qx = 0.f, qy = 0.f, qz = 0.f;
  if (active) { qx = new_xyz[(int64_t)q * 3 + 0]; qy = new_xyz[(int64_t)q * 3 + 1]; qz = new_xyz[(int64_t)q * 3 + 2]; }
  float ld[S];
  int li[S];
#pragma unroll
  for (int s = 0; s < S; s++) { ld[s] = 1e10f; li[s] = start; }
  float tau = 1e10f;  // current K1-th best (strict '<' admits, knnquery_cuda_kernel.cu:97)

  const int n = end - start;
  const float* __restrict__ seg = xyz + (int64_t)start * 3;
  const bool aligned16 = ((((uintptr_t)seg) & 15) == 0);
  const int ntiles = (n + kKnnTile - 1) / kKnnTile;

  auto stage = [&](int buf, int tI) {
    const int base = tI * kKnnTile;
    const int nfl = min(kKnnTile, n - base) * 3;
    const float* src = seg + (int64_t)base * 3;
    float* dst = tile[buf];
    if (aligned16) {  // tile bases are multiples of 1024 points = 12288 B, so alignment carries over
      const int n4 = nfl >> 2;
      for (int i = threadIdx.x; i < n4; i += 1024) knn_cp_async16(dst + 4 * i, src + 4 * i);
      for (int i = (n4 << 2) + threadIdx.x; i < nfl; i += 1024) knn_cp_async4(dst + i, src + i);
    } else {
      for (int i = threadIdx.x; i < nfl; i += 1024) knn_cp_async4(dst + i, src + i);
    }
    asm volatile("cp.async.commit_group;");
  };

  if (ntiles > 0) stage(0, 0);
  for (int tI = 0; tI < ntiles; tI++) {
    const int buf = tI & 1;
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    if (tI + 1 < ntiles) stage(buf ^ 1, tI + 1);
    if (active) {
      const float* __restrict__ sm = tile[buf];
      const int base = tI * kKnnTile;
      const int cnt = min(kKnnTile, n - base);
      for (int i = 0; i < cnt; i += 32) {
        const int p = i + lane;
        const float d = dist2_ref(qx, qy, qz, sm[3 * p + 0], sm[3 * p + 1], sm[3 * p + 2]);
        unsigned mask = __ballot_sync(0xffffffffu, (p < cnt) && (d < tau));
        while (mask) {
          const int src = __ffs(mask) - 1;
          mask &= mask - 1;
          const float nd = __shfl_sync(0xffffffffu, d, src);
          if (!(nd < tau)) continue;  // tau shrank while draining this ballot
          const int ni = start + base + i + src;
          // insertion position = number of list elements <= nd (stable: after equals)
          int pos = 0;
#pragma unroll
          for (int s = 0; s < S; s++) pos += __popc(__ballot_sync(0xffffffffu, (s * 32 + lane < K1) && (ld[s] <= nd)));
#pragma unroll
          for (int s = S - 1; s >= 0; s--) {
            float pd = __shfl_up_sync(0xffffffffu, ld[s], 1);
            int pi = __shfl_up_sync(0xffffffffu, li[s], 1);
            if (s > 0) {
              const float wd = __shfl_sync(0xffffffffu, ld[s - 1], 31);
              const int wi = __shfl_sync(0xffffffffu, li[s - 1], 31);
              if (lane == 0) { pd = wd; pi = wi; }
            }
            const int j = s * 32 + lane;
            if (j > pos) { ld[s] = pd; li[s] = pi; }
            else if (j == pos) { ld[s] = nd; li[s] = ni; }
          }
          float tv = ld[0];
#pragma unroll
          for (int s = 1; s < S; s++) tv = (((K1 - 1) >> 5) == s) ? ld[s] : tv;
          tau = __shfl_sync(0xffffffffu, tv, (K1 - 1) & 31);
        }
      }
    }
    __syncthreads();
  }

  if (active) {
    // tie check over adjacent pairs (j, j+1), j < nsample, both real (< 1e10)
    bool tie = false;
#pragma unroll
    for (int s = 0; s < S; s++) {
      float nx = __shfl_down_sync(0xffffffffu, ld[s], 1);
      if (s + 1 < S) {
        const float wrap = __shfl_sync(0xffffffffu, ld[s + 1], 0);
        if (lane == 31) nx = wrap;
      }
      const int j = s * 32 + lane;
      if (j + 1 < K1 && (s + 1 < S || lane < 31) && ld[s] == nx && nx < 1e10f) tie = true;
    }
    tie = __any_sync(0xffffffffu, tie);
    if (lane == 0) flags[q] = tie ? 1 : 0;
#pragma unroll
    for (int s = 0; s < S; s++) {
      const int j = s * 32 + lane;
      if (j < nsample) {
        idx[(int64_t)q * nsample + j] = li[s];
        dist2[(int64_t)q * nsample + j] = ld[s];
      }
    }
  }
}

// Literal heap algorithm (knnquery_cuda_kernel.cu:21-48,86-107), one thread per query; run for the
// queries the warp kernel flagged (flags != nullptr) or for every query (flags == nullptr).
template <int CAP>
__device__ __forceinline__ void heap_reheap(float (&dist)[CAP], int (&id)[CAP], int k) {
  int root = 0, child = 1;
  while (child < k) {
    if (child + 1 < k && dist[child + 1] > dist[child]) child++;
    if (dist[root] > dist[child]) return;
    const float tf = dist[root]; dist[root] = dist[child]; dist[child] = tf;
    const int ti = id[root]; id[root] = id[child]; id[child] = ti;
    root = child;
    child = root * 2 + 1;
  }
}

template <int CAP>
__global__ void __launch_bounds__(128)
knn_heap_kernel(int b, int m, int nsample, const float* __restrict__ xyz,
                const float* __restrict__ new_xyz, const int* __restrict__ offset,
                const int* __restrict__ new_offset, int n_per, int m_per, int* __restrict__ idx,
                float* __restrict__ dist2, const unsigned char* __restrict__ flags) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= m) return;
  if (flags != nullptr && flags[q] == 0) return;
  int start, end;
  if (offset == nullptr) {
    const int seg = q / m_per;
    start = seg * n_per; end = start + n_per;
  } else {
    int seg = 0;
    while (seg < b - 1 && q >= new_offset[seg]) seg++;
    start = seg ? offset[seg - 1] : 0; end = offset[seg];
  }
  const float qx = new_xyz[(int64_t)q * 3 + 0], qy = new_xyz[(int64_t)q * 3 + 1], qz = new_xyz[(int64_t)q * 3 + 2];
  float bd[CAP];
  int bi[CAP];
  for (int i = 0; i < nsample; i++) { bd[i] = 1e10f; bi[i] = start; }
  for (int i = start; i < end; i++) {
    const float d2 = dist2_ref(qx, qy, qz, xyz[(int64_t)i * 3 + 0], xyz[(int64_t)i * 3 + 1], xyz[(int64_t)i * 3 + 2]);
    if (d2 < bd[0]) { bd[0] = d2; bi[0] = i; heap_reheap<CAP>(bd, bi, nsample); }
  }
  for (int i = nsample - 1; i > 0; i--) {  // heap_sort :39-48
    const float tf = bd[0]; bd[0] = bd[i]; bd[i] = tf;
    const int ti = bi[0]; bi[0] = bi[i]; bi[i] = ti;
    heap_reheap<CAP>(bd, bi, i);
  }
  for (int i = 0; i < nsample; i++) {
    idx[(int64_t)q * nsample + i] = bi[i];
    dist2[(int64_t)q * nsample + i] = bd[i];
  }
}

}  // namespace geot

using namespace geot;

GEOT_API size_t geot_knnquery_workspace_bytes(int m) { return m > 0 ? (((size_t)m + 255) & ~(size_t)255) : 0; }

GEOT_API int geot_knnquery(int b, int m, int nsample, const float* xyz, const float* new_xyz,
                           const int* offset, const int* new_offset, int n_per, int m_per, int* idx,
                           float* dist2, void* workspace, size_t workspace_bytes, geot_stream_t stream) {
  GEOT_REQUIRE(b >= 0 && m >= 0, "knnquery: bad sizes");
  GEOT_REQUIRE(nsample >= 1 && nsample <= 1000, "knnquery: nsample=%d outside [1,1000] (reference heap size, knnquery_cuda_kernel.cu:86)", nsample);
  if (b == 0 || m == 0) return GEOT_OK;
  GEOT_REQUIRE(xyz && new_xyz && idx && dist2, "knnquery: NULL pointer");
  GEOT_REQUIRE((offset == nullptr) == (new_offset == nullptr), "knnquery: offset and new_offset must both be given or both NULL");
  GEOT_REQUIRE(m_per >= 1 && b <= 65535, "knnquery: m_per=%d must bound the queries per segment; b=%d", m_per, b);
  if (offset == nullptr) GEOT_REQUIRE(n_per >= 0 && (int64_t)b * m_per == m, "knnquery: batched layout needs b*m_per == m");
  cudaStream_t s = as_stream(stream);
  if (nsample + 1 <= 128) {
    if (workspace == nullptr || workspace_bytes < geot_knnquery_workspace_bytes(m)) {
      set_error("knnquery: workspace %zu < %zu bytes", workspace_bytes, geot_knnquery_workspace_bytes(m));
      return GEOT_ERR_WORKSPACE;
    }
    unsigned char* flags = reinterpret_cast<unsigned char*>(workspace);
    dim3 grid((m_per + 31) / 32, b);
    const int S = (nsample + 1 + 31) / 32;
    switch (S) {
      case 1: knn_warp_kernel<1><<<grid, 1024, 0, s>>>(nsample, xyz, new_xyz, offset, new_offset, n_per, m_per, idx, dist2, flags); break;
      case 2: knn_warp_kernel<2><<<grid, 1024, 0, s>>>(nsample, xyz, new_xyz, offset, new_offset, n_per, m_per, idx, dist2, flags); break;
      case 3: knn_warp_kernel<3><<<grid, 1024, 0, s>>>(nsample, xyz, new_xyz, offset, new_offset, n_per, m_per, idx, dist2, flags); break;
      default: knn_warp_kernel<4><<<grid, 1024, 0, s>>>(nsample, xyz, new_xyz, offset, new_offset, n_per, m_per, idx, dist2, flags); break;
    }
    count_launch(1);
    int rc = check_launch("knn_warp_kernel");
    if (rc) return rc;
    knn_heap_kernel<128><<<(m + 127) / 128, 128, 0, s>>>(b, m, nsample, xyz, new_xyz, offset, new_offset, n_per, m_per, idx, dist2, flags);
    count_launch(1);
    return check_launch("knn_heap_kernel<128>");
  }
  knn_heap_kernel<1000><<<(m + 127) / 128, 128, 0, s>>>(b, m, nsample, xyz, new_xyz, offset, new_offset, n_per, m_per, idx, dist2, nullptr);
  count_launch(1);
  return check_launch("knn_heap_kernel<1000>");
}
