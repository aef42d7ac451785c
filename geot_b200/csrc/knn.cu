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

// Sorted-list insertion shared by the two fast kernels.  The list (K1 = k + 1 entries across the lanes, S per lane) is
// kept in (distance, index) order, so the K1 best it ends with do not depend on the order in which the points are
// visited: equal distances rank by index.  That IS the result under GEOT_KNN_TIES_BY_INDEX; under the default (the
// reference's heap semantics) a query whose best K1 contain an exact tie is replayed by knn_heap_warp_kernel.
// A point is admitted when it ranks before the current last entry (tau, tau_i).
// LEX = false (the default semantics: a tie anywhere among the best K1 sends the query to the heap replay, so the order
// of equal entries here is irrelevant): admit on distance alone, insert after equals - three operations fewer per
// insertion and no index of the last entry to carry (the insertion is the kernels' inner loop: 0.36 vs 0.47 ms on the
// config-2 kNN).  LEX = true (GEOT_KNN_TIES_BY_INDEX): (distance, index) order, independent of the visiting order.
template <bool LEX>
__device__ __forceinline__ bool knn_admit(float nd, int ni, float tau, int tau_i) {
  return LEX ? (nd < tau || (nd == tau && ni < tau_i)) : (nd < tau);
}
template <int S, bool LEX>
__device__ __forceinline__ void knn_list_insert(float (&ld)[S], int (&li)[S], int K1, int lane, float nd, int ni, float& tau, int& tau_i) {
  int pos = 0;  // entries that rank before the new one
#pragma unroll
  for (int s = 0; s < S; s++) {
    const bool before = LEX ? (ld[s] < nd || (ld[s] == nd && li[s] < ni)) : (ld[s] <= nd);
    pos += __popc(__ballot_sync(0xffffffffu, (s * 32 + lane < K1) && before));
  }
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
  int ti = li[0];
#pragma unroll
  for (int s = 1; s < S; s++) {
    tv = (((K1 - 1) >> 5) == s) ? ld[s] : tv;
    if (LEX) ti = (((K1 - 1) >> 5) == s) ? li[s] : ti;
  }
  tau = __shfl_sync(0xffffffffu, tv, (K1 - 1) & 31);
  if (LEX) tau_i = __shfl_sync(0xffffffffu, ti, (K1 - 1) & 31);
}

template <int S, bool LEX>
__global__ void __launch_bounds__(1024, 1)
knn_warp_kernel(int nsample, const float* __restrict__ xyz, const float* __restrict__ new_xyz,
                const int* __restrict__ offset, const int* __restrict__ new_offset, int n_per,
                int m_per, int* __restrict__ idx, float* __restrict__ dist2,
                unsigned char* __restrict__ flags, const unsigned char* __restrict__ seg_grid) {
  __shared__ __align__(16) float tile[2][kKnnTile * 3];
  if (seg_grid != nullptr && seg_grid[blockIdx.y]) return;  // this segment is served by knn_grid_kernel
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
  float tau = 1e10f;  // current K1-th best (distance, index)
  int tau_i = start;

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
        unsigned mask = __ballot_sync(0xffffffffu, (p < cnt) && (LEX ? d <= tau : d < tau));
        while (mask) {
          const int src = __ffs(mask) - 1;
          mask &= mask - 1;
          const float nd = __shfl_sync(0xffffffffu, d, src);
          const int ni = start + base + i + src;
          if (!knn_admit<LEX>(nd, ni, tau, tau_i)) continue;  // the list moved while draining this ballot
          knn_list_insert<S, LEX>(ld, li, K1, lane, nd, ni, tau, tau_i);
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

// ---------------------------------------------------------------------------------------------
// Grid path (segments of >= kKnnGridMinN points, nsample < 128).  The brute-force kernel above evaluates
// n distances per query; here every segment is binned once into a uniform grid sized so that one cell holds
// ~0.65*(k+1) points (a surface patch: G = sqrt(n / that) cells per axis, at most 32), and a query searches the
// 3x3x3 block around its cell, then - only if needed - the 5x5x5 shell.  After each ring the search is complete
// as soon as the current (k+1)-th distance is smaller than the distance from the query to the nearest face of
// the searched block that still has unsearched cells behind it (computed in cell units with a 2e-4 margin for
// the fp32 binning arithmetic): no unsearched point can then enter the list.  Queries that are still open after
// the second ring are flagged and redone by the literal heap kernel, exactly like exact-tie queries; so is any
// query whose k+1 best contain an equal pair (the candidate order of the grid is not the index order).
// ---------------------------------------------------------------------------------------------
struct KnnGridHeader {
  int use_grid;
  int dx, dy, dz;
  float lox, loy, loz, inv;
  float cell;
  int pad[7];
};
constexpr int kKnnHdrBytes = 64;
constexpr int kKnnMaxDim = 256;       // cells per axis
constexpr int kKnnMaxCells = 32768;   // cells in total (the build keeps the histogram in shared memory)
constexpr int kKnnGridMinN = 3072;
__host__ __device__ inline size_t knn_seg_ws_bytes() { return kKnnHdrBytes + (((size_t)(kKnnMaxCells + 1) * 4 + 15) & ~(size_t)15); }

__device__ __forceinline__ int knn_cell(float p, float lo, float inv) { return __float2int_rd((p - lo) * inv); }

// One CTA per segment: bounding box -> grid -> counting sort of {x, y, z, local index} into `sorted`.
__global__ void __launch_bounds__(1024, 1)
knn_build_kernel(int nsample, float alpha, const float* __restrict__ xyz, const int* __restrict__ offset, int n_per,
                 unsigned char* __restrict__ seg_grid, unsigned char* __restrict__ seg_ws,
                 float4* __restrict__ sorted_all) {
  extern __shared__ int knn_hist[];
  __shared__ float s_red[6][32];
  __shared__ int s_scan[32];
  __shared__ int s_occupied;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int seg = blockIdx.x;
  const int start = offset ? (seg ? offset[seg - 1] : 0) : seg * n_per;
  const int end = offset ? offset[seg] : start + n_per;
  const int N = end - start;
  // Slot = start / kKnnGridMinN: segments of >= kKnnGridMinN points start that far apart, so slots are distinct.
  const size_t slot = (size_t)(start / kKnnGridMinN) * knn_seg_ws_bytes();
  KnnGridHeader* hdr = reinterpret_cast<KnnGridHeader*>(seg_ws + slot);
  int* cell_start = reinterpret_cast<int*>(seg_ws + slot + kKnnHdrBytes);
  if (N < kKnnGridMinN) {  // small segment: the brute-force kernel is as fast
    if (t == 0) seg_grid[seg] = 0;
    return;
  }
  const float* __restrict__ src = xyz + (int64_t)start * 3;
  float4* __restrict__ sorted = sorted_all + start;
  const float kInf = __int_as_float(0x7f800000);
  float lo[3] = {kInf, kInf, kInf}, hi[3] = {-kInf, -kInf, -kInf};
#pragma unroll 4
  for (int k = t; k < N; k += 1024) {
#pragma unroll
    for (int a = 0; a < 3; a++) {
      const float v = src[(int64_t)k * 3 + a];
      lo[a] = fminf(lo[a], v); hi[a] = fmaxf(hi[a], v);
    }
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
    lo[a] = l; hi[a] = h;
    ext = fmaxf(ext, h - l);
  }
  if (!(ext > 0.f) || !(ext < kInf)) {  // degenerate cloud (one point repeated, NaN/inf): brute force handles it
    if (t == 0) seg_grid[seg] = 0;
    return;
  }
  // Cell size.  Target: `per_cell` points per OCCUPIED cell, so that the k+1 nearest usually lie within one cell
  // of the query (ring 1).  Occupancy depends on the shape (surface ~ G^2 cells, volume ~ G^3), so it is measured:
  // histogram with a first guess, rescale G with the dimension estimated from the previous two passes.
  const float per_cell = fmaxf(1.5f, alpha * (float)(nsample + 1));
  int G = min(kKnnMaxDim, max(2, (int)ceilf(sqrtf((float)N / per_cell))));
  int G_prev = 0;
  float occ_prev = 0.f;
  float cell = 0.f, inv = 0.f;
  int dims[3], ncell = 0;
  auto code_of = [&](float x, float y, float z) -> int {
    const int cx = min(max(knn_cell(x, lo[0], inv), 0), dims[0] - 1);
    const int cy = min(max(knn_cell(y, lo[1], inv), 0), dims[1] - 1);
    const int cz = min(max(knn_cell(z, lo[2], inv), 0), dims[2] - 1);
    return (cz * dims[1] + cy) * dims[0] + cx;
  };
#pragma unroll 1
  for (int pass = 0; pass < 3; pass++) {
    bool capped = false;
    for (;;) {  // G cells along the longest axis; shrink until the whole grid fits the cell table
      cell = ext / ((float)G - 0.5f);
      inv = 1.0f / cell;
#pragma unroll
      for (int a = 0; a < 3; a++) dims[a] = min(kKnnMaxDim, max(1, knn_cell(hi[a], lo[a], inv) + 1));
      ncell = dims[0] * dims[1] * dims[2];
      if (ncell <= kKnnMaxCells || G <= 2) break;
      G = max(2, min(G - 1, (int)((float)G * cbrtf((float)kKnnMaxCells / (float)ncell))));
      capped = true;
    }
    if (t == 0) s_occupied = 0;
    for (int i = t; i < ncell; i += 1024) knn_hist[i] = 0;
    __syncthreads();
#pragma unroll 4
    for (int k = t; k < N; k += 1024)
      atomicAdd(&knn_hist[code_of(src[(int64_t)k * 3], src[(int64_t)k * 3 + 1], src[(int64_t)k * 3 + 2])], 1);
    __syncthreads();
    if (pass == 2) break;
    int occ_cnt = 0;
    for (int i = t; i < ncell; i += 1024) occ_cnt += knn_hist[i] != 0;
    occ_cnt = __reduce_add_sync(0xffffffffu, occ_cnt);
    if (lane == 0 && occ_cnt) atomicAdd(&s_occupied, occ_cnt);
    __syncthreads();
    const float occ = (float)N / (float)max(s_occupied, 1);
    __syncthreads();  // s_occupied is cleared at the top of the next pass
    if (occ >= 0.75f * per_cell && occ <= 1.4f * per_cell) break;
    float d = 2.5f;
    if (pass == 1) d = fminf(fmaxf(__logf(occ_prev / occ) / __logf((float)G / (float)G_prev), 1.5f), 3.f);
    const int Gn = min(kKnnMaxDim, max(2, (int)rintf((float)G * __powf(occ / per_cell, 1.f / d))));
    if (Gn == G || (capped && Gn > G)) break;
    G_prev = G; occ_prev = occ; G = Gn;
  }
  {
    const int PER = (ncell + 1023) / 1024;
    const int i0 = t * PER;
    int sum = 0;
    for (int i = i0; i < min(i0 + PER, ncell); i++) sum += knn_hist[i];
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
    for (int i = i0; i < min(i0 + PER, ncell); i++) {
      const int v = knn_hist[i];
      knn_hist[i] = run;
      cell_start[i] = run;
      run += v;
    }
    if (t == 0) cell_start[ncell] = N;
  }
  __syncthreads();
#pragma unroll 4
  for (int k = t; k < N; k += 1024) {
    const float x = src[(int64_t)k * 3], y = src[(int64_t)k * 3 + 1], z = src[(int64_t)k * 3 + 2];
    sorted[atomicAdd(&knn_hist[code_of(x, y, z)], 1)] = make_float4(x, y, z, __int_as_float(k));
  }
  if (t == 0) {
    seg_grid[seg] = 1;
    hdr->use_grid = 1;
    hdr->dx = dims[0]; hdr->dy = dims[1]; hdr->dz = dims[2];
    hdr->lox = lo[0]; hdr->loy = lo[1]; hdr->loz = lo[2]; hdr->inv = inv; hdr->cell = cell;
  }
}

// Warp per query, 8 queries per CTA.
template <int S, bool LEX>
__global__ void __launch_bounds__(256)
knn_grid_kernel(int nsample, const float* __restrict__ new_xyz, const int* __restrict__ offset,
                const int* __restrict__ new_offset, int n_per, int m_per, int* __restrict__ idx,
                float* __restrict__ dist2, unsigned char* __restrict__ flags,
                const unsigned char* __restrict__ seg_grid, const unsigned char* __restrict__ seg_ws,
                const float4* __restrict__ sorted_all) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int seg = blockIdx.y;
  int start, end, qstart, qend;
  knn_segment(offset, new_offset, n_per, m_per, seg, start, end, qstart, qend);
  if (!seg_grid[seg]) return;
  const int q = qstart + blockIdx.x * 8 + warp;
  if (q >= qend) return;
  const size_t slot = (size_t)(start / kKnnGridMinN) * knn_seg_ws_bytes();
  const KnnGridHeader* hdr = reinterpret_cast<const KnnGridHeader*>(seg_ws + slot);
  const int* __restrict__ cell_start = reinterpret_cast<const int*>(seg_ws + slot + kKnnHdrBytes);
  const float4* __restrict__ sorted = sorted_all + start;
  const int K1 = nsample + 1;
  const int dx = hdr->dx, dy = hdr->dy, dz = hdr->dz;
  const float inv = hdr->inv, cell = hdr->cell;
  const float qx = new_xyz[(int64_t)q * 3 + 0], qy = new_xyz[(int64_t)q * 3 + 1], qz = new_xyz[(int64_t)q * 3 + 2];
  const float ux = (qx - hdr->lox) * inv, uy = (qy - hdr->loy) * inv, uz = (qz - hdr->loz) * inv;
  // clamped so the integer arithmetic below cannot overflow for far-away / non-finite queries
  const int cx = (int)fminf(fmaxf(floorf(ux), -4.f), (float)dx + 3.f);
  const int cy = (int)fminf(fmaxf(floorf(uy), -4.f), (float)dy + 3.f);
  const int cz = (int)fminf(fmaxf(floorf(uz), -4.f), (float)dz + 3.f);

  float ld[S];
  int li[S];
#pragma unroll
  for (int s = 0; s < S; s++) { ld[s] = 1e10f; li[s] = start; }
  float tau = 1e10f;
  int tau_i = start;

  auto scan_run = [&](int s0, int e0) {
    for (int i0 = s0; i0 < e0; i0 += 32) {
      const int p = i0 + lane;
      float d = 0.f;
      int pi = 0;
      if (p < e0) {
        const float4 v = __ldg(sorted + p);
        d = dist2_ref(qx, qy, qz, v.x, v.y, v.z);
        pi = __float_as_int(v.w);
      }
      unsigned mask = __ballot_sync(0xffffffffu, (p < e0) && (LEX ? d <= tau : d < tau));
      while (mask) {
        const int src = __ffs(mask) - 1;
        mask &= mask - 1;
        const float nd = __shfl_sync(0xffffffffu, d, src);
        const int ni = start + __shfl_sync(0xffffffffu, pi, src);
        if (!knn_admit<LEX>(nd, ni, tau, tau_i)) continue;
        knn_list_insert<S, LEX>(ld, li, K1, lane, nd, ni, tau, tau_i);
      }
    }
  };
  // x-run [xa, xb] of row (y, z), clipped to the grid; empty when the row is outside
  auto run_of = [&](int xa, int xb, int y, int z, int& s0, int& e0) {
    s0 = 0; e0 = 0;
    if (y < 0 || y >= dy || z < 0 || z >= dz) return;
    xa = max(xa, 0); xb = min(xb, dx - 1);
    if (xa > xb) return;
    const int row = (z * dy + y) * dx;
    s0 = __ldg(cell_start + row + xa);
    e0 = __ldg(cell_start + row + xb + 1);
  };
  // squared distance below which nothing outside the (2R+1)^3 block can lie
  auto safe_d2 = [&](int R) -> float {
    float rho = __int_as_float(0x7f800000);
    if (cx - R > 0) rho = fminf(rho, ux - (float)(cx - R));
    if (cx + R + 1 < dx) rho = fminf(rho, (float)(cx + R + 1) - ux);
    if (cy - R > 0) rho = fminf(rho, uy - (float)(cy - R));
    if (cy + R + 1 < dy) rho = fminf(rho, (float)(cy + R + 1) - uy);
    if (cz - R > 0) rho = fminf(rho, uz - (float)(cz - R));
    if (cz + R + 1 < dz) rho = fminf(rho, (float)(cz + R + 1) - uz);
    rho = fmaxf(rho - 2e-4f * (1.f + fabsf(ux) + fabsf(uy) + fabsf(uz)) * 0.03125f - 2e-4f, 0.f) * cell;
    return rho * rho * 0.9999f;
  };

  bool done = false;
  {  // ring 1: the 3x3x3 block as 9 x-runs
    int s0 = 0, e0 = 0;
    if (lane < 9) run_of(cx - 1, cx + 1, cy + lane % 3 - 1, cz + lane / 3 - 1, s0, e0);
    scan_run(__shfl_sync(0xffffffffu, s0, 4), __shfl_sync(0xffffffffu, e0, 4));  // the query's own row first: tau drops early
#pragma unroll 1
    for (int r = 0; r < 9; r++)
      if (r != 4) scan_run(__shfl_sync(0xffffffffu, s0, r), __shfl_sync(0xffffffffu, e0, r));
    done = tau < safe_d2(1);
  }
  if (!done) {  // ring 2: the 5x5x5 shell; border rows are full runs, interior rows contribute their two end cells
    int s0 = 0, e0 = 0, s1 = 0, e1 = 0;
    if (lane < 25) {
      const int oy = lane % 5 - 2, oz = lane / 5 - 2;
      if (oy == -2 || oy == 2 || oz == -2 || oz == 2) run_of(cx - 2, cx + 2, cy + oy, cz + oz, s0, e0);
      else { run_of(cx - 2, cx - 2, cy + oy, cz + oz, s0, e0); run_of(cx + 2, cx + 2, cy + oy, cz + oz, s1, e1); }
    }
#pragma unroll 1
    for (int r = 0; r < 25; r++) {
      scan_run(__shfl_sync(0xffffffffu, s0, r), __shfl_sync(0xffffffffu, e0, r));
      scan_run(__shfl_sync(0xffffffffu, s1, r), __shfl_sync(0xffffffffu, e1, r));
    }
    done = tau < safe_d2(2);
  }
  if (!done) {  // still open (query far from the cloud, or a very uneven density): scan the whole segment
#pragma unroll
    for (int s = 0; s < S; s++) { ld[s] = 1e10f; li[s] = start; }
    tau = 1e10f;
    tau_i = start;
    scan_run(0, end - start);
  }
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

// Same literal algorithm for the flagged queries of the fast paths, one WARP per query: the lanes evaluate 32
// points per step and ballot those below the heap top; lane 0 then applies them in index order (re-checking
// against the moving top) to a heap kept in shared memory.  Insert sequence and heap layout are exactly the
// sequential ones, so the result is; only the ~n/32 filter steps and the few hundred heap updates remain serial.
// The reference's sift-down (knnquery_cuda_kernel.cu:15-37: the larger child, the right one only when strictly
// larger; stop only when the root is STRICTLY larger, so equal keys swap) with the sinking entry kept in registers:
// the chain of swaps moves one value down, so writing the children up and the value once at the end leaves the same
// arrays - three shared-memory round trips per level instead of five.
__device__ __forceinline__ void heap_reheap_smem(float* dist, int* id, int k) {
  int root = 0, child = 1;
  const float v = dist[0];
  const int vi = id[0];
  while (child < k) {
    float dc = dist[child];
    if (child + 1 < k) {
      const float dr = dist[child + 1];
      if (dr > dc) { child++; dc = dr; }
    }
    if (v > dc) break;
    dist[root] = dc;
    id[root] = id[child];
    root = child;
    child = root * 2 + 1;
  }
  dist[root] = v;
  id[root] = vi;
}

__global__ void __launch_bounds__(256)
knn_heap_warp_kernel(int b, int m, int nsample, const float* __restrict__ xyz, const float* __restrict__ new_xyz,
                     const int* __restrict__ offset, const int* __restrict__ new_offset, int n_per, int m_per,
                     int* __restrict__ idx, float* __restrict__ dist2, const unsigned char* __restrict__ flags) {
  __shared__ float s_bd[8][128];
  __shared__ int s_bi[8][128];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int q = blockIdx.x * 8 + warp;
  if (q >= m || flags[q] == 0) return;
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
  float* bd = s_bd[warp];
  int* bi = s_bi[warp];
  for (int i = lane; i < nsample; i += 32) { bd[i] = 1e10f; bi[i] = start; }
  __syncwarp();
  float top = 1e10f;
  // Flagged queries are rare (1 of 32000 on the cfg4 clouds) but each walks its whole segment: with one 32-point chunk per
  // round trip a single replay took 0.37 ms - longer than the grid search of all the others.  Eight chunks of
  // distances are in flight; they are then offered to the heap chunk by chunk, in index order, each against the
  // threshold left by the previous one: the reference's sequence of insertions, unchanged.
  constexpr int kChunks = 8;
  auto load_group = [&](int i0, float (&d)[kChunks]) {
#pragma unroll
    for (int c = 0; c < kChunks; c++) {
      const int i = i0 + 32 * c + lane;
      d[c] = 0.f;
      if (i < end) d[c] = dist2_ref(qx, qy, qz, xyz[(int64_t)i * 3 + 0], xyz[(int64_t)i * 3 + 1], xyz[(int64_t)i * 3 + 2]);
    }
  };
  float nxt[kChunks];
  load_group(start, nxt);
  for (int i0 = start; i0 < end; i0 += 32 * kChunks) {
    float d2c[kChunks];
#pragma unroll
    for (int c = 0; c < kChunks; c++) d2c[c] = nxt[c];
    load_group(i0 + 32 * kChunks, nxt);  // the next group's loads travel while this one is offered to the heap
#pragma unroll
    for (int c = 0; c < kChunks; c++) {
      const int ib = i0 + 32 * c;
      unsigned mask = __ballot_sync(0xffffffffu, ib + lane < end && d2c[c] < top);
      if (mask == 0) continue;
      while (mask) {
        const int src = __ffs(mask) - 1;
        mask &= mask - 1;
        const float nd = __shfl_sync(0xffffffffu, d2c[c], src);
        if (lane == 0 && nd < bd[0]) { bd[0] = nd; bi[0] = ib + src; heap_reheap_smem(bd, bi, nsample); }
      }
      __syncwarp();
      top = bd[0];
    }
  }
  if (lane == 0) {
    for (int i = nsample - 1; i > 0; i--) {  // heap_sort :39-48
      const float tf = bd[0]; bd[0] = bd[i]; bd[i] = tf;
      const int ti = bi[0]; bi[0] = bi[i]; bi[i] = ti;
      heap_reheap_smem(bd, bi, i);
    }
  }
  __syncwarp();
  for (int i = lane; i < nsample; i += 32) {
    idx[(int64_t)q * nsample + i] = bi[i];
    dist2[(int64_t)q * nsample + i] = bd[i];
  }
}

}  // namespace geot

using namespace geot;

namespace geot {
struct KnnWs { size_t flags, seg_grid, slots, sorted, total; };
static KnnWs knn_ws_layout(int b, int n, int m) {
  auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
  KnnWs w;
  w.flags = 0;
  w.seg_grid = up((size_t)max(m, 0));
  w.slots = w.seg_grid + up((size_t)max(b, 0));
  const bool grid = n >= kKnnGridMinN;
  w.sorted = w.slots + (grid ? up(((size_t)n / kKnnGridMinN + 1) * knn_seg_ws_bytes()) : 0);
  w.total = w.sorted + (grid ? up((size_t)n * sizeof(float4)) : 0);
  return w;
}
}  // namespace geot

GEOT_API size_t geot_knnquery_workspace_bytes(int b, int n, int m) { return (m > 0 && b > 0) ? knn_ws_layout(b, n, m).total : 0; }

GEOT_API int geot_knnquery(int b, int n, int m, int nsample, const float* xyz, const float* new_xyz,
                           const int* offset, const int* new_offset, int n_per, int m_per, int* idx,
                           float* dist2, void* workspace, size_t workspace_bytes, int flags_arg, geot_stream_t stream) {
  GEOT_REQUIRE(b >= 0 && m >= 0 && n >= 0, "knnquery: bad sizes");
  GEOT_REQUIRE(nsample >= 1 && nsample <= 1000, "knnquery: nsample=%d outside [1,1000] (reference heap size, knnquery_cuda_kernel.cu:86)", nsample);
  if (b == 0 || m == 0) return GEOT_OK;
  GEOT_REQUIRE(xyz && new_xyz && idx && dist2, "knnquery: NULL pointer");
  GEOT_REQUIRE((offset == nullptr) == (new_offset == nullptr), "knnquery: offset and new_offset must both be given or both NULL");
  GEOT_REQUIRE(m_per >= 1 && b <= 65535, "knnquery: m_per=%d must bound the queries per segment; b=%d", m_per, b);
  if (offset == nullptr) GEOT_REQUIRE(n_per >= 0 && (int64_t)b * m_per == m && (int64_t)b * n_per == n, "knnquery: batched layout needs b*m_per == m and b*n_per == n");
  cudaStream_t s = as_stream(stream);
  if (nsample + 1 <= 128) {
    const KnnWs w = knn_ws_layout(b, n, m);
    if (workspace == nullptr || workspace_bytes < w.total) {
      set_error("knnquery: workspace %zu < %zu bytes", workspace_bytes, w.total);
      return GEOT_ERR_WORKSPACE;
    }
    unsigned char* base = reinterpret_cast<unsigned char*>(workspace);
    unsigned char* flags = base + w.flags;
    // grid path: only when some segment can reach kKnnGridMinN points
    const bool grid = !(flags_arg & GEOT_KNN_NO_GRID) && n >= kKnnGridMinN && (offset != nullptr || n_per >= kKnnGridMinN);
    unsigned char* seg_grid = grid ? base + w.seg_grid : nullptr;
    const int S = (nsample + 1 + 31) / 32;
    if (grid) {
      static DynSmem ds_build;
      const int smem = kKnnMaxCells * (int)sizeof(int);
      if (int rc = ds_build.need(knn_build_kernel, smem, "knnquery")) return rc;
      // points per occupied cell = a * (k+1).  Measured (self-kNN on tooth clouds, 2 x 16000 and 8 x 16384): flat between
      // 0.55 and 1.0 at k = 16; 0.55-0.7 at k = 33 (0.40 ms against 0.50 at 0.38: fewer queries need the 5x5x5 shell);
      // 0.3-0.55 at k = 64
      const float a = fminf(fmaxf(18.f / (float)(nsample + 1), 0.3f), 0.8f);
      knn_build_kernel<<<b, 1024, smem, s>>>(nsample, a, xyz, offset, n_per, seg_grid, base + w.slots, reinterpret_cast<float4*>(base + w.sorted));
      count_launch(1);
      int rc = check_launch("knn_build_kernel");
      if (rc) return rc;
    }
    dim3 grid_w((m_per + 31) / 32, b);
    const bool lex = (flags_arg & GEOT_KNN_TIES_BY_INDEX) != 0;
#define GEOT_KNN_WARP(SS, LL) knn_warp_kernel<SS, LL><<<grid_w, 1024, 0, s>>>(nsample, xyz, new_xyz, offset, new_offset, n_per, m_per, idx, dist2, flags, seg_grid)
    switch (S) {
      case 1: if (lex) GEOT_KNN_WARP(1, true); else GEOT_KNN_WARP(1, false); break;
      case 2: if (lex) GEOT_KNN_WARP(2, true); else GEOT_KNN_WARP(2, false); break;
      case 3: if (lex) GEOT_KNN_WARP(3, true); else GEOT_KNN_WARP(3, false); break;
      default: if (lex) GEOT_KNN_WARP(4, true); else GEOT_KNN_WARP(4, false); break;
    }
#undef GEOT_KNN_WARP
    count_launch(1);
    int rc = check_launch("knn_warp_kernel");
    if (rc) return rc;
    if (grid) {
      dim3 grid_g((m_per + 7) / 8, b);
      const unsigned char* slots = base + w.slots;
      const float4* sorted = reinterpret_cast<const float4*>(base + w.sorted);
#define GEOT_KNN_GRID(SS, LL) knn_grid_kernel<SS, LL><<<grid_g, 256, 0, s>>>(nsample, new_xyz, offset, new_offset, n_per, m_per, idx, dist2, flags, seg_grid, slots, sorted)
      switch (S) {
        case 1: if (lex) GEOT_KNN_GRID(1, true); else GEOT_KNN_GRID(1, false); break;
        case 2: if (lex) GEOT_KNN_GRID(2, true); else GEOT_KNN_GRID(2, false); break;
        case 3: if (lex) GEOT_KNN_GRID(3, true); else GEOT_KNN_GRID(3, false); break;
        default: if (lex) GEOT_KNN_GRID(4, true); else GEOT_KNN_GRID(4, false); break;
      }
#undef GEOT_KNN_GRID
      count_launch(1);
      rc = check_launch("knn_grid_kernel");
      if (rc) return rc;
    }
    // ties by index: the fast kernels' (distance, index) order is the answer; otherwise the queries with an exact tie
    // among their k + 1 best are replayed with the reference's heap (its order of equal distances depends on the heap's history)
    if (flags_arg & GEOT_KNN_TIES_BY_INDEX) return GEOT_OK;
    knn_heap_warp_kernel<<<(m + 7) / 8, 256, 0, s>>>(b, m, nsample, xyz, new_xyz, offset, new_offset, n_per, m_per, idx, dist2, flags);
    count_launch(1);
    return check_launch("knn_heap_warp_kernel");
  }
  knn_heap_kernel<1000><<<(m + 127) / 128, 128, 0, s>>>(b, m, nsample, xyz, new_xyz, offset, new_offset, n_per, m_per, idx, dist2, nullptr);
  count_launch(1);
  return check_launch("knn_heap_kernel<1000>");
}
