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
//
// Grid path (used when the caller passes a workspace and the ball is small against the cloud).  The scan
// above visits points in index order until nsample hits are found: ~N*nsample/H points for H hits in the
// ball (26-40 % of the cloud at the bench shape).  bq_build_kernel bins each cloud once into a uniform grid
// of cell size >= 1.001*radius (counting sort in shared memory -> sorted {x,y,z,index} + cell offsets in the
// workspace); bq_grid_kernel then tests only the 27 cells around the query (9 contiguous runs, ~3 H
// candidates) with the same dist2_ref, sets one bit per hit in a per-warp bitmap over the point indices, and
// reads the first nsample set bits back in ascending index order - the reference's "first nsample by index"
// rule, independent of the order candidates were met.  The 0.1 % cell margin keeps the 27-cell cover
// conservative under fp32 rounding of the cell arithmetic.  bq_build_kernel decides per cloud which path is
// cheaper (dense balls finish the index-order scan after a few dozen points) and each kernel skips the
// clouds of the other.
#include <atomic>

#include "common.cuh"

namespace geot {

void count_launch(int n);

struct BqHeader {  // grid path (below): per-cloud header at the start of the cloud's workspace slice
  int use_grid;
  int dx, dy, dz;
  float lox, loy, loz, inv;
  int nlev, split;  // index levels: points with index in [l*split, (l+1)*split) are binned separately, searched in order
  int pad[6];
};
constexpr int kBqHdrBytes = 64;
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
                         const float* __restrict__ xyz, int* __restrict__ idx, int aligned16,
                         const unsigned char* __restrict__ ws, size_t ws_stride) {
  __shared__ __align__(16) float tile[2][kBqTile * 3];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = blockIdx.y;
  if (ws != nullptr && reinterpret_cast<const BqHeader*>(ws + (size_t)b * ws_stride)->use_grid) return;
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

// ---------------------------------------------------------------------------------------------
// Grid path.  Workspace per cloud: [BqHeader | cell_start (kBqMaxCells+1 ints) | sorted float4 (N)].
// The reference keeps the nsample LOWEST indices in the ball, so when the ball holds several times nsample
// points only low indices matter: the points are binned in two index levels (index < N/2 first); a query
// scans the upper level only if the lower one gave fewer than nsample hits.
// ---------------------------------------------------------------------------------------------
constexpr int kBqMaxDim = 32;
constexpr int kBqMaxCells = kBqMaxDim * kBqMaxDim * kBqMaxDim;
constexpr int kBqMaxN = 65536;  // bitmap = N bits per warp
constexpr int kBqMaxLevels = 8;
constexpr int kBqMinN = 4096;    // at or below: always the index-order scan (GEOT_BQ_FORCE_GRID overrides)

__host__ __device__ inline size_t bq_cells_bytes() { return (((size_t)(kBqMaxCells + 1) * 4) + 15) & ~(size_t)15; }
static inline size_t bq_ws_stride(int N) { return kBqHdrBytes + bq_cells_bytes() + (size_t)N * 16; }

__device__ __forceinline__ int bq_cell(float p, float lo, float inv) { return __float2int_rd((p - lo) * inv); }

// CACHED: N <= 16 * 1024 - a thread keeps its 16 points in registers after the first pass (the kernel makes four passes
// over the cloud: bounding box, occupancy, histogram, scatter).
constexpr int kBqRegPts = 16;
template <bool CACHED>
__global__ void __launch_bounds__(1024, 1)
bq_build_kernel(int N, float radius, int nsample, const float* __restrict__ xyz, unsigned char* __restrict__ ws,
                size_t ws_stride, int force_grid, int force_levels) {
  extern __shared__ int bq_hist[];  // kBqMaxCells counters
  __shared__ float s_red[6][32];
  __shared__ int s_scan[32];
  __shared__ int s_nonempty;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int b = blockIdx.x;
  const float* __restrict__ src = xyz + (int64_t)b * N * 3;
  BqHeader* hdr = reinterpret_cast<BqHeader*>(ws + (size_t)b * ws_stride);
  int* cell_start = reinterpret_cast<int*>(ws + (size_t)b * ws_stride + kBqHdrBytes);
  float4* sorted = reinterpret_cast<float4*>(ws + (size_t)b * ws_stride + kBqHdrBytes + bq_cells_bytes());
  const float kInf = __int_as_float(0x7f800000);

  float lo[3] = {kInf, kInf, kInf}, hi[3] = {-kInf, -kInf, -kInf};
  float px[CACHED ? kBqRegPts : 1], py[CACHED ? kBqRegPts : 1], pz[CACHED ? kBqRegPts : 1];
  if constexpr (CACHED) {
#pragma unroll
    for (int j = 0; j < kBqRegPts; j++) {
      const int k = t + j * 1024;
      px[j] = py[j] = pz[j] = 0.f;
      if (k < N) { px[j] = src[(int64_t)k * 3]; py[j] = src[(int64_t)k * 3 + 1]; pz[j] = src[(int64_t)k * 3 + 2]; }
    }
  }
  auto for_points = [&](auto&& fn) {  // fn(index, x, y, z) for the points of this thread
    if constexpr (CACHED) {
#pragma unroll
      for (int j = 0; j < kBqRegPts; j++) {
        const int k = t + j * 1024;
        if (k < N) fn(k, px[j], py[j], pz[j]);
      }
    } else {
#pragma unroll 4
      for (int k = t; k < N; k += 1024) fn(k, src[(int64_t)k * 3], src[(int64_t)k * 3 + 1], src[(int64_t)k * 3 + 2]);
    }
  };
  for_points([&](int, float x, float y, float z) {
    lo[0] = fminf(lo[0], x); hi[0] = fmaxf(hi[0], x);
    lo[1] = fminf(lo[1], y); hi[1] = fmaxf(hi[1], y);
    lo[2] = fminf(lo[2], z); hi[2] = fmaxf(hi[2], z);
  });
#pragma unroll
  for (int a = 0; a < 3; a++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
      hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
    }
    if (lane == 0) { s_red[a][warp] = lo[a]; s_red[3 + a][warp] = hi[a]; }
  }
  if (t == 0) s_nonempty = 0;
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
  // cell edge: a hair more than the radius (fp32-safe 27-cell cover), grown if the cloud would need > 32 cells
  float cell = fmaxf(radius * 1.001f, ext / 31.5f);
  const bool ok = (cell > 0.f) && (cell < kInf) && (ext < kInf) && (ext == ext) && N > 0;
  if (!ok) {  // degenerate input (zero radius, NaN/inf coordinates): index-order scan handles it
    if (t == 0) hdr->use_grid = 0;
    return;
  }
  const float inv = 1.0f / cell;
  int dims[3];
#pragma unroll
  for (int a = 0; a < 3; a++) dims[a] = min(kBqMaxDim, max(1, bq_cell(hi[a], lo[a], inv) + 1));
  const int ncells1 = dims[0] * dims[1] * dims[2];
  // index levels (1, 2, 4 or 8): level l holds the points with index in [l*split, (l+1)*split), binned separately;
  // a query walks the levels in order and stops as soon as it has nsample hits.  The histogram is taken at the finest
  // admissible count; the count actually used follows from the occupancy and the fine levels are merged down to it.
  int maxlev = 1;
  while (maxlev < kBqMaxLevels && 2 * maxlev * ncells1 <= kBqMaxCells && N >= 2048 * maxlev) maxlev *= 2;
  if (force_levels > 0) maxlev = min(maxlev, force_levels);
  const int split_f = maxlev > 1 ? (((N + maxlev - 1) / maxlev + 31) & ~31) : N;
  auto cell_of = [&](float x, float y, float z) -> int {
    const int cx = min(max(bq_cell(x, lo[0], inv), 0), dims[0] - 1);
    const int cy = min(max(bq_cell(y, lo[1], inv), 0), dims[1] - 1);
    const int cz = min(max(bq_cell(z, lo[2], inv), 0), dims[2] - 1);
    return (cz * dims[1] + cy) * dims[0] + cx;
  };
  // k / d for k < 65536 and d a multiple of 32 up to 65536 as a multiply-high (checked exhaustively)
  auto level_of = [](int k, unsigned magic) -> int { return (int)__umulhi((unsigned)k, magic); };
  const unsigned magic_f = 0xFFFFFFFFu / (unsigned)max(split_f, 1) + 1u;
  for (int i = t; i < maxlev * ncells1; i += 1024) bq_hist[i] = 0;
  __syncthreads();
  for_points([&](int k, float x, float y, float z) {
    atomicAdd(&bq_hist[(maxlev > 1 ? level_of(k, magic_f) * ncells1 : 0) + cell_of(x, y, z)], 1);
  });
  __syncthreads();
  {  // occupied cells (any level) for the cost model and the level count
    int ne1 = 0;
    for (int i = t; i < ncells1; i += 1024) {
      int any = 0;
      for (int l = 0; l < maxlev; l++) any |= bq_hist[l * ncells1 + i];
      ne1 += any > 0;
    }
    ne1 = __reduce_add_sync(0xffffffffu, ne1);
    if (lane == 0 && ne1) atomicAdd(&s_nonempty, ne1);
  }
  __syncthreads();
  const float occ = (float)N / (float)max(s_nonempty, 1);
  const float hits = 3.f * occ;  // points of a ball as the occupancy suggests (a surface: ~pi cells of area)
  // as many levels as leave ~nsample hits in the first one (measured at 64 x 16384, r = 0.1, nsample = 32:
  // 1 / 2 / 4 / 8 levels 0.157 / 0.126 / 0.115 / 0.137 ms)
  int nlev = 1;
  while (2 * nlev <= maxlev && hits >= (float)(2 * nlev) * (float)nsample) nlev *= 2;
  if (force_levels > 0) nlev = maxlev;
  const int ratio = maxlev / nlev;
  const int split = nlev > 1 ? split_f * ratio : N;
  const int ncell = nlev * ncells1;  // sort keys: level-major, then cell
  if (ratio > 1) {  // merge `ratio` consecutive fine levels into one: sums first, then the writes
    int sums[kBqMaxCells / 1024 / 2];
    int cnt = 0;
    for (int i = t; i < ncell; i += 1024, cnt++) {
      const int l = i / ncells1, c = i - l * ncells1;
      int v = 0;
      for (int r = 0; r < ratio; r++) v += bq_hist[(l * ratio + r) * ncells1 + c];
      sums[cnt] = v;
    }
    __syncthreads();
    cnt = 0;
    for (int i = t; i < ncell; i += 1024, cnt++) bq_hist[i] = sums[cnt];
    __syncthreads();
  }
  const unsigned magic = 0xFFFFFFFFu / (unsigned)max(split, 1) + 1u;
  auto code_of = [&](float x, float y, float z, int k) -> int { return (nlev > 1 ? level_of(k, magic) * ncells1 : 0) + cell_of(x, y, z); };
  // exclusive scan over ncell counters, 1024 threads x PER consecutive entries
  {
    const int PER = (ncell + 1023) / 1024;
    const int i0 = t * PER;
    int sum = 0;
    for (int i = i0; i < min(i0 + PER, ncell); i++) sum += bq_hist[i];
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
      const int v = bq_hist[i];
      bq_hist[i] = run;
      cell_start[i] = run;
      run += v;
    }
    if (t == 0) cell_start[ncell] = N;
  }
  __syncthreads();
  for_points([&](int k, float x, float y, float z) {
    const int pos = atomicAdd(&bq_hist[code_of(x, y, z, k)], 1);
    sorted[pos] = make_float4(x, y, z, __int_as_float(k));
  });
  if (t == 0) {
    // cost model in units of "points visited by the index-order scan" (calibrated on B200, tools/bq_ab.py): the
    // 27 cells hold ~34*occ candidate slots (occ = points per occupied cell) at ~1.5x the per-point cost of the
    // scan, plus ~600 for the bitmap read-back; the scan visits ~N*nsample/H points for H ~ 3*occ hits
    // Cost model, both sides in "points visited by the index-order scan" (0.0023 ns each on B200; tools/bq_ab.py at
    // 16 x 16384 x 4096 and 64 x 16384 x 1024, r = 0.05 .. 0.2, nsample = 16 .. 64).  The balls hold ~2.5x what the
    // occupancy suggests (measured 108 / 473 / 1900 points at r = 0.05 / 0.1 / 0.2 against 45 / 177 / 819).  Scan:
    // 390 fixed + the points visited until nsample hits.  Grid: 400 fixed (bitmap clear and read-back, 9 runs) + ~6 per
    // point of an occupied cell and walked level.
    const float per_level = hits / (float)nlev;
    const float walked = fminf((float)nlev, ceilf(1.3f * (float)nsample / fmaxf(per_level, 1e-3f)));
    const float grid_cost = 400.f + 6.f * occ * walked / (float)nlev + 60.f * (walked - 1.f);
    const float scan_cost = 390.f + fminf((float)N, (float)N * (float)nsample / fmaxf(2.5f * hits, (float)nsample));
    hdr->use_grid = (force_grid || grid_cost < scan_cost) ? 1 : 0;
    hdr->dx = dims[0]; hdr->dy = dims[1]; hdr->dz = dims[2];
    hdr->lox = lo[0]; hdr->loy = lo[1]; hdr->loz = lo[2]; hdr->inv = inv;
    hdr->nlev = nlev; hdr->split = split;
  }
}

// One warp per query; 8 queries per CTA.  Dynamic smem: 8 bitmaps of wpl*32 words (wpl words per lane, power of 2).
__global__ void __launch_bounds__(256)
bq_grid_kernel(int N, int M, float radius, int nsample, const float* __restrict__ new_xyz,
               const unsigned char* __restrict__ ws, size_t ws_stride, int* __restrict__ idx, int wpl_log2) {
  extern __shared__ uint4 bq_bitmap[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = blockIdx.y;
  const BqHeader hdr = *reinterpret_cast<const BqHeader*>(ws + (size_t)b * ws_stride);
  if (!hdr.use_grid) return;
  const int q = blockIdx.x * 8 + warp;
  if (q >= M) return;
  const int* __restrict__ cell_start = reinterpret_cast<const int*>(ws + (size_t)b * ws_stride + kBqHdrBytes);
  const float4* __restrict__ sorted = reinterpret_cast<const float4*>(ws + (size_t)b * ws_stride + kBqHdrBytes + bq_cells_bytes());
  // bitmap over point indices, word w = index >> 5; step j of the read-back covers words 32j .. 32j+31
  const int wpl = 1 << wpl_log2, nvec = wpl >> 2;  // words per lane, uint4 per lane
  uint4* bm = bq_bitmap + (size_t)warp * nvec * 32;
  unsigned* bmw = reinterpret_cast<unsigned*>(bm);
  for (int j = 0; j < nvec; j++) bm[j * 32 + lane] = make_uint4(0u, 0u, 0u, 0u);

  const float* qp = new_xyz + ((int64_t)b * M + q) * 3;
  const float qx = qp[0], qy = qp[1], qz = qp[2];
  const float r2 = __fmul_rn(radius, radius);
  const int cx = bq_cell(qx, hdr.lox, hdr.inv), cy = bq_cell(qy, hdr.loy, hdr.inv), cz = bq_cell(qz, hdr.loz, hdr.inv);
  // 9 runs per level (3 x-adjacent cells are contiguous in the sorted order); lanes 0..8 fetch the runs of an even
  // level and lanes 16..24 those of the next one in the same round trip
  const int l9 = lane & 15;
  int c9 = -1, w9 = 0;  // first cell of this lane's run (within a level) and its width
  if (l9 < 9) {
    const int z = cz + l9 / 3 - 1, y = cy + l9 % 3 - 1;
    const int x0 = max(cx - 1, 0), x1 = min(cx + 1, hdr.dx - 1);
    if (z >= 0 && z < hdr.dz && y >= 0 && y < hdr.dy && x0 <= x1) { c9 = (z * hdr.dy + y) * hdr.dx + x0; w9 = x1 - x0 + 1; }
  }
  const int ncells1 = hdr.dx * hdr.dy * hdr.dz;
  auto test = [&](const float4& p) {
    const float d2 = dist2_ref(qx, qy, qz, p.x, p.y, p.z);
    if (d2 < r2) {
      const unsigned k = (unsigned)__float_as_int(p.w);
      atomicOr(bmw + (k >> 5), 1u << (k & 31u));
    }
  };
  int rs = 0, re = 0;
  for (int lev = 0; lev < hdr.nlev; lev++) {
    if (lev > 0) {  // later levels only if the earlier ones left the row short
      __syncwarp();
      int c0 = 0;
      for (int j = 0; (j << 10) < lev * hdr.split; j++) c0 += __popc(bmw[j * 32 + lane]);
      if (__reduce_add_sync(0xffffffffu, c0) >= nsample) break;
    }
    if ((lev & 1) == 0) {
      const int mine = lev + (lane >> 4);
      rs = re = 0;
      if (c9 >= 0 && mine < hdr.nlev) {
        rs = __ldg(cell_start + mine * ncells1 + c9);
        re = __ldg(cell_start + mine * ncells1 + c9 + w9);
      }
    }
#pragma unroll 1
    for (int r = 0; r < 9; r++) {
      const int s = __shfl_sync(0xffffffffu, rs, r + 16 * (lev & 1)), e = __shfl_sync(0xffffffffu, re, r + 16 * (lev & 1));
      int i = s + lane;
      for (; i + 32 < e; i += 64) {  // two candidates per lane in flight
        const float4 p0 = __ldg(sorted + i), p1 = __ldg(sorted + i + 32);
        test(p0);
        test(p1);
      }
      if (i < e) test(__ldg(sorted + i));
    }
  }
  __syncwarp();
  // read the set bits back in ascending index order, 1024 indices per step, until nsample are out
  int* row = idx + ((int64_t)b * M + q) * nsample;
  int total = 0, first = 0;
  for (int j = 0; j < wpl && total < nsample; j++) {
    unsigned word = bmw[j * 32 + lane];
    const int cnt = __popc(word);
    int inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += y;
    }
    int slot = total + inc - cnt;
    const int base = (j * 32 + lane) << 5;
    if (total == 0) {  // the very first hit fills the unused slots (ball_query_gpu.cu:38-42)
      const unsigned have = __ballot_sync(0xffffffffu, word != 0u);
      if (have) first = __shfl_sync(0xffffffffu, base + __ffs(word) - 1, __ffs(have) - 1);
    }
    while (word && slot < nsample) {
      row[slot++] = base + __ffs(word) - 1;
      word &= word - 1;
    }
    total += __shfl_sync(0xffffffffu, inc, 31);
  }
  for (int l = min(total, nsample) + lane; l < nsample; l += 32) row[l] = first;
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

GEOT_API size_t geot_ball_query_workspace_bytes(int B, int N) {
  if (B <= 0 || N <= 0 || N > kBqMaxN) return 0;
  return (size_t)B * bq_ws_stride(N);
}

GEOT_API int geot_ball_query(int B, int N, int M, float radius, int nsample, const float* new_xyz,
                             const float* xyz, int* idx, void* workspace, size_t workspace_bytes,
                             int flags, geot_stream_t stream) {
  GEOT_REQUIRE(B >= 0 && N >= 0 && M >= 0 && nsample >= 0, "ball_query: bad sizes");
  const bool build_only = (flags & GEOT_BQ_BUILD_ONLY) != 0, prebuilt = (flags & GEOT_BQ_PREBUILT) != 0;
  GEOT_REQUIRE(!(build_only && prebuilt), "ball_query: GEOT_BQ_BUILD_ONLY and GEOT_BQ_PREBUILT exclude each other");
  if (B == 0 || nsample == 0 || (M == 0 && !build_only)) return GEOT_OK;
  GEOT_REQUIRE(xyz && (build_only || (new_xyz && idx)), "ball_query: NULL pointer");
  GEOT_REQUIRE(!(build_only || prebuilt) || workspace != nullptr, "ball_query: GEOT_BQ_BUILD_ONLY / GEOT_BQ_PREBUILT need the workspace");
  GEOT_REQUIRE(B <= 65535, "ball_query: B=%d exceeds grid.y", B);
  cudaStream_t s = as_stream(stream);
  const unsigned char* ws = nullptr;
  size_t stride = 0;
  // small clouds: the build (~20 us per launch) costs more than the scan saves (measured at N = 4096, r = 0.2 / 0.4:
  // scan 0.034 / 0.013 ms, grid 0.045 / 0.046 ms)
  if (workspace != nullptr && N > 0 && (N > kBqMinN || (flags & GEOT_BQ_FORCE_GRID)) && N <= kBqMaxN) {
    stride = bq_ws_stride(N);
    if (workspace_bytes < (size_t)B * stride) {
      set_error("ball_query: workspace %zu < %zu bytes", workspace_bytes, (size_t)B * stride);
      return GEOT_ERR_WORKSPACE;
    }
    ws = reinterpret_cast<const unsigned char*>(workspace);
    static DynSmem ds_build, ds_build_c, ds_grid;
    if (int rc = ds_grid.need(bq_grid_kernel, 8 * (kBqMaxN / 8), "ball_query")) return rc;
    unsigned char* wsw = reinterpret_cast<unsigned char*>(workspace);
    const int force_grid = (flags & GEOT_BQ_FORCE_GRID) ? 1 : 0, force_levels = GEOT_BQ_LEVELS_OF(flags);
    int rc = GEOT_OK;
    if (!prebuilt) {  // the binning depends on (xyz, radius, nsample) only: a caller may run it ahead of the queries
      if (N <= kBqRegPts * 1024) {
        if (int rc2 = ds_build_c.need(bq_build_kernel<true>, kBqMaxCells * 4, "ball_query")) return rc2;
        bq_build_kernel<true><<<B, 1024, kBqMaxCells * 4, s>>>(N, radius, nsample, xyz, wsw, stride, force_grid, force_levels);
      } else {
        if (int rc2 = ds_build.need(bq_build_kernel<false>, kBqMaxCells * 4, "ball_query")) return rc2;
        bq_build_kernel<false><<<B, 1024, kBqMaxCells * 4, s>>>(N, radius, nsample, xyz, wsw, stride, force_grid, force_levels);
      }
      count_launch(1);
      rc = check_launch("bq_build_kernel");
      if (rc) return rc;
    }
    if (build_only) return GEOT_OK;
    int wpl_log2 = 2;  // words per lane: power of two >= N/1024, at least 4
    while ((1024 << wpl_log2) < N) wpl_log2++;
    dim3 gq((M + 7) / 8, B);
    bq_grid_kernel<<<gq, 256, (size_t)8 * (128u << wpl_log2), s>>>(N, M, radius, nsample, new_xyz, ws, stride, idx, wpl_log2);
    count_launch(1);
    rc = check_launch("bq_grid_kernel");
    if (rc) return rc;
  }
  if (build_only) return GEOT_OK;  // a cloud size the grid path does not take: nothing to prepare
  const int aligned16 = (((uintptr_t)xyz & 15) == 0) && ((N * 3) % 4 == 0);
  dim3 grid((M + 31) / 32, B);
  ball_query_shared_kernel<<<grid, 1024, 0, s>>>(N, M, radius, nsample, new_xyz, xyz, idx, aligned16, ws, stride);
  count_launch(1);
  return check_launch("ball_query_shared_kernel");
}

// ---------------------------------------------------------------------------------------------
// Allocation-free variant: the grid-binning scratch lives in a file-scope device array instead of a caller
// workspace (for callers that cannot allocate: CUDA graph capture, serving loops).  No counterpart in the
// reference.  Every call on a device shares the one array: calls that overlap on that GPU race on it.
// Nothing in it has to start at zero: bq_build_kernel rewrites the header, the cell starts and the sorted
// points of every cloud it is given before bq_grid_kernel / ball_query_shared_kernel read them, and the
// counters that do start at zero live in shared memory.
// ---------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_bq_static_scratch[GEOT_BALL_QUERY_STATIC_SCRATCH_BYTES];

// In host code `g_bq_static_scratch` is only a handle for the runtime; the device address comes from
// cudaGetSymbolAddress, looked up once per device with that device current.
static int bq_static_scratch_address(unsigned char** out) {
  static std::atomic<unsigned char*> per_device[64];
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) {
    set_error("ball_query_static: cudaGetDevice: %s", cudaGetErrorString(e));
    return GEOT_ERR_CUDA;
  }
  const bool cached = dev >= 0 && dev < 64;
  unsigned char* p = cached ? per_device[dev].load(std::memory_order_acquire) : nullptr;
  if (p == nullptr) {
    void* sym = nullptr;
    e = cudaGetSymbolAddress(&sym, g_bq_static_scratch);
    if (e != cudaSuccess || sym == nullptr) {
      set_error("ball_query_static: cudaGetSymbolAddress(scratch) on device %d: %s", dev, cudaGetErrorString(e));
      return GEOT_ERR_CUDA;
    }
    p = static_cast<unsigned char*>(sym);
    if (cached) per_device[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return GEOT_OK;
}

GEOT_API int geot_ball_query_static(int B, int N, int M, float radius, int nsample, const float* new_xyz,
                                    const float* xyz, int* idx, int flags, geot_stream_t stream) {
  GEOT_REQUIRE(B >= 0 && N >= 0 && M >= 0 && nsample >= 0, "ball_query_static: bad sizes");
  if (B == 0 || M == 0 || nsample == 0) return GEOT_OK;
  GEOT_REQUIRE(N >= 1, "ball_query_static: N=%d (need at least one point per cloud)", N);
  if (N > kBqMaxN) {
    set_error("ball_query_static: N=%d exceeds the binned path's limit of %d points per cloud", N, kBqMaxN);
    return GEOT_ERR_WORKSPACE;
  }
  const size_t need = (size_t)B * bq_ws_stride(N);
  if (need > (size_t)GEOT_BALL_QUERY_STATIC_SCRATCH_BYTES) {
    set_error("ball_query_static: B=%d x N=%d needs %zu bytes of scratch, the built-in array holds %zu "
              "(use geot_ball_query with a caller workspace)",
              B, N, need, (size_t)GEOT_BALL_QUERY_STATIC_SCRATCH_BYTES);
    return GEOT_ERR_WORKSPACE;
  }
  unsigned char* scratch = nullptr;
  if (int rc = bq_static_scratch_address(&scratch)) return rc;
  return geot_ball_query(B, N, M, radius, nsample, new_xyz, xyz, idx, scratch, (size_t)GEOT_BALL_QUERY_STATIC_SCRATCH_BYTES,
                         flags, stream);
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
