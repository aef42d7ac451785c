// geot_b200/csrc/interp.cu — three_nn and three_interpolate (batched, channel-major) for sm_100a.
//
// Replaces
//   openpoints/cpp/pointnet2_batch/src/interpolate_gpu.cu:16-59   three_nn (bit-exact idx and dist2)
//   openpoints/cpp/pointnet2_batch/src/interpolate_gpu.cu:84-104  three_interpolate fwd
//   openpoints/cpp/pointnet2_batch/src/interpolate_gpu.cu:127-149 three_interpolate bwd (3 float atomics)
//   pointnet2/_ext_src/src/interpolate_gpu.cu:12-62,75-104,119-146 (same results)
//
// three_nn: FP32-ALU bound.  The reference streams all m known points from L2 once per THREAD; here a
// CTA stages tiles of the known cloud in shared memory (float4-padded, one broadcast LDS.128 per
// point) and each thread keeps two unknown points in flight so every staged point feeds two distance
// evaluations.  The reference compares a float d against double-typed bests initialised to 1e40; all
// values involved are floats or +inf after the first insertions, so float bests initialised to +inf
// give identical decisions and identical outputs (float(1e40) == +inf for the m < 3 slots).
//
// three_interpolate fwd: HBM-bound on the (B,C,n) write; idx/weight (24 B per target) are loaded once
// per thread and reused across a chunk of channels, sources come from an L1/L2-resident (m) row.
// The backward is the deterministic segmented reduction of group.cu over a scatter plan of idx.
#include <algorithm>
#include <cstdlib>
#include "common.cuh"

namespace geot {

void count_launch(int n);
bool direct_kernels_forced();  // geot_group_points_config(1) (group.cu)
bool staged_kernels_forced();  // geot_group_points_config(2)
bool rows4_disabled();         // geot_group_points_config(3)
bool three_nn_queue_off();     // geot_group_points_config(14)
bool persistent_interp_disabled();  // geot_group_points_config(4) (or 3)
int three_interpolate_grad_impl(int B, int C, int n, int m, const float* grad_out, const float* weight,
                                const void* plan, float* grad_points, cudaStream_t s);

constexpr int kNnTile = 2048;    // known points per shared tile (32 KB as float4)
constexpr int kNnThreads = 256;
constexpr int kNnPerThread = 2;  // unknown points per thread

struct Top3 {
  float d1, d2, d3;
  int i1, i2, i3;
};
__device__ __forceinline__ void top3_push(Top3& t, float d, int k) {
  // interpolate_gpu.cu:44-55 — strict '<' keeps the lower index on ties.
  if (d < t.d3) {
    if (d < t.d1) {
      t.d3 = t.d2; t.i3 = t.i2; t.d2 = t.d1; t.i2 = t.i1; t.d1 = d; t.i1 = k;
    } else if (d < t.d2) {
      t.d3 = t.d2; t.i3 = t.i2; t.d2 = d; t.i2 = k;
    } else {
      t.d3 = d; t.i3 = k;
    }
  }
}

// grid (ceil(n / (256*2)), B)
__global__ void __launch_bounds__(kNnThreads)
three_nn_kernel(int n, int m, const float* __restrict__ unknown, const float* __restrict__ known,
                float* __restrict__ dist2, int* __restrict__ idx) {
  __shared__ float4 tile[kNnTile];
  const int b = blockIdx.y;
  const float* __restrict__ kn = known + (int64_t)b * m * 3;
  const float inf = __int_as_float(0x7f800000);

  int p[kNnPerThread];
  float ux[kNnPerThread], uy[kNnPerThread], uz[kNnPerThread];
  Top3 best[kNnPerThread];
#pragma unroll
  for (int u = 0; u < kNnPerThread; u++) {
    p[u] = (blockIdx.x * kNnPerThread + u) * kNnThreads + threadIdx.x;
    const int pc = min(p[u], n - 1);
    const float* up = unknown + ((int64_t)b * n + pc) * 3;
    ux[u] = up[0]; uy[u] = up[1]; uz[u] = up[2];
    best[u] = Top3{inf, inf, inf, 0, 0, 0};
  }

  for (int base = 0; base < m; base += kNnTile) {
    const int cnt = min(kNnTile, m - base);
    __syncthreads();
    for (int i = threadIdx.x; i < cnt; i += kNnThreads) {
      const float* s = kn + (int64_t)(base + i) * 3;
      tile[i] = make_float4(s[0], s[1], s[2], 0.f);
    }
    __syncthreads();
#pragma unroll 4
    for (int i = 0; i < cnt; i++) {
      const float4 k4 = tile[i];
#pragma unroll
      for (int u = 0; u < kNnPerThread; u++) {
        const float d = dist2_ref(ux[u], uy[u], uz[u], k4.x, k4.y, k4.z);
        top3_push(best[u], d, base + i);
      }
    }
  }
#pragma unroll
  for (int u = 0; u < kNnPerThread; u++) {
    if (p[u] < n) {
      float* od = dist2 + ((int64_t)b * n + p[u]) * 3;
      int* oi = idx + ((int64_t)b * n + p[u]) * 3;
      od[0] = best[u].d1; od[1] = best[u].d2; od[2] = best[u].d3;
      oi[0] = best[u].i1; oi[1] = best[u].i2; oi[2] = best[u].i3;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// three_nn over a uniform grid (default for 64 <= m <= kNgMaxM = 9216).  The brute-force kernel above evaluates all m
// known points per unknown point (n*m distances per cloud).  The known cloud is small (it is the SAMPLED level
// of a feature-propagation stage), so every CTA bins it into shared memory itself - bounding box, cell size for
// ~1.5 points per occupied cell (measured, one re-binning if the first guess is off), counting sort of
// {x, y, z, index} - and each of its 1024 threads then searches the 3x3x3 cells around its unknown point, the
// 5x5x5 shell if the third-best distance does not yet beat the distance to the searched block's open faces, and
// the whole cloud otherwise.  Selection is by (distance, index) lexicographic order, which is exactly what the
// reference's strict-'<' scan in ascending index order produces (interpolate_gpu.cu:44-55), so the result does
// not depend on the order candidates are visited in: bit-identical idx and dist2, no fallback pass.
// ---------------------------------------------------------------------------------------------
constexpr int kNgMaxM = 9216;
constexpr int kNgMaxDim = 64;      // cells per axis
constexpr int kNgMaxCells = 8000;  // cells in total (one int table of this size lives in shared memory)
constexpr int kNgThreads = 1024;
constexpr size_t kNgSmemMax = 226 * 1024;  // dynamic shared memory the grid kernel may ask for
constexpr int kNgWarpQueue = 64;           // slots of a warp's ring-2 queue

__device__ __forceinline__ bool lex_lt(float d, int k, float D, int K) { return d < D || (d == D && k < K); }
__device__ __forceinline__ void top3_push_lex(Top3& t, float d, int k) {
  if (lex_lt(d, k, t.d3, t.i3)) {  // rare after the first few candidates; the body is select-only
    const bool l1 = lex_lt(d, k, t.d1, t.i1), l2 = lex_lt(d, k, t.d2, t.i2);
    t.d3 = l2 ? t.d2 : d; t.i3 = l2 ? t.i2 : k;
    t.d2 = l1 ? t.d1 : (l2 ? d : t.d2); t.i2 = l1 ? t.i1 : (l2 ? k : t.i2);
    t.d1 = l1 ? d : t.d1; t.i1 = l1 ? k : t.i1;
  }
}
__device__ __forceinline__ int ng_cell(float p, float lo, float inv) { return __float2int_rd((p - lo) * inv); }

__global__ void __launch_bounds__(kNgThreads, 1)
three_nn_grid_kernel(int n, int m, int m_total, int base, int accumulate, int use_queue, const float* __restrict__ unknown,
                     const float* __restrict__ known, float* __restrict__ dist2, int* __restrict__ idx) {
  // m = points of THIS part of the known cloud (known[b, base : base + m]); m_total = the cloud's stride.  Known clouds
  // longer than kNgMaxM are searched part by part: a later part starts from the top-3 the earlier parts left in
  // dist2 / idx (accumulate != 0).  Selection is lexicographic in (distance, GLOBAL index), so the top-3 of the union
  // does not depend on how the cloud was cut, and the ring termination bound stays valid with a pre-filled top-3.
  extern __shared__ __align__(16) unsigned char ng_smem[];
  __shared__ float s_red[6][32];
  __shared__ int s_scan[32];
  __shared__ int s_occupied;
  float4* sorted = reinterpret_cast<float4*>(ng_smem);               // m entries
  int* cells = reinterpret_cast<int*>(ng_smem + (size_t)m * 16);     // kNgMaxCells + 1: histogram -> cursor
  unsigned* runs = reinterpret_cast<unsigned*>(cells + kNgMaxCells + 1);  // 9 x kNgThreads: {start:16 | length:16}
  unsigned* queue = runs + 9 * kNgThreads;                            // per warp 7 x kNgWarpQueue: queries waiting for ring 2 {i, d1..d3, i1..i3}
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int b = blockIdx.y;
  const float* __restrict__ kn = known + ((int64_t)b * m_total + base) * 3;
  const float kInf = __int_as_float(0x7f800000);

  // ---- bin the known cloud ----
  float lo[3] = {kInf, kInf, kInf}, hi[3] = {-kInf, -kInf, -kInf};
  for (int k = t; k < m; k += kNgThreads) {
#pragma unroll
    for (int a = 0; a < 3; a++) {
      const float v = __ldg(kn + (int64_t)k * 3 + a);
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
  // degenerate cloud (all points equal, non-finite coordinates): one cell, every query scans everything
  const bool flat = !(ext > 0.f) || !(ext < kInf);
  const float per_cell = 1.5f;
  int G = flat ? 1 : min(kNgMaxDim, max(2, (int)ceilf(sqrtf((float)m / per_cell))));
  float cell = 1.f, inv = 0.f;
  int dims[3] = {1, 1, 1}, ncell = 1;
  bool capped = false;
  auto code_of = [&](float x, float y, float z) -> int {
    const int cx = min(max(ng_cell(x, lo[0], inv), 0), dims[0] - 1);
    const int cy = min(max(ng_cell(y, lo[1], inv), 0), dims[1] - 1);
    const int cz = min(max(ng_cell(z, lo[2], inv), 0), dims[2] - 1);
    return (cz * dims[1] + cy) * dims[0] + cx;
  };
#pragma unroll 1
  for (int pass = 0; pass < 2; pass++) {
    if (!flat) {
      capped = false;
      for (;;) {  // G cells along the longest axis; shrink until the whole grid fits the tables
        cell = ext / ((float)G - 0.5f);
        inv = 1.0f / cell;
#pragma unroll
        for (int a = 0; a < 3; a++) dims[a] = min(kNgMaxDim, max(1, ng_cell(hi[a], lo[a], inv) + 1));
        ncell = dims[0] * dims[1] * dims[2];
        if (ncell <= kNgMaxCells || G <= 2) break;
        G = max(2, min(G - 1, (int)((float)G * cbrtf((float)kNgMaxCells / (float)ncell))));
        capped = true;
      }
    }
    if (t == 0) s_occupied = 0;
    for (int i = t; i < ncell; i += kNgThreads) cells[i] = 0;
    __syncthreads();
    for (int k = t; k < m; k += kNgThreads)
      atomicAdd(&cells[flat ? 0 : code_of(__ldg(kn + (int64_t)k * 3), __ldg(kn + (int64_t)k * 3 + 1), __ldg(kn + (int64_t)k * 3 + 2))], 1);
    __syncthreads();
    if (pass == 1 || flat) break;
    int occ_cnt = 0;
    for (int i = t; i < ncell; i += kNgThreads) occ_cnt += cells[i] != 0;
    occ_cnt = __reduce_add_sync(0xffffffffu, occ_cnt);
    if (lane == 0 && occ_cnt) atomicAdd(&s_occupied, occ_cnt);
    __syncthreads();
    const float occ = (float)m / (float)max(s_occupied, 1);
    __syncthreads();
    if (occ >= 0.75f * per_cell && occ <= 1.6f * per_cell) break;
    const int Gn = min(kNgMaxDim, max(2, (int)rintf((float)G * __powf(occ / per_cell, 0.4f))));
    if (Gn == G || (capped && Gn > G)) break;
    G = Gn;
  }
  {  // exclusive scan of the histogram: cells = scatter cursor; after the scatter cells[i] = end of cell i = start of cell i+1
    const int PER = (ncell + kNgThreads - 1) / kNgThreads;
    const int i0 = t * PER;
    int sum = 0;
    for (int i = i0; i < min(i0 + PER, ncell); i++) sum += cells[i];
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
      const int v = cells[i];
      cells[i] = run;
      run += v;
    }
  }
  __syncthreads();
  for (int k = t; k < m; k += kNgThreads) {
    const float x = __ldg(kn + (int64_t)k * 3), y = __ldg(kn + (int64_t)k * 3 + 1), z = __ldg(kn + (int64_t)k * 3 + 2);
    sorted[atomicAdd(&cells[flat ? 0 : code_of(x, y, z)], 1)] = make_float4(x, y, z, __int_as_float(k + base));
  }
  __syncthreads();
  // ---- one unknown point per thread ----
  // The 3x3x3 block settles about 70 % of the queries on a surface cloud; the rest need the 5x5x5 shell, which costs
  // three times the first ring.  Left in place, nearly every warp has such a lane and walks the shell at a third of
  // its width.  So a query that is not settled after ring 1 parks its {index, top-3} in its WARP's shared-memory queue
  // and whenever that holds 32 entries (and at the end) the warp finishes them, one per lane: no CTA barrier, warps keep
  // running independently.  Long known clouds that leave no room for the queues finish every query in place.  A
  // query's result does not depend on who finishes it or when.
  auto load_first = [&](int i) -> Top3 {
    if (!accumulate) return Top3{kInf, kInf, kInf, 0, 0, 0};
    const float* pd = dist2 + ((int64_t)b * n + i) * 3;
    const int* pi = idx + ((int64_t)b * n + i) * 3;
    return Top3{pd[0], pd[1], pd[2], pi[0], pi[1], pi[2]};
  };
  auto write_out = [&](int i, const Top3& best) {
    float* od = dist2 + ((int64_t)b * n + i) * 3;
    int* oi = idx + ((int64_t)b * n + i) * 3;
    od[0] = best.d1; od[1] = best.d2; od[2] = best.d3;
    oi[0] = best.i1; oi[1] = best.i2; oi[2] = best.i3;
  };
  // phase 0: ring 1 from `best` (false = not settled yet; degenerate grids scan everything and are settled);
  // phase 1: ring 2 from the state ring 1 left, then the whole cloud if that does not settle it either
  auto search = [&](int i, int phase, Top3& best) -> bool {
    const float* up = unknown + ((int64_t)b * n + i) * 3;
    const float qx = up[0], qy = up[1], qz = up[2];
    auto visit = [&](int p) {
      const float4 v = sorted[p];
      top3_push_lex(best, dist2_ref(qx, qy, qz, v.x, v.y, v.z), __float_as_int(v.w));
    };
    bool done = false;
    if (!flat) {
      const int dx = dims[0], dy = dims[1], dz = dims[2];
      const float ux = (qx - lo[0]) * inv, uy = (qy - lo[1]) * inv, uz = (qz - lo[2]) * inv;
      const int cx = (int)fminf(fmaxf(floorf(ux), -4.f), (float)dx + 3.f);
      const int cy = (int)fminf(fmaxf(floorf(uy), -4.f), (float)dy + 3.f);
      const int cz = (int)fminf(fmaxf(floorf(uz), -4.f), (float)dz + 3.f);
      // squared distance below which no point outside the (2R+1)^3 block can lie (cell units, margin for the
      // fp32 binning arithmetic; faces at the grid boundary have nothing behind them)
      auto safe_d2 = [&](int R) -> float {
        float rho = kInf;
        if (cx - R > 0) rho = fminf(rho, ux - (float)(cx - R));
        if (cx + R + 1 < dx) rho = fminf(rho, (float)(cx + R + 1) - ux);
        if (cy - R > 0) rho = fminf(rho, uy - (float)(cy - R));
        if (cy + R + 1 < dy) rho = fminf(rho, (float)(cy + R + 1) - uy);
        if (cz - R > 0) rho = fminf(rho, uz - (float)(cz - R));
        if (cz + R + 1 < dz) rho = fminf(rho, (float)(cz + R + 1) - uz);
        rho = fmaxf(rho - 2e-4f * (1.f + fabsf(ux) + fabsf(uy) + fabsf(uz)) * 0.03125f - 2e-4f, 0.f) * cell;
        return rho * rho * 0.9999f;
      };
      auto run_of = [&](int xa, int xb, int y, int z, int& p, int& e) {
        p = 0; e = 0;
        if (y < 0 || y >= dy || z < 0 || z >= dz) return;
        xa = max(xa, 0); xb = min(xb, dx - 1);
        if (xa > xb) return;
        const int row = (z * dy + y) * dx;
        p = (row + xa) ? cells[row + xa - 1] : 0; e = cells[row + xb];
      };
      if (phase == 0) {
        // ring 1: the non-empty x-runs of the 3x3 rows go to a per-thread list in shared memory (converged code),
        // then they are walked as ONE flat candidate sequence: a warp's cost is its longest list, not the sum
        // over rows of the longest run
        int nr = 0;
#pragma unroll
        for (int r = 0; r < 9; r++) {
          int p, e;
          run_of(cx - 1, cx + 1, cy + r % 3 - 1, cz + r / 3 - 1, p, e);
          if (e > p) { runs[nr * kNgThreads + t] = ((unsigned)p << 16) | (unsigned)(e - p); nr++; }
        }
        int r = 0, p = 0, e = 0;
        for (;;) {
          if (p >= e) {
            if (r >= nr) break;
            const unsigned u = runs[r * kNgThreads + t];
            r++;
            p = (int)(u >> 16); e = p + (int)(u & 0xffffu);
          }
          visit(p++);
        }
        return best.d3 < safe_d2(1);
      }
      {  // ring 2: border rows are full runs, interior rows contribute their two end cells
        int r = 0, p = 0, e = 0;
        for (;;) {
          if (p < e) { visit(p++); continue; }
          if (r == 50) break;
          const int row = r >> 1, half = r & 1;
          const int oy = row % 5 - 2, oz = row / 5 - 2;
          const bool border = oy == -2 || oy == 2 || oz == -2 || oz == 2;
          if (border) { if (half == 0) run_of(cx - 2, cx + 2, cy + oy, cz + oz, p, e); }
          else run_of(half ? cx + 2 : cx - 2, half ? cx + 2 : cx - 2, cy + oy, cz + oz, p, e);
          r++;
        }
        done = best.d3 < safe_d2(2);
      }
    }
    if (!done) {  // far from the known cloud / degenerate grid: every known point
      best = load_first(i);
      for (int p = 0; p < m; p++) visit(p);
    }
    return true;
  };
  const int stride = gridDim.x * kNgThreads;
  const int iters = (n + stride - 1) / stride;  // the same for every lane of a warp: the loop holds warp-wide votes
  unsigned* wq = queue + (size_t)warp * 7 * kNgWarpQueue;  // this warp's queue, one array of kNgWarpQueue slots per field
  int wqn = 0;                                             // entries in it (warp-uniform)
  auto finish32 = [&](int first_slot, int count) {  // lanes 0..count-1 finish one queued query each
    if (lane < count) {
      const int e = first_slot + lane;
      const int qi = (int)wq[e];
      Top3 best = Top3{__uint_as_float(wq[kNgWarpQueue + e]), __uint_as_float(wq[2 * kNgWarpQueue + e]), __uint_as_float(wq[3 * kNgWarpQueue + e]),
                       (int)wq[4 * kNgWarpQueue + e], (int)wq[5 * kNgWarpQueue + e], (int)wq[6 * kNgWarpQueue + e]};
      search(qi, 1, best);
      write_out(qi, best);
    }
  };
#pragma unroll 1
  for (int it = 0; it < iters; it++) {
    const int i = it * stride + blockIdx.x * kNgThreads + t;
    bool parked = false;
    Top3 best = Top3{kInf, kInf, kInf, 0, 0, 0};
    if (i < n) {
      best = load_first(i);
      if (search(i, 0, best)) write_out(i, best);
      else if (use_queue) parked = true;
      else { search(i, 1, best); write_out(i, best); }
    }
    if (use_queue) {
      const unsigned mask = __ballot_sync(0xffffffffu, parked);
      if (parked) {
        const int e = wqn + __popc(mask & ((1u << lane) - 1u));
        wq[e] = (unsigned)i;
        wq[kNgWarpQueue + e] = __float_as_uint(best.d1); wq[2 * kNgWarpQueue + e] = __float_as_uint(best.d2); wq[3 * kNgWarpQueue + e] = __float_as_uint(best.d3);
        wq[4 * kNgWarpQueue + e] = (unsigned)best.i1; wq[5 * kNgWarpQueue + e] = (unsigned)best.i2; wq[6 * kNgWarpQueue + e] = (unsigned)best.i3;
      }
      wqn += __popc(mask);
      __syncwarp();
      if (wqn >= 32) {  // a full warp's worth: take the newest 32 (at most 31 stay behind, so 64 slots always suffice)
        wqn -= 32;
        finish32(wqn, 32);
        __syncwarp();
      }
    }
  }
  if (use_queue && wqn > 0) finish32(0, wqn);
}

constexpr int kTiCCH = 16;
// grid (ceil(n/256), ceil(C/CCH), B): thread = one target j, CCH channels.
__global__ void __launch_bounds__(256)
three_interpolate_kernel(int C, int m, int n, const float* __restrict__ points,
                         const int* __restrict__ idx, const float* __restrict__ weight,
                         float* __restrict__ out) {
  const int b = blockIdx.z;
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const int c0 = blockIdx.y * kTiCCH, c1 = min(C, c0 + kTiCCH);
  const int* id = idx + ((int64_t)b * n + j) * 3;
  const float* w = weight + ((int64_t)b * n + j) * 3;
  const int i0 = __ldg(id), i1 = __ldg(id + 1), i2 = __ldg(id + 2);
  const float w0 = __ldg(w), w1 = __ldg(w + 1), w2 = __ldg(w + 2);
  for (int c = c0; c < c1; c++) {
    const float* __restrict__ row = points + ((int64_t)b * C + c) * m;
    // reference SASS order: FMUL w1*p1 ; FFMA w0*p0 + . ; FFMA w2*p2 + .
    const float v = __fmaf_rn(w2, __ldg(row + i2), __fmaf_rn(w0, __ldg(row + i0), __fmul_rn(w1, __ldg(row + i1))));
    __stcs(out + ((int64_t)b * C + c) * n + j, v);
  }
}

// Row-staged variant (default): one CTA stages G consecutive feature rows of one cloud (G*m floats, one bulk
// async copy) and produces G x n_tile outputs: idx/weight are read once per target and re-used over the G
// channels, the 3 gathers per output hit shared memory instead of pulling 32-byte sectors through L1/L2.
__device__ __forceinline__ void ti_mbar_init(uint64_t* bar) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void ti_bulk_load(void* dst, const void* src, unsigned bytes, uint64_t* bar) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void ti_mbar_wait(uint64_t* bar) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n"
      "@!p bra WAIT_%=;\n"
      "}\n" ::"r"(smem_u32(bar)) : "memory");
}

constexpr int kTiThreads = 512;
// grid (ceil(n/n_tile), ceil(C/G), B), kTiThreads threads.
__global__ void __launch_bounds__(kTiThreads)
three_interpolate_rows_kernel(int C, int m, int n, int G, int n_tile, const float* __restrict__ points,
                              const int* __restrict__ idx, const float* __restrict__ weight,
                              float* __restrict__ out) {
  extern __shared__ __align__(16) float ti_rows[];
  __shared__ uint64_t bar;
  const int t = threadIdx.x, b = blockIdx.z;
  const int c0 = blockIdx.y * G, gc = min(G, C - c0);
  if (t == 0) {
    ti_mbar_init(&bar);
    ti_bulk_load(ti_rows, points + ((int64_t)b * C + c0) * m, (unsigned)gc * (unsigned)m * 4u, &bar);
  }
  const int j0 = blockIdx.x * n_tile, j1 = min(n, j0 + n_tile);
  __syncthreads();  // the barrier init is visible to the waiters
  int j = j0 + t;
  int i0 = 0, i1 = 0, i2 = 0;
  float w0 = 0.f, w1 = 0.f, w2 = 0.f;
  auto fetch = [&](int jj) {
    const int* id = idx + ((int64_t)b * n + jj) * 3;
    const float* w = weight + ((int64_t)b * n + jj) * 3;
    i0 = __ldg(id); i1 = __ldg(id + 1); i2 = __ldg(id + 2);
    w0 = __ldg(w); w1 = __ldg(w + 1); w2 = __ldg(w + 2);
  };
  if (j < j1) fetch(j);
  ti_mbar_wait(&bar);
  for (; j < j1; j += kTiThreads) {
    const int a0 = i0, a1 = i1, a2 = i2;
    const float v0 = w0, v1 = w1, v2 = w2;
    if (j + kTiThreads < j1) fetch(j + kTiThreads);  // next target's indices/weights while this one is computed
    float* o = out + ((int64_t)b * C + c0) * n + j;
#pragma unroll 4
    for (int g = 0; g < gc; g++) {
      const float* row = ti_rows + g * m;
      // reference SASS order: FMUL w1*p1 ; FFMA w0*p0 + . ; FFMA w2*p2 + .
      const float v = __fmaf_rn(v2, row[a2], __fmaf_rn(v0, row[a0], __fmul_rn(v1, row[a1])));
      __stcs(o + (int64_t)g * n, v);
    }
  }
}

// Channel-interleaved variant (default when n % 4 == 0 and at least 4 rows fit): the row-staged kernel above is bound
// by shared-memory wavefronts - three 4-byte gathers per output, ~3.5-way bank conflicts on random indices (0.8 of the
// 1 wavefront/cycle/SM limit at 49 % of HBM peak).  Here four feature rows are interleaved in shared memory as float4
// per source point, so ONE 16-byte gather per neighbour serves four channels (a quarter-warp per wavefront: ~2.3-way
// conflicts -> 0.22 instead of 0.33 wavefronts per output), a thread owns 4 consecutive targets (idx / weight arrive as
// three 16-byte loads each) and every result row leaves as 16-byte streaming stores.  Same FMA order: same bits.
constexpr int kTi4Threads = 256;
// grid (ceil(n / n_tile), ceil(C / G), B); G a multiple of 4; smem = (G/4) * m float4
__global__ void __launch_bounds__(kTi4Threads)
three_interpolate_rows4_kernel(int C, int m, int n, int G, int n_tile, const float* __restrict__ points,
                               const int* __restrict__ idx, const float* __restrict__ weight, float* __restrict__ out) {
  extern __shared__ __align__(16) float4 ti4_rows[];
  const int t = threadIdx.x, b = blockIdx.z;
  const int c0 = blockIdx.y * G, gc = min(G, C - c0);
  const int ng = (gc + 3) >> 2;
  for (int g4 = 0; g4 < ng; g4++) {
    const float* r0 = points + ((int64_t)b * C + c0 + 4 * g4) * m;
    const int live = min(4, gc - 4 * g4);
#pragma unroll 4
    for (int i = t; i < m; i += kTi4Threads) {  // unrolled: 16 row loads in flight per thread before the first store
      float4 v;
      v.x = __ldg(r0 + i);
      v.y = live > 1 ? __ldg(r0 + (int64_t)m + i) : 0.f;
      v.z = live > 2 ? __ldg(r0 + (int64_t)2 * m + i) : 0.f;
      v.w = live > 3 ? __ldg(r0 + (int64_t)3 * m + i) : 0.f;
      ti4_rows[(size_t)g4 * m + i] = v;
    }
  }
  __syncthreads();
  const int j0 = blockIdx.x * n_tile, j1 = min(n, j0 + n_tile);  // multiples of 4
  int4 ia, ib, ic;
  float4 wa, wb, wc;
  auto fetch = [&](int jj) {
    const int4* ip = reinterpret_cast<const int4*>(idx + ((int64_t)b * n + jj) * 3);
    const float4* wp = reinterpret_cast<const float4*>(weight + ((int64_t)b * n + jj) * 3);
    ia = __ldg(ip); ib = __ldg(ip + 1); ic = __ldg(ip + 2);
    wa = __ldg(wp); wb = __ldg(wp + 1); wc = __ldg(wp + 2);
  };
  int j = j0 + 4 * t;
  if (j < j1) fetch(j);
  for (; j < j1; j += 4 * kTi4Threads) {
    const int id[12] = {ia.x, ia.y, ia.z, ia.w, ib.x, ib.y, ib.z, ib.w, ic.x, ic.y, ic.z, ic.w};
    const float w[12] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w, wc.x, wc.y, wc.z, wc.w};
    if (j + 4 * kTi4Threads < j1) fetch(j + 4 * kTi4Threads);  // the next quad's indices / weights while this one is computed
    for (int g4 = 0; g4 < ng; g4++) {
      const float4* rows = ti4_rows + (size_t)g4 * m;
      float o[4][4];  // [channel][target]
#pragma unroll
      for (int q = 0; q < 4; q++) {
        const float4 p0 = rows[id[3 * q]], p1 = rows[id[3 * q + 1]], p2 = rows[id[3 * q + 2]];
        const float w0 = w[3 * q], w1 = w[3 * q + 1], w2 = w[3 * q + 2];
        // reference SASS order: FMUL w1*p1 ; FFMA w0*p0 + . ; FFMA w2*p2 + .
        o[0][q] = __fmaf_rn(w2, p2.x, __fmaf_rn(w0, p0.x, __fmul_rn(w1, p1.x)));
        o[1][q] = __fmaf_rn(w2, p2.y, __fmaf_rn(w0, p0.y, __fmul_rn(w1, p1.y)));
        o[2][q] = __fmaf_rn(w2, p2.z, __fmaf_rn(w0, p0.z, __fmul_rn(w1, p1.z)));
        o[3][q] = __fmaf_rn(w2, p2.w, __fmaf_rn(w0, p0.w, __fmul_rn(w1, p1.w)));
      }
      const int live = min(4, gc - 4 * g4);
      float* ob = out + ((int64_t)b * C + c0 + 4 * g4) * n + j;
#pragma unroll
      for (int ch = 0; ch < 4; ch++)
        if (ch < live) __stcs(reinterpret_cast<float4*>(ob + (int64_t)ch * n), make_float4(o[ch][0], o[ch][1], o[ch][2], o[ch][3]));
    }
  }
}

__device__ __forceinline__ void ti_mbar_wait_phase(uint64_t* bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAITP_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@!p bra WAITP_%=;\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}

// Persistent form of the channel-interleaved kernel (default for large launches).  The grid-per-(cloud, channel group)
// kernel above leaves the 148 SMs unevenly loaded (256 CTAs at the bench shape: 108 SMs get two, 40 get one) and every
// CTA stages its rows through registers before its first output.  Here the work - (cloud, channel group) PAIRS x quads
// of 4 consecutive targets - is cut into one contiguous, equally long run per CTA, one CTA per SM.  A pair's G rows are
// contiguous in memory: they arrive by ONE bulk async copy into a raw buffer, already issued while the previous pair is
// being computed, and are interleaved smem -> smem (conflict-free LDS.32 x4 + STS.128).  Same FMA order as every
// other variant: same bits.
constexpr int kTipThreads = 512;
// grid (#CTAs), smem = 2 * (G/4) * m * 16 bytes: [interleaved float4 (G/4)*m | raw rows G*m floats]
__global__ void __launch_bounds__(kTipThreads, 1)
three_interpolate_persist_kernel(int B, int C, int m, int n, int G, const float* __restrict__ points,
                                 const int* __restrict__ idx, const float* __restrict__ weight, float* __restrict__ out) {
  extern __shared__ __align__(128) unsigned char tip_smem[];
  __shared__ uint64_t bar;
  float4* inter = reinterpret_cast<float4*>(tip_smem);
  float* raw = reinterpret_cast<float*>(tip_smem + (size_t)(G / 4) * m * 16);
  const int t = threadIdx.x;
  const int NG = (C + G - 1) / G, Q = n >> 2;  // quads per pair (n % 4 == 0)
  const int64_t W = (int64_t)B * NG * Q;
  const int64_t lo = W * blockIdx.x / gridDim.x, hi = W * (blockIdx.x + 1) / gridDim.x;
  if (lo >= hi) return;
  const int pair_first = (int)(lo / Q), pair_last = (int)((hi - 1) / Q);
  auto load_rows = [&](int pair) {  // thread 0 only
    const int b = pair / NG, c0 = (pair - b * NG) * G, gc = min(G, C - c0);
    ti_bulk_load(raw, points + ((int64_t)b * C + c0) * m, (unsigned)gc * (unsigned)m * 4u, &bar);
  };
  if (t == 0) {
    ti_mbar_init(&bar);
    load_rows(pair_first);
  }
  __syncthreads();  // the barrier init is visible to the waiters
  unsigned parity = 0;
  for (int pair = pair_first; pair <= pair_last; pair++) {
    const int b = pair / NG, c0 = (pair - b * NG) * G, gc = min(G, C - c0), ng = (gc + 3) >> 2;
    const int q0 = (int)(max(lo, (int64_t)pair * Q) - (int64_t)pair * Q);
    const int q1 = (int)(min(hi, (int64_t)(pair + 1) * Q) - (int64_t)pair * Q);
    int4 ia, ib, ic;
    float4 wa, wb, wc;
    auto fetch = [&](int q) {
      const int4* ip = reinterpret_cast<const int4*>(idx + ((int64_t)b * n + 4 * (int64_t)q) * 3);
      const float4* wp = reinterpret_cast<const float4*>(weight + ((int64_t)b * n + 4 * (int64_t)q) * 3);
      ia = __ldg(ip); ib = __ldg(ip + 1); ic = __ldg(ip + 2);
      wa = __ldg(wp); wb = __ldg(wp + 1); wc = __ldg(wp + 2);
    };
    int q = q0 + t;
    if (q < q1) fetch(q);  // the first quad's indices / weights travel while the rows arrive and are interleaved
    ti_mbar_wait_phase(&bar, parity);
    parity ^= 1u;
    for (int g4 = 0; g4 < ng; g4++) {
      const int live = min(4, gc - 4 * g4);
      const float* r0 = raw + (size_t)4 * g4 * m;
      for (int i = t; i < m; i += kTipThreads) {
        float4 v;
        v.x = r0[i];
        v.y = live > 1 ? r0[m + i] : 0.f;
        v.z = live > 2 ? r0[2 * m + i] : 0.f;
        v.w = live > 3 ? r0[3 * m + i] : 0.f;
        inter[(size_t)g4 * m + i] = v;
      }
    }
    __syncthreads();  // interleaved rows complete; the raw buffer is free again
    if (t == 0 && pair < pair_last) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic reads of `raw` before the async write
      load_rows(pair + 1);
    }
    for (; q < q1; q += kTipThreads) {
      const int id[12] = {ia.x, ia.y, ia.z, ia.w, ib.x, ib.y, ib.z, ib.w, ic.x, ic.y, ic.z, ic.w};
      const float w[12] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w, wc.x, wc.y, wc.z, wc.w};
      if (q + kTipThreads < q1) fetch(q + kTipThreads);  // the next quad's indices / weights while this one is computed
      for (int g4 = 0; g4 < ng; g4++) {
        const float4* rows = inter + (size_t)g4 * m;
        float o[4][4];  // [channel][target]
#pragma unroll
        for (int k = 0; k < 4; k++) {
          const float4 p0 = rows[id[3 * k]], p1 = rows[id[3 * k + 1]], p2 = rows[id[3 * k + 2]];
          const float w0 = w[3 * k], w1 = w[3 * k + 1], w2 = w[3 * k + 2];
          // reference SASS order: FMUL w1*p1 ; FFMA w0*p0 + . ; FFMA w2*p2 + .
          o[0][k] = __fmaf_rn(w2, p2.x, __fmaf_rn(w0, p0.x, __fmul_rn(w1, p1.x)));
          o[1][k] = __fmaf_rn(w2, p2.y, __fmaf_rn(w0, p0.y, __fmul_rn(w1, p1.y)));
          o[2][k] = __fmaf_rn(w2, p2.z, __fmaf_rn(w0, p0.z, __fmul_rn(w1, p1.z)));
          o[3][k] = __fmaf_rn(w2, p2.w, __fmaf_rn(w0, p0.w, __fmul_rn(w1, p1.w)));
        }
        const int live = min(4, gc - 4 * g4);
        float* ob = out + ((int64_t)b * C + c0 + 4 * g4) * n + 4 * (int64_t)q;
#pragma unroll
        for (int ch = 0; ch < 4; ch++)
          if (ch < live) __stcs(reinterpret_cast<float4*>(ob + (int64_t)ch * n), make_float4(o[ch][0], o[ch][1], o[ch][2], o[ch][3]));
      }
    }
    __syncthreads();  // every thread is done with the interleaved rows before the next pair overwrites them
  }
}

}  // namespace geot

using namespace geot;

GEOT_API int geot_three_nn(int B, int n, int m, const float* unknown, const float* known, float* dist2,
                           int* idx, geot_stream_t stream) {
  GEOT_REQUIRE(B >= 0 && n >= 0 && m >= 0, "three_nn: bad sizes");
  if (B == 0 || n == 0) return GEOT_OK;
  GEOT_REQUIRE(unknown && dist2 && idx && (m == 0 || known), "three_nn: NULL pointer");
  GEOT_REQUIRE(B <= 65535, "three_nn: B exceeds grid limits");
  // the grid costs a fixed ~50 us (every CTA bins the known cloud first); below ~5e7 distance evaluations the
  // brute-force kernel finishes sooner (measured: B=6, n=8192, m=512 brute 45 us vs grid 80 us; B=16, n=4096, m=1024 grid 43 vs 78)
  // known clouds longer than the kernel's shared-memory capacity (validation: 16000 sampled points against ~1e5
  // mesh vertices, train.py:793) are cut into P equal parts by index and searched part by part
  const int P = (m + kNgMaxM - 1) / kNgMaxM;
  if (m >= 64 && P <= 8 && ((double)B * n * m >= 5e7 || staged_kernels_forced()) && !direct_kernels_forced()) {
    const int part = (m + P - 1) / P;
    static DynSmem ds;
    if (int rc = ds.need(three_nn_grid_kernel, kNgSmemMax, "three_nn")) return rc;
    // CTAs per cloud: every CTA re-bins the known cloud (~4x the cost of one 1024-query sweep), so prefer few
    // CTAs with several sweeps each, in whole waves of the 148 SMs
    const int max_cb = (n + kNgThreads - 1) / kNgThreads;
    int cb = 1;
    double best_cost = 1e30;
    for (int k = 1; k <= max_cb; k++) {
      const double waves = std::ceil((double)k * B / num_sms());
      const double cost = waves * (4.0 + std::ceil((double)n / ((double)k * kNgThreads)));
      if (cost < best_cost) { best_cost = cost; cb = k; }
    }
    dim3 gg(cb, B);
    for (int p = 0; p < P; p++) {
      const int base = p * part, mp = std::min(part, m - base);
      if (mp <= 0) break;
      const size_t fixed = (size_t)mp * 16 + (size_t)(kNgMaxCells + 1) * 4 + (size_t)9 * kNgThreads * 4;
      // the warps' queues of queries waiting for ring 2, when shared memory has room for them next to the known cloud
      const size_t qbytes = (size_t)(kNgThreads / 32) * 7 * kNgWarpQueue * 4;
      const int use_queue = !three_nn_queue_off() && fixed + qbytes <= kNgSmemMax;
      const size_t smem = fixed + (use_queue ? qbytes : 0);
      three_nn_grid_kernel<<<gg, kNgThreads, smem, as_stream(stream)>>>(n, mp, m, base, p > 0, use_queue, unknown, known, dist2, idx);
      count_launch(1);
      const int rc = check_launch("three_nn_grid_kernel");
      if (rc) return rc;
    }
    return GEOT_OK;
  }
  dim3 grid((n + kNnThreads * kNnPerThread - 1) / (kNnThreads * kNnPerThread), B);
  three_nn_kernel<<<grid, kNnThreads, 0, as_stream(stream)>>>(n, m, unknown, known, dist2, idx);
  count_launch(1);
  return check_launch("three_nn_kernel");
}

GEOT_API int geot_three_interpolate(int B, int C, int m, int n, const float* points, const int* idx,
                                    const float* weight, float* out, geot_stream_t stream) {
  GEOT_REQUIRE(B >= 0 && C >= 0 && m >= 0 && n >= 0, "three_interpolate: bad sizes");
  if (B == 0 || C == 0 || n == 0) return GEOT_OK;
  GEOT_REQUIRE(points && idx && weight && out, "three_interpolate: NULL pointer");
  GEOT_REQUIRE(B <= 65535 && (C + kTiCCH - 1) / kTiCCH <= 65535, "three_interpolate: B or C exceeds grid limits");
  // rows staged in shared memory whenever a row (m floats) is not much longer than what it produces (n outputs): the
  // direct kernel pays a 32-byte L2 sector per 4-byte gather (measured 3.6x slower at cfg4's m=8192 -> n=16000)
  // interleaved rows: n % 4 == 0 (16-byte idx / weight / output accesses), at least 4 channels, 16 rows within 64 KB
  // (measured against the plain row-staged kernel: m = 512 -> 8192: 0.026 vs 0.029 ms, bench shape m = 1024: 0.093 vs
  // 0.097 ms; with fewer rows per CTA (m = 4096: 4 rows) the plain kernel is faster, 0.027 vs 0.031 ms)
  if (m > 0 && n % 4 == 0 && C >= 4 && (size_t)m * 64 <= 64 * 1024 && ((((uintptr_t)idx | (uintptr_t)weight | (uintptr_t)out) & 15) == 0) &&
      (8 * (int64_t)n >= m || staged_kernels_forced()) && !direct_kernels_forced() && !rows4_disabled()) {
    int G = std::max(4, (int)std::min<size_t>(16, ((size_t)64 * 1024) / ((size_t)m * 4)) & ~3);
    G = std::min(G, (C + 3) & ~3);
    const int NGp = (C + G - 1) / G;
    const int64_t Wq = (int64_t)B * NGp * (n / 4);
    // persistent kernel where a pair is long against its staging
    // (graph-timed A/B, tools/ti_ab.py: 0.079 vs 0.094 ms at the bench shape, 0.023 vs 0.028 at B=16, 0.018 vs 0.027 at
    // 6 x 128 x 1000 -> 16000; short pairs (n < 4096) lose to the grid kernel's co-resident CTAs: 0.0088 vs 0.0050 ms at
    // 16 x 128 x 256 -> 1024)
    if (m % 4 == 0 && (((uintptr_t)points & 15) == 0) && !persistent_interp_disabled() &&
        ((n >= 4096 && Wq >= 16 * kTipThreads) || staged_kernels_forced())) {
      const int ctas = (int)std::min<int64_t>((Wq + kTipThreads - 1) / kTipThreads, num_sms());
      const size_t smem = (size_t)2 * (G / 4) * m * 16;
      static DynSmem dsp;
      if (int rc = dsp.need(three_interpolate_persist_kernel, 128 * 1024, "three_interpolate", true)) return rc;
      three_interpolate_persist_kernel<<<ctas, kTipThreads, smem, as_stream(stream)>>>(B, C, m, n, G, points, idx, weight, out);
      count_launch(1);
      return check_launch("three_interpolate_persist_kernel");
    }
    int n_tile = 16384;  // measured: long tiles amortise the register-staged rows (0.120 ms at 4096, 0.093 at 16384; bench shape)
    if (n_tile > n) n_tile = n;
    // few pairs (measured at 8 x 64 x 1024 -> 16384: 32 CTAs 0.045 ms, the plain row kernel 0.015): shorter tiles until a wave is full
    while ((int64_t)B * NGp * ((n + n_tile - 1) / n_tile) < num_sms() && n_tile >= 4096 && n_tile % 8 == 0) n_tile /= 2;
    const size_t smem = (size_t)(G / 4) * m * 16;
    static DynSmem ds4;
    if (int rc = ds4.need(three_interpolate_rows4_kernel, 64 * 1024, "three_interpolate", true)) return rc;
    GEOT_REQUIRE((C + G - 1) / G <= 65535, "three_interpolate: C exceeds grid limits");
    dim3 grid((n + n_tile - 1) / n_tile, (C + G - 1) / G, B);
    three_interpolate_rows4_kernel<<<grid, kTi4Threads, smem, as_stream(stream)>>>(C, m, n, G, n_tile, points, idx, weight, out);
    count_launch(1);
    return check_launch("three_interpolate_rows4_kernel");
  }
  if (m > 0 && m % 4 == 0 && (((uintptr_t)points & 15) == 0) && (size_t)m * 4 <= 96 * 1024 && (8 * (int64_t)n >= m || staged_kernels_forced()) && !direct_kernels_forced()) {
    int G = (int)std::min<size_t>(16, (64 * 1024) / ((size_t)m * 4));
    if (G < 1) G = 1;
    if (G > C) G = C;
    int n_tile = std::max(4 * m, 2048);
    if (n_tile > n) n_tile = n;
    const size_t smem = (size_t)G * m * 4;
    static DynSmem ds;
    if (int rc = ds.need(three_interpolate_rows_kernel, 96 * 1024, "three_interpolate")) return rc;
    GEOT_REQUIRE((C + G - 1) / G <= 65535, "three_interpolate: C exceeds grid limits");
    dim3 grid((n + n_tile - 1) / n_tile, (C + G - 1) / G, B);
    three_interpolate_rows_kernel<<<grid, kTiThreads, smem, as_stream(stream)>>>(C, m, n, G, n_tile, points, idx, weight, out);
    count_launch(1);
    return check_launch("three_interpolate_rows_kernel");
  }
  dim3 grid((n + 255) / 256, (C + kTiCCH - 1) / kTiCCH, B);
  three_interpolate_kernel<<<grid, 256, 0, as_stream(stream)>>>(C, m, n, points, idx, weight, out);
  count_launch(1);
  return check_launch("three_interpolate_kernel");
}

GEOT_API int geot_three_interpolate_grad(int B, int C, int n, int m, const float* grad_out,
                                         const float* weight, const void* plan, float* grad_points,
                                         geot_stream_t stream) {
  GEOT_REQUIRE(B >= 0 && C >= 0 && m >= 0 && n >= 0, "three_interpolate_grad: bad sizes");
  GEOT_REQUIRE(weight != nullptr || n == 0, "three_interpolate_grad: NULL weight");
  return three_interpolate_grad_impl(B, C, n, m, grad_out, weight, plan, grad_points, as_stream(stream));
}
