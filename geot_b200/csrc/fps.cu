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
#include "fps_common.cuh"

namespace geot {

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
  if (n > CH * 4096) {  // the caller understated n_max (the Python layer computes it from offset): mark, do not leave garbage
    for (int j = t; j < m; j += blockDim.x) idx[start_m + j] = -1;
    return;
  }
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
// Default path: bucket-pruned, warp-specialised FPS, one CTA per cloud, n <= 16384 points.
//
// The M-1 rounds are sequential, but a new sample only lowers the min-distance of points near it.  The cloud
// is therefore binned once (12-bit Morton cell of the cloud's bounding cube, counting sort in shared memory)
// into buckets of 128 spatially close points, 4 points per lane of the owning warp, min-distances in registers.
// Each bucket has a bounding box and a cached candidate (its current maximum min-distance and who holds it).
// Per round the distance from the new sample to every box is evaluated with the SAME rounded operations as
// dist2_ref; every rounded step is monotone in |gap|, so  box_distance <= d(sample, p)  holds in fp32 for every
// p in the bucket and a bucket with box_distance >= bucket_max cannot change (fminf(d, mind) == mind): it is
// skipped, exactly.  On tooth clouds ~7 % of the buckets survive the test, so a round is a few dependent
// LDS/REDUX steps instead of N distance evaluations.  Internal point order never reaches the result: ties are
// resolved by the reference's rank of the ORIGINAL index.
//
// Warp specialisation: warp 0 is a MANAGER, warps 1..31 are WORKERS (if every warp repeated the box test and
// the winner reduction the round would be issue-bound on redundant work: measured 0.78 ms vs 0.55 ms).
//   * the manager keeps the boxes of all buckets in registers (4 per lane), reads the per-bucket candidate keys,
//     picks the global winner (2 REDUX), runs the exact box test for all buckets against the new sample and
//     publishes {sample, active flags}; then bar.arrive(A);
//   * a worker waits on A, updates only its ACTIVE buckets, republishes the candidate of each bucket it touched,
//     then bar.arrive(B); an idle worker issues ~10 instructions per round;
//   * the manager waits on B.  Two named barriers in producer/consumer form, nobody waits twice.
// ---------------------------------------------------------------------------------------------
#ifdef GEOT_FPS_PROF
// Debug build only (tools/fps_prof.py): per-round SM-clock stamps of block 0 -> a caller buffer.
// layout: stamp[(round * 32 + warp) * 8 + k]
__device__ long long* g_fps_prof = nullptr;
#define FPS_PROF_INIT long long* const prof_buf = (blockIdx.x == 0 && (threadIdx.x & 31) == 0) ? g_fps_prof : nullptr
#define FPS_STAMP(round, k) do { if (prof_buf) prof_buf[((round) * 32 + warp) * 8 + (k)] = clock64(); } while (0)
#else
#define FPS_PROF_INIT do {} while (0)
#define FPS_STAMP(round, k) do {} while (0)
#endif

template <int NBK>
struct MgrCfg {
  static constexpr int NT = 1024, NWK = 31;
  static constexpr int CHW = (NBK + NWK - 1) / NWK;  // buckets per worker
  static constexpr int BPL = NBK / 32;               // buckets per manager lane
  static constexpr int CAP = NBK * 128;
  static constexpr int PLANE_BYTES = NBK * 512;
  static constexpr int HIST_OFF = 3 * PLANE_BYTES;  // 4096 cell counters (setup only)
  static constexpr int KEY_OFF = HIST_OFF + 4096 * 4;
  static constexpr int BOX_OFF = KEY_OFF + NBK * 8;
  static constexpr int ACT_OFF = BOX_OFF + NBK * 24;
  static constexpr int CTRL_OFF = ACT_OFF + 512;  // 2 buffers of 8 flag bytes per worker
  static constexpr int SMEM_BYTES = CTRL_OFF + 64;  // 2 x {sample (16 B), speculative winner key (8 B), pad}
};

template <int NBK, bool SPEC>
__global__ void __launch_bounds__(1024, 1)
fps_mgr_kernel(const float* __restrict__ xyz, const int* __restrict__ offset,
               const int* __restrict__ new_offset, int N, int M, int* __restrict__ idx, int log2bs,
               int mag_skip) {
  using Cfg = MgrCfg<NBK>;
  constexpr int NT = Cfg::NT, NWK = Cfg::NWK, CHW = Cfg::CHW, BPL = Cfg::BPL;
  extern __shared__ float4 fps_smem[];
  unsigned char* smem_raw = reinterpret_cast<unsigned char*>(fps_smem);
  int* hist = reinterpret_cast<int*>(smem_raw + Cfg::HIST_OFF);
  __shared__ float s_red[6][32];
  __shared__ int s_scan[32];

  #ifdef GEOT_FPS_PROF
  const int abl = mag_skip >> 8;  // timing ablations (results are wrong): 1 = no bucket updates, 2 = updates without REDUX / key store
  mag_skip &= 0xff;
#endif
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  int64_t start_n, start_m;
  int n, m, idx_base;
  fps_segment(offset, new_offset, N, M, blockIdx.x, start_n, n, start_m, m, idx_base);
  if (m <= 0) return;
  if (n > Cfg::CAP || n <= 0) {  // the caller understated n_max (the Python layer computes it from offset): mark, do not leave garbage
    for (int j = t; j < m; j += blockDim.x) idx[start_m + j] = -1;
    return;
  }
  const int q_count = (n + (1 << log2bs) - 1) >> log2bs;
  const float* __restrict__ src = xyz + start_n * 3;
  const float kInf = __int_as_float(0x7f800000);

  // ---- 1. the cloud into registers (one coalesced pass over global memory), bounding cube -> 16^3 cells ----------
  constexpr int PPT = Cfg::CAP / NT;  // points per thread during setup
  float px[PPT], py[PPT], pz[PPT];
  float lo[3] = {kInf, kInf, kInf}, hi[3] = {-kInf, -kInf, -kInf};
#pragma unroll
  for (int i = 0; i < PPT; i++) {
    const int k = t + i * NT;
    px[i] = py[i] = pz[i] = 0.f;
    if (k < n) {
      px[i] = src[(int64_t)k * 3 + 0]; py[i] = src[(int64_t)k * 3 + 1]; pz[i] = src[(int64_t)k * 3 + 2];
      lo[0] = fminf(lo[0], px[i]); hi[0] = fmaxf(hi[0], px[i]);
      lo[1] = fminf(lo[1], py[i]); hi[1] = fmaxf(hi[1], py[i]);
      lo[2] = fminf(lo[2], pz[i]); hi[2] = fmaxf(hi[2], pz[i]);
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
  for (int i = t; i < 4096; i += NT) hist[i] = 0;
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
  auto cell_code = [&](float x, float y, float z) -> unsigned {  // NaN -> cell 0: partitioning only, never results
    const int qx = min(max(__float2int_rd((x - lo[0]) * scale), 0), 15);
    const int qy = min(max(__float2int_rd((y - lo[1]) * scale), 0), 15);
    const int qz = min(max(__float2int_rd((z - lo[2]) * scale), 0), 15);
    return morton4((unsigned)qx) | (morton4((unsigned)qy) << 1) | (morton4((unsigned)qz) << 2);
  };

  // ---- 2. counting sort by cell: histogram, exclusive scan, then every thread scatters ITS points (still in
  //         registers) straight to their sorted slots: x and y planes now, z after the ranks have been collected
  //         (the rank table borrows the z plane) -------------------------------------------------------------------
#pragma unroll
  for (int i = 0; i < PPT; i++)
    if (t + i * NT < n) atomicAdd(&hist[cell_code(px[i], py[i], pz[i])], 1);
  __syncthreads();
  {
    constexpr int PER = 4096 / NT;
    int v[PER], sum = 0;
#pragma unroll
    for (int i = 0; i < PER; i++) { v[i] = hist[t * PER + i]; sum += v[i]; }
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
    const int wbase = __shfl_sync(0xffffffffu, winc - wtot, warp);
    int run = wbase + inc - sum;
#pragma unroll
    for (int i = 0; i < PER; i++) { hist[t * PER + i] = run; run += v[i]; }
  }
  __syncthreads();

  unsigned a_base;  // opaque to the optimiser, or it re-derives the window base (S2UR SR_CgaCtaId) inside the loop
  asm volatile("mov.u32 %0, %1;" : "=r"(a_base) : "r"(smem_u32(smem_raw)));
  const unsigned a_key = a_base + Cfg::KEY_OFF, a_box = a_base + Cfg::BOX_OFF;
  const unsigned a_act = a_base + Cfg::ACT_OFF, a_ctrl = a_base + Cfg::CTRL_OFF;
  const unsigned a_rank = a_base + 2 * Cfg::PLANE_BYTES;  // u16 rank table, lives in the z plane until z is scattered
  const int wk = warp - 1;  // worker id (warp 0 = manager)
  const int rounds = m - 1;
  const float x0 = src[0], y0 = src[1], z0 = src[2];  // first sample = segment start (sampling_gpu.cu:120-122)
  if (t == 0) idx[start_m] = idx_base;

  unsigned short spos[PPT];
#pragma unroll
  for (int i = 0; i < PPT; i++) {
    const int k = t + i * NT;
    spos[i] = 0;
    if (k < n) {
      const int pos = atomicAdd(&hist[cell_code(px[i], py[i], pz[i])], 1);
      spos[i] = (unsigned short)pos;
      sts32(a_base + 4u * pos, __float_as_uint(px[i]));
      sts32(a_base + Cfg::PLANE_BYTES + 4u * pos, __float_as_uint(py[i]));
      sts_u16(a_rank + 2u * pos, tie_rank(k, log2bs, q_count));  // n <= 16384: rank < 2^15
    }
  }
  for (int pos = n + t; pos < Cfg::CAP; pos += NT) {  // padding slots
    sts32(a_base + 4u * pos, 0u);
    sts32(a_base + Cfg::PLANE_BYTES + 4u * pos, 0u);
  }
  __syncthreads();

  // ---- 3. workers take their buckets -----------------------------------------------------------------------------
  // Running min-distance, registers only.  -1 marks "never a candidate": padding beyond n and, for the
  // pointnet2/_ext variant, points the reference skips (|p|^2 <= 1e-3, _ext_src sampling_gpu.cu:103-104).
  float mind[CHW * 4];
  unsigned kreg[CHW * 2];  // per bucket: the reference tie ranks of its 4 points, packed u16
#pragma unroll
  for (int c = 0; c < CHW; c++) {
    const int b = c * NWK + wk;
    uint2 kk = make_uint2(0u, 0u);
    if (wk >= 0 && b < NBK) kk = lds64(a_rank + 2u * (unsigned)(b * 128 + 4 * lane));
    kreg[2 * c] = kk.x; kreg[2 * c + 1] = kk.y;
  }
  __syncthreads();  // the rank table is in registers: the z plane may be written
#pragma unroll
  for (int i = 0; i < PPT; i++)
    if (t + i * NT < n) sts32(a_base + 2 * Cfg::PLANE_BYTES + 4u * spos[i], __float_as_uint(pz[i]));
  for (int pos = n + t; pos < Cfg::CAP; pos += NT) sts32(a_base + 2 * Cfg::PLANE_BYTES + 4u * pos, 0u);
  __syncthreads();
#pragma unroll
  for (int c = 0; c < CHW; c++) {
    const int b = c * NWK + wk;
#pragma unroll
    for (int e = 0; e < 4; e++) mind[c * 4 + e] = -1.0f;
    if (wk >= 0 && b < NBK) {
      const int pos0 = b * 128 + 4 * lane;
      const unsigned a_pt = a_base + 16u * (unsigned)(b * 32 + lane);
      const float4 X = lds128(a_pt), Y = lds128(a_pt + Cfg::PLANE_BYTES), Z = lds128(a_pt + 2 * Cfg::PLANE_BYTES);
      const float xs[4] = {X.x, X.y, X.z, X.w}, ys[4] = {Y.x, Y.y, Y.z, Y.w}, zs[4] = {Z.x, Z.y, Z.z, Z.w};
      const unsigned rk[4] = {kreg[2 * c] & 0xffffu, kreg[2 * c] >> 16, kreg[2 * c + 1] & 0xffffu, kreg[2 * c + 1] >> 16};
      float l3[3] = {kInf, kInf, kInf}, h3[3] = {-kInf, -kInf, -kInf};
      unsigned r = 0xffffffffu;
#pragma unroll
      for (int e = 0; e < 4; e++) {
        bool live = pos0 + e < n;
        if (live && mag_skip) {
          const float mag = __fmaf_rn(zs[e], zs[e], __fmaf_rn(xs[e], xs[e], __fmul_rn(ys[e], ys[e])));
          live = !((double)mag <= 1e-3);
        }
        if (live) {
          mind[c * 4 + e] = 1e10f;
          l3[0] = fminf(l3[0], xs[e]); h3[0] = fmaxf(h3[0], xs[e]);
          l3[1] = fminf(l3[1], ys[e]); h3[1] = fmaxf(h3[1], ys[e]);
          l3[2] = fminf(l3[2], zs[e]); h3[2] = fmaxf(h3[2], zs[e]);
          r = min(r, (rk[e] << 2) | (unsigned)e);
        }
      }
#pragma unroll
      for (int a = 0; a < 3; a++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          l3[a] = fminf(l3[a], __shfl_xor_sync(0xffffffffu, l3[a], o));
          h3[a] = fmaxf(h3[a], __shfl_xor_sync(0xffffffffu, h3[a], o));
        }
      }
      if (lane == 0) {
        float* bx = reinterpret_cast<float*>(smem_raw + Cfg::BOX_OFF) + b * 6;
        bx[0] = l3[0]; bx[1] = l3[1]; bx[2] = l3[2]; bx[3] = h3[0]; bx[4] = h3[1]; bx[5] = h3[2];
      }
      // all live points start at 1e10: the bucket's first candidate is its live point of least rank
      const unsigned wr = __reduce_min_sync(0xffffffffu, r);
      if (wr == 0xffffffffu) {
        if (lane == 0) sts64(a_key + 8u * b, 0u, (unsigned)__float_as_int(-1.0f));
      } else if (r == wr) {
        sts64(a_key + 8u * b, ((0x7fffu - (wr >> 2)) << 14) | (unsigned)(pos0 + (int)(wr & 3u)), (unsigned)__float_as_int(1e10f));
      }
    }
  }
  if (t < 128) sts32(a_act + 4u * t, 0u);  // flag bytes of buckets that do not exist stay 0 (both buffers)
  __syncthreads();
  if (rounds <= 0) return;
  FPS_PROF_INIT;

  if constexpr (SPEC) {
    // ================= speculative variant: the manager runs ONE ROUND AHEAD of the workers =================
    // While the workers apply sample r, the manager already guesses sample r+1 = the best candidate among the buckets
    // the current sample does NOT touch (their keys cannot change in this round), fetches its coordinates and runs the
    // box tests for it.  The guess is the true winner iff no bucket updated in round r ends up with a larger key -
    // each worker checks that for its own buckets after the round (one 64-bit compare), one __syncthreads_or
    // collects the verdict.  Valid in 92-95 % of the rounds on tooth clouds (profiles/r02_fps_study.md); then the
    // manager's whole chain (pick, coordinates, box tests: ~300-450 cycles) is off the critical path.  Otherwise the
    // manager picks the true winner from the final keys as before (one extra barrier).  Box tests of a guessed round
    // use the bucket maxima from BEFORE the running round (larger or equal): a superset of the exact active set, and
    // updating a bucket that cannot change is a no-op, so the result is bit-identical in both cases.
    int* out = idx + start_m;
    if (warp == 0) {
      float box[BPL][6];
      unsigned o_flag[BPL];  // byte offset of a bucket's flag inside a flag buffer
      uint2 key[BPL];
#pragma unroll
      for (int i = 0; i < BPL; i++) {
        const int b = BPL * lane + i;
#pragma unroll
        for (int a = 0; a < 6; a++) box[i][a] = lds32f(a_box + 4u * (unsigned)(b * 6 + a));
        o_flag[i] = 8u * (unsigned)(b % NWK) + (unsigned)(b / NWK);
        key[i] = lds64(a_key + 8u * (unsigned)b);
      }
      auto box_flags = [&](float sx, float sy, float sz) -> unsigned {
        unsigned f = 0;
#pragma unroll
        for (int i = 0; i < BPL; i++) {
          const float gx = fmaxf(fmaxf(__fsub_rn(box[i][0], sx), __fsub_rn(sx, box[i][3])), 0.f);
          const float gy = fmaxf(fmaxf(__fsub_rn(box[i][1], sy), __fsub_rn(sy, box[i][4])), 0.f);
          const float gz = fmaxf(fmaxf(__fsub_rn(box[i][2], sz), __fsub_rn(sz, box[i][5])), 0.f);
          const float lb = __fmaf_rn(gz, gz, __fmaf_rn(gx, gx, __fmul_rn(gy, gy)));
          f |= (lb < __uint_as_float(key[i].y) ? 1u : 0u) << i;
        }
        return f;
      };
      auto publish = [&](int buf, unsigned f, float sx, float sy, float sz, unsigned klo, int khi) {
#pragma unroll
        for (int i = 0; i < BPL; i++) sts_u8(a_act + 256u * buf + o_flag[i], (f >> i) & 1u);
        if (lane == 0) {
          sts128(a_ctrl + 32u * buf, make_float4(0.f, sx, sy, sz));
          sts64(a_ctrl + 32u * buf + 16u, klo, (unsigned)khi);
        }
      };
      // arg-max of the 64-bit keys over the buckets whose bit in `skip` is clear
      auto pick = [&](unsigned skip, int& ghi, unsigned& glo) {
        unsigned long long best = 0ull;  // below every real key (value bits ^ sign >= 0x3f800000 for -1.0f ... )
#pragma unroll
        for (int i = 0; i < BPL; i++) {
          const unsigned long long k64 = ((unsigned long long)(key[i].y ^ 0x80000000u) << 32) | key[i].x;
          if (!((skip >> i) & 1u)) best = k64 > best ? k64 : best;
        }
        const unsigned bh = (unsigned)(best >> 32);
        const unsigned gh = __reduce_max_sync(0xffffffffu, bh);
        glo = __reduce_max_sync(0xffffffffu, bh == gh ? (unsigned)best : 0u);
        ghi = gh == 0u ? (int)0x80000000 : (int)(gh ^ 0x80000000u);  // nothing eligible -> below every value
      };
      unsigned fprev = 0u, fcur = box_flags(x0, y0, z0);
      publish(0, fcur, x0, y0, z0, 0u, 0);
      __syncthreads();
      for (int r = 0; r < rounds; r++) {
        const int nb = (r + 1) & 1;
        // keys of the buckets of round r-1 are final: refresh them, then guess the winner outside the running round
#pragma unroll
        for (int i = 0; i < BPL; i++)
          if ((fprev >> i) & 1u) key[i] = lds64(a_key + 8u * (unsigned)(BPL * lane + i));
        int ghi;
        unsigned glo;
        pick(fcur, ghi, glo);
        const bool spec_ok = ghi >= 0;
        float sx = x0, sy = y0, sz = z0;
        if (spec_ok) {
          const unsigned a_w = a_base + 4u * (glo & 0x3fffu);
          sx = lds32f(a_w); sy = lds32f(a_w + Cfg::PLANE_BYTES); sz = lds32f(a_w + 2 * Cfg::PLANE_BYTES);
        }
        unsigned fnext = box_flags(sx, sy, sz);
        publish(nb, fnext, sx, sy, sz, glo, ghi);
        __syncthreads();                                   // A: round r applied everywhere, the guess is published
        const int bad = __syncthreads_or(spec_ok ? 0 : 1);  // B: did any updated bucket beat the guess?
        if (bad) {
          // the true winner, from the final keys of round r (nobody writes keys between B and C)
#pragma unroll
          for (int i = 0; i < BPL; i++)
            if ((fcur >> i) & 1u) key[i] = lds64(a_key + 8u * (unsigned)(BPL * lane + i));
          pick(0u, ghi, glo);
          sx = x0; sy = y0; sz = z0;  // no candidate at all (every point skipped): the reference keeps index 0
          if (ghi >= 0) {
            const unsigned a_w = a_base + 4u * (glo & 0x3fffu);
            sx = lds32f(a_w); sy = lds32f(a_w + Cfg::PLANE_BYTES); sz = lds32f(a_w + 2 * Cfg::PLANE_BYTES);
          }
          fnext = box_flags(sx, sy, sz);
          publish(nb, fnext, sx, sy, sz, glo, ghi);
          __syncthreads();                                 // C
        }
        // the winner's (~rank) for now: un-ranking (an integer division) leaves the manager's serial chain and is done
        // by all threads after the last round
        if (lane == 0) out[r + 1] = ghi >= 0 ? (int)(glo >> 14) : -1;
        fprev = bad ? 0u : fcur;
        fcur = fnext;
      }
    } else {
      const unsigned a_my = a_act + 8u * (unsigned)wk;
      __syncthreads();
      for (int r = 0; r < rounds; r++) {
        const unsigned buf = (unsigned)(r & 1);
        uint2 fl = lds64(a_my + 256u * buf);  // one flag byte per owned bucket
#ifdef GEOT_FPS_PROF
        if (abl == 1) fl = make_uint2(0u, 0u);
#endif
        const float4 smp = lds128(a_ctrl + 32u * buf);
        int my_hi = (int)0x80000000;                 // the largest key this warp produced in this round
        unsigned my_lo = 0u;
        if (fl.x | fl.y) {
#pragma unroll
          for (int c = 0; c < CHW; c++) {
            const unsigned f = c < 4 ? (fl.x >> (8 * c)) & 0xffu : (fl.y >> (8 * (c - 4))) & 0xffu;
            const int b = c * NWK + wk;
            if (f && b < NBK) {
              const unsigned a_pt = a_base + 16u * (unsigned)(b * 32 + lane);
              const float4 X = lds128(a_pt), Y = lds128(a_pt + Cfg::PLANE_BYTES), Z = lds128(a_pt + 2 * Cfg::PLANE_BYTES);
              const float xs[4] = {X.x, X.y, X.z, X.w}, ys[4] = {Y.x, Y.y, Y.z, Y.w}, zs[4] = {Z.x, Z.y, Z.z, Z.w};
              unsigned k01 = kreg[2 * c], k23 = kreg[2 * c + 1];
              asm volatile("" : "+r"(k01), "+r"(k23));
              const unsigned rk[4] = {((k01 & 0xffffu) << 2) | 0u, ((k01 >> 16) << 2) | 1u,
                                      ((k23 & 0xffffu) << 2) | 2u, ((k23 >> 16) << 2) | 3u};
              float cm = -1.0f;
#pragma unroll
              for (int e = 0; e < 4; e++) {
                const float d = dist2_ref(xs[e], ys[e], zs[e], smp.y, smp.z, smp.w);
                const float v = fminf(d, mind[c * 4 + e]);
                mind[c * 4 + e] = v;
                cm = fmaxf(cm, v);
              }
              const int wm = __reduce_max_sync(0xffffffffu, __float_as_int(cm));
              unsigned rr = 0xffffffffu;
#pragma unroll
              for (int e = 0; e < 4; e++) rr = min(rr, (__float_as_int(mind[c * 4 + e]) == wm) ? rk[e] : 0xffffffffu);
              const unsigned wr = __reduce_min_sync(0xffffffffu, rr);
              const unsigned hi_bits = (0x7fffu - (wr >> 2)) << 14;  // (~rank): decides between equal values; ranks are unique
              if (rr == wr) sts64(a_key + 8u * b, hi_bits | (unsigned)(b * 128 + 4 * lane) | (wr & 3u), (unsigned)wm);
              if (wm > my_hi || (wm == my_hi && hi_bits > my_lo)) { my_hi = wm; my_lo = hi_bits; }
            }
          }
        }
        __syncthreads();                                   // A
        const uint2 g = lds64(a_ctrl + 32u * (buf ^ 1u) + 16u);  // the guessed winner's key {lo, hi}
        const int bad_here = (my_hi > (int)g.y) || (my_hi == (int)g.y && my_lo > (g.x & ~0x3fffu));
        if (__syncthreads_or(bad_here)) __syncthreads();   // B (, C)
      }
    }
    __syncthreads();  // out[1..rounds] hold (~rank) or -1: rank 0 <-> index 0
    for (int j = 1 + t; j <= rounds; j += NT) {
      const int v = out[j];
      out[j] = idx_base + (v >= 0 ? tie_unrank(0x7fffu - (unsigned)v, log2bs, q_count) : 0);
    }
    return;
  }

  if (warp == 0) {
    // =========================== manager ===========================
    float box[BPL][6];
    unsigned a_flag[BPL];  // where worker (b % NWK) reads the active flag of its chunk (b / NWK)
#pragma unroll
    for (int i = 0; i < BPL; i++) {
      const int b = BPL * lane + i;
#pragma unroll
      for (int a = 0; a < 6; a++) box[i][a] = lds32f(a_box + 4u * (unsigned)(b * 6 + a));
      a_flag[i] = a_act + 8u * (unsigned)(b % NWK) + (unsigned)(b / NWK);
    }
    int* out = idx + start_m;
    // publish {sample, per-worker active flags}: exact box test of every bucket against the sample, rounded like
    // dist2_ref; a bucket's current maximum is the value word of its candidate key
    // key = {lo: (~rank << 14) | sorted position, hi: value bits}
    auto publish = [&](const float4& smp, const uint2 (&key)[BPL]) {
#pragma unroll
      for (int i = 0; i < BPL; i++) {
        const float gx = fmaxf(fmaxf(__fsub_rn(box[i][0], smp.y), __fsub_rn(smp.y, box[i][3])), 0.f);
        const float gy = fmaxf(fmaxf(__fsub_rn(box[i][1], smp.z), __fsub_rn(smp.z, box[i][4])), 0.f);
        const float gz = fmaxf(fmaxf(__fsub_rn(box[i][2], smp.w), __fsub_rn(smp.w, box[i][5])), 0.f);
        const float lb = __fmaf_rn(gz, gz, __fmaf_rn(gx, gx, __fmul_rn(gy, gy)));
        sts_u8(a_flag[i], lb < __uint_as_float(key[i].y) ? 1u : 0u);
      }
      if (lane == 0) sts128(a_ctrl, smp);
      bar_arrive_named(1);
    };
    // global winner: max value, then max of (~rank | bucket) among the holders
    auto pick = [&](const uint2 (&key)[BPL], int& ghi, unsigned& glo) {
      unsigned long long best = ((unsigned long long)(key[0].y ^ 0x80000000u) << 32) | key[0].x;
#pragma unroll
      for (int i = 1; i < BPL; i++) {
        const unsigned long long k64 = ((unsigned long long)(key[i].y ^ 0x80000000u) << 32) | key[i].x;
        best = k64 > best ? k64 : best;
      }
      const int bh = (int)((unsigned)(best >> 32) ^ 0x80000000u);
      ghi = __reduce_max_sync(0xffffffffu, bh);
      glo = __reduce_max_sync(0xffffffffu, bh == ghi ? (unsigned)best : 0u);
    };
    uint2 key[BPL];
#pragma unroll
    for (int i = 0; i < BPL; i++) key[i] = lds64(a_key + 8u * (unsigned)(BPL * lane + i));
    publish(make_float4(0.f, x0, y0, z0), key);
    for (int j = 1; j <= rounds; j++) {
      bar_sync_named(2);  // workers have applied sample j-1
      FPS_STAMP(j, 0);
#pragma unroll
      for (int i = 0; i < BPL; i++) key[i] = lds64(a_key + 8u * (unsigned)(BPL * lane + i));
      int ghi;
      unsigned glo;
      pick(key, ghi, glo);
      FPS_STAMP(j, 1);
      // no candidate at all (every point skipped): the reference keeps index 0
      float4 smp = make_float4(0.f, x0, y0, z0);
      if (ghi >= 0) {
        const unsigned a_w = a_base + 4u * (glo & 0x3fffu);  // the winner's coordinates, by sorted position
        smp.y = lds32f(a_w); smp.z = lds32f(a_w + Cfg::PLANE_BYTES); smp.w = lds32f(a_w + 2 * Cfg::PLANE_BYTES);
      }
      FPS_STAMP(j, 2);
      if (j < rounds) publish(smp, key);
      FPS_STAMP(j, 3);
      // off the critical path: un-rank the winner (rank 0 <-> index 0)
      if (lane == 0) out[j] = idx_base + (ghi >= 0 ? tie_unrank(0x7fffu - (glo >> 14), log2bs, q_count) : 0);
    }
  } else {
    // =========================== workers ===========================
    const unsigned a_my = a_act + 8u * (unsigned)wk;
    for (int j = 0; j < rounds; j++) {
      bar_sync_named(1);  // manager has published the sample and the active flags of round j
      FPS_STAMP(j, 0);
      uint2 fl = lds64(a_my);  // one flag byte per owned bucket
#ifdef GEOT_FPS_PROF
      if (abl == 1) fl = make_uint2(0u, 0u);
#endif
      if (fl.x | fl.y) {
        FPS_STAMP(j, 1);
        const float4 smp = lds128(a_ctrl);
#pragma unroll
        for (int c = 0; c < CHW; c++) {
          const unsigned f = c < 4 ? (fl.x >> (8 * c)) & 0xffu : (fl.y >> (8 * (c - 4))) & 0xffu;
          const int b = c * NWK + wk;
          if (f && b < NBK) {
            const unsigned a_pt = a_base + 16u * (unsigned)(b * 32 + lane);
            const float4 X = lds128(a_pt), Y = lds128(a_pt + Cfg::PLANE_BYTES), Z = lds128(a_pt + 2 * Cfg::PLANE_BYTES);
            const float xs[4] = {X.x, X.y, X.z, X.w}, ys[4] = {Y.x, Y.y, Y.z, Y.w}, zs[4] = {Z.x, Z.y, Z.z, Z.w};
            unsigned k01 = kreg[2 * c], k23 = kreg[2 * c + 1];
            asm volatile("" : "+r"(k01), "+r"(k23));  // keep the unpacking in the loop: hoisted it costs 20 registers
            const unsigned rk[4] = {((k01 & 0xffffu) << 2) | 0u, ((k01 >> 16) << 2) | 1u,
                                    ((k23 & 0xffffu) << 2) | 2u, ((k23 >> 16) << 2) | 3u};
            float cm = -1.0f;
#pragma unroll
            for (int e = 0; e < 4; e++) {
              const float d = dist2_ref(xs[e], ys[e], zs[e], smp.y, smp.z, smp.w);
              const float v = fminf(d, mind[c * 4 + e]);
              mind[c * 4 + e] = v;
              cm = fmaxf(cm, v);
            }
            if (cm > -5.f) FPS_STAMP(j, 4);
#ifdef GEOT_FPS_PROF
            if (abl == 2) continue;
#endif
            // candidates are >= 0 or exactly -1, so the int view of the float bits is order preserving
            const int wm = __reduce_max_sync(0xffffffffu, __float_as_int(cm));
            unsigned r = 0xffffffffu;
#pragma unroll
            for (int e = 0; e < 4; e++) r = min(r, (__float_as_int(mind[c * 4 + e]) == wm) ? rk[e] : 0xffffffffu);
            if (r != 7u) FPS_STAMP(j, 5);
            const unsigned wr = __reduce_min_sync(0xffffffffu, r);
            // exactly one lane publishes: ranks are unique (an active bucket has a live point, so wm >= 0)
            if (r == wr) sts64(a_key + 8u * b, ((0x7fffu - (wr >> 2)) << 14) | (unsigned)(b * 128 + 4 * lane) | (wr & 3u), (unsigned)wm);
            if (wr != 9u) FPS_STAMP(j, 6);
          }
        }
        FPS_STAMP(j, 2);
      }
      bar_arrive_named(2);
      FPS_STAMP(j, 3);
    }
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
  static DynSmem ds;  // per instantiation
  if (int rc = ds.need(fps_smem_kernel<CH>, smem, "fps")) return rc;
  fps_smem_kernel<CH><<<blocks, 1024, smem, s>>>(xyz, offset, new_offset, N, M, idx, log2bs, mag_skip);
  count_launch(1);
  return check_launch("fps_smem_kernel");
}

template <int NBK>
static int launch_mgr(int blocks, const float* xyz, const int* offset, const int* new_offset, int N,
                      int M, int* idx, int log2bs, int mag_skip, bool spec, cudaStream_t s) {
  constexpr int smem = MgrCfg<NBK>::SMEM_BYTES;
  static DynSmem ds, ds_spec;  // per instantiation
  if (spec) {
    if (int rc = ds_spec.need(fps_mgr_kernel<NBK, true>, smem, "fps")) return rc;
    fps_mgr_kernel<NBK, true><<<blocks, 1024, smem, s>>>(xyz, offset, new_offset, N, M, idx, log2bs, mag_skip);
  } else {
    if (int rc = ds.need(fps_mgr_kernel<NBK, false>, smem, "fps")) return rc;
    fps_mgr_kernel<NBK, false><<<blocks, 1024, smem, s>>>(xyz, offset, new_offset, N, M, idx, log2bs, mag_skip);
  }
  count_launch(1);
  return check_launch("fps_mgr_kernel");
}

static int fps_dispatch(int blocks, int n_max, const float* xyz, const int* offset, const int* new_offset,
                        int N, int M, const float* weights, float* temp, int* idx, int flags,
                        cudaStream_t s) {
  const int cap = (flags & GEOT_FPS_CAP512) ? 512 : 1024;
  const int log2bs = ilog2(ref_opt_n_threads(n_max, cap));
  int mag_skip = (flags & GEOT_FPS_MAGSKIP) ? 1 : 0;
#ifdef GEOT_FPS_PROF
  mag_skip |= flags & ~0xff;  // ablation selector of the profiling build
#endif
  const bool generic = (flags & GEOT_FPS_FORCE_GENERIC) || weights != nullptr || n_max > 4 * 4096;
  if (!generic && !(flags & GEOT_FPS_NO_PRUNE)) {
    const int ch = (n_max + 4095) / 4096;
    const bool spec = (flags & GEOT_FPS_SPEC) != 0;
    switch (ch) {
      case 1: return launch_mgr<32>(blocks, xyz, offset, new_offset, N, M, idx, log2bs, mag_skip, spec, s);
      case 2: return launch_mgr<64>(blocks, xyz, offset, new_offset, N, M, idx, log2bs, mag_skip, spec, s);
      case 3: return launch_mgr<96>(blocks, xyz, offset, new_offset, N, M, idx, log2bs, mag_skip, spec, s);
      default: return launch_mgr<128>(blocks, xyz, offset, new_offset, N, M, idx, log2bs, mag_skip, spec, s);
    }
  }
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

#ifdef GEOT_FPS_PROF
GEOT_API int geot_debug_fps_prof(long long* buf) {
  return (int)cudaMemcpyToSymbol(g_fps_prof, &buf, sizeof(buf));
}
#endif

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
