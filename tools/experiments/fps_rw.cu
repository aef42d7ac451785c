// NOT COMPILED INTO THE LIBRARY.  A measured alternative to fps.cu's manager/worker kernel, kept as evidence (profiles/r02_fps_study.md):
// bit-exact on every FPS test case, but 2.4x SLOWER at N=16384 (1.36 vs 0.54 us per round) - redundant pick + box test in every warp
// costs more dependent-issue latency than the barrier hop it removes.  To rebuild it: copy into geot_b200/csrc/ and restore the
// dispatch hook of commit history (GEOT_FPS_MGR / GEOT_FPS_NW16 flags).
// geot_b200/csrc/fps_rw.cu - bucket-pruned FPS with ONE barrier per round ("redundant warps"), sm_100a.
//
// Replaces (bit-exact indices) the same three reference kernels as fps.cu:
//   openpoints/cpp/pointnet2_batch/src/sampling_gpu.cu:101-216, pointops/src/sampling/sampling_cuda_kernel.cu:15-129,
//   pointnet2/_ext_src/src/sampling_gpu.cu:73-176.
//
// FPS is a chain of M-1 dependent rounds; with the bucket pruning of fps.cu (exact box test, ~7 % of the 128-point
// buckets survive) a round is no longer throughput- but LATENCY-bound.  The manager/worker kernel of fps.cu spends
// ~1000 cycles per round on: manager pick + box tests (~290) -> named barrier -> slowest worker (~640) -> named
// barrier.  This kernel removes one of the two hand-offs and most of the waiting:
//   * NW (8) warps per cloud instead of 32.  EVERY warp keeps the candidate keys of all buckets in registers
//     (lane l holds buckets l, l+32, l+64, l+96), picks the global winner itself (2 REDUX), fetches the winner's
//     coordinates, runs the exact box test of all buckets (4 per lane) and ballots the active set - redundantly,
//     which with 8 warps costs ~1200 warp-instructions per round over 4 schedulers instead of a barrier hop
//     (with 32 warps the same redundancy was issue-bound: measured 0.78 ms against 0.55 ms in round 1);
//   * warp w owns buckets b = w (mod NW): min-distances in registers, it updates its active buckets and publishes
//     their new candidate keys;
//   * ONE __syncthreads per round.  Keys are double-buffered by round parity: round j writes buf[j&1], readers
//     refresh only the buckets that were active in round j from buf[j&1], so a late reader can never see a key of
//     round j+1 (those go to the other buffer; skew is bounded by the barrier).
// Setup (Morton counting sort into buckets, boxes, tie ranks) re-reads the cloud from L2 in three passes instead of
// holding it in registers (256 threads: 64 points each would be 192 registers).
#include "fps_common.cuh"

namespace geot {

template <int NBK, int NW>
struct RwCfg {
  static constexpr int NT = NW * 32;
  static constexpr int CHW = NBK / NW;   // buckets owned by a warp
  static constexpr int BPL = NBK / 32;   // buckets per lane in the redundant pick / box test
  static constexpr int GPW = 32 / NW;    // owned buckets per 32-bucket ballot group
  static constexpr int CAP = NBK * 128;
  static constexpr int PPT = CAP / NT;   // points per thread during setup
  static constexpr int PLANE_BYTES = NBK * 512;
  static constexpr int HIST_OFF = 3 * PLANE_BYTES;  // 4096 cell counters (setup only)
  static constexpr int KEY_OFF = HIST_OFF + 4096 * 4;  // 2 parity buffers of NBK keys
  static constexpr int BOX_OFF = KEY_OFF + 2 * NBK * 8;
  static constexpr int SMEM_BYTES = BOX_OFF + NBK * 24;
  static_assert(NBK % 32 == 0 && NBK % NW == 0 && (NW == 8 || NW == 16), "bucket / warp layout");
};

template <int NBK, int NW>
__global__ void __launch_bounds__(NW * 32, 1)
fps_rw_kernel(const float* __restrict__ xyz, const int* __restrict__ offset, const int* __restrict__ new_offset,
              int N, int M, int* __restrict__ idx, int log2bs, int mag_skip) {
  using Cfg = RwCfg<NBK, NW>;
  constexpr int NT = Cfg::NT, CHW = Cfg::CHW, BPL = Cfg::BPL, GPW = Cfg::GPW, PPT = Cfg::PPT;
  extern __shared__ float4 fps_smem[];
  unsigned char* smem_raw = reinterpret_cast<unsigned char*>(fps_smem);
  int* hist = reinterpret_cast<int*>(smem_raw + Cfg::HIST_OFF);
  __shared__ float s_red[6][32];
  __shared__ int s_scan[32];

  #ifdef GEOT_FPS_PROF
  const int abl = mag_skip >> 8;  // timing ablations (results are wrong): 1 = no bucket updates, 2 = updates without REDUX / key store
  mag_skip &= 0xff;
#endif
  const int t = threadIdx.x, lane = t & 31, w = t >> 5;
  int64_t start_n, start_m;
  int n, m, idx_base;
  fps_segment(offset, new_offset, N, M, blockIdx.x, start_n, n, start_m, m, idx_base);
  if (m <= 0) return;
  if (n > Cfg::CAP || n <= 0) return;  // the host validates n_max; nothing to do for an empty segment
  const int q_count = (n + (1 << log2bs) - 1) >> log2bs;
  const float* __restrict__ src = xyz + start_n * 3;
  const float kInf = __int_as_float(0x7f800000);

  // ---- 1. bounding cube -> 16^3 Morton cells (pass 1 over the cloud) ---------------------------------------------
  float lo[3] = {kInf, kInf, kInf}, hi[3] = {-kInf, -kInf, -kInf};
  for (int k = t; k < n; k += NT) {
    const float x = src[(int64_t)k * 3 + 0], y = src[(int64_t)k * 3 + 1], z = src[(int64_t)k * 3 + 2];
    lo[0] = fminf(lo[0], x); hi[0] = fmaxf(hi[0], x);
    lo[1] = fminf(lo[1], y); hi[1] = fmaxf(hi[1], y);
    lo[2] = fminf(lo[2], z); hi[2] = fmaxf(hi[2], z);
  }
#pragma unroll
  for (int a = 0; a < 3; a++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
      hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
    }
    if (lane == 0) { s_red[a][w] = lo[a]; s_red[3 + a][w] = hi[a]; }
  }
  for (int i = t; i < 4096; i += NT) hist[i] = 0;
  __syncthreads();
  float ext = 0.f;
#pragma unroll
  for (int a = 0; a < 3; a++) {
    float l = lane < NW ? s_red[a][lane] : kInf, h = lane < NW ? s_red[3 + a][lane] : -kInf;
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

  // ---- 2. counting sort by cell: histogram (pass 2), exclusive scan ----------------------------------------------
  for (int k = t; k < n; k += NT)
    atomicAdd(&hist[cell_code(src[(int64_t)k * 3 + 0], src[(int64_t)k * 3 + 1], src[(int64_t)k * 3 + 2])], 1);
  __syncthreads();
  {
    constexpr int PER = 4096 / NT;
    int sum = 0;
#pragma unroll
    for (int i = 0; i < PER; i++) sum += hist[t * PER + i];
    int inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += y;
    }
    if (lane == 31) s_scan[w] = inc;
    __syncthreads();
    int wtot = lane < NW ? s_scan[lane] : 0, winc = wtot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o) winc += y;
    }
    const int wbase = __shfl_sync(0xffffffffu, winc - wtot, w);
    int run = wbase + inc - sum;
#pragma unroll
    for (int i = 0; i < PER; i++) { const int v = hist[t * PER + i]; hist[t * PER + i] = run; run += v; }
  }
  __syncthreads();

  unsigned a_base;  // opaque to the optimiser, or it re-derives the window base inside the round loop
  asm volatile("mov.u32 %0, %1;" : "=r"(a_base) : "r"(smem_u32(smem_raw)));
  const unsigned a_key = a_base + Cfg::KEY_OFF, a_box = a_base + Cfg::BOX_OFF;
  const unsigned a_rank = a_base + 2 * Cfg::PLANE_BYTES;  // u16 rank table, lives in the z plane until z is scattered
  const int rounds = m - 1;
  const float x0 = src[0], y0 = src[1], z0 = src[2];  // first sample = segment start (sampling_gpu.cu:120-122)
  if (t == 0) idx[start_m] = idx_base;

  // ---- 3. scatter (pass 3): x, y and the tie rank now; z after the ranks have been collected -------------------
  unsigned spos[PPT / 2];  // sorted slot of each of this thread's points, u16 pairs
#pragma unroll
  for (int i = 0; i < PPT; i++) {
    const int k = t + i * NT;
    unsigned pos = 0;
    if (k < n) {
      const float x = src[(int64_t)k * 3 + 0], y = src[(int64_t)k * 3 + 1], z = src[(int64_t)k * 3 + 2];
      pos = (unsigned)atomicAdd(&hist[cell_code(x, y, z)], 1);
      sts32(a_base + 4u * pos, __float_as_uint(x));
      sts32(a_base + Cfg::PLANE_BYTES + 4u * pos, __float_as_uint(y));
      sts_u16(a_rank + 2u * pos, tie_rank(k, log2bs, q_count));  // n <= 16384: rank < 2^15
    }
    if (i & 1) spos[i >> 1] |= pos << 16; else spos[i >> 1] = pos;
  }
  for (int pos = n + t; pos < Cfg::CAP; pos += NT) {  // padding slots
    sts32(a_base + 4u * pos, 0u);
    sts32(a_base + Cfg::PLANE_BYTES + 4u * pos, 0u);
  }
  __syncthreads();

  // ---- 4. every warp takes its buckets: tie ranks, min-distances (registers), boxes, first candidates -----------
  float mind[CHW * 4];
  unsigned kreg[CHW * 2];  // per owned bucket: the reference tie ranks of the lane's 4 points, packed u16
#pragma unroll
  for (int c = 0; c < CHW; c++) {
    const int b = c * NW + w;
    const uint2 kk = lds64(a_rank + 2u * (unsigned)(b * 128 + 4 * lane));
    kreg[2 * c] = kk.x; kreg[2 * c + 1] = kk.y;
  }
  __syncthreads();  // the rank table is in registers: the z plane may be written
#pragma unroll
  for (int i = 0; i < PPT; i++) {
    const int k = t + i * NT;
    if (k < n) {
      const unsigned pos = (i & 1) ? spos[i >> 1] >> 16 : spos[i >> 1] & 0xffffu;
      sts32(a_base + 2 * Cfg::PLANE_BYTES + 4u * pos, __float_as_uint(src[(int64_t)k * 3 + 2]));
    }
  }
  for (int pos = n + t; pos < Cfg::CAP; pos += NT) sts32(a_base + 2 * Cfg::PLANE_BYTES + 4u * pos, 0u);
  __syncthreads();
#pragma unroll
  for (int c = 0; c < CHW; c++) {
    const int b = c * NW + w;
    const int pos0 = b * 128 + 4 * lane;
    const unsigned a_pt = a_base + 16u * (unsigned)(b * 32 + lane);
    const float4 X = lds128(a_pt), Y = lds128(a_pt + Cfg::PLANE_BYTES), Z = lds128(a_pt + 2 * Cfg::PLANE_BYTES);
    const float xs[4] = {X.x, X.y, X.z, X.w}, ys[4] = {Y.x, Y.y, Y.z, Y.w}, zs[4] = {Z.x, Z.y, Z.z, Z.w};
    const unsigned rk[4] = {kreg[2 * c] & 0xffffu, kreg[2 * c] >> 16, kreg[2 * c + 1] & 0xffffu, kreg[2 * c + 1] >> 16};
    float l3[3] = {kInf, kInf, kInf}, h3[3] = {-kInf, -kInf, -kInf};
    unsigned r = 0xffffffffu;
#pragma unroll
    for (int e = 0; e < 4; e++) {
      // -1 marks "never a candidate": padding beyond n and, for the pointnet2/_ext variant, points the reference
      // skips (|p|^2 <= 1e-3, _ext_src sampling_gpu.cu:103-104)
      mind[c * 4 + e] = -1.0f;
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
    // key = {lo: (~rank << 14) | sorted position, hi: value bits}
    const unsigned wr = __reduce_min_sync(0xffffffffu, r);
    if (wr == 0xffffffffu) {
      if (lane == 0) sts64(a_key + 8u * b, 0u, (unsigned)__float_as_int(-1.0f));
    } else if (r == wr) {
      sts64(a_key + 8u * b, ((0x7fffu - (wr >> 2)) << 14) | (unsigned)(pos0 + (int)(wr & 3u)), (unsigned)__float_as_int(1e10f));
    }
  }
  __syncthreads();
  if (rounds <= 0) return;

  // ---- 5. rounds --------------------------------------------------------------------------------------------------
  float box[BPL][6];
  uint2 key[BPL];
#pragma unroll
  for (int i = 0; i < BPL; i++) {
    const int b = lane + 32 * i;
#pragma unroll
    for (int a = 0; a < 6; a++) box[i][a] = lds32f(a_box + 4u * (unsigned)(b * 6 + a));
    key[i] = lds64(a_key + 8u * (unsigned)b);
  }
  int* out = idx + start_m;
  unsigned act[BPL];
  // exact box test of every bucket against the sample, rounded like dist2_ref (monotone in |gap|, so a true lower
  // bound in fp32); a bucket whose bound is not below its current maximum cannot change
  auto box_test = [&](float sx, float sy, float sz) {
#pragma unroll
    for (int i = 0; i < BPL; i++) {
      const float gx = fmaxf(fmaxf(__fsub_rn(box[i][0], sx), __fsub_rn(sx, box[i][3])), 0.f);
      const float gy = fmaxf(fmaxf(__fsub_rn(box[i][1], sy), __fsub_rn(sy, box[i][4])), 0.f);
      const float gz = fmaxf(fmaxf(__fsub_rn(box[i][2], sz), __fsub_rn(sz, box[i][5])), 0.f);
      const float lb = __fmaf_rn(gz, gz, __fmaf_rn(gx, gx, __fmul_rn(gy, gy)));
      act[i] = __ballot_sync(0xffffffffu, lb < __uint_as_float(key[i].y));
    }
  };
  float sx = x0, sy = y0, sz = z0;
  box_test(sx, sy, sz);
  for (int j = 1; j <= rounds; j++) {
    const unsigned a_kw = a_key + (unsigned)(j & 1) * (NBK * 8);
    // ---- apply sample j-1 to my active buckets, publish their new candidates into buf[j&1]
#pragma unroll
    for (int g = 0; g < BPL; g++) {
      unsigned mine = act[g] >> w;  // bit (q*NW) <-> my q-th bucket of this 32-bucket group
#ifdef GEOT_FPS_PROF
      if (abl == 1) mine = 0;
#endif
      if (mine & (NW == 8 ? 0x01010101u : 0x00010001u)) {
#pragma unroll
        for (int q = 0; q < GPW; q++) {
          if ((mine >> (q * NW)) & 1u) {
            const int c = g * GPW + q;
            const int b = c * NW + w;  // = 32 g + q NW + w
            const unsigned a_pt = a_base + 16u * (unsigned)(b * 32 + lane);
            const float4 X = lds128(a_pt), Y = lds128(a_pt + Cfg::PLANE_BYTES), Z = lds128(a_pt + 2 * Cfg::PLANE_BYTES);
            const float xs[4] = {X.x, X.y, X.z, X.w}, ys[4] = {Y.x, Y.y, Y.z, Y.w}, zs[4] = {Z.x, Z.y, Z.z, Z.w};
            unsigned k01 = kreg[2 * c], k23 = kreg[2 * c + 1];
            asm volatile("" : "+r"(k01), "+r"(k23));  // keep the unpacking here: hoisted it costs registers
            const unsigned rk[4] = {((k01 & 0xffffu) << 2) | 0u, ((k01 >> 16) << 2) | 1u,
                                    ((k23 & 0xffffu) << 2) | 2u, ((k23 >> 16) << 2) | 3u};
            float cm = -1.0f;
#pragma unroll
            for (int e = 0; e < 4; e++) {
              const float d = dist2_ref(xs[e], ys[e], zs[e], sx, sy, sz);
              const float v = fminf(d, mind[c * 4 + e]);
              mind[c * 4 + e] = v;
              cm = fmaxf(cm, v);
            }
#ifdef GEOT_FPS_PROF
            if (abl == 2) continue;
#endif
            // candidates are >= 0 or exactly -1, so the int view of the float bits is order preserving
            const int wm = __reduce_max_sync(0xffffffffu, __float_as_int(cm));
            unsigned r = 0xffffffffu;
#pragma unroll
            for (int e = 0; e < 4; e++) r = min(r, (__float_as_int(mind[c * 4 + e]) == wm) ? rk[e] : 0xffffffffu);
            const unsigned wr = __reduce_min_sync(0xffffffffu, r);
            // exactly one lane publishes: ranks are unique (an active bucket has a live point, so wm >= 0)
            if (r == wr) sts64(a_kw + 8u * b, ((0x7fffu - (wr >> 2)) << 14) | (unsigned)(b * 128 + 4 * lane) | (wr & 3u), (unsigned)wm);
          }
        }
      }
    }
    __syncthreads();
    // ---- refresh the keys of the buckets that were active, pick the global winner (every warp, redundantly)
#pragma unroll
    for (int i = 0; i < BPL; i++)
      if ((act[i] >> lane) & 1u) key[i] = lds64(a_kw + 8u * (unsigned)(lane + 32 * i));
    unsigned long long best = ((unsigned long long)(key[0].y ^ 0x80000000u) << 32) | key[0].x;
#pragma unroll
    for (int i = 1; i < BPL; i++) {
      const unsigned long long k64 = ((unsigned long long)(key[i].y ^ 0x80000000u) << 32) | key[i].x;
      best = k64 > best ? k64 : best;
    }
    const int bh = (int)((unsigned)(best >> 32) ^ 0x80000000u);
    const int ghi = __reduce_max_sync(0xffffffffu, bh);
    const unsigned glo = __reduce_max_sync(0xffffffffu, bh == ghi ? (unsigned)best : 0u);
    // no candidate at all (every point skipped): the reference keeps index 0
    sx = x0; sy = y0; sz = z0;
    if (ghi >= 0) {
      const unsigned a_w = a_base + 4u * (glo & 0x3fffu);  // the winner's coordinates, by sorted position
      sx = lds32f(a_w); sy = lds32f(a_w + Cfg::PLANE_BYTES); sz = lds32f(a_w + 2 * Cfg::PLANE_BYTES);
    }
    if (j < rounds) box_test(sx, sy, sz);
    // off the critical path: un-rank the winner (rank 0 <-> index 0)
    if (t == 0) out[j] = idx_base + (ghi >= 0 ? tie_unrank(0x7fffu - (glo >> 14), log2bs, q_count) : 0);
  }
}

template <int NBK, int NW>
static int launch_rw(int blocks, const float* xyz, const int* offset, const int* new_offset, int N, int M,
                     int* idx, int log2bs, int mag_skip, cudaStream_t s) {
  constexpr int smem = RwCfg<NBK, NW>::SMEM_BYTES;
  // per device, not per process: the attribute belongs to the device's context
  cudaError_t e = cudaFuncSetAttribute(fps_rw_kernel<NBK, NW>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) { set_error("fps: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return GEOT_ERR_CUDA; }
  fps_rw_kernel<NBK, NW><<<blocks, NW * 32, smem, s>>>(xyz, offset, new_offset, N, M, idx, log2bs, mag_skip);
  count_launch(1);
  return check_launch("fps_rw_kernel");
}

template <int NW>
static int launch_rw_nbk(int nbk32, int blocks, const float* xyz, const int* offset, const int* new_offset, int N, int M,
                         int* idx, int log2bs, int mag_skip, cudaStream_t s) {
  switch (nbk32) {
    case 1: return launch_rw<32, NW>(blocks, xyz, offset, new_offset, N, M, idx, log2bs, mag_skip, s);
    case 2: return launch_rw<64, NW>(blocks, xyz, offset, new_offset, N, M, idx, log2bs, mag_skip, s);
    case 3: return launch_rw<96, NW>(blocks, xyz, offset, new_offset, N, M, idx, log2bs, mag_skip, s);
    default: return launch_rw<128, NW>(blocks, xyz, offset, new_offset, N, M, idx, log2bs, mag_skip, s);
  }
}

// n_max <= 16384.  nw: warps per cloud (8 default; 16 for A-B timing).
int fps_rw_dispatch(int blocks, int n_max, int nw, const float* xyz, const int* offset, const int* new_offset, int N,
                    int M, int* idx, int log2bs, int mag_skip, cudaStream_t s) {
  const int nbk32 = (n_max + 4095) / 4096;
  if (nw == 16) return launch_rw_nbk<16>(nbk32, blocks, xyz, offset, new_offset, N, M, idx, log2bs, mag_skip, s);
  return launch_rw_nbk<8>(nbk32, blocks, xyz, offset, new_offset, N, M, idx, log2bs, mag_skip, s);
}

}  // namespace geot
