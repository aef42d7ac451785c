// NOT COMPILED INTO THE LIBRARY.  A second measured alternative to fps.cu's manager/worker kernel, kept as evidence
// (profiles/r02_fps_study.md section 6): bit-exact on all 27 FPS parity cases, but SLOWER - 0.787 vs 0.557 ms at 64 x 16384 -> 1024,
// 2.92 vs 2.01 ms at 16 x 16384 -> 4096, 5.68 vs 3.89 ms at 6 x 16000 -> 8192 (profiles/r02b_fps_own_ab.jsonl).  32 warps that each
// issue ~100 instructions per round (box test, best-of-4 keys, publish, barrier, 32-key pick, coordinates) keep the four schedulers
// busy for ~800 cycles per round; the manager/worker kernel's idle workers issue ~10.  To rebuild: copy into geot_b200/csrc/ and call
// fps_own_dispatch from fps_dispatch (flag GEOT_FPS_OWN = 32 in the history of this file's commit).
// fps_own.cu - bucket-pruned FPS, every warp owns its buckets: ONE barrier per round.  sm_100a.
//
// Replaces (bit-exact indices) the same three reference kernels as fps.cu:
//   openpoints/cpp/pointnet2_batch/src/sampling_gpu.cu:101-216, pointops/src/sampling/sampling_cuda_kernel.cu:15-129,
//   pointnet2/_ext_src/src/sampling_gpu.cu:73-176.
//
// fps.cu's manager/worker kernel spends a round on: manager pick + box tests of all buckets -> named barrier -> the
// slowest worker's bucket updates -> named barrier (~1060 cycles at N=16384, M=1024; profiles/r02_fps_study.md).  The
// first one-barrier alternative (tools/experiments/fps_rw.cu) let every warp repeat the pick and the box tests of ALL
// buckets and lost to the redundant work.  Here the work is partitioned instead of repeated:
//   * 32 warps, warp w owns buckets b = c*32 + w (c < NBK/32 <= 4; Morton-adjacent buckets go to different warps):
//     their min-distances (registers, 4 points per lane per bucket), tie ranks, bounding boxes (lane l keeps the box of
//     owned bucket l % CHW) and candidate keys (every lane keeps all CHW of them);
//   * per round a warp box-tests ONLY its own buckets against the new sample (one ballot), updates the ones that can
//     change, takes the maximum of its CHW keys in registers and publishes ONE 64-bit key;
//   * ONE __syncthreads; then every warp reads the 32 published keys (one per lane), takes their maximum with two
//     REDUX and fetches the winner's coordinates itself (three broadcast LDS) - 32 keys instead of 128, no box tests
//     for foreign buckets, no second hand-off.
// Published keys are double-buffered by round parity: round j writes buf[j&1], the barrier of round j+1 separates the
// last read of buf[j&1] from its next write in round j+2.
// The pruning is the exact one of fps.cu: box distance rounded like dist2_ref (monotone in |gap|), so a bucket whose
// bound is not below its current maximum cannot change.  Ties: the reference's rank of the ORIGINAL index.
#include "fps_common.cuh"

namespace geot {

template <int NBK>
struct OwnCfg {
  static constexpr int NW = 32, NT = 1024;
  static constexpr int CHW = NBK / NW;   // buckets owned by a warp
  static constexpr int CAP = NBK * 128;
  static constexpr int PPT = CAP / NT;   // points per thread during setup
  static constexpr int PLANE_BYTES = NBK * 512;
  static constexpr int HIST_OFF = 3 * PLANE_BYTES;     // 4096 cell counters (setup only)
  static constexpr int KEY_OFF = HIST_OFF + 4096 * 4;  // 2 parity buffers of 32 warp keys
  static constexpr int SMEM_BYTES = KEY_OFF + 2 * NW * 8;
  static_assert(NBK % 32 == 0 && NBK >= 32 && NBK <= 128, "bucket / warp layout");
};

template <int NBK>
__global__ void __launch_bounds__(1024, 1)
fps_own_kernel(const float* __restrict__ xyz, const int* __restrict__ offset, const int* __restrict__ new_offset,
               int N, int M, int* __restrict__ idx, int log2bs, int mag_skip) {
  using Cfg = OwnCfg<NBK>;
  constexpr int NT = Cfg::NT, NW = Cfg::NW, CHW = Cfg::CHW, PPT = Cfg::PPT;
  extern __shared__ float4 fps_own_smem[];
  unsigned char* smem_raw = reinterpret_cast<unsigned char*>(fps_own_smem);
  int* hist = reinterpret_cast<int*>(smem_raw + Cfg::HIST_OFF);
  __shared__ float s_red[6][32];
  __shared__ int s_scan[32];

  const int t = threadIdx.x, lane = t & 31, w = t >> 5;
  int64_t start_n, start_m;
  int n, m, idx_base;
  fps_segment(offset, new_offset, N, M, blockIdx.x, start_n, n, start_m, m, idx_base);
  if (m <= 0) return;
  if (n > Cfg::CAP || n <= 0) {  // the caller understated n_max: mark, do not leave garbage
    for (int j = t; j < m; j += NT) idx[start_m + j] = -1;
    return;
  }
  const int q_count = (n + (1 << log2bs) - 1) >> log2bs;
  const float* __restrict__ src = xyz + start_n * 3;
  const float kInf = __int_as_float(0x7f800000);

  // ---- 1. the cloud into registers (one coalesced pass), bounding cube -> 16^3 Morton cells ------------------------
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
    if (lane == 0) { s_red[a][w] = lo[a]; s_red[3 + a][w] = hi[a]; }
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

  // ---- 2. counting sort by cell: histogram, exclusive scan, scatter from registers ---------------------------------
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
    if (lane == 31) s_scan[w] = inc;
    __syncthreads();
    int wtot = s_scan[lane], winc = wtot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o) winc += y;
    }
    const int wbase = __shfl_sync(0xffffffffu, winc - wtot, w);
    int run = wbase + inc - sum;
#pragma unroll
    for (int i = 0; i < PER; i++) { hist[t * PER + i] = run; run += v[i]; }
  }
  __syncthreads();

  unsigned a_base;  // opaque to the optimiser, or it re-derives the window base inside the round loop
  asm volatile("mov.u32 %0, %1;" : "=r"(a_base) : "r"(smem_u32(smem_raw)));
  const unsigned a_key = a_base + Cfg::KEY_OFF;
  const unsigned a_rank = a_base + 2 * Cfg::PLANE_BYTES;  // u16 rank table, lives in the z plane until z is scattered
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

  // ---- 3. every warp takes its buckets: tie ranks, min-distances (registers), box, first candidates --------------
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
  for (int i = 0; i < PPT; i++)
    if (t + i * NT < n) sts32(a_base + 2 * Cfg::PLANE_BYTES + 4u * spos[i], __float_as_uint(pz[i]));
  for (int pos = n + t; pos < Cfg::CAP; pos += NT) sts32(a_base + 2 * Cfg::PLANE_BYTES + 4u * pos, 0u);
  __syncthreads();
  const int c_mine = lane % CHW;  // the owned bucket whose box this lane keeps and tests
  float box[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  uint2 kown[CHW];                // candidate key of every owned bucket {lo: (~rank << 14) | sorted position, hi: value bits}
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
    if (c == c_mine) { box[0] = l3[0]; box[1] = l3[1]; box[2] = l3[2]; box[3] = h3[0]; box[4] = h3[1]; box[5] = h3[2]; }
    // all live points start at 1e10: the bucket's first candidate is its live point of least rank
    const unsigned wr = __reduce_min_sync(0xffffffffu, r);
    if (wr == 0xffffffffu) {
      kown[c] = make_uint2(0u, (unsigned)__float_as_int(-1.0f));
    } else {
      const int lw = __ffs(__ballot_sync(0xffffffffu, r == wr)) - 1;
      kown[c] = make_uint2(((0x7fffu - (wr >> 2)) << 14) | (unsigned)(b * 128 + 4 * lw) | (wr & 3u), (unsigned)__float_as_int(1e10f));
    }
  }
  if (rounds <= 0) return;

  // ---- 4. rounds --------------------------------------------------------------------------------------------------
  int* out = idx + start_m;
  float sx = x0, sy = y0, sz = z0;
#pragma unroll 1
  for (int j = 1; j <= rounds; j++) {
    // exact box test of MY buckets against the sample, rounded like dist2_ref (a true lower bound in fp32); a bucket
    // whose bound is not below its current maximum cannot change.  A bucket without candidates has maximum -1.
    unsigned kmax = kown[0].y;
#pragma unroll
    for (int c = 1; c < CHW; c++) kmax = c == c_mine ? kown[c].y : kmax;
    const float gx = fmaxf(fmaxf(__fsub_rn(box[0], sx), __fsub_rn(sx, box[3])), 0.f);
    const float gy = fmaxf(fmaxf(__fsub_rn(box[1], sy), __fsub_rn(sy, box[4])), 0.f);
    const float gz = fmaxf(fmaxf(__fsub_rn(box[2], sz), __fsub_rn(sz, box[5])), 0.f);
    const float lb = __fmaf_rn(gz, gz, __fmaf_rn(gx, gx, __fmul_rn(gy, gy)));
    const unsigned act = __ballot_sync(0xffffffffu, lb < __uint_as_float(kmax)) & ((1u << CHW) - 1u);  // lanes 0..CHW-1 hold distinct buckets
    if (act) {
#pragma unroll
      for (int c = 0; c < CHW; c++) {
        if ((act >> c) & 1u) {
          const int b = c * NW + w;
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
          // candidates are >= 0 or exactly -1, so the int view of the float bits is order preserving
          const int wm = __reduce_max_sync(0xffffffffu, __float_as_int(cm));
          unsigned r = 0xffffffffu;
#pragma unroll
          for (int e = 0; e < 4; e++) r = min(r, (__float_as_int(mind[c * 4 + e]) == wm) ? rk[e] : 0xffffffffu);
          const unsigned wr = __reduce_min_sync(0xffffffffu, r);
          // ranks are unique, so exactly one lane holds wr (an active bucket has a live point: wm >= 0)
          const int lw = __ffs(__ballot_sync(0xffffffffu, r == wr)) - 1;
          kown[c] = make_uint2(((0x7fffu - (wr >> 2)) << 14) | (unsigned)(b * 128 + 4 * lw) | (wr & 3u), (unsigned)wm);
        }
      }
    }
    // my best candidate -> buf[j & 1]; value bits with the sign flipped order -1 below every distance
    unsigned long long best = ((unsigned long long)(kown[0].y ^ 0x80000000u) << 32) | kown[0].x;
#pragma unroll
    for (int c = 1; c < CHW; c++) {
      const unsigned long long k64 = ((unsigned long long)(kown[c].y ^ 0x80000000u) << 32) | kown[c].x;
      best = k64 > best ? k64 : best;
    }
    const unsigned a_kw = a_key + (unsigned)(j & 1) * (NW * 8);
    if (lane == 0) sts64(a_kw + 8u * (unsigned)w, (unsigned)best, (unsigned)(best >> 32));
    __syncthreads();
    // every warp: the maximum of the 32 published keys, the winner's coordinates
    const uint2 pk = lds64(a_kw + 8u * (unsigned)lane);
    const int bh = (int)(pk.y ^ 0x80000000u);
    const int ghi = __reduce_max_sync(0xffffffffu, bh);
    const unsigned glo = __reduce_max_sync(0xffffffffu, bh == ghi ? pk.x : 0u);
    // no candidate at all (every point skipped): the reference keeps index 0
    sx = x0; sy = y0; sz = z0;
    if (ghi >= 0) {
      const unsigned a_w = a_base + 4u * (glo & 0x3fffu);  // by sorted position
      sx = lds32f(a_w); sy = lds32f(a_w + Cfg::PLANE_BYTES); sz = lds32f(a_w + 2 * Cfg::PLANE_BYTES);
    }
    // the winner's (~rank) for now: un-ranking (an integer division) stays out of the round chain
    if (t == 0) out[j] = ghi >= 0 ? (int)(glo >> 14) : -1;
  }
  __syncthreads();
  for (int j = 1 + t; j <= rounds; j += NT) {
    const int v = out[j];
    out[j] = idx_base + (v >= 0 ? tie_unrank(0x7fffu - (unsigned)v, log2bs, q_count) : 0);
  }
}

template <int NBK>
static int launch_own(int blocks, const float* xyz, const int* offset, const int* new_offset, int N, int M,
                      int* idx, int log2bs, int mag_skip, cudaStream_t s) {
  constexpr int smem = OwnCfg<NBK>::SMEM_BYTES;
  static DynSmem ds;  // per instantiation
  if (int rc = ds.need(fps_own_kernel<NBK>, smem, "fps")) return rc;
  fps_own_kernel<NBK><<<blocks, 1024, smem, s>>>(xyz, offset, new_offset, N, M, idx, log2bs, mag_skip);
  count_launch(1);
  return check_launch("fps_own_kernel");
}

// n_max <= 16384
int fps_own_dispatch(int blocks, int n_max, const float* xyz, const int* offset, const int* new_offset, int N, int M,
                     int* idx, int log2bs, int mag_skip, cudaStream_t s) {
  switch ((n_max + 4095) / 4096) {
    case 1: return launch_own<32>(blocks, xyz, offset, new_offset, N, M, idx, log2bs, mag_skip, s);
    case 2: return launch_own<64>(blocks, xyz, offset, new_offset, N, M, idx, log2bs, mag_skip, s);
    case 3: return launch_own<96>(blocks, xyz, offset, new_offset, N, M, idx, log2bs, mag_skip, s);
    default: return launch_own<128>(blocks, xyz, offset, new_offset, N, M, idx, log2bs, mag_skip, s);
  }
}

}  // namespace geot
