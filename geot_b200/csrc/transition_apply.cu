// geot_b200/csrc/transition_apply.cu — per-point application of the blended transition matrix, sm_100a.
//
// Replaces examples/segmentation/train.py:549-552 and its autograd:
//     newT = F.normalize(lam * ema_t_corr + (1 - lam) * insT, p=1, dim=2)          (BN,C,C) temp #1, #2
//     out  = bmm(logits.permute(0,2,1).view(-1,C).unsqueeze(1), newT).squeeze(1)    batched 1xC . CxC
//     out  = out.view(B,N,C).permute(0,2,1).contiguous()
// The reference streams the (BN,C,C) tensor through HBM four times forward (blend, norm, divide, bmm) and
// keeps two full-size temporaries for autograd.  Here one warp owns a point: T_n is read ONCE (coalesced,
// 4*C*C contiguous bytes), blended into a per-warp shared tile, row L1 norms and the 1xC . CxC product come
// from that tile, and the (B,C,N) <-> point-major permutes are folded into a 32-point CTA tile so that
// logits / grad_out loads and out / grad_logits stores are 128-byte rows.  HBM-bound: 4*C*C bytes per point
// forward, 8*C*C backward (read insT, write grad_insT); nothing else of that size is materialised.
#include <algorithm>
#include "common.cuh"

namespace geot {

void count_launch(int n);
bool direct_kernels_forced();  // geot_group_points_config (group.cu): A/B switch, also selects the warp-per-point kernel here
static inline bool apply_force_warp_kernel() { return direct_kernels_forced(); }

constexpr int kApplyWarps = 8;
constexpr int kApplyThreads = kApplyWarps * 32;
constexpr int kApplyTile = 32;  // points per CTA tile (4 per warp)

struct ApplySmem {
  // floats: E[C*C] | lt[C][33] | gt[C][33] | ot[C][33] | per warp { sv[nbuf][C*Cp] | ca[64] | cb[64] }
  int C, Cp, nbuf;  // nbuf = 2: the next point's matrix is prefetched while this one is used (needs 2 tiles per warp)
  __host__ __device__ ApplySmem(int c, int nb) : C(c), Cp(c | 1), nbuf(nb) {}
  __host__ __device__ size_t e_off() const { return 0; }
  __host__ __device__ size_t lt_off() const { return ((size_t)C * C + 3) & ~(size_t)3; }
  __host__ __device__ size_t gt_off() const { return lt_off() + (size_t)C * 33; }
  __host__ __device__ size_t ot_off() const { return gt_off() + (size_t)C * 33; }
  __host__ __device__ size_t warp_off(int w) const { return ot_off() + (size_t)C * 33 + (size_t)w * ((size_t)nbuf * C * Cp + 128); }
  __host__ __device__ size_t bytes() const { return warp_off(kApplyWarps) * sizeof(float); }
};

// BWD == false: computes out.  BWD == true: computes grad_logits and/or grad_insT from grad_out.
template <bool BWD>
__global__ void __launch_bounds__(kApplyThreads)
transition_apply_kernel(int B, int C, int N, int nbuf, const float* __restrict__ logits, const float* __restrict__ ema_t,
                        const float* __restrict__ insT, float lam, const float* __restrict__ grad_out,
                        float* __restrict__ out_or_gl, float* __restrict__ grad_insT) {
  extern __shared__ __align__(16) float apply_smem[];
  const ApplySmem L(C, nbuf);
  const int Cp = L.Cp, C2 = C * C;
  float* E_s = apply_smem + L.e_off();    // lam * ema_t
  float* lt = apply_smem + L.lt_off();    // logits tile      [kk][pt]
  float* gt = apply_smem + L.gt_off();    // grad_out tile    [j][pt]            (BWD)
  float* ot = apply_smem + L.ot_off();    // out / grad_logits tile [c][pt]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* svbuf = apply_smem + L.warp_off(warp);  // two T_n buffers: the next point's matrix lands while this one is used
  float* ca = svbuf + (size_t)nbuf * C * Cp;
  float* cb = ca + 64;
  // T_n (4*C*C contiguous bytes, only 4-byte aligned for odd C) -> padded [kk][Cp] layout, asynchronously
  auto prefetch = [&](int64_t g, float* buf) {
    const float* __restrict__ T = insT + g * C2;
    int kk = 0, j = lane;
    while (j >= C) { j -= C; kk++; }
    for (int f = lane; f < C2; f += 32) {
      const unsigned dst = (unsigned)__cvta_generic_to_shared(buf + kk * Cp + j);
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(T + f) : "memory");
      j += 32;
      while (j >= C) { j -= C; kk++; }
    }
  };
  const int64_t BN = (int64_t)B * N;
  const float oml = 1.0f - lam;

  for (int e = threadIdx.x; e < C2; e += kApplyThreads) E_s[e] = lam * ema_t[e];

  const int64_t ntiles = (BN + kApplyTile - 1) / kApplyTile;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t g0 = tile * kApplyTile;
    if (nbuf == 2 && g0 + warp < BN) prefetch(g0 + warp, svbuf);  // flies under the tile loads below
    asm volatile("cp.async.commit_group;" ::: "memory");
    __syncthreads();  // previous tile's ot fully stored; E_s visible on the first pass
    for (int e = threadIdx.x; e < C * kApplyTile; e += kApplyThreads) {
      const int c = e >> 5, pt = e & 31;
      const int64_t g = g0 + pt;
      float lv = 0.f, gv = 0.f;
      if (g < BN) {
        const int64_t b = g / N, n = g - b * N;
        lv = __ldg(logits + (b * C + c) * N + n);
        if (BWD) gv = __ldg(grad_out + (b * C + c) * N + n);
      }
      lt[c * 33 + pt] = lv;
      if (BWD) gt[c * 33 + pt] = gv;
    }
    __syncthreads();

    for (int pt = warp, it = 0; pt < kApplyTile; pt += kApplyWarps, it++) {
      const int64_t g = g0 + pt;
      if (g >= BN) break;  // warp-uniform
      float* sv = svbuf + (nbuf == 2 ? (size_t)(it & 1) * C * Cp : 0);
      if (nbuf == 2) {
        if (pt + kApplyWarps < kApplyTile && g + kApplyWarps < BN) prefetch(g + kApplyWarps, svbuf + (size_t)((it + 1) & 1) * C * Cp);
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 1;" ::: "memory");  // this point's matrix has landed (own copies; same lane blends them)
      } else {  // large C: one tile per warp, no overlap
        prefetch(g, sv);
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 0;" ::: "memory");
      }
      {  // blend in place: sv[kk][j] = lam*E + (1-lam)*T
        int kk = 0, j = lane;
        while (j >= C) { j -= C; kk++; }
        for (int f = lane; f < C2; f += 32) {
          sv[kk * Cp + j] = __fmaf_rn(oml, sv[kk * Cp + j], E_s[f]);
          j += 32;
          while (j >= C) { j -= C; kk++; }
        }
      }
      __syncwarp();
      for (int kk = lane; kk < C; kk += 32) {  // row L1 norms (F.normalize: x / max(||x||_1, 1e-12))
        const float* row = sv + kk * Cp;
        float s = 0.f, d = 0.f;
        for (int j = 0; j < C; j++) {
          const float v = row[j];
          s += fabsf(v);
          if (BWD) d = __fmaf_rn(gt[j * 33 + pt], v, d);
        }
        const float inv = 1.0f / fmaxf(s, 1e-12f);
        const float l = lt[kk * 33 + pt];
        if (BWD) {
          ot[kk * 33 + pt] = d * inv;      // grad_logits[kk] = sum_j go[j] * newT[kk][j]
          ca[kk] = oml * l * inv;          // d newT -> d insT scale
          cb[kk] = d * inv;                // sum_m go[m] * newT[kk][m]
        } else {
          ca[kk] = l * inv;
        }
      }
      __syncwarp();
      if (!BWD) {
        for (int j = lane; j < C; j += 32) {
          float o = 0.f;
          for (int kk = 0; kk < C; kk++) o = __fmaf_rn(ca[kk], sv[kk * Cp + j], o);
          ot[j * 33 + pt] = o;
        }
      } else if (grad_insT != nullptr) {
        float* __restrict__ gT = grad_insT + g * C2;
        int kk = 0, j = lane;
        while (j >= C) { j -= C; kk++; }
        for (int f = lane; f < C2; f += 32) {
          const float v = sv[kk * Cp + j];
          const float sgn = v > 0.f ? 1.0f : (v < 0.f ? -1.0f : 0.f);
          gT[f] = ca[kk] * (gt[j * 33 + pt] - sgn * cb[kk]);
          j += 32;
          while (j >= C) { j -= C; kk++; }
        }
      }
      __syncwarp();
    }
    __syncthreads();
    if (out_or_gl != nullptr) {
      for (int e = threadIdx.x; e < C * kApplyTile; e += kApplyThreads) {
        const int c = e >> 5, pt = e & 31;
        const int64_t g = g0 + pt;
        if (g < BN) {
          const int64_t b = g / N, n = g - b * N;
          out_or_gl[(b * C + c) * N + n] = ot[c * 33 + pt];
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Tile kernel (default for the class counts it is instantiated for: odd C, 16-byte aligned insT).
// The warp-per-point kernel above spends ~1100 warp instructions per point on index bookkeeping, 33-wide loops
// on 32 lanes and shared-memory round trips, and runs at 15-20 % of the HBM rate.  Here a LANE owns a point:
// a persistent CTA (8 warps) pulls tiles of 16 consecutive points - 16*C*C contiguous floats - through a
// double-buffered shared tile by bulk async copies (mbarrier completion), lane (p, h) walks rows kk of point p
// with the blended row in REGISTERS (C is a template parameter): row L1 norm, the 1xC . CxC partial product
// (forward) or the dot with grad_out and the grad_insT row (backward, written IN PLACE into the tile, which then
// leaves by one bulk async store).  Point stride C*C is odd, so the 16 points of a half-warp hit 16 different
// banks and the two halves (rows 16 apart) the other 16.  The sixteen row groups' partial outputs are summed in a
// fixed order.  HBM-bound: 4*C*C bytes per point forward, 8*C*C backward.
// ---------------------------------------------------------------------------------------------
constexpr int kTileP = 16;      // points per tile
constexpr int kTileThreads = 256;
constexpr int kTileGroups = 2 * kTileThreads / 32;  // row groups: (warp, half-warp)
__device__ __forceinline__ void ta_mbar_init(uint64_t* bar) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void ta_mbar_wait(uint64_t* bar, unsigned parity) {
  asm volatile(
      "{\n.reg .pred p;\nTA_WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@!p bra TA_WAIT_%=;\n}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void ta_bulk_load(float* dst, const float* src, unsigned bytes, uint64_t* bar) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
  for (unsigned o = 0; o < bytes; o += 8192u) {
    const unsigned n = min(8192u, bytes - o);
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(reinterpret_cast<const unsigned char*>(dst) + o)), "l"(reinterpret_cast<const unsigned char*>(src) + o),
                   "r"(n), "r"(smem_u32(bar)) : "memory");
  }
}
__device__ __forceinline__ void ta_bulk_store(float* dst, const float* src, unsigned bytes) {
  for (unsigned o = 0; o < bytes; o += 8192u) {
    const unsigned n = min(8192u, bytes - o);
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                 ::"l"(reinterpret_cast<unsigned char*>(dst) + o), "r"(smem_u32(reinterpret_cast<const unsigned char*>(src) + o)), "r"(n) : "memory");
  }
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

template <int C>
struct TileSmem {
  static constexpr int C2 = C * C;
  static constexpr int kTile = ((kTileP * C2 + 3) / 4) * 4;             // floats per tile buffer
  static constexpr int kRowT = C * (kTileP + 1);                        // one [C][17] tile
  static constexpr int kOffE = 2 * kTile;                               // lam * ema_t
  static constexpr int kOffL = kOffE + ((C2 + 3) / 4) * 4;              // lt[2][C][17]
  static constexpr int kOffG = kOffL + 2 * kRowT;                       // gt[2][C][17]
  static constexpr int kOffO = kOffG + 2 * kRowT;                       // ot[C][17]
  static constexpr int kOffP = kOffO + kRowT;                           // part[16][16][C]  (forward)
  static constexpr int kFloats = kOffP + kTileGroups * kTileP * C;
  static constexpr size_t kBytes = (size_t)kFloats * sizeof(float);
};

template <int C, bool BWD>
__global__ void __launch_bounds__(kTileThreads, 1)
transition_apply_tile_kernel(int B, int N, const float* __restrict__ logits, const float* __restrict__ ema_t,
                             const float* __restrict__ insT, float lam, const float* __restrict__ grad_out,
                             float* __restrict__ out_or_gl, float* __restrict__ grad_insT) {
  using L = TileSmem<C>;
  constexpr int C2 = C * C, LP = kTileP + 1;
  extern __shared__ __align__(128) float ta_smem[];
  __shared__ uint64_t bar[2];
  float* E_s = ta_smem + L::kOffE;
  float* ot = ta_smem + L::kOffO;
  float* part = ta_smem + L::kOffP;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5, p = lane & 15, h = lane >> 4;
  const int64_t BN = (int64_t)B * N;
  const int64_t ntiles = (BN + kTileP - 1) / kTileP;
  const float oml = 1.0f - lam;
  if (t == 0) {
    ta_mbar_init(&bar[0]);
    ta_mbar_init(&bar[1]);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (int e = t; e < C2; e += kTileThreads) E_s[e] = lam * ema_t[e];
  __syncthreads();
  // Tile `tile` -> buffers `buf`: the matrices by bulk async copy (thread 0; full tiles only, 16*C*C*4 bytes is a
  // multiple of 16), the logits / grad_out rows of its 16 points by 4-byte async copies (all threads; thread t
  // always serves point t & 15, so the (b, n) split of the point index is computed once per tile).
  auto issue = [&](int64_t tile, int buf) {
    if (t == 0 && (tile + 1) * kTileP <= BN)
      ta_bulk_load(ta_smem + buf * L::kTile, insT + tile * kTileP * C2, (unsigned)(kTileP * C2 * 4), &bar[buf]);
    const int pt = t & 15;
    const int64_t g = tile * kTileP + pt;
    if (g < BN) {
      const int64_t b = g / N, n = g - b * N;
      float* lt = ta_smem + L::kOffL + buf * L::kRowT;
      float* gt = ta_smem + L::kOffG + buf * L::kRowT;
      for (int c = t >> 4; c < C; c += kTileThreads / 16) {
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(lt + c * LP + pt)), "l"(logits + (b * C + c) * N + n) : "memory");
        if (BWD) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(gt + c * LP + pt)), "l"(grad_out + (b * C + c) * N + n) : "memory");
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  int64_t tile = blockIdx.x;
  if (tile < ntiles) issue(tile, 0);
  for (int it = 0; tile < ntiles; tile += gridDim.x, it++) {
    const int cur = it & 1;
    float* Tb = ta_smem + cur * L::kTile;
    const float* lt = ta_smem + L::kOffL + cur * L::kRowT;
    const float* gt = ta_smem + L::kOffG + cur * L::kRowT;
    const int64_t g0 = tile * kTileP;
    const int npts = (int)min((int64_t)kTileP, BN - g0);
    const bool full = npts == kTileP;
    const int64_t nxt = tile + gridDim.x;
    if (nxt < ntiles) {
      if (BWD && t == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // the store that still reads the other buffer
      issue(nxt, cur ^ 1);
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    if (!full)  // ragged last tile: its byte count need not be a multiple of 16
      for (int e = t; e < npts * C2; e += kTileThreads) Tb[e] = __ldg(insT + g0 * C2 + e);
    __syncthreads();
    if (full) ta_mbar_wait(&bar[cur], (unsigned)((it >> 1) & 1));

    if (p < npts) {
      float* Tp = Tb + p * C2;
      float acc[C];  // forward: partial out[j]; backward: grad_out[j] of this point
#pragma unroll
      for (int j = 0; j < C; j++) acc[j] = BWD ? gt[j * LP + p] : 0.f;
#pragma unroll 1
      for (int t4 = 0; t4 < 2 * ((C + 31) / 32); t4++) {
        const int kk = (t4 >> 1) * 32 + warp + 8 * (t4 & 1) + 16 * h;
        if (kk < C) {
          float v[C];
          float s = 0.f, d = 0.f;
#pragma unroll
          for (int j = 0; j < C; j++) {
            v[j] = __fmaf_rn(oml, Tp[kk * C + j], E_s[kk * C + j]);
            s += fabsf(v[j]);
            if (BWD) d = __fmaf_rn(acc[j], v[j], d);
          }
          const float inv = 1.0f / fmaxf(s, 1e-12f);  // F.normalize: x / max(||x||_1, 1e-12)
          const float l = lt[kk * LP + p];
          if (!BWD) {
            const float ca = l * inv;
#pragma unroll
            for (int j = 0; j < C; j++) acc[j] = __fmaf_rn(ca, v[j], acc[j]);
          } else {
            const float cb = d * inv;  // sum_m go[m] * newT[kk][m] = grad_logits[kk]
            ot[kk * LP + p] = cb;
            if (grad_insT != nullptr) {
              const float ca = oml * l * inv;
#pragma unroll
              for (int j = 0; j < C; j++) {
                const float sgn = v[j] > 0.f ? 1.0f : (v[j] < 0.f ? -1.0f : 0.f);
                Tp[kk * C + j] = ca * (acc[j] - sgn * cb);  // in place: this lane is the only user of the row
              }
            }
          }
        }
      }
      if (!BWD) {
#pragma unroll
        for (int j = 0; j < C; j++) part[((warp * 2 + h) * kTileP + p) * C + j] = acc[j];
      }
    }
    if (BWD && grad_insT != nullptr) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (!BWD || out_or_gl != nullptr) {
      const int pt = t & 15;
      if (pt < npts) {
        const int64_t g = g0 + pt, b = g / N, n = g - b * N;
        for (int c = t >> 4; c < C; c += kTileThreads / 16) {
          float o;
          if (BWD) o = ot[c * LP + pt];
          else {
            o = 0.f;
#pragma unroll
            for (int q = 0; q < kTileGroups; q++) o += part[(q * kTileP + pt) * C + c];
          }
          out_or_gl[(b * C + c) * N + n] = o;
        }
      }
    }
    if (BWD && grad_insT != nullptr) {
      if (full) {
        if (t == 0) ta_bulk_store(grad_insT + g0 * C2, Tb, (unsigned)(kTileP * C2 * 4));
      } else {
        for (int e = t; e < npts * C2; e += kTileThreads) grad_insT[g0 * C2 + e] = Tb[e];
      }
    }
    __syncthreads();  // ot / part are rewritten by the next tile
  }
  if (BWD && t == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <int C, bool BWD>
static int launch_apply_tile(int B, int N, const float* logits, const float* ema_t, const float* insT, float lam,
                             const float* grad_out, float* out_or_gl, float* grad_insT, cudaStream_t s) {
  static DynSmem ds;  // per instantiation
  if (int rc = ds.need(transition_apply_tile_kernel<C, BWD>, TileSmem<C>::kBytes, "transition_apply")) return rc;
  const int64_t tiles = ((int64_t)B * N + kTileP - 1) / kTileP;
  const int per_sm = TileSmem<C>::kBytes > 110 * 1024 ? 1 : (TileSmem<C>::kBytes > 72 * 1024 ? 2 : 3);
  const int grid = (int)std::min<int64_t>(tiles, (int64_t)num_sms() * per_sm);
  transition_apply_tile_kernel<C, BWD><<<grid, kTileThreads, TileSmem<C>::kBytes, s>>>(B, N, logits, ema_t, insT, lam, grad_out, out_or_gl, grad_insT);
  count_launch(1);
  return check_launch(BWD ? "transition_apply_tile_kernel<bwd>" : "transition_apply_tile_kernel<fwd>");
}

template <bool BWD>
static int launch_apply(int B, int C, int N, const float* logits, const float* ema_t, const float* insT,
                        float lam, const float* grad_out, float* out_or_gl, float* grad_insT, cudaStream_t s) {
  if (!apply_force_warp_kernel() && (((uintptr_t)insT) & 15) == 0 && (grad_insT == nullptr || (((uintptr_t)grad_insT) & 15) == 0)) {
    switch (C) {  // the class counts of the reference configs (17), BASELINE config 4 (33) and the small test shape
      case 5: return launch_apply_tile<5, BWD>(B, N, logits, ema_t, insT, lam, grad_out, out_or_gl, grad_insT, s);
      case 17: return launch_apply_tile<17, BWD>(B, N, logits, ema_t, insT, lam, grad_out, out_or_gl, grad_insT, s);
      case 33: return launch_apply_tile<33, BWD>(B, N, logits, ema_t, insT, lam, grad_out, out_or_gl, grad_insT, s);
      default: break;
    }
  }
  int nbuf = 2;
  if (ApplySmem(C, 2).bytes() > (size_t)112 * 1024) nbuf = 1;  // keep two CTAs per SM
  const ApplySmem L(C, nbuf);
  const size_t smem = L.bytes();
  static DynSmem ds;  // per instantiation
  if (int rc = ds.need(transition_apply_kernel<BWD>, smem, "transition_apply")) return rc;
  const int64_t tiles = ((int64_t)B * N + kApplyTile - 1) / kApplyTile;
  const int per_sm = smem > 110 * 1024 ? 1 : (smem > 72 * 1024 ? 2 : (smem > 54 * 1024 ? 3 : 4));
  const int64_t cap = (int64_t)num_sms() * per_sm;
  const int grid = (int)(tiles < cap ? tiles : cap);
  transition_apply_kernel<BWD><<<grid, kApplyThreads, smem, s>>>(B, C, N, nbuf, logits, ema_t, insT, lam, grad_out, out_or_gl, grad_insT);
  count_launch(1);
  return check_launch(BWD ? "transition_apply_kernel<bwd>" : "transition_apply_kernel<fwd>");
}

}  // namespace geot

using namespace geot;

GEOT_API int geot_transition_apply_forward(int B, int C, int N, const float* logits, const float* ema_t,
                                           const float* insT, float lam, float* out, geot_stream_t stream) {
  GEOT_REQUIRE(B >= 0 && N >= 0 && C >= 2 && C <= 64, "transition_apply_forward: bad sizes B=%d C=%d N=%d (2 <= C <= 64)", B, C, N);
  if (B == 0 || N == 0) return GEOT_OK;
  GEOT_REQUIRE(logits && ema_t && insT && out, "transition_apply_forward: NULL pointer");
  return launch_apply<false>(B, C, N, logits, ema_t, insT, lam, nullptr, out, nullptr, as_stream(stream));
}

GEOT_API int geot_transition_apply_backward(int B, int C, int N, const float* logits, const float* ema_t,
                                            const float* insT, float lam, const float* grad_out,
                                            float* grad_logits, float* grad_insT, geot_stream_t stream) {
  GEOT_REQUIRE(B >= 0 && N >= 0 && C >= 2 && C <= 64, "transition_apply_backward: bad sizes B=%d C=%d N=%d (2 <= C <= 64)", B, C, N);
  if (B == 0 || N == 0 || (grad_logits == nullptr && grad_insT == nullptr)) return GEOT_OK;
  GEOT_REQUIRE(logits && ema_t && insT && grad_out, "transition_apply_backward: NULL pointer");
  return launch_apply<true>(B, C, N, logits, ema_t, insT, lam, grad_out, grad_logits, grad_insT, as_stream(stream));
}
