// geot_b200/csrc/group.cu — grouping / gather (channel-major, batched) and the deterministic
// scatter plan used by every backward pass.
//
// Replaces
//   openpoints/cpp/pointnet2_batch/src/group_points_gpu.cu:53-72 (fwd), :14-31 (bwd, atomicAdd :30)
//   openpoints/cpp/pointnet2_batch/src/sampling_gpu.cu:15-31 (gather fwd), :53-70 (bwd, atomicAdd :69)
//   pointnet2/_ext_src/src/group_points_gpu.cu:11-31,46-67 and sampling_gpu.cu:11-23,37-50 (same results)
//
// Forward: HBM-bound on the (B,C,M,ns) write.  The reference launches one thread per output element
// and one grid row per channel, so idx is re-read C times and every access is 4 bytes.  Here a thread
// owns 4 consecutive (j,s) entries: one int4 idx load kept in registers across a chunk of channels,
// four cached 4-byte gathers per channel (a (N) feature row is L1/L2 resident) and one 16-byte
// streaming store; 64-bit flat offsets throughout.
//
// Backward: the reference scatter-adds with float atomics (run-to-run different sums).  Here a
// "scatter plan" — CSR of source positions by target, sources in ascending order — is built once per
// idx with integer atomics + a stable ranking pass (bit-reproducible), and each target sums its
// sources in that fixed order.
#include <algorithm>
#include "common.cuh"

namespace geot {

void count_launch(int n);

constexpr int kGroupCCH = 16;  // channels per thread in the forward gather
// Kernel-selection override for tests / A-B timing, set by geot_group_points_config().  THREAD-LOCAL, like the error
// slot: it only affects calls made by the thread that set it, so the library stays re-entrant (no process-wide state).
static thread_local bool g_group_force_direct = false;  // mode 1
static thread_local bool g_force_staged = false;        // mode 2: ignore the work-size heuristics
static thread_local bool g_rows4_off = false;           // mode 3: three_interpolate without the channel-interleaved kernel (A/B)
static thread_local bool g_tip_off = false;             // mode 4: three_interpolate without the persistent interleaved kernel (A/B)
static thread_local bool g_mtile_off = false;           // mode 6: manifold loss on the warp-per-point kernels (A/B)
static thread_local bool g_mtile_on = false;            // mode 7: manifold loss on the tile kernels whatever C (tests)
static thread_local int g_pr_variant = -1;              // grouping backward: bit 0 half-row pipeline, bit 1 L2 prefetch of the next row, bit 2 grid of (channel group, cloud) CTAs instead of one CTA per SM; -1 = chosen per launch; modes 9 / 10 / 11 / 12 / 13 = 4 / 3 / 1 / 6 / 2 (A/B)
static thread_local bool g_nn_queue_off = false;        // mode 14: three_nn grid kernel finishing every query in place (A/B)
static thread_local bool g_plan_ordered_off = false;    // mode 15: scatter plan without the ordered builder (A/B)
static thread_local int g_tib_stream = 0;               // mode 8: three_interpolate backward streamed whatever the launch size (tests); mode 5: never (A/B)

// VEC = 4: E % 4 == 0 and 16-byte aligned idx/out.  grid (ceil(E/VEC/256), ceil(C/CCH), B).
template <int VEC>
__global__ void __launch_bounds__(256)
group_points_kernel(int C, int N, int64_t E, int cch, const float* __restrict__ points,
                    const int* __restrict__ idx, float* __restrict__ out) {
  const int b = blockIdx.z;
  const int c0 = blockIdx.y * cch;
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // entry group
  if (g * VEC >= E) return;
  const int c1 = min(C, c0 + cch);
  if (VEC == 4) {
    const int4 id = __ldg(reinterpret_cast<const int4*>(idx + (int64_t)b * E) + g);
    for (int c = c0; c < c1; c++) {
      const float* __restrict__ row = points + ((int64_t)b * C + c) * N;
      float4 o;
      o.x = __ldg(row + id.x); o.y = __ldg(row + id.y); o.z = __ldg(row + id.z); o.w = __ldg(row + id.w);
      __stcs(reinterpret_cast<float4*>(out + ((int64_t)b * C + c) * E) + g, o);
    }
  } else {
    const int id = __ldg(idx + (int64_t)b * E + g);
    for (int c = c0; c < c1; c++) {
      const float* __restrict__ row = points + ((int64_t)b * C + c) * N;
      __stcs(out + ((int64_t)b * C + c) * E + g, __ldg(row + id));
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Row-staged forward (the default when a feature row fits shared memory and is re-used: E >= 2N).
// The gather above pulls a 32-byte sector through L2/L1 for every 4-byte element: at C=64, ns=32 that is
// 8x the algorithmic source traffic and the kernel sits on the L1/L2 sector rate (42 % of HBM peak).
// Here one CTA owns one (b, c) row: the row (N floats) is brought into shared memory once by a single
// bulk async copy (cp.async.bulk, completion on an mbarrier), idx is streamed as int4 (L2 resident: it is
// shared by the C rows of the cloud), the 4-byte gathers hit shared memory and the output leaves as 16-byte
// streaming stores.  Up to three CTAs per SM, so one row's load overlaps the neighbours' gather/store phases.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@!p bra WAIT_%=;\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// One logical row copy issued as several bulk copies (all completing on the same mbarrier): a single large
// cp.async.bulk keeps too few lines in flight to pull a 128 KB row at the SM's share of HBM bandwidth.
__device__ __forceinline__ void bulk_g2s_chunked(float* dst, const float* src, unsigned bytes, uint64_t* bar) {
  constexpr unsigned kPiece = 8192;
  mbar_expect_tx(bar, bytes);
  for (unsigned o = 0; o < bytes; o += kPiece)
    bulk_g2s(reinterpret_cast<unsigned char*>(dst) + o, reinterpret_cast<const unsigned char*>(src) + o,
             min(kPiece, bytes - o), bar);
}

// Ask L2 for a row that a later bulk copy will read (no shared memory, no completion to wait for): the HBM read then
// overlaps whatever the CTA does meanwhile and the copy itself runs at L2 speed.
__device__ __forceinline__ void bulk_prefetch_l2(const float* src, unsigned bytes) {
  constexpr unsigned kPiece = 16384;
  for (unsigned o = 0; o < bytes; o += kPiece)
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;"
                 ::"l"(reinterpret_cast<const unsigned char*>(src) + o), "r"(min(kPiece, bytes - o)) : "memory");
}

constexpr int kRowThreads = 512;
__global__ void __launch_bounds__(kRowThreads)
group_rows_kernel(int C, int N, int64_t E, const float* __restrict__ points, const int* __restrict__ idx,
                  float* __restrict__ out) {
  extern __shared__ __align__(16) float g_row[];
  __shared__ uint64_t bar;
  const int t = threadIdx.x;
  const int c = blockIdx.x, b = blockIdx.y;
  const int64_t rowi = (int64_t)b * C + c;
  if (t == 0) {
    mbar_init(&bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (t == 0) {
    bulk_g2s_chunked(g_row, points + rowi * N, (unsigned)N * 4u, &bar);
  }
  const int4* __restrict__ idp = reinterpret_cast<const int4*>(idx + (int64_t)b * E);
  float4* __restrict__ op = reinterpret_cast<float4*>(out + rowi * E);
  const int64_t n4 = E >> 2;
  constexpr int U = 4;
  int4 id[U];
#pragma unroll
  for (int u = 0; u < U; u++) {
    const int64_t g = t + (int64_t)u * kRowThreads;
    id[u] = g < n4 ? __ldg(idp + g) : make_int4(0, 0, 0, 0);
  }
  mbar_wait(&bar, 0);
  for (int64_t g0 = t; g0 < n4; g0 += (int64_t)U * kRowThreads) {
    int4 cur[U];
#pragma unroll
    for (int u = 0; u < U; u++) cur[u] = id[u];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int64_t g = g0 + (int64_t)(U + u) * kRowThreads;
      id[u] = g < n4 ? __ldg(idp + g) : make_int4(0, 0, 0, 0);
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int64_t g = g0 + (int64_t)u * kRowThreads;
      if (g < n4) {
        float4 o;
        o.x = g_row[cur[u].x]; o.y = g_row[cur[u].y]; o.z = g_row[cur[u].z]; o.w = g_row[cur[u].w];
        __stcs(op + g, o);
      }
    }
  }
}

static int launch_group(int B, int C, int N, int64_t E, const float* points, const int* idx, float* out,
                        cudaStream_t s, const char* what) {
  if (B == 0 || C == 0 || E == 0) return GEOT_OK;
  GEOT_REQUIRE(points && idx && out, "%s: NULL pointer", what);
  GEOT_REQUIRE(B <= 65535 && (C + kGroupCCH - 1) / kGroupCCH <= 65535, "%s: B or C exceeds grid limits", what);
  const bool vec = (E % 4 == 0) && (((uintptr_t)idx & 15) == 0) && (((uintptr_t)out & 15) == 0);
  const size_t row_bytes = (size_t)N * 4;
  if (vec && !g_group_force_direct && E >= 2 * (int64_t)N && N % 4 == 0 && (((uintptr_t)points & 15) == 0) &&
      row_bytes <= 200 * 1024 && C <= 65535 * 1 && N > 0) {
    static DynSmem ds;
    if (int rc = ds.need(group_rows_kernel, 200 * 1024, what)) return rc;
    dim3 grid(C, B);
    group_rows_kernel<<<grid, kRowThreads, row_bytes, s>>>(C, N, E, points, idx, out);
    count_launch(1);
    return check_launch(what);
  }
  // channels per thread: kGroupCCH amortises the idx load, but a short index list (gather_operation: E = M) then
  // leaves most SMs without work - halve it until the grid has a few waves
  const int64_t xb = vec ? (E / 4 + 255) / 256 : (E + 255) / 256;
  int cch = kGroupCCH;
  while (cch > 1 && xb * ((C + cch - 1) / cch) * B < 4 * (int64_t)num_sms() && (C + cch / 2 - 1) / (cch / 2) <= 65535) cch >>= 1;
  // a sparse gather (gather_operation: 1 of 16 points of every row) is bound by DRAM sector fetches, not by the idx
  // loads: one channel per thread keeps the most independent 4-byte gathers in flight (the reference kernel's mapping,
  // sampling_gpu.cu:15-31, plus the int4 idx load and the 16-byte store)
  if (2 * E <= (int64_t)N && C <= 65535) cch = 1;
  dim3 grid((unsigned)xb, (C + cch - 1) / cch, B);
  if (vec) group_points_kernel<4><<<grid, 256, 0, s>>>(C, N, E, cch, points, idx, out);
  else group_points_kernel<1><<<grid, 256, 0, s>>>(C, N, E, cch, points, idx, out);
  count_launch(1);
  return check_launch(what);
}

// ---------------------------------------------------------------------------------------------
// Scatter plan:  idx (B,E) -> row_ptr (B,N+1), perm (B,E)   [workspace = row_ptr | perm]
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
plan_count_kernel(int N, int64_t E, const int* __restrict__ idx, int* __restrict__ row_ptr) {
  const int b = blockIdx.y;
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  const int k = idx[(int64_t)b * E + e];
  if ((unsigned)k < (unsigned)N) atomicAdd(row_ptr + (int64_t)b * (N + 1) + k + 1, 1);  // integer: order-free
}

// In-place inclusive scan of each (N+1)-row (counts were stored shifted by one => exclusive offsets).
__global__ void __launch_bounds__(1024)
plan_scan_kernel(int N, int* __restrict__ row_ptr, int* __restrict__ cursor) {
  __shared__ int s_warp[32];
  __shared__ int s_carry;
  int* row = row_ptr + (int64_t)blockIdx.x * (N + 1);
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  if (t == 0) s_carry = 0;
  __syncthreads();
  for (int base = 0; base <= N; base += 1024) {
    const int i = base + t;
    int v = (i <= N) ? row[i] : 0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += u;
    }
    if (lane == 31) s_warp[warp] = v;
    __syncthreads();
    int w = s_warp[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += u;
    }
    const int warp_prefix = __shfl_sync(0xffffffffu, w, max(warp - 1, 0));
    const int carry = s_carry;
    v += carry + (warp ? warp_prefix : 0);
    if (i <= N) { row[i] = v; cursor[(int64_t)blockIdx.x * (N + 1) + i] = v; }
    __syncthreads();
    if (t == 1023) s_carry = v;
    __syncthreads();
  }
}

// Fill.  The counting-sort scatter itself uses integer atomics on a cursor copy of row_ptr, so the order of the
// sources inside a target's segment is arbitrary after it (into the scratch list `tmp`); a second pass puts
// every segment into ascending source order, which makes the plan - and every sum computed from it - bit
// reproducible.  Segments are short (mean E/N, a few entries): a thread sorts a segment of up to kPlanSmall
// entries by insertion in local memory.  Longer segments (the first-hit padding of ball_query, the all-zero
// rows of empty balls) go to a work list and are sorted by a whole CTA through a bitmap over the source
// universe [0,E): set one bit per source, read the bits back in order - O(L + E/32), no quadratic case.
constexpr int kPlanSmall = 64;
__global__ void __launch_bounds__(256)
plan_fill_kernel(int N, int64_t E, const int* __restrict__ idx, int* __restrict__ cursor, int* __restrict__ tmp) {
  const int b = blockIdx.y;
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  const int k = idx[(int64_t)b * E + e];
  if ((unsigned)k < (unsigned)N) tmp[(int64_t)b * E + atomicAdd(cursor + (int64_t)b * (N + 1) + k, 1)] = (int)e;
}

__global__ void __launch_bounds__(256)
plan_sort_small_kernel(int N, int64_t E, const int* __restrict__ row_ptr, const int* __restrict__ tmp,
                       int* __restrict__ perm, int* __restrict__ heavy_count, int* __restrict__ heavy_list,
                       int heavy_cap) {
  const int b = blockIdx.y;
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= N) return;
  const int* __restrict__ rp = row_ptr + (int64_t)b * (N + 1);
  const int s = rp[k], L = rp[k + 1] - s;
  if (L <= 0) return;
  const int* __restrict__ src = tmp + (int64_t)b * E + s;
  int* __restrict__ dst = perm + (int64_t)b * E + s;
  if (L == 1) { dst[0] = src[0]; return; }
  if (L > kPlanSmall) {
    const int slot = atomicAdd(heavy_count, 1);  // order of the work list does not matter
    if (slot < heavy_cap) heavy_list[slot] = b * N + k;
    return;
  }
  int v[kPlanSmall];
  for (int i = 0; i < L; i++) {
    const int x = src[i];
    int j = i;
    while (j > 0 && v[j - 1] > x) { v[j] = v[j - 1]; j--; }
    v[j] = x;
  }
  for (int i = 0; i < L; i++) dst[i] = v[i];
}

constexpr int kPlanHeavyBits = 1 << 20;  // universe chunk: 1 Mi sources = 128 KB of shared memory
__global__ void __launch_bounds__(1024, 1)
plan_sort_heavy_kernel(int N, int64_t E, const int* __restrict__ row_ptr, const int* __restrict__ tmp,
                       int* __restrict__ perm, const int* __restrict__ heavy_count,
                       const int* __restrict__ heavy_list, int heavy_cap) {
  extern __shared__ unsigned plan_bm[];
  __shared__ int s_warp[32];
  __shared__ int s_written;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int count = min(*heavy_count, heavy_cap);
  for (int item = blockIdx.x; item < count; item += gridDim.x) {
    const int bk = heavy_list[item], b = bk / N, k = bk % N;
    const int* __restrict__ rp = row_ptr + (int64_t)b * (N + 1);
    const int s = rp[k], L = rp[k + 1] - s;
    const int* __restrict__ src = tmp + (int64_t)b * E + s;
    int* __restrict__ dst = perm + (int64_t)b * E + s;
    if (t == 0) s_written = 0;
    for (int64_t u0 = 0; u0 < E; u0 += kPlanHeavyBits) {
      const int nbits = (int)min((int64_t)kPlanHeavyBits, E - u0);
      const int nwords = (nbits + 31) >> 5;
      const int wpt = (nwords + 1023) >> 10;  // words per thread, consecutive: thread order = source order
      for (int i = t; i < wpt * 1024; i += 1024) plan_bm[i] = 0u;
      __syncthreads();
      for (int i = t; i < L; i += 1024) {
        const int64_t v = (int64_t)src[i] - u0;
        if (v >= 0 && v < nbits) atomicOr(&plan_bm[v >> 5], 1u << (v & 31));
      }
      __syncthreads();
      int cnt = 0;
      for (int i = 0; i < wpt; i++) cnt += __popc(plan_bm[t * wpt + i]);
      int inc = cnt;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += y;
      }
      if (lane == 31) s_warp[warp] = inc;
      __syncthreads();
      int w = s_warp[lane], winc = w;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, winc, o);
        if (lane >= o) winc += y;
      }
      const int base = s_written + __shfl_sync(0xffffffffu, winc - w, warp) + inc - cnt;
      const int chunk_total = __shfl_sync(0xffffffffu, winc, 31);
      int o = base;
      for (int i = 0; i < wpt; i++) {
        unsigned word = plan_bm[t * wpt + i];
        while (word) {
          dst[o++] = (int)(u0 + (((int64_t)(t * wpt + i)) << 5) + __ffs(word) - 1);
          word &= word - 1;
        }
      }
      __syncthreads();
      if (t == 0) s_written += chunk_total;
      __syncthreads();
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Fused plan build: the whole CSR of one cloud by ONE CTA in shared memory (counts -> scan -> fill -> per-segment
// sort), one launch instead of two memsets and five kernels.  In training every idx tensor is new each step and the
// reference API has no plan argument (the facades rebuild the plan per backward call), so the build is on the
// critical path of every backward: 0.20-0.28 ms with the multi-kernel path at the bench shape, launch- and
// global-atomic-bound.  Here the counters are shared-memory atomics (N+1 ints), the scatter goes straight into `perm`
// and every segment is then sorted in place: short ones by a thread (insertion, <= kPlanSmall entries), long ones
// (first-hit padding of ball_query, empty balls) by the whole CTA through a bitmap over the source universe [0,E).
// Same output as the multi-kernel path, bit for bit (ascending sources per target).  Limits: N <= kFusedMaxN targets
// per cloud, E <= kFusedMaxE sources per cloud; larger problems (the offset-format ops with one 10^5-point segment)
// keep the multi-kernel path, where the work amortises the launches.
// ---------------------------------------------------------------------------------------------
constexpr int kFusedMaxN = 40960;          // (N+1) counters: 160 KB
constexpr int kFusedMaxE = 1 << 19;        // bitmap: 64 KB
constexpr int kFusedThread = 2;            // a pair is ordered by its own thread; anything longer goes to the warp sorter
                                           // (an insertion sort in local memory costs ~60 cycles per step and the whole
                                           // warp waits for its longest segment: measured 70 us of the 85 us build)
constexpr int kFusedWarp = 512;            // the warp rank sort holds up to this many values (16 per lane)
#ifndef GEOT_PLAN_WARP_CUT
#define GEOT_PLAN_WARP_CUT 256
#endif
// segments up to this length: rank sort by one warp (L*L/32 shuffle steps); longer: the whole CTA through a bitmap (a
// few microseconds whatever the length).  Measured (tools/plan_ab.py): 128 / 512 are within 10 % of each other on the
// grouping plans of ball-query lists (pieces of <= 65535 sources); an UNSPLIT dense-ball list (every query lists the
// same ~80 lowest indices, ~800 sources each; 16 x 4096, E = 65536) takes 1.89 ms at 512 and 0.42 ms at 128.
constexpr int kFusedWarpCut = GEOT_PLAN_WARP_CUT;

// Warp-cooperative sort of one segment of L <= kFusedWarp DISTINCT values, in place: every lane holds up to 16 of
// them, rank(x) = #{y < x} by broadcasting every value to every lane; cost ~ (L/32)^2 * 32 shuffles.
#ifdef GEOT_FPS_PROF
__device__ long long g_plan_prof[16];
#define PLAN_STAMP(k) do { if (blockIdx.x == 0 && threadIdx.x == 0) g_plan_prof[k] = clock64(); } while (0)
#else
#define PLAN_STAMP(k) do {} while (0)
#endif

// Warp-cooperative sort of one segment of L <= kFusedWarp DISTINCT values, in place: every lane holds NV of them,
// rank(x) = #{y < x} by broadcasting every value to every lane; NV * 32 shuffles, NV * NV * 32 compares.
template <int NV, typename T>
__device__ __forceinline__ void plan_warp_sort_t(T* __restrict__ seg, int L, int lane) {
  int v[NV], r[NV];
#pragma unroll
  for (int i = 0; i < NV; i++) {
    v[i] = (lane + 32 * i < L) ? (int)seg[lane + 32 * i] : 0x7fffffff;
    r[i] = 0;
  }
#pragma unroll
  for (int j = 0; j < NV; j++) {
    const int live = min(32, L - 32 * j);  // lanes that hold a value of row j (uniform): short segments cost 2 L instructions
#pragma unroll 4
    for (int src = 0; src < live; src++) {
      const int y = __shfl_sync(0xffffffffu, v[j], src);
#pragma unroll
      for (int i = 0; i < NV; i++) r[i] += (y < v[i]) ? 1 : 0;
    }
  }
  __syncwarp();
#pragma unroll
  for (int i = 0; i < NV; i++)
    if (lane + 32 * i < L) seg[r[i]] = (T)v[i];
}
template <typename T>
__device__ __noinline__ void plan_warp_sort(T* __restrict__ seg, int L, int lane) {
  const int nv = (L + 31) >> 5;  // values per lane, uniform across the warp
  if (nv <= 1) plan_warp_sort_t<1>(seg, L, lane);
  else if (nv <= 2) plan_warp_sort_t<2>(seg, L, lane);
  else if (nv <= 3) plan_warp_sort_t<3>(seg, L, lane);
  else if (nv <= 4) plan_warp_sort_t<4>(seg, L, lane);
  else if (nv <= 6) plan_warp_sort_t<6>(seg, L, lane);
  else if (nv <= 8) plan_warp_sort_t<8>(seg, L, lane);
  else if (nv <= 12) plan_warp_sort_t<12>(seg, L, lane);
  else plan_warp_sort_t<16>(seg, L, lane);
}

// One thread sorts one segment of 3..16 values in registers: Batcher's odd-even merge network (63 compare-exchanges, no
// local memory, no divergence inside a warp - every lane runs the same 126 min/max instructions on its own segment).
template <typename T>
__device__ __forceinline__ void plan_thread_sort16(T* __restrict__ seg, int L) {
  int v[16];
#pragma unroll
  for (int i = 0; i < 16; i++) v[i] = i < L ? (int)seg[i] : 0x7fffffff;
#define PLAN_CE(a, b) { const int lo_ = min(v[a], v[b]), hi_ = max(v[a], v[b]); v[a] = lo_; v[b] = hi_; }
  PLAN_CE(0,1); PLAN_CE(2,3); PLAN_CE(0,2); PLAN_CE(1,3); PLAN_CE(1,2); PLAN_CE(4,5);
  PLAN_CE(6,7); PLAN_CE(4,6); PLAN_CE(5,7); PLAN_CE(5,6); PLAN_CE(0,4); PLAN_CE(2,6);
  PLAN_CE(2,4); PLAN_CE(1,5); PLAN_CE(3,7); PLAN_CE(3,5); PLAN_CE(1,2); PLAN_CE(3,4);
  PLAN_CE(5,6); PLAN_CE(8,9); PLAN_CE(10,11); PLAN_CE(8,10); PLAN_CE(9,11); PLAN_CE(9,10);
  PLAN_CE(12,13); PLAN_CE(14,15); PLAN_CE(12,14); PLAN_CE(13,15); PLAN_CE(13,14); PLAN_CE(8,12);
  PLAN_CE(10,14); PLAN_CE(10,12); PLAN_CE(9,13); PLAN_CE(11,15); PLAN_CE(11,13); PLAN_CE(9,10);
  PLAN_CE(11,12); PLAN_CE(13,14); PLAN_CE(0,8); PLAN_CE(4,12); PLAN_CE(4,8); PLAN_CE(2,10);
  PLAN_CE(6,14); PLAN_CE(6,10); PLAN_CE(2,4); PLAN_CE(6,8); PLAN_CE(10,12); PLAN_CE(1,9);
  PLAN_CE(5,13); PLAN_CE(5,9); PLAN_CE(3,11); PLAN_CE(7,15); PLAN_CE(7,11); PLAN_CE(3,5);
  PLAN_CE(7,9); PLAN_CE(11,13); PLAN_CE(1,2); PLAN_CE(3,4); PLAN_CE(5,6); PLAN_CE(7,8);
  PLAN_CE(9,10); PLAN_CE(11,12); PLAN_CE(13,14);
#undef PLAN_CE
#pragma unroll
  for (int i = 0; i < 16; i++)
    if (i < L) seg[i] = (T)v[i];
}

// Chunked builder for FEW targets with LONG source lists (three_interpolate's plan: 1024 coarse points, 48 sources
// each).  The fused builder below fills every segment through an atomic cursor and sorts it afterwards - one warp per
// segment, rank sort over shuffles, cost ~ L * ceil(L/32) - which is most of its 0.088 ms at the bench shape.  Here one
// CTA per cloud cuts the E sources into 32 contiguous chunks, one per warp, and keeps a 16-bit counter per
// (chunk, target) in shared memory (64 bytes per target), i.e. it counting-sorts by (target, chunk):
//   1. warp w counts its chunk (shared-memory atomics on packed 16-bit pairs);
//   2. a thread per target turns the 32 counts of its column into cursors inside the target's segment and the segment
//      lengths into row_ptr (block scan);
//   3. warp w fills its chunk's sources through the (w, t) cursors: the pieces of a segment are in chunk order, only
//      the 1-2 sources inside one piece (48 / 32 on average) are in arbitrary order;
//   4. a thread per target puts every piece in ascending order (insertion sort, pieces of 0-3 entries) - the segment
//      is then ascending as a whole: the order the multi-kernel and the fused builder produce, bit for bit;
//   5. the 16-bit list leaves by one coalesced pass.
// (A first version ranked equal targets inside a warp with match.any instead of atomics + step 4: correct, but
// match.any costs ~0.5 us per call with ~20 distinct values in the warp - 0.117 ms, slower than the fused builder.)
constexpr int kOrdMaxN = 2048;
constexpr size_t kOrdBudget = (size_t)220 * 1024;
constexpr int kOrdThreadSort = 16;    // longest (chunk, target) piece a single thread sorts
constexpr int kOrdBitmapWords = 66;   // a chunk has at most ceil(ceil(65535 / 32) / 32) * 32 = 2048 + 32 positions
constexpr int kOrdHeavyCap = 4096;    // list of longer pieces (E / (kOrdThreadSort + 1) <= 3855 of them can exist)
static size_t plan_ordered_smem(int N, int64_t E) {
  return (size_t)32 * ((N + 1) & ~1) * 2 + (size_t)(N + 1) * 4 + (size_t)((E + 1) / 2) * 4 + (size_t)32 * kOrdBitmapWords * 4 +
         (size_t)kOrdHeavyCap * 4;
}
__global__ void __launch_bounds__(1024, 1)
plan_build_ordered_kernel(int N, int E, const int* __restrict__ idx_all, int* __restrict__ row_ptr_all, int* __restrict__ perm_all) {
  extern __shared__ __align__(16) unsigned char po_smem[];
  __shared__ int s_scan[32];
  __shared__ int s_heavy;
  constexpr int heavy_cap = kOrdHeavyCap;
  const int Np = (N + 1) & ~1;  // counters per chunk row (even: a row is whole 32-bit words)
  unsigned short* cnt = reinterpret_cast<unsigned short*>(po_smem);               // [32][Np]
  int* rp = reinterpret_cast<int*>(po_smem + (size_t)32 * Np * 2);                // N + 1
  unsigned short* perm16 = reinterpret_cast<unsigned short*>(rp + N + 1);         // E
  unsigned* bitmap = reinterpret_cast<unsigned*>(perm16 + (((size_t)E + 1) & ~(size_t)1));  // 32 x kOrdBitmapWords
  unsigned* heavy = bitmap + 32 * kOrdBitmapWords;                                // kOrdHeavyCap {chunk:16 | target:16}
  const int t = threadIdx.x, lane = t & 31, w = t >> 5;
  const int b = blockIdx.x;
  const int* __restrict__ idx = idx_all + (int64_t)b * E;
  const int groups = (E + 31) >> 5, gpw = (groups + 31) >> 5;
  const int p0 = min(w * gpw * 32, E), p1 = min((w + 1) * gpw * 32, E);  // this warp's chunk of sources
  unsigned* my = reinterpret_cast<unsigned*>(cnt + (size_t)w * Np);
  for (int i = t; i < 16 * Np; i += 1024) reinterpret_cast<unsigned*>(cnt)[i] = 0u;
  __syncthreads();
  // ---- 1. counts per (chunk, target); an index outside [0, N) is not a target and is left out (the other builders
  //         require valid indices too)
#pragma unroll 4
  for (int p = p0 + lane; p < p1; p += 32) {
    const int tg = __ldg(idx + p);
    if ((unsigned)tg < (unsigned)N) atomicAdd(&my[tg >> 1], 1u << ((tg & 1) << 4));
  }
  __syncthreads();
  // ---- 2. column sums -> cursors inside the segment; segment lengths -> row_ptr
  {
    const int PER = (N + 1023) >> 10;  // consecutive targets per thread (<= 2)
    int len[2] = {0, 0}, sum = 0;
    for (int k = 0; k < PER; k++) {
      const int tt = t * PER + k;
      if (tt < N) {
        int run = 0;
#pragma unroll 8
        for (int ww = 0; ww < 32; ww++) {
          const int c = cnt[(size_t)ww * Np + tt];
          cnt[(size_t)ww * Np + tt] = (unsigned short)run;
          run += c;
        }
        len[k] = run;
        sum += run;
      }
    }
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
    int run = __shfl_sync(0xffffffffu, winc - wtot, w) + inc - sum;
    for (int k = 0; k < PER; k++) {
      const int tt = t * PER + k;
      if (tt < N) { rp[tt] = run; run += len[k]; }
    }
    if (t == 1023) rp[N] = run;  // = E when every index is a valid target
  }
  __syncthreads();
  int* __restrict__ row_ptr = row_ptr_all + (int64_t)b * (N + 1);
  for (int i = t; i <= N; i += 1024) row_ptr[i] = rp[i];
  // ---- 3. fill through the (chunk, target) cursors
#pragma unroll 4
  for (int p = p0 + lane; p < p1; p += 32) {
    const int tg = __ldg(idx + p);
    if ((unsigned)tg < (unsigned)N) {
      const int sh = (tg & 1) << 4;
      const unsigned old = atomicAdd(&my[tg >> 1], 1u << sh);
      perm16[rp[tg] + (int)((old >> sh) & 0xffffu)] = (unsigned short)p;
    }
  }
  __syncthreads();
  // ---- 4. every (chunk, target) piece in ascending order: the cursor of chunk w now stands at the start of chunk w+1.
  //         Pieces of up to kOrdThreadSort entries by the target's thread (insertion sort); longer ones (many sources of
  //         one chunk on one target: not a three_nn list, but it must not take seconds) go to a list and are sorted by
  //         a warp each through a bitmap over the chunk's positions: cost ~ piece length + chunk length / 32.
  if (t == 0) s_heavy = 0;
  __syncthreads();
  for (int tt = t; tt < N; tt += 1024) {
    unsigned short* seg = perm16 + rp[tt];
    int s0 = 0;
    for (int ww = 0; ww < 32; ww++) {
      const int e0 = cnt[(size_t)ww * Np + tt];
      if (e0 - s0 > kOrdThreadSort) {
        const int slot = atomicAdd(&s_heavy, 1);
        if (slot < heavy_cap) { heavy[slot] = ((unsigned)ww << 16) | (unsigned)tt; s0 = e0; continue; }
      }
      for (int i = s0 + 1; i < e0; i++) {
        const unsigned short v = seg[i];
        int j = i;
        while (j > s0 && seg[j - 1] > v) { seg[j] = seg[j - 1]; j--; }
        seg[j] = v;
      }
      s0 = e0;
    }
  }
  __syncthreads();
  {
    const int nh = min(s_heavy, heavy_cap);
    unsigned* bm = bitmap + w * kOrdBitmapWords;  // this warp's bitmap over one chunk (gpw * 32 <= 2080 positions)
    for (int h = w; h < nh; h += 32) {
      const int ww = (int)(heavy[h] >> 16), tt = (int)(heavy[h] & 0xffffu);
      const int s0 = ww ? cnt[(size_t)(ww - 1) * Np + tt] : 0, e0 = cnt[(size_t)ww * Np + tt];
      unsigned short* seg = perm16 + rp[tt];
      const int cbase = ww * gpw * 32;
      for (int i = lane; i < kOrdBitmapWords; i += 32) bm[i] = 0u;
      __syncwarp();
      for (int i = s0 + lane; i < e0; i += 32) {
        const int rel = (int)seg[i] - cbase;
        atomicOr(&bm[rel >> 5], 1u << (rel & 31));
      }
      __syncwarp();
      // lane l owns words [3l, 3l+3): exclusive prefix of their population counts, then the set bits in order
      unsigned wd[3];
      int pc = 0;
#pragma unroll
      for (int k = 0; k < 3; k++) { wd[k] = 3 * lane + k < kOrdBitmapWords ? bm[3 * lane + k] : 0u; pc += __popc(wd[k]); }
      int inc = pc;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += y;
      }
      int out = s0 + inc - pc;
      __syncwarp();  // every lane has read the piece before it is rewritten
#pragma unroll
      for (int k = 0; k < 3; k++) {
        unsigned m = wd[k];
        while (m) {
          const int bit = __ffs(m) - 1;
          m &= m - 1;
          seg[out++] = (unsigned short)(cbase + (3 * lane + k) * 32 + bit);
        }
      }
      __syncwarp();
    }
  }
  __syncthreads();
  // ---- 5. write-out
  int* __restrict__ perm = perm_all + (int64_t)b * E;
  for (int i = t; i < E; i += 1024) perm[i] = perm16[i];
}

// T = unsigned short: the cloud's whole source list lives in shared memory while it is filled and sorted (E <= 65535)
// and leaves by one coalesced pass; T = int: it is filled and sorted in place in global memory (longer lists).
template <typename T>
__global__ void __launch_bounds__(1024, 1)
plan_build_fused_kernel(int Nall, int E, int list_cap, int nmax, const int* __restrict__ idx_all, int* __restrict__ row_ptr_all, int* __restrict__ perm_all) {
  constexpr bool kSmemList = sizeof(T) == 2;
  // grid (B, S): with few clouds the targets of a cloud are split into S ranges, one CTA each.  Every CTA scans ALL
  // sources of the cloud (L2-resident) but keeps only those whose target falls into [t_lo, t_hi); the number of sources
  // with a smaller target - its offset into the cloud's list - is counted in the same scan, so the CTAs of a cloud
  // never talk to each other.  N below is the LOCAL target count, targets are re-based by t_lo.
  const int t_lo = (int)((int64_t)blockIdx.y * Nall / gridDim.y), t_hi = (int)((int64_t)(blockIdx.y + 1) * Nall / gridDim.y);
  const int N = t_hi - t_lo;
  extern __shared__ int plan_sm[];
  int* cnt = plan_sm;                                   // N+1 counters, later cursors
  unsigned* bm = reinterpret_cast<unsigned*>(plan_sm + nmax + 1);   // E bits (very long segments only)
  const int bm_words = ((((E + 31) >> 5) + 1023) >> 10) << 10;
  __shared__ int s_warp[32];
  __shared__ int s_carry, s_heavy_n;
  __shared__ int s_heavy[256];
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5, b = blockIdx.x;
  const int* __restrict__ idx = idx_all + (int64_t)b * E;
  int* __restrict__ row_ptr = row_ptr_all + (int64_t)b * (Nall + 1) + t_lo;
  int* __restrict__ perm_out = perm_all + (int64_t)b * E;   // + the range's offset, known after the count pass
  __shared__ int s_below[32];
  // work lists of the sort phase (target ids): segments of 3..16 entries / of 17..kFusedWarp entries
  int* list16 = reinterpret_cast<int*>(bm + bm_words) + (kSmemList ? ((E + 1) >> 1) : 0);
  int* listw = list16 + list_cap;
  __shared__ int s_n16, s_nw;
  PLAN_STAMP(0);
  for (int i = t; i <= N; i += 1024) cnt[i] = 0;
  if (t == 0) { s_carry = 0; s_heavy_n = 0; s_n16 = 0; s_nw = 0; }
  __syncthreads();
  PLAN_STAMP(1);
  constexpr int U = 8;  // loads in flight per thread: the edge loops are L2-latency bound otherwise
  int below = 0;        // sources of this cloud whose (valid) target lies before this CTA's range
  for (int e0 = t; e0 < E; e0 += 1024 * U) {
    int k[U];
#pragma unroll
    for (int u = 0; u < U; u++) k[u] = (e0 + u * 1024 < E) ? __ldg(idx + e0 + u * 1024) : -1;
#pragma unroll
    for (int u = 0; u < U; u++) {
      if ((unsigned)(k[u] - t_lo) < (unsigned)N) atomicAdd(&cnt[k[u] - t_lo + 1], 1);   // integer: order-free
      else if ((unsigned)k[u] < (unsigned)t_lo) below++;
    }
  }
  below = __reduce_add_sync(0xffffffffu, below);
  if (lane == 0) s_below[warp] = below;
  __syncthreads();
  const int list_base = __reduce_add_sync(0xffffffffu, s_below[lane]);   // identical in every thread
  perm_out += list_base;
  T* __restrict__ perm = kSmemList ? reinterpret_cast<T*>(bm + bm_words) : reinterpret_cast<T*>(perm_out);
  PLAN_STAMP(2);
  // inclusive scan of cnt[0..N] (counts stored shifted by one => exclusive start offsets): every thread owns P
  // consecutive counters (P odd: conflict-free strides), one block-wide scan of the thread totals
  {
    const int P = (((N + 1) + 1023) >> 10) | 1;
    const int lo = min(t * P, N + 1), hi = min(lo + P, N + 1);
    int tot = 0;
    for (int i = lo; i < hi; i++) tot += cnt[i];
    int inc = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += u;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    int w = s_warp[lane], winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o) winc += u;
    }
    int run = __shfl_sync(0xffffffffu, winc - w, warp) + inc - tot;
    for (int i = lo; i < hi; i++) { run += cnt[i]; cnt[i] = run; }
    __syncthreads();
    // row_ptr[t_lo + i]; the entry at t_hi belongs to the next range (its own offset + 0), the last range also writes row_ptr[Nall]
    for (int i = t; i < N + (t_hi == Nall ? 1 : 0); i += 1024) row_ptr[i] = list_base + cnt[i];
    __syncthreads();  // the fill below turns the counters into cursors: nobody may still be copying them out
  }
  PLAN_STAMP(3);
  // fill: cnt[k] is the cursor of target k; afterwards cnt[k] = start of segment k+1
  for (int e0 = t; e0 < E; e0 += 1024 * U) {
    int k[U];
#pragma unroll
    for (int u = 0; u < U; u++) k[u] = (e0 + u * 1024 < E) ? __ldg(idx + e0 + u * 1024) : -1;
#pragma unroll
    for (int u = 0; u < U; u++)
      if ((unsigned)(k[u] - t_lo) < (unsigned)N) perm[atomicAdd(&cnt[k[u] - t_lo], 1)] = (T)(e0 + u * 1024);
  }
  __syncthreads();  // the CTA's own global writes are visible to it after the barrier
  PLAN_STAMP(4);
  // every segment into ascending source order.  Only ~20 % of the targets of a ball query are non-empty, so the
  // segments are first classified into dense work lists: 3..16 entries -> one THREAD each (sorting network in
  // registers), 17..kFusedWarp -> one WARP each (rank sort), longer -> the whole CTA through a bitmap.
  for (int base = warp * 32; base < N; base += 1024) {
    const int k = base + lane;
    int s = 0, L = 0;
    if (k < N) { s = k ? cnt[k - 1] : 0; L = cnt[k] - s; }
    if (L == 2) {
      T* __restrict__ seg = perm + s;
      const T x0 = seg[0], x1 = seg[1];
      if (x0 > x1) { seg[0] = x1; seg[1] = x0; }
    }
    if (L > kFusedWarpCut) {
      const int slot = atomicAdd(&s_heavy_n, 1);
      if (slot < 256) s_heavy[slot] = k;
    }
    const bool is16 = L >= 3 && L <= 16, isw = L > 16 && L <= kFusedWarpCut;
    const unsigned m16 = __ballot_sync(0xffffffffu, is16), mw = __ballot_sync(0xffffffffu, isw);
    int p16 = 0, pw = 0;
    if (lane == 0) {
      if (m16) p16 = atomicAdd(&s_n16, __popc(m16));
      if (mw) pw = atomicAdd(&s_nw, __popc(mw));
    }
    p16 = __shfl_sync(0xffffffffu, p16, 0) + __popc(m16 & lanemask_lt());
    pw = __shfl_sync(0xffffffffu, pw, 0) + __popc(mw & lanemask_lt());
    bool overflow = false;
    if (is16) { if (p16 < list_cap) list16[p16] = k; else overflow = true; }
    if (isw) { if (pw < list_cap) listw[pw] = k; else overflow = true; }
    unsigned now = __ballot_sync(0xffffffffu, overflow);  // lists full (tiny shared-memory budget): sort these right here
    while (now) {
      const int src = __ffs(now) - 1;
      now &= now - 1;
      plan_warp_sort(perm + __shfl_sync(0xffffffffu, s, src), __shfl_sync(0xffffffffu, L, src), lane);
    }
  }
  __syncthreads();
  {
    const int n16 = min(s_n16, list_cap), nw = min(s_nw, list_cap);
    for (int i = t; i < n16; i += 1024) {
      const int k = list16[i], s = k ? cnt[k - 1] : 0;
      plan_thread_sort16(perm + s, cnt[k] - s);
    }
    for (int i = warp; i < nw; i += 32) {
      const int k = listw[i], s = k ? cnt[k - 1] : 0;
      plan_warp_sort(perm + s, cnt[k] - s, lane);
    }
  }
  __syncthreads();
  PLAN_STAMP(5);
  const int heavy_n = s_heavy_n;
#ifdef GEOT_FPS_PROF
  if (blockIdx.x == 0 && threadIdx.x == 0) g_plan_prof[7] = heavy_n;
#endif
  const int n_local = N > 0 ? cnt[N - 1] : 0;  // sources whose target lies in this range
  if (heavy_n == 0) {
    if (kSmemList) for (int i = t; i < n_local; i += 1024) perm_out[i] = (int)perm[i];
    return;
  }
  const int nwords = (E + 31) >> 5;
  const int wpt = (nwords + 1023) >> 10;  // words per thread, consecutive: thread order = source order
  // more than 256 very long segments cannot be listed: rescan the targets instead (same work, no list)
  const int items = heavy_n <= 256 ? heavy_n : N;
  for (int item = 0; item < items; item++) {
    const int k = heavy_n <= 256 ? s_heavy[item] : item;
    const int s = k ? cnt[k - 1] : 0, L = cnt[k] - s;
    if (L <= kFusedWarpCut) continue;  // (only reachable in rescan mode; uniform across the CTA)
    T* __restrict__ seg = perm + s;
    for (int i = t; i < wpt * 1024; i += 1024) bm[i] = 0u;
    __syncthreads();
    for (int i = t; i < L; i += 1024) { const int v = (int)seg[i]; atomicOr(&bm[v >> 5], 1u << (v & 31)); }
    __syncthreads();
    int c = 0;
    for (int i = 0; i < wpt; i++) c += __popc(bm[t * wpt + i]);
    int inc = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += y;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    int w = s_warp[lane], winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o) winc += y;
    }
    int o = __shfl_sync(0xffffffffu, winc - w, warp) + inc - c;
    for (int i = 0; i < wpt; i++) {
      unsigned word = bm[t * wpt + i];
      while (word) {
        seg[o++] = (T)((((t * wpt + i)) << 5) + __ffs(word) - 1);
        word &= word - 1;
      }
    }
    __syncthreads();
  }
  if (kSmemList) for (int i = t; i < n_local; i += 1024) perm_out[i] = (int)perm[i];
  PLAN_STAMP(6);
}

#ifdef GEOT_FPS_PROF
}  // namespace geot
GEOT_API int geot_debug_plan_prof(long long* host16) { return (int)cudaMemcpyFromSymbol(host16, geot::g_plan_prof, 16 * sizeof(long long)); }
namespace geot {
#endif

static inline size_t plan_align(size_t b) { return (b + 255) & ~(size_t)255; }
struct PlanView {
  const int* row_ptr;
  const int* perm;
};
static inline size_t plan_rowptr_bytes(int B, int N) {
  size_t b = (size_t)B * (N + 1) * sizeof(int);
  return (b + 255) & ~(size_t)255;
}
static inline PlanView plan_view(const void* plan, int B, int N) {
  PlanView v;
  v.row_ptr = reinterpret_cast<const int*>(plan);
  v.perm = reinterpret_cast<const int*>(reinterpret_cast<const char*>(plan) + plan_rowptr_bytes(B, N));
  return v;
}

// ---------------------------------------------------------------------------------------------
// Deterministic segmented reduction (channel-major):  dst[b,c,t] = sum_{p in seg(b,t)} g[b,c,perm[p]/DIV] (* w[b,perm[p]])
// DIV = 1: group/gather grads.  DIV = 3, weighted: three_interpolate grad.
// One thread per target, kSegCCH channels at a time, sources in plan (ascending) order.
// ---------------------------------------------------------------------------------------------
constexpr int kSegCCH = 8;
template <int DIV, bool WEIGHTED>
__global__ void __launch_bounds__(256)
seg_reduce_cmajor_kernel(int C, int N, int64_t E, const float* __restrict__ g,
                         const float* __restrict__ w, const int* __restrict__ row_ptr,
                         const int* __restrict__ perm, float* __restrict__ dst) {
  const int b = blockIdx.z;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= N) return;
  const int c0 = blockIdx.y * kSegCCH;
  const int nc = min(kSegCCH, C - c0);
  const int64_t Eg = E / DIV;
  const int* __restrict__ rp = row_ptr + (int64_t)b * (N + 1);
  const int* __restrict__ pm = perm + (int64_t)b * E;
  const float* __restrict__ gb = g + ((int64_t)b * C + c0) * Eg;
  const int s = rp[t], e = rp[t + 1];
  float acc[kSegCCH];
#pragma unroll
  for (int i = 0; i < kSegCCH; i++) acc[i] = 0.f;
  for (int p = s; p < e; p++) {
    const int src = __ldg(pm + p);
    const int64_t pos = DIV == 1 ? src : src / DIV;
    const float wt = WEIGHTED ? __ldg(w + (int64_t)b * E + src) : 1.f;
#pragma unroll
    for (int i = 0; i < kSegCCH; i++) {
      if (i < nc) {
        const float v = __ldg(gb + (int64_t)i * Eg + pos);
        acc[i] = __fadd_rn(acc[i], WEIGHTED ? __fmul_rn(v, wt) : v);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < kSegCCH; i++)
    if (i < nc) dst[((int64_t)b * C + c0 + i) * N + t] = acc[i];
}

// Plan-resident variant for the unweighted channel-major backward (grouping_operation grad) when the whole
// cloud's plan and one gradient row fit shared memory together (E <= 65535):  a CTA keeps the plan of its cloud
// in shared memory as 16-bit entries (perm: 2 B per source, row_ptr: 2 B per target - half the bytes of the
// global plan and read from L2 once per 4-8 channels instead of once per channel) and pulls the gradient rows
// g[b,c,:] of its channels through one shared buffer by bulk async copy.  Every gather of the reduction then
// hits shared memory: no 32-byte L2 sector per 4-byte element, no global latency inside the segment loops.
// Summation order = plan order (ascending source), exactly as in the direct kernel: same bits.
constexpr int kPrTaskCap = 5456;  // non-empty targets the compact task list can hold (6 bytes each in 32 KB)
// one segment of the 16-bit plan, summed in ascending plan order (the order of the direct kernel: same bits).  The plan
// entries are read two per 32-bit word: the gather loop sits on the shared-memory instruction queue (MIO throttle,
// two LDS per source), this makes it three per two sources.  GEOT_SEG_SUM16=0 keeps the one-entry-per-load loop (A/B).
#ifndef GEOT_SEG_SUM16
#define GEOT_SEG_SUM16 1
#endif
__device__ __forceinline__ float seg_sum16(const float* __restrict__ row, const unsigned short* __restrict__ perm16, int s0, int e0,
                                           float acc = 0.f) {
#if GEOT_SEG_SUM16
  int p = s0;
  if ((p & 1) && p < e0) { acc = __fadd_rn(acc, row[perm16[p]]); p++; }
  const unsigned* pp = reinterpret_cast<const unsigned*>(perm16 + p);  // perm16 is 16-byte aligned, p even
  for (; p + 1 < e0; p += 2) {
    const unsigned u = *pp++;
    acc = __fadd_rn(acc, row[u & 0xffffu]);
    acc = __fadd_rn(acc, row[u >> 16]);
  }
  if (p < e0) acc = __fadd_rn(acc, row[perm16[p]]);
#else
  for (int p = s0; p < e0; p++) acc = __fadd_rn(acc, row[perm16[p]]);
#endif
  return acc;
}
constexpr int kPrSplitCap = 4096;  // tasks the list holds when it also keeps every segment's cut at the half row (8 bytes each)
constexpr int kPrPerThread = kPrSplitCap / 1024;  // tasks of the split list one of the 1024 threads owns
__global__ void __launch_bounds__(1024, 1)
seg_reduce_planrow_kernel(int B, int C, int N, int E, int Q, int cpc, int variant, const float* __restrict__ g, const int* __restrict__ row_ptr,
                          const int* __restrict__ perm, float* __restrict__ dst) {
  extern __shared__ __align__(16) unsigned char pr_smem[];
  __shared__ uint64_t bar[2];
  __shared__ int s_bin[64], s_ne;
  float* row = reinterpret_cast<float*>(pr_smem);                                        // E floats
  unsigned short* perm16 = reinterpret_cast<unsigned short*>(pr_smem + (size_t)E * 4);  // E entries
  // 32 KB that hold EITHER row_ptr as 16-bit (general case) OR the compact task list {target, start, length} of
  // the non-empty targets ordered by descending length (sparse case, see below)
  unsigned short* rp16 = perm16 + ((E + 7) & ~7);
  const int t = threadIdx.x, lane = t & 31;
  // rows (cloud, channel) r0 .. r1 of this CTA, cloud-major.  cpc == 0: one CTA per SM takes an equal share of all B*C
  // rows, so the plan setup below runs once or twice per CTA (its rows span at most two clouds when the share is below
  // C) and no wave is left partly filled; cpc > 0: the CTA (x, y) takes cpc channels of cloud y (A/B, mode 12).
  int r0, r1;
  if (cpc > 0) {
    r0 = blockIdx.y * C + blockIdx.x * cpc;
    r1 = blockIdx.y * C + min(C, (int)(blockIdx.x + 1) * cpc);
  } else {
    const int64_t items = (int64_t)B * C;
    r0 = (int)(items * blockIdx.x / gridDim.x);
    r1 = (int)(items * (blockIdx.x + 1) / gridDim.x);
  }
  const int64_t Etot = (int64_t)E * Q;  // a gradient row has Q pieces of E values; the plan has one CSR per (cloud, piece)
  // A gradient row arrives as two halves on two mbarriers.  The compact path sums a segment's entries below `half` as
  // soon as the first half is there and lets the next row's first half load while it sums the rest out of the second
  // half (and the other way round): the copy of a row hides behind the sums of the row before it.  A segment's entries
  // ascend, so cutting the sum at `half` keeps the order of the additions.
  const int half = (E / 2) & ~3;
  const bool halves = variant & 1, prefetch = variant & 2;
  if (t == 0) {
    mbar_init(&bar[0], 1);
    mbar_init(&bar[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  unsigned loads = 0;  // rows waited for so far (mbarrier phase parity, the same for both halves)
#pragma unroll 1
  for (int r = r0, nc = 0; r < r1; r += nc) {
  const int b = r / C, c_first = r - b * C;
  nc = min(C - c_first, r1 - r);
#pragma unroll 1
  for (int q = 0; q < Q; q++) {
    __syncthreads();  // previous piece: every thread is done with perm16 / the task list / s_bin
    if (t == 0) s_ne = 0;
    if (t < 64) s_bin[t] = 0;
    __syncthreads();
    if (t == 0) {
      const float* src = g + ((int64_t)b * C + c_first) * Etot + (int64_t)q * E;
      bulk_g2s_chunked(row, src, (unsigned)half * 4u, &bar[0]);
      bulk_g2s_chunked(row + half, src + half, (unsigned)(E - half) * 4u, &bar[1]);
      if (prefetch && nc > 1) bulk_prefetch_l2(src + Etot, (unsigned)E * 4u);  // the second channel's row, during the setup
    }
    const int* __restrict__ pm = perm + ((int64_t)b * Q + q) * E;
    const int* __restrict__ rp = row_ptr + ((int64_t)b * Q + q) * (N + 1);
    {  // the cloud's plan entries as 16 bits, four per 16-byte load (perm is 256-byte aligned, E % 4 == 0)
      const int4* __restrict__ pm4 = reinterpret_cast<const int4*>(pm);
      uint2* p16 = reinterpret_cast<uint2*>(perm16);
      for (int i = t; i < E / 4; i += 1024) {
        const int4 v = __ldg(pm4 + i);
        p16[i] = make_uint2((unsigned)v.x | ((unsigned)v.y << 16), (unsigned)v.z | ((unsigned)v.w << 16));
      }
    }
    // ball_query keeps the LOWEST indices of every ball, so most points are never a neighbour (80 % of the targets
    // are empty at the bench shape) and the rest have ~10 sources.  When the non-empty targets fit the task list
    // they are bucketed by segment length (64 bins, longest first): a warp then works on 32 segments of (almost)
    // equal length - no empty targets, no divergence on the segment loop.
    // A thread reads the row_ptr pairs of its targets once ({start, length} packed in a register each, all loads in
    // flight together) when the cloud has at most 16 targets per thread; larger clouds re-read them in the second pass.
    constexpr int kOwn = 16;
    const bool owned = N <= kOwn * 1024;
    unsigned sl[kOwn];
    if (owned) {
#pragma unroll
      for (int j = 0; j < kOwn; j++) {
        const int k = t + j * 1024;
        const int s0 = k < N ? __ldg(rp + k) : 0, e0 = k < N ? __ldg(rp + k + 1) : 0;
        sl[j] = (unsigned)s0 | ((unsigned)(e0 - s0) << 16);
      }
#pragma unroll
      for (int j = 0; j < kOwn; j++) {
        const int len = (int)(sl[j] >> 16);
        if (len > 0) atomicAdd(&s_bin[63 - min(len, 63)], 1);
      }
    } else {
      for (int k = t; k < N; k += 1024) {
        const int len = __ldg(rp + k + 1) - __ldg(rp + k);
        if (len > 0) atomicAdd(&s_bin[63 - min(len, 63)], 1);
      }
    }
    __syncthreads();
    if (t < 32) {  // exclusive scan of the 64 bins (bin 0 = longest)
      const int a0 = s_bin[2 * lane], a1 = s_bin[2 * lane + 1];
      int inc = a0 + a1;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += y;
      }
      s_bin[2 * lane] = inc - a0 - a1;
      s_bin[2 * lane + 1] = inc - a1;
      if (lane == 31) s_ne = inc;
    }
    __syncthreads();
    const int ne = s_ne;
    const bool compact = ne <= kPrTaskCap;
    const bool split = halves && ne <= kPrSplitCap;  // room for a fourth 16-bit field: where the segment crosses `half`
    const int cap = split ? kPrSplitCap : kPrTaskCap;
    unsigned short* t_tgt = rp16;
    unsigned short* t_start = rp16 + cap;
    unsigned short* t_len = rp16 + 2 * cap;
    unsigned short* t_mid = rp16 + 3 * cap;
    auto add_task = [&](int k, int s0, int len) {
      const int pos = atomicAdd(&s_bin[63 - min(len, 63)], 1);  // order inside a bin is free: sums do not depend on it
      t_tgt[pos] = (unsigned short)k; t_start[pos] = (unsigned short)s0; t_len[pos] = (unsigned short)len;
      if (split) {  // first entry >= half (entries ascend; perm16 is complete since the barrier after the bins)
        int lo = s0, hi = s0 + len;
        while (lo < hi) {
          const int mid = (lo + hi) >> 1;
          if (perm16[mid] < half) lo = mid + 1; else hi = mid;
        }
        t_mid[pos] = (unsigned short)lo;
      }
    };
    if (compact && owned) {
#pragma unroll
      for (int j = 0; j < kOwn; j++)
        if (sl[j] >> 16) add_task(t + j * 1024, (int)(sl[j] & 0xffffu), (int)(sl[j] >> 16));
    } else if (compact) {
      for (int k = t; k < N; k += 1024) {
        const int s0 = __ldg(rp + k), len = __ldg(rp + k + 1) - s0;
        if (len > 0) add_task(k, s0, len);
      }
    } else if (owned) {
#pragma unroll
      for (int j = 0; j < kOwn; j++)
        if (t + j * 1024 < N) rp16[t + j * 1024] = (unsigned short)(sl[j] & 0xffffu);
      if (t == 0) rp16[N] = (unsigned short)__ldg(rp + N);
    } else {
      for (int i = t; i <= N; i += 1024) rp16[i] = (unsigned short)__ldg(rp + i);
    }
    __syncthreads();
    for (int ci = 0; ci < nc; ci++) {
      float* __restrict__ out = dst + ((int64_t)b * C + c_first + ci) * N;
      const float* next = g + ((int64_t)b * C + c_first + ci + 1) * Etot + (int64_t)q * E;
      // piece 0 writes every target, later pieces add to it: the same thread order every run, so the sum
      // ((piece 0 + piece 1) + ...) is fixed
      if (split) {
        if (q == 0)
          for (int k = t; k < N; k += 1024) out[k] = 0.f;  // empty targets; the others are overwritten below
        float acc[kPrPerThread];
        mbar_wait(&bar[0], loads & 1u);
#pragma unroll
        for (int j = 0; j < kPrPerThread; j++) {
          const int i = t + j * 1024;
          acc[j] = i < ne ? seg_sum16(row, perm16, t_start[i], t_mid[i]) : 0.f;
        }
        __syncthreads();  // every thread is done with the first half (and the zero fill is ordered before the sums)
        if (t == 0 && ci + 1 < nc) bulk_g2s_chunked(row, next, (unsigned)half * 4u, &bar[0]);
        mbar_wait(&bar[1], loads & 1u);
#pragma unroll
        for (int j = 0; j < kPrPerThread; j++) {
          const int i = t + j * 1024;
          if (i < ne) {
            const float a = seg_sum16(row, perm16, t_mid[i], t_start[i] + t_len[i], acc[j]);
            const int k = t_tgt[i];
            out[k] = q == 0 ? a : __fadd_rn(out[k], a);
          }
        }
        loads++;
        __syncthreads();  // every thread is done with the second half (and, for later pieces, its writes to out are ordered)
        if (t == 0 && ci + 1 < nc) {
          bulk_g2s_chunked(row + half, next + half, (unsigned)(E - half) * 4u, &bar[1]);
          if (prefetch && ci + 2 < nc) bulk_prefetch_l2(next + Etot, (unsigned)E * 4u);
        }
        continue;
      }
      if (compact) {
        if (q == 0)
          for (int k = t; k < N; k += 1024) out[k] = 0.f;
        mbar_wait(&bar[0], loads & 1u);
        mbar_wait(&bar[1], loads & 1u);
        __syncthreads();  // zero fill ordered before the sums of this CTA
        for (int i = t; i < ne; i += 1024) {
          const int s0 = t_start[i], e0 = s0 + t_len[i];
          float acc = seg_sum16(row, perm16, s0, e0);
          const int k = t_tgt[i];
          out[k] = q == 0 ? acc : __fadd_rn(out[k], acc);
        }
      } else {
        mbar_wait(&bar[0], loads & 1u);
        mbar_wait(&bar[1], loads & 1u);
        for (int k = t; k < N; k += 1024) {
          const int s0 = rp16[k], e0 = rp16[k + 1];
          float acc = seg_sum16(row, perm16, s0, e0);
          out[k] = q == 0 ? acc : __fadd_rn(out[k], acc);
        }
      }
      loads++;
      __syncthreads();  // every thread is done with the row buffer (and, for later pieces, its writes to out are ordered)
      if (t == 0 && ci + 1 < nc) {
        bulk_g2s_chunked(row, next, (unsigned)half * 4u, &bar[0]);
        bulk_g2s_chunked(row + half, next + half, (unsigned)(E - half) * 4u, &bar[1]);
        if (prefetch && ci + 2 < nc) bulk_prefetch_l2(next + Etot, (unsigned)E * 4u);  // lands while this row is copied and summed
      }
    }
  }
  }
}

// Pieces plan without the shared-memory path (unaligned gradient): one thread per target, pieces in order.
__global__ void __launch_bounds__(256)
seg_reduce_pieces_direct_kernel(int C, int N, int E, int Q, const float* __restrict__ g, const int* __restrict__ row_ptr,
                                const int* __restrict__ perm, float* __restrict__ dst) {
  const int b = blockIdx.z, c = blockIdx.y;
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= N) return;
  const float* __restrict__ gr = g + ((int64_t)b * C + c) * E * Q;
  float total = 0.f;
  for (int q = 0; q < Q; q++) {
    const int* __restrict__ rp = row_ptr + ((int64_t)b * Q + q) * (N + 1);
    const int* __restrict__ pm = perm + ((int64_t)b * Q + q) * E;
    float acc = 0.f;
    for (int p = rp[k]; p < rp[k + 1]; p++) acc = __fadd_rn(acc, __ldg(gr + (int64_t)q * E + __ldg(pm + p)));
    total = q == 0 ? acc : __fadd_rn(total, acc);
  }
  dst[((int64_t)b * C + c) * N + k] = total;
}

static size_t planrow_smem(int N, int64_t E) {
  return (size_t)E * 4 + (size_t)((E + 7) & ~7) * 2 + std::max((size_t)(N + 1) * 2, std::max((size_t)kPrTaskCap * 6, (size_t)kPrSplitCap * 8)) + 16;
}
static int launch_planrow(int B, int C, int N, int E, int Q, const float* g, const int* row_ptr, const int* perm, float* dst,
                          cudaStream_t s, const char* what) {
  static DynSmem ds;
  if (int rc = ds.need(seg_reduce_planrow_kernel, 226 * 1024, what)) return rc;
  // default: the next row prefetched into L2 when every SM has rows to pull (it costs a few per cent on a part-filled
  // machine), the half-row pipeline only for plans in pieces (measured: +5 % there, -7 % on single-piece plans)
  const int64_t rows = (int64_t)B * C;
  const int variant = g_pr_variant >= 0 ? (g_pr_variant & 3) : ((rows >= 8 * (int64_t)num_sms() ? 2 : 0) | (Q > 1 ? 1 : 0));
  if (g_pr_variant < 0 || !(g_pr_variant & 4)) {
    // one CTA per SM (at least four rows each): the plan setup costs about two rows' worth of time
    const int64_t items = (int64_t)B * C;
    const int grid = (int)std::min<int64_t>(num_sms(), (items + 3) / 4);
    seg_reduce_planrow_kernel<<<grid, 1024, planrow_smem(N, E), s>>>(B, C, N, E, Q, 0, variant, g, row_ptr, perm, dst);
  } else {
    // A/B: a grid of (channel group, cloud) CTAs, channels per CTA chosen against the last wave's fill
    int cpc = 4;
    double best = 0.0;
    for (int k = 4; k <= 8; k++) {
      const double ctas = (double)B * ((C + k - 1) / k);
      const double waves = std::ceil(ctas / num_sms());
      const double score = ctas / (waves * num_sms()) * k / (k + 0.8);
      if (score > best) { best = score; cpc = k; }
    }
    dim3 grid((C + cpc - 1) / cpc, B);
    seg_reduce_planrow_kernel<<<grid, 1024, planrow_smem(N, E), s>>>(B, C, N, E, Q, cpc, variant, g, row_ptr, perm, dst);
  }
  count_launch(1);
  return check_launch(what);
}

template <int DIV, bool WEIGHTED>
static int launch_seg_reduce(int B, int C, int N, int64_t E, const float* g, const float* w,
                             const void* plan, float* dst, cudaStream_t s, const char* what) {
  if (B == 0 || C == 0 || N == 0) return GEOT_OK;
  GEOT_REQUIRE(g && plan && dst, "%s: NULL pointer", what);
  GEOT_REQUIRE(B <= 65535 && (C + kSegCCH - 1) / kSegCCH <= 65535, "%s: B or C exceeds grid limits", what);
  const PlanView pv = plan_view(plan, B, N);
  if (DIV == 1 && !WEIGHTED && !g_group_force_direct && E <= 65535 && E % 4 == 0 && E >= N && (((uintptr_t)g & 15) == 0) &&
      planrow_smem(N, E) <= (size_t)226 * 1024)
    return launch_planrow(B, C, N, (int)E, 1, g, pv.row_ptr, pv.perm, dst, s, what);
  dim3 grid((N + 255) / 256, (C + kSegCCH - 1) / kSegCCH, B);
  seg_reduce_cmajor_kernel<DIV, WEIGHTED><<<grid, 256, 0, s>>>(C, N, E, g, w, pv.row_ptr, pv.perm, dst);
  count_launch(1);
  return check_launch(what);
}

// ---------------------------------------------------------------------------------------------
// three_interpolate backward, row-resident:  dst[b,c,k] = sum_{p in seg(b,k)} w[b,perm[p]] * g[b,c,perm[p]/3].
// The direct kernel above gathers g and w per edge from L2 (one 32-byte sector per 4-byte value, ~5 % of HBM
// peak at the bench shape).  Here (1) the weights are put into plan order once per call (wsort, into the plan
// workspace's build scratch), so a segment's {perm, w} are two contiguous runs, and (2) a CTA keeps R gradient
// rows g[b,c..c+R,:] in shared memory (one bulk async copy), walks the m segments with 8 lanes per target
// - four targets in flight per warp, edges streamed coalesced from L2 once per R channels - and gathers from
// shared memory.  Per-target sums: lane-strided partial sums then a fixed xor tree, i.e. a fixed order
// (run-to-run identical); results leave through a shared staging row as coalesced stores.
// ---------------------------------------------------------------------------------------------
// One packed record per edge in plan order: {source point = perm / 3, weight bits}.  The reduce kernel then needs ONE
// 8-byte load per edge instead of two 4-byte loads and a division by 3 (it is instruction-bound: ~16 instructions per
// edge and R <= 3 channels before, round-1 ncu: 150 M warp instructions for 201 M edge-channels).
__global__ void __launch_bounds__(256)
plan_wsort_kernel(int64_t E, const float* __restrict__ w, const int* __restrict__ perm, int2* __restrict__ edges) {
  const int64_t e = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (e >= E) return;
  const int64_t base = (int64_t)blockIdx.y * E;
  const int p = __ldg(perm + base + e);
  edges[base + e] = make_int2(p / 3, __float_as_int(__ldg(w + base + p)));
}

// LPT = lanes per target: 8 for long segments (~48 edges when n/m = 16), 2 when a target has only a handful of
// edges (n/m = 2: six on average) - then 512 targets are in flight per CTA instead of 128 and the per-target
// round trip to L2, not the edge count, is what is being hidden.
template <int R, int LPT, int NB>
__global__ void __launch_bounds__(1024, 1)
seg_reduce_wrows_kernel(int C, int n, int m, const float* __restrict__ g, const int* __restrict__ row_ptr,
                        const int2* __restrict__ edges, float* __restrict__ dst) {
  extern __shared__ __align__(16) unsigned char wr_smem[];
  __shared__ uint64_t bar;
  float* rows = reinterpret_cast<float*>(wr_smem);          // R x n
  float* outs = rows + (size_t)R * n;                       // R x m
  int* rp_s = reinterpret_cast<int*>(outs + (size_t)R * m); // m + 1
  const int t = threadIdx.x;
  const int b = blockIdx.y, c0 = blockIdx.x * R;
  const int nc = min(R, C - c0);
  const int64_t E = (int64_t)3 * n;
  if (t == 0) {
    mbar_init(&bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (t == 0) bulk_g2s_chunked(rows, g + ((int64_t)b * C + c0) * n, (unsigned)(nc * n) * 4u, &bar);
  const int* __restrict__ rp = row_ptr + (int64_t)b * (m + 1);
  for (int i = t; i <= m; i += 1024) rp_s[i] = __ldg(rp + i);
  __syncthreads();
  mbar_wait(&bar, 0);
  const int2* __restrict__ ed = edges + (int64_t)b * E;
  const int sub = t % LPT;
  // the loop bound is WARP-uniform (kb = first target of the warp): the xor tree below is a full-mask shuffle, so a
  // warp whose last targets fall beyond m (m not a multiple of 32 / LPT) must still run it with all 32 lanes -
  // lanes that left the loop early would wait at the __syncthreads below and the shuffle would never complete
  for (int kb = (t >> 5) * (32 / LPT); kb < m; kb += 1024 / LPT) {
    const int k = kb + (t & 31) / LPT;
    const bool valid = k < m;
    const int s0 = valid ? rp_s[k] : 0, e0 = valid ? rp_s[k + 1] : 0;
    float acc[R];
#pragma unroll
    for (int r = 0; r < R; r++) acc[r] = 0.f;
    // up to 64 edges of the segment per round: all 8 {perm, w} pairs of a lane are requested before the first
    // is used, so a typical segment (~48 edges) costs one L2 round trip instead of one per pair
    for (int p0 = s0 + sub; p0 < e0; p0 += NB * LPT) {
      int2 e8[NB];
#pragma unroll
      for (int j = 0; j < NB; j++) {
        const int p = p0 + LPT * j;
        e8[j] = p < e0 ? __ldg(ed + p) : make_int2(0, 0);
      }
#pragma unroll
      for (int j = 0; j < NB; j++) {
        if (p0 + LPT * j < e0) {
          const float wt = __int_as_float(e8[j].y);
#pragma unroll
          for (int r = 0; r < R; r++) acc[r] = __fadd_rn(acc[r], __fmul_rn(rows[(size_t)r * n + e8[j].x], wt));
        }
      }
    }
#pragma unroll
    for (int r = 0; r < R; r++) {
#pragma unroll
      for (int o = LPT / 2; o > 0; o >>= 1) acc[r] = __fadd_rn(acc[r], __shfl_xor_sync(0xffffffffu, acc[r], o));
    }
    if (sub == 0 && valid) {
#pragma unroll
      for (int r = 0; r < R; r++) outs[(size_t)r * m + k] = acc[r];
    }
  }
  __syncthreads();
  float* __restrict__ out = dst + ((int64_t)b * C + c0) * m;  // nc rows of m floats are contiguous
  for (int i = t; i < nc * m; i += 1024) out[i] = outs[i];
}

template <int R, int LPT, int NB>
static int launch_wrows_t(int B, int C, int n, int m, size_t smem, const float* g, const PlanView& pv, const int2* wsorted,
                          float* dst, cudaStream_t s) {
  static DynSmem ds;  // per instantiation
  if (int rc = ds.need(seg_reduce_wrows_kernel<R, LPT, NB>, 226 * 1024, "three_interpolate_grad")) return rc;
  dim3 grid((C + R - 1) / R, B);
  seg_reduce_wrows_kernel<R, LPT, NB><<<grid, 1024, smem, s>>>(C, n, m, g, pv.row_ptr, wsorted, dst);
  count_launch(1);
  return check_launch("seg_reduce_wrows_kernel");
}
template <int R>
static int launch_wrows(int B, int C, int n, int m, size_t smem, const float* g, const PlanView& pv, const int2* wsorted,
                        float* dst, cudaStream_t s) {
  // average segment = 3n / m edges
  if ((int64_t)3 * n < (int64_t)24 * m) return launch_wrows_t<R, 2, 8>(B, C, n, m, smem, g, pv, wsorted, dst, s);
  return launch_wrows_t<R, 8, 8>(B, C, n, m, smem, g, pv, wsorted, dst, s);
}

// ---------------------------------------------------------------------------------------------
// three_interpolate backward, streamed (m <= 1024): a THREAD owns a target and keeps CG channel sums in registers; the
// CG gradient rows of its CTA pass through shared memory in tiles of TS consecutive source points (bulk async copies,
// double-buffered), and because the plan lists a target's sources in ascending order every thread simply walks its own
// segment of packed {source, weight} edges forward, consuming the edges whose source lies in the current tile.  No lane
// groups, no shuffles, no per-round slot bookkeeping: one edge load and CG x (LDS, FMUL, FADD) per edge.  The sum of a
// target runs strictly in plan (ascending source) order - the order of the direct kernel: same bits as mode 1.
// ---------------------------------------------------------------------------------------------
constexpr int kTsThreads = 1024;
// smem: [raw tile, 2 buffers of CG x TS floats (bulk copies) | interleaved tile: CG/4 planes of TS float4]
template <int CG, int TS>
__global__ void __launch_bounds__(kTsThreads, 1)
ti_bwd_stream_kernel(int C, int n, int m, const float* __restrict__ g, const int* __restrict__ row_ptr,
                     const int2* __restrict__ edges, float* __restrict__ dst) {
  static_assert(CG % 4 == 0 && TS == kTsThreads, "one source column per thread in the interleave pass");
  extern __shared__ __align__(16) float ts_smem[];
  __shared__ uint64_t bar[2];
  float4* inter = reinterpret_cast<float4*>(ts_smem + (size_t)2 * CG * TS);
  const int t = threadIdx.x;
  const int b = blockIdx.y, c0 = blockIdx.x * CG;
  const int nc = min(CG, C - c0);
  const int64_t E = (int64_t)3 * n;
  const int ntiles = (n + TS - 1) / TS;
  const float* __restrict__ g0 = g + ((int64_t)b * C + c0) * n;
  auto issue = [&](int tile) {  // thread 0: the nc row pieces of this tile
    const int buf = tile & 1, j0 = tile * TS, len = min(TS, n - j0);
    mbar_expect_tx(&bar[buf], (unsigned)(nc * len) * 4u);
    for (int c = 0; c < nc; c++) bulk_g2s(ts_smem + (size_t)buf * CG * TS + (size_t)c * TS, g0 + (int64_t)c * n + j0, (unsigned)len * 4u, &bar[buf]);
  };
  if (t == 0) {
    mbar_init(&bar[0], 1);
    mbar_init(&bar[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (t == 0) issue(0);
  const int2* __restrict__ ed = edges + (int64_t)b * E;
  int cur = 0, end = 0;
  if (t < m) { cur = __ldg(row_ptr + (int64_t)b * (m + 1) + t); end = __ldg(row_ptr + (int64_t)b * (m + 1) + t + 1); }
  const int2 kNone = make_int2(0x7fffffff, 0);
  int2 e = cur < end ? __ldg(ed + cur) : kNone;          // the edge being consumed and the two behind it: loads stay
  int2 e1 = cur + 1 < end ? __ldg(ed + cur + 1) : kNone;  // two edges ahead of the sums
  int2 e2 = cur + 2 < end ? __ldg(ed + cur + 2) : kNone;
  float acc[CG];
#pragma unroll
  for (int c = 0; c < CG; c++) acc[c] = 0.f;
  for (int tile = 0; tile < ntiles; tile++) {
    const int buf = tile & 1;
    if (t == 0 && tile + 1 < ntiles) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // the other raw buffer was read one tile ago (barriers below)
      issue(tile + 1);
    }
    mbar_wait(&bar[buf], (unsigned)((tile >> 1) & 1));
    {  // interleave: thread t takes source column t of the raw tile, four channels per float4 (rows beyond nc: zeros)
      const float* __restrict__ raw = ts_smem + (size_t)buf * CG * TS + t;
#pragma unroll
      for (int q = 0; q < CG / 4; q++) {
        float4 v;
        v.x = 4 * q + 0 < nc ? raw[(4 * q + 0) * TS] : 0.f;
        v.y = 4 * q + 1 < nc ? raw[(4 * q + 1) * TS] : 0.f;
        v.z = 4 * q + 2 < nc ? raw[(4 * q + 2) * TS] : 0.f;
        v.w = 4 * q + 3 < nc ? raw[(4 * q + 3) * TS] : 0.f;
        inter[q * TS + t] = v;
      }
    }
    __syncthreads();  // the interleaved tile is complete (and this raw buffer may be refilled from the next iteration on)
    const int j0 = tile * TS, j1 = j0 + TS;
    while (e.x < j1) {
      const float wt = __int_as_float(e.y);
      const float4* __restrict__ col = inter + (e.x - j0);
      cur++;
      const int2 nx = cur + 2 < end ? __ldg(ed + cur + 2) : kNone;
#pragma unroll
      for (int q = 0; q < CG / 4; q++) {
        const float4 v = col[q * TS];
        acc[4 * q + 0] = __fadd_rn(acc[4 * q + 0], __fmul_rn(v.x, wt));
        acc[4 * q + 1] = __fadd_rn(acc[4 * q + 1], __fmul_rn(v.y, wt));
        acc[4 * q + 2] = __fadd_rn(acc[4 * q + 2], __fmul_rn(v.z, wt));
        acc[4 * q + 3] = __fadd_rn(acc[4 * q + 3], __fmul_rn(v.w, wt));
      }
      e = e1; e1 = e2; e2 = nx;
    }
    __syncthreads();  // everyone is done with the interleaved tile before the next one overwrites it
  }
  if (t < m) {
#pragma unroll
    for (int c = 0; c < CG; c++)
      if (c < nc) dst[((int64_t)b * C + c0 + c) * m + t] = acc[c];
  }
}

template <int CG, int TS>
static int launch_ti_bwd_stream(int B, int C, int n, int m, const float* g, const PlanView& pv, const int2* wsorted, float* dst, cudaStream_t s) {
  static DynSmem ds;
  const size_t smem = (size_t)3 * CG * TS * 4;
  if (int rc = ds.need(ti_bwd_stream_kernel<CG, TS>, smem, "three_interpolate_grad")) return rc;  // no carve-out preference: L1 serves the edge walks
  dim3 grid((C + CG - 1) / CG, B);
  ti_bwd_stream_kernel<CG, TS><<<grid, kTsThreads, smem, s>>>(C, n, m, g, pv.row_ptr, wsorted, dst);
  count_launch(1);
  return check_launch("ti_bwd_stream_kernel");
}

static int three_interpolate_grad_rows(int B, int C, int n, int m, const float* g, const float* w, const void* plan,
                                       float* dst, cudaStream_t s, bool* taken) {
  *taken = false;
  const int64_t E = (int64_t)3 * n;
  if (g_group_force_direct || n % 4 != 0 || (((uintptr_t)g) & 15) != 0 || n < 1024 || E < m || B > 65535) return GEOT_OK;
  auto smem_of = [&](int R) { return (size_t)R * n * 4 + (size_t)R * m * 4 + (size_t)(m + 1) * 4 + 16; };
  // rows per CTA: the edge stream (8 B per edge = 24 B per row element) is re-read once per R channels
  static const int kR[6] = {1, 2, 3, 4, 6, 8};
  int R = 0;
  double best = 0.0;
  for (int i = 0; i < 6; i++) {
    const int r = kR[i];
    if (smem_of(r) > (size_t)226 * 1024 || (r > 1 && r > C)) continue;
    const double ctas = (double)B * ((C + r - 1) / r);
    const double eff = ctas / (std::ceil(ctas / num_sms()) * num_sms());
    const double score = eff / (24.0 / r + 4.0);
    if (score > best) { best = score; R = r; }
  }
  if (R == 0) return GEOT_OK;
  // plan workspace: [row_ptr | perm | cursor | tmp(B*E ints) | heavy list | spare(B*E ints)]; everything from tmp on is
  // scratch of the plan build (at least 2 x B*E ints, geot_scatter_plan_bytes) and holds the packed edges here
  const PlanView pv = plan_view(plan, B, m);
  const size_t perm_b = plan_align((size_t)B * (size_t)E * sizeof(int));
  int2* wsorted = reinterpret_cast<int2*>(const_cast<char*>(reinterpret_cast<const char*>(plan)) + plan_rowptr_bytes(B, m) + perm_b + plan_rowptr_bytes(B, m));
  dim3 gw((unsigned)((E + 255) / 256), B);
  plan_wsort_kernel<<<gw, 256, 0, s>>>(E, w, pv.perm, wsorted);
  count_launch(1);
  int rc = check_launch("plan_wsort_kernel");
  if (rc) return rc;
  *taken = true;
  // streamed kernel (thread per target, 16 channels per CTA): a CTA walks the ceil(n/1024) source tiles one after the
  // other, so it needs the machine filled by CTAs, not by tiles.  Measured (tools/tibwd_ab.py, row-resident kernel in
  // brackets): 64 x 64 x 1024 <- 16384 0.242 (0.264) ms, 16 x 128 x 1024 <- 4096 0.033 (0.038), 6 x 384 x 512 <- 2048
  // 0.021 (0.025); with 32 CTAs (8 x 64 x 1024 <- 16384) 0.111 against 0.045
  if (m <= kTsThreads && (C + 15) / 16 <= 65535 && g_tib_stream >= 0 &&
      ((int64_t)B * ((C + 15) / 16) >= 96 || g_tib_stream > 0))
    return launch_ti_bwd_stream<16, 1024>(B, C, n, m, g, pv, wsorted, dst, s);
  const size_t smem = smem_of(R);
  switch (R) {
    case 1: return launch_wrows<1>(B, C, n, m, smem, g, pv, wsorted, dst, s);
    case 2: return launch_wrows<2>(B, C, n, m, smem, g, pv, wsorted, dst, s);
    case 3: return launch_wrows<3>(B, C, n, m, smem, g, pv, wsorted, dst, s);
    case 4: return launch_wrows<4>(B, C, n, m, smem, g, pv, wsorted, dst, s);
    case 6: return launch_wrows<6>(B, C, n, m, smem, g, pv, wsorted, dst, s);
    default: return launch_wrows<8>(B, C, n, m, smem, g, pv, wsorted, dst, s);
  }
}

// used by interp.cu
bool direct_kernels_forced() { return g_group_force_direct; }
bool staged_kernels_forced() { return g_force_staged; }
bool rows4_disabled() { return g_rows4_off; }
bool three_nn_queue_off() { return g_nn_queue_off; }
bool persistent_interp_disabled() { return g_tip_off || g_rows4_off; }
bool manifold_tiles_disabled() { return g_mtile_off; }
bool manifold_tiles_forced() { return g_mtile_on; }
int three_interpolate_grad_impl(int B, int C, int n, int m, const float* grad_out, const float* weight,
                                const void* plan, float* grad_points, cudaStream_t s) {
  if (B > 0 && C > 0 && m > 0 && n > 0 && grad_out && plan && grad_points && weight) {
    bool taken = false;
    const int rc = three_interpolate_grad_rows(B, C, n, m, grad_out, weight, plan, grad_points, s, &taken);
    if (rc || taken) return rc;
  }
  return launch_seg_reduce<3, true>(B, C, m, (int64_t)3 * n, grad_out, weight, plan, grad_points, s,
                                    "three_interpolate_grad");
}

}  // namespace geot

using namespace geot;

GEOT_API int geot_group_points(int B, int C, int N, int M, int nsample, const float* points,
                               const int* idx, float* out, geot_stream_t stream) {
  GEOT_REQUIRE(B >= 0 && C >= 0 && N >= 0 && M >= 0 && nsample >= 0, "group_points: bad sizes");
  return launch_group(B, C, N, (int64_t)M * nsample, points, idx, out, as_stream(stream), "group_points");
}

namespace geot {
// ---------------------------------------------------------------------------------------------
// Fused xyz grouping of QueryAndGroup / KNNGroup (group.py:235-250, :297-312):
//     out[b,c,j,s] = (xyz[b, idx[b,j,s], c] - new_xyz[b,j,c]) (/ divisor)
// straight from the point-major (B,N,3) cloud: no (B,3,N) transpose copy, no separate subtract and divide passes
// over the (B,3,M,ns) tensor.  Same fp32 operations as the composite (one subtraction, one multiplication by the
// reciprocal of the divisor - what torch does for tensor / scalar), so the result is bit-identical to it.  One thread per 4 consecutive neighbours of a query (int4 idx, float4 stores).
// ---------------------------------------------------------------------------------------------
template <int VEC>
__global__ void __launch_bounds__(256)
group_xyz_rel_kernel(int N, int M, int ns, const float* __restrict__ xyz, const float* __restrict__ new_xyz,
                     const int* __restrict__ idx, int relative, float divisor, float* __restrict__ out) {
  const int b = blockIdx.y;
  const int64_t E = (int64_t)M * ns;
  const int64_t e0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * VEC;
  if (e0 >= E) return;
  const int j = (int)(e0 / ns);  // VEC divides ns: the VEC neighbours belong to one query
  const float* __restrict__ cloud = xyz + (int64_t)b * N * 3;
  float cx = 0.f, cy = 0.f, cz = 0.f;
  if (relative) {
    const float* __restrict__ q = new_xyz + ((int64_t)b * M + j) * 3;
    cx = __ldg(q); cy = __ldg(q + 1); cz = __ldg(q + 2);
  }
  // torch divides a tensor by a Python scalar as a multiplication by the fp32 reciprocal (BinaryDivTrueKernel.cu,
  // is_cpu_scalar branch); group.py:249 `grouped_xyz /= self.radius` therefore is x * (1.0f / radius): same here
  const float inv = divisor != 0.f ? __fdiv_rn(1.0f, divisor) : 1.0f;
  int id[VEC];
  if (VEC == 4) {
    const int4 v = __ldg(reinterpret_cast<const int4*>(idx + (int64_t)b * E + e0));
    id[0] = v.x; id[1] = v.y; id[2] = v.z; id[3] = v.w;
  } else {
    id[0] = __ldg(idx + (int64_t)b * E + e0);
  }
  float ox[VEC], oy[VEC], oz[VEC];
#pragma unroll
  for (int v = 0; v < VEC; v++) {
    const float* __restrict__ p = cloud + (int64_t)id[v] * 3;
    ox[v] = __fsub_rn(__ldg(p), cx); oy[v] = __fsub_rn(__ldg(p + 1), cy); oz[v] = __fsub_rn(__ldg(p + 2), cz);
    if (divisor != 0.f) { ox[v] = __fmul_rn(ox[v], inv); oy[v] = __fmul_rn(oy[v], inv); oz[v] = __fmul_rn(oz[v], inv); }
  }
  float* __restrict__ o = out + (int64_t)b * 3 * E + e0;
  if (VEC == 4) {
    __stcs(reinterpret_cast<float4*>(o), make_float4(ox[0], ox[1], ox[2], ox[3]));
    __stcs(reinterpret_cast<float4*>(o + E), make_float4(oy[0], oy[1], oy[2], oy[3]));
    __stcs(reinterpret_cast<float4*>(o + 2 * E), make_float4(oz[0], oz[1], oz[2], oz[3]));
  } else {
    o[0] = ox[0]; o[E] = oy[0]; o[2 * E] = oz[0];
  }
}

}  // namespace geot

GEOT_API int geot_group_xyz_relative(int B, int N, int M, int nsample, const float* xyz, const float* new_xyz,
                                     const int* idx, int relative, float divisor, float* out, geot_stream_t stream) {
  GEOT_REQUIRE(B >= 0 && N >= 0 && M >= 0 && nsample >= 0, "group_xyz_relative: bad sizes");
  const int64_t E = (int64_t)M * nsample;
  if (B == 0 || E == 0) return GEOT_OK;
  GEOT_REQUIRE(xyz && idx && out && (!relative || new_xyz), "group_xyz_relative: NULL pointer");
  GEOT_REQUIRE(B <= 65535, "group_xyz_relative: B exceeds grid limits");
  cudaStream_t s = as_stream(stream);
  const bool vec = nsample % 4 == 0 && (((uintptr_t)idx | (uintptr_t)out) & 15) == 0;
  if (vec) {
    dim3 grid((unsigned)((E / 4 + 255) / 256), B);
    group_xyz_rel_kernel<4><<<grid, 256, 0, s>>>(N, M, nsample, xyz, new_xyz, idx, relative, divisor, out);
  } else {
    dim3 grid((unsigned)((E + 255) / 256), B);
    group_xyz_rel_kernel<1><<<grid, 256, 0, s>>>(N, M, nsample, xyz, new_xyz, idx, relative, divisor, out);
  }
  count_launch(1);
  return check_launch("group_xyz_rel_kernel");
}

namespace geot {
// Point-major gather: out[b,j,:] = src[b, idx[b,j], :] for D floats per point (D = 3: the sampled centres
// new_xyz = xyz[fps_idx]).  The reference reaches this through transpose(1,2).contiguous() -> gather_operation ->
// transpose(1,2).contiguous() (e.g. pointnet2_modules / subsample.py:108-144 callers); same values, no copies.
__global__ void __launch_bounds__(256)
gather_rows_kernel(int N, int M, int D, const float* __restrict__ src, const int* __restrict__ idx, float* __restrict__ out) {
  const int b = blockIdx.y;
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // (j, d) flattened
  if (e >= (int64_t)M * D) return;
  const int j = (int)(e / D), d = (int)(e - (int64_t)j * D);
  out[(int64_t)b * M * D + e] = __ldg(src + ((int64_t)b * N + __ldg(idx + (int64_t)b * M + j)) * D + d);
}
}  // namespace geot

GEOT_API int geot_gather_rows(int B, int N, int M, int D, const float* src, const int* idx, float* out, geot_stream_t stream) {
  GEOT_REQUIRE(B >= 0 && N >= 0 && M >= 0 && D >= 1, "gather_rows: bad sizes");
  if (B == 0 || M == 0) return GEOT_OK;
  GEOT_REQUIRE(src && idx && out, "gather_rows: NULL pointer");
  GEOT_REQUIRE(B <= 65535, "gather_rows: B exceeds grid limits");
  dim3 grid((unsigned)(((int64_t)M * D + 255) / 256), B);
  gather_rows_kernel<<<grid, 256, 0, as_stream(stream)>>>(N, M, D, src, idx, out);
  count_launch(1);
  return check_launch("gather_rows_kernel");
}

/* testing / A-B timing: direct != 0 disables the row-staged forward kernel (process-wide). */
GEOT_API void geot_group_points_config(int mode) {
  g_group_force_direct = mode == 1; g_force_staged = mode == 2; g_rows4_off = mode == 3; g_tip_off = mode == 4; g_mtile_off = mode == 6; g_mtile_on = mode == 7; g_pr_variant = mode == 9 ? 4 : (mode == 10 ? 3 : (mode == 11 ? 1 : (mode == 12 ? 6 : (mode == 13 ? 2 : -1)))); g_tib_stream = mode == 8 ? 1 : (mode == 5 ? -1 : 0); g_nn_queue_off = mode == 14; g_plan_ordered_off = mode == 15;
}

GEOT_API int geot_gather_points(int B, int C, int N, int M, const float* points, const int* idx,
                                float* out, geot_stream_t stream) {
  GEOT_REQUIRE(B >= 0 && C >= 0 && N >= 0 && M >= 0, "gather_points: bad sizes");
  return launch_group(B, C, N, (int64_t)M, points, idx, out, as_stream(stream), "gather_points");
}

// Workspace: [row_ptr | perm | cursor | tmp | heavy_count | heavy_list]; consumers read the first two only.
static inline int plan_heavy_cap(int B, int64_t E) { return (int)std::min<int64_t>((int64_t)B * (E / kPlanSmall) + 1, 1 << 28); }
GEOT_API size_t geot_scatter_plan_bytes(int B, int N, int64_t E) {
  if (B < 0 || N < 0 || E < 0) return 0;
  const size_t perm_b = plan_align((size_t)B * (size_t)E * sizeof(int));
  // [row_ptr | perm] = the plan; [cursor | tmp | heavy count | heavy list | spare] = scratch of the build, reused by
  // three_interpolate_grad for its packed {source, weight} edges (8 bytes per edge: tmp + spare)
  return plan_rowptr_bytes(B, N) + perm_b + plan_rowptr_bytes(B, N) + perm_b + 256 +
         plan_align((size_t)plan_heavy_cap(B, E) * sizeof(int)) + perm_b;
}

GEOT_API int geot_scatter_plan_build(int B, int N, int64_t E, const int* idx, void* plan,
                                     size_t plan_bytes, geot_stream_t stream) {
  GEOT_REQUIRE(B >= 0 && N >= 0 && E >= 0, "scatter_plan_build: bad sizes");
  if (B == 0) return GEOT_OK;
  GEOT_REQUIRE(plan != nullptr && (E == 0 || idx != nullptr), "scatter_plan_build: NULL pointer");
  GEOT_REQUIRE(B <= 65535, "scatter_plan_build: B exceeds grid limits");
  if (plan_bytes < geot_scatter_plan_bytes(B, N, E)) {
    set_error("scatter_plan_build: workspace %zu < %zu bytes", plan_bytes, geot_scatter_plan_bytes(B, N, E));
    return GEOT_ERR_WORKSPACE;
  }
  cudaStream_t s = as_stream(stream);
  int* row_ptr = reinterpret_cast<int*>(plan);
  int* perm = reinterpret_cast<int*>(reinterpret_cast<char*>(plan) + plan_rowptr_bytes(B, N));
  // with few clouds, S CTAs per cloud (one target range each): the sort phase - most of the build - divides by S,
  // the two scans of the cloud's sources do not
  int S = 1;
  while (S < 8 && (int64_t)B * S * 2 <= num_sms() && N / (S * 2) >= 256) S *= 2;
  const int nmax = (N + S - 1) / S + 1;  // targets per CTA (upper bound)
  // every CTA scans all E sources of its cloud twice: ~0.43 ns per source, whatever the number of clouds (one wave);
  // the multi-kernel path costs ~0.067 ms of launches + 0.038 ns per source of the WHOLE batch (tools/plan_ab.py on
  // kNN lists: B=2, E=512000 fused 0.222 ms / multi-kernel 0.105; B=8 0.222 / 0.221; B=16, E=131072 0.064 / 0.200)
  const double waves = std::ceil((double)B * S / num_sms());
  const bool fused_pays = 0.43e-6 * (double)E * waves < 0.067 + 0.038e-6 * (double)E * B;
  // few targets with long source lists (three_interpolate's m = 1024 coarse points, 48 sources each): counting sort by
  // (target, chunk).  Measured against the fused builder (tools/plan_ab.py): 0.046 vs 0.087 ms at 48 sources per target,
  // 0.028 vs 0.044 at 48 with 256 targets, but 0.017 vs 0.014 and 0.030 vs 0.015 at 12 (segments the fused builder sorts
  // in registers)
  if (N >= 1 && N <= kOrdMaxN && E >= 24 * (int64_t)N && E <= 65535 && plan_ordered_smem(N, E) <= kOrdBudget && !g_force_staged &&
      !g_plan_ordered_off) {
    static DynSmem ds_ord;
    if (int rc = ds_ord.need(plan_build_ordered_kernel, kOrdBudget, "scatter_plan_build")) return rc;
    plan_build_ordered_kernel<<<B, 1024, plan_ordered_smem(N, E), s>>>(N, (int)E, idx, row_ptr, perm);
    count_launch(1);
    return check_launch("plan_build_ordered_kernel");
  }
  if (N >= 1 && nmax <= kFusedMaxN && E >= 1 && E <= kFusedMaxE && fused_pays && B <= 65535 && !g_force_staged) {
    // the whole CSR of a cloud (or of one target range of it) in shared memory (see plan_build_fused_kernel)
    const size_t kBudget = (size_t)220 * 1024;  // + 1.3 KB of static shared memory in the kernel
    const size_t smem = (size_t)(nmax + 1) * 4 + (size_t)((((E + 31) >> 5) + 1023) / 1024) * 1024 * 4;
    const size_t smem16 = smem + (size_t)((E + 1) / 2) * 4;  // + the source list as 16-bit entries
    static DynSmem ds_fused16, ds_fused32;
    // the two work lists of the sort phase take what is left, at most one entry per target each
    auto lists = [&](size_t base) { return (int)std::min<size_t>((size_t)nmax, (kBudget - base) / 8); };
    if (E <= 65535 && smem16 + 8 * 256 <= kBudget) {
      const int cap = lists(smem16);
      if (int rc = ds_fused16.need(plan_build_fused_kernel<unsigned short>, kBudget, "scatter_plan_build")) return rc;
      plan_build_fused_kernel<unsigned short><<<dim3(B, S), 1024, smem16 + (size_t)cap * 8, s>>>(N, (int)E, cap, nmax, idx, row_ptr, perm);
      count_launch(1);
      return check_launch("plan_build_fused_kernel<u16>");
    }
    if (smem + 8 * 256 <= kBudget) {
      const int cap = lists(smem);
      if (int rc = ds_fused32.need(plan_build_fused_kernel<int>, kBudget, "scatter_plan_build")) return rc;
      plan_build_fused_kernel<int><<<dim3(B, S), 1024, smem + (size_t)cap * 8, s>>>(N, (int)E, cap, nmax, idx, row_ptr, perm);
      count_launch(1);
      return check_launch("plan_build_fused_kernel<int>");
    }
  }
  cudaError_t err = cudaMemsetAsync(row_ptr, 0, (size_t)B * (N + 1) * sizeof(int), s);
  if (err != cudaSuccess) { set_error("scatter_plan_build: memset: %s", cudaGetErrorString(err)); return GEOT_ERR_CUDA; }
  if (E > 0) {
    dim3 gc((unsigned)((E + 255) / 256), B);
    plan_count_kernel<<<gc, 256, 0, s>>>(N, E, idx, row_ptr);
    count_launch(1);
    int rc = check_launch("plan_count_kernel");
    if (rc) return rc;
  }
  const size_t perm_b = plan_align((size_t)B * (size_t)E * sizeof(int));
  char* base = reinterpret_cast<char*>(plan) + plan_rowptr_bytes(B, N) + perm_b;
  int* cursor = reinterpret_cast<int*>(base);
  int* tmp = reinterpret_cast<int*>(base + plan_rowptr_bytes(B, N));
  int* heavy_count = reinterpret_cast<int*>(base + plan_rowptr_bytes(B, N) + perm_b);
  int* heavy_list = heavy_count + 64;
  const int heavy_cap = plan_heavy_cap(B, E);
  plan_scan_kernel<<<B, 1024, 0, s>>>(N, row_ptr, cursor);
  count_launch(1);
  int rc = check_launch("plan_scan_kernel");
  if (rc) return rc;
  if (E > 0 && N > 0) {
    err = cudaMemsetAsync(heavy_count, 0, sizeof(int), s);
    if (err != cudaSuccess) { set_error("scatter_plan_build: memset: %s", cudaGetErrorString(err)); return GEOT_ERR_CUDA; }
    dim3 gf((unsigned)((E + 255) / 256), B);
    plan_fill_kernel<<<gf, 256, 0, s>>>(N, E, idx, cursor, tmp);
    dim3 gs((N + 255) / 256, B);
    plan_sort_small_kernel<<<gs, 256, 0, s>>>(N, E, row_ptr, tmp, perm, heavy_count, heavy_list, heavy_cap);
    static DynSmem ds_heavy;
    if (int rc = ds_heavy.need(plan_sort_heavy_kernel, kPlanHeavyBits / 8, "scatter_plan_build")) return rc;
    const size_t bm_bytes = (size_t)std::min<int64_t>(kPlanHeavyBits, ((E + 32767) / 32768) * 32768) / 8;
    plan_sort_heavy_kernel<<<2 * num_sms(), 1024, bm_bytes, s>>>(N, E, row_ptr, tmp, perm, heavy_count, heavy_list, heavy_cap);
    count_launch(3);
    rc = check_launch("plan_fill/sort kernels");
  }
  return rc;
}

GEOT_API int geot_group_points_grad(int B, int C, int N, int M, int nsample, const float* grad_out,
                                    const void* plan, float* grad_points, geot_stream_t stream) {
  GEOT_REQUIRE(B >= 0 && C >= 0 && N >= 0 && M >= 0 && nsample >= 0, "group_points_grad: bad sizes");
  return launch_seg_reduce<1, false>(B, C, N, (int64_t)M * nsample, grad_out, nullptr, plan, grad_points,
                                     as_stream(stream), "group_points_grad");
}

GEOT_API int geot_scatter_plan_pieces(int N, int64_t E) {
  if (N <= 0 || E <= 0) return 1;
  for (int Q = 1; Q <= 64; Q++) {
    if (E % Q) continue;
    const int64_t e = E / Q;
    if (e <= 65535 && e % 4 == 0 && e >= N / 2 && planrow_smem(N, e) <= (size_t)226 * 1024) return Q;
  }
  return 1;
}

GEOT_API int geot_group_points_grad_pieces(int B, int C, int N, int M, int nsample, int pieces, const float* grad_out,
                                           const void* plan, float* grad_points, geot_stream_t stream) {
  GEOT_REQUIRE(B >= 0 && C >= 0 && N >= 0 && M >= 0 && nsample >= 0 && pieces >= 1, "group_points_grad_pieces: bad sizes");
  const int64_t E = (int64_t)M * nsample;
  if (pieces == 1) return geot_group_points_grad(B, C, N, M, nsample, grad_out, plan, grad_points, stream);
  GEOT_REQUIRE(E % pieces == 0 && E / pieces <= 65535 && (int64_t)B * pieces <= 65535, "group_points_grad_pieces: M*nsample=%lld is not divisible into %d pieces of <= 65535 edges", (long long)E, pieces);
  if (B == 0 || C == 0 || N == 0) return GEOT_OK;
  GEOT_REQUIRE(grad_out && plan && grad_points, "group_points_grad_pieces: NULL pointer");
  const int Ep = (int)(E / pieces);
  const PlanView pv = plan_view(plan, B * pieces, N);
  cudaStream_t s = as_stream(stream);
  if (!g_group_force_direct && Ep % 4 == 0 && (((uintptr_t)grad_out & 15) == 0) && planrow_smem(N, Ep) <= (size_t)226 * 1024)
    return launch_planrow(B, C, N, Ep, pieces, grad_out, pv.row_ptr, pv.perm, grad_points, s, "group_points_grad_pieces");
  GEOT_REQUIRE(C <= 65535 && B <= 65535, "group_points_grad_pieces: B or C exceeds grid limits");
  dim3 grid((N + 255) / 256, C, B);
  seg_reduce_pieces_direct_kernel<<<grid, 256, 0, s>>>(C, N, Ep, pieces, grad_out, pv.row_ptr, pv.perm, grad_points);
  count_launch(1);
  return check_launch("group_points_grad_pieces");
}

GEOT_API int geot_gather_points_grad(int B, int C, int N, int M, const float* grad_out, const void* plan,
                                     float* grad_points, geot_stream_t stream) {
  GEOT_REQUIRE(B >= 0 && C >= 0 && N >= 0 && M >= 0, "gather_points_grad: bad sizes");
  return launch_seg_reduce<1, false>(B, C, N, (int64_t)M, grad_out, nullptr, plan, grad_points,
                                     as_stream(stream), "gather_points_grad");
}
