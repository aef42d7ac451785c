// geot_b200/csrc/fps_common.cuh - helpers shared by the FPS kernels (fps.cu, fps_rw.cu): the reference's tie order in
// closed form, segment lookup for the batched / offset formats, Morton codes, narrow shared-memory accesses.
#pragma once
#include "common.cuh"

namespace geot {

void count_launch(int n);

struct TieRule {
  int log2bs;  // BS = 1 << log2bs, the reference's block size for this n
};

__device__ __forceinline__ unsigned tie_rank(int k, int log2bs, int q_count) {
  const unsigned tid = (unsigned)k & ((1u << log2bs) - 1u);
  const unsigned q = (unsigned)k >> log2bs;
  const unsigned br = log2bs ? (__brev(tid) >> (32 - log2bs)) : 0u;
  return br * (unsigned)q_count + q;
}
__device__ __forceinline__ int tie_unrank(unsigned rank, int log2bs, int q_count) {
  const unsigned br = rank / (unsigned)q_count, q = rank % (unsigned)q_count;
  const unsigned tid = log2bs ? (__brev(br) >> (32 - log2bs)) : 0u;
  return (int)((q << log2bs) | tid);
}

// Segment of block b.  Batched: offset == nullptr, indices local.  Offset format: global indices.
__device__ __forceinline__ void fps_segment(const int* __restrict__ offset, const int* __restrict__ new_offset,
                                            int N, int M, int b, int64_t& start_n, int& n,
                                            int64_t& start_m, int& m, int& idx_base) {
  if (offset == nullptr) {
    start_n = (int64_t)b * N; n = N; start_m = (int64_t)b * M; m = M; idx_base = 0;
  } else {
    const int s_n = b ? offset[b - 1] : 0, s_m = b ? new_offset[b - 1] : 0;
    start_n = s_n; n = offset[b] - s_n; start_m = s_m; m = new_offset[b] - s_m; idx_base = s_n;
  }
}

__device__ __forceinline__ unsigned morton4(unsigned v) {  // 4 bits -> every third bit
  return (v & 1u) | ((v & 2u) << 2) | ((v & 4u) << 4) | ((v & 8u) << 6);
}

__device__ __forceinline__ void bar_sync_named(int id) { asm volatile("bar.sync %0, 1024;" ::"r"(id) : "memory"); }
__device__ __forceinline__ void bar_arrive_named(int id) { asm volatile("bar.arrive %0, 1024;" ::"r"(id) : "memory"); }
__device__ __forceinline__ unsigned lds_u8(unsigned a) {
  unsigned v;
  asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts_u8(unsigned a, unsigned v) { asm volatile("st.shared.u8 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ void sts_u16(unsigned a, unsigned v) { asm volatile("st.shared.u16 [%0], %1;" ::"r"(a), "h"((unsigned short)v) : "memory"); }
__device__ __forceinline__ void sts32(unsigned a, unsigned v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }

}  // namespace geot
