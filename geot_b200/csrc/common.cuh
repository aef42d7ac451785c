// geot_b200/csrc/common.cuh — shared helpers for the sm_100a kernels behind include/geot_b200.h.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <cmath>
#include "../../include/geot_b200.h"

#define GEOT_API extern "C" __attribute__((visibility("default")))

namespace geot {

void set_error(const char* fmt, ...);

// Launch-error plumbing: the reference exit(-1)s on a failed launch (sampling_gpu.cu:255-259);
// the C ABI returns a code and keeps a per-thread message instead.
inline int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return GEOT_ERR_CUDA;
  }
  return GEOT_OK;
}

#define GEOT_REQUIRE(cond, ...)          \
  do {                                   \
    if (!(cond)) {                       \
      ::geot::set_error(__VA_ARGS__);    \
      return GEOT_ERR_INVALID;           \
    }                                    \
  } while (0)

inline cudaStream_t as_stream(geot_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

// The reference's block-size rule (cuda_utils.h:10-14): largest power of two <= n, capped.
// It decides the FPS tie order, so it is reproduced with the same double-log expression.
inline int ref_opt_n_threads(int work_size, int cap) {
  const int pow_2 = (int)(std::log(static_cast<double>(work_size)) / std::log(2.0));
  int v = 1 << pow_2;
  if (v > cap) v = cap;
  if (v < 1) v = 1;
  return v;
}

// Squared distance in the evaluation order nvcc gives the reference kernels on sm_100
// (FMUL dy*dy ; FFMA dx*dx+. ; FFMA dz*dz+.) — explicit _rn intrinsics are never re-contracted.
__device__ __forceinline__ float dist2_ref(float ax, float ay, float az, float bx, float by, float bz) {
  const float dx = __fsub_rn(ax, bx), dy = __fsub_rn(ay, by), dz = __fsub_rn(az, bz);
  return __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
}

__device__ __forceinline__ unsigned lanemask_lt() {
  unsigned m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

// Explicit shared-memory accesses by 32-bit shared address (volatile: never cached across a barrier).
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ float4 lds128(unsigned a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ uint2 lds64(unsigned a) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a));
  return v;
}
__device__ __forceinline__ float lds32f(unsigned a) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ unsigned lds_u16(unsigned a) {
  unsigned short v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts64(unsigned a, unsigned x, unsigned y) {
  asm volatile("st.shared.v2.u32 [%0], {%1,%2};" ::"r"(a), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ void sts128(unsigned a, float4 v) {
  asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(a), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// SM count, asked from the driver once per process (the device current at the first call; 148 on B200).  It sizes
// grids and picks kernel variants, no result depends on it; the library is built for sm_100a only, so the devices of
// one process share it.  (A per-call cudaGetDevice here showed up as +2.5 us on the 8-us gather launches.)
inline int num_sms() {
  static const int cached = [] {
    int dev = 0, v = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) return 148;
    return v;
  }();
  return cached;
}

// cudaFuncAttributeMaxDynamicSharedMemorySize belongs to the function IN ONE DEVICE'S CONTEXT, not to the process: a
// launcher keeps one static DynSmem per kernel and calls need() before every launch; the attribute is (re)set when the
// CURRENT device has not been configured for that many bytes yet (one cudaGetDevice, thread-local, per launch).
struct DynSmem {
  size_t configured[64] = {};
  template <typename K>
  int need(K kernel, size_t bytes, const char* what, bool prefer_smem_carveout = false) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e == cudaSuccess && dev >= 0 && dev < 64 && bytes <= configured[dev]) return GEOT_OK;
    if (e == cudaSuccess) e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e == cudaSuccess && prefer_smem_carveout) e = cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
    if (e != cudaSuccess) {
      set_error("%s: cudaFuncSetAttribute(%zu bytes of dynamic shared memory): %s", what, bytes, cudaGetErrorString(e));
      return GEOT_ERR_CUDA;
    }
    if (dev >= 0 && dev < 64) configured[dev] = bytes;  // racing threads write the same value
    return GEOT_OK;
  }
};

}  // namespace geot
