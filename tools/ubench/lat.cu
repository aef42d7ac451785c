// Dev microbenchmark (GPU box): dependent-chain latencies of the warp primitives the FPS round is made of.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o lat lat.cu && ./lat
#include <cstdio>
#include <cuda_runtime.h>
#define REP 256
__global__ void k(long long* out, int seed) {
  __shared__ float sm[4096];
  __shared__ unsigned long long keys[128];
  const int lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 4096; i += blockDim.x) sm[i] = (float)((i * 37 + seed) & 1023);
  for (int i = threadIdx.x; i < 128; i += blockDim.x) keys[i] = i * 977 + seed;
  __syncthreads();
  long long t0, t1;
  int v = lane * 31 + seed;
  // 1. REDUX max (s32) dependent chain
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < REP; i++) v = __reduce_max_sync(0xffffffffu, v ^ i) + lane;
  t1 = clock64();
  if (threadIdx.x == 0) out[0] = (t1 - t0) / REP;
  // 2. REDUX min (u32)
  unsigned u = v;
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < REP; i++) u = __reduce_min_sync(0xffffffffu, u ^ (unsigned)i) + lane;
  t1 = clock64();
  if (threadIdx.x == 0) out[1] = (t1 - t0) / REP;
  // 3. shuffle butterfly max (5 SHFL + 5 MNMX)
  float f = (float)u;
  t0 = clock64();
#pragma unroll 4
  for (int i = 0; i < REP; i++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) f = fmaxf(f, __shfl_xor_sync(0xffffffffu, f, o));
    f += (float)lane;
  }
  t1 = clock64();
  if (threadIdx.x == 0) out[2] = (t1 - t0) / REP;
  // 4. single SHFL
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < REP; i++) f = __shfl_xor_sync(0xffffffffu, f, 1) + 1.0f;
  t1 = clock64();
  if (threadIdx.x == 0) out[3] = (t1 - t0) / REP;
  // 5. ballot
  unsigned b = (unsigned)f;
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < REP; i++) b = __ballot_sync(0xffffffffu, (b >> lane) & 1u) + i;
  t1 = clock64();
  if (threadIdx.x == 0) out[4] = (t1 - t0) / REP;
  // 6. LDS.32 pointer chase
  int p = (b & 1023);
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < REP; i++) p = ((int)sm[p] + lane) & 1023;
  t1 = clock64();
  if (threadIdx.x == 0) out[5] = (t1 - t0) / REP;
  // 7. LDS.128 chase
  const float4* s4 = reinterpret_cast<const float4*>(sm);
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < REP; i++) { float4 q = s4[(p + lane) & 1023]; p = (int)(q.x + q.w) & 1023; }
  t1 = clock64();
  if (threadIdx.x == 0) out[6] = (t1 - t0) / REP;
  // 8. FFMA dependent chain
  float a = (float)p;
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < REP; i++) a = __fmaf_rn(a, 1.0001f, 0.5f);
  t1 = clock64();
  if (threadIdx.x == 0) out[7] = (t1 - t0) / REP;
  // 9. FMNMX chain
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < REP; i++) a = fminf(a, (float)i) + 0.f * a;
  t1 = clock64();
  if (threadIdx.x == 0) out[8] = (t1 - t0) / REP;
  // 10. __syncthreads round trip (all warps)
  t0 = clock64();
#pragma unroll 4
  for (int i = 0; i < REP; i++) __syncthreads();
  t1 = clock64();
  if (threadIdx.x == 0) out[9] = (t1 - t0) / REP;
  // 11. STS -> barrier -> LDS of another warp's value (hand-off through smem)
  t0 = clock64();
  int h = p;
#pragma unroll 4
  for (int i = 0; i < REP; i++) {
    if (lane == 0) sm[threadIdx.x >> 5] = (float)(h + i);
    __syncthreads();
    h = (int)sm[((threadIdx.x >> 5) + 1) % (blockDim.x >> 5)];
    __syncthreads();
  }
  t1 = clock64();
  if (threadIdx.x == 0) out[10] = (t1 - t0) / REP;
  // 12. match_any / u64 max via two REDUX
  unsigned long long k64 = keys[lane] + h;
  t0 = clock64();
#pragma unroll 8
  for (int i = 0; i < REP; i++) {
    const unsigned hi = (unsigned)(k64 >> 32), lo = (unsigned)k64;
    const unsigned gh = __reduce_max_sync(0xffffffffu, hi);
    const unsigned gl = __reduce_max_sync(0xffffffffu, hi == gh ? lo : 0u);
    k64 = (((unsigned long long)gh << 32) | gl) + lane + i;
  }
  t1 = clock64();
  if (threadIdx.x == 0) out[11] = (t1 - t0) / REP;
  // 13. redux via shfl on 64-bit keys (5 x 2 SHFL + compare)
  t0 = clock64();
#pragma unroll 4
  for (int i = 0; i < REP; i++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { const unsigned long long other = __shfl_xor_sync(0xffffffffu, k64, o); k64 = other > k64 ? other : k64; }
    k64 += lane + i;
  }
  t1 = clock64();
  if (threadIdx.x == 0) out[12] = (t1 - t0) / REP;
  if (threadIdx.x == 0) out[13] = (long long)(v + u + f + b + p + a + h + (int)k64);
}
int main() {
  long long* d; cudaMalloc(&d, 16 * 8);
  const char* names[] = {"REDUX.max s32 (+add)", "REDUX.min u32 (+add)", "5x SHFL+FMNMX butterfly (+add)", "SHFL (+add)", "ballot (+shift,and,add)",
                         "LDS.32 chase (+cvt,add,and)", "LDS.128 chase", "FFMA", "FMNMX+FFMA", "__syncthreads", "STS->bar->LDS->bar", "2x REDUX u64 max", "shfl u64 max butterfly"};
  for (int nt : {32, 256, 1024}) {
    k<<<1, nt>>>(d, 3); cudaDeviceSynchronize();
    k<<<1, nt>>>(d, 5); cudaDeviceSynchronize();
    long long h[16]; cudaMemcpy(h, d, 16 * 8, cudaMemcpyDeviceToHost);
    printf("threads=%d\n", nt);
    for (int i = 0; i < 13; i++) printf("  %-34s %4lld cycles\n", names[i], h[i]);
  }
  printf("err=%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
