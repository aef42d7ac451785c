/*
 * oracle/cpu_ref.c — CPU restatement of the reference's CUDA kernel bodies for the GeoT
 * sampling / grouping / interpolation hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library; nothing under geot_b200/ may.  The reference
 * ships NO CPU implementation of these ops ("CPU not supported", pointnet2/_ext_src/src/
 * ball_query.cpp:31) so this file restates each __global__ body in plain C, thread loop by
 * thread loop, with the floating-point evaluation order nvcc -O2 (default -fmad=true) gives the
 * reference on sm_100 — read off the SASS of oracle/_ref:
 *
 *     d2 = fma(dz, dz, fma(dx, dx, round(dy*dy)))            FMUL ; FFMA ; FFMA
 *
 * in furthest_point_sampling_kernel, ball_query_kernel_fast, three_nn_kernel_fast and
 * knnquery_cuda_kernel alike.  Compile with -ffp-contract=off so only the explicit fmaf()s fuse.
 *
 * Pinning: the reference has no golden vectors for these ops.  This restatement is
 * pinned against the reference's own kernels, compiled unchanged into oracle/_ref and run on a
 * B200 (tests/test_ref_parity_gpu.py; fixtures generated there live in tests/golden/).
 *
 * Every function cites the reference file:line it follows (paths relative to /root/reference).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define API __attribute__((visibility("default")))

/* (a - b) squared distance in the reference's contracted order. */
static inline float d2_ref(float ax, float ay, float az, float bx, float by, float bz) {
  float dx = ax - bx, dy = ay - by, dz = az - bz;
  return fmaf(dz, dz, fmaf(dx, dx, dy * dy));
}

API int oracle_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
/* torchrun exports OMP_NUM_THREADS=1; the timed CPU baseline asks for all host threads explicitly. */
API void oracle_set_num_threads(int n) { if (n > 0) omp_set_num_threads(n); }


/* openpoints/cpp/pointnet2_batch/src/cuda_utils.h:10-14 (cap 1024);
 * pointnet2/_ext_src/include/cuda_utils.h:15-21 (cap 512). */
API int oracle_opt_n_threads(int work_size, int cap) {
  const int pow_2 = (int)(log((double)work_size) / log(2.0));
  int v = 1 << pow_2;
  if (v > cap) v = cap;
  if (v < 1) v = 1;
  return v;
}

/* ------------------------------------------------------------------------------------------
 * Furthest point sampling — one "block" (cloud / segment).
 * Follows openpoints/cpp/pointnet2_batch/src/sampling_gpu.cu:101-216 (batched, first = 0),
 * pointops/src/sampling/sampling_cuda_kernel.cu:15-129 (offset form, first = segment start),
 * pointnet2/_ext_src/src/sampling_gpu.cu:73-176 (|p|^2 <= 1e-3 skip :103-104),
 * pointops/src/sampling/sampling_cuda_kernel.cu:175-307 (weights :222).
 * The thread loop and the shared-memory tree (__update, sampling_gpu.cu:93-98) are emulated
 * literally so the tie behaviour is the reference's by construction.
 * xyz/tmp are the *global* arrays; [start_n, end_n) is this block's slice; out receives
 * m_count global indices.
 * ------------------------------------------------------------------------------------------ */
static void fps_block(const float* xyz, int start_n, int end_n, int m_count, int BS, int mag_skip,
                      const float* weights, float* tmp, int* out, float* dists, int* dists_i) {
  if (m_count <= 0) return;
  int old = start_n;
  out[0] = old;
  for (int j = 1; j < m_count; j++) {
    const float x1 = xyz[old * 3 + 0], y1 = xyz[old * 3 + 1], z1 = xyz[old * 3 + 2];
    for (int tid = 0; tid < BS; tid++) {
      int besti = start_n;
      float best = -1.0f;
      for (int k = start_n + tid; k < end_n; k += BS) {
        const float x2 = xyz[k * 3 + 0], y2 = xyz[k * 3 + 1], z2 = xyz[k * 3 + 2];
        if (mag_skip) {
          const float mag = fmaf(z2, z2, fmaf(x2, x2, y2 * y2));
          if ((double)mag <= 1e-3) continue;
        }
        float d = d2_ref(x2, y2, z2, x1, y1, z1);
        if (weights) d = (float)((double)d * fmax((double)weights[k], 1e-12));
        const float d2 = fminf(d, tmp[k]);
        tmp[k] = d2;
        besti = d2 > best ? k : besti;
        best = d2 > best ? d2 : best;
      }
      dists[tid] = best;
      dists_i[tid] = besti;
    }
    for (int s = BS / 2; s >= 1; s >>= 1) {
      for (int tid = 0; tid < s; tid++) {
        const float v1 = dists[tid], v2 = dists[tid + s];
        const int i1 = dists_i[tid], i2 = dists_i[tid + s];
        dists[tid] = fmaxf(v1, v2);
        dists_i[tid] = v2 > v1 ? i2 : i1;
      }
    }
    old = dists_i[0];
    out[j] = old;
  }
}

/* Batched FPS: xyz (B,N,3), temp (B,N) pre-filled by the caller with 1e10, idx (B,M).
 * cap = 1024 (pointnet2_batch) or 512 (pointnet2/_ext); mag_skip = 1 for pointnet2/_ext. */
API void oracle_fps_batched(int B, int N, int M, const float* xyz, float* temp, int* idx, int cap,
                            int mag_skip) {
  const int BS = oracle_opt_n_threads(N, cap);
#pragma omp parallel
  {
    float* dists = (float*)malloc(sizeof(float) * BS);
    int* dists_i = (int*)malloc(sizeof(int) * BS);
#pragma omp for schedule(dynamic, 1)
    for (int b = 0; b < B; b++)
      fps_block(xyz + (size_t)b * N * 3, 0, N, M, BS, mag_skip, NULL, temp + (size_t)b * N,
                idx + (size_t)b * M, dists, dists_i);
    free(dists);
    free(dists_i);
  }
}

/* Offset-format FPS (global indices): xyz (n,3), offset/new_offset cumulative (b), tmp (n) = 1e10.
 * n_max only chooses the block size (sampling_cuda_kernel.cu:131-133). weights may be NULL. */
API void oracle_furthestsampling(int b, int n_max, const float* xyz, const int* offset,
                                 const int* new_offset, const float* weights, float* tmp, int* idx) {
  const int BS = oracle_opt_n_threads(n_max, 1024);
#pragma omp parallel
  {
    float* dists = (float*)malloc(sizeof(float) * BS);
    int* dists_i = (int*)malloc(sizeof(int) * BS);
#pragma omp for schedule(dynamic, 1)
    for (int bid = 0; bid < b; bid++) {
      const int start_n = bid ? offset[bid - 1] : 0, end_n = offset[bid];
      const int start_m = bid ? new_offset[bid - 1] : 0, end_m = new_offset[bid];
      fps_block(xyz, start_n, end_n, end_m - start_m, BS, 0, weights, tmp, idx + start_m, dists,
                dists_i);
    }
    free(dists);
    free(dists_i);
  }
}

/* ------------------------------------------------------------------------------------------
 * Ball query — openpoints/cpp/pointnet2_batch/src/ball_query_gpu.cu:15-51 (twin
 * pointnet2/_ext_src/src/ball_query_gpu.cu:12-47).  idx must be zero-filled by the caller
 * (group.py:194) — rows without a hit are left untouched.
 * ------------------------------------------------------------------------------------------ */
static inline void ball_row(const float* xyz, int start, int end, float nx, float ny, float nz,
                            float radius2, int nsample, int* idx) {
  int cnt = 0;
  for (int k = start; k < end; ++k) {
    const float d2 = d2_ref(nx, ny, nz, xyz[k * 3 + 0], xyz[k * 3 + 1], xyz[k * 3 + 2]);
    if (d2 < radius2) {
      if (cnt == 0)
        for (int l = 0; l < nsample; ++l) idx[l] = k;
      idx[cnt] = k;
      ++cnt;
      if (cnt >= nsample) break;
    }
  }
}

API void oracle_ball_query(int B, int N, int M, float radius, int nsample, const float* new_xyz,
                           const float* xyz, int* idx) {
  const float radius2 = radius * radius;
#pragma omp parallel for collapse(2) schedule(static)
  for (int b = 0; b < B; b++)
    for (int j = 0; j < M; j++) {
      const float* q = new_xyz + ((size_t)b * M + j) * 3;
      ball_row(xyz + (size_t)b * N * 3, 0, N, q[0], q[1], q[2], radius2, nsample,
               idx + ((size_t)b * M + j) * nsample);
    }
}

/* get_bt_idx: pointops/src/knnquery/knnquery_cuda_kernel.cu:51-62 */
static inline int bt_idx(int i, const int* offset) {
  int b = 0;
  while (!(i < offset[b])) b++;
  return b;
}

/* Offset-format ball query — openpoints/cpp/pointops/src/ballquery/ballquery_cuda_kernel.cu:26-78. */
API void oracle_ballquery_offset(int m, float radius, int nsample, const float* xyz,
                                 const float* new_xyz, const int* offset, const int* new_offset,
                                 int* idx) {
  const float radius2 = radius * radius;
#pragma omp parallel for schedule(static)
  for (int p = 0; p < m; p++) {
    const int bt = bt_idx(p, new_offset);
    const int start = bt ? offset[bt - 1] : 0, end = offset[bt];
    ball_row(xyz, start, end, new_xyz[p * 3 + 0], new_xyz[p * 3 + 1], new_xyz[p * 3 + 2], radius2,
             nsample, idx + (size_t)p * nsample);
  }
}

/* ------------------------------------------------------------------------------------------
 * kNN query — pointops/src/knnquery/knnquery_cuda_kernel.cu:21-108 (reheap :21-36,
 * heap_sort :39-48, kernel :65-108).  Max-heap of size nsample seeded with (1e10, start);
 * strict '<' insert at the root; final heap sort → ascending d2.  Emulated literally so that
 * membership and order under exact distance ties are the reference's.
 * ------------------------------------------------------------------------------------------ */
static void reheap(float* dist, int* idx, int k) {
  int root = 0, child = 1;
  while (child < k) {
    if (child + 1 < k && dist[child + 1] > dist[child]) child++;
    if (dist[root] > dist[child]) return;
    float tf = dist[root]; dist[root] = dist[child]; dist[child] = tf;
    int ti = idx[root]; idx[root] = idx[child]; idx[child] = ti;
    root = child;
    child = root * 2 + 1;
  }
}

static void heap_sort(float* dist, int* idx, int k) {
  for (int i = k - 1; i > 0; i--) {
    float tf = dist[0]; dist[0] = dist[i]; dist[i] = tf;
    int ti = idx[0]; idx[0] = idx[i]; idx[i] = ti;
    reheap(dist, idx, i);
  }
}

API void oracle_knnquery(int m, int nsample, const float* xyz, const float* new_xyz,
                         const int* offset, const int* new_offset, int* idx, float* dist2) {
#pragma omp parallel
  {
    float* bd = (float*)malloc(sizeof(float) * (nsample > 0 ? nsample : 1));
    int* bi = (int*)malloc(sizeof(int) * (nsample > 0 ? nsample : 1));
#pragma omp for schedule(static)
    for (int p = 0; p < m; p++) {
      const int bt = bt_idx(p, new_offset);
      const int start = bt ? offset[bt - 1] : 0, end = offset[bt];
      const float nx = new_xyz[p * 3 + 0], ny = new_xyz[p * 3 + 1], nz = new_xyz[p * 3 + 2];
      for (int i = 0; i < nsample; i++) { bd[i] = 1e10f; bi[i] = start; }
      for (int i = start; i < end; i++) {
        const float d2 = d2_ref(nx, ny, nz, xyz[i * 3 + 0], xyz[i * 3 + 1], xyz[i * 3 + 2]);
        if (d2 < bd[0]) { bd[0] = d2; bi[0] = i; reheap(bd, bi, nsample); }
      }
      heap_sort(bd, bi, nsample);
      for (int i = 0; i < nsample; i++) {
        idx[(size_t)p * nsample + i] = bi[i];
        dist2[(size_t)p * nsample + i] = bd[i];
      }
    }
    free(bd);
    free(bi);
  }
}

/* ------------------------------------------------------------------------------------------
 * three_nn — openpoints/cpp/pointnet2_batch/src/interpolate_gpu.cu:16-59 (twin
 * pointnet2/_ext_src/src/interpolate_gpu.cu:12-62).  double-typed bests vs float d (exact).
 * ------------------------------------------------------------------------------------------ */
API void oracle_three_nn(int B, int n, int m, const float* unknown, const float* known,
                         float* dist2, int* idx) {
#pragma omp parallel for collapse(2) schedule(static)
  for (int b = 0; b < B; b++)
    for (int p = 0; p < n; p++) {
      const float* u = unknown + ((size_t)b * n + p) * 3;
      const float* kn = known + (size_t)b * m * 3;
      double best1 = 1e40, best2 = 1e40, best3 = 1e40;
      int besti1 = 0, besti2 = 0, besti3 = 0;
      for (int k = 0; k < m; ++k) {
        const float d = d2_ref(u[0], u[1], u[2], kn[k * 3 + 0], kn[k * 3 + 1], kn[k * 3 + 2]);
        if (d < best1) {
          best3 = best2; besti3 = besti2; best2 = best1; besti2 = besti1; best1 = d; besti1 = k;
        } else if (d < best2) {
          best3 = best2; besti3 = besti2; best2 = d; besti2 = k;
        } else if (d < best3) {
          best3 = d; besti3 = k;
        }
      }
      float* o = dist2 + ((size_t)b * n + p) * 3;
      int* oi = idx + ((size_t)b * n + p) * 3;
      o[0] = (float)best1; o[1] = (float)best2; o[2] = (float)best3;
      oi[0] = besti1; oi[1] = besti2; oi[2] = besti3;
    }
}

/* three_interpolate fwd — interpolate_gpu.cu:84-104; SASS order fma(w2,p2,fma(w0,p0,w1*p1)). */
API void oracle_three_interpolate(int B, int c, int m, int n, const float* points, const int* idx,
                                  const float* weight, float* out) {
#pragma omp parallel for collapse(2) schedule(static)
  for (int b = 0; b < B; b++)
    for (int ci = 0; ci < c; ci++) {
      const float* p = points + ((size_t)b * c + ci) * m;
      for (int j = 0; j < n; j++) {
        const int* id = idx + ((size_t)b * n + j) * 3;
        const float* w = weight + ((size_t)b * n + j) * 3;
        out[((size_t)b * c + ci) * n + j] = fmaf(w[2], p[id[2]], fmaf(w[0], p[id[0]], w[1] * p[id[1]]));
      }
    }
}

/* three_interpolate bwd — interpolate_gpu.cu:127-149.  The reference's float atomics have no
 * defined order, so the oracle sums in double (order-free to ~1e-16) and rounds once. */
API void oracle_three_interpolate_grad(int B, int c, int n, int m, const float* grad_out,
                                       const int* idx, const float* weight, float* grad_points) {
#pragma omp parallel
  {
    double* acc = (double*)malloc(sizeof(double) * (m > 0 ? m : 1));
#pragma omp for collapse(2) schedule(static)
    for (int b = 0; b < B; b++)
      for (int ci = 0; ci < c; ci++) {
        memset(acc, 0, sizeof(double) * m);
        const float* g = grad_out + ((size_t)b * c + ci) * n;
        for (int j = 0; j < n; j++) {
          const int* id = idx + ((size_t)b * n + j) * 3;
          const float* w = weight + ((size_t)b * n + j) * 3;
          for (int t = 0; t < 3; t++) acc[id[t]] += (double)(g[j] * w[t]);
        }
        float* o = grad_points + ((size_t)b * c + ci) * m;
        for (int k = 0; k < m; k++) o[k] = (float)acc[k];
      }
    free(acc);
  }
}

/* group_points fwd — group_points_gpu.cu:53-72. */
API void oracle_group_points(int B, int c, int n, int npoints, int nsample, const float* points,
                             const int* idx, float* out) {
#pragma omp parallel for collapse(2) schedule(static)
  for (int b = 0; b < B; b++)
    for (int ci = 0; ci < c; ci++) {
      const float* p = points + ((size_t)b * c + ci) * n;
      const int* id = idx + (size_t)b * npoints * nsample;
      float* o = out + ((size_t)b * c + ci) * npoints * nsample;
      for (size_t e = 0; e < (size_t)npoints * nsample; e++) o[e] = p[id[e]];
    }
}

/* group_points bwd — group_points_gpu.cu:14-31 (atomicAdd :30) → double accumulation. */
API void oracle_group_points_grad(int B, int c, int n, int npoints, int nsample,
                                  const float* grad_out, const int* idx, float* grad_points) {
#pragma omp parallel
  {
    double* acc = (double*)malloc(sizeof(double) * (n > 0 ? n : 1));
#pragma omp for collapse(2) schedule(static)
    for (int b = 0; b < B; b++)
      for (int ci = 0; ci < c; ci++) {
        memset(acc, 0, sizeof(double) * n);
        const float* g = grad_out + ((size_t)b * c + ci) * npoints * nsample;
        const int* id = idx + (size_t)b * npoints * nsample;
        for (size_t e = 0; e < (size_t)npoints * nsample; e++) acc[id[e]] += (double)g[e];
        float* o = grad_points + ((size_t)b * c + ci) * n;
        for (int k = 0; k < n; k++) o[k] = (float)acc[k];
      }
    free(acc);
  }
}

/* gather_points fwd — sampling_gpu.cu:15-31;  bwd — :53-70 (atomicAdd :69). */
API void oracle_gather_points(int B, int c, int n, int npoints, const float* points, const int* idx,
                              float* out) {
  oracle_group_points(B, c, n, npoints, 1, points, idx, out);
}
API void oracle_gather_points_grad(int B, int c, int n, int npoints, const float* grad_out,
                                   const int* idx, float* grad_points) {
  oracle_group_points_grad(B, c, n, npoints, 1, grad_out, idx, grad_points);
}

/* ------------------------------------------------------------------------------------------
 * Offset-format (channel-last) ops — openpoints/cpp/pointops/src/ (all sub-directories).
 * ------------------------------------------------------------------------------------------ */
/* grouping fwd — grouping/grouping_cuda_kernel.cu:5-14 */
API void oracle_grouping_forward(int m, int nsample, int c, const float* input, const int* idx,
                                 float* output) {
#pragma omp parallel for schedule(static)
  for (long e = 0; e < (long)m * nsample; e++)
    memcpy(output + (size_t)e * c, input + (size_t)idx[e] * c, sizeof(float) * c);
}
/* grouping bwd — :16-25 (atomicAdd :24); n = rows of grad_input. */
API void oracle_grouping_backward(int m, int nsample, int c, int n, const float* grad_output,
                                  const int* idx, float* grad_input) {
  double* acc = (double*)calloc((size_t)n * c, sizeof(double));
  for (long e = 0; e < (long)m * nsample; e++)
    for (int ci = 0; ci < c; ci++) acc[(size_t)idx[e] * c + ci] += (double)grad_output[(size_t)e * c + ci];
  for (size_t t = 0; t < (size_t)n * c; t++) grad_input[t] = (float)acc[t];
  free(acc);
}
/* interpolation fwd — interpolation/interpolation_cuda_kernel.cu:5-18: output pre-zeroed by the
 * caller, out = fma(in, w, out) for i = 0..k-1 in order. */
API void oracle_interpolation_forward(int n, int c, int k, const float* input, const int* idx,
                                      const float* weight, float* output) {
#pragma omp parallel for schedule(static)
  for (int p = 0; p < n; p++)
    for (int ci = 0; ci < c; ci++) {
      float o = output[(size_t)p * c + ci];
      for (int i = 0; i < k; i++)
        o = fmaf(input[(size_t)idx[p * k + i] * c + ci], weight[p * k + i], o);
      output[(size_t)p * c + ci] = o;
    }
}
/* interpolation bwd — :20-33 (atomicAdd :31); m = rows of grad_input. */
API void oracle_interpolation_backward(int n, int c, int k, int m, const float* grad_output,
                                       const int* idx, const float* weight, float* grad_input) {
  double* acc = (double*)calloc((size_t)m * c, sizeof(double));
  for (int p = 0; p < n; p++)
    for (int i = 0; i < k; i++)
      for (int ci = 0; ci < c; ci++)
        acc[(size_t)idx[p * k + i] * c + ci] += (double)(grad_output[(size_t)p * c + ci] * weight[p * k + i]);
  for (size_t t = 0; t < (size_t)m * c; t++) grad_input[t] = (float)acc[t];
  free(acc);
}
/* subtraction fwd — subtraction/subtraction_cuda_kernel.cu:5-16 */
API void oracle_subtraction_forward(int n, int nsample, int c, const float* input1,
                                    const float* input2, const int* idx, float* output) {
#pragma omp parallel for schedule(static)
  for (int p = 0; p < n; p++)
    for (int s = 0; s < nsample; s++)
      for (int ci = 0; ci < c; ci++)
        output[((size_t)p * nsample + s) * c + ci] =
            input1[(size_t)p * c + ci] - input2[(size_t)idx[p * nsample + s] * c + ci];
}
/* subtraction bwd — :18-30 (two atomicAdds) */
API void oracle_subtraction_backward(int n, int nsample, int c, const int* idx,
                                     const float* grad_output, float* grad_input1,
                                     float* grad_input2) {
  double* a1 = (double*)calloc((size_t)n * c, sizeof(double));
  double* a2 = (double*)calloc((size_t)n * c, sizeof(double));
  for (int p = 0; p < n; p++)
    for (int s = 0; s < nsample; s++)
      for (int ci = 0; ci < c; ci++) {
        const double g = (double)grad_output[((size_t)p * nsample + s) * c + ci];
        a1[(size_t)p * c + ci] += g;
        a2[(size_t)idx[p * nsample + s] * c + ci] -= g;
      }
  for (size_t t = 0; t < (size_t)n * c; t++) { grad_input1[t] = (float)a1[t]; grad_input2[t] = (float)a2[t]; }
  free(a1);
  free(a2);
}
/* aggregation fwd — aggregation/aggregation_cuda_kernel.cu:5-20: out = fma(in+pos, w, out). */
API void oracle_aggregation_forward(int n, int nsample, int c, int w_c, const float* input,
                                    const float* position, const float* weight, const int* idx,
                                    float* output) {
#pragma omp parallel for schedule(static)
  for (int p = 0; p < n; p++)
    for (int ci = 0; ci < c; ci++) {
      float o = output[(size_t)p * c + ci];
      for (int s = 0; s < nsample; s++) {
        const float a = input[(size_t)idx[p * nsample + s] * c + ci] +
                        position[((size_t)p * nsample + s) * c + ci];
        o = fmaf(a, weight[((size_t)p * nsample + s) * w_c + ci % w_c], o);
      }
      output[(size_t)p * c + ci] = o;
    }
}
/* aggregation bwd — :22-39 */
API void oracle_aggregation_backward(int n, int nsample, int c, int w_c, const float* input,
                                     const float* position, const float* weight, const int* idx,
                                     const float* grad_output, float* grad_input,
                                     float* grad_position, float* grad_weight) {
  double* gi = (double*)calloc((size_t)n * c, sizeof(double));
  double* gw = (double*)calloc((size_t)n * nsample * w_c, sizeof(double));
  for (int p = 0; p < n; p++)
    for (int ci = 0; ci < c; ci++)
      for (int s = 0; s < nsample; s++) {
        const size_t ii = (size_t)idx[p * nsample + s] * c + ci;
        const size_t pi = ((size_t)p * nsample + s) * c + ci;
        const size_t wi = ((size_t)p * nsample + s) * w_c + ci % w_c;
        const float g = grad_output[(size_t)p * c + ci];
        gi[ii] += (double)(g * weight[wi]);
        grad_position[pi] = g * weight[wi];
        gw[wi] += (double)(g * (input[ii] + position[pi]));
      }
  for (size_t t = 0; t < (size_t)n * c; t++) grad_input[t] = (float)gi[t];
  for (size_t t = 0; t < (size_t)n * nsample * w_c; t++) grad_weight[t] = (float)gw[t];
  free(gi);
  free(gw);
}
