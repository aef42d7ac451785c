// geot_b200/csrc/stats_exchange.cu - the one exchange of the multi-GPU path, fused around ONE collective.
//
// The reference is single-process and reduces nothing (examples/segmentation/train.py:7); with clouds sharded by rank
// the per-step transition-matrix statistics have to be combined: the class-anchor rows of train.py:509-526 (an ARG-MAX
// pick, not a sum), the count-weighted class means of cal_mean_feature (train.py:884-895) and the T-predictor
// gradients.  Instead of three collectives and ~50 tiny tensor ops per step, every rank packs ONE record
//     [ best_val (C) | class_T (C*C) | cm * c_num (C*C) | c_num (C) | grads (P) ]
// (geot_stats_pack), the records are all-gathered once (NCCL over NVLink, W * 297 KB at C=33), and every rank
// combines them itself (geot_stats_combine): arg-max with the lowest rank winning ties - identical to the
// single-process first-max over the concatenated batch, because ranks hold contiguous blocks of clouds in rank order -
// and sums in FIXED rank order, so every rank computes bit-identical results without a broadcast.
#include "common.cuh"

namespace geot {
void count_launch(int n);

__global__ void stats_pack_kernel(int C, int64_t P, const float* __restrict__ class_T, const float* __restrict__ best_val,
                                  const float* __restrict__ cm, const float* __restrict__ c_num,
                                  const float* __restrict__ grads, float* __restrict__ rec) {
  const int64_t CC = (int64_t)C * C, L = 2 * CC + 2 * C + P;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < L; i += (int64_t)gridDim.x * blockDim.x) {
    float v;
    if (i < C) v = best_val[i];
    else if (i < C + CC) v = class_T[i - C];
    else if (i < C + 2 * CC) { const int64_t k = i - C - CC; v = cm[k] * c_num[k / C]; }
    else if (i < 2 * C + 2 * CC) v = c_num[i - C - 2 * CC];
    else v = grads ? grads[i - 2 * C - 2 * CC] : 0.f;
    rec[i] = v;
  }
}

__global__ void stats_combine_kernel(int W, int C, int64_t P, const float* __restrict__ all, float grad_scale,
                                     float* __restrict__ class_T, float* __restrict__ cm, float* __restrict__ c_num,
                                     float* __restrict__ grads) {
  const int64_t CC = (int64_t)C * C, L = 2 * CC + 2 * C + P;
  const int64_t total = 2 * CC + C + P;  // outputs: class_T | cm | c_num | grads
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    if (i < CC) {  // anchor row of class c: the rank with the largest score, lowest rank on ties
      const int c = (int)(i / C);
      int win = 0;
      float best = all[c];
      for (int r = 1; r < W; r++) {
        const float v = all[r * L + c];
        if (v > best) { best = v; win = r; }
      }
      class_T[i] = all[win * L + C + i];
    } else if (i < 2 * CC) {  // count-weighted class means
      const int64_t k = i - CC;
      const int c = (int)(k / C);
      float num = 0.f, tot = 0.f;
      for (int r = 0; r < W; r++) { num += all[r * L + C + CC + k]; tot += all[r * L + C + 2 * CC + c]; }
      if (tot > 0.f) cm[k] = num / tot;  // a class no rank has seen keeps this rank's value
    } else if (i < 2 * CC + C) {
      const int c = (int)(i - 2 * CC);
      float tot = 0.f;
      for (int r = 0; r < W; r++) tot += all[r * L + C + 2 * CC + c];
      c_num[c] = tot;
    } else if (grads) {
      const int64_t k = i - 2 * CC - C;
      float s = 0.f;
      for (int r = 0; r < W; r++) s += all[r * L + 2 * C + 2 * CC + k];
      grads[k] = s * grad_scale;
    }
  }
}
}  // namespace geot

using namespace geot;

GEOT_API int64_t geot_stats_record_floats(int C, int64_t P) { return 2 * (int64_t)C * C + 2 * C + P; }

GEOT_API int geot_stats_pack(int C, int64_t P, const float* class_T, const float* best_val, const float* cm,
                             const float* c_num, const float* grads, float* record, geot_stream_t stream) {
  GEOT_REQUIRE(C >= 1 && P >= 0, "stats_pack: bad sizes C=%d P=%lld", C, (long long)P);
  GEOT_REQUIRE(class_T && best_val && cm && c_num && record, "stats_pack: NULL pointer");
  const int64_t L = geot_stats_record_floats(C, P);
  const int grid = (int)std::min<int64_t>((L + 255) / 256, (int64_t)num_sms() * 8);
  stats_pack_kernel<<<grid, 256, 0, as_stream(stream)>>>(C, P, class_T, best_val, cm, c_num, grads, record);
  count_launch(1);
  return check_launch("stats_pack_kernel");
}

GEOT_API int geot_stats_combine(int W, int C, int64_t P, const float* gathered, float grad_scale, float* class_T,
                                float* cm, float* c_num, float* grads, geot_stream_t stream) {
  GEOT_REQUIRE(W >= 1 && C >= 1 && P >= 0, "stats_combine: bad sizes W=%d C=%d P=%lld", W, C, (long long)P);
  GEOT_REQUIRE(gathered && class_T && cm && c_num, "stats_combine: NULL pointer");
  const int64_t total = 2 * (int64_t)C * C + C + P;
  const int grid = (int)std::min<int64_t>((total + 255) / 256, (int64_t)num_sms() * 8);
  stats_combine_kernel<<<grid, 256, 0, as_stream(stream)>>>(W, C, P, gathered, grad_scale, class_T, cm, c_num, grads);
  count_launch(1);
  return check_launch("stats_combine_kernel");
}
