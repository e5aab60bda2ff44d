// K4: fused output-activation + loss + metrics + gradient kernel.
//
// One pass over the last layer's pre-activations z (fp32, written by the GEMM epilogue) produces
//   * the model output a = act(z) (softmax or any elementwise activation), fp32
//   * the per-sample Keras loss (epsilon 1e-7 clipping, mean over the last axis where Keras does)
//   * up to four metrics (categorical / binary accuracy, mae, mse, mape)
//   * dL/dz in bf16, already scaled by 1/batch (what the dgrad / wgrad GEMMs consume)
// and accumulates sum(loss), sum(metric_k), count into a small fp64 statistics buffer, so a whole
// epoch (or a distributed evaluate) needs a single 64-byte device->host read.
//
// Absorbs Keras loss/metric evaluation and its backward pass (reference call sites:
// elephas/worker.py:42,108,123; elephas/spark_model.py:291; tests/integration/test_end_to_end.py:40,92).
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "../include/elephas_b200.h"
#include "host.h"
#include "math.cuh"

namespace eb {

constexpr float kEps = 1e-7f;

struct LossParams {
  const float* z; long ldz;
  const float* y; long ldy;       // targets, fp32 [M][ldy]; ldy == 0: one scalar label per row (y[m])
  float* out; long ldo;           // optional activated output
  __nv_bfloat16* dz; long lddz;   // optional gradient wrt z
  double* stats;                  // [0]=sum loss, [1..4]=sum metrics, [7]=rows
  int rows, cols;
  int loss, act;                  // EB_LOSS_*, EB_ACT_* or EB_ACT_SOFTMAX
  int metrics[4];
  float grad_scale;               // 1 / (rows of the whole batch)
};

// Reductions over the LANES (power of two) consecutive lanes that share a row.
template <int LANES>
__device__ __forceinline__ float group_sum(float v) {
#pragma unroll
  for (int o = LANES / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

template <int LANES>
__device__ __forceinline__ float group_max(float v) {
#pragma unroll
  for (int o = LANES / 2; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ float sgn(float v) { return v > 0.f ? 1.f : (v < 0.f ? -1.f : 0.f); }

// Per-element loss term and d(term)/da for elementwise losses; `inv_n` = 1/cols.
__device__ __forceinline__ void loss_elem(int loss, float a, float y, float inv_n, float& term, float& dterm) {
  switch (loss) {
    case EB_LOSS_MSE: { float d = a - y; term = d * d * inv_n; dterm = 2.f * d * inv_n; break; }
    case EB_LOSS_MAE: { float d = a - y; term = fabsf(d) * inv_n; dterm = sgn(d) * inv_n; break; }
    case EB_LOSS_MAPE: {
      float den = fmaxf(fabsf(y), kEps);
      float d = (y - a) / den;
      term = 100.f * fabsf(d) * inv_n;
      dterm = -100.f * sgn(d) / den * inv_n;
      break;
    }
    case EB_LOSS_MSLE: {
      float ac = fmaxf(a, kEps), yc = fmaxf(y, kEps);
      float d = logf(ac + 1.f) - logf(yc + 1.f);
      term = d * d * inv_n;
      dterm = a > kEps ? 2.f * d / (ac + 1.f) * inv_n : 0.f;
      break;
    }
    case EB_LOSS_LOGCOSH: {
      float d = a - y;
      float ad = fabsf(d);
      term = (ad + log1pf(expf(-2.f * ad)) - 0.6931471805599453f) * inv_n;
      dterm = tanhf(d) * inv_n;
      break;
    }
    case EB_LOSS_BCE: {
      float p = fminf(fmaxf(a, kEps), 1.f - kEps);
      term = -(y * logf(p) + (1.f - y) * logf(1.f - p)) * inv_n;
      dterm = (a > kEps && a < 1.f - kEps) ? (-y / p + (1.f - y) / (1.f - p)) * inv_n : 0.f;
      break;
    }
    default: term = 0.f; dterm = 0.f;
  }
}

// LANES lanes of a warp share one row (32 / LANES rows per warp): narrow heads (10 classes, one regression output)
// keep every lane busy instead of leaving 22 - 31 of them idle, and the reductions shrink to log2(LANES) shuffles.
template <int LANES>
__global__ void __launch_bounds__(256) loss_kernel(const LossParams p) {
  __shared__ double block_stats[8][5];
  constexpr int RPW = 32 / LANES;  // rows per warp
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int sub = lane % LANES;
  double acc[5] = {0, 0, 0, 0, 0};
  const float inv_n = 1.f / static_cast<float>(p.cols);
  for (int base = (blockIdx.x * 8 + warp) * RPW; base < p.rows; base += gridDim.x * 8 * RPW) {
    const int row = base + lane / LANES;
    const int cols = row < p.rows ? p.cols : 0;  // lanes of a missing row run the shuffles with empty loops
    const int srow = row < p.rows ? row : base;
    const float* z = p.z + static_cast<long>(srow) * p.ldz;
    const float* y = p.ldy ? p.y + static_cast<long>(srow) * p.ldy : nullptr;
    const float y_scalar = p.ldy ? 0.f : p.y[srow];
    float* out = p.out ? p.out + static_cast<long>(srow) * p.ldo : nullptr;
    __nv_bfloat16* dz = p.dz ? p.dz + static_cast<long>(srow) * p.lddz : nullptr;
    float loss_part = 0.f, m_abs = 0.f, m_sq = 0.f, m_ape = 0.f, m_bin = 0.f;
    float best_a = -INFINITY, best_y = -INFINITY;
    int arg_a = 0, arg_y = 0;

    if (p.act == EB_ACT_SOFTMAX) {
      float mx = -INFINITY;
      for (int j = sub; j < cols; j += LANES) mx = fmaxf(mx, z[j]);
      mx = group_max<LANES>(mx);
      float se = 0.f, sy = 0.f;
      for (int j = sub; j < cols; j += LANES) { se += expf(z[j] - mx); sy += y[j]; }
      se = group_sum<LANES>(se);
      sy = group_sum<LANES>(sy);
      const float inv_se = 1.f / se;
      for (int j = sub; j < cols; j += LANES) {
        const float a = expf(z[j] - mx) * inv_se, yy = y[j];
        if (out) out[j] = a;
        if (p.loss == EB_LOSS_CCE) {
          loss_part -= yy * logf(fminf(fmaxf(a, kEps), 1.f - kEps));
          if (dz) dz[j] = __float2bfloat16_rn((a * sy - yy) * p.grad_scale);
        }
        m_abs += fabsf(a - yy); m_sq += (a - yy) * (a - yy);
        m_ape += fabsf((yy - a) / fmaxf(fabsf(yy), kEps));
        if (a > best_a) { best_a = a; arg_a = j; }
        if (yy > best_y) { best_y = yy; arg_y = j; }
      }
    } else {
      for (int j = sub; j < cols; j += LANES) {
        const float a = act_apply(p.act, z[j]);
        const float yy = y ? y[j] : y_scalar;
        if (out) out[j] = a;
        float term, dterm;
        loss_elem(p.loss, a, yy, inv_n, term, dterm);
        loss_part += term;
        if (dz) dz[j] = __float2bfloat16_rn(dterm * act_grad_from_out(p.act, a) * p.grad_scale);
        m_abs += fabsf(a - yy); m_sq += (a - yy) * (a - yy);
        m_ape += fabsf((yy - a) / fmaxf(fabsf(yy), kEps));
        m_bin += ((a > 0.5f ? 1.f : 0.f) == yy) ? 1.f : 0.f;
        if (a > best_a) { best_a = a; arg_a = j; }
        if (yy > best_y) { best_y = yy; arg_y = j; }
      }
    }
    loss_part = group_sum<LANES>(loss_part);
    m_abs = group_sum<LANES>(m_abs); m_sq = group_sum<LANES>(m_sq); m_ape = group_sum<LANES>(m_ape); m_bin = group_sum<LANES>(m_bin);
    // argmax across lanes; ties resolve to the lowest index, like np.argmax
#pragma unroll
    for (int o = LANES / 2; o > 0; o >>= 1) {
      float oa = __shfl_xor_sync(0xffffffffu, best_a, o); int ia = __shfl_xor_sync(0xffffffffu, arg_a, o);
      if (oa > best_a || (oa == best_a && ia < arg_a)) { best_a = oa; arg_a = ia; }
      float oy = __shfl_xor_sync(0xffffffffu, best_y, o); int iy = __shfl_xor_sync(0xffffffffu, arg_y, o);
      if (oy > best_y || (oy == best_y && iy < arg_y)) { best_y = oy; arg_y = iy; }
    }
    if (sub == 0 && row < p.rows) {
      acc[0] += loss_part;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float v = 0.f;
        switch (p.metrics[k]) {
          case EB_METRIC_CATEGORICAL_ACC: v = arg_a == arg_y ? 1.f : 0.f; break;
          case EB_METRIC_BINARY_ACC: v = m_bin * inv_n; break;
          case EB_METRIC_MAE: v = m_abs * inv_n; break;
          case EB_METRIC_MSE: v = m_sq * inv_n; break;
          case EB_METRIC_MAPE: v = 100.f * m_ape * inv_n; break;
          default: break;
        }
        acc[1 + k] += v;
      }
    }
  }
  // fold the rows of a warp (lanes 0, LANES, 2*LANES, ...) into lane 0
#pragma unroll
  for (int k = 0; k < 5; ++k) {
#pragma unroll
    for (int o = 16; o >= LANES; o >>= 1) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], o);
  }
  if (lane == 0)
    for (int k = 0; k < 5; ++k) block_stats[warp][k] = acc[k];
  __syncthreads();
  if (p.stats && threadIdx.x < 5) {
    double s = 0;
    for (int w = 0; w < 8; ++w) s += block_stats[w][threadIdx.x];
    if (s != 0.0) atomicAdd(p.stats + threadIdx.x, s);
  }
  if (p.stats && blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(p.stats + 7, static_cast<double>(p.rows));
}

}  // namespace eb

extern "C" int eb_loss_fwd_bwd(const EbLossArgs* a, void* stream) {
  using namespace eb;
  if (a == nullptr) return fail("eb_loss_fwd_bwd: null args");
  if (a->rows <= 0 || a->cols <= 0) return 0;
  if (a->act == EB_ACT_SOFTMAX && a->ldy == 0) return fail("eb_loss_fwd_bwd: softmax needs dense targets");
  if (a->act == EB_ACT_SOFTMAX && a->dz && a->loss != EB_LOSS_CCE)
    return fail("eb_loss_fwd_bwd: softmax output is only differentiated together with categorical_crossentropy");
  if (a->act != EB_ACT_SOFTMAX && a->loss == EB_LOSS_CCE)
    return fail("eb_loss_fwd_bwd: categorical_crossentropy needs a softmax output on the native path");
  LossParams p;
  p.z = a->z; p.ldz = a->ldz; p.y = a->y; p.ldy = a->ldy; p.out = a->out; p.ldo = a->ldo;
  p.dz = static_cast<__nv_bfloat16*>(a->dz); p.lddz = a->lddz; p.stats = a->stats;
  p.rows = a->rows; p.cols = a->cols; p.loss = a->loss; p.act = a->act;
  for (int k = 0; k < 4; ++k) p.metrics[k] = a->metrics[k];
  p.grad_scale = a->grad_scale;
  // lanes per row: at most ~4 columns per lane
  const int lanes = a->cols > 64 ? 32 : a->cols > 32 ? 16 : a->cols > 16 ? 8 : a->cols > 4 ? 4 : a->cols > 1 ? 2 : 1;
  const int rows_per_block = 8 * (32 / lanes);
  int blocks = (a->rows + rows_per_block - 1) / rows_per_block;
  const int cap = num_sms() * 8;
  if (blocks > cap) blocks = cap;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (lanes) {
    case 32: loss_kernel<32><<<blocks, 256, 0, s>>>(p); break;
    case 16: loss_kernel<16><<<blocks, 256, 0, s>>>(p); break;
    case 8: loss_kernel<8><<<blocks, 256, 0, s>>>(p); break;
    case 4: loss_kernel<4><<<blocks, 256, 0, s>>>(p); break;
    case 2: loss_kernel<2><<<blocks, 256, 0, s>>>(p); break;
    default: loss_kernel<1><<<blocks, 256, 0, s>>>(p); break;
  }
  EB_LAUNCH_OK("loss_kernel");
  return 0;
}
