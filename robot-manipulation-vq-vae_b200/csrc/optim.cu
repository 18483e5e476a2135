// Fused AdamW over one flat fp32 buffer (parameters, gradients and both moments are contiguous slabs).
// Follows torch.optim.AdamW's single-tensor update rule; the reference trains the tokenizer with
// AdamW(lr=1e-3, weight_decay=1e-4) (robomimic/algo/icl.py:1437-1441, icl_mamba.py:884-888).
#include "common.cuh"

namespace lipvq {
namespace {

__global__ void __launch_bounds__(256)
    adamw_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                 int64_t n, int step_host, const int32_t* __restrict__ step_dev, float lr, float beta1, float beta2,
                 float eps, float weight_decay, float grad_scale) {
  grid_dep_wait();  // programmatic dependent launch: nothing above touches memory
  grid_dep_launch();
  __shared__ float s_c[3];
  if (threadIdx.x == 0) {
    const int step = step_dev ? *step_dev : step_host;
    const double bc1 = 1.0 - pow((double)beta1, (double)step);
    const double bc2 = 1.0 - pow((double)beta2, (double)step);
    s_c[0] = (float)((double)lr / bc1);          // step_size
    s_c[1] = (float)sqrt(bc2);                   // bias_correction2_sqrt
    s_c[2] = (float)(1.0 - (double)lr * (double)weight_decay);
  }
  __syncthreads();
  const float step_size = s_c[0], bc2s = s_c[1], decay = s_c[2];
  const float w1 = 1.f - beta1, w2 = 1.f - beta2;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float gi = g[i] * grad_scale;
    float pi = p[i] * decay;                       // param.mul_(1 - lr*wd)
    const float mi = fmaf(w1, gi - m[i], m[i]);    // exp_avg.lerp_(grad, 1-beta1)
    const float vi = fmaf(w2 * gi, gi, v[i] * beta2);  // exp_avg_sq.mul_(beta2).addcmul_(g, g, value=1-beta2)
    const float denom = sqrtf(vi) / bc2s + eps;
    pi = fmaf(-step_size, mi / denom, pi);         // param.addcdiv_(exp_avg, denom, value=-step_size)
    p[i] = pi;
    m[i] = mi;
    v[i] = vi;
  }
}

__global__ void counter_add_kernel(int32_t* ctr, int delta) {
  grid_dep_wait();
  grid_dep_launch();
  *ctr += delta;
}

}  // namespace
}  // namespace lipvq

using namespace lipvq;

extern "C" {

int lipvq_adamw_step_f32(float* p, const float* g, float* m, float* v, int64_t n, int step, const int32_t* step_dev,
                         float lr, float beta1, float beta2, float eps, float weight_decay, float grad_scale,
                         void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(n >= 0, "adamw: bad n");
  if (n == 0) return LIPVQ_OK;
  LIPVQ_REQUIRE(p && g && m && v, "adamw: null pointer");
  LIPVQ_REQUIRE(step_dev != nullptr || step >= 1, "adamw: step must be >= 1");
  int64_t blocks = ceil_div64(n, 256);
  const int64_t cap = (int64_t)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  LIPVQ_CUDA(launch_pdl(adamw_kernel, dim3((unsigned)blocks), dim3(256), 0, stream, p, g, m, v, n, step, step_dev, lr, beta1, beta2, eps, weight_decay, grad_scale));
  LIPVQ_LAUNCHED("adamw_kernel");
  return LIPVQ_OK;
}

int lipvq_counter_add_i32(int32_t* ctr, int delta, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(ctr, "counter_add: null pointer");
  LIPVQ_CUDA(launch_pdl(counter_add_kernel, dim3(1), dim3(1), 0, stream, ctr, delta));
  LIPVQ_LAUNCHED("counter_add_kernel");
  return LIPVQ_OK;
}

}  // extern "C"
