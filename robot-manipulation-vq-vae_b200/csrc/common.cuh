// Shared helpers for liblipvq.so (sm_100a).  Error plumbing, launch accounting, small device utilities.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "lipvq.h"

namespace lipvq {

void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);
void count_launch(int n = 1);
int sm_count();                 // of the CURRENT device
constexpr int kMaxDevices = 64;  // per-device caches (SM count, shared-memory opt-in of each kernel) are indexed by this
int current_device();

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// tcgen05 linear layers (linear_tc.cu): mode 0 = forward (bias + act_out), mode 1 = backward-x (* act_in'(Xpre)).
// B[n][k] = W[n * w_rs + k * w_cs].
// dual (forward only): C2 also receives act_out'(pre) -- one activation evaluation serves every later consumer.
bool linear_tc_supported(int mode, int64_t N, int Ka, int Nb, bool dual = false);
int linear_tc_launch(int mode, const float* A, int64_t N, int Ka, int64_t lda, int actA, const float* W, int64_t w_rs,
                     int64_t w_cs, int Nb, float* C, int64_t ldc, const float* bias, int act_out, const float* Xpre,
                     int64_t ldx, int act_in, cudaStream_t stream, float* C2 = nullptr);
// weight gradient dW += dY^T act(X), db += colsum(dY) with the reduction over tokens on the tensor cores
bool linear_tc_bwd_w_supported(int64_t N, int out_dim, int in_dim, int actX, const float* dY, const float* X);
int linear_tc_bwd_w_launch(const float* dY, int64_t N, int out_dim, const float* X, int in_dim, int actX, float* dW, float* db,
                           cudaStream_t stream);


inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }

}  // namespace lipvq

#define LIPVQ_REQUIRE(cond, ...)      \
  do {                                \
    if (!(cond)) {                    \
      lipvq::set_error(__VA_ARGS__);  \
      return LIPVQ_ERR_INVALID;       \
    }                                 \
  } while (0)

#define LIPVQ_CUDA(expr)                                             \
  do {                                                               \
    cudaError_t _e = (expr);                                         \
    if (_e != cudaSuccess) return lipvq::cuda_fail(_e, #expr);       \
  } while (0)

// after a <<<>>> launch: count it and surface launch-configuration errors
#define LIPVQ_LAUNCHED(name)                                         \
  do {                                                               \
    lipvq::count_launch();                                           \
    cudaError_t _e = cudaGetLastError();                             \
    if (_e != cudaSuccess) return lipvq::cuda_fail(_e, name);        \
  } while (0)

namespace lipvq {

// Programmatic dependent launch: a kernel launched with launch_pdl may start (prologue, barrier / TMEM setup) while the
// kernel before it in the stream is still draining; it calls grid_dep_wait() before touching anything that kernel
// wrote.  grid_dep_launch() in the earlier kernel lets the dependent grid be scheduled as soon as SMs free up.
__device__ __forceinline__ void grid_dep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void grid_dep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
bool pdl_enabled();  // LIPVQ_PDL=0 turns the attribute off (A/B runs)
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

__device__ __forceinline__ float act_apply(int act, float x) {
  switch (act) {
    case LIPVQ_ACT_RELU: return x > 0.f ? x : 0.f;
    case LIPVQ_ACT_GELU: return 0.5f * x * (1.f + erff(x * 0.70710678118654752440f));
    case LIPVQ_ACT_SIGMOID: return 1.f / (1.f + expf(-x));
    default: return x;
  }
}

// derivative of act at pre-activation x
__device__ __forceinline__ float act_grad(int act, float x) {
  switch (act) {
    case LIPVQ_ACT_RELU: return x > 0.f ? 1.f : 0.f;
    case LIPVQ_ACT_GELU: {
      // d/dx [x Phi(x)] = Phi(x) + x phi(x)
      const float cdf = 0.5f * (1.f + erff(x * 0.70710678118654752440f));
      const float pdf = 0.39894228040143267794f * expf(-0.5f * x * x);
      return cdf + x * pdf;
    }
    case LIPVQ_ACT_SIGMOID: {
      const float y = 1.f / (1.f + expf(-x));
      return y * (1.f - y);
    }
    case LIPVQ_ACT_GIVEN: return x;  // the caller stored act'(pre) itself
    default: return 1.f;
  }
}

// GELU (erf form, nn.GELU() default) and its derivative from ONE short evaluation, for the tensor-core epilogues, where the
// library erff + expf (~77 instructions per element for value and derivative) was the bound of the whole layer:
//   erf(t) = sign(t) (1 - P(k) exp(-t^2)),  k = 1 / (1 + 0.3275911 |t|),  P degree 5 (Abramowitz & Stegun 7.1.26,
//   |error| <= 1.5e-7 absolute), exp through ex2.approx.  exp(-t^2) = exp(-x^2/2) is also the Gaussian density the
//   derivative needs.  ~18 instructions.  Absolute error of value and derivative <= 2e-7 (|x| + 1), far inside the
//   1e-4 relative parity gate on losses / reconstructions (tests compare against the reference's exact erf).
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// 17 instructions, two of them on the special-function pipe (MUFU.RCP, MUFU.EX2: 1 and 2 ulp) -- no range fix-ups, no
// Newton step, no slow-path call (what __frcp_rn / __expf expand to)
__device__ __forceinline__ void gelu_value_grad_fast(float x, float& val, float& grad) {
  const float k = rcp_approx(fmaf(0.3275911f * 0.70710678118654752440f, fabsf(x), 1.f));
  float poly = fmaf(1.061405429f, k, -1.453152027f);
  poly = fmaf(poly, k, 1.421413741f);
  poly = fmaf(poly, k, -0.284496736f);
  poly = fmaf(poly, k, 0.254829592f);
  poly *= k;
  const float e = ex2_approx((x * x) * (-0.5f * 1.44269504088896340736f));  // exp(-x^2 / 2)
  const float erf_abs = fmaf(-poly, e, 1.f);
  const float cdf = fmaf(copysignf(0.5f, x), erf_abs, 0.5f);
  val = x * cdf;
  grad = fmaf(x * 0.39894228040143267794f, e, cdf);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// block-wide sum of doubles -> one atomicAdd(double) by thread 0.  blockDim.x <= 1024.
__device__ __forceinline__ void block_atomic_add_double(double v, double* dst) {
  __shared__ double s_part[32];
  v = warp_sum_d(v);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) s_part[w] = v;
  __syncthreads();
  if (w == 0) {
    const int nw = (blockDim.x + 31) >> 5;
    double t = lane < nw ? s_part[lane] : 0.0;
    t = warp_sum_d(t);
    if (lane == 0) atomicAdd(dst, t);
  }
}

}  // namespace lipvq
