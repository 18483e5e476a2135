// Callers and data formats either side of the tokenizer -- HBM-bound index / gather work.
//
//   window_gather     action-chunk input side: SequenceDataset.get_item's window + pad + normalise semantics
//                     (robomimic/utils/dataset.py:499-632, obs_utils.normalize_dict :464-501) over demonstrations that are
//                     resident in HBM as one [L_total, A] array: one launch builds a [B, T, A] batch (+ pad mask) from B
//                     sampled sequence indices -- the reference does this per sample in Python/NumPy on the host.
//   bin_embed_*       AdaptiveBinActionEmbedding (robomimic/models/bin_action/backbone.py:5-89), the A/B baseline tokenizer
//                     behind the same boundary (obs_nets.py:1214-1217,1346): running min/max, torch.linspace boundaries,
//                     torch.bucketize, per-dimension nn.Embedding gather + concat (forward) and the scatter-add of the
//                     embedding gradient (backward).  Its two Linear+GELU layers run through lipvq_linear_*.
#include "common.cuh"

namespace lipvq {
namespace {

// one thread per output element [b, t, a]; rows are gathered with the reference's clamping:
//   index_in_demo = index - demo_start_seq + (pad_frame_stack ? 0 : n_frame_stack - 1)
//   window rows   = index_in_demo - (n_frame_stack - 1) ... index_in_demo + seq_length - 1, clamped to [0, L) ("pad_same")
__global__ void __launch_bounds__(256)
    window_gather_kernel(const float* __restrict__ data, const int64_t* __restrict__ demo_row0, const int32_t* __restrict__ demo_len,
                         const int64_t* __restrict__ demo_seq0, const int32_t* __restrict__ sample_demo,
                         const int64_t* __restrict__ sample_index, int64_t B, int A, int frames_back, int seq_length,
                         int index_offset, const float* __restrict__ offset, const float* __restrict__ scale,
                         float* __restrict__ out, uint8_t* __restrict__ pad_mask) {
  const int T = frames_back + seq_length;
  const int64_t total = B * (int64_t)T * A;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int a = (int)(e % A);
    const int64_t bt = e / A;
    const int t = (int)(bt % T);
    const int64_t b = bt / T;
    const int d = sample_demo[b];
    const int L = demo_len[d];
    const int64_t idx = sample_index[b] - demo_seq0[d] + index_offset;  // index_in_demo
    const int64_t want = idx - frames_back + t;
    const int64_t row = want < 0 ? 0 : (want >= L ? (int64_t)L - 1 : want);
    float v = __ldg(data + (demo_row0[d] + row) * A + a);
    if (offset != nullptr) v = (v - __ldg(offset + a)) / __ldg(scale + a);  // (x - offset) / scale, obs_utils.py:499
    out[e] = v;
    if (a == 0 && pad_mask != nullptr) pad_mask[bt] = (want >= 0 && want < L) ? 1 : 0;
  }
}

// running_min/max <- min/max(running, column-wise min/max of actions [N, A]); one CTA per action dimension
__global__ void __launch_bounds__(256)
    bin_stats_kernel(const float* __restrict__ actions, int64_t N, int A, float* __restrict__ rmin, float* __restrict__ rmax) {
  __shared__ float s_lo[8], s_hi[8];
  const int a = blockIdx.x;
  float lo = INFINITY, hi = -INFINITY;
  for (int64_t n = threadIdx.x; n < N; n += blockDim.x) {
    const float v = __ldg(actions + n * A + a);
    lo = fminf(lo, v);
    hi = fmaxf(hi, v);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  if ((threadIdx.x & 31) == 0) {
    s_lo[threadIdx.x >> 5] = lo;
    s_hi[threadIdx.x >> 5] = hi;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int i = 1; i < 8; ++i) {
      lo = fminf(lo, s_lo[i]);
      hi = fmaxf(hi, s_hi[i]);
    }
    rmin[a] = fminf(rmin[a], lo);
    rmax[a] = fmaxf(rmax[a], hi);
  }
}

// torch.linspace(lo, hi, steps) element i in fp32, with torch's two-sided rule: fma(step, i, start) below the middle,
// fma(-step, steps-1-i, end) above -- one rounding per element, as the CPU kernel the oracle runs (checked bit for bit
// against torch.linspace in tests/test_callers_cpu.py)
__device__ __forceinline__ float linspace_at(float lo, float hi, int steps, int i) {
  const float step = __fdiv_rn(__fsub_rn(hi, lo), (float)(steps - 1));
  return (i < steps / 2) ? __fmaf_rn(step, (float)i, lo) : __fmaf_rn(-step, (float)(steps - 1 - i), hi);
}

// one warp per token: lane -> (dimension, slice of the embedding row).  idx = clamp(bucketize(x, boundaries) - 1, 0, nb-1),
// bucketize(right=False) = number of boundaries strictly below x.
__global__ void __launch_bounds__(256)
    bin_embed_fwd_kernel(const float* __restrict__ actions, int64_t N, int A, int nb, int E, const float* __restrict__ rmin,
                         const float* __restrict__ rmax, const float* __restrict__ tables, float* __restrict__ out,
                         int64_t* __restrict__ idx_out) {
  const int64_t n = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (n >= N) return;
  for (int a = 0; a < A; ++a) {
    const float x = __ldg(actions + n * A + a);
    const float lo = __ldg(rmin + a), hi = __ldg(rmax + a);
    int below = 0;
    for (int i = 0; i <= nb; ++i) below += (linspace_at(lo, hi, nb + 1, i) < x) ? 1 : 0;  // nb <= 64: a scan, no search
    int idx = below - 1;
    idx = idx < 0 ? 0 : (idx > nb - 1 ? nb - 1 : idx);
    if (lane == 0 && idx_out != nullptr) idx_out[n * A + a] = idx;
    const float* row = tables + ((size_t)a * nb + idx) * E;
    float* dst = out + (n * A + a) * (int64_t)E;
    for (int e = lane; e < E; e += 32) dst[e] = __ldg(row + e);
  }
}

__global__ void __launch_bounds__(256)
    bin_embed_bwd_kernel(const float* __restrict__ dout, const int64_t* __restrict__ idx, int64_t N, int A, int nb, int E,
                         float* __restrict__ dtables) {
  const int64_t total = N * A * (int64_t)E;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int e = (int)(i % E);
    const int64_t na = i / E;
    const int a = (int)(na % A);
    const int64_t k = idx[na];
    if (k >= 0 && k < nb) atomicAdd(dtables + ((size_t)a * nb + k) * E + e, dout[i]);
  }
}

}  // namespace
}  // namespace lipvq

using namespace lipvq;

extern "C" {

int lipvq_window_gather_f32(const float* data, const int64_t* demo_row0, const int32_t* demo_len, const int64_t* demo_seq0,
                            const int32_t* sample_demo, const int64_t* sample_index, int64_t B, int A, int n_frame_stack,
                            int seq_length, int pad_frame_stack, const float* offset, const float* scale, float* out,
                            uint8_t* pad_mask, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(B >= 0 && A > 0 && n_frame_stack >= 1 && seq_length >= 1, "window_gather: bad shape");
  LIPVQ_REQUIRE((offset == nullptr) == (scale == nullptr), "window_gather: offset and scale come together");
  if (B == 0) return LIPVQ_OK;
  LIPVQ_REQUIRE(data && demo_row0 && demo_len && demo_seq0 && sample_demo && sample_index && out, "window_gather: null pointer");
  const int frames_back = n_frame_stack - 1;
  const int64_t total = B * (int64_t)(frames_back + seq_length) * A;
  int64_t blocks = ceil_div64(total, 256);
  const int64_t cap = (int64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  window_gather_kernel<<<(unsigned)blocks, 256, 0, stream>>>(data, demo_row0, demo_len, demo_seq0, sample_demo, sample_index, B, A,
                                                            frames_back, seq_length, pad_frame_stack ? 0 : frames_back, offset,
                                                            scale, out, pad_mask);
  LIPVQ_LAUNCHED("window_gather_kernel");
  return LIPVQ_OK;
}

int lipvq_bin_embed_fwd_f32(const float* actions, int64_t N, int A, int num_bins, int E, int update_stats, float* running_min,
                            float* running_max, const float* tables, float* out, int64_t* idx_out, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(N >= 0 && A > 0 && num_bins >= 1 && num_bins <= 4096 && E > 0, "bin_embed_fwd: bad shape");
  if (N == 0) return LIPVQ_OK;
  LIPVQ_REQUIRE(actions && running_min && running_max && tables && out, "bin_embed_fwd: null pointer");
  if (update_stats) {
    bin_stats_kernel<<<A, 256, 0, stream>>>(actions, N, A, running_min, running_max);
    LIPVQ_LAUNCHED("bin_stats_kernel");
  }
  bin_embed_fwd_kernel<<<(unsigned)ceil_div64(N * 32, 256), 256, 0, stream>>>(actions, N, A, num_bins, E, running_min, running_max,
                                                                             tables, out, idx_out);
  LIPVQ_LAUNCHED("bin_embed_fwd_kernel");
  return LIPVQ_OK;
}

int lipvq_bin_embed_bwd_f32(const float* dout, const int64_t* idx, int64_t N, int A, int num_bins, int E, float* dtables,
                            void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(N >= 0 && A > 0 && num_bins >= 1 && E > 0, "bin_embed_bwd: bad shape");
  if (N == 0) return LIPVQ_OK;
  LIPVQ_REQUIRE(dout && idx && dtables, "bin_embed_bwd: null pointer");
  int64_t blocks = ceil_div64(N * A * (int64_t)E, 256);
  const int64_t cap = (int64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  bin_embed_bwd_kernel<<<(unsigned)blocks, 256, 0, stream>>>(dout, idx, N, A, num_bins, E, dtables);
  LIPVQ_LAUNCHED("bin_embed_bwd_kernel");
  return LIPVQ_OK;
}

}  // extern "C"
