// K8: MS-TCN dilated temporal convolution layers of the hierarchical tokenizer (T2), LayerNorm, normalize backward.
//
// Reference ops replaced (robomimic/algo/heirarchical_vqvae.py):
//   DilatedTemporalConvLayer.forward :39-57   out = dropout(relu(Conv1d(E,E,3,dilation=d,padding=d)(x))) + x
//   ActionEncoder/Decoder.forward    :139-162, 194-214   (transposes to [B,E,T] and back, LayerNorm(E))
//   F.normalize backward             :441-459, 652-667   (cosine quantisation / normalised commitment losses)
//
// Layout: the reference transposes [B,T,E] -> [B,E,T] for cuDNN; here activations stay token-major [B*T, E] (E
// contiguous) and the convolution is ONE contraction over k = (tap j, channel i):
//   pre[n,o] = b[o] + sum_{j,i} X[n + (j-1)d, i] * W[o,i,j]        rows whose time index leaves [0,T) contribute zero
// Taps that can only ever read padding (d >= T: j = 0 and j = 2) are not multiplied at all -- with T = 10..16 and
// dilations up to 2^9 that removes 2/3 of the MACs of most layers.  fp32 FFMA tiles (128x64x16, 8x4 per thread).
// The ReLU/dropout keep-mask is written as one byte per element and applied by the backward kernels while they load
// dOUT, so d(pre) is never materialised.
#include "common.cuh"

namespace lipvq {
namespace {

constexpr int BK = 16;
constexpr int NT = 256;
constexpr int PAD = 4;
constexpr int BN = 64;
enum { CONV_FWD = 0, CONV_BWD_X = 1, CONV_BWD_W = 2 };

struct ConvArgs {
  const float* X;          // FWD/BWD_W: layer input [N,Ei]
  const float* dOUT;       // BWD_*: gradient at the layer output [N,Eo]
  const uint8_t* mask;     // BWD_*: keep mask [N,Eo]
  const float* W;          // FWD: Wf [Eo][3][Ei]   BWD_X: Wb [3][Eo][Ei]
  const float* bias;       // FWD
  float* OUT;              // FWD: [N,Eo]   BWD_X: dX [N,Ei]   BWD_W: dWf [Eo][3][Ei] (accumulated)
  uint8_t* mask_out;       // FWD
  float* db;               // BWD_W (accumulated), may be null
  int64_t N;               // tokens
  int T, Ei, Eo, dil;
  int jlo, jhi;            // live taps [jlo, jhi)
  float scale;             // 1/(1-p)
  uint32_t drop_thr;       // drop when hash < thr (0: no dropout)
  uint32_t seed;
  int64_t k_chunk;         // BWD_W: tokens per grid.z slice (multiple of BK)
};

__device__ __forceinline__ uint32_t mix32(uint32_t a, uint32_t b) {
  uint32_t h = a * 0x9E3779B1u ^ (b + 0x7F4A7C15u + (a << 6) + (a >> 2));
  h ^= h >> 16; h *= 0x85EBCA6Bu; h ^= h >> 13; h *= 0xC2B2AE35u; h ^= h >> 16;
  return h;
}

__device__ __forceinline__ float4 ld4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }

__device__ __forceinline__ float4 masked(const float* g, const uint8_t* m, float scale) {
  float4 v = ld4(g);
  const uchar4 k = *reinterpret_cast<const uchar4*>(m);
  v.x = k.x ? v.x * scale : 0.f; v.y = k.y ? v.y * scale : 0.f;
  v.z = k.z ? v.z * scale : 0.f; v.w = k.w ? v.w * scale : 0.f;
  return v;
}

// C[m,n] = sum_k A(m,k) B(k,n); see the mode table in the file header of how (m,n,k) map to tokens/channels/taps.
template <int MODE, int BM>
__global__ void __launch_bounds__(NT, 2) conv_tile_kernel(const ConvArgs g) {
  constexpr int TM = BM / 16, TN = BN / 16;  // 8x4 or 4x4 outputs per thread
  constexpr int UA = BM * 4 / NT;            // float4 loads of A per thread and chunk
  __shared__ __align__(16) float As[BK][BM + PAD];
  __shared__ __align__(16) float Bs[BK][BN + PAD];
  __shared__ float rs_s[MODE == CONV_BWD_W ? BM : 1];

  const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
  const int64_t m0 = (int64_t)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  const int Ein = g.Ei, Eout = g.Eo, T = g.T, dil = g.dil;

  // reduction range
  int64_t kbeg, kend;
  if (MODE == CONV_FWD) { kbeg = (int64_t)g.jlo * Ein; kend = (int64_t)g.jhi * Ein; }
  else if (MODE == CONV_BWD_X) { kbeg = (int64_t)g.jlo * Eout; kend = (int64_t)g.jhi * Eout; }
  else { kbeg = (int64_t)blockIdx.z * g.k_chunk; kend = kbeg + g.k_chunk < g.N ? kbeg + g.k_chunk : g.N; }

  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;
  float4 rsum[UA];
#pragma unroll
  for (int u = 0; u < UA; ++u) rsum[u] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (MODE == CONV_BWD_W && t < BM) rs_s[t] = 0.f;

  float4 ra[UA], rb;

  auto load_a = [&](int64_t k0) {
#pragma unroll
    for (int u = 0; u < UA; ++u) {
      const int f = t + NT * u;
      float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
      if (MODE == CONV_BWD_W) {  // A(m=o, k=token): o contiguous
        const int kk = f / (BM / 4), rq = f % (BM / 4);
        const int64_t tok = k0 + kk;
        const int64_t o = m0 + 4 * rq;
        if (tok < kend && o < Eout) x = masked(g.dOUT + tok * Eout + o, g.mask + tok * Eout + o, g.scale);
      } else {                   // A(m=token, k=(j,c)): c contiguous
        const int r = f >> 2, kq = f & 3;
        const int64_t row = m0 + r, k = k0 + 4 * kq;
        if (row < g.N && k < kend) {
          const int E = (MODE == CONV_FWD) ? Ein : Eout;
          const int j = (int)(k / E), c = (int)(k - (int64_t)j * E);
          const int sh = (MODE == CONV_FWD) ? (j - 1) * dil : (1 - j) * dil;
          const int tt = (int)(row % T) + sh;
          if (tt >= 0 && tt < T) {
            const int64_t src = (row + sh) * E + c;
            x = (MODE == CONV_FWD) ? ld4(g.X + src) : masked(g.dOUT + src, g.mask + src, g.scale);
          }
        }
      }
      ra[u] = x;
    }
  };
  auto load_b = [&](int64_t k0) {
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (MODE == CONV_FWD) {         // B(k=(j,i), n=o) = Wf[o][j][i]: k contiguous
      const int r = t >> 2, kq = t & 3;
      const int o = n0 + r;
      const int64_t k = k0 + 4 * kq;
      if (o < Eout && k < kend) x = ld4(g.W + (int64_t)o * 3 * Ein + k);
    } else if (MODE == CONV_BWD_X) {  // B(k=(j,o), n=i) = Wb[j][o][i]: n contiguous
      const int kk = t / (BN / 4), rq = t % (BN / 4);
      const int64_t k = k0 + kk;
      const int i = n0 + 4 * rq;
      if (k < kend && i < Ein) x = ld4(g.W + k * Ein + i);
    } else {                          // B(k=token, n=(j,i)) = X[token + (j-1)d][i]: n contiguous
      const int kk = t / (BN / 4), rq = t % (BN / 4);
      const int64_t tok = k0 + kk;
      const int n = n0 + 4 * rq;
      if (tok < kend && n < 3 * Ein) {
        const int j = n / Ein, i = n - j * Ein;
        const int sh = (j - 1) * dil;
        const int tt = (int)(tok % T) + sh;
        if (j >= g.jlo && j < g.jhi && tt >= 0 && tt < T) x = ld4(g.X + (tok + sh) * Ein + i);
      }
    }
    rb = x;
  };
  auto store_tiles = [&]() {
#pragma unroll
    for (int u = 0; u < UA; ++u) {
      const int f = t + NT * u;
      if (MODE == CONV_BWD_W) {
        const int kk = f / (BM / 4), rq = f % (BM / 4);
        *reinterpret_cast<float4*>(&As[kk][4 * rq]) = ra[u];
        rsum[u].x += ra[u].x; rsum[u].y += ra[u].y; rsum[u].z += ra[u].z; rsum[u].w += ra[u].w;
      } else {
        const int r = f >> 2, kq = f & 3;
        As[4 * kq + 0][r] = ra[u].x; As[4 * kq + 1][r] = ra[u].y; As[4 * kq + 2][r] = ra[u].z; As[4 * kq + 3][r] = ra[u].w;
      }
    }
    if (MODE == CONV_FWD) {
      const int r = t >> 2, kq = t & 3;
      Bs[4 * kq + 0][r] = rb.x; Bs[4 * kq + 1][r] = rb.y; Bs[4 * kq + 2][r] = rb.z; Bs[4 * kq + 3][r] = rb.w;
    } else {
      const int kk = t / (BN / 4), rq = t % (BN / 4);
      *reinterpret_cast<float4*>(&Bs[kk][4 * rq]) = rb;
    }
  };

  if (kbeg < kend) { load_a(kbeg); load_b(kbeg); }
  for (int64_t k0 = kbeg; k0 < kend; k0 += BK) {
    __syncthreads();
    store_tiles();
    __syncthreads();
    if (k0 + BK < kend) { load_a(k0 + BK); load_b(k0 + BK); }
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[TM], b[TN];
      const float4 a0 = *reinterpret_cast<const float4*>(&As[kk][4 * ty]);
      a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w;
      if (TM == 8) {
        const float4 a1 = *reinterpret_cast<const float4*>(&As[kk][BM / 2 + 4 * ty]);
        a[TM - 4] = a1.x; a[TM - 3] = a1.y; a[TM - 2] = a1.z; a[TM - 1] = a1.w;
      }
      const float4 b0 = *reinterpret_cast<const float4*>(&Bs[kk][4 * tx]);
      b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w;
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
  }

  // ---- epilogue ----
  const int col = n0 + 4 * tx;
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const int64_t row = m0 + ((i < 4) ? (4 * ty + i) : (BM / 2 + 4 * ty + (i - 4)));
    if (MODE == CONV_FWD) {
      if (row >= g.N || col >= Eout) continue;
      const float4 res = ld4(g.X + row * Ein + col);  // residual (Ei == Eo)
      const float4 bb = ld4(g.bias + col);
      const float pre[4] = {acc[i][0] + bb.x, acc[i][1] + bb.y, acc[i][2] + bb.z, acc[i][3] + bb.w};
      const float rr[4] = {res.x, res.y, res.z, res.w};
      float o4[4];
      uint8_t k4[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        bool keep = pre[j] > 0.f;
        if (g.drop_thr) keep = keep && (mix32(g.seed, (uint32_t)(row * Eout + col + j)) >= g.drop_thr);
        k4[j] = keep ? 1 : 0;
        o4[j] = (keep ? pre[j] * g.scale : 0.f) + rr[j];
      }
      *reinterpret_cast<float4*>(g.OUT + row * Eout + col) = make_float4(o4[0], o4[1], o4[2], o4[3]);
      *reinterpret_cast<uchar4*>(g.mask_out + row * Eout + col) = make_uchar4(k4[0], k4[1], k4[2], k4[3]);
    } else if (MODE == CONV_BWD_X) {
      if (row >= g.N || col >= Ein) continue;
      const float4 res = ld4(g.dOUT + row * Eout + col);  // residual branch of the layer
      *reinterpret_cast<float4*>(g.OUT + row * Ein + col) =
          make_float4(acc[i][0] + res.x, acc[i][1] + res.y, acc[i][2] + res.z, acc[i][3] + res.w);
    } else {
      if (row >= Eout || col >= 3 * Ein) continue;
      float* cp = g.OUT + row * 3 * Ein + col;
#pragma unroll
      for (int j = 0; j < 4; ++j) atomicAdd(cp + j, acc[i][j]);
    }
  }
  if (MODE == CONV_BWD_W && g.db != nullptr && blockIdx.y == 0) {
    __syncthreads();
#pragma unroll
    for (int u = 0; u < UA; ++u) {
      const int rq = (t + NT * u) % (BM / 4);
      atomicAdd(&rs_s[4 * rq + 0], rsum[u].x); atomicAdd(&rs_s[4 * rq + 1], rsum[u].y);
      atomicAdd(&rs_s[4 * rq + 2], rsum[u].z); atomicAdd(&rs_s[4 * rq + 3], rsum[u].w);
    }
    __syncthreads();
    if (t < BM && m0 + t < Eout) atomicAdd(g.db + m0 + t, rs_s[t]);
  }
}

// W [Eo][Ei][3] (nn.Conv1d layout) -> Wf [Eo][3][Ei] (forward / dW layout) and Wb [3][Eo][Ei] (dX layout)
__global__ void __launch_bounds__(256) conv_pack_kernel(const float* __restrict__ W, int Eo, int Ei, float* __restrict__ Wf,
                                                        float* __restrict__ Wb) {
  const int64_t total = (int64_t)Eo * Ei * 3;
  for (int64_t it = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; it < total; it += (int64_t)gridDim.x * blockDim.x) {
    const int j = (int)(it % 3);
    const int i = (int)((it / 3) % Ei);
    const int o = (int)(it / (3 * (int64_t)Ei));
    const float v = W[it];
    if (Wf) Wf[((int64_t)o * 3 + j) * Ei + i] = v;
    if (Wb) Wb[((int64_t)j * Eo + o) * Ei + i] = v;
  }
}

// dW [Eo][Ei][3] += dWf [Eo][3][Ei]
__global__ void __launch_bounds__(256) conv_unpack_grad_kernel(const float* __restrict__ dWf, int Eo, int Ei,
                                                               float* __restrict__ dW) {
  const int64_t total = (int64_t)Eo * Ei * 3;
  for (int64_t it = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; it < total; it += (int64_t)gridDim.x * blockDim.x) {
    const int j = (int)(it % 3);
    const int i = (int)((it / 3) % Ei);
    const int o = (int)(it / (3 * (int64_t)Ei));
    dW[it] += dWf[((int64_t)o * 3 + j) * Ei + i];
  }
}

// ---- LayerNorm over the last dim (nn.LayerNorm(E), eps inside the sqrt; heirarchical_vqvae.py:137,159) --------------
// one warp per row; rows are re-read from L1/L2 for the second pass
__global__ void __launch_bounds__(256)
    layernorm_fwd_kernel(const float* __restrict__ x, int64_t rows, int E, const float* __restrict__ w,
                         const float* __restrict__ b, float eps, float* __restrict__ y, float* __restrict__ mean_out,
                         float* __restrict__ rstd_out) {
  const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const float* xr = x + row * E;
  float s = 0.f;
  for (int d = lane; d < E; d += 32) s += xr[d];
  const float mean = warp_sum(s) / (float)E;
  float v = 0.f;
  for (int d = lane; d < E; d += 32) {
    const float c = xr[d] - mean;
    v = fmaf(c, c, v);
  }
  const float rstd = rsqrtf(warp_sum(v) / (float)E + eps);
  for (int d = lane; d < E; d += 32) y[row * E + d] = (xr[d] - mean) * rstd * w[d] + b[d];
  if (lane == 0) {
    mean_out[row] = mean;
    rstd_out[row] = rstd;
  }
}

// dx = rstd * (g*w - mean_E(g*w) - xhat * mean_E(g*w*xhat)); dw += sum_rows g*xhat; db += sum_rows g
__global__ void __launch_bounds__(256)
    layernorm_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ mean,
                         const float* __restrict__ rstd, const float* __restrict__ w, int64_t rows, int E,
                         float* __restrict__ dx, float* __restrict__ dw, float* __restrict__ db) {
  extern __shared__ float s_acc[];  // [2][E]
  for (int d = threadIdx.x; d < 2 * E; d += blockDim.x) s_acc[d] = 0.f;
  __syncthreads();
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  for (int64_t row = (int64_t)blockIdx.x * wpb + wib; row < rows; row += (int64_t)gridDim.x * wpb) {
    const float* xr = x + row * E;
    const float* gr = dy + row * E;
    const float mu = mean[row], rs = rstd[row];
    float s1 = 0.f, s2 = 0.f;
    for (int d = lane; d < E; d += 32) {
      const float gw = gr[d] * w[d];
      const float xh = (xr[d] - mu) * rs;
      s1 += gw;
      s2 = fmaf(gw, xh, s2);
    }
    s1 = warp_sum(s1) / (float)E;
    s2 = warp_sum(s2) / (float)E;
    for (int d = lane; d < E; d += 32) {
      const float gv = gr[d];
      const float xh = (xr[d] - mu) * rs;
      dx[row * E + d] = rs * (gv * w[d] - s1 - xh * s2);
      atomicAdd(&s_acc[d], gv * xh);
      atomicAdd(&s_acc[E + d], gv);
    }
  }
  __syncthreads();
  for (int d = threadIdx.x; d < E; d += blockDim.x) {
    if (dw) atomicAdd(dw + d, s_acc[d]);
    if (db) atomicAdd(db + d, s_acc[E + d]);
  }
}

// backward of y = x / max(||x||, eps):  dx = (g - y (y.g)) / max(||x||, eps)   (rows with ||x|| < eps: dx = g/eps)
__global__ void __launch_bounds__(256)
    l2_normalize_rows_bwd_kernel(const float* __restrict__ x, const float* __restrict__ dy, int64_t rows, int D,
                                 float* __restrict__ dx) {
  const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const float* xr = x + row * (int64_t)D;
  const float* gr = dy + row * (int64_t)D;
  float n2 = 0.f, xg = 0.f;
  for (int d = lane; d < D; d += 32) {
    const float v = xr[d];
    n2 = fmaf(v, v, n2);
    xg = fmaf(v, gr[d], xg);
  }
  n2 = warp_sum(n2);
  xg = warp_sum(xg);
  const float nrm = sqrtf(n2);
  float* o = dx + row * (int64_t)D;
  if (nrm > 1e-12f) {
    const float inv = 1.f / nrm, c = xg / (n2 * nrm);
    for (int d = lane; d < D; d += 32) o[d] = gr[d] * inv - xr[d] * c;
  } else {
    for (int d = lane; d < D; d += 32) o[d] = gr[d] * 1e12f;
  }
}

__global__ void __launch_bounds__(256) affine_kernel(const float* __restrict__ x, float a, float b, int64_t n,
                                                     float* __restrict__ y) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    y[i] = fmaf(a, x[i], b);
}

inline int live_taps(int T, int dil, int* jlo, int* jhi) {
  if (dil >= T) { *jlo = 1; *jhi = 2; } else { *jlo = 0; *jhi = 3; }
  return *jhi - *jlo;
}

}  // namespace
}  // namespace lipvq

using namespace lipvq;

namespace lipvq {  // conv_tc.cu
bool conv_tc_supported(int64_t N, int T, int Ei, int Eo);
int conv_tc_fwd_launch(const float* X, int64_t N, int T, int E, const float* Wf, const float* bias, int dilation, int jlo, int jhi,
                       float scale, uint32_t drop_thr, uint32_t seed, float* OUT, uint8_t* mask, cudaStream_t stream);
int conv_tc_bwd_x_launch(const float* dOUT, const uint8_t* mask, int64_t N, int T, int E, const float* Wb, int dilation, int jlo,
                         int jhi, float scale, float* dX, cudaStream_t stream);
int conv_tc_bwd_w_launch(const float* dOUT, const uint8_t* mask, const float* X, int64_t N, int T, int E, int dilation, int jlo,
                         int jhi, float scale, float* dWf, float* db, cudaStream_t stream);
}  // namespace lipvq

static inline bool conv_shape_ok(int64_t N, int T, int Ei, int Eo) {
  return N >= 0 && T > 0 && Ei > 0 && Eo > 0 && (N % T) == 0 && (Ei % 4) == 0 && (Eo % 4) == 0;
}

extern "C" {

int lipvq_conv_pack_weights_f32(const float* W, int Eo, int Ei, float* Wf, float* Wb, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(W && (Wf || Wb) && Eo > 0 && Ei > 0, "conv_pack_weights: bad argument");
  const int64_t total = (int64_t)Eo * Ei * 3;
  int64_t blocks = ceil_div64(total, 256);
  if (blocks > sm_count() * 8) blocks = sm_count() * 8;
  conv_pack_kernel<<<(unsigned)blocks, 256, 0, stream>>>(W, Eo, Ei, Wf, Wb);
  LIPVQ_LAUNCHED("conv_pack_kernel");
  return LIPVQ_OK;
}

int lipvq_conv_unpack_grad_f32(const float* dWf, int Eo, int Ei, float* dW, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(dWf && dW && Eo > 0 && Ei > 0, "conv_unpack_grad: bad argument");
  const int64_t total = (int64_t)Eo * Ei * 3;
  int64_t blocks = ceil_div64(total, 256);
  if (blocks > sm_count() * 8) blocks = sm_count() * 8;
  conv_unpack_grad_kernel<<<(unsigned)blocks, 256, 0, stream>>>(dWf, Eo, Ei, dW);
  LIPVQ_LAUNCHED("conv_unpack_grad_kernel");
  return LIPVQ_OK;
}

int lipvq_dilated_conv_fwd_f32(const float* X, int64_t N, int T, int Ei, const float* Wf, const float* bias, int Eo,
                               int dilation, float drop_p, uint32_t seed, float* OUT, uint8_t* mask, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(conv_shape_ok(N, T, Ei, Eo), "dilated_conv_fwd: bad shape N=%lld T=%d Ei=%d Eo=%d (N %% T == 0, E %% 4 == 0)",
                (long long)N, T, Ei, Eo);
  if (Ei != Eo) {
    set_error("dilated_conv_fwd: in_channels != out_channels (1x1 residual conv) is not on the reference's path");
    return LIPVQ_ERR_UNSUPPORTED;
  }
  LIPVQ_REQUIRE(dilation >= 1 && drop_p >= 0.f && drop_p < 1.f, "dilated_conv_fwd: bad dilation/dropout");
  if (N == 0) return LIPVQ_OK;
  LIPVQ_REQUIRE(X && Wf && bias && OUT && mask, "dilated_conv_fwd: null pointer");
  LIPVQ_REQUIRE(aligned16(X) && aligned16(Wf) && aligned16(bias) && aligned16(OUT) && ((uintptr_t)mask & 3u) == 0,
                "dilated_conv_fwd: pointers must be 16-byte aligned");
  ConvArgs g{};
  g.X = X; g.W = Wf; g.bias = bias; g.OUT = OUT; g.mask_out = mask;
  g.N = N; g.T = T; g.Ei = Ei; g.Eo = Eo; g.dil = dilation;
  live_taps(T, dilation, &g.jlo, &g.jhi);
  g.scale = 1.f / (1.f - drop_p);
  g.drop_thr = drop_p > 0.f ? (uint32_t)((double)drop_p * 4294967296.0) : 0u;
  g.seed = seed;
  // wide layers with enough tokens: the contraction on the tensor cores (conv_tc.cu), same values and the same mask
  if (conv_tc_supported(N, T, Ei, Eo) && ((uintptr_t)mask & 7u) == 0)
    return conv_tc_fwd_launch(X, N, T, Ei, Wf, bias, dilation, g.jlo, g.jhi, g.scale, g.drop_thr, seed, OUT, mask, stream);
  dim3 grid((unsigned)ceil_div64(N, 128), (unsigned)((Eo + BN - 1) / BN));
  conv_tile_kernel<CONV_FWD, 128><<<grid, NT, 0, stream>>>(g);
  LIPVQ_LAUNCHED("conv_tile_kernel<fwd>");
  return LIPVQ_OK;
}

int lipvq_dilated_conv_bwd_x_f32(const float* dOUT, const uint8_t* mask, int64_t N, int T, int Eo, const float* Wb, int Ei,
                                 int dilation, float drop_p, float* dX, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(conv_shape_ok(N, T, Ei, Eo) && Ei == Eo && dilation >= 1, "dilated_conv_bwd_x: bad shape");
  if (N == 0) return LIPVQ_OK;
  LIPVQ_REQUIRE(dOUT && mask && Wb && dX, "dilated_conv_bwd_x: null pointer");
  LIPVQ_REQUIRE(aligned16(dOUT) && aligned16(Wb) && aligned16(dX) && ((uintptr_t)mask & 3u) == 0,
                "dilated_conv_bwd_x: pointers must be 16-byte aligned");
  ConvArgs g{};
  g.dOUT = dOUT; g.mask = mask; g.W = Wb; g.OUT = dX;
  g.N = N; g.T = T; g.Ei = Ei; g.Eo = Eo; g.dil = dilation;
  live_taps(T, dilation, &g.jlo, &g.jhi);
  g.scale = 1.f / (1.f - drop_p);
  if (conv_tc_supported(N, T, Ei, Eo) && ((uintptr_t)mask & 7u) == 0)
    return conv_tc_bwd_x_launch(dOUT, mask, N, T, Ei, Wb, dilation, g.jlo, g.jhi, g.scale, dX, stream);
  dim3 grid((unsigned)ceil_div64(N, 128), (unsigned)((Ei + BN - 1) / BN));
  conv_tile_kernel<CONV_BWD_X, 128><<<grid, NT, 0, stream>>>(g);
  LIPVQ_LAUNCHED("conv_tile_kernel<bwd_x>");
  return LIPVQ_OK;
}

int lipvq_dilated_conv_bwd_w_f32(const float* dOUT, const uint8_t* mask, const float* X, int64_t N, int T, int Eo, int Ei,
                                 int dilation, float drop_p, float* dWf, float* db, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(conv_shape_ok(N, T, Ei, Eo) && dilation >= 1, "dilated_conv_bwd_w: bad shape");
  if (N == 0) return LIPVQ_OK;
  LIPVQ_REQUIRE(dOUT && mask && X && dWf, "dilated_conv_bwd_w: null pointer");
  LIPVQ_REQUIRE(aligned16(dOUT) && aligned16(X) && ((uintptr_t)mask & 3u) == 0,
                "dilated_conv_bwd_w: pointers must be 16-byte aligned");
  ConvArgs g{};
  g.dOUT = dOUT; g.mask = mask; g.X = X; g.OUT = dWf; g.db = db;
  g.N = N; g.T = T; g.Ei = Ei; g.Eo = Eo; g.dil = dilation;
  live_taps(T, dilation, &g.jlo, &g.jhi);
  g.scale = 1.f / (1.f - drop_p);
  if (conv_tc_supported(N, T, Ei, Eo) && aligned16(dWf) && ((uintptr_t)mask & 7u) == 0)
    return conv_tc_bwd_w_launch(dOUT, mask, X, N, T, Ei, dilation, g.jlo, g.jhi, g.scale, dWf, db, stream);
  const int64_t tiles = ceil_div64(Eo, 64) * ((3 * Ei + BN - 1) / BN);
  int64_t splits = ceil_div64(2 * (int64_t)sm_count(), tiles);
  const int64_t max_splits = ceil_div64(N, 64);
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  g.k_chunk = ceil_div64(ceil_div64(N, splits), BK) * BK;
  splits = ceil_div64(N, g.k_chunk);
  LIPVQ_REQUIRE(splits <= 65535, "dilated_conv_bwd_w: too many token slices");
  dim3 grid((unsigned)ceil_div64(Eo, 64), (unsigned)((3 * Ei + BN - 1) / BN), (unsigned)splits);
  conv_tile_kernel<CONV_BWD_W, 64><<<grid, NT, 0, stream>>>(g);
  LIPVQ_LAUNCHED("conv_tile_kernel<bwd_w>");
  return LIPVQ_OK;
}

int lipvq_layernorm_fwd_f32(const float* X, int64_t rows, int E, const float* w, const float* b, float eps, float* Y,
                            float* mean, float* rstd, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(rows >= 0 && E > 0, "layernorm_fwd: bad shape");
  if (rows == 0) return LIPVQ_OK;
  LIPVQ_REQUIRE(X && w && b && Y && mean && rstd, "layernorm_fwd: null pointer");
  layernorm_fwd_kernel<<<(unsigned)ceil_div64(rows * 32, 256), 256, 0, stream>>>(X, rows, E, w, b, eps, Y, mean, rstd);
  LIPVQ_LAUNCHED("layernorm_fwd_kernel");
  return LIPVQ_OK;
}

int lipvq_layernorm_bwd_f32(const float* dY, const float* X, const float* mean, const float* rstd, const float* w,
                            int64_t rows, int E, float* dX, float* dw, float* db, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(rows >= 0 && E > 0 && E <= 4096, "layernorm_bwd: bad shape");
  if (rows == 0) return LIPVQ_OK;
  LIPVQ_REQUIRE(dY && X && mean && rstd && w && dX, "layernorm_bwd: null pointer");
  int64_t blocks = ceil_div64(rows, 8);
  if (blocks > 2 * sm_count()) blocks = 2 * sm_count();
  layernorm_bwd_kernel<<<(unsigned)blocks, 256, 2 * E * sizeof(float), stream>>>(dY, X, mean, rstd, w, rows, E, dX, dw, db);
  LIPVQ_LAUNCHED("layernorm_bwd_kernel");
  return LIPVQ_OK;
}

int lipvq_l2_normalize_rows_bwd_f32(const float* x, const float* dy, int64_t rows, int D, float* dx, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(rows >= 0 && D > 0, "l2_normalize_rows_bwd: bad shape");
  if (rows == 0) return LIPVQ_OK;
  LIPVQ_REQUIRE(x && dy && dx, "l2_normalize_rows_bwd: null pointer");
  l2_normalize_rows_bwd_kernel<<<(unsigned)ceil_div64(rows * 32, 256), 256, 0, stream>>>(x, dy, rows, D, dx);
  LIPVQ_LAUNCHED("l2_normalize_rows_bwd_kernel");
  return LIPVQ_OK;
}

int lipvq_affine_f32(const float* x, float a, float b, int64_t numel, float* y, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(numel >= 0, "affine: bad numel");
  if (numel == 0) return LIPVQ_OK;
  LIPVQ_REQUIRE(x && y, "affine: null pointer");
  int64_t blocks = ceil_div64(numel, 256);
  if (blocks > 8 * sm_count()) blocks = 8 * sm_count();
  affine_kernel<<<(unsigned)blocks, 256, 0, stream>>>(x, a, b, numel, y);
  LIPVQ_LAUNCHED("affine_kernel");
  return LIPVQ_OK;
}

}  // extern "C"
