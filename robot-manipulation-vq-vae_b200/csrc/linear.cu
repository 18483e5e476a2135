// K1/K2/K7: fused linear layers of the encoder / Lipschitz / decoder MLPs (fp32 FFMA tiles).
//
// One templated tile kernel  C[m,n] (+)= sum_k A(m,k) B(k,n)  serves the three passes; operands are described by a
// layout (which index is contiguous in memory) and an activation applied while the tile is loaded, so a layer's
// stored pre-activation is consumed directly:
//   forward   Y  = act_out(act_in(X) W^T + b)          A = X  (k contiguous)  B = W (k contiguous)
//   backward  dX = (dY W) * act_in'(X)                  A = dY (k contiguous)  B = W (n contiguous)
//   backward  dW += dY^T act_in(X),  db += colsum(dY)   A = dY (m contiguous)  B = X (n contiguous), K = tokens, split
//
// Reference ops replaced: nn.Linear + nn.GELU chains (backbone_lfqvae_v5.py:54-68,71,75-76), nn.Linear + ReLU
// (backbone.py:17-32), LipschitzMLP.forward / normalization() (backbone_lfqvae_v5.py:6-24) and their autograd.
#include "common.cuh"

namespace lipvq {
namespace {

constexpr int BK = 16;
constexpr int NT = 256;
constexpr int PAD = 4;
enum { L_KC = 0, L_MC = 1 };                      // K-contiguous (transposing load) / MN-contiguous (direct load)
enum { MODE_FWD = 0, MODE_BWD_X = 1, MODE_BWD_W = 2 };

struct GemmArgs {
  const float* A; int64_t lda; int actA; int vecA;
  const float* B; int64_t ldb; int actB; int vecB;
  float* C; int64_t ldc; int vecC;
  int64_t M; int N; int64_t K;
  int64_t k_chunk;          // reduction range per grid.z slice (multiple of BK)
  const float* bias; int act_out;                  // MODE_FWD
  const float* Xpre; int64_t ldx; int act_in;      // MODE_BWD_X: multiply by act_in'(Xpre)
  float* rowsum;                                   // MODE_BWD_W: db (may be null)
};

template <int ROWS>
struct TileRegs { float4 v[(ROWS * 4 + NT - 1) / NT]; };

// Load a [ROWS x BK] operand tile into registers (zero outside the matrix), activation applied to valid elements.
//   L_KC: element (r,k) at src[(r0+r)*ld + k0+k]     L_MC: element (r,k) at src[(k0+k)*ld + r0+r]
template <int ROWS, int LAYOUT>
__device__ __forceinline__ void ldg_tile(const float* __restrict__ src, int64_t ld, int64_t r0, int64_t rmax, int64_t k0,
                                         int64_t kmax, int act, bool vec, TileRegs<ROWS>& t) {
  constexpr int CNT = ROWS * 4;
#pragma unroll
  for (int u = 0; u < (CNT + NT - 1) / NT; ++u) {
    const int f = threadIdx.x + NT * u;
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (f < CNT) {
      if (LAYOUT == L_KC) {
        const int r = f >> 2, kq = f & 3;
        const int64_t row = r0 + r, k = k0 + 4 * kq;
        if (row < rmax && k < kmax) {
          const float* p = src + row * ld + k;
          if (vec && k + 3 < kmax) {
            x = __ldg(reinterpret_cast<const float4*>(p));
            x.x = act_apply(act, x.x); x.y = act_apply(act, x.y); x.z = act_apply(act, x.z); x.w = act_apply(act, x.w);
          } else {
            x.x = act_apply(act, __ldg(p));
            if (k + 1 < kmax) x.y = act_apply(act, __ldg(p + 1));
            if (k + 2 < kmax) x.z = act_apply(act, __ldg(p + 2));
            if (k + 3 < kmax) x.w = act_apply(act, __ldg(p + 3));
          }
        }
      } else {
        const int k = f / (ROWS / 4), rq = f % (ROWS / 4);
        const int64_t kk = k0 + k, row = r0 + 4 * rq;
        if (kk < kmax && row < rmax) {
          const float* p = src + kk * ld + row;
          if (vec && row + 3 < rmax) {
            x = __ldg(reinterpret_cast<const float4*>(p));
            x.x = act_apply(act, x.x); x.y = act_apply(act, x.y); x.z = act_apply(act, x.z); x.w = act_apply(act, x.w);
          } else {
            x.x = act_apply(act, __ldg(p));
            if (row + 1 < rmax) x.y = act_apply(act, __ldg(p + 1));
            if (row + 2 < rmax) x.z = act_apply(act, __ldg(p + 2));
            if (row + 3 < rmax) x.w = act_apply(act, __ldg(p + 3));
          }
        }
      }
    }
    t.v[u] = x;
  }
}

// registers -> shared [k][row]
template <int ROWS, int LAYOUT>
__device__ __forceinline__ void sts_tile(float (*S)[ROWS + PAD], const TileRegs<ROWS>& t) {
  constexpr int CNT = ROWS * 4;
#pragma unroll
  for (int u = 0; u < (CNT + NT - 1) / NT; ++u) {
    const int f = threadIdx.x + NT * u;
    if (f < CNT) {
      if (LAYOUT == L_KC) {
        const int r = f >> 2, kq = f & 3;
        S[4 * kq + 0][r] = t.v[u].x;
        S[4 * kq + 1][r] = t.v[u].y;
        S[4 * kq + 2][r] = t.v[u].z;
        S[4 * kq + 3][r] = t.v[u].w;
      } else {
        const int k = f / (ROWS / 4), rq = f % (ROWS / 4);
        *reinterpret_cast<float4*>(&S[k][4 * rq]) = t.v[u];
      }
    }
  }
}

template <int BM, int BN, int LA, int LB, int MODE>
__global__ void __launch_bounds__(NT, 2) gemm_tile_kernel(const GemmArgs g) {
  constexpr int TM = BM / 16, TN = BN / 16;
  static_assert(TM == 4 || TM == 8, "BM must be 64 or 128");
  static_assert(TN == 1 || TN == 4 || TN == 8, "BN must be 16, 64 or 128");
  __shared__ __align__(16) float As[BK][BM + PAD];
  __shared__ __align__(16) float Bs[BK][BN + PAD];
  __shared__ float rs_s[MODE == MODE_BWD_W ? BM : 1];

  const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
  const int64_t m0 = (int64_t)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  const int64_t kbeg = (int64_t)blockIdx.z * g.k_chunk;
  const int64_t kend = (kbeg + g.k_chunk < g.K) ? kbeg + g.k_chunk : g.K;

  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  // MODE_BWD_W: per-thread partial of db over the A (= dY) elements this thread loads (fixed rows for every chunk)
  float4 rsum[(BM * 4 + NT - 1) / NT];
#pragma unroll
  for (int u = 0; u < (BM * 4 + NT - 1) / NT; ++u) rsum[u] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (MODE == MODE_BWD_W) {
    if (t < BM) rs_s[t] = 0.f;
  }

  TileRegs<BM> ra;
  TileRegs<BN> rb;
  if (kbeg < kend) {
    ldg_tile<BM, LA>(g.A, g.lda, m0, g.M, kbeg, kend, g.actA, g.vecA != 0, ra);
    ldg_tile<BN, LB>(g.B, g.ldb, n0, g.N, kbeg, kend, g.actB, g.vecB != 0, rb);
  }
  for (int64_t k0 = kbeg; k0 < kend; k0 += BK) {
    __syncthreads();
    sts_tile<BM, LA>(As, ra);
    sts_tile<BN, LB>(Bs, rb);
    if (MODE == MODE_BWD_W) {
#pragma unroll
      for (int u = 0; u < (BM * 4 + NT - 1) / NT; ++u) {
        rsum[u].x += ra.v[u].x; rsum[u].y += ra.v[u].y; rsum[u].z += ra.v[u].z; rsum[u].w += ra.v[u].w;
      }
    }
    __syncthreads();
    if (k0 + BK < kend) {
      ldg_tile<BM, LA>(g.A, g.lda, m0, g.M, k0 + BK, kend, g.actA, g.vecA != 0, ra);
      ldg_tile<BN, LB>(g.B, g.ldb, n0, g.N, k0 + BK, kend, g.actB, g.vecB != 0, rb);
    }
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[TM], b[TN];
      {
        const float4 a0 = *reinterpret_cast<const float4*>(&As[kk][4 * ty]);
        a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w;
        if (TM == 8) {
          const float4 a1 = *reinterpret_cast<const float4*>(&As[kk][(BM / 2) + 4 * ty]);
          a[TM - 4] = a1.x; a[TM - 3] = a1.y; a[TM - 2] = a1.z; a[TM - 1] = a1.w;
        }
      }
      if (TN == 1) {
        b[0] = Bs[kk][tx];
      } else {
        const float4 b0 = *reinterpret_cast<const float4*>(&Bs[kk][4 * tx]);
        b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w;
        if (TN == 8) {
          const float4 b1 = *reinterpret_cast<const float4*>(&Bs[kk][(BN / 2) + 4 * tx]);
          b[TN - 4] = b1.x; b[TN - 3] = b1.y; b[TN - 2] = b1.z; b[TN - 1] = b1.w;
        }
      }
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
  }

  // ---- epilogue ----
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const int64_t row = m0 + ((i < 4) ? (4 * ty + i) : ((BM / 2) + 4 * ty + (i - 4)));
    if (row >= g.M) continue;
#pragma unroll
    for (int jg = 0; jg < (TN + 3) / 4; ++jg) {  // groups of up to 4 consecutive columns
      const int colbase = n0 + ((TN == 1) ? tx : (jg == 0 ? 4 * tx : (BN / 2) + 4 * tx));
      constexpr int GW = (TN == 1) ? 1 : 4;
      float v[GW];
#pragma unroll
      for (int j = 0; j < GW; ++j) v[j] = acc[i][jg * 4 + j];
      if (MODE == MODE_FWD) {
#pragma unroll
        for (int j = 0; j < GW; ++j) {
          const int col = colbase + j;
          if (col < g.N) v[j] = act_apply(g.act_out, v[j] + (g.bias ? __ldg(g.bias + col) : 0.f));
        }
      } else if (MODE == MODE_BWD_X) {
#pragma unroll
        for (int j = 0; j < GW; ++j) {
          const int col = colbase + j;
          if (col < g.N && g.act_in != LIPVQ_ACT_NONE) v[j] *= act_grad(g.act_in, __ldg(g.Xpre + row * g.ldx + col));
        }
      }
      float* cp = g.C + row * g.ldc + colbase;
      if (MODE == MODE_BWD_W) {
#pragma unroll
        for (int j = 0; j < GW; ++j)
          if (colbase + j < g.N) atomicAdd(cp + j, v[j]);
      } else if (GW == 4 && g.vecC && colbase + 3 < g.N) {
        *reinterpret_cast<float4*>(cp) = make_float4(v[0], v[1], v[2], v[3]);
      } else {
#pragma unroll
        for (int j = 0; j < GW; ++j)
          if (colbase + j < g.N) cp[j] = v[j];
      }
    }
  }

  if (MODE == MODE_BWD_W) {
    if (g.rowsum != nullptr && blockIdx.y == 0) {  // uniform per CTA
      __syncthreads();                              // rs_s zero-init visible
      constexpr int CNT = BM * 4;
#pragma unroll
      for (int u = 0; u < (CNT + NT - 1) / NT; ++u) {
        const int f = t + NT * u;
        if (f < CNT) {
          const int rq = f % (BM / 4);  // L_MC mapping: rows 4rq..4rq+3
          atomicAdd(&rs_s[4 * rq + 0], rsum[u].x);
          atomicAdd(&rs_s[4 * rq + 1], rsum[u].y);
          atomicAdd(&rs_s[4 * rq + 2], rsum[u].z);
          atomicAdd(&rs_s[4 * rq + 3], rsum[u].w);
        }
      }
      __syncthreads();
      if (t < BM && m0 + t < g.M) atomicAdd(g.rowsum + m0 + t, rs_s[t]);
    }
  }
}

template <int BM, int BN, int LA, int LB, int MODE>
int launch_gemm(const GemmArgs& g, int splits, cudaStream_t stream) {
  dim3 grid((unsigned)ceil_div64(g.M, BM), (unsigned)((g.N + BN - 1) / BN), (unsigned)splits);
  gemm_tile_kernel<BM, BN, LA, LB, MODE><<<grid, NT, 0, stream>>>(g);
  LIPVQ_LAUNCHED("gemm_tile_kernel");
  return LIPVQ_OK;
}

template <int LA, int LB, int MODE>
int launch_gemm_bn(const GemmArgs& g, cudaStream_t stream) {  // token-major passes: BM = 128, pick BN by width
  if (g.N <= 16) return launch_gemm<128, 16, LA, LB, MODE>(g, 1, stream);
  if (g.N <= 64) return launch_gemm<128, 64, LA, LB, MODE>(g, 1, stream);
  return launch_gemm<128, 128, LA, LB, MODE>(g, 1, stream);
}

// ---- Lipschitz weight clamp -------------------------------------------------------------------------------------
__device__ __forceinline__ float softplus_f(float x) {
  // F.softplus (beta=1, threshold=20): x if x > 20 else log1p(exp(x))
  return x > 20.f ? x : log1pf(expf(x));
}

__global__ void __launch_bounds__(256)
    lipschitz_fwd_kernel(const float* __restrict__ W, const float* __restrict__ ci, int out_dim, int in_dim,
                         float* __restrict__ scale, float* __restrict__ Wn) {
  grid_dep_wait();  // programmatic dependent launch: nothing above touches memory
  grid_dep_launch();
  const int o = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (o >= out_dim) return;
  const float* w = W + (size_t)o * in_dim;
  float l1 = 0.f;
  for (int i = lane; i < in_dim; i += 32) l1 += fabsf(w[i]);
  l1 = warp_sum(l1);
  const float s = fminf(1.f, softplus_f(ci[o]) / l1);
  if (lane == 0 && scale) scale[o] = s;
  for (int i = lane; i < in_dim; i += 32) Wn[(size_t)o * in_dim + i] = w[i] * s;
}

__global__ void __launch_bounds__(256)
    lipschitz_bwd_kernel(const float* __restrict__ G, const float* __restrict__ W, const float* __restrict__ ci,
                         int out_dim, int in_dim, float* __restrict__ dW, float* __restrict__ dci) {
  grid_dep_wait();  // programmatic dependent launch: nothing above touches memory
  grid_dep_launch();
  const int o = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (o >= out_dim) return;
  const float* w = W + (size_t)o * in_dim;
  const float* gr = G + (size_t)o * in_dim;
  float l1 = 0.f, gw = 0.f;
  for (int i = lane; i < in_dim; i += 32) {
    l1 += fabsf(w[i]);
    gw = fmaf(gr[i], w[i], gw);
  }
  l1 = warp_sum(l1);
  gw = warp_sum(gw);
  const float c = ci[o];
  const float sp = softplus_f(c);
  const float ratio = sp / l1;
  // torch.minimum(1, ratio): gradient goes to ratio where ratio < 1, is split evenly on an exact tie
  const float gate = ratio < 1.f ? 1.f : (ratio == 1.f ? 0.5f : 0.f);
  const float s = fminf(1.f, ratio);
  const float dratio = gw * gate;               // dL/d ratio
  const float dl1 = -dratio * sp / (l1 * l1);   // dL/d l1
  for (int i = lane; i < in_dim; i += 32) {
    const float wv = w[i];
    const float sgn = (wv > 0.f) ? 1.f : (wv < 0.f ? -1.f : 0.f);
    dW[(size_t)o * in_dim + i] += fmaf(gr[i], s, dl1 * sgn);
  }
  if (lane == 0) {
    const float sig = 1.f / (1.f + expf(-c));   // d softplus / d c (threshold branch: 1)
    dci[o] += dratio / l1 * (c > 20.f ? 1.f : sig);
  }
}

__global__ void __launch_bounds__(256)
    act_bwd_out_kernel(const float* __restrict__ dy, const float* __restrict__ y, int act, int64_t n,
                       float* __restrict__ dx) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float yv = y[i];
    float d = 1.f;
    if (act == LIPVQ_ACT_SIGMOID) d = yv * (1.f - yv);
    else if (act == LIPVQ_ACT_RELU) d = yv > 0.f ? 1.f : 0.f;
    dx[i] = dy[i] * d;
  }
}

}  // namespace
// y = act(pre), pre_io = act'(pre)   (fallback of the dual forward on the FFMA path)
__global__ void __launch_bounds__(256)
    act_dual_kernel(float* __restrict__ pre_io, float* __restrict__ y, int act, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float pre = pre_io[i];
    y[i] = act_apply(act, pre);
    pre_io[i] = act_grad(act, pre);
  }
}

}  // namespace lipvq

using namespace lipvq;

extern "C" {

int lipvq_linear_fwd_f32(const float* X, int64_t N, int in_dim, int act_in, const float* W, const float* bias,
                         int out_dim, int act_out, float* Y, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(N >= 0 && in_dim > 0 && out_dim > 0, "linear_fwd: bad shape");
  if (N == 0) return LIPVQ_OK;
  LIPVQ_REQUIRE(X && W && Y, "linear_fwd: null pointer");
  if (linear_tc_supported(0, N, in_dim, out_dim))
    return linear_tc_launch(0, X, N, in_dim, in_dim, act_in, W, in_dim, 1, out_dim, Y, out_dim, bias, act_out, nullptr, 0,
                            LIPVQ_ACT_NONE, stream);
  GemmArgs g{};
  g.A = X; g.lda = in_dim; g.actA = act_in; g.vecA = (in_dim % 4 == 0) && aligned16(X);
  g.B = W; g.ldb = in_dim; g.actB = LIPVQ_ACT_NONE; g.vecB = (in_dim % 4 == 0) && aligned16(W);
  g.C = Y; g.ldc = out_dim; g.vecC = (out_dim % 4 == 0) && aligned16(Y);
  g.M = N; g.N = out_dim; g.K = in_dim; g.k_chunk = (in_dim + BK - 1) / BK * BK;
  g.bias = bias; g.act_out = act_out;
  return launch_gemm_bn<L_KC, L_KC, MODE_FWD>(g, stream);
}

int lipvq_linear_fwd_dual_f32(const float* X, int64_t N, int in_dim, int act_in, const float* W, const float* bias,
                              int out_dim, int act_out, float* Y, float* dYdpre, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(N >= 0 && in_dim > 0 && out_dim > 0, "linear_fwd_dual: bad shape");
  if (N == 0) return LIPVQ_OK;
  LIPVQ_REQUIRE(X && W && Y && dYdpre, "linear_fwd_dual: null pointer");
  LIPVQ_REQUIRE(act_out >= LIPVQ_ACT_NONE && act_out <= LIPVQ_ACT_SIGMOID, "linear_fwd_dual: bad activation %d", act_out);
  if (linear_tc_supported(0, N, in_dim, out_dim, true))
    return linear_tc_launch(0, X, N, in_dim, in_dim, act_in, W, in_dim, 1, out_dim, Y, out_dim, bias, act_out, nullptr, 0,
                            LIPVQ_ACT_NONE, stream, dYdpre);
  // FFMA tiles: pre-activation into dYdpre, then one elementwise pass
  const int rc = lipvq_linear_fwd_f32(X, N, in_dim, act_in, W, bias, out_dim, LIPVQ_ACT_NONE, dYdpre, stream_);
  if (rc != LIPVQ_OK) return rc;
  const int64_t numel = N * (int64_t)out_dim;
  int64_t blocks = ceil_div64(numel, 256);
  if (blocks > 8 * sm_count()) blocks = 8 * sm_count();
  act_dual_kernel<<<(unsigned)blocks, 256, 0, stream>>>(dYdpre, Y, act_out, numel);
  LIPVQ_LAUNCHED("act_dual_kernel");
  return LIPVQ_OK;
}

int lipvq_linear_bwd_x_f32(const float* dY, int64_t N, int out_dim, const float* W, int in_dim, const float* X,
                           int act_in, float* dX, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(N >= 0 && in_dim > 0 && out_dim > 0, "linear_bwd_x: bad shape");
  if (N == 0) return LIPVQ_OK;
  LIPVQ_REQUIRE(dY && W && dX && (X || act_in == LIPVQ_ACT_NONE), "linear_bwd_x: null pointer");
  if (linear_tc_supported(1, N, out_dim, in_dim))
    return linear_tc_launch(1, dY, N, out_dim, out_dim, LIPVQ_ACT_NONE, W, 1, in_dim, in_dim, dX, in_dim, nullptr,
                            LIPVQ_ACT_NONE, X, in_dim, act_in, stream);
  GemmArgs g{};
  g.A = dY; g.lda = out_dim; g.actA = LIPVQ_ACT_NONE; g.vecA = (out_dim % 4 == 0) && aligned16(dY);
  g.B = W; g.ldb = in_dim; g.actB = LIPVQ_ACT_NONE; g.vecB = (in_dim % 4 == 0) && aligned16(W);
  g.C = dX; g.ldc = in_dim; g.vecC = (in_dim % 4 == 0) && aligned16(dX);
  g.M = N; g.N = in_dim; g.K = out_dim; g.k_chunk = (out_dim + BK - 1) / BK * BK;
  g.Xpre = X; g.ldx = in_dim; g.act_in = act_in;
  return launch_gemm_bn<L_KC, L_MC, MODE_BWD_X>(g, stream);
}

int lipvq_linear_bwd_w_f32(const float* dY, int64_t N, int out_dim, const float* X, int in_dim, int act_in, float* dW,
                           float* db, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(N >= 0 && in_dim > 0 && out_dim > 0, "linear_bwd_w: bad shape");
  if (N == 0) return LIPVQ_OK;
  LIPVQ_REQUIRE(dY && X && dW, "linear_bwd_w: null pointer");
  if (linear_tc_bwd_w_supported(N, out_dim, in_dim, act_in, dY, X))
    return linear_tc_bwd_w_launch(dY, N, out_dim, X, in_dim, act_in, dW, db, stream);
  GemmArgs g{};
  g.A = dY; g.lda = out_dim; g.actA = LIPVQ_ACT_NONE; g.vecA = (out_dim % 4 == 0) && aligned16(dY);
  g.B = X; g.ldb = in_dim; g.actB = act_in; g.vecB = (in_dim % 4 == 0) && aligned16(X);
  g.C = dW; g.ldc = in_dim; g.vecC = 0;
  g.M = out_dim; g.N = in_dim; g.K = N;
  g.rowsum = db;
  // One wave of CTAs: every slice of tokens costs out*in atomics at the end, so use as few slices as fill the SMs.
  const bool narrow = in_dim <= 16;
  const int bn = narrow ? 16 : 64;
  const int64_t tiles = ceil_div64(out_dim, 64) * ((in_dim + bn - 1) / bn);
  int64_t splits = ceil_div64(sm_count(), tiles);
  const int64_t max_splits = ceil_div64(N, 64);
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  g.k_chunk = ceil_div64(ceil_div64(N, splits), BK) * BK;
  splits = ceil_div64(N, g.k_chunk);
  LIPVQ_REQUIRE(splits <= 65535, "linear_bwd_w: too many token slices");
  if (narrow) return launch_gemm<64, 16, L_MC, L_MC, MODE_BWD_W>(g, (int)splits, stream);
  return launch_gemm<64, 64, L_MC, L_MC, MODE_BWD_W>(g, (int)splits, stream);
}

int lipvq_act_bwd_from_output_f32(const float* dY, const float* Y, int act, int64_t numel, float* dX, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(numel >= 0, "act_bwd: bad numel");
  if (numel == 0) return LIPVQ_OK;
  LIPVQ_REQUIRE(dY && Y && dX, "act_bwd: null pointer");
  LIPVQ_REQUIRE(act == LIPVQ_ACT_NONE || act == LIPVQ_ACT_RELU || act == LIPVQ_ACT_SIGMOID,
                "act_bwd: activation %d has no from-output derivative", act);
  int64_t blocks = ceil_div64(numel, 256);
  const int64_t cap = (int64_t)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  act_bwd_out_kernel<<<(unsigned)blocks, 256, 0, stream>>>(dY, Y, act, numel, dX);
  LIPVQ_LAUNCHED("act_bwd_out_kernel");
  return LIPVQ_OK;
}

int lipvq_lipschitz_scale_fwd_f32(const float* W, const float* ci, int out_dim, int in_dim, float* scale,
                                  float* W_norm, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(W && ci && W_norm && out_dim > 0 && in_dim > 0, "lipschitz_scale_fwd: bad argument");
  LIPVQ_CUDA(launch_pdl(lipschitz_fwd_kernel, dim3((unsigned)ceil_div64((int64_t)out_dim * 32, 256)), dim3(256), 0, stream, W, ci, out_dim, in_dim, scale, W_norm));
  LIPVQ_LAUNCHED("lipschitz_fwd_kernel");
  return LIPVQ_OK;
}

int lipvq_lipschitz_scale_bwd_f32(const float* G, const float* W, const float* ci, int out_dim, int in_dim, float* dW,
                                  float* dci, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(G && W && ci && dW && dci && out_dim > 0 && in_dim > 0, "lipschitz_scale_bwd: bad argument");
  LIPVQ_CUDA(launch_pdl(lipschitz_bwd_kernel, dim3((unsigned)ceil_div64((int64_t)out_dim * 32, 256)), dim3(256), 0, stream, G, W, ci, out_dim, in_dim, dW, dci));
  LIPVQ_LAUNCHED("lipschitz_bwd_kernel");
  return LIPVQ_OK;
}

}  // extern "C"
