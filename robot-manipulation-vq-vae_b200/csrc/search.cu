// K3: fused distance + argmin (nearest code) -- fp32 SIMT contraction, top-2 merge/flag, exact rescore.
//
// Replaces LFQQuantizer.forward's [N,K,D] broadcast + torch.norm + torch.argmin
// (reference robomimic/models/vq_vae/backbone_lfqvae_v5.py:39-46; hq_vae.py:92-97,247-255),
// VQVAE.quantize's squared form (backbone.py:57-63) and HierarchicalVQVAE's matmul+argmax
// (robomimic/algo/heirarchical_vqvae.py:441-449).
//
// Pipeline (all on one stream, no host sync):
//   1. code_sqnorm_kernel   ||e_k||^2 and max_k ||e_k||^2
//   2. contraction          score[n,k] = ||e_k||^2 - 2 z_n.e_k, per-row running top-2 (never stored as [N,K])
//                           -- search_simt_kernel (here); the tcgen05 path is search_fused.cu
//   3. finalize_kernel      merge the code slices; if second-best - best > error bound the best is provably the
//                           reference's winner, else the row is queued
//   4. rescore_kernel       queued rows: exact fp32 direct-difference scan in the reference's domain
//                           (sqrt for METRIC_L2, first-min tie rule)
#include <cstdlib>

#include "search.cuh"

namespace lipvq {
namespace {

constexpr int BM = 128;  // z rows per CTA
constexpr int BN = 64;   // codes per tile
constexpr int BK = 16;   // latent-dim chunk
constexpr int PAD = 4;
constexpr int NT = 256;
constexpr int kTargetCtas = 2 * 148 * 2;  // two resident CTAs per SM, two waves

// Centre of the data cloud: mu = mean of up to 32 evenly spaced rows of z (zero for the dot metric).  Distances are
// translation invariant, but the expanded form ||e||^2 - 2 z.e loses |z||e|*eps of absolute accuracy; when z and the
// live codes sit in a tight cluster away from the origin (sigmoid latents around 0.5, an EMA codebook tracking them)
// that would push almost every row into the exact rescore.  Subtracting a nearby point from both sides first makes
// the error scale with the spread of the cloud instead.  Any mu is valid; it only affects how tight the bound is.
__global__ void __launch_bounds__(256) center_kernel(const float* __restrict__ z, int64_t N, int D, int metric,
                                                     float* __restrict__ mu, SearchHdr* hdr) {
  __shared__ float s_part[8];
  constexpr int RMAX = 32;
  const int R = (int)(N < RMAX ? N : RMAX);
  const int64_t stride = N / R;
  float nrm = 0.f;
  for (int d = threadIdx.x; d < D; d += blockDim.x) {
    float v[RMAX];
#pragma unroll
    for (int i = 0; i < RMAX; ++i)  // independent loads, all in flight together
      v[i] = (metric != LIPVQ_METRIC_DOT && i < R) ? __ldg(z + (int64_t)i * stride * D + d) : 0.f;
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < RMAX; ++i) s += v[i];
    const float m = s / (float)R;
    mu[d] = m;
    nrm = fmaf(m, m, nrm);
  }
  nrm = warp_sum(nrm);
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = nrm;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) t += s_part[i];
    hdr->mu_norm = sqrtf(t);
  }
}

__global__ void __launch_bounds__(256) code_sqnorm_kernel(const float* __restrict__ cb, int K, int D, int metric,
                                                          const float* __restrict__ mu, float* __restrict__ enorm,
                                                          SearchHdr* hdr) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= K) return;
  const float* row = cb + (size_t)warp * D;
  float s = 0.f;
  for (int d = lane; d < D; d += 32) {
    const float v = __ldg(row + d) - mu[d];
    s = fmaf(v, v, s);
  }
  s = warp_sum(s);
  if (lane == 0) {
    enorm[warp] = (metric == LIPVQ_METRIC_DOT) ? 0.f : s;
    atomicMax(&hdr->emax2_bits, __float_as_uint(s));
  }
}

// global -> registers for a [ROWS x BK] tile of a row-major matrix whose reduction dim is contiguous.
template <int ROWS>
__device__ __forceinline__ void ldg_tile(const float* __restrict__ src, const float* __restrict__ mu, int64_t r0,
                                         int64_t rmax, int k0, int kmax, int ld, bool vec,
                                         float4 (&v)[ROWS * BK / 4 / NT]) {
#pragma unroll
  for (int u = 0; u < ROWS * BK / 4 / NT; ++u) {
    const int f = threadIdx.x + NT * u;
    const int r = f >> 2, kq = f & 3;
    const int64_t row = r0 + r;
    const int k = k0 + 4 * kq;
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row < rmax && k < kmax) {
      const float* p = src + row * (int64_t)ld + k;
      if (vec) {
        x = __ldg(reinterpret_cast<const float4*>(p));
        const float4 m = *reinterpret_cast<const float4*>(mu + k);
        x.x -= m.x; x.y -= m.y; x.z -= m.z; x.w -= m.w;
      } else {
        x.x = __ldg(p) - mu[k];
        if (k + 1 < kmax) x.y = __ldg(p + 1) - mu[k + 1];
        if (k + 2 < kmax) x.z = __ldg(p + 2) - mu[k + 2];
        if (k + 3 < kmax) x.w = __ldg(p + 3) - mu[k + 3];
      }
    }
    v[u] = x;
  }
}

// registers -> smem, transposed to [k][row]
template <int ROWS>
__device__ __forceinline__ void sts_tile(float (*S)[ROWS + PAD], const float4 (&v)[ROWS * BK / 4 / NT]) {
#pragma unroll
  for (int u = 0; u < ROWS * BK / 4 / NT; ++u) {
    const int f = threadIdx.x + NT * u;
    const int r = f >> 2, kq = f & 3;
    S[4 * kq + 0][r] = v[u].x;
    S[4 * kq + 1][r] = v[u].y;
    S[4 * kq + 2][r] = v[u].z;
    S[4 * kq + 3][r] = v[u].w;
  }
}

// score[n,k] = enorm[k] - 2 z_n.e_k on fp32 FFMA tiles; per-row running (best, second, argbest).
// grid = (ceil(N/BM), ksplit).  Thread (ty,tx) owns rows {4ty+i, 64+4ty+i} x codes {4tx+j}.
__global__ void __launch_bounds__(NT, 2)
    search_simt_kernel(const float* __restrict__ z, int64_t N, int D, const float* __restrict__ cb, int K,
                       const float* __restrict__ mu, const float* __restrict__ enorm, int codes_per_split, float* __restrict__ pbest,
                       float* __restrict__ psecond, int* __restrict__ pidx, int vec) {
  __shared__ __align__(16) float As[BK][BM + PAD];
  __shared__ __align__(16) float Bs[BK][BN + PAD];
  __shared__ float en_s[BN];

  const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
  const int64_t m0 = (int64_t)blockIdx.x * BM;
  const int kbeg = blockIdx.y * codes_per_split;
  const int kend = min(K, kbeg + codes_per_split);

  float best[8], second[8];
  int bidx[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    best[i] = INFINITY;
    second[i] = INFINITY;
    bidx[i] = kbeg;
  }

  for (int kt = kbeg; kt < kend; kt += BN) {
    float acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    float4 ra[2], rb[1];
    ldg_tile<BM>(z, mu, m0, N, 0, D, D, vec != 0, ra);
    ldg_tile<BN>(cb, mu, kt, kend, 0, D, D, vec != 0, rb);

    for (int d0 = 0; d0 < D; d0 += BK) {
      __syncthreads();  // everyone is done with the previous chunk (and the previous tile's epilogue)
      sts_tile<BM>(As, ra);
      sts_tile<BN>(Bs, rb);
      if (d0 == 0 && t < BN) en_s[t] = (kt + t < kend) ? __ldg(enorm + kt + t) : INFINITY;
      __syncthreads();
      if (d0 + BK < D) {
        ldg_tile<BM>(z, mu, m0, N, d0 + BK, D, D, vec != 0, ra);
        ldg_tile<BN>(cb, mu, kt, kend, d0 + BK, D, D, vec != 0, rb);
      }
#pragma unroll
      for (int kk = 0; kk < BK; ++kk) {
        const float4 a0 = *reinterpret_cast<const float4*>(&As[kk][4 * ty]);
        const float4 a1 = *reinterpret_cast<const float4*>(&As[kk][64 + 4 * ty]);
        const float4 b0 = *reinterpret_cast<const float4*>(&Bs[kk][4 * tx]);
        const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
        const float b[4] = {b0.x, b0.y, b0.z, b0.w};
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
      }
    }

    // epilogue of this code tile: fold the 4 scores of every row into its running top-2
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int col = 4 * tx + j;
      const float en = en_s[col];  // +inf for codes past the slice end -> never selected
      const int k = kt + col;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float s = fmaf(-2.f, acc[i][j], en);
        const bool lt = s < best[i];
        second[i] = lt ? best[i] : fminf(second[i], s);
        bidx[i] = lt ? k : bidx[i];
        best[i] = lt ? s : best[i];
      }
    }
  }

  // merge the 16 threads (tx) that share a row: lanes differ only in the low 4 bits
#pragma unroll
  for (int i = 0; i < 8; ++i) {
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) {
      const float ob = __shfl_xor_sync(0xffffffffu, best[i], o);
      const float os = __shfl_xor_sync(0xffffffffu, second[i], o);
      const int oi = __shfl_xor_sync(0xffffffffu, bidx[i], o);
      const bool take = (ob < best[i]) || (ob == best[i] && oi < bidx[i]);
      const float loser = take ? best[i] : ob;
      second[i] = fminf(fminf(second[i], os), loser);
      best[i] = take ? ob : best[i];
      bidx[i] = take ? oi : bidx[i];
    }
  }
  if (tx == 0) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int64_t row = m0 + (i < 4 ? 4 * ty + i : 64 + 4 * ty + (i - 4));
      if (row < N) {
        const int64_t o = (int64_t)blockIdx.y * N + row;
        pbest[o] = best[i];
        psecond[o] = second[i];
        pidx[o] = bidx[i];
      }
    }
  }
}

// One warp per row: merge code slices, decide whether the approximate winner is provably exact.
__global__ void __launch_bounds__(256)
    finalize_kernel(const float* __restrict__ z, int64_t N, int D, const float* __restrict__ cb,
                    const float* __restrict__ mu, int metric, int ksplit, const float* __restrict__ pbest, const float* __restrict__ psecond,
                    const int* __restrict__ pidx, float gamma_en, float gamma_dot, float gamma_ref, SearchHdr* hdr,
                    int* __restrict__ work, int64_t* __restrict__ idx_out, float* __restrict__ best_out) {
  const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= N) return;

  float b = INFINITY, s2 = INFINITY;
  int bi = 0;
  for (int s = 0; s < ksplit; ++s) {  // uniform across the warp
    const float ob = pbest[(int64_t)s * N + row], os = psecond[(int64_t)s * N + row];
    const int oi = pidx[(int64_t)s * N + row];
    const bool take = (ob < b) || (ob == b && oi < bi) || s == 0;
    const float loser = take ? b : ob;
    s2 = fminf(fminf(s2, os), loser);
    b = take ? ob : b;
    bi = take ? oi : bi;
  }

  // ||z||^2 and the exact value of the candidate winner, lanes across the latent dim
  const float* zr = z + row * (int64_t)D;
  const float* er = cb + (int64_t)bi * D;
  float zn2 = 0.f, ex = 0.f, zo2 = 0.f;  // zn2: centred ||z - mu||^2 (what the contraction saw); zo2: ||z||^2
  for (int d = lane; d < D; d += 32) {
    const float zv = __ldg(zr + d), ev = __ldg(er + d);
    const float zc = zv - mu[d];
    zn2 = fmaf(zc, zc, zn2);
    zo2 = fmaf(zv, zv, zo2);
    if (metric == LIPVQ_METRIC_DOT) {
      ex = fmaf(zv, ev, ex);
    } else {
      const float df = zv - ev;
      ex = fmaf(df, df, ex);
    }
  }
  zn2 = warp_sum(zn2);
  zo2 = warp_sum(zo2);
  ex = warp_sum(ex);
  if (lane != 0) return;

  const float emax2 = __uint_as_float(hdr->emax2_bits);
  const float zn = sqrtf(zn2);
  float emax = sqrtf(emax2);
  // Only codes that can still beat the candidate matter, and by the triangle inequality those satisfy
  // ||e_k|| <= ||z|| + d_k <= ||z|| + d_best.  This keeps the bound tight when some codes are huge (an EMA codebook
  // divides by ~eps for codes that were never hit, hq_vae.py:277-279), which a global max-norm bound would not.
  if (metric != LIPVQ_METRIC_DOT) emax = fminf(emax, zn + 1.02f * sqrtf(fmaxf(b + zn2, 0.f)) + 1e-30f);
  // both competitors' scores can be off by gamma_en*||e||^2 (the separately rounded norm) + gamma_dot*2||z|| ||e|| each
  const float margin_alg = 2.f * (gamma_en * emax * emax + gamma_dot * 2.f * zn * emax);
  // the reference's own fp32 rounding of d^2 (or of the similarity): stay out of its noise too
  const float ref_mag = (metric == LIPVQ_METRIC_DOT) ? 2.f * zn * emax : fmaxf(b + zn2, 0.f);
  // rounding of the subtraction of mu itself: each vector moves by at most eps*(|x| + |mu|), which perturbs d^2 by
  // 2 d delta + delta^2 per competitor
  const float dhat = sqrtf(fmaxf(b + zn2, 0.f));
  const float delta = 5.9604645e-8f * (sqrtf(zo2) + emax + 3.f * hdr->mu_norm);
  const float margin_c = (metric == LIPVQ_METRIC_DOT) ? 0.f : 2.f * (2.f * dhat * delta + delta * delta);
  const float margin = fmaxf(margin_alg + margin_c, gamma_ref * ref_mag);
  if (s2 - b > margin) {
    idx_out[row] = (int64_t)bi;
    if (best_out) best_out[row] = (metric == LIPVQ_METRIC_L2) ? sqrtf(ex) : ex;
  } else {
    const int pos = atomicAdd(&hdr->flagged, 1);
    work[pos] = (int)row;
  }
}

// Exact scan for queued rows: one CTA (8 warps) per batch of R rows, grid-stride over the queue.  A group of LPC lanes
// owns one code at a time and walks its row with coalesced float4 loads; each code's fp32 direct-difference sum (the
// reference's expression, backbone_lfqvae_v5.py:41-45) is reduced over the group in a fixed butterfly order, so the
// result is deterministic and does not depend on UN or R.  Winner = smallest value, lowest index on exact ties
// (torch.argmin's first-min rule).  Every row needs the WHOLE codebook: with one row per CTA the scan was bound by L2
// bandwidth (3000 queued rows x 128 KB = 393 MB -> 100 us); R rows share each code load.
#ifndef LIPVQ_RS_UN
#define LIPVQ_RS_UN 2
#endif
constexpr int kRescoreWideR = 16;  // rows per batch of the long-queue form (LIPVQ_RS_UN codes in flight per lane group)

// long-queue form at D = 128 NJ with a whole warp per code: every chunk of the UN code rows in flight before the first use
// (the generic loop exposes one L2 round trip per 128 elements); same sums in the same order as the loop it replaces
template <int UN, int R, int NJ>
__device__ __forceinline__ void rescore_unrolled(float (&a)[UN][R], const float* zs, const float* __restrict__ cb, int K, int k0,
                                                 int lane, int metric) {
  constexpr int D = 128 * NJ;
  float4 e4[UN][NJ];
#pragma unroll
  for (int u = 0; u < UN; ++u) {
    const int k = k0 + u * 8;
#pragma unroll
    for (int j = 0; j < NJ; ++j)
      e4[u][j] = (k < K) ? __ldg(reinterpret_cast<const float4*>(cb + (int64_t)k * D + 4 * lane + 128 * j)) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
#pragma unroll
  for (int j = 0; j < NJ; ++j) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const float4 z4 = *reinterpret_cast<const float4*>(zs + r * D + 4 * lane + 128 * j);
#pragma unroll
      for (int u = 0; u < UN; ++u) {
        if (metric == LIPVQ_METRIC_DOT) {
          a[u][r] = fmaf(z4.x, e4[u][j].x, a[u][r]); a[u][r] = fmaf(z4.y, e4[u][j].y, a[u][r]);
          a[u][r] = fmaf(z4.z, e4[u][j].z, a[u][r]); a[u][r] = fmaf(z4.w, e4[u][j].w, a[u][r]);
        } else {
          float df;
          df = z4.x - e4[u][j].x; a[u][r] = fmaf(df, df, a[u][r]);
          df = z4.y - e4[u][j].y; a[u][r] = fmaf(df, df, a[u][r]);
          df = z4.z - e4[u][j].z; a[u][r] = fmaf(df, df, a[u][r]);
          df = z4.w - e4[u][j].w; a[u][r] = fmaf(df, df, a[u][r]);
        }
      }
    }
  }
}

template <int UN, int R>
__device__ __forceinline__ void rescore_body(float* zs, float (*s_bv)[8], int (*s_bk)[8], const float* __restrict__ z, int D,
                                             const float* __restrict__ cb, int K, int metric, int nflag,
                                             const int* __restrict__ work, int64_t* __restrict__ idx_out,
                                             float* __restrict__ best_out, int vec, int lpc) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int sub = lane & (lpc - 1), grp = lane / lpc, groups = 32 / lpc;
  for (int w0 = blockIdx.x * R; w0 < nflag; w0 += gridDim.x * R) {
    const int nr = (nflag - w0 < R) ? nflag - w0 : R;
    __syncthreads();  // previous batch fully consumed
    for (int i = threadIdx.x; i < R * D; i += blockDim.x) {
      const int r = i / D, d = i - r * D;
      zs[i] = (r < nr) ? __ldg(z + (int64_t)work[w0 + r] * D + d) : 0.f;
    }
    __syncthreads();
    float bv[R];
    int bk[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      bv[r] = INFINITY;
      bk[r] = 0x7fffffff;
    }
    // UN codes per group and iteration, independent accumulators: the loads of all of them are in flight together
    for (int k0 = warp * groups; k0 < K; k0 += 8 * groups * UN) {
      float a[UN][R];
#pragma unroll
      for (int u = 0; u < UN; ++u)
#pragma unroll
        for (int r = 0; r < R; ++r) a[u][r] = 0.f;
      if (R > 1 && vec && lpc == 32 && D == 512) {
        rescore_unrolled<UN, R, 4>(a, zs, cb, K, k0, lane, metric);
      } else if (R > 1 && vec && lpc == 32 && D == 256) {
        rescore_unrolled<UN, R, 2>(a, zs, cb, K, k0, lane, metric);
      } else if (vec) {
        for (int d = 4 * sub; d < D; d += 4 * lpc) {
          float4 e4[UN];
#pragma unroll
          for (int u = 0; u < UN; ++u) {
            const int k = k0 + grp + u * 8 * groups;
            e4[u] = (k < K) ? __ldg(reinterpret_cast<const float4*>(cb + (int64_t)k * D + d)) : make_float4(0.f, 0.f, 0.f, 0.f);
          }
#pragma unroll
          for (int r = 0; r < R; ++r) {
            const float4 z4 = *reinterpret_cast<const float4*>(zs + r * D + d);
#pragma unroll
            for (int u = 0; u < UN; ++u) {
              if (metric == LIPVQ_METRIC_DOT) {
                a[u][r] = fmaf(z4.x, e4[u].x, a[u][r]); a[u][r] = fmaf(z4.y, e4[u].y, a[u][r]);
                a[u][r] = fmaf(z4.z, e4[u].z, a[u][r]); a[u][r] = fmaf(z4.w, e4[u].w, a[u][r]);
              } else {
                float df;
                df = z4.x - e4[u].x; a[u][r] = fmaf(df, df, a[u][r]);
                df = z4.y - e4[u].y; a[u][r] = fmaf(df, df, a[u][r]);
                df = z4.z - e4[u].z; a[u][r] = fmaf(df, df, a[u][r]);
                df = z4.w - e4[u].w; a[u][r] = fmaf(df, df, a[u][r]);
              }
            }
          }
        }
      } else {
        for (int d = sub; d < D; d += lpc) {
          float ev[UN];
#pragma unroll
          for (int u = 0; u < UN; ++u) {
            const int k = k0 + grp + u * 8 * groups;
            ev[u] = (k < K) ? __ldg(cb + (int64_t)k * D + d) : 0.f;
          }
#pragma unroll
          for (int r = 0; r < R; ++r) {
            const float zv = zs[r * D + d];
#pragma unroll
            for (int u = 0; u < UN; ++u) {
              if (metric == LIPVQ_METRIC_DOT) {
                a[u][r] = fmaf(zv, ev[u], a[u][r]);
              } else {
                const float df = zv - ev[u];
                a[u][r] = fmaf(df, df, a[u][r]);
              }
            }
          }
        }
      }
#pragma unroll
      for (int u = 0; u < UN; ++u) {  // ascending k within the group, so a strict < keeps the lowest index on ties
        const int k = k0 + grp + u * 8 * groups;
#pragma unroll
        for (int r = 0; r < R; ++r) {
          float av = a[u][r];
          for (int o = lpc >> 1; o > 0; o >>= 1) av += __shfl_xor_sync(0xffffffffu, av, o);
          const float val = (metric == LIPVQ_METRIC_L2) ? sqrtf(av) : (metric == LIPVQ_METRIC_DOT ? -av : av);
          if (k < K && val < bv[r]) {
            bv[r] = val;
            bk[r] = k;
          }
        }
      }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, bv[r], o);
        const int ok = __shfl_xor_sync(0xffffffffu, bk[r], o);
        if (ov < bv[r] || (ov == bv[r] && ok < bk[r])) {
          bv[r] = ov;
          bk[r] = ok;
        }
      }
      if (lane == 0) {
        s_bv[r][warp] = bv[r];
        s_bk[r][warp] = bk[r];
      }
    }
    __syncthreads();
    if (threadIdx.x < nr) {
      const int r = threadIdx.x;
      float v = s_bv[r][0];
      int k = s_bk[r][0];
      for (int i = 1; i < 8; ++i)
        if (s_bv[r][i] < v || (s_bv[r][i] == v && s_bk[r][i] < k)) {
          v = s_bv[r][i];
          k = s_bk[r][i];
        }
      if (k == 0x7fffffff) k = 0;  // all-NaN row: torch would return the NaN's index; we return 0
      const int64_t row = work[w0 + r];
      idx_out[row] = (int64_t)k;
      if (best_out) best_out[row] = (metric == LIPVQ_METRIC_DOT) ? -v : v;
    }
  }
}

// Long-queue form proper (D = 128 NJ, a whole warp per code): per iteration a warp holds 2 codes x 16 rows = 32 partial sums
// per lane.  Reducing each over the warp costs five shuffles and five adds per sum -- more issue slots than the FMAs.  A
// TRANSPOSED butterfly reduces all 32 at once with 31 shuffles: at the level that pairs lanes (l, l ^ 16) every lane keeps
// half of its values and sends the other half, and so on; lane l ends with the complete sum number l (code l >> 4, row
// l & 15).  Every level adds the same two operands as the plain butterfly (16, 8, 4, 2, 1), so the sums are bit-identical.
template <int NJ>
__device__ __forceinline__ void rescore_wide32(float* zs, float (*s_bv)[8], int (*s_bk)[8], const float* __restrict__ z,
                                               const float* __restrict__ cb, int K, int metric, int nflag,
                                               const int* __restrict__ work, int64_t* __restrict__ idx_out,
                                               float* __restrict__ best_out) {
  constexpr int R = kRescoreWideR, D = 128 * NJ;
  static_assert(R == 16, "the transposed butterfly is written for 2 codes x 16 rows");
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4, b1 = lane & 2, b0 = lane & 1;
  for (int w0 = blockIdx.x * R; w0 < nflag; w0 += gridDim.x * R) {
    const int nr = (nflag - w0 < R) ? nflag - w0 : R;
    __syncthreads();  // previous batch fully consumed
    for (int i = threadIdx.x; i < R * (D / 4); i += blockDim.x) {
      const int r = i / (D / 4), d4 = i - r * (D / 4);
      reinterpret_cast<float4*>(zs)[i] = (r < nr) ? __ldg(reinterpret_cast<const float4*>(z + (int64_t)work[w0 + r] * D) + d4)
                                                  : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    __syncthreads();
    float bv = INFINITY;
    int bk = 0x7fffffff;
    for (int k0 = warp; k0 < K; k0 += 16) {
      float a[2][R];
#pragma unroll
      for (int u = 0; u < 2; ++u)
#pragma unroll
        for (int r = 0; r < R; ++r) a[u][r] = 0.f;
      rescore_unrolled<2, R, NJ>(a, zs, cb, K, k0, lane, metric);
      float w16[16], w8[8], w4[4], w2[2];
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const float keep = b4 ? a[1][i] : a[0][i], send = b4 ? a[0][i] : a[1][i];
        w16[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float keep = b3 ? w16[i + 8] : w16[i], send = b3 ? w16[i] : w16[i + 8];
        w8[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float keep = b2 ? w8[i + 4] : w8[i], send = b2 ? w8[i] : w8[i + 4];
        w4[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
      }
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const float keep = b1 ? w4[i + 2] : w4[i], send = b1 ? w4[i] : w4[i + 2];
        w2[i] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
      }
      const float keep = b0 ? w2[1] : w2[0], send = b0 ? w2[0] : w2[1];
      const float av = keep + __shfl_xor_sync(0xffffffffu, send, 1);
      const int k = k0 + (lane >> 4) * 8;   // ascending per lane: a strict < keeps the lowest index on ties
      const float val = (metric == LIPVQ_METRIC_L2) ? sqrtf(av) : (metric == LIPVQ_METRIC_DOT ? -av : av);
      if (k < K && val < bv) {
        bv = val;
        bk = k;
      }
    }
    {  // lanes l and l ^ 16 hold the same row, disjoint code sets
      const float ov = __shfl_xor_sync(0xffffffffu, bv, 16);
      const int ok = __shfl_xor_sync(0xffffffffu, bk, 16);
      if (ov < bv || (ov == bv && ok < bk)) {
        bv = ov;
        bk = ok;
      }
    }
    if (lane < 16) {
      s_bv[lane][warp] = bv;
      s_bk[lane][warp] = bk;
    }
    __syncthreads();
    if (threadIdx.x < nr) {
      const int r = threadIdx.x;
      float v = s_bv[r][0];
      int k = s_bk[r][0];
      for (int i = 1; i < 8; ++i)
        if (s_bv[r][i] < v || (s_bv[r][i] == v && s_bk[r][i] < k)) {
          v = s_bv[r][i];
          k = s_bk[r][i];
        }
      if (k == 0x7fffffff) k = 0;
      const int64_t row = work[w0 + r];
      idx_out[row] = (int64_t)k;
      if (best_out) best_out[row] = (metric == LIPVQ_METRIC_DOT) ? -v : v;
    }
  }
}

// A long queue (a collapsed codebook: every row a near-tie, e.g. the cosine quantisers of a HierarchicalVQVAE a few steps
// into training -- 65 536 rows x 1024 codes x 512 took 12 ms with one row per batch, every row re-reading the 2 MB codebook
// from L2) is scanned sixteen rows per batch: the same sums in the same order (see above), a sixteenth of the L2 traffic.
// The queue length is only known on the device, so both forms live in one kernel; wide = 0 (host: the rows of a batch do
// not fit the dynamic shared memory) keeps the one-row form.
template <int UN, int R>
__global__ void __launch_bounds__(256)
    rescore_kernel(const float* __restrict__ z, int D, const float* __restrict__ cb, int K, int metric,
                   const int* __restrict__ count, const int* __restrict__ first_count, const int* __restrict__ work,
                   const int* __restrict__ work_first, int skip_rows, int64_t* __restrict__ idx_out, float* __restrict__ best_out, int32_t* __restrict__ stats_out, int vec,
                   int lpc, int wide) {
  extern __shared__ __align__(16) float zs[];  // [R][D] ([kRescoreWideR][D] with wide)
  __shared__ float s_bv[kRescoreWideR][8];
  __shared__ int s_bk[kRescoreWideR][8];
  grid_dep_wait();  // launched as a programmatic dependent of the last search pass
  int nflag = *count;
  if (work_first != nullptr && *first_count <= skip_rows) {  // the 3-term pass stepped aside: its input queue is ours
    nflag = *first_count;
    work = work_first;
  }
  if (stats_out && blockIdx.x == 0 && threadIdx.x == 0) {
    stats_out[0] = nflag;         // rows decided by the exact scan
    stats_out[1] = *first_count;  // rows the first (fastest) level could not certify
  }
  if (wide && nflag >= kRescoreWideR * (int)gridDim.x / 4) {
    if (vec && D == 512) rescore_wide32<4>(zs, s_bv, s_bk, z, cb, K, metric, nflag, work, idx_out, best_out);
    else if (vec && D == 256) rescore_wide32<2>(zs, s_bv, s_bk, z, cb, K, metric, nflag, work, idx_out, best_out);
    else rescore_body<LIPVQ_RS_UN, kRescoreWideR>(zs, s_bv, s_bk, z, D, cb, K, metric, nflag, work, idx_out, best_out, vec, lpc);
  } else
    rescore_body<UN, R>(zs, s_bv, s_bk, z, D, cb, K, metric, nflag, work, idx_out, best_out, vec, lpc);
}

// Exact scan, small codebooks (D % 4 == 0, D <= 256, K * (D + 4) * 4 bytes <= 200 KB): the fp32 codebook is staged in
// shared memory ONCE per CTA and every queued row of the CTA scans it there, one LANE per code.
//   * With the codebook in global memory every row re-read all of it from L2, and because every SM reads the SAME
//     K * D * 4 bytes the requests pile up on the few L2 slices that own those lines: 3000 queued rows took 137 us
//     (~2.9 TB/s) however the scan was arranged (group butterfly, sixteen codes in flight, eight rows per batch).
//   * Rows are padded to D + 4 floats: lane k reads 16 bytes at k * (D + 4) * 4, an odd multiple of 16 bytes apart, so
//     the eight lanes of a quarter-warp request hit distinct banks.
//   * No cross-lane reduction per code; each code's sum runs over four interleaved partial sums in a fixed order, so the
//     result is deterministic.  Winner = smallest value, lowest index on exact ties (torch.argmin's first-min rule).
__global__ void __launch_bounds__(256, 1)
    rescore_smem_kernel(const float* __restrict__ z, int D, const float* __restrict__ cb, int K, int metric,
                        const int* __restrict__ count, const int* __restrict__ first_count, const int* __restrict__ work,
                        const int* __restrict__ work_first, int skip_rows, int64_t* __restrict__ idx_out,
                        float* __restrict__ best_out, int32_t* __restrict__ stats_out) {
  extern __shared__ __align__(16) float cs[];  // [K][D + 4] codebook, then [D] the row
  __shared__ float s_bv[8];
  __shared__ int s_bk[8];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nv = D >> 2, ldc = D + 4;
  float* zs = cs + (size_t)K * ldc;
  grid_dep_wait();
  int nflag = *count;
  if (work_first != nullptr && *first_count <= skip_rows) {
    nflag = *first_count;
    work = work_first;
  }
  if (stats_out && blockIdx.x == 0 && threadIdx.x == 0) {
    stats_out[0] = nflag;
    stats_out[1] = *first_count;
  }
  if ((int)blockIdx.x >= nflag) return;  // the common case: nothing queued, nothing staged
  {  // every CTA stages the same lines: each starts at its own offset so the L2 slices are not hit in lock step
    const int total = K * nv;
    const int start = (int)(((int64_t)total * blockIdx.x) / gridDim.x) & ~255;
    for (int i0 = threadIdx.x; i0 < total; i0 += 8 * 256) {  // eight independent loads in flight per thread
      float4 t[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        int i = i0 + u * 256 + start;
        if (i >= total) i -= total;
        t[u] = (i0 + u * 256 < total) ? __ldg(reinterpret_cast<const float4*>(cb) + i) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        int i = i0 + u * 256 + start;
        if (i >= total) i -= total;
        const int k = i / nv, v = i - k * nv;
        if (i0 + u * 256 < total) *reinterpret_cast<float4*>(cs + (size_t)k * ldc + 4 * v) = t[u];
      }
    }
  }
  // the row of the NEXT trip is requested before the current one is scanned (index, then row: two dependent loads)
  float znext = ((int)threadIdx.x < D) ? __ldg(z + (int64_t)work[blockIdx.x] * D + threadIdx.x) : 0.f;
  for (int w = blockIdx.x; w < nflag; w += gridDim.x) {
    const int64_t row = work[w];
    __syncthreads();  // codebook staged / previous row consumed
    if ((int)threadIdx.x < D) zs[threadIdx.x] = znext;
    __syncthreads();
    if (w + (int)gridDim.x < nflag && (int)threadIdx.x < D)
      znext = __ldg(z + (int64_t)work[w + gridDim.x] * D + threadIdx.x);
    float bv = INFINITY;
    int bk = 0x7fffffff;
    // two codes per trip (k and k + 256) share every load of the row; ascending k per thread, so a strict < keeps the
    // lowest index on ties
    for (int k = threadIdx.x; k < K; k += 512) {
      const bool two = k + 256 < K;
      const float4* e0 = reinterpret_cast<const float4*>(cs + (size_t)k * ldc);
      const float4* e1 = reinterpret_cast<const float4*>(cs + (size_t)(two ? k + 256 : k) * ldc);
      float a[4] = {0.f, 0.f, 0.f, 0.f}, b[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
      for (int v = 0; v < nv; ++v) {
        const float4 x = e0[v], y = e1[v];
        const float4 z4 = *reinterpret_cast<const float4*>(zs + 4 * v);
        float& accA = a[v & 3];
        float& accB = b[v & 3];
        if (metric == LIPVQ_METRIC_DOT) {
          accA = fmaf(z4.x, x.x, accA); accA = fmaf(z4.y, x.y, accA); accA = fmaf(z4.z, x.z, accA); accA = fmaf(z4.w, x.w, accA);
          accB = fmaf(z4.x, y.x, accB); accB = fmaf(z4.y, y.y, accB); accB = fmaf(z4.z, y.z, accB); accB = fmaf(z4.w, y.w, accB);
        } else {
          float df;
          df = z4.x - x.x; accA = fmaf(df, df, accA);
          df = z4.y - x.y; accA = fmaf(df, df, accA);
          df = z4.z - x.z; accA = fmaf(df, df, accA);
          df = z4.w - x.w; accA = fmaf(df, df, accA);
          df = z4.x - y.x; accB = fmaf(df, df, accB);
          df = z4.y - y.y; accB = fmaf(df, df, accB);
          df = z4.z - y.z; accB = fmaf(df, df, accB);
          df = z4.w - y.w; accB = fmaf(df, df, accB);
        }
      }
      const float av = (a[0] + a[1]) + (a[2] + a[3]), bvv = (b[0] + b[1]) + (b[2] + b[3]);
      const float val = (metric == LIPVQ_METRIC_L2) ? sqrtf(av) : (metric == LIPVQ_METRIC_DOT ? -av : av);
      const float val2 = (metric == LIPVQ_METRIC_L2) ? sqrtf(bvv) : (metric == LIPVQ_METRIC_DOT ? -bvv : bvv);
      if (val < bv) {
        bv = val;
        bk = k;
      }
      if (two && val2 < bv) {
        bv = val2;
        bk = k + 256;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const int ok = __shfl_xor_sync(0xffffffffu, bk, o);
      if (ov < bv || (ov == bv && ok < bk)) {
        bv = ov;
        bk = ok;
      }
    }
    if (lane == 0) {
      s_bv[warp] = bv;
      s_bk[warp] = bk;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int i = 1; i < 8; ++i)
        if (s_bv[i] < bv || (s_bv[i] == bv && s_bk[i] < bk)) {
          bv = s_bv[i];
          bk = s_bk[i];
        }
      if (bk == 0x7fffffff) bk = 0;  // all-NaN row: torch would return the NaN's index; we return 0
      idx_out[row] = (int64_t)bk;
      if (best_out) best_out[row] = (metric == LIPVQ_METRIC_DOT) ? -bv : bv;
    }
  }
}

// Exact scan, the tokenizer's own shape class (K <= 512, D <= 64, D % 4 == 0): thread k keeps code k in REGISTERS for
// the whole kernel and every queued row of the CTA is scanned against them -- per row one broadcast read of z from
// shared memory and ~130 arithmetic instructions per thread (the shared-memory variant above re-reads 128 KB of codes
// per row and runs at ~2 us per row and CTA).  Same fixed summation order as rescore_smem_kernel.
__global__ void __launch_bounds__(512, 1)
    rescore_reg_kernel(const float* __restrict__ z, int D, const float* __restrict__ cb, int K, int metric,
                       const int* __restrict__ count, const int* __restrict__ first_count, const int* __restrict__ work,
                       const int* __restrict__ work_first, int skip_rows, int64_t* __restrict__ idx_out,
                       float* __restrict__ best_out, int32_t* __restrict__ stats_out) {
  __shared__ __align__(16) float zs[8][64];
  __shared__ int s_rows[32];
  __shared__ float s_bv[4][16];
  __shared__ int s_bk[4][16];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int k = threadIdx.x;
  grid_dep_wait();
  int nflag = *count;
  if (work_first != nullptr && *first_count <= skip_rows) {
    nflag = *first_count;
    work = work_first;
  }
  if (stats_out && blockIdx.x == 0 && threadIdx.x == 0) {
    stats_out[0] = nflag;
    stats_out[1] = *first_count;
  }
  if ((int)blockIdx.x >= nflag) return;  // the common case: nothing queued
  // (loading the codes above the dependency wait was tried: the 128 KB per CTA are then read even when nothing is queued,
  //  and the search without queued rows went from 29.7 to 37.9 us)
  const int nv = D >> 2;
  float4 e[16];
#pragma unroll
  for (int v = 0; v < 16; ++v)
    e[v] = (k < K && v < nv) ? __ldg(reinterpret_cast<const float4*>(cb + (size_t)k * D) + v) : make_float4(0.f, 0.f, 0.f, 0.f);
  // Rows are scanned FOUR per trip (the per-row chain -- shared-memory reads, FMA chain, ten shuffles, two barriers, the
  // final warp -- is latency, so four independent rows share it) and reach shared memory through cp.async one trip
  // ahead; the CTA's row numbers are fetched 32 at a time so a row is one L2 round trip, not two dependent ones.
  constexpr int RB = 4;
  for (int i = threadIdx.x; i < 2 * RB * 64; i += 512) (&zs[0][0])[i] = 0.f;  // padded dims read as zero
  const int mine = (nflag - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;  // rows of this CTA
  for (int base = 0; base < mine; base += 32) {
    const int cnt = (mine - base < 32) ? mine - base : 32;
    __syncthreads();  // previous batch done with s_rows and the ring; the zero fill above is visible
    if ((int)threadIdx.x < cnt) s_rows[threadIdx.x] = __ldg(work + blockIdx.x + (size_t)(base + threadIdx.x) * gridDim.x);
    __syncthreads();
    auto issue = [&](int t) {  // trip t: rows 4 t .. 4 t + 3 of this batch -> ring half t & 1
      const int r = (int)threadIdx.x >> 6, d = (int)threadIdx.x & 63, j = RB * t + r;
      if (r < RB && j < cnt && d < D) {
        const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&zs[(t & 1) * RB + r][d]);
        const float* src = z + (int64_t)s_rows[j] * D + d;
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
      }
      asm volatile("cp.async.commit_group;" ::: "memory");  // (possibly empty: keeps the group count uniform)
    };
    const int trips = (cnt + RB - 1) / RB;
    issue(0);
    for (int t = 0; t < trips; ++t) {
      issue(t + 1);
      asm volatile("cp.async.wait_group 1;" ::: "memory");  // trip t has landed (this thread's part)
      __syncthreads();                                       // ... everyone's part; s_bv / s_bk of the previous trip consumed
      float a[RB][4];
#pragma unroll
      for (int r = 0; r < RB; ++r)
#pragma unroll
        for (int q = 0; q < 4; ++q) a[r][q] = 0.f;
#pragma unroll
      for (int v = 0; v < 16; ++v) {  // padded dims hold zeros on both sides
#pragma unroll
        for (int r = 0; r < RB; ++r) {
          const float4 z4 = *reinterpret_cast<const float4*>(&zs[(t & 1) * RB + r][4 * v]);
          float& acc = a[r][v & 3];
          if (metric == LIPVQ_METRIC_DOT) {
            acc = fmaf(z4.x, e[v].x, acc); acc = fmaf(z4.y, e[v].y, acc); acc = fmaf(z4.z, e[v].z, acc); acc = fmaf(z4.w, e[v].w, acc);
          } else {
            float df;
            df = z4.x - e[v].x; acc = fmaf(df, df, acc);
            df = z4.y - e[v].y; acc = fmaf(df, df, acc);
            df = z4.z - e[v].z; acc = fmaf(df, df, acc);
            df = z4.w - e[v].w; acc = fmaf(df, df, acc);
          }
        }
      }
      float bv[RB];
      int bk[RB];
#pragma unroll
      for (int r = 0; r < RB; ++r) {
        const float av = (a[r][0] + a[r][1]) + (a[r][2] + a[r][3]);
        bv[r] = (metric == LIPVQ_METRIC_L2) ? sqrtf(av) : (metric == LIPVQ_METRIC_DOT ? -av : av);
        bk[r] = k;
        if (k >= K || !(bv[r] == bv[r])) {  // NaN never wins (as in the other scans: val < best is false for NaN)
          bv[r] = INFINITY;
          bk[r] = 0x7fffffff;
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int r = 0; r < RB; ++r) {
          const float ov = __shfl_xor_sync(0xffffffffu, bv[r], o);
          const int ok = __shfl_xor_sync(0xffffffffu, bk[r], o);
          if (ov < bv[r] || (ov == bv[r] && ok < bk[r])) {
            bv[r] = ov;
            bk[r] = ok;
          }
        }
      }
      if (lane == 0) {
#pragma unroll
        for (int r = 0; r < RB; ++r) {
          s_bv[r][warp] = bv[r];
          s_bk[r][warp] = bk[r];
        }
      }
      __syncthreads();  // partial winners visible; every thread is done reading this ring half (refilled by issue(t + 2))
      if (warp < RB && RB * t + warp < cnt) {  // warp r finishes row r of the trip
        float fv = lane < 16 ? s_bv[warp][lane] : INFINITY;
        int fk = lane < 16 ? s_bk[warp][lane] : 0x7fffffff;
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) {
          const float ov = __shfl_xor_sync(0xffffffffu, fv, o);
          const int ok = __shfl_xor_sync(0xffffffffu, fk, o);
          if (ov < fv || (ov == fv && ok < fk)) {
            fv = ov;
            fk = ok;
          }
        }
        if (lane == 0) {
          if (fk == 0x7fffffff) fk = 0;  // all-NaN row: torch would return the NaN's index; we return 0
          const int64_t row = s_rows[RB * t + warp];
          idx_out[row] = (int64_t)fk;
          if (best_out) best_out[row] = (metric == LIPVQ_METRIC_DOT) ? -fv : fv;
        }
      }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
  }
}

__global__ void __launch_bounds__(256)
    l2_normalize_rows_kernel(const float* __restrict__ x, int64_t rows, int D, float* __restrict__ y) {
  const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const float* xr = x + row * (int64_t)D;
  float s = 0.f;
  for (int d = lane; d < D; d += 32) {
    const float v = xr[d];
    s = fmaf(v, v, s);
  }
  s = warp_sum(s);
  const float denom = fmaxf(sqrtf(s), 1e-12f);  // F.normalize: x / max(||x||_2, eps)
  float* yr = y + row * (int64_t)D;
  for (int d = lane; d < D; d += 32) yr[d] = xr[d] / denom;
}

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

}  // namespace

SearchPlan make_search_plan(int64_t N, int D, int K, int metric, int algo) {
  (void)metric;
  SearchPlan p{};
  if (algo == LIPVQ_ALGO_AUTO) algo = lipvq_nn_search_pick_algo(N, D, K, metric);
  p.algo = algo;
  p.Np = ceil_div64(N > 0 ? N : 1, 128) * 128;
  if (algo == LIPVQ_ALGO_TCGEN05) {
    int Kp, Dp;
    search_fused_plan(D, K, &Kp, &Dp);
    p.codes_per_split = Kp;
    p.ksplit = 0;  // the fused kernel keeps the running top-2 in registers: no per-slice partials
    p.Dp = Dp;
    p.Kp = Kp;
  } else {
    const int64_t row_tiles = ceil_div64(N > 0 ? N : 1, BM);
    const int code_tiles = (K + BN - 1) / BN;
    int ks = (int)ceil_div64(kTargetCtas, row_tiles);
    if (ks > code_tiles) ks = code_tiles;
    if (ks < 1) ks = 1;
    const int tiles_per_split = (code_tiles + ks - 1) / ks;
    p.codes_per_split = tiles_per_split * BN;
    p.ksplit = (K + p.codes_per_split - 1) / p.codes_per_split;
    p.Dp = (D + 63) / 64 * 64;
    p.Kp = (K + 127) / 128 * 128;
  }
  size_t o = 0;
  p.off_hdr = o;      o = align_up(o + sizeof(SearchHdr), 256);
  p.off_mu = o;       o = align_up(o + sizeof(float) * 512, 256);
  p.off_enorm = o;    o = align_up(o + sizeof(float) * (size_t)p.Kp, 256);
  p.off_pbest = o;    o = align_up(o + sizeof(float) * (size_t)p.ksplit * N, 256);
  p.off_psecond = o;  o = align_up(o + sizeof(float) * (size_t)p.ksplit * N, 256);
  p.off_pidx = o;     o = align_up(o + sizeof(int) * (size_t)p.ksplit * N, 256);
  if (algo == LIPVQ_ALGO_TCGEN05) {  // prepared codebook state: before every N-dependent array, so prepare() needs no N
    p.off_prep = o;   o = align_up(o + (search_p1_fits(N > 0 ? N : 1, D, K) ? search_p1_state_bytes(D, K) : 0), 1024);
  }
  const int64_t work_rows = (algo == LIPVQ_ALGO_TCGEN05 && search_p1_fits(N > 0 ? N : 1, D, K)) ? search_p1_queue_rows(N) : (N > 0 ? N : 1);
  p.off_work = o;     o = align_up(o + sizeof(int) * (size_t)work_rows, 256);
  if (algo == LIPVQ_ALGO_TCGEN05) {
    const size_t eb = sizeof(uint16_t) * (size_t)p.Kp * p.Dp;
    p.off_work2 = o; o = align_up(o + sizeof(int) * 2 * (size_t)work_rows, 256);  // (single-launch pass: two floats per queued row)
    p.off_ehi = o;  o = align_up(o + eb, 1024);
    p.off_elo = o;  o = align_up(o + eb, 1024);
    p.off_img = o;  o = align_up(o + (search_p1_fits(N, D, K) ? search_p1_image_bytes(D, K) : 0), 1024);
  }
  p.total = o;
  return p;
}

}  // namespace lipvq

using namespace lipvq;

extern "C" {

int lipvq_nn_search_pick_algo(int64_t N, int D, int K, int metric) {
  (void)metric;
  // LIPVQ_SEARCH_ALGO=simt|tc overrides the heuristic (A/B runs in bench.py and the profiling scripts)
  if (const char* f = getenv("LIPVQ_SEARCH_ALGO")) {
    if (f[0] == 's') return LIPVQ_ALGO_SIMT_F32;
    if (f[0] == 't' && search_fused_supported(N, D, K)) return LIPVQ_ALGO_TCGEN05;
  }
  // The tensor path pays a bf16 split pass over z and the codebook and 128x128 tiles; it wins once there is
  // enough contraction per row (profiles/ has the crossover).  Tiny problems stay on the FFMA kernel.
  if (search_fused_supported(N, D, K) && (int64_t)N * K >= (int64_t)1 << 20) return LIPVQ_ALGO_TCGEN05;
  return LIPVQ_ALGO_SIMT_F32;
}

size_t lipvq_nn_search_workspace_bytes(int64_t N, int D, int K, int metric, int algo) {
  if (N < 0 || D <= 0 || K <= 0) return 0;
  return make_search_plan(N, D, K, metric, algo).total;
}

int lipvq_nn_search_f32(const float* z, int64_t N, int D, const float* codebook, int K, int metric, int algo,
                        int64_t* idx_out, float* best_out, int32_t* stats_out, void* workspace,
                        size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(N >= 0 && D > 0 && K > 0, "nn_search: bad shape N=%lld D=%d K=%d", (long long)N, D, K);
  LIPVQ_REQUIRE(N < ((int64_t)1 << 31), "nn_search: N=%lld exceeds int32 row ids", (long long)N);
  LIPVQ_REQUIRE(metric >= LIPVQ_METRIC_L2 && metric <= LIPVQ_METRIC_DOT, "nn_search: bad metric %d", metric);
  LIPVQ_REQUIRE(algo >= LIPVQ_ALGO_AUTO && algo <= LIPVQ_ALGO_TCGEN05, "nn_search: bad algo %d", algo);
  if (N == 0) {
    if (stats_out) LIPVQ_CUDA(cudaMemsetAsync(stats_out, 0, 4 * sizeof(int32_t), stream));
    return LIPVQ_OK;
  }
  LIPVQ_REQUIRE(z && codebook && idx_out && workspace, "nn_search: null pointer");
  if (algo == LIPVQ_ALGO_TCGEN05)
    LIPVQ_REQUIRE(search_fused_supported(N, D, K), "nn_search: tcgen05 path does not support N=%lld D=%d K=%d",
                  (long long)N, D, K);
  const SearchPlan plan = make_search_plan(N, D, K, metric, algo);
  LIPVQ_REQUIRE(workspace_bytes >= plan.total, "nn_search: workspace %zu < required %zu", workspace_bytes, plan.total);
  LIPVQ_REQUIRE(aligned16(workspace), "nn_search: workspace must be 16-byte aligned");

  char* ws = (char*)workspace;
  SearchHdr* hdr = (SearchHdr*)(ws + plan.off_hdr);
  float* enorm = (float*)(ws + plan.off_enorm);
  float* pbest = (float*)(ws + plan.off_pbest);
  float* psecond = (float*)(ws + plan.off_psecond);
  int* pidx = (int*)(ws + plan.off_pidx);
  int* work = (int*)(ws + plan.off_work);

  LIPVQ_CUDA(cudaMemsetAsync(hdr, 0, sizeof(SearchHdr), stream));
  if (stats_out) LIPVQ_CUDA(cudaMemsetAsync(stats_out, 0, 4 * sizeof(int32_t), stream));

  float* mu = (float*)(ws + plan.off_mu);
  const int vec = (D % 4 == 0) && aligned16(z) && aligned16(codebook);
  const int rs_blocks = 4 * sm_count();
  int lpc = 1;  // lanes per code: enough to cover the row with one float4 (or one float) each, at most a warp
  while (lpc < 32 && lpc * (vec ? 4 : 1) < D) lpc <<= 1;
  // small codebooks: staged in shared memory, one lane per code (rescore_smem_kernel)
  const size_t rs_wide_smem = ((size_t)K * (D + 4) + D) * sizeof(float);
  const bool rs_wide = vec && D <= 256 && rs_wide_smem <= 200 * 1024;
  const char* rs_env = getenv("LIPVQ_RESCORE_WIDE");  // =0: one row per batch whatever the queue length (A/B runs)
  const int rs_multi = ((size_t)kRescoreWideR * D * sizeof(float) <= 48 * 1024 && !(rs_env && rs_env[0] == '0')) ? 1 : 0;
  const size_t rs_smem = (size_t)(rs_multi ? kRescoreWideR : 1) * D * sizeof(float);
  const bool rs_reg = vec && D <= 64 && K <= 512;  // one code per thread, held in registers (rescore_reg_kernel)
  if (rs_wide) {
    static size_t smem_set[kMaxDevices] = {};  // the attribute is per device
    const int dev_ = current_device();
    if (rs_wide_smem > smem_set[dev_]) {
      LIPVQ_CUDA(cudaFuncSetAttribute(rescore_smem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rs_wide_smem));
      smem_set[dev_] = rs_wide_smem;
    }
  }
  LIPVQ_REQUIRE(rs_smem <= 48 * 1024, "nn_search: D=%d too large for the rescore kernel", D);

  if (plan.algo == LIPVQ_ALGO_TCGEN05) {
    // prep + fp16 pass + 3-term pass (search_fused.cu); whatever is still queued gets the exact scan
    const int* queue = nullptr;
    const int* count = nullptr;
    const int rc = search_fused_launch(z, N, D, codebook, K, metric, plan, ws, idx_out, stream, &queue, &count, stats_out);
    if (rc != LIPVQ_OK) return rc;
    if (count == nullptr) {  // single-launch pass with its own exact scan: every row is settled
      if (best_out) return search_best_values(z, N, D, codebook, metric, idx_out, best_out, stream);
      return LIPVQ_OK;
    }
    static const bool dbg_no_tails = getenv("LIPVQ_DEBUG_NO_TAILS") != nullptr;  // timing experiments only
    if (dbg_no_tails) return LIPVQ_OK;
    if (rs_reg)
      LIPVQ_CUDA(launch_pdl(rescore_reg_kernel, dim3(sm_count()), dim3(512), 0, stream, z, D, codebook, K, metric, count,
                            (const int*)&hdr->flagged, queue, (const int*)(queue == work ? nullptr : work),
                            search_skip_rows(N, D, K), idx_out, (float*)nullptr, stats_out));
    else if (rs_wide)
      LIPVQ_CUDA(launch_pdl(rescore_smem_kernel, dim3(sm_count()), dim3(256), rs_wide_smem, stream, z, D, codebook, K, metric, count,
                            (const int*)&hdr->flagged, queue, (const int*)(queue == work ? nullptr : work),
                            search_skip_rows(N, D, K), idx_out, (float*)nullptr, stats_out));
    else
      LIPVQ_CUDA(launch_pdl(rescore_kernel<4, 1>, dim3(rs_blocks), dim3(256), rs_smem, stream, z, D, codebook, K, metric, count,
                            (const int*)&hdr->flagged, queue, (const int*)(queue == work ? nullptr : work),
                            search_skip_rows(N, D, K), idx_out, (float*)nullptr, stats_out, vec, lpc, rs_multi));
    LIPVQ_LAUNCHED("rescore_kernel");
    if (best_out) return search_best_values(z, N, D, codebook, metric, idx_out, best_out, stream);
    return LIPVQ_OK;
  }

  center_kernel<<<1, 256, 0, stream>>>(z, N, D, metric, mu, hdr);
  LIPVQ_LAUNCHED("center_kernel");
  code_sqnorm_kernel<<<(unsigned)ceil_div64((int64_t)K * 32, 256), 256, 0, stream>>>(codebook, K, D, metric, mu, enorm,
                                                                                     hdr);
  LIPVQ_LAUNCHED("code_sqnorm_kernel");

  const float gamma_en = (float)(D + 4) * 5.9604645e-8f;  // fp32 FMA chain of ||e||^2 plus the final add
  dim3 grid((unsigned)ceil_div64(N, BM), (unsigned)plan.ksplit);
  search_simt_kernel<<<grid, NT, 0, stream>>>(z, N, D, codebook, K, mu, enorm, plan.codes_per_split, pbest, psecond, pidx,
                                              vec);
  LIPVQ_LAUNCHED("search_simt_kernel");
  // fp32 FMA chain of length D plus the separately rounded ||e||^2 and the final fma: (D+4) unit roundoffs
  const float gamma = (float)(D + 4) * 5.9604645e-8f;
  const float gamma_ref = 4.f * (float)(D + 4) * 5.9604645e-8f;

  finalize_kernel<<<(unsigned)ceil_div64(N * 32, 256), 256, 0, stream>>>(z, N, D, codebook, mu, metric, plan.ksplit, pbest,
                                                                         psecond, pidx, gamma_en, gamma, gamma_ref, hdr, work,
                                                                         idx_out, best_out);
  LIPVQ_LAUNCHED("finalize_kernel");
  if (rs_reg)
    rescore_reg_kernel<<<sm_count(), 512, 0, stream>>>(z, D, codebook, K, metric, &hdr->flagged, &hdr->flagged, work,
                                                      (const int*)nullptr, 0, idx_out, best_out, stats_out);
  else if (rs_wide)
    rescore_smem_kernel<<<sm_count(), 256, rs_wide_smem, stream>>>(z, D, codebook, K, metric, &hdr->flagged, &hdr->flagged, work,
                                                      (const int*)nullptr, 0, idx_out, best_out, stats_out);
  else
    rescore_kernel<4, 1><<<rs_blocks, 256, rs_smem, stream>>>(z, D, codebook, K, metric, &hdr->flagged, &hdr->flagged, work,
                                                          (const int*)nullptr, 0, idx_out, best_out, stats_out, vec, lpc, rs_multi);
  LIPVQ_LAUNCHED("rescore_kernel");
  return LIPVQ_OK;
}

int lipvq_nn_search_prepared_supported(int64_t N, int D, int K, int metric) {
  if (N < 1 || D <= 0 || K <= 0) return 0;
  if (const char* e = getenv("LIPVQ_SEARCH_PREPARED"))  // A/B switch: 0 = callers fall back to lipvq_nn_search_f32
    if (e[0] == '0') return 0;
  return lipvq_nn_search_pick_algo(N, D, K, metric) == LIPVQ_ALGO_TCGEN05 && search_p1_supported(N, D, K) ? 1 : 0;
}

int lipvq_nn_search_prepare_f32(const float* codebook, int K, int D, int metric, const float* center_rows,
                                int64_t n_center_rows, void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(D > 0 && K > 0, "nn_search_prepare: bad shape D=%d K=%d", D, K);
  LIPVQ_REQUIRE(metric >= LIPVQ_METRIC_L2 && metric <= LIPVQ_METRIC_DOT, "nn_search_prepare: bad metric %d", metric);
  LIPVQ_REQUIRE(codebook && workspace, "nn_search_prepare: null pointer");
  LIPVQ_REQUIRE(center_rows == nullptr || n_center_rows >= 1, "nn_search_prepare: center_rows given with %lld rows",
                (long long)n_center_rows);
  if (!search_p1_fits(1, D, K)) {
    set_error("nn_search_prepare: codebook K=%d D=%d is streamed, not resident: use lipvq_nn_search_f32", K, D);
    return LIPVQ_ERR_UNSUPPORTED;
  }
  const SearchPlan plan = make_search_plan(1, D, K, metric, LIPVQ_ALGO_TCGEN05);
  LIPVQ_REQUIRE(workspace_bytes >= plan.off_prep + search_p1_state_bytes(D, K), "nn_search_prepare: workspace %zu too small",
                workspace_bytes);
  LIPVQ_REQUIRE(aligned16(workspace), "nn_search_prepare: workspace must be 16-byte aligned");
  return search_prep_launch(codebook, K, D, metric, center_rows, n_center_rows, (char*)workspace + plan.off_prep, stream);
}

int lipvq_nn_search_prepared_f32(const float* z, int64_t N, int D, const float* codebook, int K, int metric,
                                 int64_t* idx_out, int32_t* stats_out, void* workspace, size_t workspace_bytes,
                                 void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(N >= 1 && D > 0 && K > 0, "nn_search_prepared: bad shape N=%lld D=%d K=%d", (long long)N, D, K);
  LIPVQ_REQUIRE(N < ((int64_t)1 << 31), "nn_search_prepared: N=%lld exceeds int32 row ids", (long long)N);
  LIPVQ_REQUIRE(metric >= LIPVQ_METRIC_L2 && metric <= LIPVQ_METRIC_DOT, "nn_search_prepared: bad metric %d", metric);
  LIPVQ_REQUIRE(z && codebook && idx_out && workspace, "nn_search_prepared: null pointer");
  if (!search_p1_fits(N, D, K)) {
    set_error("nn_search_prepared: codebook K=%d D=%d is streamed, not resident: use lipvq_nn_search_f32", K, D);
    return LIPVQ_ERR_UNSUPPORTED;
  }
  const SearchPlan plan = make_search_plan(N, D, K, metric, LIPVQ_ALGO_TCGEN05);
  LIPVQ_REQUIRE(workspace_bytes >= plan.total, "nn_search_prepared: workspace %zu < required %zu", workspace_bytes, plan.total);
  LIPVQ_REQUIRE(aligned16(workspace), "nn_search_prepared: workspace must be 16-byte aligned");
  char* ws = (char*)workspace;
  return search_p1_launch(z, N, D, codebook, K, metric, 0, nullptr, nullptr, nullptr, nullptr, nullptr,
                          (SearchHdr*)(ws + plan.off_hdr), (int*)(ws + plan.off_work), nullptr, idx_out, stream,
                          ws + plan.off_prep, 1, stats_out, (float*)(ws + plan.off_work2));
}

int lipvq_l2_normalize_rows_f32(const float* x, int64_t rows, int D, float* y, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(rows >= 0 && D > 0, "l2_normalize_rows: bad shape");
  if (rows == 0) return LIPVQ_OK;
  LIPVQ_REQUIRE(x && y, "l2_normalize_rows: null pointer");
  l2_normalize_rows_kernel<<<(unsigned)ceil_div64(rows * 32, 256), 256, 0, stream>>>(x, rows, D, y);
  LIPVQ_LAUNCHED("l2_normalize_rows_kernel");
  return LIPVQ_OK;
}

}  // extern "C"
