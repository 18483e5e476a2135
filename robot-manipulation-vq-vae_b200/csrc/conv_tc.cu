// K8 on the tensor cores: the E -> E dilated convolution of an MS-TCN layer (kernel 3) as ONE contraction
//   pre[n, o] = sum_{j, c} X[n + (j - 1) d, c] * Wf[o][j][c]          (rows outside the sequence of n read as zero)
// i.e. a token-major linear layer whose input row is the concatenation of three shifted rows of X and whose weight is
// the packed [Eo][3 Ei] matrix of lipvq_conv_pack_weights_f32 -- followed by the layer's epilogue
//   OUT = (keep ? relu(pre + b) / (1 - p) : 0) + X,   mask = keep
// (reference robomimic/algo/heirarchical_vqvae.py:39-57: Conv1d(E, E, 3, dilation = d, padding = d) -> ReLU -> Dropout
// -> + x).  At the shipped width E = 512 a layer is 103 GFLOP per 65 536 tokens; the fp32 FFMA tiles of conv.cu need
// ~3 ms for it.
//
// fp32 accuracy as in linear_tc.cu: both operands split into three bf16 terms, six tcgen05.mma chains, fp32 accumulation
// in TMEM.  Tiles of 128 tokens x 128 outputs; the reduction (3 E, dead taps skipped) streams through a two-stage ring of
// 64-wide chunks: sixteen converter warps turn fp32 rows of X (shifted per tap, zero outside the sequence) AND fp32 rows
// of Wf into swizzled K-major bf16 tiles (the weights of a layer are 3 MB and stay in L2: no prepared copy, no workspace),
// one thread issues the 24 MMAs of a chunk, eight epilogue warps drain one of two accumulators while the next tile's
// chunks already flow.  Persistent CTAs; consecutive tiles share a row block (its X rows are read from L2, not HBM).
#include <cuda.h>
#include <cuda_bf16.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace lipvq {
namespace {

using namespace tc;

constexpr int kBM = 128, kBN = 128, kChunk = 64;
constexpr int kTile = kBM * kChunk * 2;      // one term of a 128-row, 64-wide bf16 tile: 16 KB
constexpr int kStage = 6 * kTile;            // A (h, m, l) + B (h, m, l): 96 KB
constexpr int kEpiWarps = 8, kCvtWarps = 16;
constexpr int kWarpMma = kEpiWarps + kCvtWarps;
constexpr int kThreads = 32 * (kWarpMma + 1);  // 800

enum { kFwd = 0, kBwdX = 1 };

struct ConvTcParams {
  const float* X; int64_t N; int T, E, dil, jlo, jhi;   // kBwdX: X = dOUT (masked by gmask, scaled) and OUT = dX
  const float* Wf; const float* bias;                   // kBwdX: Wf = Wb [3][Eo][Ei]
  float* OUT; uint8_t* mask;
  const uint8_t* gmask;
  float scale; uint32_t drop_thr, seed;
  int terms;   // 3: fp32 accuracy (three bf16 terms per operand, six chains); 1: bf16 mode (plain bf16 operands, one chain)
  int split;   // s > 0: one tile = (row block, column tile, tap, one of s ranges of the tap's chunks); raw partial sums are
               // added into OUT (see conv_tc_*_launch)
};

__device__ __forceinline__ uint32_t mix32(uint32_t a, uint32_t b) {  // (conv.cu: the same dropout stream)
  uint32_t h = a * 0x9E3779B1u ^ (b + 0x7F4A7C15u + (a << 6) + (a >> 2));
  h ^= h >> 16; h *= 0x85EBCA6Bu; h ^= h >> 13; h *= 0xC2B2AE35u; h ^= h >> 16;
  return h;
}
__device__ __forceinline__ uint32_t bf16x2_rn(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ void split3_pair(float x0, float x1, uint32_t& h, uint32_t& m, uint32_t& l) {
  h = bf16x2_rn(x0, x1);
  const float r0 = x0 - __uint_as_float(h << 16), r1 = x1 - __uint_as_float(h & 0xFFFF0000u);
  m = bf16x2_rn(r0, r1);
  l = bf16x2_rn(r0 - __uint_as_float(m << 16), r1 - __uint_as_float(m & 0xFFFF0000u));
}
// eight consecutive values of tile row r, 16-byte chunk ch: the three terms, kTile bytes apart
__device__ __forceinline__ void store_split8(uint8_t* base, int r, int ch, const float4& a, const float4& b, int terms = 3) {
  uint8_t* d = base + (uint32_t)(r * 128 + ((ch ^ (r & 7)) << 4));
  if (terms == 1) {  // bf16 mode: the rounded value alone
    *reinterpret_cast<uint4*>(d) = make_uint4(bf16x2_rn(a.x, a.y), bf16x2_rn(a.z, a.w), bf16x2_rn(b.x, b.y), bf16x2_rn(b.z, b.w));
    return;
  }
  uint32_t hh[4], mm[4], ll[4];
  split3_pair(a.x, a.y, hh[0], mm[0], ll[0]);
  split3_pair(a.z, a.w, hh[1], mm[1], ll[1]);
  split3_pair(b.x, b.y, hh[2], mm[2], ll[2]);
  split3_pair(b.z, b.w, hh[3], mm[3], ll[3]);
  *reinterpret_cast<uint4*>(d) = make_uint4(hh[0], hh[1], hh[2], hh[3]);
  *reinterpret_cast<uint4*>(d + kTile) = make_uint4(mm[0], mm[1], mm[2], mm[3]);
  *reinterpret_cast<uint4*>(d + 2 * kTile) = make_uint4(ll[0], ll[1], ll[2], ll[3]);
}

// MN-major, SWIZZLE_128B: LBO = distance between 64-element blocks of the M/N dimension, SBO = between 8-token groups
__device__ __forceinline__ uint64_t desc_mn_sw128(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((smem_addr & 0x3FFFF) >> 4) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) |
         (1ull << 46) | (2ull << 61);
}
constexpr int kWBlk = 64 * 128;  // one 64-channel block of one term of a 64-token chunk: 8 KB

// eight consecutive channels (16-byte chunk c of 16) of token tk of the chunk: the three terms, kTile bytes apart
__device__ __forceinline__ void store_split8_mn(uint8_t* base, int tk, int c, const float4& a, const float4& b, int terms = 3) {
  uint8_t* d = base + (uint32_t)((c >> 3) * kWBlk + (tk >> 3) * 1024 + (tk & 7) * 128 + (((c & 7) ^ (tk & 7)) << 4));
  if (terms == 1) {
    *reinterpret_cast<uint4*>(d) = make_uint4(bf16x2_rn(a.x, a.y), bf16x2_rn(a.z, a.w), bf16x2_rn(b.x, b.y), bf16x2_rn(b.z, b.w));
    return;
  }
  uint32_t hh[4], mm[4], ll[4];
  split3_pair(a.x, a.y, hh[0], mm[0], ll[0]);
  split3_pair(a.z, a.w, hh[1], mm[1], ll[1]);
  split3_pair(b.x, b.y, hh[2], mm[2], ll[2]);
  split3_pair(b.z, b.w, hh[3], mm[3], ll[3]);
  *reinterpret_cast<uint4*>(d) = make_uint4(hh[0], hh[1], hh[2], hh[3]);
  *reinterpret_cast<uint4*>(d + kTile) = make_uint4(mm[0], mm[1], mm[2], mm[3]);
  *reinterpret_cast<uint4*>(d + 2 * kTile) = make_uint4(ll[0], ll[1], ll[2], ll[3]);
}

__device__ __forceinline__ void mask8(float4& a, float4& b, const uint8_t* m, float scale) {
  const uint2 k = __ldg(reinterpret_cast<const uint2*>(m));
  a.x = (k.x & 0xFFu) ? a.x * scale : 0.f; a.y = (k.x & 0xFF00u) ? a.y * scale : 0.f;
  a.z = (k.x & 0xFF0000u) ? a.z * scale : 0.f; a.w = (k.x & 0xFF000000u) ? a.w * scale : 0.f;
  b.x = (k.y & 0xFFu) ? b.x * scale : 0.f; b.y = (k.y & 0xFF00u) ? b.y * scale : 0.f;
  b.z = (k.y & 0xFF0000u) ? b.z * scale : 0.f; b.w = (k.y & 0xFF000000u) ? b.w * scale : 0.f;
}

// MODE kFwd: the layer forward.  MODE kBwdX: its input gradient
//   dX[n, i] = dOUT[n, i] + sum_{j, o} masked(dOUT)[n + (1 - j) d, o] * Wb[j][o][i]
// -- the same contraction with the shifts negated; the B tile (64 outputs o of the reduction x 128 inputs i) keeps Wb's
// i-contiguous rows: an MN-major operand (layout of the weight-gradient kernel below), read with coalesced float4 loads.
template <int MODE, int TERMS>
__global__ void __launch_bounds__(kThreads, 1) conv_tc_kernel(const ConvTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 2 * kStage);
  const uint32_t bar_full = smem_u32(bars + 0), bar_empty = smem_u32(bars + 2);
  const uint32_t bar_acc_full = smem_u32(bars + 4), bar_acc_empty = smem_u32(bars + 6);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row_blocks = (int)((p.N + kBM - 1) / kBM), nct = p.E / kBN;
  const int ntap = p.split ? (p.jhi - p.jlo) * p.split : 1;  // split: the live taps (and ranges of their chunks) of a tile go to different CTAs
  const int tiles = row_blocks * nct * ntap;
  const int ntile = ((int)blockIdx.x < tiles) ? (tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  if (ntile == 0) return;
  const int cpt = p.E / kChunk;               // chunks per tap
  const int nk = p.split ? cpt / p.split : (p.jhi - p.jlo) * cpt;  // chunks per tile (dead taps skipped)

  if (threadIdx.x == 0) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_full + 8 * s, kCvtWarps);
      mbar_init(bar_empty + 8 * s, 1);
      mbar_init(bar_acc_full + 8 * s, 1);
      mbar_init(bar_acc_empty + 8 * s, kEpiWarps);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kWarpMma) tmem_alloc(smem_u32(tmem_slot), 4 * kBN);
  grid_dep_wait();
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == kWarpMma) {
    // ===== MMA issuer =====
    if (lane == 0) {
      // kBwdX: the weight tile is MN-major (input channels contiguous, see the converters), A stays K-major
      const uint32_t idesc = idesc_bf16(kBM, kBN) | (MODE == kBwdX ? (1u << 16) : 0u);
      const int ta[6] = {1, 2, 0, 1, 0, 0}, tb[6] = {1, 0, 2, 0, 1, 0};  // term 0 = h, 1 = m, 2 = l; smallest products first
      uint32_t a_lo[2][3], b_lo[2][3];
      uint64_t b_mn[2][3];
      for (int s = 0; s < 2; ++s)
        for (int t = 0; t < 3; ++t) {
          a_lo[s][t] = desc_sw128_lo(smem_u32(smem + s * kStage + t * kTile));
          b_lo[s][t] = desc_sw128_lo(smem_u32(smem + s * kStage + (3 + t) * kTile));
          b_mn[s][t] = desc_mn_sw128(smem_u32(smem + s * kStage + (3 + t) * kTile), kWBlk, 1024);
        }
      uint32_t g = 0;  // chunks issued (ring position)
      for (int i = 0; i < ntile; ++i) {
        const uint32_t ab = (uint32_t)i & 1;
        mbar_wait(bar_acc_empty + 8 * ab, (((uint32_t)i >> 1) & 1) ^ 1);
        fence_after();
        // Two accumulators per tile.  The tensor core TRUNCATES the fp32 accumulator on every accumulating MMA (measured: a
        // toward-zero bias growing linearly with the number of MMAs, 6.5e-6 at E = 512 with one accumulator), so the h*h
        // chain, whose partial sums have the result's magnitude, gets an accumulator of its own (a sixth of the truncation
        // events) and the five small chains (2^-8 and below) share the other, where a truncation is 2^-8 as large.
        const uint32_t d_hh = tmem_base + ab * (uint32_t)(2 * kBN), d_small = d_hh + (uint32_t)kBN;
        uint32_t acc_flag = 0, acc_small = 0;
        for (int kc = 0; kc < nk; ++kc, ++g) {
          const uint32_t s = g & 1;
          mbar_wait(bar_full + 8 * s, (g >> 1) & 1);
          fence_after();
#pragma unroll
          for (int combo = 0; combo < 6; ++combo) {
            if (TERMS == 1 && combo < 5) continue;  // bf16 mode: the h * h chain alone
            const uint32_t al = s ? a_lo[1][ta[combo]] : a_lo[0][ta[combo]];
            const uint32_t bl = s ? b_lo[1][tb[combo]] : b_lo[0][tb[combo]];
#pragma unroll
            for (int k = 0; k < kChunk / 16; ++k) {
              const uint32_t dst = combo < 5 ? d_small : d_hh;
              const uint32_t fl = combo < 5 ? acc_small : acc_flag;
              if (MODE == kBwdX) {  // 16 outputs o of the reduction = two 8-row atoms of the MN-major tile
                const uint64_t bm = s ? b_mn[1][tb[combo]] : b_mn[0][tb[combo]];
                umma_f16(dst, (((uint64_t)kDescSw128Hi) << 32 | (uint64_t)al) + (uint64_t)(2 * k), bm + (uint64_t)(128 * k), idesc, fl);
              } else {
                umma_f16_lo(dst, al + 2 * k, bl + 2 * k, kDescSw128Hi, idesc, fl);
              }
              if (combo < 5) acc_small = 1; else acc_flag = 1;
            }
          }
          umma_commit(bar_empty + 8 * s);
        }
        umma_commit(bar_acc_full + 8 * ab);
      }
    }
  } else if (warp >= kEpiWarps) {
    // ===== converters: per chunk 1024 items of A (shifted rows of X) and 1024 of B (rows of Wf), four per thread =====
    const int tcv = (int)threadIdx.x - kEpiWarps * 32;  // 0..511
    const int ch = tcv & 7, r0 = tcv >> 3;              // 16-byte chunk of the tile row; rows r0 and r0 + 64
    uint32_t g = 0;
    for (int i = 0; i < ntile; ++i) {
      const int tile = (int)blockIdx.x + i * (int)gridDim.x;
      const int t2 = tile / ntap, part = tile - t2 * ntap;
      const int jbase = p.jlo + (p.split ? part / p.split : 0), cbase = p.split ? (part % p.split) * nk : 0;
      const int rb = t2 / nct, ct = t2 - rb * nct;
      const int64_t row0 = (int64_t)rb * kBM;
      int64_t arow[2];
      int at[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        arow[u] = row0 + r0 + 64 * u;
        at[u] = (arow[u] < p.N) ? (int)(arow[u] % p.T) : -1;  // time index inside the sequence (-1: no such row)
      }
      const float* wrow[2] = {p.Wf + (size_t)(ct * kBN + r0) * 3 * p.E, p.Wf + (size_t)(ct * kBN + r0 + 64) * 3 * p.E};
      const int bc = tcv & 15, bo = tcv >> 4;  // kBwdX: B items (16-byte chunk bc of a Wb row, rows bo and bo + 32 of the chunk)
      for (int kc = 0; kc < nk; ++kc, ++g) {
        const int kq = kc + cbase;  // position in the tile's full chunk sequence (tap-major)
        const int tap = jbase + kq / cpt, c0 = (kq % cpt) * kChunk + ch * 8;
        const int shift = (MODE == kFwd) ? (tap - 1) * p.dil : (1 - tap) * p.dil;
        float4 a[2][2], b[2][2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int ts = at[u] + shift;
          if (at[u] >= 0 && ts >= 0 && ts < p.T) {
            const int64_t off = (arow[u] + shift) * p.E + c0;
            const float4* src = reinterpret_cast<const float4*>(p.X + off);
            a[u][0] = __ldg(src);
            a[u][1] = __ldg(src + 1);
            if (MODE == kBwdX) mask8(a[u][0], a[u][1], p.gmask + off, p.scale);
          } else {
            a[u][0] = a[u][1] = make_float4(0.f, 0.f, 0.f, 0.f);
          }
          if (MODE == kFwd) {
            const float4* ws = reinterpret_cast<const float4*>(wrow[u] + tap * p.E + c0);
            b[u][0] = __ldg(ws);
            b[u][1] = __ldg(ws + 1);
          } else {  // row o = bo + 32 u of the chunk of Wb[tap], input channels 8 bc .. 8 bc + 7 of the column tile
            const float4* ws = reinterpret_cast<const float4*>(p.Wf + ((size_t)tap * p.E + (kq % cpt) * kChunk + bo + 32 * u) * p.E +
                                                               ct * kBN + 8 * bc);
            b[u][0] = __ldg(ws);
            b[u][1] = __ldg(ws + 1);
          }
        }
        const uint32_t s = g & 1;
        mbar_wait(bar_empty + 8 * s, ((g >> 1) & 1) ^ 1);  // the loads above are in flight while this waits
        uint8_t* st = smem + s * kStage;
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          store_split8(st, r0 + 64 * u, ch, a[u][0], a[u][1], TERMS);
          if (MODE == kFwd) store_split8(st + 3 * kTile, r0 + 64 * u, ch, b[u][0], b[u][1], TERMS);
          else store_split8_mn(st + 3 * kTile, bo + 32 * u, bc, b[u][0], b[u][1], TERMS);
        }
        fence_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_full + 8 * s);
      }
    }
  } else {
    // ===== epilogue warps 0-7: TMEM lane quadrant q = warp & 3, 8-column groups part, part + 2, ... =====
    const int q = warp & 3, part = warp >> 2;
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    for (int i = 0; i < ntile; ++i) {
      const int tile = (int)blockIdx.x + i * (int)gridDim.x;
      const int t2 = tile / ntap;
      const int rb = t2 / nct, ct = t2 - rb * nct;
      const int64_t row = (int64_t)rb * kBM + q * 32 + lane;
      const uint32_t ab = (uint32_t)i & 1;
      mbar_wait(bar_acc_full + 8 * ab, ((uint32_t)i >> 1) & 1);
      fence_after();
      const uint32_t trow = tmem_base + ab * (uint32_t)(2 * kBN) + lane_off;
#pragma unroll 1
      for (int c0 = 8 * part; c0 < kBN; c0 += 16) {
        uint32_t v[8], w[8];
        tmem_ld8(trow + c0, v);
        if (TERMS == 3) tmem_ld8(trow + kBN + c0, w);
        tmem_ld_wait8(v);
        if (TERMS == 3) {
          tmem_ld_wait8(w);
#pragma unroll
          for (int e = 0; e < 8; ++e) v[e] = __float_as_uint(__uint_as_float(v[e]) + __uint_as_float(w[e]));
        }
        if (c0 + 16 >= kBN) {  // this warp's last load of the accumulator: hand it back before the arithmetic
          fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_acc_empty + 8 * ab);
        }
        if (p.split) {  // a tap's share of the sums: added to OUT, which the host side initialised
          if (row < p.N) {
            float* dst = p.OUT + row * p.E + ct * kBN + c0;
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(__uint_as_float(v[0])),
                         "f"(__uint_as_float(v[1])), "f"(__uint_as_float(v[2])), "f"(__uint_as_float(v[3])) : "memory");
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + 4), "f"(__uint_as_float(v[4])),
                         "f"(__uint_as_float(v[5])), "f"(__uint_as_float(v[6])), "f"(__uint_as_float(v[7])) : "memory");
          }
        } else if (MODE == kBwdX) {
          if (row < p.N) {
            const int col = ct * kBN + c0;
            const float4 x0 = __ldg(reinterpret_cast<const float4*>(p.X + row * p.E + col));  // residual branch of the layer
            const float4 x1 = __ldg(reinterpret_cast<const float4*>(p.X + row * p.E + col) + 1);
            float* dst = p.OUT + row * p.E + col;
            *reinterpret_cast<float4*>(dst) = make_float4(__uint_as_float(v[0]) + x0.x, __uint_as_float(v[1]) + x0.y,
                                                          __uint_as_float(v[2]) + x0.z, __uint_as_float(v[3]) + x0.w);
            *reinterpret_cast<float4*>(dst + 4) = make_float4(__uint_as_float(v[4]) + x1.x, __uint_as_float(v[5]) + x1.y,
                                                              __uint_as_float(v[6]) + x1.z, __uint_as_float(v[7]) + x1.w);
          }
        } else if (row < p.N) {
          const int col = ct * kBN + c0;
          const float4 b0 = __ldg(reinterpret_cast<const float4*>(p.bias + col)), b1 = __ldg(reinterpret_cast<const float4*>(p.bias + col) + 1);
          const float4 x0 = __ldg(reinterpret_cast<const float4*>(p.X + row * p.E + col));
          const float4 x1 = __ldg(reinterpret_cast<const float4*>(p.X + row * p.E + col) + 1);
          const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
          const float rr[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
          float o[8];
          uint32_t kb[2] = {0u, 0u};
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const float pre = __uint_as_float(v[e]) + bb[e];
            bool keep = pre > 0.f;
            if (p.drop_thr) keep = keep && (mix32(p.seed, (uint32_t)(row * p.E + col + e)) >= p.drop_thr);
            o[e] = (keep ? pre * p.scale : 0.f) + rr[e];
            kb[e >> 2] |= (keep ? 1u : 0u) << (8 * (e & 3));
          }
          float* dst = p.OUT + row * p.E + col;
          *reinterpret_cast<float4*>(dst) = make_float4(o[0], o[1], o[2], o[3]);
          *reinterpret_cast<float4*>(dst + 4) = make_float4(o[4], o[5], o[6], o[7]);
          *reinterpret_cast<uint2*>(p.mask + row * p.E + col) = make_uint2(kb[0], kb[1]);
        }
      }
    }
  }

  fence_before();
  __syncthreads();
  grid_dep_launch();
  if (warp == kWarpMma) {
    fence_after();
    tmem_dealloc(tmem_base, 4 * kBN);
  }
}

constexpr size_t kSmem = 1024 + 2 * (size_t)kStage + 128;

// Few tokens (a grid far below the SM count, e.g. the 160 tokens of the reference's own batch: 8 tiles, each a chain of 24
// chunks ~ 50 us): the live taps of a tile go to different CTAs (a third of the chain), which add their partial sums into
// OUT -- zeroed beforehand (forward; this kernel then applies the layer's epilogue in place) or preset to dOUT (input
// gradient: nothing left to do).
__global__ void __launch_bounds__(256) conv_split_epilogue_kernel(const ConvTcParams p) {
  grid_dep_wait();
  const int64_t n4 = p.N * p.E / 4;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t o = 4 * i;
    const int col = (int)(o % p.E);
    const float4 a = *reinterpret_cast<const float4*>(p.OUT + o);
    const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + col));
    const float4 x = __ldg(reinterpret_cast<const float4*>(p.X + o));
    const float pre[4] = {a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w};
    const float rr[4] = {x.x, x.y, x.z, x.w};
    float out[4];
    uint32_t kb = 0;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      bool keep = pre[e] > 0.f;
      if (p.drop_thr) keep = keep && (mix32(p.seed, (uint32_t)(o + e)) >= p.drop_thr);
      out[e] = (keep ? pre[e] * p.scale : 0.f) + rr[e];
      kb |= (keep ? 1u : 0u) << (8 * e);
    }
    *reinterpret_cast<float4*>(p.OUT + o) = make_float4(out[0], out[1], out[2], out[3]);
    *reinterpret_cast<uint32_t*>(p.mask + o) = kb;
  }
}

// LIPVQ_CONV_PRECISION=bf16: the bf16 mode of the three kernels (plain bf16 operands, ONE MMA chain, fp32 accumulation /
// bias / activation / I/O) -- values within 1e-2 of the fp32 evaluation (north_star's bf16 gate); default: fp32 accuracy
inline int conv_terms() {
  const char* f = getenv("LIPVQ_CONV_PRECISION");
  return (f && f[0] == 'b') ? 1 : 3;
}

// split by tap when that still leaves the grid inside one wave (LIPVQ_CONV_SPLIT=0: never)
inline int conv_split(int64_t N, int E, int jlo, int jhi) {
  if (const char* f = getenv("LIPVQ_CONV_SPLIT"))
    if (f[0] == '0') return 0;
  const int64_t tiles = ceil_div64(N, kBM) * (E / kBN) * (jhi - jlo);
  const int cpt = E / kChunk;
  if (jhi - jlo <= 1 && tiles * 2 > sm_count()) return 0;
  for (int s = 4; s >= 1; s >>= 1)  // each tap's chunks in s ranges while the grid stays inside one wave
    if (cpt % s == 0 && tiles * s <= sm_count() && (s > 1 || jhi - jlo > 1)) return s;
  return 0;
}

// ---- weight gradient ------------------------------------------------------------------------------------------------
//   dWf[o][j][i] += sum_tok masked(dOUT)[tok, o] * X[tok + (j - 1) d, i],     db[o] += sum_tok masked(dOUT)[tok, o]
// The reduction runs over TOKENS while both operands are stored channel-contiguous: they go into MN-major tiles (canonical
// SWIZZLE_128B layout: 8 tokens x 128 bytes = one 1024-byte atom, 16-byte chunk ^= token & 7; atoms of consecutive 8-token
// groups SBO = 1024 bytes apart, the two 64-channel blocks LBO = 8 KB apart -- the layout of linear_tc_bwd_w_kernel), so the
// converters read rows with the same coalesced float4 loads as the forward and nothing is transposed.  (A first version
// gathered K-major tiles with scalar loads: 0.99 ms per layer; with a branch per element even 3.09 ms, the sixteen round
// trips of a chunk serialised.)  One CTA = one 128 x 128 tile of dWf and one slice of the tokens (tiles x slices ~ the SM
// count); four accumulators: the five small chains in one, the h*h chain dealt over three by chunk (the accumulator
// truncates per MMA, see above: a long reduction into one accumulator would lose ~2e-5 of the slice's magnitude); the
// slice's result is added to dWf with vector reductions.
struct ConvTcWParams {
  const float* G; const uint8_t* gmask; const float* X;
  int64_t N; int T, E, dil, jlo, jhi;
  float scale;
  float* dWf; float* db;
  int chunks_per_slice, n_tiles;   // n_tiles = (jhi - jlo) * E / 128 column tiles
  int terms;                       // 3 / 1 as in ConvTcParams
};

__device__ __forceinline__ void red_add_v4(float* dst, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

template <int TERMS>
__global__ void __launch_bounds__(kThreads, 1) conv_tc_bwd_w_kernel(const ConvTcWParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 2 * kStage);
  const uint32_t bar_full = smem_u32(bars + 0), bar_empty = smem_u32(bars + 2), bar_acc_full = smem_u32(bars + 4);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);
  float* rs_s = reinterpret_cast<float*>(bars + 10);  // [128] bias gradient of the slice

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tile = (int)blockIdx.x, slice = (int)blockIdx.y;
  const int mt = tile / p.n_tiles, nt = tile - mt * p.n_tiles;
  const int tpt = p.E / kBN;                       // column tiles per tap
  const int tap = p.jlo + nt / tpt, i0 = (nt % tpt) * kBN, o0 = mt * kBM;
  const int64_t total_chunks = (p.N + kChunk - 1) / kChunk;
  const int64_t cbeg = (int64_t)slice * p.chunks_per_slice;
  const int nk = (int)((cbeg + p.chunks_per_slice <= total_chunks ? p.chunks_per_slice : total_chunks - cbeg));
  if (nk <= 0) return;

  if (threadIdx.x == 0) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_full + 8 * s, kCvtWarps);
      mbar_init(bar_empty + 8 * s, 1);
    }
    mbar_init(bar_acc_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kWarpMma) tmem_alloc(smem_u32(tmem_slot), 4 * kBN);
  if (threadIdx.x < 128) rs_s[threadIdx.x] = 0.f;
  grid_dep_wait();
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == kWarpMma) {
    if (lane == 0) {
      const uint32_t idesc = idesc_bf16(kBM, kBN) | (1u << 15) | (1u << 16);  // A and B MN-major
      const int ta[6] = {1, 2, 0, 1, 0, 0}, tb[6] = {1, 0, 2, 0, 1, 0};
      uint64_t da[2][3], db_[2][3];  // descriptors of (stage, term) at token 0; a 16-token step = two atoms = 2048 bytes
      for (int s = 0; s < 2; ++s)
        for (int t = 0; t < 3; ++t) {
          da[s][t] = desc_mn_sw128(smem_u32(smem + s * kStage + t * kTile), kWBlk, 1024);
          db_[s][t] = desc_mn_sw128(smem_u32(smem + s * kStage + (3 + t) * kTile), kWBlk, 1024);
        }
      uint32_t acc_small = 0, hh_used = 0;
      int rot = 0;
      for (int kc = 0; kc < nk; ++kc) {
        const uint32_t s = (uint32_t)kc & 1;
        mbar_wait(bar_full + 8 * s, ((uint32_t)kc >> 1) & 1);
        fence_after();
        const uint32_t d_hh = tmem_base + (uint32_t)((1 + rot) * kBN);
        uint32_t acc_hh = (hh_used >> rot) & 1u;
#pragma unroll
        for (int combo = 0; combo < 6; ++combo) {
          if (TERMS == 1 && combo < 5) continue;
          const uint64_t ad = s ? da[1][ta[combo]] : da[0][ta[combo]];
          const uint64_t bd = s ? db_[1][tb[combo]] : db_[0][tb[combo]];
#pragma unroll
          for (int k = 0; k < kChunk / 16; ++k) {
            if (combo < 5) {
              umma_f16(tmem_base, ad + (uint64_t)(128 * k), bd + (uint64_t)(128 * k), idesc, acc_small);
              acc_small = 1;
            } else {
              umma_f16(d_hh, ad + (uint64_t)(128 * k), bd + (uint64_t)(128 * k), idesc, acc_hh);
              acc_hh = 1;
            }
          }
        }
        hh_used |= 1u << rot;
        rot = rot == 2 ? 0 : rot + 1;
        umma_commit(bar_empty + 8 * s);
      }
      umma_commit(bar_acc_full);
    }
  } else if (warp >= kEpiWarps) {
    // ===== converters: item = (token of the chunk, 16-byte chunk of eight channels); a warp covers two token rows =====
    const int tcv = (int)threadIdx.x - kEpiWarps * 32;
    const int c = tcv & 15, tkb = tcv >> 4;          // channels 8c .. 8c+7 of the tile; tokens tkb and tkb + 32 of the chunk
    const int sh = (tap - 1) * p.dil;
    const float* gcol = p.G + o0 + 8 * c;
    const uint8_t* mcol = p.gmask + o0 + 8 * c;
    const float* xcol = p.X + i0 + 8 * c;
    float rsum[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) rsum[e] = 0.f;
    int t0[2];  // time index of this thread's two tokens, advanced by 64 per chunk
#pragma unroll
    for (int u = 0; u < 2; ++u) t0[u] = (int)((cbeg * kChunk + tkb + 32 * u) % p.T);
    const int tstep = kChunk % p.T;
    const int last = (int)p.N - 1;
    for (int kc = 0; kc < nk; ++kc) {
      const int k0 = (int)(cbeg + kc) * kChunk;   // (N E < 2^32: token indices and element offsets fit 32 bits)
      // loads unconditional (rows clamped into the arrays), conditions applied afterwards as selects
      float4 a[2][2], b[2][2];
      uint2 mk[2];
      bool live[2], inx[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int tok = k0 + tkb + 32 * u;
        const int tc = min(tok, last), xr = max(0, min(tc + sh, last));
        const uint32_t go = (uint32_t)tc * (uint32_t)p.E;
        a[u][0] = __ldg(reinterpret_cast<const float4*>(gcol + go));
        a[u][1] = __ldg(reinterpret_cast<const float4*>(gcol + go) + 1);
        mk[u] = __ldg(reinterpret_cast<const uint2*>(mcol + go));
        const float4* xs = reinterpret_cast<const float4*>(xcol + (uint32_t)xr * (uint32_t)p.E);
        b[u][0] = __ldg(xs);
        b[u][1] = __ldg(xs + 1);
        live[u] = tok <= last;
        const int ts = t0[u] + sh;
        inx[u] = live[u] && ts >= 0 && ts < p.T;
        t0[u] += tstep;
        if (t0[u] >= p.T) t0[u] -= p.T;
      }
      const uint32_t s = (uint32_t)kc & 1;
      mbar_wait(bar_empty + 8 * s, (((uint32_t)kc >> 1) & 1) ^ 1);
      uint8_t* st = smem + s * kStage;
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const float sc = live[u] ? p.scale : 0.f;
        float4 g0 = a[u][0], g1 = a[u][1];
        g0.x = (mk[u].x & 0xFFu) ? g0.x * sc : 0.f; g0.y = (mk[u].x & 0xFF00u) ? g0.y * sc : 0.f;
        g0.z = (mk[u].x & 0xFF0000u) ? g0.z * sc : 0.f; g0.w = (mk[u].x & 0xFF000000u) ? g0.w * sc : 0.f;
        g1.x = (mk[u].y & 0xFFu) ? g1.x * sc : 0.f; g1.y = (mk[u].y & 0xFF00u) ? g1.y * sc : 0.f;
        g1.z = (mk[u].y & 0xFF0000u) ? g1.z * sc : 0.f; g1.w = (mk[u].y & 0xFF000000u) ? g1.w * sc : 0.f;
        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
        const int tk = tkb + 32 * u;
        store_split8_mn(st, tk, c, g0, g1, TERMS);
        store_split8_mn(st + 3 * kTile, tk, c, inx[u] ? b[u][0] : z4, inx[u] ? b[u][1] : z4, TERMS);
        rsum[0] += g0.x; rsum[1] += g0.y; rsum[2] += g0.z; rsum[3] += g0.w;
        rsum[4] += g1.x; rsum[5] += g1.y; rsum[6] += g1.z; rsum[7] += g1.w;
      }
      fence_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_full + 8 * s);
    }
    if (p.db != nullptr && nt == 0) {
#pragma unroll
      for (int e = 0; e < 8; ++e) atomicAdd(rs_s + 8 * c + e, rsum[e]);
    }
  } else {
    // ===== epilogue warps: the slice's tile, summed over the four accumulators, added to dWf =====
    const int q = warp & 3, part = warp >> 2;
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    mbar_wait(bar_acc_full, 0);
    fence_after();
    const bool three = nk >= 3, two = nk >= 2;   // accumulators that were written at all
    float* drow = p.dWf + ((size_t)(o0 + q * 32 + lane) * 3 + tap) * p.E + i0;
#pragma unroll 1
    for (int c0 = 8 * part; c0 < kBN; c0 += 16) {
      uint32_t v[8], w[8];
      float o[8];
      if (TERMS == 3) tmem_ld8(tmem_base + lane_off + c0, v);   // (bf16 mode: the small-chain accumulator was never written)
      tmem_ld8(tmem_base + lane_off + kBN + c0, w);
      if (TERMS == 3) tmem_ld_wait8(v);
      tmem_ld_wait8(w);
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = (TERMS == 3 ? __uint_as_float(v[e]) : 0.f) + __uint_as_float(w[e]);
      if (two) {
        tmem_ld8(tmem_base + lane_off + 2 * kBN + c0, v);
        tmem_ld_wait8(v);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] += __uint_as_float(v[e]);
      }
      if (three) {
        tmem_ld8(tmem_base + lane_off + 3 * kBN + c0, v);
        tmem_ld_wait8(v);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] += __uint_as_float(v[e]);
      }
      red_add_v4(drow + c0, o[0], o[1], o[2], o[3]);
      red_add_v4(drow + c0 + 4, o[4], o[5], o[6], o[7]);
    }
  }

  fence_before();
  __syncthreads();
  grid_dep_launch();
  if (p.db != nullptr && nt == 0 && threadIdx.x < 128)
    atomicAdd(p.db + o0 + threadIdx.x, rs_s[threadIdx.x]);
  if (warp == kWarpMma) {
    fence_after();
    tmem_dealloc(tmem_base, 4 * kBN);
  }
}

constexpr size_t kSmemW = 1024 + 2 * (size_t)kStage + 128 + 4 * 128 * sizeof(float);

}  // namespace

// Shapes the tensor path takes: widths in whole tiles, any number of tokens from LIPVQ_CONV_TC_MIN_TOKENS (default 32) up --
// with a few hundred tokens only a handful of CTAs have work, but each walks its K = 3E in ~25 us where the FFMA tiles (as
// few CTAs, one 16-wide step at a time) need 140-190 us.  LIPVQ_CONV_ALGO=simt forces the FFMA tiles of conv.cu (A/B runs).
bool conv_tc_supported(int64_t N, int T, int Ei, int Eo) {
  if (const char* f = getenv("LIPVQ_CONV_ALGO"))
    if (f[0] == 's') return false;
  int64_t min_tokens = 32;
  if (const char* f = getenv("LIPVQ_CONV_TC_MIN_TOKENS")) min_tokens = atoll(f);
  return N >= min_tokens && T >= 1 && Ei == Eo && Ei % kBN == 0 && Ei >= kBN && Ei <= 4096 && N * (int64_t)Ei < ((int64_t)1 << 32);
}

int conv_tc_fwd_launch(const float* X, int64_t N, int T, int E, const float* Wf, const float* bias, int dilation, int jlo, int jhi,
                       float scale, uint32_t drop_thr, uint32_t seed, float* OUT, uint8_t* mask, cudaStream_t stream) {
  ConvTcParams p{};
  p.X = X; p.N = N; p.T = T; p.E = E; p.dil = dilation; p.jlo = jlo; p.jhi = jhi;
  p.Wf = Wf; p.bias = bias; p.OUT = OUT; p.mask = mask;
  p.scale = scale; p.drop_thr = drop_thr; p.seed = seed;
  static bool smem_set[kMaxDevices] = {};  // the attribute is per device
  const int dev_ = current_device();
  if (!smem_set[dev_]) {
    LIPVQ_CUDA(cudaFuncSetAttribute(conv_tc_kernel<kFwd, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem));
    LIPVQ_CUDA(cudaFuncSetAttribute(conv_tc_kernel<kFwd, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem));
    smem_set[dev_] = true;
  }
  p.split = conv_split(N, E, jlo, jhi);
  p.terms = conv_terms();
  const int64_t tiles = ceil_div64(N, kBM) * (E / kBN) * (p.split ? (jhi - jlo) * p.split : 1);
  const int sms = sm_count();
  if (p.split) LIPVQ_CUDA(cudaMemsetAsync(OUT, 0, (size_t)N * E * sizeof(float), stream));
  if (p.terms == 1) LIPVQ_CUDA(launch_pdl(conv_tc_kernel<kFwd, 1>, dim3((unsigned)(tiles < sms ? tiles : sms)), dim3(kThreads), kSmem, stream, p));
  else LIPVQ_CUDA(launch_pdl(conv_tc_kernel<kFwd, 3>, dim3((unsigned)(tiles < sms ? tiles : sms)), dim3(kThreads), kSmem, stream, p));
  LIPVQ_LAUNCHED("conv_tc_kernel<fwd>");
  if (p.split) {
    int64_t blocks = ceil_div64(N * E / 4, 256);
    if (blocks > 4 * sms) blocks = 4 * sms;
    LIPVQ_CUDA(launch_pdl(conv_split_epilogue_kernel, dim3((unsigned)blocks), dim3(256), 0, stream, p));
    LIPVQ_LAUNCHED("conv_split_epilogue_kernel");
  }
  return LIPVQ_OK;
}

int conv_tc_bwd_x_launch(const float* dOUT, const uint8_t* mask, int64_t N, int T, int E, const float* Wb, int dilation, int jlo,
                         int jhi, float scale, float* dX, cudaStream_t stream) {
  ConvTcParams p{};
  p.X = dOUT; p.gmask = mask; p.N = N; p.T = T; p.E = E; p.dil = dilation; p.jlo = jlo; p.jhi = jhi;
  p.Wf = Wb; p.OUT = dX; p.scale = scale;
  static bool smem_set[kMaxDevices] = {};
  const int dev_ = current_device();
  if (!smem_set[dev_]) {
    LIPVQ_CUDA(cudaFuncSetAttribute(conv_tc_kernel<kBwdX, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem));
    LIPVQ_CUDA(cudaFuncSetAttribute(conv_tc_kernel<kBwdX, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem));
    smem_set[dev_] = true;
  }
  p.split = conv_split(N, E, jlo, jhi);
  p.terms = conv_terms();
  const int64_t tiles = ceil_div64(N, kBM) * (E / kBN) * (p.split ? (jhi - jlo) * p.split : 1);
  const int sms = sm_count();
  if (p.split) LIPVQ_CUDA(cudaMemcpyAsync(dX, dOUT, (size_t)N * E * sizeof(float), cudaMemcpyDeviceToDevice, stream));
  if (p.terms == 1) LIPVQ_CUDA(launch_pdl(conv_tc_kernel<kBwdX, 1>, dim3((unsigned)(tiles < sms ? tiles : sms)), dim3(kThreads), kSmem, stream, p));
  else LIPVQ_CUDA(launch_pdl(conv_tc_kernel<kBwdX, 3>, dim3((unsigned)(tiles < sms ? tiles : sms)), dim3(kThreads), kSmem, stream, p));
  LIPVQ_LAUNCHED("conv_tc_kernel<bwd_x>");
  return LIPVQ_OK;
}

int conv_tc_bwd_w_launch(const float* dOUT, const uint8_t* mask, const float* X, int64_t N, int T, int E, int dilation, int jlo,
                         int jhi, float scale, float* dWf, float* db, cudaStream_t stream) {
  ConvTcWParams p{};
  p.G = dOUT; p.gmask = mask; p.X = X; p.N = N; p.T = T; p.E = E; p.dil = dilation; p.jlo = jlo; p.jhi = jhi;
  p.scale = scale; p.dWf = dWf; p.db = db;
  p.n_tiles = (jhi - jlo) * (E / kBN);
  p.terms = conv_terms();
  const int tiles = (E / kBM) * p.n_tiles;
  const int64_t total_chunks = ceil_div64(N, kChunk);
  int64_t slices = sm_count() / tiles;
  if (slices < 1) slices = 1;
  if (slices > total_chunks) slices = total_chunks;
  p.chunks_per_slice = (int)ceil_div64(total_chunks, slices);
  slices = ceil_div64(total_chunks, p.chunks_per_slice);
  static bool smem_set[kMaxDevices] = {};
  const int dev_ = current_device();
  if (!smem_set[dev_]) {
    LIPVQ_CUDA(cudaFuncSetAttribute(conv_tc_bwd_w_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemW));
    LIPVQ_CUDA(cudaFuncSetAttribute(conv_tc_bwd_w_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemW));
    smem_set[dev_] = true;
  }
  if (p.terms == 1) LIPVQ_CUDA(launch_pdl(conv_tc_bwd_w_kernel<1>, dim3((unsigned)tiles, (unsigned)slices), dim3(kThreads), kSmemW, stream, p));
  else LIPVQ_CUDA(launch_pdl(conv_tc_bwd_w_kernel<3>, dim3((unsigned)tiles, (unsigned)slices), dim3(kThreads), kSmemW, stream, p));
  LIPVQ_LAUNCHED("conv_tc_bwd_w_kernel");
  return LIPVQ_OK;
}

}  // namespace lipvq
