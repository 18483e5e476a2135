// K3, third generation of the first (fp16) pass: ONE launch that prepares the codebook itself and folds the scores with
// ~1.7 ALU-pipe instructions per score instead of 5.
//
// Used when the whole prepared codebook is resident in shared memory (K * Dp * 2 B + 32 K bytes next to the z tiles;
// e.g. every tokenizer codebook of the reference: 512 x 64, 1024 x 208 -> no, that one streams).  Larger codebooks keep
// the streamed kernel of search_fused.cu.  Differences from search_fused_kernel<1>:
//
//   * no preparation kernel, no header memset: every CTA derives mu (mean of <= 32 sampled rows), -2(e-mu) in fp16 and
//     ||e-mu||^2 from the fp32 codebook in its prologue (128 KB from L2 for 512 x 64) and writes the fp16 tiles straight
//     into their swizzled UMMA layout; CTAs also publish slices of those arrays for the (rarely needed) later passes;
//   * ||e-mu||^2 reaches the accumulator through the tensor core -- one extra K=16 MMA per code tile: a constant
//     [1,1,1,0..] slab of A times a [hi,mid,lo,0..] fp16 split of the norm -- instead of a tcgen05.st pre-store by the
//     epilogue warps (900 cycles per tile in the round-1 kernel, and the accumulator hand-back no longer waits for it);
//   * the fold tracks the BEST score only, as a packed key (score with its 5 low mantissa bits replaced by the column
//     within the 32-column chunk): one LOP3 + half an FMNMX3 per score on the ALU pipe.  Whether any OTHER score of
//     the chunk lies within the certification margin of the chunk's best is counted on the FMA pipe, which the fold
//     leaves idle: u = sat((best + margin - s) * (1/margin)) is 1 for the best, 0 for every clearly larger score and
//     fractional inside the margin; the row is certified only if the count of its winning chunk is exactly 1.0 and the
//     best of every other chunk (a chunk-level top-2, a handful of instructions per 32 scores) is beyond the margin.
//     Rigorous in the same sense as before: a row leaves this pass with an index only when the index is proven.
//
// Reference op replaced: [N,K,D] broadcast + norm + argmin of LFQQuantizer.forward
// (robomimic/models/vq_vae/backbone_lfqvae_v5.py:39-46) and the variants listed in search.cu.
#include <cuda.h>
#include <cuda_fp16.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "search.cuh"
#include "tc_common.cuh"

namespace lipvq {
namespace {

using namespace tc;

constexpr int kBM = 128;    // z rows per block (UMMA_M)
constexpr int kChunk = 64;  // latent elements per swizzled tile row
constexpr int kBN = 256;    // codes per tile (UMMA_N); TMEM holds two tiles
constexpr int kEpiWarps = 8, kCvtWarps = 4;
constexpr int kWarpMma = kEpiWarps + kCvtWarps;
constexpr int kThreads = 32 * (kWarpMma + 1);  // 416
constexpr int kZStage = kBM * kChunk * 4;      // staged fp32 z tile
constexpr int kStatSlots = 4;
constexpr float kF16Safe = 3.0e4f;
constexpr float kEnormMax = 6.0e4f;        // ||e-mu||^2 must fit an fp16 (it travels through the norm slab)
constexpr float kSentinel = 65504.f;       // padding / excluded codes: 3 x 65504 in the slab, zeros in the tile

struct P1Params {
  const float* z;
  const float* cb;
  int64_t N;
  int D, K, Kp, NC, T;
  int nbufA, zstages;
  int metric;
  int Kp_pub;     // rows of the published arrays (the later passes' padded code count)
  uint32_t kmask; // 0xFFFFFFE0 (see pack_key)
  int slab_swap;  // debug: swap LBO/SBO of the no-swizzle descriptors
  // outputs for the later passes (search_fused_kernel<3>, rescore): published by slices of the grid
  float* mu_g;
  float* enorm_g;
  __half* ehi_g;
  __half* elo_g;
  SearchHdr* hdr;
  int* queue_out;
  int* queue_count;
  int64_t* idx_out;
  float gamma_en, gamma_dot, gamma_ref;
};

__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait(uint32_t (&v)[32]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]), "+r"(v[8]),
                 "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15]), "+r"(v[16]),
                 "+r"(v[17]), "+r"(v[18]), "+r"(v[19]), "+r"(v[20]), "+r"(v[21]), "+r"(v[22]), "+r"(v[23]), "+r"(v[24]),
                 "+r"(v[25]), "+r"(v[26]), "+r"(v[27]), "+r"(v[28]), "+r"(v[29]), "+r"(v[30]), "+r"(v[31])
               :
               : "memory");
}
// A = B = f16 (format 0), D = f32, both K-major
__device__ __forceinline__ uint32_t idesc_f16(int M, int N) {
  return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// K-major, no swizzle: 8-row x 16-byte core matrices; lbo = byte step between core matrices along K, sbo = between
// 8-row groups (canonical layout ((8,m),(T,2k)):((1T,SBO),(1,LBO)) in 16-byte units)
__device__ __forceinline__ uint64_t desc_noswz(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((smem_addr & 0x3FFFF) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46);
}
// byte offset of element (row r, column c < 16) in a no-swizzle slab of 16 fp16 columns
__device__ __forceinline__ uint32_t slab_off(int r, int c) {
  return (uint32_t)((r >> 3) * 256 + (c >> 3) * 128 + (r & 7) * 16 + (c & 7) * 2);
}

__device__ __forceinline__ float fmin3(float a, float b, float c) {
  float d;
  asm("min.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}
__device__ __forceinline__ float ffma_sat(float a, float b, float c) {
  float d;
  asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}
// (s & mask) | col : the packed key of one score
__device__ __forceinline__ float pack_key(uint32_t s, uint32_t mask, uint32_t col) {
  uint32_t d;
  asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(d) : "r"(s), "r"(mask), "r"(col));
  return __uint_as_float(d);
}

struct RowBest {  // partial result of one epilogue group for one row, handed to the group that certifies the row
  float best, second, ma, mb;  // ma/mb: count (FOLD 0) or the two 16-bit near-best masks (FOLD 1) of the winning chunk
  int base;
  int pad[3];
};

// Warp roles (416 threads, one CTA per SM):
//   0-7   epilogue: warp w owns TMEM lanes 32*(w&3).. (= rows); group w>>2 drains every second code tile
//   8-11  converters: fp32 z rows -> centred fp16 -> swizzled A tiles (+ per-row ||z-mu||^2, ||z||^2)
//   12    TMEM allocation + single-thread tcgen05.mma issue
template <int FOLD>
__global__ void __launch_bounds__(kThreads, 1) search_p1_kernel(const __grid_constant__ CUtensorMap tm_z, const P1Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int a_tile = kBM * kChunk * 2;  // 16 KB
  const int a_bytes = p.NC * a_tile;
  const int b_tile = kBN * kChunk * 2;  // 32 KB
  uint8_t* sA = smem;
  uint8_t* sB = sA + p.nbufA * a_bytes;                 // [T][NC] resident code tiles
  uint8_t* sZ = sB + p.T * p.NC * b_tile;               // [zstages][128][64] fp32
  uint8_t* sE = sZ + p.zstages * kZStage;               // norm slab [Kp x 16] fp16, no swizzle
  uint8_t* sA1 = sE + p.Kp * 32;                        // ones slab [128 x 16] fp16
  float2* rowstat = reinterpret_cast<float2*>(sA1 + kBM * 32);             // [kStatSlots][128]
  RowBest* xchg = reinterpret_cast<RowBest*>(rowstat + kStatSlots * kBM);  // [2][128]
  float* mu_s = reinterpret_cast<float*>(xchg + 2 * kBM);                  // [256]
  float* red_s = mu_s + 256;                                               // [32] block reductions
  uint64_t* bars = reinterpret_cast<uint64_t*>(red_s + 32);
  const uint32_t bar_a_full = smem_u32(bars + 0), bar_a_empty = smem_u32(bars + 2);
  const uint32_t bar_acc_full = smem_u32(bars + 4), bar_acc_empty = smem_u32(bars + 6);
  const uint32_t bar_z_full = smem_u32(bars + 8), bar_z_empty = smem_u32(bars + 12);
  const uint32_t bar_x_full = smem_u32(bars + 16), bar_x_empty = smem_u32(bars + 18);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 20);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row_blocks = (int)((p.N + kBM - 1) / kBM);
  const int nblk = ((int)blockIdx.x < row_blocks) ? (row_blocks - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  const int total_tiles = nblk * p.T;
  const int Dp = p.NC * kChunk;

  if (threadIdx.x == 0) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_a_full + 8 * s, kCvtWarps);
      mbar_init(bar_a_empty + 8 * s, 1);
      mbar_init(bar_acc_full + 8 * s, 1);
      mbar_init(bar_acc_empty + 8 * s, kEpiWarps / 2);
      mbar_init(bar_x_full + 8 * s, kEpiWarps / 2);
      mbar_init(bar_x_empty + 8 * s, kEpiWarps / 2);
    }
    for (int s = 0; s < 4; ++s) {
      mbar_init(bar_z_full + 8 * s, 1);
      mbar_init(bar_z_empty + 8 * s, kCvtWarps);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kWarpMma) tmem_alloc(smem_u32(tmem_slot), 512);
  grid_dep_wait();  // z (and, in a training step, the codebook) are the previous kernels' outputs
  __syncthreads();  // barriers initialised before the first bulk copy names one

  // staged z rows: the first tiles start flowing now and land while the codebook is being prepared
  const int total_units = nblk * p.NC;
  auto z_issue = [&](int unit) {
    const int i2 = unit / p.NC, c2 = unit - i2 * p.NC;
    const uint32_t st = (uint32_t)unit % (uint32_t)p.zstages;
    const int64_t l0 = ((int64_t)blockIdx.x + (int64_t)i2 * gridDim.x) * kBM;
    mbar_expect_tx(bar_z_full + 8 * st, (uint32_t)kZStage);
    tma_load_2d(smem_u32(sZ + st * kZStage), &tm_z, bar_z_full + 8 * st, c2 * kChunk, (int)l0);
  };
  if (p.zstages > 0 && threadIdx.x == kEpiWarps * 32)
    for (int u = 0; u < p.zstages && u < total_units; ++u) z_issue(u);

  // ---- prologue: mu, then the whole codebook -> fp16 tiles + norm slab (all threads) ----------------------------------
  {
    constexpr int RMAX = 32;
    const int R = (int)(p.N < RMAX ? p.N : RMAX);
    const int64_t stride = p.N / R;
    for (int d = threadIdx.x; d < 256; d += kThreads) {
      float m = 0.f;
      if (d < p.D && p.metric != LIPVQ_METRIC_DOT) {
        float v[RMAX];
#pragma unroll
        for (int i = 0; i < RMAX; ++i) v[i] = (i < R) ? __ldg(p.z + (int64_t)i * stride * p.D + d) : 0.f;
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < RMAX; ++i) s += v[i];  // same order as fused_prep_kernel / center_kernel
        m = s / (float)R;
      }
      mu_s[d] = m;
    }
    // ones slab: columns 0..2 = 1.0 for every row
    for (int r = threadIdx.x; r < kBM; r += kThreads) {
      *reinterpret_cast<uint4*>(sA1 + slab_off(r, 0)) = make_uint4(0x3C003C00u, 0x00003C00u, 0u, 0u);  // 1, 1, 1, 0 ...
      *reinterpret_cast<uint4*>(sA1 + slab_off(r, 8)) = make_uint4(0u, 0u, 0u, 0u);
    }
    __syncthreads();
    if (warp == 0) {  // ||mu||
      float t = 0.f;
      for (int d = lane; d < 256; d += 32) t = fmaf(mu_s[d], mu_s[d], t);
      t = warp_sum(t);
      if (lane == 0) red_s[31] = t;
    }
  }
  float emax2_loc = 0.f, excl_inv_loc = 0.f;  // per-thread partial maxima
  {
    // item = (code k, 16-byte output chunk ch8 of its row): 8 latent dims = two float4 loads.  Items of one row are
    // consecutive lanes (Dp/8 = 8, 16 or 32 lanes), so ||e-mu||^2 is a fixed-order shuffle reduction.
    const int ipr = Dp / 8;  // items per row: 8 (D<=64), 16, 24 (-> handled as 32 with idle lanes), 32
    const int ipr_p2 = ipr <= 8 ? 8 : (ipr <= 16 ? 16 : 32);
    const int rows_per_pass = kThreads / ipr_p2;  // whole rows per pass over the CTA (416/8 = 52, /16 = 26, /32 = 13)
    const int sub = (int)threadIdx.x % ipr_p2, rslot = (int)threadIdx.x / ipr_p2;
    const bool vec = (p.D % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.cb) & 15) == 0);
    const bool publish_all = (gridDim.x == 1);
    constexpr int UNR = 4;  // rows per thread in flight: the prologue is one or two round trips to L2, not ten
    for (int k00 = 0; k00 < p.Kp; k00 += rows_per_pass * UNR) {
      float x[UNR][8];
      const int d0 = sub * 8;
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        const int k = k00 + u * rows_per_pass + rslot;
#pragma unroll
        for (int e = 0; e < 8; ++e) x[u][e] = 0.f;
        if (k < p.K && sub < ipr && d0 < p.D) {
          const float* src = p.cb + (size_t)k * p.D + d0;
          if (vec && d0 + 8 <= p.D) {
            const float4 a = __ldg(reinterpret_cast<const float4*>(src)), b = __ldg(reinterpret_cast<const float4*>(src) + 1);
            x[u][0] = a.x; x[u][1] = a.y; x[u][2] = a.z; x[u][3] = a.w; x[u][4] = b.x; x[u][5] = b.y; x[u][6] = b.z; x[u][7] = b.w;
          } else {
#pragma unroll
            for (int e = 0; e < 8; ++e) x[u][e] = (d0 + e < p.D) ? __ldg(src + e) : 0.f;
          }
        }
      }
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        const int k = k00 + u * rows_per_pass + rslot;
        const bool live = k < p.K && sub < ipr && d0 < p.D;
        float s = 0.f, amax = 0.f;
        bool finite = true;
        float cdiff[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float v = (live && d0 + e < p.D) ? x[u][e] - mu_s[d0 + e] : 0.f;
          cdiff[e] = v;
          s = fmaf(v, v, s);
          amax = fmaxf(amax, fabsf(v));
          finite = finite && (fabsf(v) <= 3.0e38f);
        }
        for (int o = 1; o < ipr_p2; o <<= 1) {
          s += __shfl_xor_sync(0xffffffffu, s, o);
          amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
          finite = finite && (__shfl_xor_sync(0xffffffffu, (int)finite, o) != 0);
        }
        if (k >= p.Kp) continue;  // (after the shuffles: every lane of the warp takes part in them)
        const bool valid = k < p.K;
        const bool excluded = valid && (!finite || amax >= kF16Safe || (p.metric != LIPVQ_METRIC_DOT && !(s <= kEnormMax)));
        const bool zero_row = !valid || excluded;
        const bool publish = k < p.Kp_pub && (publish_all || (k % (int)gridDim.x) == (int)blockIdx.x);
        if (sub < ipr) {
          uint32_t hh[4], ll[4];
#pragma unroll
          for (int e = 0; e < 8; e += 2) {
            const float va = zero_row ? 0.f : -2.f * cdiff[e], vb = zero_row ? 0.f : -2.f * cdiff[e + 1];
            const __half2 hp = __floats2half2_rn(va, vb);
            const float2 back = __half22float2(hp);
            const __half2 lp = __floats2half2_rn(va - back.x, vb - back.y);
            hh[e >> 1] = *reinterpret_cast<const uint32_t*>(&hp);
            ll[e >> 1] = *reinterpret_cast<const uint32_t*>(&lp);
          }
          const int j = k / kBN, r = k % kBN, c = sub >> 3, ch = sub & 7;
          *reinterpret_cast<uint4*>(sB + (size_t)(j * p.NC + c) * b_tile + r * 128 + ((ch ^ (r & 7)) << 4)) =
              make_uint4(hh[0], hh[1], hh[2], hh[3]);
          // later passes read the prepared codebook from global memory: each CTA publishes a slice of the rows
          if (publish) {
            *reinterpret_cast<uint4*>(p.ehi_g + (size_t)k * Dp + sub * 8) = make_uint4(hh[0], hh[1], hh[2], hh[3]);
            *reinterpret_cast<uint4*>(p.elo_g + (size_t)k * Dp + sub * 8) = make_uint4(ll[0], ll[1], ll[2], ll[3]);
          }
        }
        if (sub == 0) {
          const float en = (p.metric == LIPVQ_METRIC_DOT) ? 0.f : s;
          float h0, h1, h2;
          if (zero_row) {
            h0 = h1 = h2 = kSentinel;
          } else {
            h0 = __half2float(__float2half_rn(en));
            h1 = __half2float(__float2half_rn(en - h0));
            h2 = __half2float(__float2half_rn(en - h0 - h1));
            emax2_loc = fmaxf(emax2_loc, s);
          }
          if (excluded) {
            const float inv = (finite && s < 3.0e38f) ? rsqrtf(s) * 1.0001f : 3.0e38f;  // 1/||e-mu|| rounded up
            excl_inv_loc = fmaxf(excl_inv_loc, inv);
          }
          uint32_t w0, w1;
          {
            const __half2 a = __floats2half2_rn(h0, h1), b = __floats2half2_rn(h2, 0.f);
            w0 = *reinterpret_cast<const uint32_t*>(&a);
            w1 = *reinterpret_cast<const uint32_t*>(&b);
          }
          *reinterpret_cast<uint4*>(sE + slab_off(k, 0)) = make_uint4(w0, w1, 0u, 0u);  // columns 0..7
          *reinterpret_cast<uint4*>(sE + slab_off(k, 8)) = make_uint4(0u, 0u, 0u, 0u);  // columns 8..15
          if (publish) p.enorm_g[k] = zero_row ? INFINITY : en;
        }
      }
    }
  }
  // block-wide maxima (every CTA holds the same values: same data, same order)
  {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      emax2_loc = fmaxf(emax2_loc, __shfl_xor_sync(0xffffffffu, emax2_loc, o));
      excl_inv_loc = fmaxf(excl_inv_loc, __shfl_xor_sync(0xffffffffu, excl_inv_loc, o));
    }
    if (lane == 0) {
      red_s[warp] = emax2_loc;
      red_s[16 + warp] = excl_inv_loc;
    }
  }
  fence_async_smem();  // generic-proxy writes of the code tiles / slabs -> visible to the tensor core
  fence_before();
  __syncthreads();
  fence_after();
  float emax2 = 0.f, excl_inv = 0.f;
  for (int w = 0; w <= kWarpMma; ++w) {
    emax2 = fmaxf(emax2, red_s[w]);
    excl_inv = fmaxf(excl_inv, red_s[16 + w]);
  }
  const float mu_norm = sqrtf(red_s[31]);
  if (blockIdx.x == 0) {  // what the later passes read
    for (int d = threadIdx.x; d < Dp; d += kThreads) p.mu_g[d] = mu_s[d];
    if (threadIdx.x == 0) {
      p.hdr->emax2_bits = __float_as_uint(emax2);
      p.hdr->excl_inv_bits = __float_as_uint(excl_inv);
      p.hdr->mu_norm = mu_norm;
    }
  }
  const uint32_t tmem_base = *tmem_slot;

  if (warp == kWarpMma) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0) {
      const uint32_t idesc = idesc_f16(kBM, kBN);
      const uint32_t lbo = p.slab_swap ? 256u : 128u, sbo = p.slab_swap ? 128u : 256u;
      const uint64_t a1desc = desc_noswz(smem_u32(sA1), lbo, sbo);
      uint32_t t = 0;
      for (int i = 0; i < nblk; ++i) {
        const int buf = i % p.nbufA;
        mbar_wait(bar_a_full + 8 * buf, (uint32_t)(i / p.nbufA) & 1);
        fence_after();
        const uint32_t a_base = smem_u32(sA + buf * a_bytes);
        for (int j = 0; j < p.T; ++j, ++t) {
          const uint32_t ab = t & 1;
          mbar_wait(bar_acc_empty + 8 * ab, ((t >> 1) & 1) ^ 1);  // the epilogue has drained the tile before last
          fence_after();
          const uint32_t d_tmem = tmem_base + ab * (uint32_t)kBN;
          // score = ||e-mu||^2 (norm slab, overwrites the accumulator) - 2 (z-mu).(e-mu)
          umma_f16(d_tmem, a1desc, desc_noswz(smem_u32(sE + (size_t)j * kBN * 32), lbo, sbo), idesc, 0u);
          for (int c = 0; c < p.NC; ++c) {
            const uint32_t aa = a_base + (uint32_t)(c * a_tile);
            const uint32_t bb = smem_u32(sB + (size_t)(j * p.NC + c) * b_tile);
#pragma unroll
            for (int k = 0; k < kChunk / 16; ++k) umma_f16(d_tmem, desc_sw128(aa + 32 * k), desc_sw128(bb + 32 * k), idesc, 1u);
          }
          umma_commit(bar_acc_full + 8 * ab);
        }
        umma_commit(bar_a_empty + 8 * buf);
      }
    }
  } else if (warp >= kEpiWarps) {
    // ===== converters: fp32 z rows -> centred fp16 A tiles =====
    const int tc_ = (int)threadIdx.x - kEpiWarps * 32;  // 0..127
    const int ch = tc_ & 7, r0 = tc_ >> 3;
    const bool fast = (p.D % 8 == 0) && ((reinterpret_cast<uintptr_t>(p.z) & 15) == 0);
    for (int i = 0; i < nblk; ++i) {
      const int buf = i % p.nbufA;
      const int64_t rb = (int64_t)blockIdx.x + (int64_t)i * gridDim.x;
      int64_t grow[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int64_t lrow = rb * kBM + r0 + 16 * u;
        grow[u] = lrow < p.N ? lrow : -1;
      }
      float zn2[8], zo2[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) zn2[u] = zo2[u] = 0.f;
      mbar_wait(bar_a_empty + 8 * buf, ((uint32_t)(i / p.nbufA) & 1) ^ 1);
      uint8_t* a_buf = sA + buf * a_bytes;
      for (int c = 0; c < p.NC; ++c) {
        const int col = c * kChunk + ch * 8;
        float4 va[8], vb[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          va[u] = make_float4(0.f, 0.f, 0.f, 0.f);
          vb[u] = va[u];
        }
        if (p.zstages > 0) {
          const uint32_t unit = (uint32_t)(i * p.NC + c);
          const uint32_t st = unit % (uint32_t)p.zstages, par = (unit / (uint32_t)p.zstages) & 1;
          mbar_wait(bar_z_full + 8 * st, par);
          const float* zt = reinterpret_cast<const float*>(sZ + st * kZStage) + ch * 8;
#pragma unroll
          for (int u = 0; u < 8; ++u)
            if (grow[u] >= 0 && col < p.D) {
              const float4* src = reinterpret_cast<const float4*>(zt + (r0 + 16 * u) * kChunk);
              va[u] = src[0];
              if (col + 4 < p.D) vb[u] = src[1];
            }
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_z_empty + 8 * st);
          if (tc_ == 0 && (int)unit + p.zstages < total_units) {
            mbar_wait(bar_z_empty + 8 * st, par);
            z_issue((int)unit + p.zstages);
          }
        } else if (fast) {
          if (col < p.D) {
#pragma unroll
            for (int u = 0; u < 8; ++u)
              if (grow[u] >= 0) {
                const float4* src = reinterpret_cast<const float4*>(p.z + grow[u] * (int64_t)p.D + col);
                va[u] = __ldg(src);
                vb[u] = __ldg(src + 1);
              }
          }
        } else {
#pragma unroll
          for (int u = 0; u < 8; ++u)
            if (grow[u] >= 0 && col < p.D) {
              const float* src = p.z + grow[u] * (int64_t)p.D + col;
              float t[8];
#pragma unroll
              for (int e = 0; e < 8; ++e) t[e] = (col + e < p.D) ? __ldg(src + e) : 0.f;
              va[u] = make_float4(t[0], t[1], t[2], t[3]);
              vb[u] = make_float4(t[4], t[5], t[6], t[7]);
            }
        }
        const float4 m0 = *reinterpret_cast<const float4*>(mu_s + col);
        const float4 m1 = *reinterpret_cast<const float4*>(mu_s + col + 4);
        const float mu8[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const float x[8] = {va[u].x, va[u].y, va[u].z, va[u].w, vb[u].x, vb[u].y, vb[u].z, vb[u].w};
          uint32_t hh[4];
          float amax = 0.f;
#pragma unroll
          for (int e = 0; e < 8; e += 2) {
            const float za = x[e] - mu8[e], zb = x[e + 1] - mu8[e + 1];
            zo2[u] = fmaf(x[e], x[e], zo2[u]);
            zo2[u] = fmaf(x[e + 1], x[e + 1], zo2[u]);
            zn2[u] = fmaf(za, za, zn2[u]);
            zn2[u] = fmaf(zb, zb, zn2[u]);
            amax = fmaxf(amax, fmaxf(fabsf(za), fabsf(zb)));
            const __half2 hp = __floats2half2_rn(za, zb);
            hh[e >> 1] = *reinterpret_cast<const uint32_t*>(&hp);
          }
          if (!(amax < 6.0e4f)) zn2[u] = INFINITY;  // does not fit fp16 (or NaN): the row takes the exact path
          const int r = r0 + 16 * u;
          const uint32_t off = (uint32_t)(r * 128 + ((ch ^ (r & 7)) << 4));
          *reinterpret_cast<uint4*>(a_buf + c * a_tile + off) = make_uint4(hh[0], hh[1], hh[2], hh[3]);
        }
      }
      float2* stat = rowstat + (i % kStatSlots) * kBM;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) {
          zn2[u] += __shfl_xor_sync(0xffffffffu, zn2[u], o);
          zo2[u] += __shfl_xor_sync(0xffffffffu, zo2[u], o);
        }
        if (ch == 0) stat[r0 + 16 * u] = make_float2(zn2[u], zo2[u]);
      }
      fence_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_a_full + 8 * buf);
    }
  } else {
    // ===== epilogue warps 0-7 =====
    const int q = warp & 3, g = warp >> 2;
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    const uint32_t tbuf = tmem_base + (uint32_t)g * (uint32_t)kBN + lane_off;
    const float emax_g = sqrtf(emax2);
    const uint32_t kmask = p.kmask;  // 0xFFFFFFE0, passed as data so that it lives in a register and the COLUMN is the LOP3 immediate

    // running state of one row (lane): best score (FOLD 0: packed key), best of the OTHER chunks, first code of the
    // winning chunk, and what is known about near-best scores inside the winning chunk
    float best = INFINITY, second = INFINITY, ma = 0.f, mb = 0.f;
    int base = 0;
    float big = 0.f, t0 = 0.f, margin_row = 0.f;
    int cur_blk = -1;
    float2 st = make_float2(0.f, 0.f);
    constexpr float kH = 1.2089258e24f;  // 2^80 (FOLD 1): (thr - s) * 2^80 saturates to exactly 0 or 1 unless |thr - s| < 2^-80
    for (int t = g; t < total_tiles; t += 2) {
      const int i = t / p.T, j = t - i * p.T;
      mbar_wait(bar_acc_full + 8 * g, (uint32_t)(t >> 1) & 1);
      fence_after();
      if (i != cur_blk) {
        // New row block.  The within-chunk margin is fixed per row BEFORE the scan, from bounds that hold for any
        // candidate (||e-mu|| <= emax, distance <= ||z-mu|| + emax); the final test uses the refined bound, which is
        // never larger.  (rowstat was written before bar_a_full completed and the MMA issuer observed that barrier
        // before issuing the MMAs whose commit we just waited for: the read is ordered.)
        cur_blk = i;
        st = rowstat[(i % kStatSlots) * kBM + q * 32 + lane];
        const float zn2 = st.x, zo2 = st.y;
        const float zn = sqrtf(zn2);
        const float dmax = zn + emax_g;
        const float sub = 2.9802322e-8f * sqrtf((float)Dp + 16.f) * (zn + 2.f * emax_g + 3.f);
        const float m_alg = 2.f * (p.gamma_en * emax_g * emax_g + p.gamma_dot * 2.f * zn * emax_g + sub);
        const float delta = 5.9604645e-8f * (sqrtf(zo2) + emax_g + 3.f * mu_norm);
        const float m_c = (p.metric == LIPVQ_METRIC_DOT) ? 0.f : 2.f * (2.f * dmax * delta + delta * delta);
        const float ref_mag = (p.metric == LIPVQ_METRIC_DOT) ? 2.f * zn * emax_g : dmax * dmax;
        const float smax = emax_g * emax_g + 2.f * zn * emax_g;  // |score| <= smax for every real code
        const float slack = (FOLD == 0) ? 4.7683716e-6f * smax : 0.f;  // key truncation (2^-18) + rounding of the count FMAs
        margin_row = (fmaxf(m_alg + m_c, p.gamma_ref * ref_mag) + 2.f * slack) * 1.0000002f + 1e-30f;
        // FOLD 0: u(s) = sat((key_best - s) * big + t0): exactly 1 for the chunk's best (its key is within `slack` of its
        // score), > 0 for every score below key_best + margin_row + slack, 0 beyond
        big = 1.f / margin_row;
        t0 = 1.f + slack * big + 1e-6f;
        if (!(margin_row < 3.0e38f) || !(big > 0.f)) {  // non-finite row statistics: nothing can be certified
          big = 0.f;
          t0 = 2.f;  // every score counts: the count is never 1
          margin_row = INFINITY;
        }
      }
      for (int c0 = 0; c0 < kBN; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(tbuf + c0, v);
        tmem_ld_wait(v);
        if (c0 + 32 >= kBN) {  // last load of this accumulator has landed: hand it back to the MMA warp
          fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_acc_empty + 8 * g);
        }
        float cm, xa, xb = 0.f;
        if (FOLD == 0) {
          // --- best of the chunk: packed keys, four chains of FMNMX3 ---
          float k0_ = fminf(pack_key(v[0], kmask, 0), pack_key(v[1], kmask, 1));
          float k1_ = fminf(pack_key(v[2], kmask, 2), pack_key(v[3], kmask, 3));
          float k2_ = fminf(pack_key(v[4], kmask, 4), pack_key(v[5], kmask, 5));
          float k3_ = fminf(pack_key(v[6], kmask, 6), pack_key(v[7], kmask, 7));
#pragma unroll
          for (int e = 8; e < 32; e += 8) {
            k0_ = fmin3(k0_, pack_key(v[e], kmask, e), pack_key(v[e + 1], kmask, e + 1));
            k1_ = fmin3(k1_, pack_key(v[e + 2], kmask, e + 2), pack_key(v[e + 3], kmask, e + 3));
            k2_ = fmin3(k2_, pack_key(v[e + 4], kmask, e + 4), pack_key(v[e + 5], kmask, e + 5));
            k3_ = fmin3(k3_, pack_key(v[e + 6], kmask, e + 6), pack_key(v[e + 7], kmask, e + 7));
          }
          cm = fmin3(k0_, k1_, fminf(k2_, k3_));
          // --- how many scores of the chunk lie within the margin of its best (FMA pipe) ---
          const float t1 = fmaf(cm, big, t0);
          const float nbig = -big;
          float n0 = ffma_sat(__uint_as_float(v[0]), nbig, t1), n1 = ffma_sat(__uint_as_float(v[1]), nbig, t1);
          float n2 = ffma_sat(__uint_as_float(v[2]), nbig, t1), n3 = ffma_sat(__uint_as_float(v[3]), nbig, t1);
#pragma unroll
          for (int e = 4; e < 32; e += 4) {
            n0 += ffma_sat(__uint_as_float(v[e]), nbig, t1);
            n1 += ffma_sat(__uint_as_float(v[e + 1]), nbig, t1);
            n2 += ffma_sat(__uint_as_float(v[e + 2]), nbig, t1);
            n3 += ffma_sat(__uint_as_float(v[e + 3]), nbig, t1);
          }
          xa = (n0 + n1) + (n2 + n3);
        } else {
          // --- best of the chunk: plain fp32 minimum (half an ALU instruction per score) ---
          float k0_ = fminf(__uint_as_float(v[0]), __uint_as_float(v[1])), k1_ = fminf(__uint_as_float(v[2]), __uint_as_float(v[3]));
          float k2_ = fminf(__uint_as_float(v[4]), __uint_as_float(v[5])), k3_ = fminf(__uint_as_float(v[6]), __uint_as_float(v[7]));
#pragma unroll
          for (int e = 8; e < 32; e += 8) {
            k0_ = fmin3(k0_, __uint_as_float(v[e]), __uint_as_float(v[e + 1]));
            k1_ = fmin3(k1_, __uint_as_float(v[e + 2]), __uint_as_float(v[e + 3]));
            k2_ = fmin3(k2_, __uint_as_float(v[e + 4]), __uint_as_float(v[e + 5]));
            k3_ = fmin3(k3_, __uint_as_float(v[e + 6]), __uint_as_float(v[e + 7]));
          }
          cm = fmin3(k0_, k1_, fminf(k2_, k3_));
          // --- WHICH scores lie below best + margin, as an exact bit mask built on the FMA pipe:
          //     u = sat((thr - s) * 2^80) is exactly 1 for s < thr and exactly 0 for s >= thr (a fractional value needs
          //     |thr - s| < 2^-80 and is caught by the integrality test below); mask += u * 2^column.
          const float tH = (cm + margin_row) * kH;
          const float nH = -kH;
          float l0 = 0.f, l1 = 0.f, h0 = 0.f, h1 = 0.f;
#pragma unroll
          for (int e = 0; e < 16; e += 2) {
            l0 = fmaf(ffma_sat(__uint_as_float(v[e]), nH, tH), (float)(1u << e), l0);
            l1 = fmaf(ffma_sat(__uint_as_float(v[e + 1]), nH, tH), (float)(1u << (e + 1)), l1);
            h0 = fmaf(ffma_sat(__uint_as_float(v[16 + e]), nH, tH), (float)(1u << e), h0);
            h1 = fmaf(ffma_sat(__uint_as_float(v[17 + e]), nH, tH), (float)(1u << (e + 1)), h1);
          }
          xa = l0 + l1;
          xb = h0 + h1;
        }
        // --- chunk-level top-2 ---
        const bool rec = cm < best;
        second = fminf(second, fmaxf(best, cm));
        base = rec ? (j * kBN + c0) : base;
        ma = rec ? xa : ma;
        mb = rec ? xb : mb;
        best = fminf(best, cm);
      }
      const bool last_of_block = (j == p.T - 1);
      const bool leaves_block = last_of_block || (j + 2 >= p.T);
      if (!leaves_block) continue;

      RowBest* xs = xchg + (i & 1) * kBM + q * 32 + lane;
      const uint32_t xpar = (uint32_t)(i >> 1) & 1;
      if (!last_of_block) {
        mbar_wait(bar_x_empty + 8 * (i & 1), xpar ^ 1);
        xs->best = best; xs->second = second; xs->ma = ma; xs->mb = mb; xs->base = base;
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_x_full + 8 * (i & 1));
        best = INFINITY; second = INFINITY; ma = 0.f; mb = 0.f; base = 0;
        continue;
      }
      if (p.T > 1) {
        mbar_wait(bar_x_full + 8 * (i & 1), xpar);
        const float ob = xs->best, os = xs->second, oa = xs->ma, obm = xs->mb;
        const int obase = xs->base;
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_x_empty + 8 * (i & 1));
        const bool take = ob < best;
        second = fminf(fminf(second, os), take ? best : ob);
        base = take ? obase : base;
        ma = take ? oa : ma;
        mb = take ? obm : mb;
        best = take ? ob : best;
      }
      const int64_t rb = (int64_t)blockIdx.x + (int64_t)i * gridDim.x;
      const int64_t row = rb * kBM + q * 32 + lane;
      if (row < p.N) {
        int bidx;
        bool single;
        if (FOLD == 0) {
          bidx = base + (int)(__float_as_uint(best) & 31u);
          single = (ma == 1.0f);
        } else {
          const int ilo = __float2int_rn(ma), ihi = __float2int_rn(mb);
          const bool exact = ((float)ilo == ma) && ((float)ihi == mb) && ilo >= 0 && ihi >= 0 && ilo < 65536 && ihi < 65536;
          const uint32_t mask = (uint32_t)ilo | ((uint32_t)ihi << 16);
          single = exact && (__popc(mask) == 1);
          bidx = base + __ffs((int)mask) - 1;
        }
        const float zn2 = st.x, zo2 = st.y;
        const float zn = sqrtf(zn2);
        const float dhat = sqrtf(fmaxf(best + zn2, 0.f));
        float emax = emax_g;
        if (p.metric != LIPVQ_METRIC_DOT) emax = fminf(emax, zn + 1.02f * dhat + 1e-30f);
        const float sub = 2.9802322e-8f * sqrtf((float)Dp + 16.f) * (zn + 2.f * emax + 3.f);
        const float margin_alg = 2.f * (p.gamma_en * emax * emax + p.gamma_dot * 2.f * zn * emax + sub);
        const float ref_mag = (p.metric == LIPVQ_METRIC_DOT) ? 2.f * zn * emax : fmaxf(best + zn2, 0.f);
        const float delta = 5.9604645e-8f * (sqrtf(zo2) + emax + 3.f * mu_norm);
        const float margin_c = (p.metric == LIPVQ_METRIC_DOT) ? 0.f : 2.f * (2.f * dhat * delta + delta * delta);
        const float smax = emax_g * emax_g + 2.f * zn * emax_g;
        const float slack = (FOLD == 0) ? 4.7683716e-6f * smax : 0.f;
        // never larger than the bound the in-chunk test used
        const float margin = fminf(fmaxf(margin_alg + margin_c, p.gamma_ref * ref_mag) + 2.f * slack, margin_row);
        bool ok = (second - best > margin) && single && (bidx >= 0) && (bidx < p.K) && (margin_row < 3.0e38f);
        if (excl_inv > 0.f) ok = ok && (p.metric != LIPVQ_METRIC_DOT) && ((zn + 1.02f * dhat) * excl_inv < 0.999f);
        if (ok) {
          p.idx_out[row] = (int64_t)bidx;
        } else {
          const int pos = atomicAdd(p.queue_count, 1);
          p.queue_out[pos] = (int)row;
        }
      }
      best = INFINITY; second = INFINITY; ma = 0.f; mb = 0.f; base = 0;
    }
  }

  fence_before();
  __syncthreads();
  grid_dep_launch();
  if (warp == kWarpMma) {
    fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* ptr = nullptr;
  cudaDriverEntryPointQueryResult qres;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) != cudaSuccess ||
      qres != cudaDriverEntryPointSuccess)
    return nullptr;
  fn = reinterpret_cast<EncodeTiledFn>(ptr);
  return fn;
}
int make_z_map(CUtensorMap* tm, const float* z, int64_t N, int D) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled entry point unavailable");
    return LIPVQ_ERR_CUDA;
  }
  const cuuint64_t gdim[2] = {(cuuint64_t)D, (cuuint64_t)N};
  const cuuint64_t gstride[1] = {(cuuint64_t)D * 4};
  const cuuint32_t box[2] = {(cuuint32_t)kChunk, (cuuint32_t)kBM};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(z), gdim, gstride, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled(z) failed with %d (N=%lld D=%d)", (int)r, (long long)N, D);
    return LIPVQ_ERR_CUDA;
  }
  return LIPVQ_OK;
}

struct P1Geom {
  int Dp, NC, Kp, T, nbufA, zstages;
  size_t smem;
  bool ok;
};

constexpr int kMisc = kStatSlots * kBM * 8 + 2 * kBM * 32 + 256 * 4 + 32 * 4 + 24 * 8 + 64;

P1Geom p1_geom(int D, int K, bool staged) {
  P1Geom g{};
  g.Dp = (D + kChunk - 1) / kChunk * kChunk;
  g.NC = g.Dp / kChunk;
  g.Kp = (K + kBN - 1) / kBN * kBN;
  g.T = g.Kp / kBN;
  g.ok = false;
  if (g.Dp > 256) return g;
  const int budget = 227 * 1024 - 1024;
  const int fixed = g.T * g.NC * kBN * kChunk * 2 + g.Kp * 32 + kBM * 32 + kMisc;
  const int a_bytes = g.NC * kBM * kChunk * 2;
  static const int opts[5][2] = {{2, 3}, {2, 2}, {1, 2}, {2, 0}, {1, 0}};  // (A buffers, z stages)
  for (int o = 0; o < 5 && !g.ok; ++o) {
    const int nbuf = opts[o][0], zs = opts[o][1];
    if (!staged && zs > 0) continue;
    const int need = fixed + nbuf * a_bytes + zs * kZStage;
    if (need > budget) continue;
    g.nbufA = nbuf;
    g.zstages = zs;
    g.smem = (size_t)need + 1024;
    g.ok = true;
  }
  return g;
}

}  // namespace

bool search_p1_supported(int64_t N, int D, int K) {
  if (N < 1 || K < 1 || D < 1) return false;
  if (const char* e = getenv("LIPVQ_SEARCH_P1"))
    if (e[0] == '0') return false;
  return p1_geom(D, K, false).ok;
}

// One launch: codebook preparation + fp16 pass over all rows.  Uncertified rows are appended to `queue` (count at
// *queue_count, which must be zero on entry); mu / enorm / ehi / elo / hdr are published for the later passes.
int search_p1_launch(const float* z, int64_t N, int D, const float* cb, int K, int metric, int Kp_pub, float* mu,
                     float* enorm, void* ehi_, void* elo_, SearchHdr* hdr, int* queue, int* queue_count, int64_t* idx_out,
                     cudaStream_t stream) {
  __half* ehi = reinterpret_cast<__half*>(ehi_);
  __half* elo = reinterpret_cast<__half*>(elo_);
  const bool staged = (D % 4 == 0) && aligned16(z);
  P1Geom g = p1_geom(D, K, staged);
  if (!g.ok) g = p1_geom(D, K, false);
  if (!g.ok) {
    set_error("search_p1: shape N=%lld D=%d K=%d does not fit", (long long)N, D, K);
    return LIPVQ_ERR_INVALID;
  }
  P1Params p{};
  p.z = z; p.cb = cb; p.N = N; p.D = D; p.K = K; p.Kp = g.Kp; p.NC = g.NC; p.T = g.T;
  p.nbufA = g.nbufA; p.zstages = g.zstages; p.metric = metric;
  p.Kp_pub = Kp_pub;
  p.kmask = 0xFFFFFFE0u;
  p.slab_swap = 0;
  if (const char* e = getenv("LIPVQ_SLAB_DESC_SWAP")) p.slab_swap = (e[0] == '1');
  p.mu_g = mu; p.enorm_g = enorm; p.ehi_g = ehi; p.elo_g = elo; p.hdr = hdr;
  p.queue_out = queue; p.queue_count = queue_count; p.idx_out = idx_out;
  const float u22 = 2.3841858e-7f;
  const float steps = (float)(g.Dp / 16 + 1);  // + the norm-slab MMA
  p.gamma_en = (float)(D + 4) * 5.9604645e-8f + (steps + 1.f) * u22;
  p.gamma_dot = 9.7751617e-4f + (steps + 1.f) * u22;
  p.gamma_ref = 4.f * (float)(D + 4) * 5.9604645e-8f;
  CUtensorMap tm_z;
  if (g.zstages > 0) {
    const int rc = make_z_map(&tm_z, z, N, D);
    if (rc != LIPVQ_OK) return rc;
  } else {
    memset(&tm_z, 0, sizeof(tm_z));
  }
  int fold = 1;  // 1: exact near-best bit mask on the FMA pipe (default); 0: packed keys + saturating count (A/B)
  if (const char* e = getenv("LIPVQ_P1_FOLD")) fold = (e[0] == '0') ? 0 : 1;
  static size_t smem_set[2][kMaxDevices] = {};
  const int dev_ = current_device();
  if (g.smem > smem_set[fold][dev_]) {
    if (fold == 0)
      LIPVQ_CUDA(cudaFuncSetAttribute(search_p1_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g.smem));
    else
      LIPVQ_CUDA(cudaFuncSetAttribute(search_p1_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g.smem));
    smem_set[fold][dev_] = g.smem;
  }
  const int row_blocks = (int)ceil_div64(N, kBM);
  const int sms = sm_count();
  const int grid = row_blocks < sms ? row_blocks : sms;
  if (fold == 0)
    LIPVQ_CUDA(launch_pdl(search_p1_kernel<0>, dim3(grid), dim3(kThreads), g.smem, stream, tm_z, p));
  else
    LIPVQ_CUDA(launch_pdl(search_p1_kernel<1>, dim3(grid), dim3(kThreads), g.smem, stream, tm_z, p));
  LIPVQ_LAUNCHED("search_p1_kernel");
  return LIPVQ_OK;
}

}  // namespace lipvq
