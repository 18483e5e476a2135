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

#ifdef LIPVQ_TRACE
// developer build: lane 0 of every warp of CTA 0 logs (clock64 << 8 | event)
constexpr int kTraceMax = 4096;
__device__ unsigned long long g_trace_p1[16][kTraceMax];
__device__ int g_trace_p1_n[16];
#define TR(ev)                                                                                             \
  do {                                                                                                     \
    if (blockIdx.x == 0 && lane == 0 && tr_n < kTraceMax)                                                  \
      g_trace_p1[warp][tr_n++] = ((unsigned long long)clock64() << 8) | (unsigned long long)(ev);          \
  } while (0)
#else
#define TR(ev) do {} while (0)
#endif

// Prepared state of one codebook (search_prep_kernel -> search_p1_kernel, "prepared" mode): everything the search needs
// that depends on the codebook only.  Lives in the search workspace, followed (at kStateHdr bytes) by the image of the
// shared-memory code tiles + norm slab.
constexpr int kStateParts = 64;
constexpr int kStateHdr = 2048;
constexpr int kStateMagic = 0x4c505631;
struct P1State {
  float mu[256];
  float emax2_part[kStateParts];     // per prep CTA: max ||e-mu||^2 over its codes
  float excl_inv_part[kStateParts];  // per prep CTA: max 1/||e-mu|| over its codes left out of the contraction (0 = none)
  float mu_norm;
  int nparts, magic, K, D, metric;
  int flagged, done;                 // per-call counters of the search kernel; the last CTA of a call clears them
  int ncand, nall;                   // statistics: candidates evaluated exactly / rows whose every code was a candidate
};
static_assert(sizeof(P1State) <= kStateHdr, "state header");

struct P1Params {
  const float* z;
  const float* cb;
  int64_t N;
  int D, K, Kp, NC, T;
  int nbufA, zstages;
  int metric;
  int Kp_pub;     // rows of the published arrays (the later passes' padded code count)
  int early_trigger;  // griddepcontrol.launch_dependents right after set-up instead of at the end
  int slab_swap;  // debug: swap LBO/SBO of the no-swizzle descriptors
  int swap_groups;  // even tiles per block: the two epilogue groups take turns at certifying (opt-in, LIPVQ_P1_SWAP=1)
  int split_last;   // two tiles per block: group 0 also drains the first columns of the second tile (opt-in, LIPVQ_P1_SPLIT=1)
  int prepared;   // the code image, mu and the code maxima come from a P1State written by search_prep_kernel
  int tail;       // rows this pass cannot certify are settled HERE (exact fp32 scan at the end of the CTA): no later launches
  int qcap;       // tail: queue_out / queue_thr are cut into per-CTA segments of qcap entries
  float* queue_thr;  // tail: per queued row, the score threshold of the candidate filter and the row's reach (2 floats)
  P1State* st;    // prepared
  int* flagged_ctr;  // tail: grid-wide count of queued rows (statistics only) and the ticket that resets both
  int* done_ctr;
  int* ncand_ctr;    // tail statistics (only when stats_out is given): exact evaluations, rows scanned in full
  int* nall_ctr;
  int32_t* stats_out;
  // outputs for the later passes (search_fused_kernel<3>, rescore): published by slices of the grid
  float* mu_g;
  float* enorm_g;
  __half* ehi_g;
  __half* elo_g;
  uint8_t* img;  // global image of the shared-memory code tiles + norm slab ([T][NC][256 x 128 B] | [Kp x 32 B])
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
__device__ __forceinline__ void bulk_load_1d(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(reinterpret_cast<uint64_t>(src)), "r"(bytes), "r"(bar)
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
// one wait for two outstanding loads; "touches" both register sets so no use can be scheduled above it
__device__ __forceinline__ void tmem_ld_wait2(uint32_t (&a)[32], uint32_t (&b)[32]) {
  tmem_ld_wait(a);
  asm volatile(""
               : "+r"(b[0]), "+r"(b[1]), "+r"(b[2]), "+r"(b[3]), "+r"(b[4]), "+r"(b[5]), "+r"(b[6]), "+r"(b[7]), "+r"(b[8]),
                 "+r"(b[9]), "+r"(b[10]), "+r"(b[11]), "+r"(b[12]), "+r"(b[13]), "+r"(b[14]), "+r"(b[15]), "+r"(b[16]),
                 "+r"(b[17]), "+r"(b[18]), "+r"(b[19]), "+r"(b[20]), "+r"(b[21]), "+r"(b[22]), "+r"(b[23]), "+r"(b[24]),
                 "+r"(b[25]), "+r"(b[26]), "+r"(b[27]), "+r"(b[28]), "+r"(b[29]), "+r"(b[30]), "+r"(b[31])
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
struct RowBest {  // partial result of one epilogue group for one row, handed to the group that certifies the row
  float best, second, ma, mb;  // ma/mb: count (FOLD 0) or the two 16-bit near-best masks (FOLD 1) of the winning chunk
  int base;
  int pad[3];
};

// Rows a CTA of search_p1_kernel could not certify are settled by the same CTA (one warp per row when there are many,
// all warps per row when there are few: p1_tail_scan / p1_tail_coop below), in two steps:
//   filter   the row's scores against ALL codes are recomputed on the CUDA cores from what is still in shared memory --
//            the fp16 code tiles -2(e-mu), the fp16 split of ||e-mu||^2 -- with the row centred in fp32.  Each score is
//            within the contraction's error bound of the true value (one operand is exact now), so every code that can be
//            the nearest -- or tie with it inside the reference's own fp32 noise -- has a score <= thr, where thr = the
//            row's best tensor-core score + 1.25 x the certification margin (written next to the row number; the margin
//            is twice the error bound of a score, which is what the argument needs).  Codes
//            that were left out of the contraction (sentinel norm) are always candidates; a non-finite thr (NaN rows,
//            unusable row statistics, a workspace that was never prepared) makes every code a candidate.
//   exact    the candidates (usually two or three) are evaluated as the stand-alone exact scan does: fp32 direct
//            differences, four interleaved partial sums in a fixed order, sqrt for L2, lowest index on exact ties
//            (torch.argmin's first-minimum rule), NaN never wins, an all-NaN row gets index 0.
// No codebook traffic: staging the fp32 codebook per CTA (128 KB from L2, the same lines for all 148 CTAs) cost 18 us for
// a few hundred rows.  Not inlined: its registers must not weigh on the fold loop of the kernel.
// score of code k for the candidate filter: ||e_k-mu||^2 (fp16 split in the norm slab) + (z-mu) . (-2(e_k-mu)) (fp16 tile
// row), fp32 accumulation in four interleaved partial sums.  Explicit shared-space loads: through generic pointers each
// 16-byte load of 32 different tile rows was handled like a scattered global access.  The tiles do not change during the
// scan, so their loads carry no ordering (volatile loads exposed one shared-memory latency per 16 bytes).
template <bool ONE>
__device__ __forceinline__ float tail_score(int k, int NC, uint32_t sB_s, uint32_t sE_s, uint32_t zc_s, const float (&zreg)[64],
                                            float& en, float& inv_x) {
  constexpr int b_tile = kBN * kChunk * 2;
  const int jt = k >> 8, r = k & 255;  // kBN = 256
  uint2 nh;
  asm("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(nh.x), "=r"(nh.y) : "r"(sE_s + slab_off(k, 0)));
  const float2 n01 = __half22float2(*reinterpret_cast<const __half2*>(&nh.x));
  const float2 n2 = __half22float2(*reinterpret_cast<const __half2*>(&nh.y));
  en = (n01.x + n01.y) + n2.x;
  inv_x = n2.y;  // excluded codes: 1/||e-mu|| (rounded up)
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
  if (ONE) {  // one 64-wide chunk (the tokenizer's own width): the centred row sits in registers
    const uint32_t trow = sB_s + (uint32_t)(jt * b_tile + r * 128);
    uint4 h[8];
#pragma unroll
    for (int q = 0; q < 8; ++q)
      asm("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(h[q].x), "=r"(h[q].y), "=r"(h[q].z), "=r"(h[q].w) : "r"(trow + ((q ^ (r & 7)) << 4)));
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const float2 f0 = __half22float2(*reinterpret_cast<const __half2*>(&h[q].x));
      const float2 f1 = __half22float2(*reinterpret_cast<const __half2*>(&h[q].y));
      const float2 f2 = __half22float2(*reinterpret_cast<const __half2*>(&h[q].z));
      const float2 f3 = __half22float2(*reinterpret_cast<const __half2*>(&h[q].w));
      a0 = fmaf(zreg[8 * q + 0], f0.x, a0); a1 = fmaf(zreg[8 * q + 1], f0.y, a1); a2 = fmaf(zreg[8 * q + 2], f1.x, a2); a3 = fmaf(zreg[8 * q + 3], f1.y, a3);
      a0 = fmaf(zreg[8 * q + 4], f2.x, a0); a1 = fmaf(zreg[8 * q + 5], f2.y, a1); a2 = fmaf(zreg[8 * q + 6], f3.x, a2); a3 = fmaf(zreg[8 * q + 7], f3.y, a3);
    }
  } else {
    for (int c = 0; c < NC; ++c) {
      const uint32_t trow = sB_s + (uint32_t)((jt * NC + c) * b_tile + r * 128);
      const uint32_t zq = zc_s + (uint32_t)(c * kChunk * 4);
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        uint4 h;
        asm("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(h.x), "=r"(h.y), "=r"(h.z), "=r"(h.w) : "r"(trow + ((q ^ (r & 7)) << 4)));
        float4 za, zb;  // (the centred row was written by this warp before a __syncwarp: ordered loads)
        lds_f4(zq + 32 * q, za);
        lds_f4(zq + 32 * q + 16, zb);
        const float2 f0 = __half22float2(*reinterpret_cast<const __half2*>(&h.x));
        const float2 f1 = __half22float2(*reinterpret_cast<const __half2*>(&h.y));
        const float2 f2 = __half22float2(*reinterpret_cast<const __half2*>(&h.z));
        const float2 f3 = __half22float2(*reinterpret_cast<const __half2*>(&h.w));
        a0 = fmaf(za.x, f0.x, a0); a1 = fmaf(za.y, f0.y, a1); a2 = fmaf(za.z, f1.x, a2); a3 = fmaf(za.w, f1.y, a3);
        a0 = fmaf(zb.x, f2.x, a0); a1 = fmaf(zb.y, f2.y, a1); a2 = fmaf(zb.z, f3.x, a2); a3 = fmaf(zb.w, f3.y, a3);
      }
    }
  }
  return en + ((a0 + a1) + (a2 + a3));
}

// exact value of code k for the row in zs (zeros past D), as the stand-alone exact scan computes it
__device__ __forceinline__ float tail_exact(int k, const float* __restrict__ cb, const float* zs, int D, int metric, bool vec) {
  const bool dotm = (metric == LIPVQ_METRIC_DOT);
  const float* er = cb + (size_t)k * D;
  float a[4] = {0.f, 0.f, 0.f, 0.f};
  if (vec && D <= 64) {  // the tokenizer's own width: the code row in two batches of eight loads in flight together
    const int nv = D >> 2;
#pragma unroll
    for (int hb = 0; hb < 2; ++hb) {
      float4 x[8];
#pragma unroll
      for (int u = 0; u < 8; ++u)
        x[u] = (8 * hb + u < nv) ? __ldg(reinterpret_cast<const float4*>(er) + 8 * hb + u) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int v = 8 * hb + u;
        if (v < nv) {
          const float4 z4 = *reinterpret_cast<const float4*>(zs + 4 * v);
          const int j = v & 3;
          if (dotm) {
            a[j] = fmaf(z4.x, x[u].x, a[j]); a[j] = fmaf(z4.y, x[u].y, a[j]); a[j] = fmaf(z4.z, x[u].z, a[j]); a[j] = fmaf(z4.w, x[u].w, a[j]);
          } else {
            float df;
            df = z4.x - x[u].x; a[j] = fmaf(df, df, a[j]);
            df = z4.y - x[u].y; a[j] = fmaf(df, df, a[j]);
            df = z4.z - x[u].z; a[j] = fmaf(df, df, a[j]);
            df = z4.w - x[u].w; a[j] = fmaf(df, df, a[j]);
          }
        }
      }
    }
  } else if (vec) {
    const int nv = D >> 2;
    for (int v0 = 0; v0 < nv; v0 += 4) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int v = v0 + j;
        if (v < nv) {
          const float4 x = __ldg(reinterpret_cast<const float4*>(er) + v);
          const float4 z4 = *reinterpret_cast<const float4*>(zs + 4 * v);
          if (dotm) {
            a[j] = fmaf(z4.x, x.x, a[j]); a[j] = fmaf(z4.y, x.y, a[j]); a[j] = fmaf(z4.z, x.z, a[j]); a[j] = fmaf(z4.w, x.w, a[j]);
          } else {
            float df;
            df = z4.x - x.x; a[j] = fmaf(df, df, a[j]);
            df = z4.y - x.y; a[j] = fmaf(df, df, a[j]);
            df = z4.z - x.z; a[j] = fmaf(df, df, a[j]);
            df = z4.w - x.w; a[j] = fmaf(df, df, a[j]);
          }
        }
      }
    }
  } else {
    for (int d0 = 0; d0 < D; d0 += 16) {  // the same partial-sum assignment: float4 v -> a[v & 3]
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int d = d0 + 4 * j + e;
          if (d < D) {
            const float zv = zs[d], ev = __ldg(er + d);
            if (dotm) {
              a[j] = fmaf(zv, ev, a[j]);
            } else {
              const float df = zv - ev;
              a[j] = fmaf(df, df, a[j]);
            }
          }
        }
    }
  }
  const float av = (a[0] + a[1]) + (a[2] + a[3]);
  return (metric == LIPVQ_METRIC_L2) ? sqrtf(av) : (dotm ? -av : av);
}

// candidate test of the filter.  sc <= thr: the score can belong to the nearest code.  Codes left out of the contraction
// (sentinel norm 3 x 65504, zero tile row) have no usable score: such a code is a candidate unless it is provably further
// than the row's best code, ||e-mu|| - ||z-mu|| > d_best, i.e. reach * (1/||e-mu||) < 1 with reach = ||z-mu|| + 1.02 d_best
// (an EMA codebook early in training holds hundreds of never-hit rows divided by ~eps: evaluating all of them exactly
// for every queued row cost 45 us per search).
__device__ __forceinline__ bool tail_is_candidate(float sc, float en, float inv_x, float thr, float reach, bool dotm) {
  if (en > 1.0e5f) return dotm || !(reach * inv_x < 0.999f);
  return (sc <= thr) || !(sc == sc);
}

constexpr int kTailWarps = kThreads / 32;      // 13: every warp of the CTA takes part
constexpr int kTailCoopRows = 8;               // up to this many queued rows a CTA works on each row with ALL its warps
constexpr int kTailList = 64;                  // candidate list entries (per warp / per row)
// per-warp scratch: centred row, row, candidate list; then the two row-level lists + counters of the cooperative form
__device__ __forceinline__ int tail_warp_floats(int Dp) { return 2 * Dp + kTailList; }

// One warp per row (many queued rows per CTA: the CTA is throughput-bound, rows are independent).
__device__ __noinline__ void p1_tail_scan(const float* __restrict__ z, const float* __restrict__ cb, const int* __restrict__ rows,
                                          const float* __restrict__ thrs, int64_t* __restrict__ idx_out, int D, int K, int Kp,
                                          int NC, int metric, const uint8_t* sB, const uint8_t* sE, const float* mu_s,
                                          float* scratch, int nq, int* ncand_ctr, int* nall_ctr) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Dp = NC * kChunk;
  int n_eval = 0, n_all = 0;
  float* zc = scratch + warp * tail_warp_floats(Dp);  // centred row, zeros past D
  float* zs = zc + Dp;                                // the row itself, zeros past D
  int* clist = reinterpret_cast<int*>(zs + Dp);       // candidates found by the filter, evaluated 32 at a time
  const uint32_t sB_s = smem_u32(sB), sE_s = smem_u32(sE), zc_s = smem_u32(zc);
  const bool vec = (D % 4 == 0) && ((reinterpret_cast<uintptr_t>(cb) & 15) == 0);
  for (int w = warp; w < nq; w += kTailWarps) {
    const int64_t row = rows[w];
    const float thr = thrs[2 * w], reach = thrs[2 * w + 1];
    const bool all = !(thr < 3.0e38f);
    const float* zr = z + row * (int64_t)D;
    for (int d = lane; d < Dp; d += 32) {
      const float v = (d < D) ? __ldg(zr + d) : 0.f;
      zs[d] = v;
      zc[d] = (d < D) ? v - mu_s[d] : 0.f;
    }
    __syncwarp();
    float zreg[64];
    if (NC == 1) {
#pragma unroll
      for (int v = 0; v < 16; ++v) {
        float4 t;
        lds_f4(zc_s + 16 * v, t);
        zreg[4 * v] = t.x; zreg[4 * v + 1] = t.y; zreg[4 * v + 2] = t.z; zreg[4 * v + 3] = t.w;
      }
    }
    float bv = INFINITY;
    int bk = 0x7fffffff;
    // exact evaluation of one candidate per lane: all lanes' code rows are in flight together (evaluating a candidate
    // where the filter finds it serialised one L2 round trip per candidate)
    auto evaluate = [&](int k) {
      const float val = tail_exact(k, cb, zs, D, metric, vec);
      if (val < bv || (val == bv && k < bk)) {  // lowest index on exact ties, whatever the order of evaluation
        bv = val;
        bk = k;
      }
    };
    n_all += all ? 1 : 0;
#ifdef LIPVQ_TAILPROF
    const long long tp0 = clock64();
#endif
    int cnt = 0;  // candidates waiting in this warp's list (warp-uniform)
    for (int k = lane; k < Kp; k += 32) {
      bool cand = all;
      if (!all) {
        float en, inv_x;
        const float sc = (NC == 1) ? tail_score<true>(k, NC, sB_s, sE_s, zc_s, zreg, en, inv_x)
                                   : tail_score<false>(k, NC, sB_s, sE_s, zc_s, zreg, en, inv_x);
        cand = tail_is_candidate(sc, en, inv_x, thr, reach, metric == LIPVQ_METRIC_DOT);
      }
      cand = cand && k < K;
      const unsigned m = __ballot_sync(0xffffffffu, cand);
      if (cand) clist[cnt + __popc(m & ((1u << lane) - 1u))] = k;
      cnt += __popc(m);
      n_eval += __popc(m);
      __syncwarp();
      if (cnt >= 32) {  // a full batch: one candidate per lane
        evaluate(clist[lane]);
        const int rest = cnt - 32;
        const int moved = (lane < rest) ? clist[32 + lane] : 0;
        __syncwarp();
        if (lane < rest) clist[lane] = moved;
        cnt = rest;
        __syncwarp();
      }
    }
#ifdef LIPVQ_TAILPROF
    const long long tp1 = clock64();
#endif
    if (lane < cnt) evaluate(clist[lane]);
#ifdef LIPVQ_TAILPROF
    const long long tp2 = clock64();
    n_eval += (int)(tp1 - tp0);  // (profiling build: stats[2] = filter cycles, stats[3] = evaluation cycles)
    n_all += (int)(tp2 - tp1);
#endif
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const int ok = __shfl_xor_sync(0xffffffffu, bk, o);
      if (ov < bv || (ov == bv && ok < bk)) {
        bv = ov;
        bk = ok;
      }
    }
    if (lane == 0) idx_out[row] = (int64_t)(bk == 0x7fffffff ? 0 : bk);
    __syncwarp();  // the row buffers are rewritten by the next trip
  }
  if (ncand_ctr != nullptr && lane == 0) {
    atomicAdd(ncand_ctr, n_eval);
    if (n_all) atomicAdd(nall_ctr, n_all);
  }
}

// A handful of queued rows per CTA (the usual case when there are any: a few hundred near-ties in a 65 536-token batch):
// with one warp per row a lone warp walks all K codes -- ~10 000 cycles of dependent work while twelve warps idle -- and
// every row pays its own chain of dependent loads (row number -> row -> candidates' code rows).  Here
//   0. the rows are fetched together into CTA-level buffers (one chain of dependent loads for all of them),
//   1. ALL warps filter every row (warp w takes codes 32 w .., 32 (w + 13) ..; the centred row is read from the shared
//      buffer into registers), candidates go to the row's list in shared memory -- no barrier between rows,
//   2. after one barrier the candidates of all rows are evaluated exactly, 32 at a time, dealt round-robin to the warps.
// Called by every thread of the CTA, nq <= kTailCoopRows.  More candidates than a list holds: every code is evaluated.
__device__ __noinline__ void p1_tail_coop(const float* __restrict__ z, const float* __restrict__ cb, const int* __restrict__ rows,
                                          const float* __restrict__ thrs, int64_t* __restrict__ idx_out, int D, int K, int Kp,
                                          int NC, int metric, const uint8_t* sB, const uint8_t* sE, const float* mu_s,
                                          float* scratch, int nq, int* ncand_ctr, int* nall_ctr) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Dp = NC * kChunk;
  float* zc_all = scratch;                          // [kTailCoopRows][Dp] centred rows, zeros past D
  float* zs_all = zc_all + kTailCoopRows * Dp;      // [kTailCoopRows][Dp] the rows themselves
  int* rlist = reinterpret_cast<int*>(zs_all + kTailCoopRows * Dp);  // [kTailCoopRows][kTailList]
  int* rcnt = rlist + kTailCoopRows * kTailList;                       // [kTailCoopRows]
  unsigned long long* rkey = reinterpret_cast<unsigned long long*>(rcnt + kTailCoopRows);  // [kTailCoopRows] (8-byte aligned)
  const uint32_t sB_s = smem_u32(sB), sE_s = smem_u32(sE);
  const bool vec = (D % 4 == 0) && ((reinterpret_cast<uintptr_t>(cb) & 15) == 0);
  if ((int)threadIdx.x < nq) {
    rcnt[threadIdx.x] = 0;
    rkey[threadIdx.x] = ~0ull;
  }
  for (int i = threadIdx.x; i < nq * Dp; i += kThreads) {
    const int w = i / Dp, d = i - w * Dp;
    const float v = (d < D) ? __ldg(z + (int64_t)rows[w] * D + d) : 0.f;
    zs_all[i] = v;
    zc_all[i] = (d < D) ? v - mu_s[d] : 0.f;
  }
  __syncthreads();
  for (int w = 0; w < nq; ++w) {
    const float thr = thrs[2 * w], reach = thrs[2 * w + 1];
    if (!(thr < 3.0e38f)) continue;  // every code is a candidate: left to the evaluation below
    const uint32_t zc_s = smem_u32(zc_all + w * Dp);
    float zreg[64];
    if (NC == 1) {
#pragma unroll
      for (int v = 0; v < 16; ++v) {
        float4 t;
        lds_f4(zc_s + 16 * v, t);
        zreg[4 * v] = t.x; zreg[4 * v + 1] = t.y; zreg[4 * v + 2] = t.z; zreg[4 * v + 3] = t.w;
      }
    }
    for (int k = warp * 32 + lane; k < Kp; k += kTailWarps * 32) {
      float en, inv_x;
      const float sc = (NC == 1) ? tail_score<true>(k, NC, sB_s, sE_s, zc_s, zreg, en, inv_x)
                                 : tail_score<false>(k, NC, sB_s, sE_s, zc_s, zreg, en, inv_x);
      if (tail_is_candidate(sc, en, inv_x, thr, reach, metric == LIPVQ_METRIC_DOT) && k < K) {
        const int pos = atomicAdd(&rcnt[w], 1);
        if (pos < kTailList) rlist[w * kTailList + pos] = k;
      }
    }
  }
  __syncthreads();
  // 2. exact evaluation, batches of 32 candidates dealt round-robin to the warps (a row whose list overflowed, or whose
  //    threshold is not finite, has every code evaluated: 16 batches at K = 512 -- one warp alone needed 37 000 cycles for
  //    such a row).  Row results meet in a 64-bit key per row: (order-preserving bits of the value, code index), so the
  //    minimum is the smallest value and, on exact ties, the lowest index.  NaN values never enter.
  {
    int item = 0;
    for (int w = 0; w < nq; ++w) {
      const bool all = !(thrs[2 * w] < 3.0e38f);
      const int cnt = rcnt[w];
      const bool full = all || cnt > kTailList;
      const int n = full ? K : cnt;
      const float* zs = zs_all + w * Dp;
      for (int b0 = 0; b0 < n; b0 += 32, ++item) {
        if (item % kTailWarps != warp) continue;
        const int e = b0 + lane;
        const int k = (e < n) ? (full ? e : rlist[w * kTailList + e]) : -1;
        unsigned long long key = ~0ull;
        if (k >= 0) {
          const float val = tail_exact(k, cb, zs, D, metric, vec) + 0.f;  // (-0 -> +0)
          if (val == val) {
            const uint32_t u = __float_as_uint(val);
            key = ((unsigned long long)((u & 0x80000000u) ? ~u : (u | 0x80000000u)) << 32) | (uint32_t)k;
          }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
          key = other < key ? other : key;
        }
        if (lane == 0 && key != ~0ull) atomicMin(&rkey[w], key);
      }
    }
  }
  __syncthreads();
  if ((int)threadIdx.x < nq) {
    const int w = threadIdx.x;
    const unsigned long long key = rkey[w];
    idx_out[rows[w]] = (key == ~0ull) ? 0 : (int64_t)(uint32_t)(key & 0xffffffffu);  // all-NaN row: index 0 (as the exact-scan kernels)
    if (ncand_ctr != nullptr) {
      const bool all = !(thrs[2 * w] < 3.0e38f);
      const int cnt = rcnt[w];
      atomicAdd(ncand_ctr, (all || cnt > kTailList) ? K : cnt);
      if (all) atomicAdd(nall_ctr, 1);
    }
  }
}

// Warp roles (416 threads, one CTA per SM):
//   0-7   epilogue: warp w owns TMEM lanes 32*(w&3).. (= rows); group w>>2 drains every second code tile
//   8-11  converters: fp32 z rows -> centred fp16 -> swizzled A tiles (+ per-row ||z-mu||^2, ||z||^2)
//   12    TMEM allocation + single-thread tcgen05.mma issue
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
  float* rowstat = reinterpret_cast<float*>(sA1 + kBM * 32);               // [kStatSlots][128]: ||z-mu||^2 per row
  RowBest* xchg = reinterpret_cast<RowBest*>(rowstat + kStatSlots * kBM);  // [2][128]
  float* mu_s = reinterpret_cast<float*>(xchg + 2 * kBM);                  // [256]
  float* red_s = mu_s + 256;                                               // [32] block reductions
  uint64_t* bars = reinterpret_cast<uint64_t*>(red_s + 32);
  const uint32_t bar_a_full = smem_u32(bars + 0), bar_a_empty = smem_u32(bars + 2);
  const uint32_t bar_acc_full = smem_u32(bars + 4), bar_acc_empty = smem_u32(bars + 6);
  // "accumulator ab holds a tile for group g": barrier ab of group 0 at bars + 4 + ab, of group 1 at bars + 24 + ab.  One
  // barrier per (buffer, draining group): every phase of a barrier is then waited for by the same group, in order (a
  // parity wait only tells phase n from phase n - 1, and with the swapped assignment below a group skips tiles of a buffer)
  const uint32_t bar_acc_full1 = smem_u32(bars + 24);
  const uint32_t bar_z_full = smem_u32(bars + 8), bar_z_empty = smem_u32(bars + 12);
  const uint32_t bar_x_full = smem_u32(bars + 16), bar_x_empty = smem_u32(bars + 18);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 20);
  const uint32_t bar_img = smem_u32(bars + 21), bar_mu = smem_u32(bars + 22);
  int* s_qn = reinterpret_cast<int*>(bars + 23);  // tail: rows this CTA has queued for its exact scan

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#ifdef LIPVQ_TRACE
  int tr_n = 0;
#endif
  TR(0);
  const int row_blocks = (int)((p.N + kBM - 1) / kBM);
  const int nblk = ((int)blockIdx.x < row_blocks) ? (row_blocks - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  const int total_tiles = nblk * p.T;
  const int Dp = p.NC * kChunk;

  if (threadIdx.x == 0) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_a_full + 8 * s, kCvtWarps);
      mbar_init(bar_a_empty + 8 * s, 1);
      mbar_init(bar_acc_full + 8 * s, 1);
      mbar_init(bar_acc_full1 + 8 * s, 1);
      // (split mode: both groups drain columns of the second tile, which always sits in buffer 1)
      mbar_init(bar_acc_empty + 8 * s, (s == 1 && p.T == 2 && p.split_last) ? kEpiWarps : kEpiWarps / 2);
      mbar_init(bar_x_full + 8 * s, kEpiWarps / 2);
      mbar_init(bar_x_empty + 8 * s, kEpiWarps / 2);
    }
    mbar_init(bar_img, 1);
    mbar_init(bar_mu, 1);
    *s_qn = 0;
    for (int s = 0; s < 4; ++s) {
      mbar_init(bar_z_full + 8 * s, 1);
      mbar_init(bar_z_empty + 8 * s, kCvtWarps);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kWarpMma) tmem_alloc(smem_u32(tmem_slot), 512);
  grid_dep_wait();  // z (and, in a training step, the codebook) are the previous kernels' outputs
  fence_before();
  __syncthreads();  // barriers initialised, TMEM allocated
  TR(6);

  fence_after();
  const uint32_t tmem_base = *tmem_slot;
  TR(1);
  // The later passes (almost always empty) may be scheduled as soon as SMs free up: their launch latency then hides
  // behind this kernel's tail.  They cannot co-reside with it (shared memory), and they wait for this grid's completion
  // (griddepcontrol.wait) before reading anything it wrote.
  if (p.early_trigger) grid_dep_launch();

  if (warp == kWarpMma) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0) {
      const uint32_t idesc = idesc_f16(kBM, kBN);
      const uint32_t lbo = p.slab_swap ? 256u : 128u, sbo = p.slab_swap ? 128u : 256u;
      const uint64_t a1desc = desc_noswz(smem_u32(sA1), lbo, sbo);
      uint32_t t = 0;
      if (p.prepared) {
        // the prepared image is copied by this thread (idle until it lands): twenty bulk copies issued from an epilogue
        // thread held up that role's prologue barrier for 3 000 cycles
        const size_t img_tiles_bytes = (size_t)p.T * p.NC * b_tile;
        // 4 KB pieces, every CTA starting at a different piece: all CTAs copy the SAME 80 KB, and 148 of them asking for
        // the same lines at the same moment queue up at the L2 slices that own those lines (measured ~3 TB/s, 4 us)
        const uint32_t slab_bytes = (uint32_t)p.Kp * 32u;
        mbar_expect_tx(bar_img, (uint32_t)img_tiles_bytes + slab_bytes);
        const uint32_t n_tile = (uint32_t)(img_tiles_bytes >> 12), n_all = n_tile + (slab_bytes >> 12);
        uint32_t j = ((uint32_t)blockIdx.x * 7u) % n_all;
  #pragma unroll 1
        for (uint32_t i = 0; i < n_all; ++i) {
          const uint32_t off = j << 12;
          const uint32_t dst = (j < n_tile) ? smem_u32(sB) + off : smem_u32(sE) + (off - (uint32_t)img_tiles_bytes);
          bulk_load_1d(dst, p.img + off, 4096u, bar_img);
          if (++j == n_all) j = 0;
        }
      }
      mbar_wait(bar_img, 0);  // code tiles + norm slab (bulk copies) are in place (the ones slab: ordered through bar_mu -> bar_a_full)
      fence_after();
      const uint32_t mma_swap = (p.T % 2 == 0 && p.swap_groups) ? 1u : 0u;
      const bool mma_split = (p.T == 2 && p.split_last);
      for (int i = 0; i < nblk; ++i) {
        const int buf = i % p.nbufA;
        mbar_wait(bar_a_full + 8 * buf, (uint32_t)(i / p.nbufA) & 1);
        fence_after();
        TR(30);
        const uint32_t a_base = smem_u32(sA + buf * a_bytes);
        for (int j = 0; j < p.T; ++j, ++t) {
          const uint32_t ab = t & 1;
          mbar_wait(bar_acc_empty + 8 * ab, ((t >> 1) & 1) ^ 1);  // the epilogue has drained the tile before last
          fence_after();
          TR(31);
          const uint32_t d_tmem = tmem_base + ab * (uint32_t)kBN;
          // score = ||e-mu||^2 (norm slab, overwrites the accumulator) - 2 (z-mu).(e-mu)
          umma_f16(d_tmem, a1desc, desc_noswz(smem_u32(sE + (size_t)j * kBN * 32), lbo, sbo), idesc, 0u);
          for (int c = 0; c < p.NC; ++c) {
            const uint32_t aa = a_base + (uint32_t)(c * a_tile);
            const uint32_t bb = smem_u32(sB + (size_t)(j * p.NC + c) * b_tile);
#pragma unroll
            for (int k = 0; k < kChunk / 16; ++k) umma_f16(d_tmem, desc_sw128(aa + 32 * k), desc_sw128(bb + 32 * k), idesc, 1u);
          }
          if (mma_split && j == 1) {  // both groups drain columns of this tile
            umma_commit(bar_acc_full + 8 * ab);
            umma_commit(bar_acc_full1 + 8 * ab);
          } else {
            umma_commit((((t + mma_swap * (uint32_t)i) & 1u) ? bar_acc_full1 : bar_acc_full) + 8 * ab);  // (the draining group's)
          }
          TR(32);
        }
        umma_commit(bar_a_empty + 8 * buf);
      }
    }
  } else if (warp >= kEpiWarps) {
    // ===== converters: fp32 z rows -> centred fp16 A tiles =====
    const int tc_ = (int)threadIdx.x - kEpiWarps * 32;  // 0..127
    const int ch = tc_ & 7, r0 = tc_ >> 3;
    const bool fast = (p.D % 8 == 0) && ((reinterpret_cast<uintptr_t>(p.z) & 15) == 0);
    // staged z rows: the first tiles start flowing now and land while the codebook is being prepared
    const int total_units = nblk * p.NC;
    auto z_issue = [&](int unit) {
      const int i2 = unit / p.NC, c2 = unit - i2 * p.NC;
      const uint32_t st = (uint32_t)unit % (uint32_t)p.zstages;
      const int64_t l0 = ((int64_t)blockIdx.x + (int64_t)i2 * gridDim.x) * kBM;
      mbar_expect_tx(bar_z_full + 8 * st, (uint32_t)kZStage);
      tma_load_2d(smem_u32(sZ + st * kZStage), &tm_z, bar_z_full + 8 * st, c2 * kChunk, (int)l0);
    };
    if (p.zstages > 0 && tc_ == 0) {
  #pragma unroll 1
      for (int u = 0; u < p.zstages && u < total_units; ++u) z_issue(u);
    }

    mbar_wait(bar_mu, 0);  // mu is ready (the codebook may still be in preparation: the converters do not need it)
    for (int i = 0; i < nblk; ++i) {
      const int buf = i % p.nbufA;
      const int64_t rb = (int64_t)blockIdx.x + (int64_t)i * gridDim.x;
      int64_t grow[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int64_t lrow = rb * kBM + r0 + 16 * u;
        grow[u] = lrow < p.N ? lrow : -1;
      }
      float zn2[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) zn2[u] = 0.f;
      TR(20);
      mbar_wait(bar_a_empty + 8 * buf, ((uint32_t)(i / p.nbufA) & 1) ^ 1);
      TR(21);
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
          TR(22);
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
#pragma unroll
          for (int e = 0; e < 8; e += 2) {
            const float za = x[e] - mu8[e], zb = x[e + 1] - mu8[e + 1];
            zn2[u] = fmaf(za, za, zn2[u]);
            zn2[u] = fmaf(zb, zb, zn2[u]);
            const __half2 hp = __floats2half2_rn(za, zb);
            hh[e >> 1] = *reinterpret_cast<const uint32_t*>(&hp);
          }
          const int r = r0 + 16 * u;
          const uint32_t off = (uint32_t)(r * 128 + ((ch ^ (r & 7)) << 4));
          *reinterpret_cast<uint4*>(a_buf + c * a_tile + off) = make_uint4(hh[0], hh[1], hh[2], hh[3]);
        }
      }
      TR(24);
      // ||z-mu||^2 per row.  It also guards the fp16 range: every centred entry is at most sqrt(zn2) in magnitude, so a
      // row with zn2 < 3.6e9 fits; anything else (huge, inf, NaN) is marked uncertifiable and takes the exact path.
      float* stat = rowstat + (i % kStatSlots) * kBM;
      {
        // 8 partial sums (rows r0 + 16u) in each of the 8 lanes of a row group -> one total per lane, 7 shuffles
        // (transposing butterfly: every step halves the values a lane still carries)
        const bool b0 = ch & 1, b1 = ch & 2, b2 = ch & 4;
        float k4[4];
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          const float send = b0 ? zn2[jj] : zn2[4 + jj];
          const float recv = __shfl_xor_sync(0xffffffffu, send, 1);
          k4[jj] = (b0 ? zn2[4 + jj] : zn2[jj]) + recv;
        }
        float k2[2];
#pragma unroll
        for (int jj = 0; jj < 2; ++jj) {
          const float send = b1 ? k4[jj] : k4[2 + jj];
          const float recv = __shfl_xor_sync(0xffffffffu, send, 2);
          k2[jj] = (b1 ? k4[2 + jj] : k4[jj]) + recv;
        }
        const float send = b2 ? k2[0] : k2[1];
        const float tot = (b2 ? k2[1] : k2[0]) + __shfl_xor_sync(0xffffffffu, send, 4);
        const int u = (b0 ? 4 : 0) + (b1 ? 2 : 0) + (b2 ? 1 : 0);
        stat[r0 + 16 * u] = (tot < 3.6e9f) ? tot : INFINITY;
      }
      TR(25);
      fence_async_smem();
      TR(26);
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_a_full + 8 * buf);
      TR(23);
    }
  } else {
    // ===== epilogue warps 0-7: first the cooperative codebook preparation, then the fold =====
    // ---- prologue: the codebook is prepared ONCE per launch, cooperatively ------------------------------------------------
    // Converting the whole codebook in every CTA cost 15 000 cycles (148 x redundant work, measured).  Instead CTA b
    // converts rows [b*R, (b+1)*R) (3-4 rows at K = 512) into a global IMAGE of the shared-memory layout (swizzled fp16
    // tiles + norm slab), the CTAs meet at an arrive counter, and everyone bulk-copies the finished image (80 KB from L2).
    // All CTAs are co-resident (grid <= number of SMs, one CTA per SM), so the spin cannot deadlock; it is bounded anyway.
    // mu (mean of <= 32 sampled z rows) is needed before the conversion and by the z converters: every CTA derives it
    // itself, with the sample loads and the slice's code rows requested together (one round trip).
    // item = (code k, 16-byte output chunk of its row): 8 latent dims = two float4 loads; the items of one row sit in
    // consecutive lanes (Dp/8 = 8, 16 or 32 of them), so ||e-mu||^2 is a fixed-order shuffle reduction.
    const size_t img_tiles_bytes = (size_t)p.T * p.NC * b_tile;
    float emax2 = 0.f, excl_inv = 0.f, mu_norm = 0.f;
    if (p.prepared) {
      // ---- prepared codebook (search_prep_kernel ran earlier in the stream, once per codebook version): the image is
      // bulk-copied straight away (by the MMA thread), mu and the code maxima are read from the state header -- no
      // grid-wide meeting
      const P1State* st = p.st;
      mu_s[threadIdx.x] = __ldg(st->mu + threadIdx.x);  // 256 epilogue threads, 256 entries (zeros past D)
      const int np = __ldg(&st->nparts);
      const bool valid = __ldg(&st->magic) == kStateMagic && __ldg(&st->K) == p.K && __ldg(&st->D) == p.D &&
                         __ldg(&st->metric) == p.metric && np >= 1 && np <= kStateParts;
      // (all kStateParts entries: the preparation zero-fills the unused ones, so no load waits for nparts -- with a cold
      //  L2 every dependent load here is a DRAM round trip in front of the first MMA)
      float e2 = fmaxf(__ldg(st->emax2_part + lane), __ldg(st->emax2_part + lane + 32));
      float xi = fmaxf(__ldg(st->excl_inv_part + lane), __ldg(st->excl_inv_part + lane + 32));
      mu_norm = __ldg(&st->mu_norm);
      for (int r = threadIdx.x; r < kBM; r += kEpiWarps * 32) {  // ones slab: columns 0..2 = 1.0 for every row
        *reinterpret_cast<uint4*>(sA1 + slab_off(r, 0)) = make_uint4(0x3C003C00u, 0x00003C00u, 0u, 0u);
        *reinterpret_cast<uint4*>(sA1 + slab_off(r, 8)) = make_uint4(0u, 0u, 0u, 0u);
      }
      fence_async_smem();  // (the ones slab, for the tensor core)
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (threadIdx.x == 0) mbar_arrive(bar_mu);  // the z converters may start
      TR(43);
  #pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        e2 = fmaxf(e2, __shfl_xor_sync(0xffffffffu, e2, o));
        xi = fmaxf(xi, __shfl_xor_sync(0xffffffffu, xi, o));
      }
      // a workspace that was never prepared for this (K, D, metric): nothing is certified, every row takes the exact scan
      emax2 = valid ? e2 : INFINITY;
      excl_inv = xi;
      TR(46);
      mbar_wait(bar_img, 0);
      TR(47);
    } else {
    const int ipr = Dp / 8;  // items per row: 8 (D <= 64), 16, 24 (-> handled as 32 with idle lanes), 32
    const int ipr_p2 = ipr <= 8 ? 8 : (ipr <= 16 ? 16 : 32);
    const int rpp = (kEpiWarps * 32) / ipr_p2;  // whole rows per pass over the 256 epilogue threads
    const int sub = (int)threadIdx.x % ipr_p2, rslot = (int)threadIdx.x / ipr_p2;
    const int d0 = sub * 8;
    const bool col_live = sub < ipr && d0 < p.D;
    const bool vec = (p.D % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.cb) & 15) == 0) && (d0 + 8 <= p.D);
    const int slice_n = (p.Kp + (int)gridDim.x - 1) / (int)gridDim.x;
    const int slice_lo = (int)blockIdx.x * slice_n, slice_hi = min(slice_lo + slice_n, p.Kp);
    auto load_item = [&](int k, float (&x)[8]) {
  #pragma unroll
      for (int e = 0; e < 8; ++e) x[e] = 0.f;
      if (k < slice_hi && k < p.K && col_live) {
        const float* src = p.cb + (size_t)k * p.D + d0;
        if (vec) {
          const float4 a = __ldg(reinterpret_cast<const float4*>(src)), b = __ldg(reinterpret_cast<const float4*>(src) + 1);
          x[0] = a.x; x[1] = a.y; x[2] = a.z; x[3] = a.w; x[4] = b.x; x[5] = b.y; x[6] = b.z; x[7] = b.w;
        } else {
  #pragma unroll
          for (int e = 0; e < 8; ++e)
            if (d0 + e < p.D) x[e] = __ldg(src + e);
        }
      }
    };
    float xfirst[8];
    {
      // mu[d] = mean over R <= 32 evenly spaced rows.  The R x D samples are spread over the threads (a few loads each,
      // all in flight together), parked in shared memory (the A buffers are idle until the converters start) and summed
      // per dimension in row order -- the same value in every CTA, which the slice owners and the z converters rely on.
      // (Straight-line code is fetched at ~8 cycles per instruction the first time: this is a loop, not 32 unrolled loads.)
      constexpr int RMAX = 32;
      const int R = (int)(p.N < RMAX ? p.N : RMAX);
      const int64_t stride = p.N / R;
      float* samp = reinterpret_cast<float*>(sA);  // [R][256]
      const bool want_mu = p.metric != LIPVQ_METRIC_DOT;
      if (want_mu) {
        const int dl = (int)threadIdx.x & 63, il = (int)threadIdx.x >> 6;  // 64 dims x 4 row lanes
        for (int db = 0; db < p.D; db += 64) {
          const int d = db + dl;
          if (d < p.D) {
            float v[8];
  #pragma unroll
            for (int n = 0; n < 8; ++n) {
              const int i = il + 4 * n;
              v[n] = (i < R) ? __ldg(p.z + (int64_t)i * stride * p.D + d) : 0.f;
            }
  #pragma unroll
            for (int n = 0; n < 8; ++n) {
              const int i = il + 4 * n;
              samp[i * 256 + d] = v[n];
            }
          }
        }
      }
      TR(40);
      load_item(slice_lo + rslot, xfirst);  // in flight together with the mu samples
      TR(41);
      // ones slab: columns 0..2 = 1.0 for every row
      for (int r = threadIdx.x; r < kBM; r += kEpiWarps * 32) {
        *reinterpret_cast<uint4*>(sA1 + slab_off(r, 0)) = make_uint4(0x3C003C00u, 0x00003C00u, 0u, 0u);  // 1, 1, 1, 0 ...
        *reinterpret_cast<uint4*>(sA1 + slab_off(r, 8)) = make_uint4(0u, 0u, 0u, 0u);
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      {
        const int d = (int)threadIdx.x;
        float sum = 0.f;
        if (want_mu && d < p.D) {
  #pragma unroll 8
          for (int i = 0; i < RMAX; ++i) sum += samp[i * 256 + d];  // rows >= R hold zeros; fixed order
        }
        mu_s[d] = sum / (float)R;
      }
      TR(42);
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (threadIdx.x == 0) mbar_arrive(bar_mu);  // the z converters may start: they need mu only
      TR(43);
      if (warp == 0) {  // ||mu||
        float t = 0.f;
        for (int dd = lane; dd < 256; dd += 32) t = fmaf(mu_s[dd], mu_s[dd], t);
        t = warp_sum(t);
        if (lane == 0) red_s[31] = t;
      }
    }
    TR(4);
    {
      float mu8[8];
  #pragma unroll
      for (int e = 0; e < 8; ++e) mu8[e] = (col_live && d0 + e < p.D) ? mu_s[d0 + e] : 0.f;
      const int ch = sub & 7;
      const bool dot = (p.metric == LIPVQ_METRIC_DOT);
      const __half2 m2 = __floats2half2_rn(-2.f, -2.f);
      float emax2_loc = 0.f, excl_inv_loc = 0.f;
      for (int k0 = slice_lo; k0 < slice_hi; k0 += rpp) {
        const int k = k0 + rslot;
        float x[8];
        if (k0 == slice_lo) {
  #pragma unroll
          for (int e = 0; e < 8; ++e) x[e] = xfirst[e];
        } else {
          load_item(k, x);
        }
        float c[8];
        float s = 0.f, am = 0.f;
  #pragma unroll
        for (int e = 0; e < 8; ++e) {
          c[e] = x[e] - mu8[e];
          s = fmaf(c[e], c[e], s);
        }
  #pragma unroll
        for (int e = 0; e < 8; e += 2) am = fmaxf(am, fmaxf(fabsf(c[e]), fabsf(c[e + 1])));
        for (int o = 1; o < ipr_p2; o <<= 1) {
          s += __shfl_xor_sync(0xffffffffu, s, o);
          am = fmaxf(am, __shfl_xor_sync(0xffffffffu, am, o));
        }
        if (k >= slice_hi || sub >= ipr) continue;  // (after the shuffles: every lane of the warp takes part in them)
        // a code whose centred entries or norm do not fit fp16 (or that holds inf / NaN) is left out of the contraction
        const bool excluded = k < p.K && (!(am < kF16Safe) || (dot ? !(s < 3.0e38f) : !(s <= kEnormMax)));
        const bool zero_row = k >= p.K || excluded;
        uint32_t hh[4], ll[4];
  #pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float va = zero_row ? 0.f : -2.f * c[2 * e], vb = zero_row ? 0.f : -2.f * c[2 * e + 1];
          const __half2 hp = __floats2half2_rn(va, vb);
          const float2 back = __half22float2(hp);
          const __half2 lp = __floats2half2_rn(va - back.x, vb - back.y);
          hh[e] = *reinterpret_cast<const uint32_t*>(&hp);
          ll[e] = *reinterpret_cast<const uint32_t*>(&lp);
        }
        (void)m2;
        const uint4 hv = make_uint4(hh[0], hh[1], hh[2], hh[3]);
        const int j = k >> 8, r = k & 255;  // kBN = 256
        // image of the shared-memory tile layout: [j][c] tiles of [256 x 64] fp16, 128-byte swizzle
        *reinterpret_cast<uint4*>(p.img + ((size_t)j * p.NC + (sub >> 3)) * b_tile + r * 128 + ((ch ^ (r & 7)) << 4)) = hv;
        if (k < p.Kp_pub) {  // linear copies for the 3-term pass (TMA applies the swizzle itself)
          *reinterpret_cast<uint4*>(p.ehi_g + (size_t)k * Dp + d0) = hv;
          *reinterpret_cast<uint4*>(p.elo_g + (size_t)k * Dp + d0) = make_uint4(ll[0], ll[1], ll[2], ll[3]);
        }
        if (sub == 0) {
          const float en = dot ? 0.f : s;
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
            const float inv = (s < 3.0e38f) ? rsqrtf(s) * 1.0001f : 3.0e38f;  // 1/||e-mu|| rounded up (NaN/inf: unusable)
            excl_inv_loc = fmaxf(excl_inv_loc, inv);
          }
          // column 3 (times a zero of the ones slab: no effect on the contraction) tells the in-kernel scan how far an
          // EXCLUDED code is: 1/||e-mu|| rounded up, finite (0 for ordinary codes and padding)
          const float inv_x = excluded ? fminf((s < 3.0e38f) ? rsqrtf(s) * 1.0001f : 65504.f, 65504.f) : 0.f;
          __half2 b = __floats2half2_rn(h2, 0.f);
          b.y = __float2half_ru(inv_x);
          const __half2 a = __floats2half2_rn(h0, h1);
          uint8_t* slab = p.img + img_tiles_bytes;
          *reinterpret_cast<uint4*>(slab + slab_off(k, 0)) =
              make_uint4(*reinterpret_cast<const uint32_t*>(&a), *reinterpret_cast<const uint32_t*>(&b), 0u, 0u);
          *reinterpret_cast<uint4*>(slab + slab_off(k, 8)) = make_uint4(0u, 0u, 0u, 0u);
          if (k < p.Kp_pub) p.enorm_g[k] = zero_row ? INFINITY : en;
        }
      }
      // maxima over all codes: non-negative floats order like their bit patterns (header zeroed by the host-side memset)
      if (emax2_loc > 0.f) atomicMax(&p.hdr->emax2_bits, __float_as_uint(emax2_loc));
      if (excl_inv_loc > 0.f) atomicMax(&p.hdr->excl_inv_bits, __float_as_uint(excl_inv_loc));
    }
    TR(5);
    fence_async_smem();  // (the ones slab, for the tensor core)
    asm volatile("bar.sync 1, 256;" ::: "memory");  // this CTA's slice is written
    TR(44);
    if (threadIdx.x == 0) {
      // arrive (release: the slice and the two maxima), then wait until every CTA has arrived (acquire)
      __threadfence();
      atomicAdd(&p.hdr->arrived, 1);
      const int want = (int)gridDim.x;
      long long t0 = 0;
      uint32_t spins = 0;
      for (;;) {
        int seen;
        asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(seen) : "l"(&p.hdr->arrived) : "memory");
        if (seen >= want) break;
        if ((++spins & 63u) == 0) {  // a launch that is not fully resident must fault, not hang
          if (t0 == 0) t0 = clock64();
          else if (clock64() - t0 > 4000000000LL) __trap();
        }
        __nanosleep(64);
      }
      TR(45);
      // the image was written through the generic proxy (by other SMs); the bulk copies read it through the async proxy
      asm volatile("fence.proxy.async;" ::: "memory");
      const uint32_t slab_bytes = (uint32_t)p.Kp * 32u;
      mbar_expect_tx(bar_img, (uint32_t)img_tiles_bytes + slab_bytes);
  #pragma unroll 1
      for (size_t off = 0; off < img_tiles_bytes; off += 32768)
        bulk_load_1d(smem_u32(sB + off), p.img + off, 32768u, bar_img);
  #pragma unroll 1
      for (uint32_t off = 0; off < slab_bytes; off += 8192)
        bulk_load_1d(smem_u32(sE + off), p.img + img_tiles_bytes + off, 8192u, bar_img);
    }
    TR(46);
    mbar_wait(bar_img, 0);  // every thread: code tiles + norm slab have landed (async-proxy writes, visible to the MMA)
    TR(47);
    {
      unsigned a, b;
      asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(a) : "l"(&p.hdr->emax2_bits) : "memory");
      asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(b) : "l"(&p.hdr->excl_inv_bits) : "memory");
      emax2 = __uint_as_float(a);
      excl_inv = __uint_as_float(b);
    }
    mu_norm = sqrtf(red_s[31]);
    if (blockIdx.x == 0) {  // what the later passes read
      for (int d = threadIdx.x; d < Dp; d += kEpiWarps * 32) p.mu_g[d] = mu_s[d];
      if (threadIdx.x == 0) p.hdr->mu_norm = mu_norm;
    }
    }  // !prepared

    // TMEM lane = row; group g = warp >> 2 drains every second code tile.  Per 32-column chunk and row:
    //   best of the chunk      plain fp32 minimum, half an ALU instruction per score (FMNMX3)
    //   near-best bit mask     u = sat((best + margin - s) * 2^80) is exactly 1 for s < best + margin and exactly 0 for
    //                          s >= best + margin (a fractional value needs |thr - s| < 2^-80 and is caught by the
    //                          integrality test at the end); mask += u * 2^column -- two FMA-pipe instructions per score,
    //                          on the pipe the minimum leaves idle
    //   chunk-level top-2      a handful of instructions per 32 scores: best, best of the OTHER chunks, and the mask of
    //                          the winning chunk
    // A row is certified when the winning chunk's mask has exactly one bit (that bit is the index) and every other
    // chunk's best is further away than the margin.  Two chunks are folded per iteration: the dependent chains of one
    // chunk (min tree, mask accumulation) are latency- not throughput-bound, the second chunk fills the gaps.
    const int q = warp & 3, g = warp >> 2;
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    const uint32_t tbuf0 = tmem_base + lane_off;  // accumulator buffer of tile t: t & 1
    const float emax_g = sqrtf(emax2);
    const bool dotm = (p.metric == LIPVQ_METRIC_DOT);
    // margin(zn) = max(m_alg + m_c, gamma_ref * ref_mag) with every term bounded for ANY candidate code
    // (||e-mu|| <= emax, distance <= zn + emax, ||z|| <= zn + ||mu||): a polynomial in zn = ||z-mu|| whose coefficients
    // are per-launch constants
    const float c_sub = 2.9802322e-8f * sqrtf((float)Dp + 16.f);
    const float u32f = 5.9604645e-8f;
    auto margin_of = [&](float zn, float emax, float dist) {
      const float m_alg = 2.f * (p.gamma_en * emax * emax + p.gamma_dot * 2.f * zn * emax + c_sub * (zn + 2.f * emax + 3.f));
      const float delta = u32f * (zn + emax + 4.f * mu_norm);
      const float m_c = dotm ? 0.f : 2.f * (2.f * dist * delta + delta * delta);
      const float ref_mag = dotm ? 2.f * zn * emax : dist * dist;
      return fmaxf(m_alg + m_c, p.gamma_ref * ref_mag);
    };

    float best = INFINITY, second = INFINITY, ma = 0.f, mb = 0.f;  // best, best of the other chunks, masks of the winning chunk
    int base = 0;                                                  // first code of the winning chunk
    float margin_row = 0.f, zn = 0.f;
    int cur_blk = -1;
    constexpr float kH = 1.2089258e24f;  // 2^80
    const float nH = -kH;

    // fold one 32-column chunk held in v: chunk minimum + near-best masks (low / high 16 columns)
    auto fold32 = [&](const uint32_t (&v)[32], float& cm, float& xa, float& xb) {
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
      const float tH = (cm + margin_row) * kH;
      float l0 = 0.f, l1 = 0.f, l2 = 0.f, l3 = 0.f, h0 = 0.f, h1 = 0.f, h2 = 0.f, h3 = 0.f;
#pragma unroll
      for (int e = 0; e < 16; e += 4) {
        l0 = fmaf(ffma_sat(__uint_as_float(v[e]), nH, tH), (float)(1u << e), l0);
        l1 = fmaf(ffma_sat(__uint_as_float(v[e + 1]), nH, tH), (float)(1u << (e + 1)), l1);
        l2 = fmaf(ffma_sat(__uint_as_float(v[e + 2]), nH, tH), (float)(1u << (e + 2)), l2);
        l3 = fmaf(ffma_sat(__uint_as_float(v[e + 3]), nH, tH), (float)(1u << (e + 3)), l3);
        h0 = fmaf(ffma_sat(__uint_as_float(v[16 + e]), nH, tH), (float)(1u << e), h0);
        h1 = fmaf(ffma_sat(__uint_as_float(v[17 + e]), nH, tH), (float)(1u << (e + 1)), h1);
        h2 = fmaf(ffma_sat(__uint_as_float(v[18 + e]), nH, tH), (float)(1u << (e + 2)), h2);
        h3 = fmaf(ffma_sat(__uint_as_float(v[19 + e]), nH, tH), (float)(1u << (e + 3)), h3);
      }
      xa = (l0 + l1) + (l2 + l3);
      xb = (h0 + h1) + (h2 + h3);
    };
    auto merge_chunk = [&](float cm, float xa, float xb, int first_code) {
      const bool rec = cm < best;
      second = fminf(second, fmaxf(best, cm));
      base = rec ? first_code : base;
      ma = rec ? xa : ma;
      mb = rec ? xb : mb;
      best = fminf(best, cm);
    };
    // Which group drains tile t.  The group that drains a block's LAST tile also merges the other group's partial and
    // certifies the block's rows (~1 300 cycles per block on top of the fold).  With an even number of tiles per block a
    // fixed assignment t & 1 gives that work to group 1 every time (its period was fold + certification = the kernel's
    // block period, group 0 had slack); so for even T the assignment is swapped on odd blocks and the two groups take
    // turns.  Accumulator buffer and barrier parities depend on t only.
    const int swap = (p.T % 2 == 0 && p.swap_groups) ? 1 : 0;
    // Split mode (opt-in; two tiles per block, the tokenizer's own shape): with one tile per group, group 1 -- fold +
    // merge + certification -- sets the block period (4 450 cycles) while group 0 waits ~1 000 cycles per block.  In
    // split mode group 0 also drains the first kSplitCols columns of the second tile (any warp may read any accumulator
    // column of its lane quadrant), group 1 the rest.  Measured: no gain (N = 1 M rows: 154-157 us vs 156-162 us) -- the
    // two groups share the schedulers, group 1's certification stretched from 1 400 to 2 900 cycles: what bounds the block
    // period is the latency of each warp's dependent chains with three warps per scheduler, not the split of the work.
    // The swapped assignment above (groups take turns at certifying) was worse: 186-205 us, the certification of block i
    // and the fold of the first tile of block i + 1 then sit on one serial chain.
    const bool split = (p.T == 2 && p.split_last);
    constexpr int kSplitCols = 64;
    uint32_t acc_par = 0;
    for (int t = 0; t < total_tiles; ++t) {
      const int i = t / p.T, j = t - i * p.T;
      int c_lo = 0, c_hi = kBN;
      if (split) {
        if (j == 0) {
          if (g != 0) continue;
        } else if (g == 0) {
          c_hi = kSplitCols;
        } else {
          c_lo = kSplitCols;
        }
      } else if (((t + swap * i) & 1) != g) {
        continue;
      }
      const uint32_t ab = (uint32_t)t & 1u;
      const uint32_t tbuf = tbuf0 + ab * (uint32_t)kBN;
      TR(10);
      mbar_wait((g ? bar_acc_full1 : bar_acc_full) + 8 * ab, (acc_par >> ab) & 1u);  // this group's own phase count of buffer ab
      acc_par ^= 1u << ab;
      fence_after();
      TR(11);
      if (i != cur_blk) {
        // New row block: the in-chunk margin is fixed per row BEFORE the scan (bounds valid for any candidate); the
        // final test may refine it, never enlarge it.  (rowstat was written before bar_a_full completed and the MMA
        // issuer observed that barrier before issuing the MMAs whose commit we just waited for: the read is ordered.)
        cur_blk = i;
        const float zn2 = rowstat[(i % kStatSlots) * kBM + q * 32 + lane];
        zn = sqrtf(zn2);
        margin_row = margin_of(zn, emax_g, zn + emax_g) * 1.0000002f + 1e-30f;
        if (!(margin_row < 3.0e38f)) margin_row = INFINITY;  // non-finite row statistics: nothing can be certified
      }
#pragma unroll 1
      for (int c0 = c_lo; c0 < c_hi; c0 += 64) {
        uint32_t va[32], vb[32];
        tmem_ld32(tbuf + c0, va);
        tmem_ld32(tbuf + c0 + 32, vb);
        tmem_ld_wait2(va, vb);
        if (c0 + 64 >= c_hi) {  // this group's last load of the accumulator has landed: hand it back to the MMA warp
          fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_acc_empty + 8 * ab);
          TR(12);
        }
        float cma, xaa, xba, cmb, xab, xbb;
        fold32(va, cma, xaa, xba);
        fold32(vb, cmb, xab, xbb);
        merge_chunk(cma, xaa, xba, j * kBN + c0);
        merge_chunk(cmb, xab, xbb, j * kBN + c0 + 32);
      }
      TR(14);
      // last_of_block: this group merges the other group's partial and certifies; otherwise it hands its partial over
      const bool last_of_block = split ? (j == 1 && g == 1) : (j == p.T - 1);
      const bool leaves_block = split ? (j == 1) : (last_of_block || (j + 2 >= p.T));
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
        const int ilo = __float2int_rn(ma), ihi = __float2int_rn(mb);
        const bool exact = ((float)ilo == ma) && ((float)ihi == mb) && ((unsigned)ilo < 65536u) && ((unsigned)ihi < 65536u);
        uint32_t mask = (uint32_t)ilo | ((uint32_t)ihi << 16);
        const int ncand = __popc(mask);
        int bidx = base + __ffs((int)mask) - 1;
        // every code OUTSIDE the winning chunk's near-best mask is provably not the nearest when (i) the best of the
        // other chunks is beyond the margin and (ii) no excluded code can interfere
        bool others_out = exact && ncand >= 1 && (margin_row < 3.0e38f);
        const float gap = second - best;
        if (others_out && !(gap > margin_row)) {
          // refine: only codes that can still beat the candidate matter, ||e-mu|| <= ||z-mu|| + d_best (triangle inequality)
          const float dhat = sqrtf(fmaxf(best + zn * zn, 0.f));
          const float emax = dotm ? emax_g : fminf(emax_g, zn + 1.02f * dhat + 1e-30f);
          others_out = gap > fminf(margin_of(zn, emax, dhat), margin_row);
        }
        if (excl_inv > 0.f) {  // codes left out of the contraction: each is further than ||z-mu|| + d_best, or the row is queued
          const float dhat = sqrtf(fmaxf(best + zn * zn, 0.f));
          others_out = others_out && !dotm && ((zn + 1.02f * dhat) * excl_inv < 0.999f);
        }
        bool ok = others_out && ncand == 1 && bidx < p.K;
        if (others_out && ncand >= 2 && ncand <= 4 && base + 32 - __clz((int)mask) <= p.K) {
          // A handful of codes of ONE chunk are within the margin of each other: settle it here, exactly -- fp32 direct
          // differences in the reference's domain (sqrt for L2, first minimum on ties), as the exact scan would, instead
          // of sending the row through two more launches (0.7 % of the rows of a training batch end up here).
          const float* zr = p.z + row * (int64_t)p.D;
          float bv = INFINITY;
          int bk = -1;
          while (mask) {  // ascending code index: a strict < keeps the lowest index on exact ties
            const int c = __ffs((int)mask) - 1;
            mask &= mask - 1;
            const float* er = p.cb + (int64_t)(base + c) * p.D;
            float acc = 0.f;
            for (int d = 0; d < p.D; ++d) {
              const float zv = __ldg(zr + d), ev = __ldg(er + d);
              if (dotm) {
                acc = fmaf(zv, ev, acc);
              } else {
                const float df = zv - ev;
                acc = fmaf(df, df, acc);
              }
            }
            const float val = (p.metric == LIPVQ_METRIC_L2) ? sqrtf(acc) : (dotm ? -acc : acc);
            if (val < bv) {
              bv = val;
              bk = base + c;
            }
          }
          if (bk >= 0) {  // (NaN distances leave bk = -1: the exact scan owns the reference's NaN semantics)
            bidx = bk;
            ok = true;
          }
        }
        if (ok) {
          p.idx_out[row] = (int64_t)bidx;
        } else {
          if (p.tail) {  // this CTA's own segment: scanned at the end of this kernel
            const int pos = atomicAdd(s_qn, 1);
            p.queue_out[(size_t)blockIdx.x * p.qcap + pos] = (int)row;
            // the margin of the codes that can still be the nearest (triangle inequality, as in the refinement above):
            // with the a-priori margin a few far-away codes (large ||e-mu||) made half the codebook a candidate
            const float dhat = sqrtf(fmaxf(best + zn * zn, 0.f));
            const float emax_r = dotm ? emax_g : fminf(emax_g, zn + 1.02f * dhat + 1e-30f);
            const float m_thr = fminf(margin_of(zn, emax_r, dhat), margin_row);
            float* qt = p.queue_thr + 2 * ((size_t)blockIdx.x * p.qcap + pos);
            qt[0] = best + 1.25f * m_thr;  // (NaN / inf: every code is a candidate)
            qt[1] = zn + 1.02f * dhat;     // how far a code can be from the centre and still be the nearest
          } else {
            const int pos = atomicAdd(p.queue_count, 1);
            p.queue_out[pos] = (int)row;
          }
        }
      }
      best = INFINITY; second = INFINITY; ma = 0.f; mb = 0.f; base = 0;
    }
  }

  TR(2);
  fence_before();
  __syncthreads();
  if (!p.early_trigger) grid_dep_launch();
  if (warp == kWarpMma) {
    fence_after();
    tmem_dealloc(tmem_base, 512);
  }
  if (p.tail) {
    // ===== the rows this CTA could not certify (usually none): candidate filter + exact evaluation, see p1_tail_scan.
    // Every MMA, TMA load and bulk copy of this CTA has completed (each was waited for by its consumer); the code tiles
    // and the norm slab are still in place, the A buffers hold the row buffers and candidate lists (13 x (2 Dp + 64) words, or 8 x (2 Dp + 65): <= 16 KB x NC).
    const int nq = *s_qn;
    if (nq > 0) {
      fence_async_smem();  // the A buffers (written by the converters, read by the MMAs) become scratch
      const int* qrows = p.queue_out + (size_t)blockIdx.x * p.qcap;
      const float* qthr = p.queue_thr + 2 * (size_t)blockIdx.x * p.qcap;  // (threshold, reach) per queued row
      if (nq <= kTailCoopRows)
        p1_tail_coop(p.z, p.cb, qrows, qthr, p.idx_out, p.D, p.K, p.Kp, p.NC, p.metric, sB, sE, mu_s, reinterpret_cast<float*>(sA),
                     nq, p.stats_out ? p.ncand_ctr : nullptr, p.nall_ctr);
      else
        p1_tail_scan(p.z, p.cb, qrows, qthr, p.idx_out, p.D, p.K, p.Kp, p.NC, p.metric, sB, sE, mu_s, reinterpret_cast<float*>(sA),
                     nq, p.stats_out ? p.ncand_ctr : nullptr, p.nall_ctr);
      __syncthreads();  // (the statistics below read what the warps added)
    }
    // statistics: the last CTA of the call publishes the grid-wide count and clears the counters for the next call
    if (threadIdx.x == 0 && p.stats_out != nullptr) {  // (two grid-wide round trips at the very end: only when asked for)
      if (nq > 0) atomicAdd(p.flagged_ctr, nq);
      __threadfence();
      const int ticket = atomicAdd(p.done_ctr, 1);
      if (ticket == (int)gridDim.x - 1) {
        __threadfence();
        const int tot = atomicExch(p.flagged_ctr, 0);
        if (p.stats_out) {
          p.stats_out[0] = tot;  // rows settled by the exact evaluation of their candidates
          p.stats_out[1] = tot;  // rows the fp16 pass left uncertified
          p.stats_out[2] = atomicExch(p.ncand_ctr, 0);  // codes evaluated exactly for those rows
          p.stats_out[3] = atomicExch(p.nall_ctr, 0);   // rows for which every code was a candidate (non-finite threshold)
        }
        atomicExch(p.done_ctr, 0);
      }
    }
  }
#ifdef LIPVQ_TRACE
  TR(3);
  if (blockIdx.x == 0 && lane == 0) g_trace_p1_n[warp] = tr_n;
#endif
}

// Codebook preparation as a launch of its own: everything search_p1_kernel needs that depends on the codebook only -- the
// centre mu, -2(e-mu) as swizzled fp16 tiles, the fp16 split of ||e-mu||^2 (norm slab), the maxima the certification
// margin uses.  Run once per codebook VERSION (bulk tokenisation: once; training: after the codebook update, off the
// search's critical path); the search itself then starts with one bulk copy instead of an 18 000-cycle prologue.
// mu = mean of <= 32 evenly spaced rows of `center` (e.g. a batch of latents; any centre is valid -- the bounds are
// evaluated for the centre that was used -- a centre near the latents certifies the most rows) or, without one, of the
// codebook itself.  Each CTA owns cpc codes; every CTA derives the same mu (fixed summation order).
__global__ void __launch_bounds__(256, 1)
    search_prep_kernel(const float* __restrict__ cb, int K, int D, int Kp, int NC, int metric, const float* __restrict__ center,
                       int64_t n_center, int cpc, P1State* st, uint8_t* img) {
  __shared__ float samp[32 * 256];
  __shared__ float mu_s[256];
  __shared__ float red_a[8], red_b[8];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Dp = NC * kChunk;
  const int b_tile = kBN * kChunk * 2;
  const int T = Kp / kBN;
  grid_dep_wait();
  {
    constexpr int RMAX = 32;
    const float* src = center ? center : cb;
    const int64_t n = center ? n_center : (int64_t)K;
    const int R = (int)(n < RMAX ? n : RMAX);
    const int64_t stride = n / R;
    const bool want_mu = metric != LIPVQ_METRIC_DOT;
    const int dl = (int)threadIdx.x & 63, il = (int)threadIdx.x >> 6;
    for (int db = 0; db < 256; db += 64) {
      const int d = db + dl;
      float v[8];
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const int i = il + 4 * q;
        v[q] = (want_mu && d < D && i < R) ? __ldg(src + (int64_t)i * stride * D + d) : 0.f;
      }
#pragma unroll
      for (int q = 0; q < 8; ++q) samp[(il + 4 * q) * 256 + d] = v[q];
    }
    __syncthreads();
    const int d = (int)threadIdx.x;
    float sum = 0.f;
#pragma unroll 8
    for (int i = 0; i < RMAX; ++i) sum += samp[i * 256 + d];  // rows >= R hold zeros; fixed order
    mu_s[d] = sum / (float)R;
    __syncthreads();
  }
  const int ipr = Dp / 8;
  const int ipr_p2 = ipr <= 8 ? 8 : (ipr <= 16 ? 16 : 32);
  const int rpp = 256 / ipr_p2;
  const int sub = (int)threadIdx.x % ipr_p2, rslot = (int)threadIdx.x / ipr_p2;
  const int d0 = sub * 8;
  const bool col_live = sub < ipr && d0 < D;
  const bool vec = (D % 4 == 0) && ((reinterpret_cast<uintptr_t>(cb) & 15) == 0) && (d0 + 8 <= D);
  const int lo = (int)blockIdx.x * cpc, hi = min(lo + cpc, Kp);
  const bool dot = (metric == LIPVQ_METRIC_DOT);
  const size_t img_tiles_bytes = (size_t)T * NC * b_tile;
  float mu8[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) mu8[e] = (col_live && d0 + e < D) ? mu_s[d0 + e] : 0.f;
  const int ch = sub & 7;
  float emax2_loc = 0.f, excl_inv_loc = 0.f;
  for (int k0 = lo; k0 < hi; k0 += rpp) {
    const int k = k0 + rslot;
    float x[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) x[e] = 0.f;
    if (k < hi && k < K && col_live) {
      const float* src = cb + (size_t)k * D + d0;
      if (vec) {
        const float4 a = __ldg(reinterpret_cast<const float4*>(src)), b = __ldg(reinterpret_cast<const float4*>(src) + 1);
        x[0] = a.x; x[1] = a.y; x[2] = a.z; x[3] = a.w; x[4] = b.x; x[5] = b.y; x[6] = b.z; x[7] = b.w;
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e)
          if (d0 + e < D) x[e] = __ldg(src + e);
      }
    }
    float c[8];
    float s = 0.f, am = 0.f;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      c[e] = x[e] - mu8[e];
      s = fmaf(c[e], c[e], s);
    }
#pragma unroll
    for (int e = 0; e < 8; e += 2) am = fmaxf(am, fmaxf(fabsf(c[e]), fabsf(c[e + 1])));
    for (int o = 1; o < ipr_p2; o <<= 1) {
      s += __shfl_xor_sync(0xffffffffu, s, o);
      am = fmaxf(am, __shfl_xor_sync(0xffffffffu, am, o));
    }
    if (k >= hi || sub >= ipr) continue;  // (after the shuffles: every lane of the warp takes part in them)
    // a code whose centred entries or norm do not fit fp16 (or that holds inf / NaN) is left out of the contraction
    const bool excluded = k < K && (!(am < kF16Safe) || (dot ? !(s < 3.0e38f) : !(s <= kEnormMax)));
    const bool zero_row = k >= K || excluded;
    uint32_t hh[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float va = zero_row ? 0.f : -2.f * c[2 * e], vb = zero_row ? 0.f : -2.f * c[2 * e + 1];
      const __half2 hp = __floats2half2_rn(va, vb);
      hh[e] = *reinterpret_cast<const uint32_t*>(&hp);
    }
    const int j = k >> 8, r = k & 255;  // kBN = 256
    *reinterpret_cast<uint4*>(img + ((size_t)j * NC + (sub >> 3)) * b_tile + r * 128 + ((ch ^ (r & 7)) << 4)) =
        make_uint4(hh[0], hh[1], hh[2], hh[3]);
    if (sub == 0) {
      const float en = dot ? 0.f : s;
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
        const float inv = (s < 3.0e38f) ? rsqrtf(s) * 1.0001f : 3.0e38f;  // 1/||e-mu|| rounded up (NaN/inf: unusable)
        excl_inv_loc = fmaxf(excl_inv_loc, inv);
      }
      // column 3 (times a zero of the ones slab: no effect on the contraction) tells the in-kernel scan how far an
      // EXCLUDED code is: 1/||e-mu|| rounded up, finite (0 for ordinary codes and padding)
      const float inv_x = excluded ? fminf((s < 3.0e38f) ? rsqrtf(s) * 1.0001f : 65504.f, 65504.f) : 0.f;
      __half2 b = __floats2half2_rn(h2, 0.f);
      b.y = __float2half_ru(inv_x);
      const __half2 a = __floats2half2_rn(h0, h1);
      uint8_t* slab = img + img_tiles_bytes;
      *reinterpret_cast<uint4*>(slab + slab_off(k, 0)) =
          make_uint4(*reinterpret_cast<const uint32_t*>(&a), *reinterpret_cast<const uint32_t*>(&b), 0u, 0u);
      *reinterpret_cast<uint4*>(slab + slab_off(k, 8)) = make_uint4(0u, 0u, 0u, 0u);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    emax2_loc = fmaxf(emax2_loc, __shfl_xor_sync(0xffffffffu, emax2_loc, o));
    excl_inv_loc = fmaxf(excl_inv_loc, __shfl_xor_sync(0xffffffffu, excl_inv_loc, o));
  }
  if (lane == 0) {
    red_a[warp] = emax2_loc;
    red_b[warp] = excl_inv_loc;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = red_a[0], b = red_b[0];
    for (int i = 1; i < 8; ++i) {
      a = fmaxf(a, red_a[i]);
      b = fmaxf(b, red_b[i]);
    }
    st->emax2_part[blockIdx.x] = a;
    st->excl_inv_part[blockIdx.x] = b;
  }
  if (blockIdx.x == 0 && (int)threadIdx.x >= (int)gridDim.x && threadIdx.x < kStateParts) {  // unused parts read as zero
    st->emax2_part[threadIdx.x] = 0.f;
    st->excl_inv_part[threadIdx.x] = 0.f;
  }
  if (blockIdx.x == 0) {
    st->mu[threadIdx.x] = (int)threadIdx.x < Dp ? mu_s[threadIdx.x] : 0.f;
    if (warp == 0) {
      float t = 0.f;
      for (int dd = lane; dd < 256; dd += 32) t = fmaf(mu_s[dd], mu_s[dd], t);
      t = warp_sum(t);
      if (lane == 0) {
        st->mu_norm = sqrtf(t);
        st->nparts = (int)gridDim.x;
        st->K = K;
        st->D = D;
        st->metric = metric;
        st->flagged = 0;
        st->done = 0;
        st->ncand = 0;
        st->nall = 0;
        st->magic = kStateMagic;
      }
    }
  }
  grid_dep_launch();
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

constexpr int kMisc = kStatSlots * kBM * 4 + 2 * kBM * 32 + 256 * 4 + 32 * 4 + 24 * 8 + 64;

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

bool search_p1_fits(int64_t N, int D, int K) {
  if (N < 1 || K < 1 || D < 1) return false;
  return p1_geom(D, K, false).ok;
}
bool search_p1_supported(int64_t N, int D, int K) {
  if (const char* e = getenv("LIPVQ_SEARCH_P1"))  // A/B switch: 0 = the round-1 prep + streamed first pass
    if (e[0] == '0') return false;
  return search_p1_fits(N, D, K);
}

bool search_p1_tail() {
  if (const char* e = getenv("LIPVQ_P1_TAIL")) return e[0] != '0';
  return true;
}

size_t search_p1_image_bytes(int D, int K) {
  const size_t Dp = (size_t)(D + kChunk - 1) / kChunk * kChunk, Kp = (size_t)(K + kBN - 1) / kBN * kBN;
  return Kp * Dp * 2 + Kp * 32;
}

size_t search_p1_state_bytes(int D, int K) { return (size_t)kStateHdr + search_p1_image_bytes(D, K); }

// rows of the per-CTA queue segments the in-kernel exact scan needs (N + one row block of slack per CTA)
int64_t search_p1_queue_rows(int64_t N) { return N + (int64_t)kBM * 257; }

// Codebook preparation for the "prepared" mode (see search_prep_kernel).  state = search_p1_state_bytes(D, K) bytes.
int search_prep_launch(const float* cb, int K, int D, int metric, const float* center, int64_t n_center, void* state,
                       cudaStream_t stream) {
  const P1Geom g = p1_geom(D, K, false);
  if (!g.ok) {
    set_error("search_prep: codebook K=%d D=%d does not stay resident in shared memory", K, D);
    return LIPVQ_ERR_UNSUPPORTED;
  }
  int cpc = 32;
  while ((g.Kp + cpc - 1) / cpc > kStateParts) cpc += 32;
  const int grid = (g.Kp + cpc - 1) / cpc;
  uint8_t* base = reinterpret_cast<uint8_t*>(state);
  LIPVQ_CUDA(launch_pdl(search_prep_kernel, dim3(grid), dim3(256), 0, stream, cb, K, D, g.Kp, g.NC, metric, center, n_center, cpc,
                        reinterpret_cast<P1State*>(base), base + kStateHdr));
  LIPVQ_LAUNCHED("search_prep_kernel");
  return LIPVQ_OK;
}

// The fp16 pass over all rows.  prepared_state == nullptr: ONE launch that also prepares the codebook (cooperatively,
// in its prologue) and publishes mu / enorm / ehi / elo / hdr for the later passes; otherwise the state written by
// search_prep_launch for this codebook.  tail != 0: rows the pass cannot certify are settled by the same kernel (exact
// scan; `queue` then holds search_p1_queue_rows(N) ints, `queue_thr` twice as many floats, and nothing is left for later passes); tail == 0: they are
// appended to `queue` (count at *queue_count, which must be zero on entry).
int search_p1_launch(const float* z, int64_t N, int D, const float* cb, int K, int metric, int Kp_pub, float* mu,
                     float* enorm, void* ehi_, void* elo_, void* img, SearchHdr* hdr, int* queue, int* queue_count, int64_t* idx_out,
                     cudaStream_t stream, void* prepared_state, int tail, int32_t* stats_out, float* queue_thr) {
  __half* ehi = reinterpret_cast<__half*>(ehi_);
  __half* elo = reinterpret_cast<__half*>(elo_);
  const bool staged = (D % 4 == 0) && aligned16(z);
  P1Geom g = p1_geom(D, K, staged);
  if (!g.ok) g = p1_geom(D, K, false);
  if (!g.ok) {
    set_error("search_p1: shape N=%lld D=%d K=%d does not fit", (long long)N, D, K);
    return LIPVQ_ERR_INVALID;
  }
  if (prepared_state != nullptr && !tail) {
    set_error("search_p1: the prepared mode settles its own rows (tail)");
    return LIPVQ_ERR_INVALID;
  }
  P1Params p{};
  p.z = z; p.cb = cb; p.N = N; p.D = D; p.K = K; p.Kp = g.Kp; p.NC = g.NC; p.T = g.T;
  p.nbufA = g.nbufA; p.zstages = g.zstages; p.metric = metric;
  p.Kp_pub = Kp_pub;
  p.slab_swap = 0;
  p.early_trigger = 1;
  if (const char* e = getenv("LIPVQ_P1_EARLY_TRIGGER")) p.early_trigger = (e[0] != '0');
  if (const char* e = getenv("LIPVQ_SLAB_DESC_SWAP")) p.slab_swap = (e[0] == '1');
  p.swap_groups = 0;
  if (const char* e = getenv("LIPVQ_P1_SWAP")) p.swap_groups = (e[0] == '1');
  p.split_last = 0;
  if (const char* e = getenv("LIPVQ_P1_SPLIT")) p.split_last = (e[0] == '1');
  if (p.swap_groups) p.split_last = 0;
  p.mu_g = mu; p.enorm_g = enorm; p.ehi_g = ehi; p.elo_g = elo; p.hdr = hdr;
  p.img = reinterpret_cast<uint8_t*>(img);
  p.queue_out = queue; p.queue_count = queue_count; p.idx_out = idx_out;
  p.prepared = prepared_state != nullptr;
  p.tail = tail;
  p.stats_out = stats_out;
  p.queue_thr = queue_thr;
  if (tail && queue_thr == nullptr) {
    set_error("search_p1: the in-kernel exact scan needs the threshold queue");
    return LIPVQ_ERR_INVALID;
  }
  if (p.prepared) {
    p.st = reinterpret_cast<P1State*>(prepared_state);
    p.img = reinterpret_cast<uint8_t*>(prepared_state) + kStateHdr;
    p.flagged_ctr = &p.st->flagged;
    p.done_ctr = &p.st->done;
    p.ncand_ctr = &p.st->ncand;
    p.nall_ctr = &p.st->nall;
  } else {
    p.flagged_ctr = &hdr->flagged;
    p.done_ctr = &hdr->done;
    p.ncand_ctr = &hdr->ncand;
    p.nall_ctr = &hdr->nall;
  }
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
  static size_t smem_set[kMaxDevices] = {};
  const int dev_ = current_device();
  if (g.smem > smem_set[dev_]) {
    LIPVQ_CUDA(cudaFuncSetAttribute(search_p1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g.smem));
    smem_set[dev_] = g.smem;
  }
  const int row_blocks = (int)ceil_div64(N, kBM);
  int sms = sm_count();
  if (sms > 256) sms = 256;  // (search_p1_queue_rows)
  const int grid = row_blocks < sms ? row_blocks : sms;
  p.qcap = (row_blocks + grid - 1) / grid * kBM;
  LIPVQ_CUDA(launch_pdl(search_p1_kernel, dim3(grid), dim3(kThreads), g.smem, stream, tm_z, p));
  LIPVQ_LAUNCHED("search_p1_kernel");
  return LIPVQ_OK;
}

}  // namespace lipvq

#ifdef LIPVQ_TRACE
extern "C" LIPVQ_API int lipvq_debug_trace_p1(unsigned long long* out, int* counts) {
  if (cudaMemcpyFromSymbol(out, lipvq::g_trace_p1, sizeof(unsigned long long) * 16 * lipvq::kTraceMax) != cudaSuccess) return -1;
  if (cudaMemcpyFromSymbol(counts, lipvq::g_trace_p1_n, sizeof(int) * 16) != cudaSuccess) return -1;
  return lipvq::kTraceMax;
}
#endif
