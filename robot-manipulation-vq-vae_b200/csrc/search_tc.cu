// K3 on the 5th-gen tensor cores: TMA-fed tcgen05.mma split-bf16 contraction with the argmin in the TMEM epilogue.
//
// score[n,k] = ||e_k||^2 - 2 z_n.e_k.  z and e' = -2e are split into bf16 (hi, lo) pairs so that
//   z.e' ~= z_hi.e'_hi + z_hi.e'_lo + z_lo.e'_hi          (dropped terms <= 3*2^-18 |z_d e'_d| per element)
// is three bf16 MMAs accumulated in fp32 in TMEM -- ~16 bits of operand precision at bf16 tensor rate, which keeps the
// "ambiguous row" set that needs the exact fp32 rescore (search.cu) at a fraction of a percent.
//
// Persistent, warp-specialised, one CTA per SM (192 threads):
//   warps 0-3  epilogue: tcgen05.ld accumulator rows (TMEM lane = z row), + ||e||^2, running top-2 / argmin per row
//   warp  4    TMA producer: the resident code slice once per job, then streams [128 x 64] z chunks (hi, lo)
//   warp  5    TMEM allocation + single-thread tcgen05.mma issue; tcgen05.commit signals smem-free / accumulator-full
// A "job" is (code slice, range of 128-row blocks).  The code slice (<= 128 KB of hi/lo bf16, SWIZZLE_128B K-major tiles)
// stays resident in shared memory; z streams through a 2-stage ring exactly once per slice, so L2->SM traffic is
// N*D*4 bytes per slice instead of one codebook tile per (row block, code tile).  Accumulators ping-pong between two
// TMEM buffers so the MMAs of tile t+1 overlap the epilogue of tile t.
//
// Reference op replaced: the [N,K,D] broadcast + norm + argmin of LFQQuantizer.forward
// (robomimic/models/vq_vae/backbone_lfqvae_v5.py:39-46) and its variants listed in search.cu.
#include <cuda.h>
#include <cuda_bf16.h>

#include "search.cuh"

namespace lipvq {
namespace {

constexpr int kBM = 128;            // z rows per block (UMMA_M, cta_group::1)
constexpr int kChunk = 64;          // latent-dim elements per smem tile row: 64 bf16 = one 128-byte swizzle atom
constexpr int kStages = 2;          // z-chunk ring depth
constexpr int kAStageBytes = 2 * kBM * kChunk * 2;  // hi + lo tile
constexpr int kBBudget = 128 * 1024;
constexpr int kThreads = 192;

struct TcGeom {
  int Dp, NC, BN, NT, codes_res, S;
};

// Resident-slice geometry for a shape: as many codes as fit the 128 KB budget, MMA N in {64,128,256}.  With D <= 64 a
// z stage holds the full row, so a slice may span two code tiles (NT = 2); that halves the z re-reads but coarsens the
// work units, so it is used only when it does not leave SMs idle (units per CTA = row blocks per part x NT).
inline int tc_units(int row_blocks, int S, int NT, int sms) {
  int parts = sms / S;
  if (parts < 1) parts = 1;
  if (parts > row_blocks) parts = row_blocks;
  const int rb_per_part = (row_blocks + parts - 1) / parts;
  const int jobs = S * ((row_blocks + rb_per_part - 1) / rb_per_part);
  const int rounds = (jobs + sms - 1) / sms;
  return rounds * rb_per_part * NT;
}

inline TcGeom tc_geom(int64_t N, int D, int K) {
  TcGeom g;
  g.Dp = (D + kChunk - 1) / kChunk * kChunk;
  g.NC = g.Dp / kChunk;
  const int fit = kBBudget / (g.Dp * 4);  // hi+lo bf16 = 4 bytes per element
  g.BN = fit >= 256 ? 256 : (fit >= 128 ? 128 : 64);
  while (g.BN > 64 && g.BN / 2 >= K) g.BN /= 2;  // no wider than the codebook
  g.NT = 1;
  if (g.NC == 1 && fit >= 2 * g.BN && K > g.BN) {
    const int row_blocks = (int)((N + kBM - 1) / kBM);
    const int S1 = (K + g.BN - 1) / g.BN, S2 = (K + 2 * g.BN - 1) / (2 * g.BN);
    if (tc_units(row_blocks, S2, 2, 148) < tc_units(row_blocks, S1, 1, 148)) g.NT = 2;
  }
  g.codes_res = g.BN * g.NT;
  g.S = (K + g.codes_res - 1) / g.codes_res;
  return g;
}

struct TcParams {
  int64_t N;
  int K, NC, BN, NT, codes_res, S;
  int row_blocks;  // ceil(N/128)
  int parts;       // row-block ranges per slice
  int rb_per_part;
  const float* enorm;
  float* pbest;
  float* psecond;
  int* pidx;
};

// ---- PTX wrappers ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// Bounded wait: a protocol bug must fault (trap) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  const long long t0 = clock64();
  do {
    asm volatile(
        "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (!done && clock64() - t0 > 4000000000LL) __trap();
  } while (!done);
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

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
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// UMMA shared-memory descriptor, K-major, SWIZZLE_128B, tile rows of 128 bytes (64 bf16), 8-row atoms 1024 B apart.
// Bit layout per cute/arch/mma_sm100_desc.hpp (UMMA::SmemDescriptor): start>>4 [0,14), LBO>>4 [16,30) (=1 for swizzled
// K-major), SBO>>4 [32,46) (=1024>>4), version [46,48) = 1, layout type [61,64) = 2 (SWIZZLE_128B).
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
  return (uint64_t)((smem_addr & 0x3FFFF) >> 4) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// Instruction descriptor (UMMA::InstrDescriptor): D=f32 (bit 4), A=B=bf16 (bits 7, 10), both K-major, N>>3 at [17,23),
// M>>4 at [24,29).
__device__ __forceinline__ uint32_t umma_idesc_bf16(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// ---- bf16 hi/lo split (+ row norms) -----------------------------------------------------------------------------
// out_hi/out_lo [rows_p][Dp] bf16, zero padded.  scale = 1 for z, -2 for the codebook.  One thread per 8 elements.
__global__ void __launch_bounds__(256)
    split_bf16_kernel(const float* __restrict__ x, const float* __restrict__ mu, int64_t rows, int D, int64_t rows_p,
                      int Dp, float scale, __nv_bfloat16* __restrict__ hi, __nv_bfloat16* __restrict__ lo) {
  const int per_row = Dp / 8;
  const int64_t total = rows_p * per_row;
  for (int64_t it = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; it < total; it += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = it / per_row;
    const int c = (int)(it - r * per_row) * 8;
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = 0.f;
    if (r < rows) {
      const float* p = x + r * (int64_t)D + c;
      if ((D & 3) == 0 && c + 8 <= D && ((reinterpret_cast<uintptr_t>(p) & 15) == 0)) {
        const float4 a = __ldg(reinterpret_cast<const float4*>(p));
        const float4 b = __ldg(reinterpret_cast<const float4*>(p + 4));
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] -= mu[c + j];
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (c + j < D) v[j] = __ldg(p + j) - mu[c + j];
      }
    }
    __align__(16) __nv_bfloat16 h[8], l[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float s = v[j] * scale;  // exact (power of two)
      h[j] = __float2bfloat16_rn(s);
      l[j] = __float2bfloat16_rn(s - __bfloat162float(h[j]));
    }
    *reinterpret_cast<uint4*>(hi + r * (int64_t)Dp + c) = *reinterpret_cast<const uint4*>(h);
    *reinterpret_cast<uint4*>(lo + r * (int64_t)Dp + c) = *reinterpret_cast<const uint4*>(l);
  }
}

// ---- the contraction ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads, 1)
    search_tc_kernel(const __grid_constant__ CUtensorMap tm_zhi, const __grid_constant__ CUtensorMap tm_zlo,
                     const __grid_constant__ CUtensorMap tm_ehi, const __grid_constant__ CUtensorMap tm_elo,
                     const TcParams p) {
  extern __shared__ uint8_t smem_raw[];
  // carve: [B hi | B lo | A stages | en | barriers], 1024-byte aligned for SWIZZLE_128B
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int b_tile_bytes = p.BN * kChunk * 2;                 // one [BN x 64] bf16 tile
  const int b_half_bytes = p.NT * p.NC * b_tile_bytes;        // all hi (or all lo) tiles of the slice
  uint8_t* sBhi = smem;
  uint8_t* sBlo = smem + b_half_bytes;
  uint8_t* sA = smem + 2 * b_half_bytes;                      // stage s: [hi 16 KB | lo 16 KB]
  float* en_s = reinterpret_cast<float*>(sA + kStages * kAStageBytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(en_s + 512);
  // barrier slots
  const uint32_t bar_b_full = smem_u32(bars + 0), bar_b_empty = smem_u32(bars + 1);
  const uint32_t bar_a_full = smem_u32(bars + 2);    // [kStages]
  const uint32_t bar_a_empty = smem_u32(bars + 4);   // [kStages]
  const uint32_t bar_acc_full = smem_u32(bars + 6);  // [2]
  const uint32_t bar_acc_empty = smem_u32(bars + 8); // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 10);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t tmem_cols = 2u * (uint32_t)p.BN;  // 128 / 256 / 512: power of two >= 32

  if (threadIdx.x == 0) {
    mbar_init(bar_b_full, 1);
    mbar_init(bar_b_empty, 1);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(bar_a_full + 8 * s, 1);
      mbar_init(bar_a_empty + 8 * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_acc_full + 8 * s, 1);
      mbar_init(bar_acc_empty + 8 * s, 4);  // one arrive per epilogue warp
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 5) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"(tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int total_jobs = p.S * p.parts;

  if (warp == 4) {
    // ===== TMA producer =====
    if (lane == 0) {
      uint32_t sidx = 0, jb = 0;
      for (int job = blockIdx.x; job < total_jobs; job += gridDim.x, ++jb) {
        const int slice = job / p.parts, part = job - slice * p.parts;
        const int rb0 = part * p.rb_per_part;
        const int rb1 = min(p.row_blocks, rb0 + p.rb_per_part);
        // resident code slice: wait until the previous job's MMAs have drained it
        mbar_wait(bar_b_empty, (jb & 1) ^ 1);
        mbar_expect_tx(bar_b_full, 2u * (uint32_t)b_half_bytes);
        for (int j = 0; j < p.NT; ++j)
          for (int c = 0; c < p.NC; ++c) {
            const int row = slice * p.codes_res + j * p.BN;
            const uint32_t off = (uint32_t)((j * p.NC + c) * b_tile_bytes);
            tma_load_2d(smem_u32(sBhi + off), &tm_ehi, bar_b_full, c * kChunk, row);
            tma_load_2d(smem_u32(sBlo + off), &tm_elo, bar_b_full, c * kChunk, row);
          }
        for (int rb = rb0; rb < rb1; ++rb)
          for (int c = 0; c < p.NC; ++c, ++sidx) {
            const uint32_t st = sidx % kStages, par = (sidx / kStages) & 1;
            mbar_wait(bar_a_empty + 8 * st, par ^ 1);
            mbar_expect_tx(bar_a_full + 8 * st, (uint32_t)kAStageBytes);
            const uint32_t dst = smem_u32(sA + st * kAStageBytes);
            tma_load_2d(dst, &tm_zhi, bar_a_full + 8 * st, c * kChunk, rb * kBM);
            tma_load_2d(dst + kAStageBytes / 2, &tm_zlo, bar_a_full + 8 * st, c * kChunk, rb * kBM);
          }
      }
    }
  } else if (warp == 5) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(kBM, p.BN);
      uint32_t sidx = 0, jb = 0, t = 0;
      for (int job = blockIdx.x; job < total_jobs; job += gridDim.x, ++jb) {
        const int slice = job / p.parts, part = job - slice * p.parts;
        const int rb0 = part * p.rb_per_part;
        const int rb1 = min(p.row_blocks, rb0 + p.rb_per_part);
        mbar_wait(bar_b_full, jb & 1);
        tc_fence_after();
        for (int rb = rb0; rb < rb1; ++rb) {
          for (int j = 0; j < p.NT; ++j, ++t) {
            const uint32_t ab = t & 1;
            mbar_wait(bar_acc_empty + 8 * ab, ((t >> 1) & 1) ^ 1);
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + ab * (uint32_t)p.BN;
            for (int c = 0; c < p.NC; ++c) {
              const uint32_t s_here = sidx + c;
              const uint32_t st = s_here % kStages, par = (s_here / kStages) & 1;
              if (j == 0) {
                mbar_wait(bar_a_full + 8 * st, par);
                tc_fence_after();
              }
              const uint32_t a_hi = smem_u32(sA + st * kAStageBytes), a_lo = a_hi + kAStageBytes / 2;
              const uint32_t boff = (uint32_t)((j * p.NC + c) * b_tile_bytes);
              const uint32_t b_hi = smem_u32(sBhi + boff), b_lo = smem_u32(sBlo + boff);
              // hi*hi + hi*lo + lo*hi, four K=16 steps each (32 bytes along the 128-byte swizzled row)
#pragma unroll
              for (int combo = 0; combo < 3; ++combo) {
                const uint32_t aa = (combo == 2) ? a_lo : a_hi;
                const uint32_t bb = (combo == 1) ? b_lo : b_hi;
#pragma unroll
                for (int k = 0; k < kChunk / 16; ++k) {
                  const uint32_t acc = (c | combo | k) != 0 ? 1u : 0u;
                  umma_bf16(d_tmem, umma_desc_sw128(aa + 32 * k), umma_desc_sw128(bb + 32 * k), idesc, acc);
                }
              }
              if (j == p.NT - 1) umma_commit(bar_a_empty + 8 * st);  // z chunk free once these MMAs retire
            }
            umma_commit(bar_acc_full + 8 * ab);
          }
          sidx += p.NC;
        }
        umma_commit(bar_b_empty);  // code slice free once every MMA of this job has retired
      }
    }
  } else {
    // ===== epilogue warps 0-3: TMEM lane = row =====
    uint32_t t = 0;
    for (int job = blockIdx.x; job < total_jobs; job += gridDim.x) {
      const int slice = job / p.parts, part = job - slice * p.parts;
      const int rb0 = part * p.rb_per_part;
      const int rb1 = min(p.row_blocks, rb0 + p.rb_per_part);
      const int k0 = slice * p.codes_res;
      asm volatile("bar.sync 1, 128;" ::: "memory");  // everyone finished reading the previous slice's norms
      for (int i = threadIdx.x; i < p.codes_res; i += 128) en_s[i] = (k0 + i < p.K) ? __ldg(p.enorm + k0 + i) : INFINITY;
      asm volatile("bar.sync 1, 128;" ::: "memory");
      for (int rb = rb0; rb < rb1; ++rb) {
        float best = INFINITY, second = INFINITY;
        int bidx = k0;
        for (int j = 0; j < p.NT; ++j, ++t) {
          const uint32_t ab = t & 1;
          mbar_wait(bar_acc_full + 8 * ab, (t >> 1) & 1);
          tc_fence_after();
          const uint32_t trow = tmem_base + ab * (uint32_t)p.BN + ((uint32_t)(warp * 32) << 16);
          for (int c0 = 0; c0 < p.BN; c0 += 32) {
            uint32_t v[32];
            tmem_ld32(trow + c0, v);
            tmem_ld_wait();
            const float* en = en_s + j * p.BN + c0;
            const int kb = k0 + j * p.BN + c0;
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              const float s = __uint_as_float(v[i]) + en[i];
              const bool lt = s < best;
              second = lt ? best : fminf(second, s);
              bidx = lt ? (kb + i) : bidx;
              best = lt ? s : best;
            }
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_acc_empty + 8 * ab);
        }
        const int64_t row = (int64_t)rb * kBM + warp * 32 + lane;
        if (row < p.N) {
          const int64_t o = (int64_t)slice * p.N + row;
          p.pbest[o] = best;
          p.psecond[o] = second;
          p.pidx[o] = bidx;
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols) : "memory");
  }
}

// ---- host side --------------------------------------------------------------------------------------------------
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

// [rows][Dp] bf16 row-major, box = [64 elements x box_rows], 128-byte swizzle
int make_map(CUtensorMap* tm, void* base, int64_t rows, int Dp, int box_rows) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled entry point unavailable");
    return LIPVQ_ERR_CUDA;
  }
  const cuuint64_t gdim[2] = {(cuuint64_t)Dp, (cuuint64_t)rows};
  const cuuint64_t gstride[1] = {(cuuint64_t)Dp * 2};
  const cuuint32_t box[2] = {(cuuint32_t)kChunk, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, gdim, gstride, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with %d (rows=%lld Dp=%d box_rows=%d)", (int)r, (long long)rows, Dp, box_rows);
    return LIPVQ_ERR_CUDA;
  }
  return LIPVQ_OK;
}

}  // namespace

bool search_tc_supported(int64_t N, int D, int K) {
  if (N < 1 || K < 1 || D < 1 || D > 512) return false;
  return true;
}

float search_tc_gamma(int D) {
  const int Dp = (D + kChunk - 1) / kChunk * kChunk;
  // 3*2^-18 for the dropped lo*lo / residual terms of the split, plus fp32 accumulation of 3*Dp products in the tensor
  // core (bounded as one unit roundoff 2^-23 of truncation per accumulated product)
  return 3.f * 3.8146973e-6f + (float)(3 * Dp + 8) * 1.1920929e-7f;
}

// Geometry hooks for make_search_plan (search.cu)
void search_tc_plan(int64_t N, int D, int K, int* codes_res, int* S, int* Dp) {
  const TcGeom g = tc_geom(N, D, K);
  *codes_res = g.codes_res;
  *S = g.S;
  *Dp = g.Dp;
}

int search_tc_launch(const float* z, int64_t N, int D, const float* cb, int K, const float* mu, const SearchPlan& plan,
                     char* ws, cudaStream_t stream) {
  const TcGeom g = tc_geom(N, D, K);
  if (plan.ksplit != g.S || plan.codes_per_split != g.codes_res || plan.Dp != g.Dp) {
    set_error("search_tc: plan/geometry mismatch");
    return LIPVQ_ERR_INVALID;
  }
  __nv_bfloat16* zhi = (__nv_bfloat16*)(ws + plan.off_zhi);
  __nv_bfloat16* zlo = (__nv_bfloat16*)(ws + plan.off_zlo);
  __nv_bfloat16* ehi = (__nv_bfloat16*)(ws + plan.off_ehi);
  __nv_bfloat16* elo = (__nv_bfloat16*)(ws + plan.off_elo);
  const int64_t Np = plan.Np;
  const int Kp = plan.Kp;

  {
    const int64_t items = Np * (g.Dp / 8);
    int64_t blocks = ceil_div64(items, 256);
    const int64_t cap = (int64_t)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    split_bf16_kernel<<<(unsigned)blocks, 256, 0, stream>>>(z, mu, N, D, Np, g.Dp, 1.f, zhi, zlo);
    LIPVQ_LAUNCHED("split_bf16_kernel(z)");
    const int64_t items_e = (int64_t)Kp * (g.Dp / 8);
    blocks = ceil_div64(items_e, 256);
    if (blocks > cap) blocks = cap;
    split_bf16_kernel<<<(unsigned)blocks, 256, 0, stream>>>(cb, mu, K, D, Kp, g.Dp, -2.f, ehi, elo);
    LIPVQ_LAUNCHED("split_bf16_kernel(codebook)");
  }

  CUtensorMap tm_zhi, tm_zlo, tm_ehi, tm_elo;
  int rc;
  if ((rc = make_map(&tm_zhi, zhi, Np, g.Dp, kBM)) != LIPVQ_OK) return rc;
  if ((rc = make_map(&tm_zlo, zlo, Np, g.Dp, kBM)) != LIPVQ_OK) return rc;
  if ((rc = make_map(&tm_ehi, ehi, Kp, g.Dp, g.BN)) != LIPVQ_OK) return rc;
  if ((rc = make_map(&tm_elo, elo, Kp, g.Dp, g.BN)) != LIPVQ_OK) return rc;

  TcParams p{};
  p.N = N; p.K = K; p.NC = g.NC; p.BN = g.BN; p.NT = g.NT; p.codes_res = g.codes_res; p.S = g.S;
  p.row_blocks = (int)(Np / kBM);
  const int sms = sm_count();
  int parts = sms / g.S;
  if (parts < 1) parts = 1;
  if (parts > p.row_blocks) parts = p.row_blocks;
  p.rb_per_part = (p.row_blocks + parts - 1) / parts;
  p.parts = (p.row_blocks + p.rb_per_part - 1) / p.rb_per_part;
  p.enorm = (const float*)(ws + plan.off_enorm);
  p.pbest = (float*)(ws + plan.off_pbest);
  p.psecond = (float*)(ws + plan.off_psecond);
  p.pidx = (int*)(ws + plan.off_pidx);

  const size_t smem = 1024 + (size_t)g.codes_res * g.Dp * 4 + (size_t)kStages * kAStageBytes + 512 * sizeof(float) + 128;
  static size_t smem_set = 0;
  if (smem > smem_set) {
    LIPVQ_CUDA(cudaFuncSetAttribute(search_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    smem_set = smem;
  }
  int grid = p.S * p.parts;
  if (grid > sms) grid = sms;
  search_tc_kernel<<<grid, kThreads, smem, stream>>>(tm_zhi, tm_zlo, tm_ehi, tm_elo, p);
  LIPVQ_LAUNCHED("search_tc_kernel");
  return LIPVQ_OK;
}

}  // namespace lipvq
