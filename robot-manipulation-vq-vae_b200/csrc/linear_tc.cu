// K1/K2/K7 on the tensor cores: token-major linear layers  C[N, Nb] = epi( act(A)[N, Ka] . B[Nb, Ka]^T )  with fp32
// accuracy.  Both operands are split into THREE bf16 terms (x = h + m + l captures all 24 mantissa bits) and six
// tcgen05.mma chains  hh + hm + mh + hl + lh + mm  accumulate in fp32 in TMEM (dropped terms <= 3*2^-24 |a||b|), so the
// result is as accurate as an fp32 FFMA GEMM while the kernel runs at the HBM rate of its activations.
//
//   forward    Y  = act_out(act_in(X) W^T + b)     A = X  [N, in],  B[n][k] = W[n][k]
//   backward   dX = (dY W) * act_in'(Xpre)         A = dY [N, out], B[n][k] = W[k][n]
//
// One persistent CTA per SM (800 threads; both software roles are dependent-instruction chains, so they get many warps):
//   warps 0-15  epilogue: tcgen05.ld (TMEM lane = token row) -> bias/activation (and its derivative) or activation
//               gradient -> per-warp swizzled scratch tile -> global stores with whole sectors per row
//   warps 16-23 converters: fp32 rows -> (activation) -> bf16 (h, m, l) -> 128B-swizzled K-major A tiles
//   warp  24    TMEM allocation + single-thread tcgen05.mma issue
// (the weight-gradient kernel below keeps 8 epilogue + 16 converter warps: all of its work is conversion)
// The weight matrix (<= 256 x 128) is split and laid out in shared memory once per CTA by all threads: no workspace,
// no second launch.  Accumulators ping-pong between two TMEM buffers; A tiles are double-buffered when they fit.
//
// Reference ops replaced: nn.Linear + GELU/ReLU/sigmoid chains and their autograd (backbone_lfqvae_v5.py:54-68,71-76;
// backbone.py:17-32).  The fp32 FFMA tiles of linear.cu remain the path for shapes outside this kernel's range.
#include <cuda.h>
#include <cuda_bf16.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace lipvq {
namespace {

using namespace tc;

constexpr int kBM = 128, kChunk = 64;
constexpr int kATile = kBM * kChunk * 2;  // 16 KB
constexpr int kEpiWarps = 8, kCvtWarps = 16;  // plenty of warps per scheduler: both roles are latency-bound chains
constexpr int kWarpMma = kEpiWarps + kCvtWarps;
constexpr int kThreads = 32 * (kWarpMma + 1);  // 800 -> 80 registers per thread
// forward / backward-x kernel: activations are evaluated in the epilogue (once per element, see the dual forward), so
// that role gets sixteen warps (four per TMEM lane quadrant, splitting the columns) and the converters eight
constexpr int kEpiF = 16, kCvtF = 8;
static_assert(kEpiF + kCvtF == kWarpMma, "same MMA warp and block size in both kernels");
constexpr int kCvtItems = (kBM * 8) / (kCvtF * 32);  // 16-byte items per converter thread and chunk (4)

#ifdef LIPVQ_TRACE
// developer build: lane 0 of every warp of CTA 0 logs (clock64 << 8 | event)
constexpr int kTraceMax = 1024;
__device__ unsigned long long g_ltrace[32][kTraceMax];
__device__ int g_ltrace_n[32];
#define TR(ev)                                                                                  \
  do {                                                                                          \
    if (blockIdx.x == 0 && lane == 0 && tr_n < kTraceMax)                                       \
      g_ltrace[warp][tr_n++] = ((unsigned long long)clock64() << 8) | (unsigned long long)(ev); \
  } while (0)
#else
#define TR(ev) do {} while (0)
#endif

struct LinTcParams {
  const float* A; int64_t N; int Ka; int64_t lda; int actA;
  const float* W; int64_t w_rs, w_cs;  // B[n][k] = W[n * w_rs + k * w_cs]
  int Nb;
  float* C; int64_t ldc;
  float* C2;                                 // mode 0, optional: act_out'(pre), same layout as C
  const float* bias; int act_out;            // mode 0
  const float* Xpre; int64_t ldx; int act_in;  // mode 1
  int mode, NC, Nbp, nbufA;
  uint32_t tmem_cols;
  int scr_banks;   // 16-warp scratch banks: 2 for the dual forward and for the staged Xpre tile of mode 1
  int prefetch;    // L2 prefetch of the next row block's operands (LIPVQ_LINEAR_PREFETCH=1)
  int tma_store;   // output tiles leave through cp.async.bulk.tensor (needs 16-byte aligned C / C2 and row pitch)
};

// Three-way bf16 split of two values at once, packed as (x0 | x1 << 16) per term.  The packed conversion
// (cvt.rn.bf16x2.f32) issues on the ordinary pipes; the scalar F2F form runs on the quarter-rate conversion unit and
// was what bounded the converters.  x - float(bf16(x)) is exact, so h + m + l reproduces all 24 mantissa bits.
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
// Run body(tag) with the activation as a compile-time constant: the element loops then contain straight-line code
// instead of a switch per element (branch latency is not hidden with this few warps per scheduler).
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* tm, uint32_t src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(tm), "r"(src), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void prefetch_l2(const void* ptr) { asm volatile("prefetch.global.L2 [%0];" ::"l"(ptr)); }
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

template <int ACT>
struct ActTag { static constexpr int value = ACT; };
template <typename F>
__device__ __forceinline__ void dispatch_act(int act, F&& body) {
  switch (act) {
    case LIPVQ_ACT_RELU: body(ActTag<LIPVQ_ACT_RELU>{}); break;
    case LIPVQ_ACT_GELU: body(ActTag<LIPVQ_ACT_GELU>{}); break;
    case LIPVQ_ACT_SIGMOID: body(ActTag<LIPVQ_ACT_SIGMOID>{}); break;
    case LIPVQ_ACT_GIVEN: body(ActTag<LIPVQ_ACT_GIVEN>{}); break;
    default: body(ActTag<LIPVQ_ACT_NONE>{}); break;
  }
}


template <bool FASTG>
__global__ void __launch_bounds__(kThreads, 1)
    linear_tc_kernel(const __grid_constant__ CUtensorMap tm_c, const __grid_constant__ CUtensorMap tm_c2, const LinTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int a_bytes = 3 * p.NC * kATile;          // [term][chunk]
  const int b_tile = p.Nbp * kChunk * 2;          // [Nbp x 64] bf16
  uint8_t* sA = smem;
  uint8_t* sB = sA + p.nbufA * a_bytes;           // [term][chunk]
  float* bias_s = reinterpret_cast<float*>(sB + 3 * p.NC * b_tile);  // [256]
  float* epi_scr = bias_s + 256;                                     // [1 or 2 banks][16 epilogue warps][32 x 16] fp32
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_scr + p.scr_banks * kEpiF * 512);
  // A-tile barriers are per (buffer, 64-column chunk): slot = buf * NC + c (nbufA * NC <= 4), so with two chunks the MMAs
  // of chunk 0 run while chunk 1 is still being converted, and chunk 0 of the next row block is refilled as soon as its
  // own MMAs retire (with one barrier per buffer the K = 128 layers ran convert -> MMA -> convert strictly in series)
  const uint32_t bar_a_full = smem_u32(bars + 0), bar_a_empty = smem_u32(bars + 4);
  const uint32_t bar_acc_full = smem_u32(bars + 8), bar_acc_empty = smem_u32(bars + 10);
  const uint32_t bar_w = smem_u32(bars + 12);  // weights and bias are in shared memory
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 13);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row_blocks = (int)((p.N + kBM - 1) / kBM);
  const int nblk = ((int)blockIdx.x < row_blocks) ? (row_blocks - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  if (nblk == 0) return;
#ifdef LIPVQ_TRACE
  int tr_n = 0;
  TR(0);
#endif

  if (threadIdx.x == 0) {
    for (int s = 0; s < 4; ++s) {
      mbar_init(bar_a_full + 8 * s, kCvtF);
      mbar_init(bar_a_empty + 8 * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_acc_full + 8 * s, 1);
      mbar_init(bar_acc_empty + 8 * s, kEpiF);
    }
    mbar_init(bar_w, kEpiF);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kWarpMma) tmem_alloc(smem_u32(tmem_slot), p.tmem_cols);
  grid_dep_wait();  // launched as a programmatic dependent: barrier / TMEM setup above overlapped the previous kernel's tail

  fence_before();
  __syncthreads();  // barriers initialised, TMEM base published
  fence_after();
  const uint32_t tmem_base = *tmem_slot;
  TR(1);

  if (warp == kWarpMma) {
    // ===== MMA issuer =====
    if (lane == 0) {
      const uint32_t idesc = idesc_bf16(kBM, p.Nbp);
      const uint32_t b_base = smem_u32(sB);
      mbar_wait(bar_w, 0);
      for (int i = 0; i < nblk; ++i) {
        const int buf = i % p.nbufA;
        const uint32_t ab = (uint32_t)i & 1;
        const uint32_t a_phase = (uint32_t)(i / p.nbufA) & 1;
        const uint32_t d_tmem = tmem_base + ab * (uint32_t)p.Nbp;
        const uint32_t a_base = smem_u32(sA + buf * a_bytes);
        uint32_t first = 0;
        for (int c = 0; c < p.NC; ++c) {
          const int slot = buf * p.NC + c;
          mbar_wait(bar_a_full + 8 * slot, a_phase);
          if (c == 0) {
            TR(30);
            mbar_wait(bar_acc_empty + 8 * ab, (((uint32_t)i >> 1) & 1) ^ 1);
            TR(31);
          }
          fence_after();
          // k-steps that hold data: a 12-wide input needs one 16-deep MMA per product, not the chunk's four
          const int krem = p.Ka - c * kChunk;
          const int ksteps = krem >= kChunk ? kChunk / 16 : (krem + 15) / 16;
          // term index: 0 = h, 1 = m, 2 = l.  smallest products first.
          const int ta[6] = {1, 2, 0, 1, 0, 0}, tb[6] = {1, 0, 2, 0, 1, 0};
          // (one thread issues every MMA: the full chunk is unrolled so its descriptors are immediates off two bases)
          if (ksteps == kChunk / 16) {
#pragma unroll
            for (int combo = 0; combo < 6; ++combo) {
              const uint32_t aa = a_base + (uint32_t)((ta[combo] * p.NC + c) * kATile);
              const uint32_t bb = b_base + (uint32_t)((tb[combo] * p.NC + c) * b_tile);
              const uint32_t a_lo = desc_sw128_lo(aa), b_lo = desc_sw128_lo(bb);
#pragma unroll
              for (int k = 0; k < kChunk / 16; ++k) {
                umma_f16_lo(d_tmem, a_lo + 2 * k, b_lo + 2 * k, kDescSw128Hi, idesc, first);
                first = 1;
              }
            }
          } else {
#pragma unroll
            for (int combo = 0; combo < 6; ++combo) {
              const uint32_t aa = a_base + (uint32_t)((ta[combo] * p.NC + c) * kATile);
              const uint32_t bb = b_base + (uint32_t)((tb[combo] * p.NC + c) * b_tile);
              for (int k = 0; k < ksteps; ++k) {
                umma_f16(d_tmem, desc_sw128(aa + 32 * k), desc_sw128(bb + 32 * k), idesc, first);
                first = 1;
              }
            }
          }
          umma_commit(bar_a_empty + 8 * slot);
        }
        umma_commit(bar_acc_full + 8 * ab);
        TR(32);
      }
    }
  } else if (warp >= kEpiF) {
    // ===== converters: fp32 rows -> (activation) -> bf16 (h, m, l) -> swizzled A tiles =====
    const int tcv = (int)threadIdx.x - kEpiF * 32;
    const int ch = tcv & 7, r0 = tcv >> 3;  // 16-byte chunk of the 128-byte tile row; rows r0 + 32 u
    const bool fast = (p.Ka % 8 == 0) && (p.lda % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.A) & 15) == 0);
    for (int i = 0; i < nblk; ++i) {
      const int buf = i % p.nbufA;
      const int64_t rb = (int64_t)blockIdx.x + (int64_t)i * gridDim.x;
      uint8_t* a_buf = sA + buf * a_bytes;
      for (int c = 0; c < p.NC; ++c) {
        const int col = c * kChunk + ch * 8;
        float4 va[kCvtItems], vb[kCvtItems];
#pragma unroll
        for (int u = 0; u < kCvtItems; ++u) {
          va[u] = make_float4(0.f, 0.f, 0.f, 0.f);
          vb[u] = va[u];
        }
        if (col < p.Ka) {
#pragma unroll
          for (int u = 0; u < kCvtItems; ++u) {
            const int64_t row = rb * kBM + r0 + 32 * u;
            if (row >= p.N) continue;
            const float* src = p.A + row * p.lda + col;
            if (fast) {
              va[u] = __ldg(reinterpret_cast<const float4*>(src));
              vb[u] = __ldg(reinterpret_cast<const float4*>(src) + 1);
            } else {
              float t[8];
#pragma unroll
              for (int e = 0; e < 8; ++e) t[e] = (col + e < p.Ka) ? __ldg(src + e) : 0.f;
              va[u] = make_float4(t[0], t[1], t[2], t[3]);
              vb[u] = make_float4(t[4], t[5], t[6], t[7]);
            }
          }
        }
        // next row block's lines of this chunk -> L2 (one request per 128-byte line), so its loads do not pay DRAM latency
        if (p.prefetch && i + 1 < nblk && col < p.Ka && (ch & 3) == 0) {
#pragma unroll
          for (int u = 0; u < kCvtItems; ++u) {
            const int64_t row = (rb + gridDim.x) * kBM + r0 + 32 * u;
            if (row < p.N) prefetch_l2(p.A + row * p.lda + col);
          }
        }
        // the loads above are already in flight while this waits for the MMAs that read this chunk's tiles
        if (c == 0) TR(20);
        mbar_wait(bar_a_empty + 8 * (buf * p.NC + c), (((uint32_t)(i / p.nbufA)) & 1) ^ 1);
        if (c == 0) TR(21);
#pragma unroll
        for (int u = 0; u < kCvtItems; ++u) {
          const float x[8] = {va[u].x, va[u].y, va[u].z, va[u].w, vb[u].x, vb[u].y, vb[u].z, vb[u].w};
          const int64_t row = rb * kBM + r0 + 32 * u;
          uint32_t hh[4], mm[4], ll[4];
          float y[8];
          dispatch_act(p.actA, [&](auto tag) {
            constexpr int ACT = decltype(tag)::value;
#pragma unroll
            for (int e = 0; e < 8; ++e)  // act(0) may be non-zero (sigmoid): padding must stay zero
              y[e] = (ACT == LIPVQ_ACT_NONE) ? x[e] : ((row < p.N && col + e < p.Ka) ? act_apply(ACT, x[e]) : 0.f);
          });
#pragma unroll
          for (int e = 0; e < 8; e += 2) {
            split3_pair(y[e], y[e + 1], hh[e >> 1], mm[e >> 1], ll[e >> 1]);
          }
          const int r = r0 + 32 * u;
          const uint32_t off = (uint32_t)(r * 128 + ((ch ^ (r & 7)) << 4));
          *reinterpret_cast<uint4*>(a_buf + (0 * p.NC + c) * kATile + off) = make_uint4(hh[0], hh[1], hh[2], hh[3]);
          *reinterpret_cast<uint4*>(a_buf + (1 * p.NC + c) * kATile + off) = make_uint4(mm[0], mm[1], mm[2], mm[3]);
          *reinterpret_cast<uint4*>(a_buf + (2 * p.NC + c) * kATile + off) = make_uint4(ll[0], ll[1], ll[2], ll[3]);
        }
        fence_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_a_full + 8 * (buf * p.NC + c));
      }
      TR(23);
    }
  } else {
    // ===== epilogue warps 0-15: TMEM lane quadrant q = warp & 3, 16-column groups part, part + 4, ... =====
    // TMEM hands each lane one token row, but a warp-wide global access with one row per lane touches 32 different
    // lines.  Every 32x16 group therefore passes through this warp's swizzled scratch tile, so global loads (Xpre) and
    // stores run with four lanes per 64-byte row segment (whole sectors).
    // ---- weights: split into (h, m, l) and store as swizzled K-major tiles; bias to smem ----
    // (done by the sixteen epilogue warps, which have nothing to drain yet, while the converters already fill the
    //  first A tile; the MMA warp waits for bar_w before its first instruction)
    // An item is one 16-byte chunk of a tile row: 8 consecutive k of one output n.  Threads follow the contiguous direction
    // of W (k for the forward layout, n for the transposed one), so the loads coalesce; the swizzle spreads the 16-byte
    // stores of consecutive n over the banks.
    {
      const int KC = p.NC * (kChunk / 8);  // chunks per row: 8 or 16
      const int items = p.Nbp * KC;
      const bool k_fast = (p.w_cs == 1);
      const bool vec = k_fast && (p.w_rs % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.W) & 15) == 0);
      for (int base = 0; base < items; base += 2 * kEpiF * 32) {
        float wv[2][8];
        int nn[2], kc[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int it = base + u * kEpiF * 32 + (int)threadIdx.x;
          nn[u] = k_fast ? it / KC : it % p.Nbp;
          kc[u] = k_fast ? it % KC : it / p.Nbp;
          const int k0 = kc[u] * 8;
          const bool live = it < items && nn[u] < p.Nb;
          const float* src = p.W + (int64_t)nn[u] * p.w_rs + (int64_t)k0 * p.w_cs;
          if (live && vec && k0 + 8 <= p.Ka) {
            const float4 a = __ldg(reinterpret_cast<const float4*>(src)), b = __ldg(reinterpret_cast<const float4*>(src) + 1);
            wv[u][0] = a.x; wv[u][1] = a.y; wv[u][2] = a.z; wv[u][3] = a.w;
            wv[u][4] = b.x; wv[u][5] = b.y; wv[u][6] = b.z; wv[u][7] = b.w;
          } else {
#pragma unroll
            for (int e = 0; e < 8; ++e) wv[u][e] = (live && k0 + e < p.Ka) ? __ldg(src + (int64_t)e * p.w_cs) : 0.f;
          }
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int it = base + u * kEpiF * 32 + (int)threadIdx.x;
          if (it >= items) continue;
          uint32_t hh[4], mm[4], ll[4];
#pragma unroll
          for (int e = 0; e < 8; e += 2) {
            split3_pair(wv[u][e], wv[u][e + 1], hh[e >> 1], mm[e >> 1], ll[e >> 1]);
          }
          const int c = kc[u] >> 3, n = nn[u];
          uint8_t* t0 = sB + c * b_tile + (uint32_t)(n * 128 + (((kc[u] & 7) ^ (n & 7)) << 4));
          *reinterpret_cast<uint4*>(t0) = make_uint4(hh[0], hh[1], hh[2], hh[3]);
          *reinterpret_cast<uint4*>(t0 + p.NC * b_tile) = make_uint4(mm[0], mm[1], mm[2], mm[3]);
          *reinterpret_cast<uint4*>(t0 + 2 * p.NC * b_tile) = make_uint4(ll[0], ll[1], ll[2], ll[3]);
        }
      }
      for (int i = threadIdx.x; i < 256; i += kEpiF * 32)
        bias_s[i] = (p.mode == 0 && p.bias != nullptr && i < p.Nb) ? __ldg(p.bias + i) : 0.f;
      fence_async_smem();
    }
    asm volatile("bar.sync 1, 512;" ::: "memory");  // bias visible to every epilogue warp
    if (lane == 0) mbar_arrive(bar_w);

    const int q = warp & 3, part = warp >> 2;
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    const bool vecC = (p.ldc % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.C) & 15) == 0) &&
                      ((reinterpret_cast<uintptr_t>(p.C2) & 15) == 0);
    const bool vecX = p.mode == 1 && p.Xpre != nullptr && (p.ldx % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.Xpre) & 15) == 0);
    const bool mulX = p.mode == 1 && p.act_in != LIPVQ_ACT_NONE;
    const bool dual = p.mode == 0 && p.C2 != nullptr;
    float* scr = epi_scr + warp * 512;           // [32 rows][4 chunks of 4 floats]; chunk j of row r sits at j ^ ((r >> 1) & 3)
    const int gr = lane >> 2, gj = lane & 3;     // coalesced side: row 8 * it + gr, chunk gj
    const int rsw = (lane >> 1) & 3;             // row side: this lane's token row, chunk j at j ^ rsw
    float* my_row = scr + lane * 16;
    auto load_x = [&](int c0, int64_t row0, float4 (&xr)[4]) {
#pragma unroll
      for (int it = 0; it < 4; ++it) {
        const int r = 8 * it + gr, col = c0 + 4 * gj;
        float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row0 + r < p.N && col < p.Nb) {
          const float* xp = p.Xpre + (row0 + r) * p.ldx + col;
          if (vecX && col + 4 <= p.Nb) {
            t = __ldg(reinterpret_cast<const float4*>(xp));
          } else {
            t.x = __ldg(xp);
            if (col + 1 < p.Nb) t.y = __ldg(xp + 1);
            if (col + 2 < p.Nb) t.z = __ldg(xp + 2);
            if (col + 3 < p.Nb) t.w = __ldg(xp + 3);
          }
        }
        xr[it] = t;
      }
    };
    // coalesced side of the transpose: shared-space reads (32-bit addresses), one 64-bit row pointer stepped by 8 rows
    const uint32_t scr_s = smem_u32(scr);
    const uint32_t tile_rd = scr_s + (uint32_t)(gr * 16) * 4u;  // row 8 * it + gr: + it * 512 bytes; chunk (gj ^ ((r >> 1) & 3))
    auto store_tile = [&](uint32_t tile_s, float* dst, int c0, int64_t row0) {
      const int col = c0 + 4 * gj;
      float* cp = dst + (row0 + gr) * p.ldc + col;
      const bool full = vecC && col + 4 <= p.Nb;
#pragma unroll
      for (int it = 0; it < 4; ++it, cp += 8 * p.ldc) {
        const int r = 8 * it + gr;
        float4 t;
        lds_f4(tile_s + (uint32_t)it * 512u + (uint32_t)((gj ^ ((r >> 1) & 3)) << 4), t);
        if (row0 + r < p.N && col < p.Nb) {
          if (full) {
            *reinterpret_cast<float4*>(cp) = t;
          } else {
            cp[0] = t.x;
            if (col + 1 < p.Nb) cp[1] = t.y;
            if (col + 2 < p.Nb) cp[2] = t.z;
            if (col + 3 < p.Nb) cp[3] = t.w;
          }
        }
      }
    };
    for (int i = 0; i < nblk; ++i) {
      const uint32_t ab = (uint32_t)i & 1;
      const int64_t rb = (int64_t)blockIdx.x + (int64_t)i * gridDim.x;
      const int64_t row0 = rb * kBM + q * 32;
      float4 xr[4];
      if (mulX && 16 * part < p.Nbp) load_x(16 * part, row0, xr);  // independent of the MMA: in flight during the wait
      if (p.prefetch && mulX && i + 1 < nblk) {  // the next row block's Xpre segments of this warp -> L2
        const int64_t rn = row0 + (int64_t)gridDim.x * kBM + lane;
        if (rn < p.N)
          for (int c0 = 16 * part; c0 < p.Nb; c0 += 64) prefetch_l2(p.Xpre + rn * p.ldx + c0);
      }
      TR(10);
      mbar_wait(bar_acc_full + 8 * ab, ((uint32_t)i >> 1) & 1);
      fence_after();
      TR(11);
      const uint32_t trow = tmem_base + ab * (uint32_t)p.Nbp + lane_off;
      if (16 * part >= p.Nbp) {  // narrow outputs: this warp has no column group
        fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_acc_empty + 8 * ab);
      }
      // The group loop is instantiated per (activation, kind) OUTSIDE the loop: inside it the activation switch was an
      // indirect branch per four columns, and together with generic-space scratch accesses and per-iteration set-up it
      // made the epilogue -- the long pole of every linear layer -- a branchy 200-instruction body per 4 columns.
      // kind: 0 forward, 1 forward + derivative (dual), 2 input gradient, 3 input gradient * act_in'(X)
      const uint32_t my_row_s = smem_u32(my_row);
      const uint32_t xs_s = scr_s + kEpiF * 2048;  // second bank: staging of the Xpre tile (kind 3)
      auto groups = [&](auto act_tag, auto kind_tag) {
        constexpr int ACT = decltype(act_tag)::value;
        constexpr int KIND = decltype(kind_tag)::value;
        for (int c0 = 16 * part; c0 < p.Nbp; c0 += 64) {
          const bool last = c0 + 64 >= p.Nbp;
          if (KIND == 3) {
#pragma unroll
            for (int it = 0; it < 4; ++it) {
              const int r = 8 * it + gr;
              sts_f4(xs_s + (uint32_t)(r * 16 + ((gj ^ ((r >> 1) & 3)) << 2)) * 4u, xr[it].x, xr[it].y, xr[it].z, xr[it].w);
            }
            __syncwarp();
            if (!last) load_x(c0 + 64, row0, xr);  // next group's, in flight during this group's math
          }
#pragma unroll
          for (int h = 0; h < 2; ++h) {  // eight columns at a time: one TMEM load + one wait per half group
            uint32_t v[8];
            tmem_ld8(trow + c0 + 8 * h, v);
            tmem_ld_wait8(v);
            if (h == 1 && last) {  // this warp's last load: release the accumulator before the math
              fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(bar_acc_empty + 8 * ab);
            }
            float o[8], d[8];
            if (KIND <= 1) {
              const float4 b0 = *reinterpret_cast<const float4*>(bias_s + c0 + 8 * h);
              const float4 b1 = *reinterpret_cast<const float4*>(bias_s + c0 + 8 * h + 4);
              const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
              for (int e = 0; e < 8; ++e) {
                const float pre = __uint_as_float(v[e]) + bb[e];
                if (ACT == LIPVQ_ACT_GELU && FASTG) {
                  gelu_value_grad_fast(pre, o[e], d[e]);
                } else {
                  o[e] = act_apply(ACT, pre);
                  if (KIND == 1) d[e] = act_grad(ACT, pre);
                }
              }
            } else {
#pragma unroll
              for (int e = 0; e < 8; ++e) o[e] = __uint_as_float(v[e]);
              if (KIND == 3) {
#pragma unroll
                for (int e4 = 0; e4 < 2; ++e4) {
                  float4 t;
                  lds_f4(my_row_s + kEpiF * 2048 + ((((2 * h + e4) ^ rsw) << 2) << 2), t);
                  const float xv[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
                  for (int e = 0; e < 4; ++e) {
                    float g;
                    if (ACT == LIPVQ_ACT_GELU && FASTG) {
                      float unused;
                      gelu_value_grad_fast(xv[e], unused, g);
                    } else {
                      g = act_grad(ACT, xv[e]);
                    }
                    o[4 * e4 + e] *= g;
                  }
                }
              }
            }
            if (h == 0 && p.tma_store) {  // the previous group's bulk store has finished READING this scratch tile
              if (lane == 0) bulk_wait_read0();
              __syncwarp();
            }
#pragma unroll
            for (int e4 = 0; e4 < 2; ++e4) {
              const uint32_t slot = my_row_s + ((((2 * h + e4) ^ rsw) << 2) << 2);  // chunk j of this lane's row at j ^ rsw
              sts_f4(slot, o[4 * e4], o[4 * e4 + 1], o[4 * e4 + 2], o[4 * e4 + 3]);
              if (KIND == 1) sts_f4(slot + kEpiF * 512 * 4, d[4 * e4], d[4 * e4 + 1], d[4 * e4 + 2], d[4 * e4 + 3]);
            }
          }
          TR(12);
          if (p.tma_store) {
            // asynchronous tile store: the warp goes on to the next group while the copy engine drains this one
            // (store_tile kept the LSU busy for ~2000 cycles per group -- as long as the math)
            fence_async_smem();
            __syncwarp();
            if (lane == 0) {
              tma_store_2d(&tm_c, scr_s, c0, (int)row0);
              if (KIND == 1) tma_store_2d(&tm_c2, scr_s + kEpiF * 2048, c0, (int)row0);
              bulk_commit();
            }
          } else {
            __syncwarp();
            store_tile(tile_rd, p.C, c0, row0);
            if (KIND == 1) store_tile(tile_rd + kEpiF * 2048, p.C2, c0, row0);
            __syncwarp();
          }
          TR(13);
        }
      };
      if (dual) {
        dispatch_act(p.act_out, [&](auto tag) { groups(tag, std::integral_constant<int, 1>{}); });
      } else if (p.mode == 0) {
        dispatch_act(p.act_out, [&](auto tag) { groups(tag, std::integral_constant<int, 0>{}); });
      } else if (mulX) {
        dispatch_act(p.act_in, [&](auto tag) { groups(tag, std::integral_constant<int, 3>{}); });
      } else {
        groups(std::integral_constant<int, LIPVQ_ACT_NONE>{}, std::integral_constant<int, 2>{});
      }
      TR(14);
    }
    if (p.tma_store && lane == 0) bulk_wait_all();  // the tile stores are complete before the CTA retires
  }

  TR(2);
  fence_before();
  __syncthreads();
  grid_dep_launch();  // every role of this CTA is done: the next kernel may be scheduled (a waiting dependent polls, so not earlier)
  if (warp == kWarpMma) {
    fence_after();
    tmem_dealloc(tmem_base, p.tmem_cols);
  }
#ifdef LIPVQ_TRACE
  TR(3);
  if (blockIdx.x == 0 && lane == 0) g_ltrace_n[warp] = tr_n;
#endif
}

// ---- weight gradient:  dW[o][k] += sum_t dY[t][o] * act(X)[t][k],  db[o] += sum_t dY[t][o] ---------------------------
// The reduction runs over tokens (K = token), and both operands sit in memory with their OTHER dimension contiguous
// (dY[t][o], X[t][k]).  That is exactly the MN-major operand layout of tcgen05 (instruction descriptor bits 15/16), so
// the converters store whole 16-byte chunks row by row -- no transposition, no sub-word stores:
//   A (dY^T): for each token a 128-byte row holds 64 consecutive outputs (bf16); outputs 64.. form a second block
//   B (X^T) : the same with input columns; column in_dim is a constant 1 (-> db falls out of the same MMA)
// Canonical MN-major SWIZZLE_128B layout: 8 tokens x 128 bytes = one 1024-byte atom (16-byte chunk ^= token & 7); atoms of
// consecutive 8-token groups are SBO = 1024 bytes apart, 64-column blocks LBO bytes apart.  When out_dim <= 64 the second
// A block does not exist: its LBO points at one shared block of zeros instead (M stays 128).
// Every CTA accumulates its share of the 64-token blocks in ONE TMEM accumulator (no per-block epilogue) and adds it to
// dW / db with fire-and-forget reductions at the end.  Double-buffered tiles; 8 epilogue + 16 converter warps.
constexpr int kWTok = 64;          // tokens per block
constexpr int kWBlk = kWTok * 128;  // one 64-column block of one term: 8 KB

struct LinTcWParams {
  const float* dY; const float* X; int64_t N; int out_dim, in_dim; int actX;
  float* dW; float* db;
  int Np;        // MMA N: pad16(in_dim + 1)
  int nbA, nbB;  // 64-column blocks of A (1 or 2) and B (1..3)
  int CA, CB;    // 16-byte chunks per token the converters write: ceil(out_dim / 8) of dY, CB of X
  int ones_sep;  // in_dim % 8 == 0: the ones column has a chunk of its own, written once at kernel start (CB = in_dim / 8)
  uint32_t tmem_cols;
};

// MN-major, SWIZZLE_128B: LBO = distance between 64-element blocks of the M/N dimension, SBO = between 8-token groups
__device__ __forceinline__ uint64_t desc_mn_sw128(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((smem_addr & 0x3FFFF) >> 4) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) |
         (1ull << 46) | (2ull << 61);
}

// Roles of the weight-gradient kernel: the epilogue runs once, at the very end, so four warps (one per TMEM lane quadrant) are
// enough (21 warps instead of 25).  Measured at the bench shape (step time): 8 epilogue warps 0.4549 ms, 4 epilogue warps
// 0.4531 ms; THREE converter register sets (two units in flight; 15 converter warps = 640 threads for 96 registers, or 16
// with 80) 0.480 / 0.492 ms, 15 converter warps with two sets 0.461 ms: the converters are bound by their issue rate, not
// by the bytes in flight -- the macros keep those variants buildable (-DLIPVQ_W_CVT / _SETS / _EPI).
#ifndef LIPVQ_W_CVT
#define LIPVQ_W_CVT 16
#endif
#ifndef LIPVQ_W_SETS
#define LIPVQ_W_SETS 2
#endif
#ifndef LIPVQ_W_EPI
#define LIPVQ_W_EPI 4
#endif
constexpr int kWEpi = LIPVQ_W_EPI, kWCvt = LIPVQ_W_CVT;  // (registers are allocated per four warps: 20 warps -> 96 per thread, 21 -> 80)
constexpr int kWMma = kWEpi + kWCvt;
constexpr int kWThreads = 32 * (kWMma + 1);  // 640

__global__ void __launch_bounds__(kWThreads, 1) linear_tc_bwd_w_kernel(const LinTcWParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int a_term = p.nbA * kWBlk, b_term = p.nbB * kWBlk;   // one term of one buffer
  const int buf_bytes = 3 * (a_term + b_term);                // [A h,m,l | B h,m,l]
  uint8_t* zero_blk = smem + 2 * buf_bytes;                   // 8 KB of zeros: A's missing second block
  uint64_t* bars = reinterpret_cast<uint64_t*>(zero_blk + kWBlk);
  const uint32_t bar_full = smem_u32(bars + 0), bar_empty = smem_u32(bars + 2), bar_done = smem_u32(bars + 4);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 6);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int blocks = (int)((p.N + kWTok - 1) / kWTok);
  const int nblk = ((int)blockIdx.x < blocks) ? (blocks - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  if (nblk == 0) return;
#ifdef LIPVQ_TRACE
  int tr_n = 0;
  TR(0);
#endif

  if (threadIdx.x == 0) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_full + 8 * s, kWCvt);
      mbar_init(bar_empty + 8 * s, 1);
    }
    mbar_init(bar_done, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kWMma) tmem_alloc(smem_u32(tmem_slot), p.tmem_cols);
  // chunks the converters never write (columns past out_dim / in_dim + 1) must read as zero
  for (int i = threadIdx.x; i < (2 * buf_bytes + kWBlk) / 16; i += kWThreads) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  if (p.ones_sep) {  // constant chunk (1, 0, ..., 0) of every token, h term, both buffers: never touched again
    __syncthreads();
    if (threadIdx.x < 2 * kWTok) {
      const int tk = threadIdx.x & (kWTok - 1), buf = threadIdx.x >> 6, c1 = p.in_dim >> 3;
      uint8_t* d = smem + buf * buf_bytes + 3 * a_term + (c1 >> 3) * kWBlk + (tk >> 3) * 1024 + (tk & 7) * 128 + (((c1 & 7) ^ (tk & 7)) << 4);
      *reinterpret_cast<uint4*>(d) = make_uint4(0x00003F80u, 0, 0, 0);  // bf16 1.0 in the low half
    }
  }
  grid_dep_wait();  // programmatic dependent launch: everything above is local to the CTA
  fence_async_smem();
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tmem_base = *tmem_slot;
  TR(1);

  if (warp == kWMma) {
    if (lane == 0) {
      const uint32_t idesc = idesc_bf16(kBM, p.Np) | (1u << 15) | (1u << 16);  // A and B MN-major
      const uint32_t zero_addr = smem_u32(zero_blk);
      // descriptors of (buffer, term) at token 0; a 16-token step advances the start address field by 2048 >> 4
      uint64_t da[2][3], dbd[2][3];
#pragma unroll
      for (int b = 0; b < 2; ++b)
#pragma unroll
        for (int t = 0; t < 3; ++t) {
          const uint32_t aa = smem_u32(smem + b * buf_bytes) + (uint32_t)(t * a_term);
          const uint32_t bb = smem_u32(smem + b * buf_bytes) + 3 * (uint32_t)a_term + (uint32_t)(t * b_term);
          da[b][t] = desc_mn_sw128(aa, (p.nbA == 2) ? (uint32_t)kWBlk : zero_addr - aa, 1024);
          dbd[b][t] = desc_mn_sw128(bb, kWBlk, 1024);
        }
      uint32_t acc = 0;
      for (int i = 0; i < nblk; ++i) {
        const int buf = i & 1;
        mbar_wait(bar_full + 8 * buf, ((uint32_t)i >> 1) & 1);
        fence_after();
        TR(30);
        const int ta[6] = {1, 2, 0, 1, 0, 0}, tb[6] = {1, 0, 2, 0, 1, 0};  // smallest products first
#pragma unroll
        for (int combo = 0; combo < 6; ++combo) {
          const uint64_t a_d = buf ? da[1][ta[combo]] : da[0][ta[combo]];
          const uint64_t b_d = buf ? dbd[1][tb[combo]] : dbd[0][tb[combo]];
#pragma unroll
          for (int k = 0; k < kWTok / 16; ++k) {  // 16 tokens = two 8-token atoms per MMA
            umma_f16(tmem_base, a_d + (uint64_t)(128 * k), b_d + (uint64_t)(128 * k), idesc, acc);
            acc = 1;
          }
        }
        umma_commit(bar_empty + 8 * buf);
        TR(32);
      }
      umma_commit(bar_done);
    }
  } else if (warp >= kWEpi) {
    // ===== converters: item = (token, 16-byte chunk); consecutive threads take consecutive chunks of one token row =====
    const int tcv = (int)threadIdx.x - kWEpi * 32;
    const int CT = p.CA + p.CB;
    const int items = kWTok * CT;
    const bool vecY = (p.out_dim % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.dY) & 15) == 0);
    const bool vecX = (p.in_dim % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.X) & 15) == 0);
    constexpr int kCT = kWCvt * 32;
    const float rcpCT = 1.f / (float)CT;
    const int nbatch = (items + 2 * kCT - 1) / (2 * kCT);  // batches of two items per thread and block
    const int total_units = nblk * nbatch;
    // loads of unit u + 1 (a batch of a block) are issued before unit u is converted: the converters do little
    // arithmetic per byte, so exposed HBM latency would dominate otherwise
    auto load_unit = [&](int unit, float (&w)[2][8]) {
#pragma unroll
      for (int u = 0; u < 2; ++u)
#pragma unroll
        for (int e = 0; e < 8; ++e) w[u][e] = 0.f;
      if (unit >= total_units) return;
      const int i2 = unit / nbatch, b2 = unit - i2 * nbatch;
      const int64_t t0 = ((int64_t)blockIdx.x + (int64_t)i2 * gridDim.x) * kWTok;
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int it = b2 * 2 * kCT + u * kCT + tcv;
        if (it >= items) continue;
        const int tk = __float2int_rd(((float)it + 0.5f) * rcpCT), c = it - tk * CT;
        const int64_t t = t0 + tk;
        if (t >= p.N) continue;
        const bool isA = c < p.CA;
        const int c8 = (isA ? c : c - p.CA) * 8;
        const int dim = isA ? p.out_dim : p.in_dim;
        const float* src = (isA ? p.dY : p.X) + t * dim + c8;
        if ((isA ? vecY : vecX) && c8 + 8 <= dim) {
          const float4 q0 = __ldg(reinterpret_cast<const float4*>(src)), q1 = __ldg(reinterpret_cast<const float4*>(src) + 1);
          w[u][0] = q0.x; w[u][1] = q0.y; w[u][2] = q0.z; w[u][3] = q0.w;
          w[u][4] = q1.x; w[u][5] = q1.y; w[u][6] = q1.z; w[u][7] = q1.w;
        } else {
#pragma unroll
          for (int e = 0; e < 8; ++e)
            if (c8 + e < dim) w[u][e] = __ldg(src + e);
        }
      }
    };
    auto convert_unit = [&](int unit, float (&w)[2][8]) {
      if (unit >= total_units) return;
      const int i = unit / nbatch, b = unit - i * nbatch;
      const int buf = i & 1;
      const int64_t t0 = ((int64_t)blockIdx.x + (int64_t)i * gridDim.x) * kWTok;
      uint8_t* a0 = smem + buf * buf_bytes;
      uint8_t* b0 = a0 + 3 * a_term;
      if (b == 0) {
        TR(20);
        mbar_wait(bar_empty + 8 * buf, (((uint32_t)i >> 1) & 1) ^ 1);
        TR(21);
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int it = b * 2 * kCT + u * kCT + tcv;
        if (it >= items) continue;
        const int tk = __float2int_rd(((float)it + 0.5f) * rcpCT), cc = it - tk * CT;
        const bool isA = cc < p.CA;
        const int c = isA ? cc : cc - p.CA;
        const bool live = t0 + tk < p.N;
        if (!isA) {
          if (p.actX != LIPVQ_ACT_NONE)
            dispatch_act(p.actX, [&](auto tag) {
              constexpr int ACT = decltype(tag)::value;
#pragma unroll
              for (int e = 0; e < 8; ++e) w[u][e] = (live && c * 8 + e < p.in_dim) ? act_apply(ACT, w[u][e]) : 0.f;
            });
          if (!p.ones_sep) {  // in_dim % 8 != 0: the ones column (db = dY^T 1; dead tokens have dY = 0) shares a data chunk
#pragma unroll
            for (int e = 0; e < 8; ++e)
              if (c * 8 + e == p.in_dim) w[u][e] = 1.f;
          }
        }
        uint32_t hh[4], mm[4], ll[4];
#pragma unroll
        for (int e = 0; e < 8; e += 2) split3_pair(w[u][e], w[u][e + 1], hh[e >> 1], mm[e >> 1], ll[e >> 1]);
        uint8_t* d = (isA ? a0 : b0) + (c >> 3) * kWBlk + (tk >> 3) * 1024 + (tk & 7) * 128 + (((c & 7) ^ (tk & 7)) << 4);
        const int term = isA ? a_term : b_term;
        *reinterpret_cast<uint4*>(d) = make_uint4(hh[0], hh[1], hh[2], hh[3]);
        *reinterpret_cast<uint4*>(d + term) = make_uint4(mm[0], mm[1], mm[2], mm[3]);
        *reinterpret_cast<uint4*>(d + 2 * term) = make_uint4(ll[0], ll[1], ll[2], ll[3]);
      }
      if (b == nbatch - 1) {
        fence_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_full + 8 * buf);
        TR(23);
      }
    };
#if LIPVQ_W_SETS == 3
    // three register sets in rotation, two units in flight (a copy at the end of the iteration would wait for the loads it
    // is meant to hide)
    float w0[2][8], w1[2][8], w2[2][8];
    load_unit(0, w0);
    load_unit(1, w1);
    for (int unit = 0; unit < total_units; unit += 3) {
      load_unit(unit + 2, w2);
      convert_unit(unit, w0);
      load_unit(unit + 3, w0);
      convert_unit(unit + 1, w1);
      load_unit(unit + 4, w1);
      convert_unit(unit + 2, w2);
    }
#else
    float w0[2][8], w1[2][8];
    load_unit(0, w0);
    for (int unit = 0; unit < total_units; unit += 2) {
      load_unit(unit + 1, w1);
      convert_unit(unit, w0);
      load_unit(unit + 2, w0);
      convert_unit(unit + 1, w1);
    }
#endif
  } else {
    // ===== epilogue warps 0-3: TMEM lane = output row o; one pass over the 16-column groups at the very end =====
    const int q = warp & 3, half = (kWEpi == 8) ? (warp >> 2) : 0;
    const int o = q * 32 + lane;
    TR(10);
    mbar_wait(bar_done, 0);
    fence_after();
    TR(11);
    const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16);
    // 16-byte vector reductions (four columns per request); every CTA starts at a different column group so the 148
    // CTAs do not walk the same addresses in lock step
    const bool vec = (p.in_dim % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.dW) & 15) == 0);
    const int ngroups = p.Np / 16, mine = (kWEpi == 8) ? (ngroups - half + 1) / 2 : ngroups;
    for (int g = 0; g < mine; ++g) {
      const int c0 = 16 * (half + (kWEpi == 8 ? 2 : 1) * ((g + (int)blockIdx.x) % mine));
      uint32_t v[16];
      tmem_ld16(trow + c0, v);
      tmem_ld_wait16(v);
      if (o < p.out_dim) {
#pragma unroll
        for (int e = 0; e < 16; e += 4) {
          const int k = c0 + e;
          float* dst = p.dW + (size_t)o * p.in_dim + k;
          if (vec && k + 4 <= p.in_dim) {
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(__uint_as_float(v[e])),
                         "f"(__uint_as_float(v[e + 1])), "f"(__uint_as_float(v[e + 2])), "f"(__uint_as_float(v[e + 3]))
                         : "memory");
          } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              if (k + j < p.in_dim) atomicAdd(dst + j, __uint_as_float(v[e + j]));
              else if (k + j == p.in_dim && p.db != nullptr) atomicAdd(p.db + o, __uint_as_float(v[e + j]));
            }
          }
        }
      }
    }
  }

  TR(2);
  fence_before();
  __syncthreads();
  grid_dep_launch();  // every role of this CTA is done: the next kernel may be scheduled (a waiting dependent polls, so not earlier)
  if (warp == kWMma) {
    fence_after();
    tmem_dealloc(tmem_base, p.tmem_cols);
  }
#ifdef LIPVQ_TRACE
  TR(3);
  if (blockIdx.x == 0 && lane == 0) g_ltrace_n[warp] = tr_n;
#endif
}
#undef TR

constexpr int kMisc = 256 * 4 + kEpiF * 2048 + 128;  // bias, epilogue scratch (one bank), barriers
constexpr int kDualScratch = kEpiF * 2048;            // second scratch bank of the dual forward

}  // namespace

// Shapes the tensor path takes: enough tokens to fill the SMs, reduction <= 128, width <= 256, everything fits smem.
// LIPVQ_LINEAR_ALGO=simt forces the FFMA tiles (A/B runs).
bool linear_tc_supported(int mode, int64_t N, int Ka, int Nb, bool dual) {
  if (const char* f = getenv("LIPVQ_LINEAR_ALGO"))
    if (f[0] == 's') return false;
  if (!(N >= 4096 && Ka >= 1 && Ka <= 128 && Nb >= 1 && Nb <= 256)) return false;
  (void)mode;
  const int NC = (Ka + kChunk - 1) / kChunk, Nbp = (Nb + 15) / 16 * 16;
  const size_t a_bytes = (size_t)3 * NC * kATile, b_bytes = (size_t)3 * NC * Nbp * kChunk * 2;
  return 1024 + a_bytes + b_bytes + kMisc + ((dual || mode == 1) ? kDualScratch : 0) <= 227 * 1024;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn lin_encode_fn() {
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
// [N, Nb] fp32 with row pitch ldc, box = one epilogue warp's tile (32 rows x 16 columns, 64-byte rows); the 64-byte
// swizzle IS the scratch layout of the epilogue (16-byte chunk j of row r at j ^ ((r >> 1) & 3)); out-of-range rows and
// columns of a box are clipped by the copy engine
static bool make_out_map(CUtensorMap* tm, float* C, int64_t N, int Nb, int64_t ldc) {
  EncodeTiledFn enc = lin_encode_fn();
  if (!enc || C == nullptr || (reinterpret_cast<uintptr_t>(C) & 15) != 0 || ldc % 4 != 0 || N >= (int64_t)1 << 31) return false;
  const cuuint64_t gdim[2] = {(cuuint64_t)Nb, (cuuint64_t)N};
  const cuuint64_t gstride[1] = {(cuuint64_t)ldc * 4};
  const cuuint32_t box[2] = {16, 32};
  const cuuint32_t estr[2] = {1, 1};
  return enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, C, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
             CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

int linear_tc_launch(int mode, const float* A, int64_t N, int Ka, int64_t lda, int actA, const float* W, int64_t w_rs,
                     int64_t w_cs, int Nb, float* C, int64_t ldc, const float* bias, int act_out, const float* Xpre,
                     int64_t ldx, int act_in, cudaStream_t stream, float* C2) {
  LinTcParams p{};
  p.A = A; p.N = N; p.Ka = Ka; p.lda = lda; p.actA = actA;
  p.W = W; p.w_rs = w_rs; p.w_cs = w_cs; p.Nb = Nb;
  p.C = C; p.ldc = ldc; p.bias = bias; p.act_out = act_out;
  p.C2 = (mode == 0) ? C2 : nullptr;
  p.scr_banks = (p.C2 != nullptr || (mode == 1 && act_in != LIPVQ_ACT_NONE)) ? 2 : 1;
  const size_t misc = kMisc + (p.scr_banks == 2 ? kDualScratch : 0);
  CUtensorMap tm_c, tm_c2;
  memset(&tm_c, 0, sizeof(tm_c));
  memset(&tm_c2, 0, sizeof(tm_c2));
  static const bool no_tma = getenv("LIPVQ_LINEAR_TMA_STORE") != nullptr && getenv("LIPVQ_LINEAR_TMA_STORE")[0] == '0';
  static const bool pre = getenv("LIPVQ_LINEAR_PREFETCH") != nullptr && getenv("LIPVQ_LINEAR_PREFETCH")[0] == '1';
  p.prefetch = pre;
  p.tma_store = !no_tma && make_out_map(&tm_c, C, N, Nb, ldc) && (p.C2 == nullptr || make_out_map(&tm_c2, p.C2, N, Nb, ldc));
  p.Xpre = Xpre; p.ldx = ldx; p.act_in = act_in;
  p.mode = mode;
  p.NC = (Ka + kChunk - 1) / kChunk;
  p.Nbp = (Nb + 15) / 16 * 16;
  uint32_t cols = 32;
  while (cols < 2u * (uint32_t)p.Nbp) cols <<= 1;
  p.tmem_cols = cols;
  const size_t a_bytes = (size_t)3 * p.NC * kATile, b_bytes = (size_t)3 * p.NC * p.Nbp * kChunk * 2;
  p.nbufA = (2 * a_bytes + b_bytes + misc + 1024 <= 227 * 1024) ? 2 : 1;
  const size_t smem = 1024 + p.nbufA * a_bytes + b_bytes + misc;
  if (smem > 227 * 1024) {
    set_error("linear_tc: shape Ka=%d Nb=%d does not fit shared memory", Ka, Nb);
    return LIPVQ_ERR_INVALID;
  }
  // GELU epilogues: the short A&S 7.1.26 form (default; |error| <= 2e-7 (|x| + 1)) or libm erff / expf (LIPVQ_EXACT_GELU=1)
  static const bool fast_gelu = !(getenv("LIPVQ_EXACT_GELU") != nullptr && getenv("LIPVQ_EXACT_GELU")[0] == '1');
  static size_t smem_set[2][kMaxDevices] = {};  // the attribute is per device
  const int dev_ = current_device();
  if (smem > smem_set[fast_gelu][dev_]) {
    if (fast_gelu)
      LIPVQ_CUDA(cudaFuncSetAttribute(linear_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    else
      LIPVQ_CUDA(cudaFuncSetAttribute(linear_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    smem_set[fast_gelu][dev_] = smem;
  }
  const int row_blocks = (int)ceil_div64(N, kBM);
  const int sms = sm_count();
  const int grid = row_blocks < sms ? row_blocks : sms;
  if (fast_gelu)
    LIPVQ_CUDA(launch_pdl(linear_tc_kernel<true>, dim3(grid), dim3(kThreads), smem, stream, tm_c, tm_c2, p));
  else
    LIPVQ_CUDA(launch_pdl(linear_tc_kernel<false>, dim3(grid), dim3(kThreads), smem, stream, tm_c, tm_c2, p));
  LIPVQ_LAUNCHED("linear_tc_kernel");
  return LIPVQ_OK;
}

static size_t bwd_w_smem(int out_dim, int in_dim) {
  const int nbA = (out_dim + 63) / 64, nbB = (in_dim + 1 + 63) / 64;
  return 1024 + 2 * (size_t)3 * (nbA + nbB) * kWBlk + kWBlk + 128;
}

bool linear_tc_bwd_w_supported(int64_t N, int out_dim, int in_dim, int actX, const float* dY, const float* X) {
  (void)actX; (void)dY; (void)X;
  if (const char* f = getenv("LIPVQ_LINEAR_ALGO"))
    if (f[0] == 's') return false;
  return N >= 4096 && out_dim >= 1 && out_dim <= 128 && in_dim >= 1 && in_dim <= 128 && bwd_w_smem(out_dim, in_dim) <= 227 * 1024;
}

int linear_tc_bwd_w_launch(const float* dY, int64_t N, int out_dim, const float* X, int in_dim, int actX, float* dW, float* db,
                           cudaStream_t stream) {
  LinTcWParams p{};
  p.dY = dY; p.X = X; p.N = N; p.out_dim = out_dim; p.in_dim = in_dim; p.actX = actX; p.dW = dW; p.db = db;
  p.Np = (in_dim + 1 + 15) / 16 * 16;
  p.nbA = (out_dim + 63) / 64;
  p.nbB = (in_dim + 1 + 63) / 64;
  p.CA = (out_dim + 7) / 8;
  p.ones_sep = (in_dim % 8 == 0) ? 1 : 0;
  p.CB = p.ones_sep ? in_dim / 8 : (in_dim + 1 + 7) / 8;
  uint32_t cols = 32;
  while (cols < (uint32_t)p.Np) cols <<= 1;
  p.tmem_cols = cols;
  const size_t smem = bwd_w_smem(out_dim, in_dim);
  static size_t smem_set[kMaxDevices] = {};  // the attribute is per device
  const int dev_ = current_device();
  if (smem > smem_set[dev_]) {
    LIPVQ_CUDA(cudaFuncSetAttribute(linear_tc_bwd_w_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    smem_set[dev_] = smem;
  }
  const int blocks = (int)ceil_div64(N, kWTok);
  const int sms = sm_count();
  LIPVQ_CUDA(launch_pdl(linear_tc_bwd_w_kernel, dim3(blocks < sms ? blocks : sms), dim3(kWThreads), smem, stream, p));
  LIPVQ_LAUNCHED("linear_tc_bwd_w_kernel");
  return LIPVQ_OK;
}

}  // namespace lipvq

#ifdef LIPVQ_TRACE
extern "C" LIPVQ_API int lipvq_debug_trace_linear(unsigned long long* out, int* counts) {
  if (cudaMemcpyFromSymbol(out, lipvq::g_ltrace, sizeof(unsigned long long) * 32 * lipvq::kTraceMax) != cudaSuccess) return -1;
  if (cudaMemcpyFromSymbol(counts, lipvq::g_ltrace_n, sizeof(int) * 32) != cudaSuccess) return -1;
  return lipvq::kTraceMax;
}
#endif
