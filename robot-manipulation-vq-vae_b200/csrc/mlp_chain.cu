// K1/K2 fused for inference: a chain of THREE token-major linear layers  Y = act3(act2(act1(X W1^T + b1) W2^T + b2) W3^T + b3)
// in one launch, hidden activations never leaving the SM.  This is the encoder of the tokenizer under torch.no_grad
// (bulk tokenisation: nn.Linear + GELU, nn.Linear + GELU, LipschitzMLP linear + sigmoid; reference
// robomimic/models/vq_vae/backbone_lfqvae_v5.py:54-62, 70-76) and, with other widths, its decoder.
//
// Same arithmetic as three calls of linear_tc_kernel (linear_tc.cu): every operand split into three bf16 terms, six
// tcgen05.mma chains hh + hm + mh + hl + lh + mm accumulating in fp32 in TMEM, bias + activation in fp32 -- but
//   * the weights of all three layers are split and laid out in shared memory ONCE per CTA (120 KB at 12-64-128-64),
//   * the epilogue of layer l turns its TMEM accumulator straight into the swizzled K-major bf16 A tiles of layer l + 1
//     (TMEM lane = token row, eight consecutive columns = one 16-byte chunk of that row's tile line: no transposition),
//   * only the last layer writes to global memory.
// Per launch the three separate kernels paid three prologues (weight staging), three pipeline fills and drains at 3.5
// row blocks per SM, and wrote / re-read 192 floats per token.
//
// One persistent CTA per SM, 800 threads: 24 worker warps in TWO pipelines of twelve (TMEM lane quadrant = warp & 3,
// column groups split three ways), warp 24 allocates TMEM and issues every MMA from one thread.  The layers of one row
// block are a serial chain (epilogue -> MMA -> epilogue ...), so each pipeline works on its own row block and the MMAs
// of one overlap the epilogue arithmetic of the other.  To fit two pipelines next to 120 KB of weights a pipeline owns
// ONE 48 KB region that holds a single 64-wide chunk of the current layer's input at a time (a 128-wide reduction is
// fed in two steps, the region recycled in between) and its own accumulators (2 x 256 TMEM columns at 12-64-128-64).
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
constexpr int kTile = kBM * kChunk * 2;  // one 64-wide chunk of one term of a 128-row A tile: 16 KB
constexpr int kWorkers = 24;
constexpr int kPipeWarps = kWorkers / 2;       // two pipelines of twelve warps: while one waits for its MMAs the other works
constexpr int kParts = kPipeWarps / 4;         // warps per TMEM lane quadrant and pipeline: they split the column groups
constexpr int kRegion = 3 * kTile;             // per pipeline: the three terms of ONE 64-wide chunk of A (48 KB)
constexpr int kThreads = 32 * (kWorkers + 1);  // 800

#ifdef LIPVQ_TRACE
// developer build: lane 0 of every warp of CTA 0 logs (clock64 << 8 | event)
constexpr int kTraceMax = 512;
__device__ unsigned long long g_mtrace[32][kTraceMax];
__device__ int g_mtrace_n[32];
#define TR(ev)                                                                                  \
  do {                                                                                          \
    if (blockIdx.x == 0 && lane == 0 && tr_n < kTraceMax)                                       \
      g_mtrace[warp][tr_n++] = ((unsigned long long)clock64() << 8) | (unsigned long long)(ev); \
  } while (0)
#else
#define TR(ev) do {} while (0)
#endif

struct Mlp3Params {
  const float* X; int64_t N; int64_t ldx;
  const float* W[3]; const float* b[3];
  int din[3], dout[3], act[3];
  int NC[3];   // 64-wide chunks of each layer's reduction
  int Np[3];   // outputs padded to a multiple of 16 (MMA N)
  int wofs[3]; // byte offset of each layer's weight tiles behind the activation region
  int acc[3];  // TMEM column of each layer's accumulator (within a pipeline's columns)
  int acc_cols; // TMEM columns per pipeline
  float* Y; int64_t ldy;
  float* H[2];  // training forward: act_l(pre_l) of the two hidden layers, [N, dout[l]] (NULL: not stored)
  float* G[2];  // ... and act_l'(pre_l): one evaluation of erf / exp serves the value and the derivative
  uint32_t tmem_cols;
  int fast_gelu;
  int terms;   // 3: fp32-accurate (three bf16 terms per operand, six MMA chains); 1: plain bf16 operands, one chain
};

__device__ __forceinline__ uint32_t bf16x2_rn(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
// three-way bf16 split of two values (x = h + m + l holds all 24 mantissa bits), packed (x0 | x1 << 16) per term
__device__ __forceinline__ void split3_pair(float x0, float x1, uint32_t& h, uint32_t& m, uint32_t& l) {
  h = bf16x2_rn(x0, x1);
  const float r0 = x0 - __uint_as_float(h << 16), r1 = x1 - __uint_as_float(h & 0xFFFF0000u);
  m = bf16x2_rn(r0, r1);
  l = bf16x2_rn(r0 - __uint_as_float(m << 16), r1 - __uint_as_float(m & 0xFFFF0000u));
}
// eight consecutive values of tile row r, 16-byte chunk ch of 64-wide chunk c: the three terms of an operand whose
// terms are `term_stride` bytes apart
__device__ __forceinline__ void store_split8(uint8_t* base, int term_stride, int r, int ch, const float (&y)[8], int terms = 3) {
  uint8_t* d = base + (uint32_t)(r * 128 + ((ch ^ (r & 7)) << 4));
  if (terms == 1) {  // bf16 mode: round to nearest, one term
    *reinterpret_cast<uint4*>(d) = make_uint4(bf16x2_rn(y[0], y[1]), bf16x2_rn(y[2], y[3]), bf16x2_rn(y[4], y[5]), bf16x2_rn(y[6], y[7]));
    return;
  }
  uint32_t hh[4], mm[4], ll[4];
#pragma unroll
  for (int e = 0; e < 8; e += 2) split3_pair(y[e], y[e + 1], hh[e >> 1], mm[e >> 1], ll[e >> 1]);
  *reinterpret_cast<uint4*>(d) = make_uint4(hh[0], hh[1], hh[2], hh[3]);
  *reinterpret_cast<uint4*>(d + term_stride) = make_uint4(mm[0], mm[1], mm[2], mm[3]);
  *reinterpret_cast<uint4*>(d + 2 * term_stride) = make_uint4(ll[0], ll[1], ll[2], ll[3]);
}
// one wait for all outstanding tcgen05.ld of this thread (the asm volatile + memory clobber keeps uses below it)
__device__ __forceinline__ void tmem_wait_all() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// one non-blocking probe of an mbarrier phase
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t parity) {
  uint32_t done;
  asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
               : "=r"(done)
               : "r"(bar), "r"(parity)
               : "memory");
  return done != 0;
}
__device__ __forceinline__ void touch8(uint32_t (&v)[8]) {
  asm volatile("" : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7])::"memory");
}
// bias + activation of eight accumulator columns, the activation a compile-time constant: the unrolled element loop is
// then straight-line code of ONE activation (with a switch per element every copy carried erff, expf and the GELU
// polynomial, and the epilogue did not fit the instruction cache)
template <int ACT, bool FAST>
__device__ __forceinline__ void act8(const uint32_t (&v)[8], const float* bl, float (&y)[8]) {
  const float4 b0 = *reinterpret_cast<const float4*>(bl), b1 = *reinterpret_cast<const float4*>(bl + 4);
  const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const float pre = __uint_as_float(v[e]) + bb[e];
    if (ACT == LIPVQ_ACT_GELU && FAST) {
      float d;
      gelu_value_grad_fast(pre, y[e], d);
    } else if (ACT == LIPVQ_ACT_SIGMOID && FAST) {
      // 1 / (1 + 2^(-x log2 e)) on the approximate ex2 / rcp units (each within 2 ulp: |error| <= 3e-7 on a value in (0, 1));
      // expf + an IEEE division made the 64-column output pass as long as two 128-column GELU steps
      y[e] = rcp_approx(1.f + ex2_approx(-1.4426950408889634f * pre));
    } else {
      y[e] = act_apply(ACT, pre);
    }
  }
}
// the same with the activation's derivative (training forward: the hidden layers store both)
template <int ACT, bool FAST>
__device__ __forceinline__ void act8g(const uint32_t (&v)[8], const float* bl, float (&y)[8], float (&d)[8]) {
  const float4 b0 = *reinterpret_cast<const float4*>(bl), b1 = *reinterpret_cast<const float4*>(bl + 4);
  const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const float pre = __uint_as_float(v[e]) + bb[e];
    if (ACT == LIPVQ_ACT_GELU && FAST) {
      gelu_value_grad_fast(pre, y[e], d[e]);
    } else {
      y[e] = act_apply(ACT, pre);
      d[e] = act_grad(ACT, pre);
    }
  }
}
template <int ACT, bool FAST>
struct ActSel { static constexpr int act = ACT; static constexpr bool fast = FAST; };
template <typename F>
__device__ __forceinline__ void dispatch_chain_act(int act, int fast_gelu, F&& body) {
  switch (act) {
    case LIPVQ_ACT_RELU: body(ActSel<LIPVQ_ACT_RELU, false>{}); break;
    case LIPVQ_ACT_GELU:
      if (fast_gelu) body(ActSel<LIPVQ_ACT_GELU, true>{});
      else body(ActSel<LIPVQ_ACT_GELU, false>{});
      break;
    case LIPVQ_ACT_SIGMOID:
      if (fast_gelu) body(ActSel<LIPVQ_ACT_SIGMOID, true>{});
      else body(ActSel<LIPVQ_ACT_SIGMOID, false>{});
      break;
    default: body(ActSel<LIPVQ_ACT_NONE, false>{}); break;
  }
}

// Step s of a row block = (layer l, 64-wide chunk c of that layer's reduction).  Per pipeline the workers write the bf16
// terms of ONE chunk into the pipeline's 48 KB region (from X for layer 0, else from the previous layer's accumulator:
// bias + activation + split), the MMA thread multiplies it into the layer's accumulator and commits `done`, which frees
// the region for the next step and -- after a layer's last chunk -- says that layer's accumulator is complete.
__global__ void __launch_bounds__(kThreads, 1) mlp3_tc_kernel(const Mlp3Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sW = smem + 2 * kRegion;   // per layer: [term][chunk] tiles of [Np x 64] bf16
  float* bias_s = reinterpret_cast<float*>(sW + p.wofs[2] + 3 * p.NC[2] * p.Np[2] * 128);  // [3][256]
  uint64_t* bars = reinterpret_cast<uint64_t*>(bias_s + 3 * 256);
  const uint32_t bar_full = smem_u32(bars + 0);  // [2 pipelines]: the chunk of the current step is written
  const uint32_t bar_done = smem_u32(bars + 2);  // [2 pipelines]: the MMAs of the current step have completed
  const uint32_t bar_w = smem_u32(bars + 4);     // weights and biases are in shared memory
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 5);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row_blocks = (int)((p.N + kBM - 1) / kBM);
  const int nblk = ((int)blockIdx.x < row_blocks) ? (row_blocks - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  if (nblk == 0) return;
  const int S = p.NC[0] + p.NC[1] + p.NC[2];  // steps per row block
#ifdef LIPVQ_TRACE
  int tr_n = 0;
#endif
  TR(0);

  if (threadIdx.x == 0) {
    for (int pl = 0; pl < 2; ++pl) {
      mbar_init(bar_full + 8 * pl, kPipeWarps);
      mbar_init(bar_done + 8 * pl, 1);
    }
    mbar_init(bar_w, kWorkers);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kWorkers) tmem_alloc(smem_u32(tmem_slot), p.tmem_cols);
  grid_dep_wait();  // X (and, in a step, the clamped weight) are the previous kernels' outputs
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tmem_base = *tmem_slot;
  TR(1);

  if (warp == kWorkers) {
    // ===== MMA issuer: ONE thread serves both pipelines, whichever has a chunk ready (the tensor pipe is one resource: with
    // an issuing warp per pipeline the two ran in lockstep, both computing and then both multiplying, and nothing
    // overlapped).  A step costs the pipe ~2 500 cycles whatever its size (6 to 24 small MMAs fed from shared memory).
    if (lane == 0) {
      mbar_wait(bar_w, 0);
      fence_after();
      const int ta[6] = {1, 2, 0, 1, 0, 0}, tb[6] = {1, 0, 2, 0, 1, 0};  // term index 0 = h, 1 = m, 2 = l; smallest products first
      uint32_t a_lo[2][3];
      for (int t = 0; t < 3; ++t) {
        a_lo[0][t] = desc_sw128_lo(smem_u32(smem) + (uint32_t)(t * kTile));
        a_lo[1][t] = desc_sw128_lo(smem_u32(smem) + (uint32_t)(kRegion + t * kTile));
      }
      // per pipeline: steps served, steps to serve, position (layer, chunk) inside the current row block
      uint32_t cnt[2] = {0, 0};
      const uint32_t total[2] = {(uint32_t)(((nblk + 1) / 2) * S), (uint32_t)((nblk / 2) * S)};
      int ll[2] = {0, 0}, cc[2] = {0, 0};
      int pl = 0;
      while (cnt[0] < total[0] || cnt[1] < total[1]) {
        // poll: the pipeline after the one served last is asked first
        bool got = false;
        for (int tr = 0; tr < 2 && !got; ++tr) {
          const int cand = pl ^ tr ^ 1;
          if (cnt[cand] < total[cand] && mbar_try(bar_full + 8 * cand, cnt[cand] & 1)) {
            pl = cand;
            got = true;
          }
        }
        if (!got) continue;
        fence_after();
        TR(31);
        const int l = pl ? ll[1] : ll[0], c = pl ? cc[1] : cc[0];
        const uint32_t idesc = idesc_bf16(kBM, p.Np[l]);
        const uint32_t d_tmem = tmem_base + (uint32_t)(pl * p.acc_cols + p.acc[l]);
        const int b_tile = p.Np[l] * 128;
        const uint32_t bt = (uint32_t)(p.NC[l] * b_tile);
        const uint32_t b0 = smem_u32(sW + p.wofs[l]) + (uint32_t)(c * b_tile);
        const uint32_t b_lo[3] = {desc_sw128_lo(b0), desc_sw128_lo(b0 + bt), desc_sw128_lo(b0 + 2 * bt)};
        const int krem = p.din[l] - c * kChunk;
        const int ksteps = krem >= kChunk ? kChunk / 16 : (krem + 15) / 16;
        uint32_t acc_flag = c > 0 ? 1u : 0u;
        if (p.terms == 1) {  // bf16 mode: the h x h chain only
          const uint32_t al = pl ? a_lo[1][0] : a_lo[0][0];
          for (int k = 0; k < ksteps; ++k) {
            umma_f16_lo(d_tmem, al + 2 * k, b_lo[0] + 2 * k, kDescSw128Hi, idesc, acc_flag);
            acc_flag = 1;
          }
        } else if (ksteps == kChunk / 16) {
#pragma unroll
          for (int combo = 0; combo < 6; ++combo) {
            const uint32_t al = pl ? a_lo[1][ta[combo]] : a_lo[0][ta[combo]], bl_ = b_lo[tb[combo]];
#pragma unroll
            for (int k = 0; k < kChunk / 16; ++k) {
              umma_f16_lo(d_tmem, al + 2 * k, bl_ + 2 * k, kDescSw128Hi, idesc, acc_flag);
              acc_flag = 1;
            }
          }
        } else {
#pragma unroll
          for (int combo = 0; combo < 6; ++combo) {
            const uint32_t al = pl ? a_lo[1][ta[combo]] : a_lo[0][ta[combo]], bl_ = b_lo[tb[combo]];
            for (int k = 0; k < ksteps; ++k) {
              umma_f16_lo(d_tmem, al + 2 * k, bl_ + 2 * k, kDescSw128Hi, idesc, acc_flag);
              acc_flag = 1;
            }
          }
        }
        umma_commit(bar_done + 8 * pl);
        TR(32);
        if (pl) {
          ++cnt[1];
          if (++cc[1] == p.NC[ll[1]]) { cc[1] = 0; ll[1] = (ll[1] + 1) % 3; }
        } else {
          ++cnt[0];
          if (++cc[0] == p.NC[ll[0]]) { cc[0] = 0; ll[0] = (ll[0] + 1) % 3; }
        }
      }
    }
  } else {
    // ===== workers =====
    // ---- weights of the three layers: (h, m, l) split, swizzled K-major tiles; biases (all sixteen warps) ----
    // One index space over the items of all three layers (an item = eight consecutive k of one output row), four items
    // per thread and pass with ALL their loads in flight before the first conversion: layer by layer, one item at a time,
    // the staging paid a dependent L2 / DRAM round trip per item and layer (6 600 cycles).
    {
      int KCl[3], itemsl[3];
      for (int l = 0; l < 3; ++l) {
        KCl[l] = p.NC[l] * (kChunk / 8);
        itemsl[l] = p.Np[l] * KCl[l];
      }
      const int total_items = itemsl[0] + itemsl[1] + itemsl[2];
      for (int base = 0; base < total_items; base += 4 * kWorkers * 32) {
        float w[4][8];
        int lay[4], nn[4], kcs[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          int it = base + u * kWorkers * 32 + (int)threadIdx.x;
          int l = -1;
          if (it < total_items) {
            l = 0;
            if (it >= itemsl[0]) { it -= itemsl[0]; l = 1; }
            if (l == 1 && it >= itemsl[1]) { it -= itemsl[1]; l = 2; }
          }
          lay[u] = l;
#pragma unroll
          for (int e = 0; e < 8; ++e) w[u][e] = 0.f;
          if (l >= 0) {
            const int KC = l == 0 ? KCl[0] : (l == 1 ? KCl[1] : KCl[2]);
            const int n = it / KC, kc = it - n * KC;
            nn[u] = n;
            kcs[u] = kc;
            const int Ka = p.din[l], Nb = p.dout[l], k0 = kc * 8;
            const float* W = p.W[l];
            const float* src = W + (int64_t)n * Ka + k0;
            if (n < Nb) {
              if ((Ka % 4 == 0) && ((reinterpret_cast<uintptr_t>(W) & 15) == 0) && k0 + 8 <= Ka) {
                const float4 a = __ldg(reinterpret_cast<const float4*>(src)), b = __ldg(reinterpret_cast<const float4*>(src) + 1);
                w[u][0] = a.x; w[u][1] = a.y; w[u][2] = a.z; w[u][3] = a.w; w[u][4] = b.x; w[u][5] = b.y; w[u][6] = b.z; w[u][7] = b.w;
              } else {
#pragma unroll
                for (int e = 0; e < 8; ++e)
                  if (k0 + e < Ka) w[u][e] = __ldg(src + e);
              }
            }
          }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int l = lay[u];
          if (l >= 0) {
            const int b_tile = p.Np[l] * 128;
            store_split8(sW + p.wofs[l] + (kcs[u] >> 3) * b_tile, p.NC[l] * b_tile, nn[u], kcs[u] & 7, w[u], p.terms);
          }
        }
      }
      for (int i = threadIdx.x; i < 3 * 256; i += kWorkers * 32) {
        const int l = i >> 8, c = i & 255;
        bias_s[i] = (p.b[l] != nullptr && c < p.dout[l]) ? __ldg(p.b[l] + c) : 0.f;
      }
    }
    fence_async_smem();
    asm volatile("bar.sync 1, %0;" ::"n"(kWorkers * 32) : "memory");  // biases visible to every worker
    if (lane == 0) mbar_arrive(bar_w);
    TR(4);

    const int pl = warp / kPipeWarps;                 // pipeline of this warp: row blocks pl, pl + 2, ... of the CTA
    const int pw = warp % kPipeWarps;                 // warp within the pipeline
    const int pt = (int)threadIdx.x - pl * kPipeWarps * 32;  // thread within the pipeline (0..255)
    const int q = warp & 3, part = pw >> 2;           // TMEM lane quadrant; column groups part, part + 2, ...
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    const int r = q * 32 + lane;                      // this lane's row of the block = its TMEM lane
    uint8_t* sA = smem + pl * kRegion;
    const uint32_t full = bar_full + 8 * pl, done = bar_done + 8 * pl;
    const uint32_t acc_base = tmem_base + (uint32_t)(pl * p.acc_cols) + lane_off;
    const bool vecX = (p.ldx % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.X) & 15) == 0);
    const bool vecY = (p.ldy % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.Y) & 15) == 0);
    uint32_t waited = 0;  // `done` phases this warp has consumed
    // (starting pipeline 1 half a row block late, so that one pipeline's output pass falls into the other's chain of steps,
    //  was tried: with two row blocks per pipeline the late start costs more than the overlap returns, 37.6 -> 42.3 us)
    for (int i = pl; i < nblk; i += 2) {
      const int64_t row0 = ((int64_t)blockIdx.x + (int64_t)i * gridDim.x) * kBM;
      for (int l = 0; l < 3; ++l) {
        for (int c = 0; c < p.NC[l]; ++c) {
          TR(10);
          if (l == 0) {
            // fp32 rows of X, columns [64 c, 64 c + 64) -> bf16 terms.  (The region is free: every earlier step of this
            // pipeline has been waited for below.)
            const int krem = p.din[0] - c * kChunk;
            const int nch = krem >= kChunk ? 8 : 2 * ((krem + 15) / 16);  // 16-byte chunks the MMAs of this step read
            for (int it = pt; it < kBM * nch; it += kPipeWarps * 32) {
              const int rr = it / nch, ch = it - rr * nch;
              const int col = c * kChunk + ch * 8;
              float x[8];
              const int64_t row = row0 + rr;
              const float* src = p.X + row * p.ldx + col;
              if (row < p.N && vecX && col + 8 <= p.din[0]) {
                const float4 a = __ldg(reinterpret_cast<const float4*>(src)), b = __ldg(reinterpret_cast<const float4*>(src) + 1);
                x[0] = a.x; x[1] = a.y; x[2] = a.z; x[3] = a.w; x[4] = b.x; x[5] = b.y; x[6] = b.z; x[7] = b.w;
              } else {
#pragma unroll
                for (int e = 0; e < 8; ++e) x[e] = (row < p.N && col + e < p.din[0]) ? __ldg(src + e) : 0.f;
              }
              store_split8(sA, kTile, rr, ch, x, p.terms);
            }
          } else {
            // columns [64 c, 64 c + 64) of the previous layer's accumulator: bias + activation -> bf16 terms
            const uint32_t trow = acc_base + (uint32_t)p.acc[l - 1];
            const float* bl = bias_s + (l - 1) * 256;
            const int act = p.act[l - 1];
            const int c_hi = min(p.Np[l - 1], (c + 1) * kChunk);
            float* Hl = p.H[l - 1];
            float* Gl = p.G[l - 1];
            const int dl = p.dout[l - 1];
            const bool vecH = (dl % 4 == 0) && (((reinterpret_cast<uintptr_t>(Hl) | reinterpret_cast<uintptr_t>(Gl)) & 15) == 0);
            dispatch_chain_act(act, p.fast_gelu, [&](auto sel) {
              using Sel = decltype(sel);
#pragma unroll 1
              for (int c0 = c * kChunk + 8 * part; c0 < c_hi; c0 += 8 * kParts) {
                uint32_t v[8];
                tmem_ld8(trow + c0, v);
                tmem_ld_wait8(v);
                float y[8];
                if (Hl != nullptr) {
                  // training forward: the hidden activation and its derivative also go to global memory (one 32-byte
                  // piece per lane and array: whole sectors)
                  float d[8];
                  act8g<Sel::act, Sel::fast>(v, bl + c0, y, d);
                  const int64_t row = row0 + r;
                  if (row < p.N && c0 < dl) {
                    float* hp = Hl + row * dl + c0;
                    float* gp = Gl + row * dl + c0;
                    if (vecH && c0 + 8 <= dl) {
                      *reinterpret_cast<float4*>(hp) = make_float4(y[0], y[1], y[2], y[3]);
                      *reinterpret_cast<float4*>(hp + 4) = make_float4(y[4], y[5], y[6], y[7]);
                      *reinterpret_cast<float4*>(gp) = make_float4(d[0], d[1], d[2], d[3]);
                      *reinterpret_cast<float4*>(gp + 4) = make_float4(d[4], d[5], d[6], d[7]);
                    } else {
#pragma unroll
                      for (int e = 0; e < 8; ++e)
                        if (c0 + e < dl) {
                          hp[e] = y[e];
                          gp[e] = d[e];
                        }
                    }
                  }
                } else {
                  act8<Sel::act, Sel::fast>(v, bl + c0, y);
                }
                // (columns past dout hold act(0): they meet zero weight columns in this layer)
                store_split8(sA, kTile, r, (c0 & 63) >> 3, y, p.terms);
              }
            });
            fence_before();  // the TMEM reads above are ordered before anything that follows the arrive
          }
          fence_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(full);
          TR(11);
          mbar_wait(done, waited & 1);  // the MMAs of this step have completed: region free, and after a layer's last chunk its accumulator is complete
          ++waited;
          fence_after();
          TR(12);
        }
      }
      // ---- last layer's accumulator -> bias + activation -> global memory
      {
        const uint32_t trow = acc_base + (uint32_t)p.acc[2];
        const float* bl = bias_s + 2 * 256;
        const int act = p.act[2];
        // A lane holds eight columns of ITS row: stored directly that is one 32-byte piece per lane, 32 different rows
        // per instruction (the output pass took 7 000 cycles per row block).  The pipeline's region is free now, so the
        // tile goes through it -- rows padded by 4 floats: the eight lanes of a quarter warp hit distinct banks -- and
        // leaves in whole rows, consecutive threads writing consecutive 16 bytes.
        const int ldo = p.Np[2] + 4;
        const bool staged = (size_t)kBM * ldo * 4 <= (size_t)kRegion && vecY && (p.dout[2] % 4 == 0);
        float* stile = reinterpret_cast<float*>(sA);
        dispatch_chain_act(act, p.fast_gelu, [&](auto sel) {
          using Sel = decltype(sel);
#pragma unroll 1
          for (int c0 = 8 * part; c0 < p.Np[2]; c0 += 8 * kParts) {
            uint32_t v[8];
            tmem_ld8(trow + c0, v);
            tmem_ld_wait8(v);
            float y[8];
            act8<Sel::act, Sel::fast>(v, bl + c0, y);
            if (staged) {
              float* d = stile + r * ldo + c0;
              *reinterpret_cast<float4*>(d) = make_float4(y[0], y[1], y[2], y[3]);
              *reinterpret_cast<float4*>(d + 4) = make_float4(y[4], y[5], y[6], y[7]);
              continue;
            }
            const int64_t row = row0 + r;
            if (row < p.N && c0 < p.dout[2]) {
              float* dst = p.Y + row * p.ldy + c0;
              if (vecY && c0 + 8 <= p.dout[2]) {
                *reinterpret_cast<float4*>(dst) = make_float4(y[0], y[1], y[2], y[3]);
                *reinterpret_cast<float4*>(dst + 4) = make_float4(y[4], y[5], y[6], y[7]);
              } else {
#pragma unroll
                for (int e = 0; e < 8; ++e)
                  if (c0 + e < p.dout[2]) dst[e] = y[e];
              }
            }
          }
        });
        if (staged) {
          asm volatile("bar.sync %0, %1;" ::"r"(2 + pl), "n"(kPipeWarps * 32) : "memory");  // the tile is complete
          const int c4n = p.dout[2] >> 2;  // float4 per output row
          for (int idx = pt; idx < kBM * c4n; idx += kPipeWarps * 32) {
            const int rr = idx / c4n, c4 = idx - rr * c4n;
            const int64_t row = row0 + rr;
            if (row < p.N)
              *reinterpret_cast<float4*>(p.Y + row * p.ldy + 4 * c4) = *reinterpret_cast<const float4*>(stile + rr * ldo + 4 * c4);
          }
          asm volatile("bar.sync %0, %1;" ::"r"(2 + pl), "n"(kPipeWarps * 32) : "memory");  // read out: the region may be rewritten
        }
        TR(13);
        fence_before();  // ordered before this warp's next arrive (the next block's first step overwrites nothing it read,
                         // but the MMAs of the next block's last layer overwrite this accumulator)
      }
    }
  }

  TR(2);
  fence_before();
  __syncthreads();
  grid_dep_launch();
  if (warp == kWorkers) {
    fence_after();
    tmem_dealloc(tmem_base, p.tmem_cols);
  }
#ifdef LIPVQ_TRACE
  TR(3);
  if (blockIdx.x == 0 && lane == 0) g_mtrace_n[warp] = tr_n;
#endif
}
#undef TR

struct Mlp3Geom {
  int NC[3], Np[3], wofs[3], acc[3], acc_cols;
  size_t smem;
  uint32_t tmem_cols;
  bool ok;
};

Mlp3Geom mlp3_geom(const int din[3], const int dout[3]) {
  Mlp3Geom g{};
  int w = 0, cols = 0;
  for (int l = 0; l < 3; ++l) {
    g.NC[l] = (din[l] + kChunk - 1) / kChunk;
    g.Np[l] = (dout[l] + 15) / 16 * 16;
    g.wofs[l] = w;
    w += 3 * g.NC[l] * g.Np[l] * 128;
    g.acc[l] = cols;
    cols += g.Np[l];
  }
  g.acc_cols = cols;
  uint32_t tc_ = 32;
  while (tc_ < 2u * (uint32_t)cols) tc_ <<= 1;
  g.tmem_cols = tc_;
  g.smem = 1024 + 2 * (size_t)kRegion + (size_t)w + 3 * 256 * 4 + 128;
  g.ok = 2 * cols <= 512 && g.smem <= 227 * 1024;
  for (int l = 0; l < 3; ++l) g.ok = g.ok && din[l] >= 1 && din[l] <= 256 && dout[l] >= 1 && dout[l] <= 256;
  return g;
}

}  // namespace
}  // namespace lipvq

using namespace lipvq;

extern "C" {

int lipvq_mlp3_fwd_supported(int64_t N, int d0, int d1, int d2, int d3) {
  if (const char* f = getenv("LIPVQ_MLP3"))  // A/B switch: 0 = callers use three lipvq_linear_fwd_f32 launches
    if (f[0] == '0') return 0;
  if (const char* f = getenv("LIPVQ_LINEAR_ALGO"))
    if (f[0] == 's') return 0;
  if (N < 4096) return 0;  // (as the tensor-core linear layers: small batches stay on the FFMA tiles)
  const int din[3] = {d0, d1, d2}, dout[3] = {d1, d2, d3};
  return mlp3_geom(din, dout).ok ? 1 : 0;
}

static int mlp3_launch(int terms, float* H1, float* G1, float* H2, float* G2, const float* X, int64_t N, int d0, const float* W1, const float* b1, int d1, int act1,
                       const float* W2, const float* b2, int d2, int act2, const float* W3, const float* b3, int d3, int act3,
                       float* Y, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(N >= 0 && d0 > 0 && d1 > 0 && d2 > 0 && d3 > 0, "mlp3_fwd: bad shape");
  if (N == 0) return LIPVQ_OK;
  LIPVQ_REQUIRE(X && W1 && W2 && W3 && Y, "mlp3_fwd: null pointer");
  const int din[3] = {d0, d1, d2}, dout[3] = {d1, d2, d3};
  const Mlp3Geom g = mlp3_geom(din, dout);
  if (!g.ok) {
    set_error("mlp3_fwd: widths %d-%d-%d-%d do not fit shared memory / TMEM: use three lipvq_linear_fwd_f32 calls", d0, d1, d2, d3);
    return LIPVQ_ERR_UNSUPPORTED;
  }
  Mlp3Params p{};
  p.X = X; p.N = N; p.ldx = d0;
  p.W[0] = W1; p.W[1] = W2; p.W[2] = W3;
  p.b[0] = b1; p.b[1] = b2; p.b[2] = b3;
  p.act[0] = act1; p.act[1] = act2; p.act[2] = act3;
  for (int l = 0; l < 3; ++l) {
    p.din[l] = din[l]; p.dout[l] = dout[l]; p.NC[l] = g.NC[l]; p.Np[l] = g.Np[l]; p.wofs[l] = g.wofs[l]; p.acc[l] = g.acc[l];
  }
  p.acc_cols = g.acc_cols;
  p.Y = Y; p.ldy = d3;
  p.tmem_cols = g.tmem_cols;
  p.fast_gelu = !(getenv("LIPVQ_EXACT_GELU") != nullptr && getenv("LIPVQ_EXACT_GELU")[0] == '1');
  p.terms = terms;
  p.H[0] = H1; p.G[0] = G1; p.H[1] = H2; p.G[1] = G2;
  static size_t smem_set[kMaxDevices] = {};  // the attribute is per device
  const int dev_ = current_device();
  if (g.smem > smem_set[dev_]) {
    LIPVQ_CUDA(cudaFuncSetAttribute(mlp3_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g.smem));
    smem_set[dev_] = g.smem;
  }
  const int row_blocks = (int)ceil_div64(N, kBM);
  const int sms = sm_count();
  LIPVQ_CUDA(launch_pdl(mlp3_tc_kernel, dim3(row_blocks < sms ? row_blocks : sms), dim3(kThreads), g.smem, stream, p));
  LIPVQ_LAUNCHED("mlp3_tc_kernel");
  return LIPVQ_OK;
}

int lipvq_mlp3_fwd_f32(const float* X, int64_t N, int d0, const float* W1, const float* b1, int d1, int act1, const float* W2,
                       const float* b2, int d2, int act2, const float* W3, const float* b3, int d3, int act3, float* Y,
                       void* stream) {
  return mlp3_launch(3, nullptr, nullptr, nullptr, nullptr, X, N, d0, W1, b1, d1, act1, W2, b2, d2, act2, W3, b3, d3, act3, Y, stream);
}

int lipvq_mlp3_fwd_bf16_f32(const float* X, int64_t N, int d0, const float* W1, const float* b1, int d1, int act1, const float* W2,
                            const float* b2, int d2, int act2, const float* W3, const float* b3, int d3, int act3, float* Y,
                            void* stream) {
  return mlp3_launch(1, nullptr, nullptr, nullptr, nullptr, X, N, d0, W1, b1, d1, act1, W2, b2, d2, act2, W3, b3, d3, act3, Y, stream);
}

int lipvq_mlp3_fwd_train_f32(const float* X, int64_t N, int d0, const float* W1, const float* b1, int d1, int act1,
                             const float* W2, const float* b2, int d2, int act2, const float* W3, const float* b3, int d3,
                             int act3, float* Y, float* H1, float* G1, float* H2, float* G2, void* stream) {
  LIPVQ_REQUIRE(H1 && G1 && H2 && G2, "mlp3_fwd_train: null pointer");
  return mlp3_launch(3, H1, G1, H2, G2, X, N, d0, W1, b1, d1, act1, W2, b2, d2, act2, W3, b3, d3, act3, Y, stream);
}

#ifdef LIPVQ_TRACE
LIPVQ_API int lipvq_debug_trace_mlp3(unsigned long long* out, int* counts) {
  if (cudaMemcpyFromSymbol(out, lipvq::g_mtrace, sizeof(unsigned long long) * 32 * lipvq::kTraceMax) != cudaSuccess) return -1;
  if (cudaMemcpyFromSymbol(counts, lipvq::g_mtrace_n, sizeof(int) * 32) != cudaSuccess) return -1;
  return lipvq::kTraceMax;
}
#endif

}  // extern "C"
