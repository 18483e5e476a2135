// Deterministic, contention-free per-code sums: the codebook gradient and the EMA statistics of the tokenizer.
//
//   sums[k, :] += sum over tokens t with idx[t] == k of v[t, :]          (hq_vae.py:262-275 embed_sum / cluster_size,
//   counts[k]  += #{t : idx[t] == k}                                      algo/heirarchical_vqvae.py:518-530, and the
//                                                                          index_add of the codebook gradient)
//
// The shared-memory scatter in quantize.cu adds with atomics (order of the float additions varies run to run) and
// flushes 148 private K x D tables into the same addresses.  Here the tokens are first put in code order by a STABLE
// counting sort (integer work only: deterministic), and every code's rows are then added in ascending token order:
//
//   1. code_hist_kernel    CTA g owns a contiguous slice of tokens; histogram of its codes in shared memory
//                          (integer atomics) -> hist[g][k]
//   2. code_scan_kernel    one warp per code: exclusive scan of hist[.][k] over the slices (in place), total[k]
//   3. code_place_kernel   seg[] = exclusive scan of total[] (every CTA, in shared memory); CTA g walks its slice in
//                          token order and gives token t the position seg[k] + hist[g][k] + (rank of t among the slice's
//                          earlier tokens of code k) -> perm[position] = t.  Ranks come from match_any inside a warp and
//                          an in-order hand-over between warps.
//   4. segsum kernels      the sorted order is cut into PIECES of 16 positions; a group of lanes (one float4 each) adds
//                          the rows of its piece in order.  A run of one code that covers the code's whole segment is
//                          written straight to the output; a run cut by a piece boundary goes to a per-piece partial
//                          (first / last run of the piece)
//   5. seg_combine_kernel  codes cut by piece boundaries: their partials are added in piece order; one writer per
//                          output row, so `out[k] += sum` is deterministic too.
//
// Bit-identical results run to run and independent of the grid size of steps 4-5 (the piece size is fixed); the grid
// of steps 1-3 only changes the integer bookkeeping.  Needs D % 4 == 0, 16-byte aligned rows, K <= 8192.
#include <stdlib.h>

#include "common.cuh"

// Scratch of lipvq_ema_stats_static_f32 (end of this file): lives in the module image, so every GPU that loads the
// library holds these 8 MiB until the process exits.  In HOST code this name is only a stand-in for the symbol -- the
// device address comes from cudaGetSymbolAddress (static_scratch_ptr below), once per device.
constexpr size_t kStaticScratchBytes = (size_t)8 << 20;
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

namespace lipvq {
namespace {

constexpr int kSortThreads = 512;
constexpr int kPiece = 16;  // sorted positions per piece

__device__ __forceinline__ int clamp_code(int64_t k, int K) { return k < 0 ? 0 : (k >= K ? K - 1 : (int)k); }

__global__ void __launch_bounds__(kSortThreads)
    code_hist_kernel(const int64_t* __restrict__ idx, int64_t N, int K, int* __restrict__ hist) {
  extern __shared__ int s_hist[];
  grid_dep_wait();
  for (int k = threadIdx.x; k < K; k += blockDim.x) s_hist[k] = 0;
  __syncthreads();
  const int64_t per = (N + gridDim.x - 1) / gridDim.x;
  const int64_t beg = (int64_t)blockIdx.x * per, end = (beg + per < N) ? beg + per : N;
  for (int64_t t = beg + threadIdx.x; t < end; t += blockDim.x) atomicAdd(&s_hist[clamp_code(__ldg(idx + t), K)], 1);
  __syncthreads();
  grid_dep_launch();
  for (int k = threadIdx.x; k < K; k += blockDim.x) hist[(size_t)k * gridDim.x + blockIdx.x] = s_hist[k];  // [k][g]: the scan reads along g
}

// one warp per code: hist[g][k] <- sum_{g' < g} hist[g'][k], total[k] = sum over all slices
__global__ void __launch_bounds__(128) code_scan_kernel(int* __restrict__ hist, int G, int K, int* __restrict__ total) {
  grid_dep_wait();
  const int k = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (k >= K) return;
  int carry = 0;
  for (int g0 = 0; g0 < G; g0 += 32) {
    const int g = g0 + lane;
    const int v = (g < G) ? hist[(size_t)k * G + g] : 0;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int n = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += n;
    }
    if (g < G) hist[(size_t)k * G + g] = carry + inc - v;
    carry += __shfl_sync(0xffffffffu, inc, 31);
  }
  if (lane == 0) total[k] = carry;
}

// exclusive scan of total[0..K) into s_seg[0..K] (s_seg[K] = N) by the whole CTA; K <= 8192
__device__ __forceinline__ void block_seg_scan(const int* __restrict__ total, int K, int* s_seg, int* s_warp) {
  const int per = (K + blockDim.x - 1) / blockDim.x;  // consecutive codes per thread
  const int k0 = threadIdx.x * per;
  int sum = 0;
  for (int i = 0; i < per; ++i)
    if (k0 + i < K) sum += total[k0 + i];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int inc = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int n = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += n;
  }
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    const int nw = blockDim.x >> 5;
    int w = lane < nw ? s_warp[lane] : 0;
    int winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int n = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o) winc += n;
    }
    if (lane < nw) s_warp[lane] = winc - w;  // exclusive
  }
  __syncthreads();
  int run = s_warp[warp] + inc - sum;
  for (int i = 0; i < per; ++i)
    if (k0 + i < K) {
      s_seg[k0 + i] = run;
      run += total[k0 + i];
    }
  if (threadIdx.x == blockDim.x - 1 || k0 + per >= K) {
    if (k0 < K && k0 + per >= K) s_seg[K] = run;
  }
  __syncthreads();
}

// CTA g places the tokens of its slice, in token order.  cursor[k] starts at seg[k] + (tokens of code k in earlier
// slices) and advances as the warps hand over in order, so positions inside one code ascend with the token index.
__global__ void __launch_bounds__(kSortThreads)
    code_place_kernel(const int64_t* __restrict__ idx, int64_t N, int K, const int* __restrict__ hist,
                      const int* __restrict__ total, int* __restrict__ seg, int* __restrict__ perm, int* __restrict__ skey) {
  extern __shared__ int s_mem[];  // cursor[K + 1]
  __shared__ int s_warp[32];
  int* cursor = s_mem;
  grid_dep_wait();
  block_seg_scan(total, K, cursor, s_warp);
  if (blockIdx.x == 0)
    for (int k = threadIdx.x; k <= K; k += blockDim.x) seg[k] = cursor[k];
  __syncthreads();
  for (int k = threadIdx.x; k < K; k += blockDim.x) cursor[k] += hist[(size_t)k * gridDim.x + blockIdx.x];
  __syncthreads();
  const int64_t per = (N + gridDim.x - 1) / gridDim.x;
  const int64_t beg = (int64_t)blockIdx.x * per, end = (beg + per < N) ? beg + per : N;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int64_t t0 = beg; t0 < end; t0 += blockDim.x) {
    const int64_t t = t0 + threadIdx.x;
    const bool live = t < end;
    const int k = live ? clamp_code(__ldg(idx + t), K) : -1 - lane;  // dead lanes: distinct keys, no partners
    const unsigned peers = __match_any_sync(0xffffffffu, k);
    const int rank = __popc(peers & ((1u << lane) - 1u));
    const bool leader = rank == 0;
    const int cnt = __popc(peers);
    int base = 0;
    for (int w = 0; w < nw; ++w) {  // warps take their turn in token order
      if (warp == w && live && leader) {
        base = cursor[k];
        cursor[k] = base + cnt;
      }
      __syncthreads();
    }
    base = __shfl_sync(0xffffffffu, base, __ffs(peers) - 1);
    if (live) {
      perm[base + rank] = (int)t;
      skey[base + rank] = k;  // the code of every sorted position: the sums never gather idx again
    }
  }
}

// ---- 4. ordered sums over the pieces ------------------------------------------------------------------------------
// F::load(t, c4) -> raw operands of token t, 16-byte column c4; F::finish(raw, t, c4) -> float4 contribution (and any
// per-token side effect, e.g. the store of dz).  Loads and stores are kept in separate loops: a store between the loads
// of two rows would pin their order (the compiler must assume the pointers alias) and every row would cost a full
// memory round trip -- measured 36 us for the quantizer backward before the split.
template <class F, int B>
__global__ void __launch_bounds__(256)
    segsum_kernel(F f, const int* __restrict__ perm, const int* __restrict__ skey, const int* __restrict__ seg, int64_t N, int D,
                  int lpr, float* __restrict__ out, float* __restrict__ part_first, float* __restrict__ part_last) {
  grid_dep_wait();
  const int gpc = blockDim.x / lpr;
  const int c4 = threadIdx.x % lpr;
  const bool active = 4 * c4 < D;
  const int64_t npieces = (N + kPiece - 1) / kPiece;
  for (int64_t p = (int64_t)blockIdx.x * gpc + threadIdx.x / lpr; p < npieces; p += (int64_t)gridDim.x * gpc) {
    const int64_t p0 = p * kPiece;
    const int n = (int)((N - p0 < kPiece) ? N - p0 : kPiece);
    int tt[kPiece], tk[kPiece];
#pragma unroll
    for (int u = 0; u < kPiece; ++u) {  // the same address across the group: one broadcast transaction each
      tt[u] = (u < n) ? __ldg(perm + p0 + u) : 0;
      tk[u] = (u < n) ? __ldg(skey + p0 + u) : -1;
    }
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    int cur = -1;
    int64_t run_start = p0;
    auto flush = [&](int k, int64_t a, int64_t b) {
      if (!active) return;
      const bool complete = (a == seg[k]) && (b == seg[k + 1]);
      float* dst = complete ? out + (size_t)k * D + 4 * c4
                            : ((a == p0) ? part_first : part_last) + (size_t)p * D + 4 * c4;
      float4 w = acc;
      if (complete) {  // one writer per output row: += keeps the caller's accumulate semantics, deterministically
        const float4 o = *reinterpret_cast<const float4*>(dst);
        w.x += o.x; w.y += o.y; w.z += o.z; w.w += o.w;
      }
      *reinterpret_cast<float4*>(dst) = w;
    };
#pragma unroll
    for (int b0 = 0; b0 < kPiece; b0 += B) {
      typename F::Raw raw[B];
#pragma unroll
      for (int u = 0; u < B; ++u)
        if (active && tk[b0 + u] >= 0) raw[u] = f.load(tt[b0 + u], c4);
#pragma unroll
      for (int u = 0; u < B; ++u) {
        const int k = tk[b0 + u];
        if (k < 0) continue;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (active) v = f.finish(raw[u], tt[b0 + u], c4);
        if (k != cur) {
          if (cur >= 0) flush(cur, run_start, p0 + b0 + u);
          cur = k;
          run_start = p0 + b0 + u;
          acc = v;
        } else {
          acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
      }
    }
    if (cur >= 0) flush(cur, run_start, p0 + n);
  }
}

// ---- 5. codes cut by piece boundaries: partials added in piece order; counts --------------------------------------
__global__ void __launch_bounds__(256)
    seg_combine_kernel(const int* __restrict__ seg, int D, int K, int lpr, const float* __restrict__ part_first,
                       const float* __restrict__ part_last, float* __restrict__ out, float* __restrict__ counts) {
  grid_dep_wait();
  const int gpc = blockDim.x / lpr;
  const int c4 = threadIdx.x % lpr;
  const int k = blockIdx.x * gpc + threadIdx.x / lpr;
  if (k >= K) return;
  const int a = seg[k], b = seg[k + 1];
  if (counts != nullptr && c4 == 0) counts[k] += (float)(b - a);
  if (b <= a || out == nullptr || 4 * c4 >= D) return;
  const int pf = a / kPiece, pl = (b - 1) / kPiece;
  if (pf == pl) return;  // whole segment inside one piece: written by segsum_kernel
  const size_t col = 4 * (size_t)c4;
  float4 acc = *reinterpret_cast<const float4*>(((a % kPiece == 0) ? part_first : part_last) + (size_t)pf * D + col);
  for (int p = pf + 1; p <= pl; p += 8) {
    float4 t[8];
#pragma unroll
    for (int u = 0; u < 8; ++u)
      t[u] = (p + u <= pl) ? *reinterpret_cast<const float4*>(part_first + (size_t)(p + u) * D + col) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int u = 0; u < 8; ++u)
      if (p + u <= pl) {
        acc.x += t[u].x; acc.y += t[u].y; acc.z += t[u].z; acc.w += t[u].w;
      }
  }
  float* dst = out + (size_t)k * D + col;
  const float4 o = *reinterpret_cast<const float4*>(dst);
  *reinterpret_cast<float4*>(dst) = make_float4(o.x + acc.x, o.y + acc.y, o.z + acc.z, o.w + acc.w);
}

struct StatsFn {
  typedef float4 Raw;
  const float* z;
  int D;
  __device__ __forceinline__ Raw load(int t, int c4) const { return __ldg(reinterpret_cast<const float4*>(z + (size_t)t * D) + c4); }
  __device__ __forceinline__ float4 finish(const Raw& r, int, int) const { return r; }
};

// gather_ste_loss_bwd (quantize.cu scatter_smem_kernel<true>): per token dz, and the code-row gradient as contribution
struct BwdFn {
  struct Raw {
    float4 z, q, u, y;
  };
  const float* z;
  const float* zq;
  const float* up;
  const float* sig;     // nullptr, or sigmoid outputs whose derivative multiplies dz
  float* dz;
  const float* gscale;  // optional device scalar multiplying the loss gradient
  float coef0, w_commit, w_codebook;
  int ste, D, sig_is_z;
  __device__ __forceinline__ Raw load(int t, int c4) const {
    const size_t o = (size_t)t * D + 4 * (size_t)c4;
    Raw r;
    r.z = __ldg(reinterpret_cast<const float4*>(z + o));
    r.q = __ldg(reinterpret_cast<const float4*>(zq + o));
    r.u = up ? __ldg(reinterpret_cast<const float4*>(up + o)) : make_float4(0.f, 0.f, 0.f, 0.f);
    r.y = (sig && !sig_is_z) ? __ldg(reinterpret_cast<const float4*>(sig + o)) : r.z;
    return r;
  }
  __device__ __forceinline__ float4 finish(const Raw& r, int t, int c4) const {
    const float coef = coef0 * (gscale ? __ldg(gscale) : 1.f);  // same expression order as scatter_smem_kernel<true>
    const float cz = w_commit * coef, cq = w_codebook * coef;
    const float4 z4 = r.z, q4 = r.q, u4 = r.u;
    float4 gz, gq;
    gz.x = cz * (z4.x - q4.x); gz.y = cz * (z4.y - q4.y); gz.z = cz * (z4.z - q4.z); gz.w = cz * (z4.w - q4.w);
    gq.x = cq * (q4.x - z4.x); gq.y = cq * (q4.y - z4.y); gq.z = cq * (q4.z - z4.z); gq.w = cq * (q4.w - z4.w);
    if (ste) { gz.x += u4.x; gz.y += u4.y; gz.z += u4.z; gz.w += u4.w; }
    else     { gq.x += u4.x; gq.y += u4.y; gq.z += u4.z; gq.w += u4.w; }
    if (sig) {
      const float4 y = r.y;
      gz.x *= y.x * (1.f - y.x); gz.y *= y.y * (1.f - y.y); gz.z *= y.z * (1.f - y.z); gz.w *= y.w * (1.f - y.w);
    }
    if (dz) *reinterpret_cast<float4*>(dz + (size_t)t * D + 4 * (size_t)c4) = gz;
    return gq;
  }
};

// ---- host side ------------------------------------------------------------------------------------------------------
// The caller owns the scratch (lipvq_code_sort_workspace_bytes): slice histograms, totals, segment starts, permutation,
// sorted codes, piece partials.  The layout is a pure function of (N, D, K) and the device's SM count.
struct DetPlan {
  int G, lpr;
  int *hist, *total, *seg, *perm, *skey;
  float *part_first, *part_last;
  size_t bytes;
};

inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

DetPlan det_plan(int64_t N, int D, int K, uint8_t* base) {
  DetPlan pl;
  const int sms = sm_count();
  int G = (int)ceil_div64(N, kSortThreads);
  if (G > sms) G = sms;
  if (G < 1) G = 1;
  const int64_t npieces = ceil_div64(N, kPiece);
  const size_t b_hist = align256((size_t)G * K * 4), b_total = align256((size_t)K * 4), b_seg = align256((size_t)(K + 1) * 4),
               b_perm = align256((size_t)N * 4), b_part = align256((size_t)npieces * D * 4);
  size_t off = 0;
  pl.G = G;
  pl.hist = (int*)(base + off); off += b_hist;
  pl.total = (int*)(base + off); off += b_total;
  pl.seg = (int*)(base + off); off += b_seg;
  pl.perm = (int*)(base + off); off += b_perm;
  pl.skey = (int*)(base + off); off += b_perm;
  pl.part_first = (float*)(base + off); off += b_part;
  pl.part_last = (float*)(base + off); off += b_part;
  pl.bytes = off;
  int lpr = 1;
  while (lpr < 32 && lpr * 4 < D) lpr <<= 1;
  pl.lpr = lpr;
  return pl;
}

template <class F, int B>
int det_sum(F f, int64_t N, int D, int K, const DetPlan& pl, float* out, float* counts, cudaStream_t stream) {
  const int gpc = 256 / pl.lpr;
  const int64_t npieces = ceil_div64(N, kPiece);
  int64_t grid = ceil_div64(npieces, gpc);
  const int64_t cap = (int64_t)sm_count() * 8;
  if (grid > cap) grid = cap;
  LIPVQ_CUDA(launch_pdl(segsum_kernel<F, B>, dim3((unsigned)grid), dim3(256), 0, stream, f, (const int*)pl.perm, (const int*)pl.skey,
                        (const int*)pl.seg, N, D, pl.lpr, out, pl.part_first, pl.part_last));
  LIPVQ_LAUNCHED("segsum_kernel");
  LIPVQ_CUDA(launch_pdl(seg_combine_kernel, dim3((unsigned)ceil_div64(K, gpc)), dim3(256), 0, stream, (const int*)pl.seg, D, K,
                        pl.lpr, (const float*)pl.part_first, (const float*)pl.part_last, out, counts));
  LIPVQ_LAUNCHED("seg_combine_kernel");
  return LIPVQ_OK;
}

bool det_shape_ok(int64_t N, int D, int K) {
  return N >= 1 && N < ((int64_t)1 << 31) && D >= 4 && D % 4 == 0 && D <= 128 && K >= 1 && K <= 8192;
}

// device address of g_static_scratch on the CURRENT device (looked up once per device, with that device current)
int static_scratch_ptr(uint8_t** out) {
  static uint8_t* ptrs[kMaxDevices] = {};
  const int dev = current_device();
  if (dev < 0 || dev >= kMaxDevices) {
    set_error("ema_stats_static: device ordinal %d out of range", dev);
    return LIPVQ_ERR_INVALID;
  }
  if (ptrs[dev] == nullptr) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_static_scratch)");
    if (p == nullptr) {
      set_error("ema_stats_static: cudaGetSymbolAddress returned a null address");
      return LIPVQ_ERR_CUDA;
    }
    ptrs[dev] = (uint8_t*)p;
  }
  *out = ptrs[dev];
  return LIPVQ_OK;
}

}  // namespace
}  // namespace lipvq

using namespace lipvq;

extern "C" {

size_t lipvq_code_sort_workspace_bytes(int64_t N, int D, int K) {
  if (!det_shape_ok(N, D, K)) return 0;
  return det_plan(N, D, K, nullptr).bytes;
}

int lipvq_code_sort_i64(const int64_t* idx, int64_t N, int D, int K, void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(det_shape_ok(N, D, K), "code_sort: needs 1 <= N < 2^31, D %% 4 == 0, 4 <= D <= 128, 1 <= K <= 8192 (N=%lld D=%d K=%d)",
                (long long)N, D, K);
  LIPVQ_REQUIRE(idx && workspace && aligned16(workspace), "code_sort: null or misaligned pointer");
  const DetPlan pl = det_plan(N, D, K, (uint8_t*)workspace);
  LIPVQ_REQUIRE(workspace_bytes >= pl.bytes, "code_sort: workspace of %zu bytes, %zu needed", workspace_bytes, pl.bytes);
  LIPVQ_CUDA(launch_pdl(code_hist_kernel, dim3(pl.G), dim3(kSortThreads), (size_t)K * 4, stream, idx, N, K, pl.hist));
  LIPVQ_LAUNCHED("code_hist_kernel");
  LIPVQ_CUDA(launch_pdl(code_scan_kernel, dim3((unsigned)ceil_div64((int64_t)K * 32, 128)), dim3(128), 0, stream, pl.hist, pl.G, K,
                        pl.total));
  LIPVQ_LAUNCHED("code_scan_kernel");
  LIPVQ_CUDA(launch_pdl(code_place_kernel, dim3(pl.G), dim3(kSortThreads), (size_t)(K + 1) * 4, stream, idx, N, K,
                        (const int*)pl.hist, (const int*)pl.total, pl.seg, pl.perm, pl.skey));
  LIPVQ_LAUNCHED("code_place_kernel");
  return LIPVQ_OK;
}

int lipvq_ema_stats_sorted_f32(const float* z, int64_t N, int D, int K, void* workspace, size_t workspace_bytes, float* counts,
                               float* sums, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(det_shape_ok(N, D, K), "ema_stats_sorted: unsupported shape N=%lld D=%d K=%d", (long long)N, D, K);
  LIPVQ_REQUIRE(z && workspace && counts && sums && aligned16(z) && aligned16(sums) && aligned16(workspace),
                "ema_stats_sorted: null or misaligned pointer");
  const DetPlan pl = det_plan(N, D, K, (uint8_t*)workspace);
  LIPVQ_REQUIRE(workspace_bytes >= pl.bytes, "ema_stats_sorted: workspace of %zu bytes, %zu needed", workspace_bytes, pl.bytes);
  StatsFn f{z, D};
  return det_sum<StatsFn, 8>(f, N, D, K, pl, sums, counts, stream);
}

int lipvq_gather_ste_loss_bwd_sorted_f32(const float* z, const float* z_q, const float* dzq_up, const float* gscale_dev,
                                         float w_commit, float w_codebook, int ste, const float* sigmoid_out, int64_t N, int D,
                                         int K, void* workspace, size_t workspace_bytes, float* dz_out, float* dcodebook,
                                         void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(det_shape_ok(N, D, K), "gather_bwd_sorted: unsupported shape N=%lld D=%d K=%d", (long long)N, D, K);
  LIPVQ_REQUIRE(z && z_q && workspace && dcodebook, "gather_bwd_sorted: null pointer");
  LIPVQ_REQUIRE(aligned16(z) && aligned16(z_q) && (!dzq_up || aligned16(dzq_up)) && (!sigmoid_out || aligned16(sigmoid_out)) &&
                    (!dz_out || aligned16(dz_out)) && aligned16(dcodebook) && aligned16(workspace),
                "gather_bwd_sorted: pointers must be 16-byte aligned");
  const DetPlan pl = det_plan(N, D, K, (uint8_t*)workspace);
  LIPVQ_REQUIRE(workspace_bytes >= pl.bytes, "gather_bwd_sorted: workspace of %zu bytes, %zu needed", workspace_bytes, pl.bytes);
  BwdFn f{z, z_q, dzq_up, sigmoid_out, dz_out, gscale_dev, (float)(2.0 / ((double)N * (double)D)), w_commit, w_codebook, ste, D,
          sigmoid_out == z ? 1 : 0};
  return det_sum<BwdFn, 4>(f, N, D, K, pl, dcodebook, (float*)nullptr, stream);
}

size_t lipvq_ema_stats_static_max_bytes(void) { return kStaticScratchBytes; }

// Sort + ordered sums in one call, scratch in g_static_scratch: no cudaMalloc / cudaFree / stream-ordered allocation,
// only kernel launches.  Nothing in the scratch has to start at zero (every word the kernels read is written earlier in
// the same call), so it is not cleared; it keeps the last call's contents, and calls that overlap on the GPU race.
int lipvq_ema_stats_static_f32(const float* z, const int64_t* idx, int64_t N, int D, int K, float* counts, float* sums,
                               void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(det_shape_ok(N, D, K), "ema_stats_static: needs 1 <= N < 2^31, D %% 4 == 0, 4 <= D <= 128, 1 <= K <= 8192 (N=%lld D=%d K=%d)",
                (long long)N, D, K);
  LIPVQ_REQUIRE(z && idx && counts && sums && aligned16(z) && aligned16(sums), "ema_stats_static: null or misaligned pointer");
  const size_t need = det_plan(N, D, K, nullptr).bytes;
  LIPVQ_REQUIRE(need <= kStaticScratchBytes,
                "ema_stats_static: N=%lld D=%d K=%d needs %zu bytes of scratch, the built-in array holds %zu; use "
                "lipvq_code_sort_i64 + lipvq_ema_stats_sorted_f32 with a caller-owned workspace",
                (long long)N, D, K, need, kStaticScratchBytes);
  uint8_t* base = nullptr;
  const int rc = static_scratch_ptr(&base);
  if (rc != LIPVQ_OK) return rc;
  const DetPlan pl = det_plan(N, D, K, base);
  LIPVQ_CUDA(launch_pdl(code_hist_kernel, dim3(pl.G), dim3(kSortThreads), (size_t)K * 4, stream, idx, N, K, pl.hist));
  LIPVQ_LAUNCHED("code_hist_kernel");
  LIPVQ_CUDA(launch_pdl(code_scan_kernel, dim3((unsigned)ceil_div64((int64_t)K * 32, 128)), dim3(128), 0, stream, pl.hist, pl.G, K,
                        pl.total));
  LIPVQ_LAUNCHED("code_scan_kernel");
  LIPVQ_CUDA(launch_pdl(code_place_kernel, dim3(pl.G), dim3(kSortThreads), (size_t)(K + 1) * 4, stream, idx, N, K,
                        (const int*)pl.hist, (const int*)pl.total, pl.seg, pl.perm, pl.skey));
  LIPVQ_LAUNCHED("code_place_kernel");
  StatsFn f{z, D};
  return det_sum<StatsFn, 8>(f, N, D, K, pl, sums, counts, stream);
}

}  // extern "C"
