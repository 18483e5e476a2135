// K4 gather / straight-through / commitment loss, K5 EMA codebook statistics + update, K6 usage bookkeeping.
// HBM-bound streaming kernels: one coalesced pass over z (float4 when the row pitch allows), code rows come from
// L2 (the codebook is re-read by every token), results leave in one coalesced pass.
//
// Reference anchors: backbone_lfqvae_v5.py:47,74,80-81 (gather, detach-copy, two MSE terms); backbone.py:66-74
// (STE); hq_vae.py:103-130,262-279 and heirarchical_vqvae.py:498-541 (EMA counts/sums/update);
// heirarchical_vqvae.py:687-706 (usage summary).
#include <stdlib.h>

#include "common.cuh"

namespace lipvq {
namespace {

// ---- K4 forward -------------------------------------------------------------------------------------------------
// VEC: D % 4 == 0 and 16-byte aligned bases -> one float4 per lane-step.  A group of `lanes_per_row` lanes owns
// a row; groups stride over rows (grid-stride), so a warp reads 32*16 contiguous bytes of z per step when D >= 128.
template <bool VEC>
__global__ void __launch_bounds__(256)
    gather_fwd_kernel(const float* __restrict__ z, const int64_t* __restrict__ idx, const float* __restrict__ cb,
                      int64_t N, int D, int K, float* __restrict__ zq, float* __restrict__ zst,
                      double* __restrict__ sqerr) {
  grid_dep_wait();  // programmatic dependent launch: nothing above touches memory
  grid_dep_launch();
  constexpr int W = VEC ? 4 : 1;
  const int64_t total = N * (int64_t)(D / W);  // work items: one vector (or scalar) element each
  const int per_row = D / W;
  double local = 0.0;
  float part = 0.f;
  int cnt = 0;
  for (int64_t it = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; it < total; it += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = it / per_row;
    const int c = (int)(it - row * per_row) * W;
    int64_t k = __ldg(idx + row);
    k = k < 0 ? 0 : (k >= K ? K - 1 : k);  // defensive clamp; indices come from our own search
    if (VEC) {
      const float4 zv = __ldg(reinterpret_cast<const float4*>(z + row * D + c));
      const float4 qv = __ldg(reinterpret_cast<const float4*>(cb + k * D + c));
      *reinterpret_cast<float4*>(zq + row * D + c) = qv;
      const float dx = qv.x - zv.x, dy = qv.y - zv.y, dz = qv.z - zv.z, dw = qv.w - zv.w;
      if (zst) {
        // z + (z_q - z): same two roundings as the reference's STE expression (backbone.py:74)
        float4 st;
        st.x = __fadd_rn(zv.x, dx); st.y = __fadd_rn(zv.y, dy); st.z = __fadd_rn(zv.z, dz); st.w = __fadd_rn(zv.w, dw);
        *reinterpret_cast<float4*>(zst + row * D + c) = st;
      }
      part = fmaf(dx, dx, part); part = fmaf(dy, dy, part); part = fmaf(dz, dz, part); part = fmaf(dw, dw, part);
    } else {
      const float zv = __ldg(z + row * D + c);
      const float qv = __ldg(cb + k * D + c);
      zq[row * D + c] = qv;
      const float dx = qv - zv;
      if (zst) zst[row * D + c] = __fadd_rn(zv, dx);
      part = fmaf(dx, dx, part);
    }
    if (++cnt == 64) {  // flush the fp32 partial into the double accumulator before it loses bits
      local += (double)part;
      part = 0.f;
      cnt = 0;
    }
  }
  local += (double)part;
  block_atomic_add_double(local, sqerr);
}

// ---- K4 backward ------------------------------------------------------------------------------------------------
// Sub-warp groups of G lanes own a row (lane gl holds float4 columns (gl + G*v)*4) and walk a contiguous run of rows.
// The codebook-gradient contribution of consecutive rows that selected the same code is summed in registers and
// flushed with one vector red per element when the code changes: consecutive action steps usually share a code, and
// a collapsed codebook (every token on a handful of codes) would otherwise serialise N*D/4 atomics on a few lines.
template <int G, int VPL>
__global__ void __launch_bounds__(256)
    gather_bwd_runs_kernel(const float* __restrict__ z, const float* __restrict__ zq, const int64_t* __restrict__ idx,
                           const float* __restrict__ dzq_up, const float* __restrict__ gscale_dev, float w_commit,
                           float w_codebook, int ste, const float* __restrict__ sig_out, int64_t N, int D, int K,
                           int64_t rows_per_group, float* __restrict__ dz, float* __restrict__ dcb) {
  const int64_t gid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
  const int gl = threadIdx.x % G;
  const int64_t r0 = gid * rows_per_group;
  if (r0 >= N) return;
  const int64_t r1 = (r0 + rows_per_group < N) ? r0 + rows_per_group : N;
  const float g = gscale_dev ? __ldg(gscale_dev) : 1.f;
  const float coef = (float)(2.0 / ((double)N * (double)D)) * g;
  const float cz = w_commit * coef, cq = w_codebook * coef;
  float4 acc[VPL];
#pragma unroll
  for (int v = 0; v < VPL; ++v) acc[v] = make_float4(0.f, 0.f, 0.f, 0.f);
  int64_t cur = -1;
  for (int64_t r = r0; r < r1; ++r) {
    int64_t k = __ldg(idx + r);
    k = k < 0 ? 0 : (k >= K ? K - 1 : k);
    if (k != cur) {
      if (cur >= 0 && dcb) {
#pragma unroll
        for (int v = 0; v < VPL; ++v) {
          const int c = (gl + G * v) * 4;
          if (c < D) atomicAdd(reinterpret_cast<float4*>(dcb + cur * D + c), acc[v]);  // red.global.add.v4.f32
          acc[v] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
      cur = k;
    }
#pragma unroll
    for (int v = 0; v < VPL; ++v) {
      const int c = (gl + G * v) * 4;
      if (c < D) {
        const int64_t o = r * D + c;
        const float4 zv = __ldg(reinterpret_cast<const float4*>(z + o));
        const float4 qv = __ldg(reinterpret_cast<const float4*>(zq + o));
        float4 up = make_float4(0.f, 0.f, 0.f, 0.f);
        if (dzq_up) up = __ldg(reinterpret_cast<const float4*>(dzq_up + o));
        float4 gz, gq;
        gz.x = cz * (zv.x - qv.x); gz.y = cz * (zv.y - qv.y); gz.z = cz * (zv.z - qv.z); gz.w = cz * (zv.w - qv.w);
        gq.x = cq * (qv.x - zv.x); gq.y = cq * (qv.y - zv.y); gq.z = cq * (qv.z - zv.z); gq.w = cq * (qv.w - zv.w);
        if (ste) { gz.x += up.x; gz.y += up.y; gz.z += up.z; gz.w += up.w; }
        else     { gq.x += up.x; gq.y += up.y; gq.z += up.z; gq.w += up.w; }
        if (sig_out) {
          const float4 y = __ldg(reinterpret_cast<const float4*>(sig_out + o));
          gz.x *= y.x * (1.f - y.x); gz.y *= y.y * (1.f - y.y); gz.z *= y.z * (1.f - y.z); gz.w *= y.w * (1.f - y.w);
        }
        if (dz) *reinterpret_cast<float4*>(dz + o) = gz;
        acc[v].x += gq.x; acc[v].y += gq.y; acc[v].z += gq.z; acc[v].w += gq.w;
      }
    }
  }
  if (cur >= 0 && dcb) {
#pragma unroll
    for (int v = 0; v < VPL; ++v) {
      const int c = (gl + G * v) * 4;
      if (c < D) atomicAdd(reinterpret_cast<float4*>(dcb + cur * D + c), acc[v]);
    }
  }
}

// scalar fallback (D % 4 != 0 or unaligned bases)
__global__ void __launch_bounds__(256)
    gather_bwd_scalar_kernel(const float* __restrict__ z, const float* __restrict__ zq,
                             const int64_t* __restrict__ idx, const float* __restrict__ dzq_up,
                             const float* __restrict__ gscale_dev, float w_commit, float w_codebook, int ste,
                             const float* __restrict__ sig_out, int64_t N, int D, int K, float* __restrict__ dz,
                             float* __restrict__ dcb) {
  const int64_t total = N * (int64_t)D;
  const float g = gscale_dev ? __ldg(gscale_dev) : 1.f;
  const float coef = (float)(2.0 / ((double)N * (double)D)) * g;
  const float cz = w_commit * coef, cq = w_codebook * coef;
  for (int64_t o = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; o < total; o += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = o / D;
    const int c = (int)(o - row * D);
    int64_t k = __ldg(idx + row);
    k = k < 0 ? 0 : (k >= K ? K - 1 : k);
    const float zv = __ldg(z + o), qv = __ldg(zq + o);
    const float up = dzq_up ? __ldg(dzq_up + o) : 0.f;
    float gz = cz * (zv - qv), gq = cq * (qv - zv);
    if (ste) gz += up; else gq += up;
    if (sig_out) {
      const float y = __ldg(sig_out + o);
      gz *= y * (1.f - y);
    }
    if (dz) dz[o] = gz;
    if (dcb) atomicAdd(dcb + k * D + c, gq);
  }
}

// ---- shared-memory privatised scatter (K4 backward / K5 statistics) ------------------------------------------------
// When the whole [K, D] table fits in shared memory (K*D <= ~50 K floats: every tokenizer codebook of the reference),
// each CTA (one per SM, a contiguous slice of rows) accumulates its rows into a private copy with shared-memory
// reductions and adds only the touched entries to global memory once at the end.  A collapsed codebook -- every token
// on a handful of codes, the normal state of a fresh EMA quantizer -- then costs ~#SM global reductions per entry
// instead of N, which is what made the run-length kernels above L2-atomic-bound (profiles/r1g_launches.md).
// Lane mapping: LPR = pow2 >= D/4 lanes own one row (one float4 each, or a strided loop for D > 128).
constexpr int kScatThreads = 512;

// Adds this CTA's table to the global one.  Every CTA holds the SAME K*D addresses, so the walk starts at a
// CTA-dependent offset (the CTAs would otherwise hit each address at the same moment) and moves 16 bytes per reduction.
__device__ __forceinline__ void scatter_flush(const float* s_tab, int64_t n, float* __restrict__ g) {
  if ((n & 3) == 0 && (reinterpret_cast<uintptr_t>(g) & 15) == 0) {
    const int64_t n4 = n >> 2;
    const int64_t start = (n4 / gridDim.x) * blockIdx.x;
    for (int64_t i = threadIdx.x; i < n4; i += blockDim.x) {
      int64_t j = i + start;
      if (j >= n4) j -= n4;
      const float4 v = reinterpret_cast<const float4*>(s_tab)[j];
      if (v.x != 0.f || v.y != 0.f || v.z != 0.f || v.w != 0.f) atomicAdd(reinterpret_cast<float4*>(g) + j, v);
    }
    return;
  }
  for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
    const float v = s_tab[i];
    if (v != 0.f) atomicAdd(g + i, v);
  }
}

template <bool BWD>
__global__ void __launch_bounds__(kScatThreads, 1)
    scatter_smem_kernel(const float* __restrict__ z, const float* __restrict__ zq, const int64_t* __restrict__ idx,
                        const float* __restrict__ dzq_up, const float* __restrict__ gscale_dev, float w_commit,
                        float w_codebook, int ste, const float* __restrict__ sig_out, int64_t N, int D, int K, int lpr,
                        float* __restrict__ dz, float* __restrict__ tab_g, float* __restrict__ counts_g) {
  grid_dep_wait();  // programmatic dependent launch: nothing above touches memory
  extern __shared__ __align__(16) float s_tab[];  // [K*D] (+ [K] counts for the statistics)
  float* s_cnt = s_tab + (size_t)K * D;
  const int64_t tab_n = (int64_t)K * D + (BWD ? 0 : K);
  for (int64_t i = threadIdx.x; i < tab_n; i += blockDim.x) s_tab[i] = 0.f;
  __syncthreads();

  float cz = 0.f, cq = 0.f;
  if (BWD) {
    const float g = gscale_dev ? __ldg(gscale_dev) : 1.f;
    const float coef = (float)(2.0 / ((double)N * (double)D)) * g;
    cz = w_commit * coef;
    cq = w_codebook * coef;
  }
  const int64_t per_cta = (N + gridDim.x - 1) / gridDim.x;
  const int64_t r_beg = (int64_t)blockIdx.x * per_cta;
  const int64_t r_end = (r_beg + per_cta < N) ? r_beg + per_cta : N;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const int rpw = 32 / lpr;                 // rows per warp pass
  const int sub = lane / lpr, c4 = lane % lpr;
  const int nvec = D >> 2;
  // UN row passes per iteration with every load issued up front: the loop is otherwise one DRAM latency per pass
  constexpr int UN = 4;
  const int64_t pass = (int64_t)nwarps * rpw;
  for (int64_t r0 = r_beg + (int64_t)warp * rpw; r0 < r_end; r0 += pass * UN) {
    for (int v0 = 0; v0 < nvec; v0 += lpr) {  // one trip for D <= 128
      const int v = v0 + c4;
      int64_t kk[UN];
      float4 zv[UN], qv[UN], up[UN], yv[UN];
      bool ok[UN];
#pragma unroll
      for (int u = 0; u < UN; ++u) {
        const int64_t r = r0 + u * pass + sub;
        ok[u] = (r < r_end) && (v < nvec);
        kk[u] = 0;
        zv[u] = qv[u] = up[u] = yv[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (ok[u]) {
          const int64_t o = r * D + 4 * v;
          kk[u] = __ldg(idx + r);
          zv[u] = __ldg(reinterpret_cast<const float4*>(z + o));
          if (BWD) {
            qv[u] = __ldg(reinterpret_cast<const float4*>(zq + o));
            if (dzq_up) up[u] = __ldg(reinterpret_cast<const float4*>(dzq_up + o));
            if (sig_out) yv[u] = __ldg(reinterpret_cast<const float4*>(sig_out + o));
          }
        }
      }
#pragma unroll
      for (int u = 0; u < UN; ++u) {
        if (!ok[u]) continue;
        const int64_t r = r0 + u * pass + sub;
        const int64_t o = r * D + 4 * v;
        const int64_t k = kk[u] < 0 ? 0 : (kk[u] >= K ? K - 1 : kk[u]);
        if (!BWD && v == 0) atomicAdd(s_cnt + k, 1.f);
        float4 add = zv[u];
        if (BWD) {
          const float4 z4 = zv[u], q4 = qv[u];
          float4 gz, gq;
          gz.x = cz * (z4.x - q4.x); gz.y = cz * (z4.y - q4.y); gz.z = cz * (z4.z - q4.z); gz.w = cz * (z4.w - q4.w);
          gq.x = cq * (q4.x - z4.x); gq.y = cq * (q4.y - z4.y); gq.z = cq * (q4.z - z4.z); gq.w = cq * (q4.w - z4.w);
          if (ste) { gz.x += up[u].x; gz.y += up[u].y; gz.z += up[u].z; gz.w += up[u].w; }
          else     { gq.x += up[u].x; gq.y += up[u].y; gq.z += up[u].z; gq.w += up[u].w; }
          if (sig_out) {
            const float4 y = yv[u];
            gz.x *= y.x * (1.f - y.x); gz.y *= y.y * (1.f - y.y); gz.z *= y.z * (1.f - y.z); gz.w *= y.w * (1.f - y.w);
          }
          if (dz) *reinterpret_cast<float4*>(dz + o) = gz;
          add = gq;
        }
        if (!BWD || tab_g) {
          float* t = s_tab + k * D + 4 * v;
          atomicAdd(t, add.x); atomicAdd(t + 1, add.y); atomicAdd(t + 2, add.z); atomicAdd(t + 3, add.w);
        }
      }
    }
  }
  __syncthreads();
  grid_dep_launch();  // only the flush is left: the next kernel may be scheduled (it waits for our completion at its top)
  if (tab_g) scatter_flush(s_tab, (int64_t)K * D, tab_g);
  if (!BWD) scatter_flush(s_cnt, K, counts_g);
}

// shared-memory table fits?  (leave room for the driver's reserved kilobyte)
inline bool scatter_fits(int K, int D, bool with_counts) {
  return ((size_t)K * D + (with_counts ? K : 0)) * sizeof(float) <= 200 * 1024;
}
inline int scatter_lpr(int D) {
  int lpr = 1;
  while (lpr < 32 && lpr * 4 < D) lpr <<= 1;
  return lpr;
}

// ---- K5 statistics ----------------------------------------------------------------------------------------------
// Same sub-warp-group walk as above: the running sum of the current code lives in registers and is flushed with one
// vector red per element (and one red for the count) only when the code changes.
template <int G, int VPL>
__global__ void __launch_bounds__(256)
    ema_stats_runs_kernel(const float* __restrict__ z, const int64_t* __restrict__ idx, int64_t N, int D, int K,
                          int64_t rows_per_group, float* __restrict__ counts, float* __restrict__ sums) {
  const int64_t gid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
  const int gl = threadIdx.x % G;
  const int64_t r0 = gid * rows_per_group;
  if (r0 >= N) return;
  const int64_t r1 = (r0 + rows_per_group < N) ? r0 + rows_per_group : N;
  float4 acc[VPL];
#pragma unroll
  for (int v = 0; v < VPL; ++v) acc[v] = make_float4(0.f, 0.f, 0.f, 0.f);
  int64_t cur = -1;
  float run = 0.f;
  for (int64_t r = r0; r < r1; ++r) {
    int64_t k = __ldg(idx + r);
    k = k < 0 ? 0 : (k >= K ? K - 1 : k);
    if (k != cur) {
      if (cur >= 0) {
#pragma unroll
        for (int v = 0; v < VPL; ++v) {
          const int c = (gl + G * v) * 4;
          if (c < D) atomicAdd(reinterpret_cast<float4*>(sums + cur * D + c), acc[v]);
          acc[v] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (gl == 0) atomicAdd(counts + cur, run);
      }
      cur = k;
      run = 0.f;
    }
    run += 1.f;
#pragma unroll
    for (int v = 0; v < VPL; ++v) {
      const int c = (gl + G * v) * 4;
      if (c < D) {
        const float4 zv = __ldg(reinterpret_cast<const float4*>(z + r * D + c));
        acc[v].x += zv.x; acc[v].y += zv.y; acc[v].z += zv.z; acc[v].w += zv.w;
      }
    }
  }
  if (cur >= 0) {
#pragma unroll
    for (int v = 0; v < VPL; ++v) {
      const int c = (gl + G * v) * 4;
      if (c < D) atomicAdd(reinterpret_cast<float4*>(sums + cur * D + c), acc[v]);
    }
    if (gl == 0) atomicAdd(counts + cur, run);
  }
}

// Picks the group width for a row of D/4 float4 vectors and the rows each group walks; calls f.template
// operator()<G, VPL>(grid, rows_per_group).
struct RunsPlan { int G, VPL; int64_t rows_per_group; unsigned grid; };
inline RunsPlan plan_runs(int64_t N, int D) {
  const int nvec = D / 4;
  RunsPlan p;
  p.G = nvec <= 8 ? 8 : (nvec <= 16 ? 16 : 32);
  p.VPL = (nvec + p.G - 1) / p.G;
  if (p.VPL == 3) p.VPL = 4;
  const int64_t resident_groups = (int64_t)sm_count() * 2048 / p.G;
  int64_t rpg = ceil_div64(N, resident_groups);
  if (rpg < 8) rpg = 8;  // at least 8 consecutive rows so run-length aggregation has something to merge
  p.rows_per_group = rpg;
  const int64_t groups = ceil_div64(N, rpg);
  p.grid = (unsigned)ceil_div64(groups * p.G, 256);
  return p;
}

// generic fallback (any D / alignment): one thread per element, plain scalar atomics
__global__ void __launch_bounds__(256)
    ema_stats_scalar_kernel(const float* __restrict__ z, const int64_t* __restrict__ idx, int64_t N, int D, int K,
                            float* __restrict__ counts, float* __restrict__ sums) {
  const int64_t total = N * (int64_t)D;
  for (int64_t it = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; it < total; it += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = it / D;
    const int c = (int)(it - row * D);
    int64_t k = __ldg(idx + row);
    k = k < 0 ? 0 : (k >= K ? K - 1 : k);
    atomicAdd(sums + k * D + c, __ldg(z + it));
    if (c == 0) atomicAdd(counts + k, 1.f);
  }
}

// ---- K5 update --------------------------------------------------------------------------------------------------
// pass 1: size <- decay*size + (1-decay)*counts ; total = sum(size) (Laplace variant needs it)
__global__ void __launch_bounds__(256)
    ema_size_kernel(const float* __restrict__ counts, int K, float decay, float omd, float* __restrict__ ema_size,
                    double* __restrict__ total) {
  grid_dep_wait();  // programmatic dependent launch: nothing above touches memory
  grid_dep_launch();
  double local = 0.0;
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < K; k += gridDim.x * blockDim.x) {
    const float s = fmaf(omd, counts[k], ema_size[k] * decay);  // mul_(decay).add_(x, alpha=1-decay)
    ema_size[k] = s;
    local += (double)s;
  }
  block_atomic_add_double(local, total);
}

// pass 2: one warp per code row
__global__ void __launch_bounds__(256)
    ema_codebook_kernel(const float* __restrict__ sums, int K, int D, float decay, float omd, float eps, int variant,
                        const float* __restrict__ ema_size, const double* __restrict__ total,
                        float* __restrict__ ema_sum, float* __restrict__ codebook) {
  grid_dep_wait();  // programmatic dependent launch: nothing above touches memory
  grid_dep_launch();
  const int k = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (k >= K) return;
  const float size = ema_size[k];
  float denom;
  if (variant == LIPVQ_EMA_LAPLACE) {
    const float n = (float)(*total);
    denom = (size + eps) / (n + (float)K * eps);
  } else {
    denom = size + eps;
  }
  float* es = ema_sum + (size_t)k * D;
  const float* sm = sums + (size_t)k * D;
  float* cbk = codebook + (size_t)k * D;
  float nrm = 0.f;
  for (int d = lane; d < D; d += 32) {
    const float e = fmaf(omd, sm[d], es[d] * decay);
    es[d] = e;
    const float c = e / denom;
    cbk[d] = c;
    nrm = fmaf(c, c, nrm);
  }
  if (variant == LIPVQ_EMA_NORMALIZE) {
    nrm = warp_sum(nrm);
    const float dn = fmaxf(sqrtf(nrm), 1e-12f);
    __syncwarp();
    for (int d = lane; d < D; d += 32) cbk[d] = cbk[d] / dn;
  }
}

// ---- K6 ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024)
    usage_stats_kernel(const float* __restrict__ counts, int K, float* __restrict__ usage_counts,
                       float* __restrict__ usage_ma, float* __restrict__ entropy_ma) {
  __shared__ double s_red[32];
  __shared__ double s_total;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  double tot = 0.0;
  for (int k = threadIdx.x; k < K; k += blockDim.x) tot += (double)counts[k];
  tot = warp_sum_d(tot);
  if (lane == 0) s_red[w] = tot;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) t += s_red[i];
    s_total = t;
  }
  __syncthreads();
  const float denom = (float)s_total + 1e-8f;
  double ent = 0.0;
  for (int k = threadIdx.x; k < K; k += blockDim.x) {
    const float c = counts[k];
    usage_counts[k] += c;
    usage_ma[k] = 0.99f * usage_ma[k] + 0.01f * (c > 0.f ? 1.f : 0.f);
    const float p = c / denom;
    ent += (double)(p * logf(p + 1e-8f));
  }
  ent = warp_sum_d(ent);
  __syncthreads();
  if (lane == 0) s_red[w] = ent;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) t += s_red[i];
    entropy_ma[0] = 0.99f * entropy_ma[0] + 0.01f * (float)(-t);
  }
}

__global__ void __launch_bounds__(256)
    codebook_usage_kernel(const float* __restrict__ size, int K, float dead_thr, int32_t* __restrict__ out2) {
  int used = 0, dead = 0;
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < K; k += gridDim.x * blockDim.x) {
    const float s = size[k];
    used += s > 0.f;
    dead += s < dead_thr;
  }
  used = __reduce_add_sync(0xffffffffu, used);
  dead = __reduce_add_sync(0xffffffffu, dead);
  if ((threadIdx.x & 31) == 0) {
    if (used) atomicAdd(out2 + 0, used);
    if (dead) atomicAdd(out2 + 1, dead);
  }
}

// ---- MSE (+ gradient) and the loss tail ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
    mse_kernel(const float* __restrict__ y, const float* __restrict__ t, int64_t n, const float* __restrict__ gscale_dev,
               int act_out, double* __restrict__ sqerr, float* __restrict__ dy) {
  grid_dep_wait();  // programmatic dependent launch: nothing above touches memory
  grid_dep_launch();
  const float g = gscale_dev ? __ldg(gscale_dev) : 1.f;
  const float coef = (float)(2.0 / (double)n) * g;
  double local = 0.0;
  float part = 0.f;
  int cnt = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float yv = y[i];
    const float d = yv - t[i];
    part = fmaf(d, d, part);
    if (dy) {
      float gr = coef * d;
      if (act_out == LIPVQ_ACT_RELU && !(yv > 0.f)) gr = 0.f;
      dy[i] = gr;
    }
    if (++cnt == 64) { local += (double)part; part = 0.f; cnt = 0; }
  }
  local += (double)part;
  block_atomic_add_double(local, sqerr);
}

__global__ void t1_loss_finalize_kernel(const double* __restrict__ sq2, double n_recon, double n_lat, float w_commit,
                                        float w_cb, float* __restrict__ out4) {
  grid_dep_wait();  // programmatic dependent launch: nothing above touches memory
  grid_dep_launch();
  const float recon = (float)(sq2[0] / n_recon);
  const float q = (float)(sq2[1] / n_lat);
  // loss = recon_loss + 0.25 * commitment_loss + 0.25 * codebook_loss, left to right (backbone_lfqvae_v5.py:83)
  const float loss = __fadd_rn(__fadd_rn(recon, __fmul_rn(w_commit, q)), __fmul_rn(w_cb, q));
  out4[0] = loss;
  out4[1] = recon;
  out4[2] = q;
  out4[3] = q;
}

inline int stream_grid(int64_t work_items, int threads, int per_sm) {
  const int64_t want = ceil_div64(work_items, threads);
  const int64_t cap = (int64_t)sm_count() * per_sm;
  return (int)(want < 1 ? 1 : (want < cap ? want : cap));
}

}  // namespace
}  // namespace lipvq

using namespace lipvq;

extern "C" {

int lipvq_gather_ste_loss_fwd_f32(const float* z, const int64_t* idx, const float* codebook, int64_t N, int D, int K,
                                  float* z_q_out, float* z_st_out, double* sqerr_sum, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(N >= 0 && D > 0 && K > 0, "gather_fwd: bad shape");
  if (N == 0) return LIPVQ_OK;
  LIPVQ_REQUIRE(z && idx && codebook && z_q_out && sqerr_sum, "gather_fwd: null pointer");
  const bool vec = (D % 4 == 0) && aligned16(z) && aligned16(codebook) && aligned16(z_q_out) &&
                   (!z_st_out || aligned16(z_st_out));
  const int64_t items = N * (int64_t)(vec ? D / 4 : D);
  const int grid = stream_grid(items, 256, 8);
  if (vec)
    LIPVQ_CUDA(launch_pdl(gather_fwd_kernel<true>, dim3(grid), dim3(256), 0, stream, z, idx, codebook, N, D, K, z_q_out, z_st_out, sqerr_sum));
  else
    LIPVQ_CUDA(launch_pdl(gather_fwd_kernel<false>, dim3(grid), dim3(256), 0, stream, z, idx, codebook, N, D, K, z_q_out, z_st_out, sqerr_sum));
  LIPVQ_LAUNCHED("gather_fwd_kernel");
  return LIPVQ_OK;
}

int lipvq_gather_ste_loss_bwd_f32(const float* z, const float* z_q, const int64_t* idx, const float* dzq_up,
                                  const float* gscale_dev, float w_commit, float w_codebook, int ste,
                                  const float* sigmoid_out, int64_t N, int D, int K, float* dz_out,
                                  float* dcodebook, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(N >= 0 && D > 0 && K > 0, "gather_bwd: bad shape");
  if (N == 0) return LIPVQ_OK;
  LIPVQ_REQUIRE(z && z_q && idx, "gather_bwd: null pointer");
  const bool vec = (D % 4 == 0) && aligned16(z) && aligned16(z_q) && (!dzq_up || aligned16(dzq_up)) &&
                   (!sigmoid_out || aligned16(sigmoid_out)) && (!dz_out || aligned16(dz_out)) &&
                   (!dcodebook || aligned16(dcodebook));
  if (vec && scatter_fits(K, D, false) && N >= 4096) {
    const size_t smem = (size_t)K * D * sizeof(float);
    static size_t smem_set[kMaxDevices] = {};  // the attribute is per device
  const int dev_ = current_device();
    if (smem > smem_set[dev_]) {
      LIPVQ_CUDA(cudaFuncSetAttribute(scatter_smem_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      smem_set[dev_] = smem;
    }
    LIPVQ_CUDA(launch_pdl(scatter_smem_kernel<true>, dim3(sm_count()), dim3(kScatThreads), smem, stream, z, z_q, idx, dzq_up, gscale_dev, w_commit, w_codebook, ste, sigmoid_out, N, D, K, scatter_lpr(D), dz_out, dcodebook, nullptr));
  } else if (vec && D <= 512) {
    const RunsPlan rp = plan_runs(N, D);
#define LIPVQ_GB(G_, V_)                                                                                          \
  gather_bwd_runs_kernel<G_, V_><<<rp.grid, 256, 0, stream>>>(z, z_q, idx, dzq_up, gscale_dev, w_commit, w_codebook, \
                                                              ste, sigmoid_out, N, D, K, rp.rows_per_group, dz_out,  \
                                                              dcodebook)
    if (rp.G == 8) LIPVQ_GB(8, 1);
    else if (rp.G == 16) LIPVQ_GB(16, 1);
    else if (rp.VPL == 1) LIPVQ_GB(32, 1);
    else if (rp.VPL == 2) LIPVQ_GB(32, 2);
    else LIPVQ_GB(32, 4);
#undef LIPVQ_GB
  } else {
    gather_bwd_scalar_kernel<<<stream_grid(N * (int64_t)D, 256, 8), 256, 0, stream>>>(
        z, z_q, idx, dzq_up, gscale_dev, w_commit, w_codebook, ste, sigmoid_out, N, D, K, dz_out, dcodebook);
  }
  LIPVQ_LAUNCHED("gather_bwd_kernel");
  return LIPVQ_OK;
}

int lipvq_ema_stats_f32(const float* z, const int64_t* idx, int64_t N, int D, int K, float* counts, float* sums,
                        void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(N >= 0 && D > 0 && K > 0, "ema_stats: bad shape");
  if (N == 0) return LIPVQ_OK;
  LIPVQ_REQUIRE(z && idx && counts && sums, "ema_stats: null pointer");
  const bool vec = (D % 4 == 0) && D <= 512 && aligned16(z) && aligned16(sums);
  if (vec && scatter_fits(K, D, true) && N >= 4096) {
    const size_t smem = ((size_t)K * D + K) * sizeof(float);
    static size_t smem_set[kMaxDevices] = {};  // the attribute is per device
  const int dev_ = current_device();
    if (smem > smem_set[dev_]) {
      LIPVQ_CUDA(cudaFuncSetAttribute(scatter_smem_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      smem_set[dev_] = smem;
    }
    LIPVQ_CUDA(launch_pdl(scatter_smem_kernel<false>, dim3(sm_count()), dim3(kScatThreads), smem, stream, z, nullptr, idx, nullptr, nullptr, 0.f, 0.f, 0, nullptr, N, D, K, scatter_lpr(D), nullptr, sums, counts));
  } else if (vec) {
    const RunsPlan rp = plan_runs(N, D);
#define LIPVQ_ES(G_, V_) \
  ema_stats_runs_kernel<G_, V_><<<rp.grid, 256, 0, stream>>>(z, idx, N, D, K, rp.rows_per_group, counts, sums)
    if (rp.G == 8) LIPVQ_ES(8, 1);
    else if (rp.G == 16) LIPVQ_ES(16, 1);
    else if (rp.VPL == 1) LIPVQ_ES(32, 1);
    else if (rp.VPL == 2) LIPVQ_ES(32, 2);
    else LIPVQ_ES(32, 4);
#undef LIPVQ_ES
  } else {
    ema_stats_scalar_kernel<<<stream_grid(N * (int64_t)D, 256, 8), 256, 0, stream>>>(z, idx, N, D, K, counts, sums);
  }
  LIPVQ_LAUNCHED("ema_stats_kernel");
  return LIPVQ_OK;
}

int lipvq_ema_apply_f32(const float* counts, const float* sums, int K, int D, float decay, float one_minus_decay,
                        float eps, int variant, float* ema_size, float* ema_sum, float* codebook, void* scratch,
                        void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(K > 0 && D > 0, "ema_apply: bad shape");
  LIPVQ_REQUIRE(counts && sums && ema_size && ema_sum && codebook && scratch, "ema_apply: null pointer");
  LIPVQ_REQUIRE(variant >= LIPVQ_EMA_SIMPLE && variant <= LIPVQ_EMA_NORMALIZE, "ema_apply: bad variant %d", variant);
  LIPVQ_REQUIRE(((uintptr_t)scratch & 7u) == 0, "ema_apply: scratch must be 8-byte aligned");
  double* total = (double*)scratch;
  LIPVQ_CUDA(cudaMemsetAsync(total, 0, sizeof(double), stream));
  // (normal launch: its predecessor in the stream is the memset above, not a kernel)
  ema_size_kernel<<<stream_grid(K, 256, 4), 256, 0, stream>>>(counts, K, decay, one_minus_decay, ema_size, total);
  LIPVQ_LAUNCHED("ema_size_kernel");
  LIPVQ_CUDA(launch_pdl(ema_codebook_kernel, dim3((unsigned)ceil_div64((int64_t)K * 32, 256)), dim3(256), 0, stream, sums, K, D, decay, one_minus_decay, eps, variant, ema_size, total, ema_sum, codebook));
  LIPVQ_LAUNCHED("ema_codebook_kernel");
  return LIPVQ_OK;
}

int lipvq_usage_stats_f32(const float* counts, int K, float* usage_counts, float* usage_ma, float* entropy_ma,
                          void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(K > 0 && counts && usage_counts && usage_ma && entropy_ma, "usage_stats: bad argument");
  usage_stats_kernel<<<1, 1024, 0, stream>>>(counts, K, usage_counts, usage_ma, entropy_ma);
  LIPVQ_LAUNCHED("usage_stats_kernel");
  return LIPVQ_OK;
}

int lipvq_codebook_usage_f32(const float* cluster_size, int K, float dead_thr, int32_t* out2, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(K > 0 && cluster_size && out2, "codebook_usage: bad argument");
  LIPVQ_CUDA(cudaMemsetAsync(out2, 0, 2 * sizeof(int32_t), stream));
  codebook_usage_kernel<<<stream_grid(K, 256, 2), 256, 0, stream>>>(cluster_size, K, dead_thr, out2);
  LIPVQ_LAUNCHED("codebook_usage_kernel");
  return LIPVQ_OK;
}

int lipvq_mse_fwd_bwd_f32(const float* Y, const float* T, int64_t numel, const float* gscale_dev, int act_out,
                          double* sqerr_sum, float* dY, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(numel >= 0, "mse: bad numel");
  if (numel == 0) return LIPVQ_OK;
  LIPVQ_REQUIRE(Y && T && sqerr_sum, "mse: null pointer");
  LIPVQ_REQUIRE(act_out == LIPVQ_ACT_NONE || act_out == LIPVQ_ACT_RELU, "mse: act_out must be NONE or RELU");
  LIPVQ_CUDA(launch_pdl(mse_kernel, dim3(stream_grid(numel, 256, 8)), dim3(256), 0, stream, Y, T, numel, gscale_dev, act_out, sqerr_sum, dY));
  LIPVQ_LAUNCHED("mse_kernel");
  return LIPVQ_OK;
}

int lipvq_t1_loss_finalize(const double* sq2, double n_recon, double n_lat, float w_commit, float w_cb,
                           float* losses_out4, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(sq2 && losses_out4 && n_recon > 0 && n_lat > 0, "t1_loss_finalize: bad argument");
  LIPVQ_CUDA(launch_pdl(t1_loss_finalize_kernel, dim3(1), dim3(1), 0, stream, sq2, n_recon, n_lat, w_commit, w_cb, losses_out4));
  LIPVQ_LAUNCHED("t1_loss_finalize_kernel");
  return LIPVQ_OK;
}

}  // extern "C"
