// One-shot sum all-reduce of the data-parallel reduction slab over NVLink peer memory.
//
// The slab ([gradients | EMA counts | EMA sums | losses], ~0.3 MB) is far below the size where a ring or tree pays:
// every rank simply reads the slab of every peer through its peer mapping (NVSwitch: full bandwidth to each peer) and
// adds the W copies in rank order, so all ranks produce the same bits.  One kernel, two flag exchanges:
//   (1) "my earlier kernels are complete"  -- CTA b of rank r raises flag[0][r][b] in every peer's buffer and waits for
//       flag[0][*][b] in its own; a CTA can only run after the kernels that filled the slab, so this is all the
//       ordering the reads need;
//   (2) "I have read your slab"            -- same exchange after the sums, so a rank leaves the kernel only when no
//       peer still reads its slab (the next kernel clears it).
// Flags carry a per-CTA call counter (monotonic, kept on the device so a captured graph replays correctly).
// The symmetric allocation and the peer pointers come from torch.distributed._symmetric_memory (host plumbing);
// the data path is this kernel only.  NCCL remains the fallback (lipvq/parallel.py).
#include <stdio.h>
#include <stdlib.h>

#include "common.cuh"

namespace lipvq {
namespace {

constexpr int kArThreads = 512;

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_peer_f4(const float4* p) {  // system-scope load: never served from a stale L1 line
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}

// flags of one rank: [2 phases][world][gridDim.x] uint32 at byte offset flag_off of its symmetric buffer
__device__ __forceinline__ void signal_peers(const uint64_t* peer_ptrs, int64_t flag_off, int phase, int rank, int world, uint32_t epoch) {
  const int G = gridDim.x, b = blockIdx.x;
  if ((int)threadIdx.x < world) {
    uint32_t* theirs = reinterpret_cast<uint32_t*>(peer_ptrs[threadIdx.x] + flag_off) + ((size_t)phase * world + rank) * G + b;
    __threadfence_system();
    st_release_sys(theirs, epoch);
  }
}
__device__ __forceinline__ void wait_peers(const uint64_t* peer_ptrs, int64_t flag_off, int phase, int rank, int world, uint32_t epoch,
                                           long long timeout_clocks) {
  const int G = gridDim.x, b = blockIdx.x;
  if ((int)threadIdx.x < world) {
    const uint32_t* mine = reinterpret_cast<const uint32_t*>(peer_ptrs[rank] + flag_off) + ((size_t)phase * world + threadIdx.x) * G + b;
    long long t0 = 0;
    uint32_t spins = 0;
    while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
      if ((++spins & 255u) == 0) {
        // A rank that NEVER arrives must not hang the box for good -- but a rank that is merely late (rank-0-only
        // checkpointing or validation, a data-loader stall, first-step warm-up) has to be waited for, as NCCL would.
        // The bound is LIPVQ_PEER_TIMEOUT_S (default 1800 s, NCCL's default watchdog); on expiry say who is missing.
        if (t0 == 0) t0 = clock64();
        else if (clock64() - t0 > timeout_clocks) {
          printf("lipvq: peer all-reduce: rank %d gave up waiting for rank %d (phase %d, call %u) after LIPVQ_PEER_TIMEOUT_S; "
                 "that rank stopped issuing steps\n", rank, (int)threadIdx.x, phase, epoch);
          __trap();
        }
      }
    }
  }
  __syncthreads();
}
__device__ __forceinline__ void exchange(const uint64_t* peer_ptrs, int64_t flag_off, int phase, int rank, int world,
                                         uint32_t epoch, long long timeout_clocks) {
  signal_peers(peer_ptrs, flag_off, phase, rank, world, epoch);
  wait_peers(peer_ptrs, flag_off, phase, rank, world, epoch, timeout_clocks);
}

// defer_release != 0: exchange (2) is split -- this kernel only SIGNALS "I have read your slab"; the wait for the peers'
// signals happens in allreduce_clear_kernel, which the next step runs before it clears the slab (by then the signals are
// ~a step old: the round trip leaves the critical path).
__global__ void __launch_bounds__(kArThreads)
    allreduce_oneshot_kernel(const uint64_t* __restrict__ peer_ptrs, int rank, int world, int64_t n4, int64_t flag_off,
                             float4* __restrict__ out, uint32_t* __restrict__ cta_epoch, long long timeout_clocks, int defer_release) {
  __shared__ uint32_t s_epoch;
  grid_dep_wait();  // launched as a programmatic dependent: the kernels that filled the slab are complete
  if (threadIdx.x == 0) s_epoch = ++cta_epoch[blockIdx.x];
  __syncthreads();
  const uint32_t epoch = s_epoch;
  exchange(peer_ptrs, flag_off, 0, rank, world, epoch, timeout_clocks);
  for (int64_t i = (int64_t)blockIdx.x * kArThreads + threadIdx.x; i < n4; i += (int64_t)gridDim.x * kArThreads) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r = 0; r < world; ++r) {  // rank order on every rank: identical bits everywhere
      const float4 v = ld_peer_f4(reinterpret_cast<const float4*>(peer_ptrs[r]) + i);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    out[i] = acc;
  }
  __syncthreads();
  if (defer_release) {
    signal_peers(peer_ptrs, flag_off, 1, rank, world, epoch);
  } else {
    exchange(peer_ptrs, flag_off, 1, rank, world, epoch, timeout_clocks);
  }
  grid_dep_launch();
}

// The other half of the deferred exchange (2), fused with the clear it protects: CTA b waits until every peer has
// signalled that it has read CTA b's share of this rank's slab in the LAST all-reduce (epoch = the call counter that
// all-reduce left behind; 0 before the first one: nothing to wait for), then zeroes that share.
__global__ void __launch_bounds__(kArThreads)
    allreduce_clear_kernel(const uint64_t* __restrict__ peer_ptrs, int rank, int world, int64_t n4, int64_t flag_off,
                           const uint32_t* __restrict__ cta_epoch, long long timeout_clocks) {
  grid_dep_wait();
  wait_peers(peer_ptrs, flag_off, 1, rank, world, cta_epoch[blockIdx.x], timeout_clocks);
  float4* mine = reinterpret_cast<float4*>(peer_ptrs[rank]);
  // (the same index -> CTA mapping as the sums above: a peer's CTA b reads exactly what this CTA clears)
  for (int64_t i = (int64_t)blockIdx.x * kArThreads + threadIdx.x; i < n4; i += (int64_t)gridDim.x * kArThreads)
    mine[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  grid_dep_launch();
}

}  // namespace
}  // namespace lipvq

using namespace lipvq;

extern "C" {

int lipvq_allreduce_oneshot_ctas(void) { return 32; }

// how long a rank waits for a peer's flag before it gives up (SM clocks; ~2 GHz)
static long long peer_timeout_clocks() {
  static long long cached = 0;
  if (cached == 0) {
    double s = 1800.0;
    if (const char* e = getenv("LIPVQ_PEER_TIMEOUT_S")) {
      const double v = atof(e);
      if (v > 0.0) s = v;
    }
    cached = (long long)(s * 2.0e9);
  }
  return cached;
}

size_t lipvq_allreduce_oneshot_flag_bytes(int world) {
  return (size_t)2 * (size_t)(world > 0 ? world : 1) * (size_t)lipvq_allreduce_oneshot_ctas() * sizeof(uint32_t);
}

static int allreduce_launch(const uint64_t* peer_ptrs_dev, int rank, int world, int64_t n, int64_t flag_offset_bytes, float* out,
                            uint32_t* cta_epoch, void* stream_, int defer_release) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(world >= 1 && world <= 64 && rank >= 0 && rank < world, "allreduce_oneshot: bad rank/world %d/%d", rank, world);
  LIPVQ_REQUIRE(n >= 0 && n % 4 == 0, "allreduce_oneshot: length must be a multiple of 4 floats");
  LIPVQ_REQUIRE(flag_offset_bytes >= n * 4 && flag_offset_bytes % 16 == 0, "allreduce_oneshot: flags must follow the slab");
  if (n == 0) return LIPVQ_OK;
  LIPVQ_REQUIRE(peer_ptrs_dev && out && cta_epoch && aligned16(out), "allreduce_oneshot: null or misaligned pointer");
  LIPVQ_CUDA(launch_pdl(allreduce_oneshot_kernel, dim3(lipvq_allreduce_oneshot_ctas()), dim3(kArThreads), 0, stream, peer_ptrs_dev, rank,
                        world, n / 4, flag_offset_bytes, reinterpret_cast<float4*>(out), cta_epoch, peer_timeout_clocks(), defer_release));
  LIPVQ_LAUNCHED("allreduce_oneshot_kernel");
  return LIPVQ_OK;
}

int lipvq_allreduce_oneshot_f32(const uint64_t* peer_ptrs_dev, int rank, int world, int64_t n, int64_t flag_offset_bytes,
                                float* out, uint32_t* cta_epoch, void* stream) {
  return allreduce_launch(peer_ptrs_dev, rank, world, n, flag_offset_bytes, out, cta_epoch, stream, 0);
}

int lipvq_allreduce_oneshot_deferred_f32(const uint64_t* peer_ptrs_dev, int rank, int world, int64_t n, int64_t flag_offset_bytes,
                                         float* out, uint32_t* cta_epoch, void* stream) {
  return allreduce_launch(peer_ptrs_dev, rank, world, n, flag_offset_bytes, out, cta_epoch, stream, 1);
}

int lipvq_allreduce_oneshot_clear_f32(const uint64_t* peer_ptrs_dev, int rank, int world, int64_t n, int64_t flag_offset_bytes,
                                      const uint32_t* cta_epoch, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  LIPVQ_REQUIRE(world >= 1 && world <= 64 && rank >= 0 && rank < world, "allreduce_clear: bad rank/world %d/%d", rank, world);
  LIPVQ_REQUIRE(n >= 0 && n % 4 == 0 && flag_offset_bytes >= n * 4 && flag_offset_bytes % 16 == 0, "allreduce_clear: bad layout");
  if (n == 0) return LIPVQ_OK;
  LIPVQ_REQUIRE(peer_ptrs_dev && cta_epoch, "allreduce_clear: null pointer");
  LIPVQ_CUDA(launch_pdl(allreduce_clear_kernel, dim3(lipvq_allreduce_oneshot_ctas()), dim3(kArThreads), 0, stream, peer_ptrs_dev, rank,
                        world, n / 4, flag_offset_bytes, cta_epoch, peer_timeout_clocks()));
  LIPVQ_LAUNCHED("allreduce_clear_kernel");
  return LIPVQ_OK;
}

}  // extern "C"
