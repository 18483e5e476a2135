// Micro-benchmark of the issue/pipe costs that bound the nearest-code search epilogue on sm_100a (B200).
// Builds stand-alone:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_pipes scripts/ubench_pipes.cu
// One CTA; W warps per scheduler (SM sub-partition).  Every body is written with asm volatile so ptxas keeps the mix.
// Prints cycles per 32-score "chunk" per warp and per scheduler for each instruction mix.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define REP 256

__device__ __forceinline__ float fmin3(float a, float b, float c) {
  float d;
  asm volatile("min.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}
__device__ __forceinline__ float fmin2(float a, float b) {
  float d;
  asm volatile("min.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b));
  return d;
}
__device__ __forceinline__ uint32_t lop_pack(uint32_t s, uint32_t mask, uint32_t c) {
  uint32_t d;
  asm volatile("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(d) : "r"(s), "r"(mask), "r"(c));  // (a & b) | c
  return d;
}
__device__ __forceinline__ float fadd(float a, float b) {
  float d;
  asm volatile("add.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b));
  return d;
}
__device__ __forceinline__ float ffma(float a, float b, float c) {
  float d;
  asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}
__device__ __forceinline__ float ffma_sat(float a, float b, float c) {
  float d;
  asm volatile("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}
__device__ __forceinline__ void top2_update(float& best, float& second, float& idxf, float s, float e, float nz, float pz) {
  asm volatile("{\n .reg .pred p;\n setp.lt.f32 p, %3, %0;\n @p add.f32 %1, %0, %5;\n @!p min.f32 %1, %1, %3;\n"
      " @p add.f32 %0, %3, %5;\n @p fma.rn.f32 %2, %3, %6, %4;\n}"
      : "+f"(best), "+f"(second), "+f"(idxf)
      : "f"(s), "f"(e), "f"(nz), "f"(pz));
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

// MODE: which mix.  One "chunk" = 32 score registers.
template <int MODE>
__global__ void __launch_bounds__(512, 1) bench(const float* in, float* out, long long* cyc, uint32_t mask, float nz, float pz) {
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t tmem_base = 0;
  if (MODE >= 20) {
    if (warp == 0) {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&tmem_slot)), "r"(512u) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    tmem_base = tmem_slot;
  }
  uint32_t v[32];
#pragma unroll
  for (int e = 0; e < 32; ++e) v[e] = __float_as_uint(in[(threadIdx.x * 32 + e) & 4095]);
  float a0 = in[lane], a1 = in[lane + 1], a2 = in[lane + 2], a3 = in[lane + 3];
  float c0 = 0.f, c1 = 0.f, c2 = 0.f, c3 = 0.f;
  float i0 = 0.f, i1 = 0.f, i2 = 0.f, i3 = 0.f;
  float s0 = a0, s1 = a1, s2 = a2, s3 = a3;
  const float thr = in[lane + 7], big = in[lane + 9];
  const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int r = 0; r < REP; ++r) {
    if (MODE >= 20) {
      tmem_ld32(taddr + (uint32_t)((r & 7) * 32), v);
      tmem_ld_wait(v);
    }
    if (MODE == 0) {  // FMNMX 2-input, 4 chains
#pragma unroll
      for (int e = 0; e < 32; e += 4) {
        a0 = fmin2(a0, __uint_as_float(v[e])); a1 = fmin2(a1, __uint_as_float(v[e + 1]));
        a2 = fmin2(a2, __uint_as_float(v[e + 2])); a3 = fmin2(a3, __uint_as_float(v[e + 3]));
      }
    } else if (MODE == 1) {  // FMNMX3, 4 chains: 16 instructions per chunk
#pragma unroll
      for (int e = 0; e < 32; e += 8) {
        a0 = fmin3(a0, __uint_as_float(v[e]), __uint_as_float(v[e + 1])); a1 = fmin3(a1, __uint_as_float(v[e + 2]), __uint_as_float(v[e + 3]));
        a2 = fmin3(a2, __uint_as_float(v[e + 4]), __uint_as_float(v[e + 5])); a3 = fmin3(a3, __uint_as_float(v[e + 6]), __uint_as_float(v[e + 7]));
      }
    } else if (MODE == 2) {  // LOP3 pack only (32 per chunk), results consumed by xor chain at the end
      uint32_t k[32];
#pragma unroll
      for (int e = 0; e < 32; ++e) k[e] = lop_pack(v[e], mask, (uint32_t)e);
#pragma unroll
      for (int e = 0; e < 32; ++e) v[e] = k[e];
    } else if (MODE == 3) {  // FADD 32 per chunk, 4 chains
#pragma unroll
      for (int e = 0; e < 32; e += 4) {
        c0 = fadd(c0, __uint_as_float(v[e])); c1 = fadd(c1, __uint_as_float(v[e + 1]));
        c2 = fadd(c2, __uint_as_float(v[e + 2])); c3 = fadd(c3, __uint_as_float(v[e + 3]));
      }
    } else if (MODE == 4) {  // FFMA 3-reg, 4 chains
#pragma unroll
      for (int e = 0; e < 32; e += 4) {
        c0 = ffma(__uint_as_float(v[e]), big, c0); c1 = ffma(__uint_as_float(v[e + 1]), big, c1);
        c2 = ffma(__uint_as_float(v[e + 2]), big, c2); c3 = ffma(__uint_as_float(v[e + 3]), big, c3);
      }
    } else if (MODE == 5) {  // FFMA.SAT then FADD: threshold count (2 FMA-pipe ops per score)
#pragma unroll
      for (int e = 0; e < 32; e += 4) {
        c0 = fadd(c0, ffma_sat(__uint_as_float(v[e]), big, thr)); c1 = fadd(c1, ffma_sat(__uint_as_float(v[e + 1]), big, thr));
        c2 = fadd(c2, ffma_sat(__uint_as_float(v[e + 2]), big, thr)); c3 = fadd(c3, ffma_sat(__uint_as_float(v[e + 3]), big, thr));
      }
    } else if (MODE == 6 || MODE == 20) {  // pack + min3: the proposed best-only fold (48 ALU ops per chunk)
#pragma unroll
      for (int e = 0; e < 32; e += 8) {
        a0 = fmin3(a0, __uint_as_float(lop_pack(v[e], mask, e)), __uint_as_float(lop_pack(v[e + 1], mask, e + 1)));
        a1 = fmin3(a1, __uint_as_float(lop_pack(v[e + 2], mask, e + 2)), __uint_as_float(lop_pack(v[e + 3], mask, e + 3)));
        a2 = fmin3(a2, __uint_as_float(lop_pack(v[e + 4], mask, e + 4)), __uint_as_float(lop_pack(v[e + 5], mask, e + 5)));
        a3 = fmin3(a3, __uint_as_float(lop_pack(v[e + 6], mask, e + 6)), __uint_as_float(lop_pack(v[e + 7], mask, e + 7)));
      }
    } else if (MODE == 7 || MODE == 21) {  // pack + min3 + threshold count (48 ALU + 64 FMA-pipe per chunk)
#pragma unroll
      for (int e = 0; e < 32; e += 8) {
        a0 = fmin3(a0, __uint_as_float(lop_pack(v[e], mask, e)), __uint_as_float(lop_pack(v[e + 1], mask, e + 1)));
        a1 = fmin3(a1, __uint_as_float(lop_pack(v[e + 2], mask, e + 2)), __uint_as_float(lop_pack(v[e + 3], mask, e + 3)));
        a2 = fmin3(a2, __uint_as_float(lop_pack(v[e + 4], mask, e + 4)), __uint_as_float(lop_pack(v[e + 5], mask, e + 5)));
        a3 = fmin3(a3, __uint_as_float(lop_pack(v[e + 6], mask, e + 6)), __uint_as_float(lop_pack(v[e + 7], mask, e + 7)));
      }
#pragma unroll
      for (int e = 0; e < 32; e += 4) {
        c0 = fadd(c0, ffma_sat(__uint_as_float(v[e]), big, thr)); c1 = fadd(c1, ffma_sat(__uint_as_float(v[e + 1]), big, thr));
        c2 = fadd(c2, ffma_sat(__uint_as_float(v[e + 2]), big, thr)); c3 = fadd(c3, ffma_sat(__uint_as_float(v[e + 3]), big, thr));
      }
    } else if (MODE == 8 || MODE == 22) {  // the round-1 top-2 update (2 ALU + 3 FMA-pipe per score)
#pragma unroll
      for (int e = 0; e < 32; e += 4) {
        top2_update(a0, s0, i0, __uint_as_float(v[e]), (float)e, nz, pz);
        top2_update(a1, s1, i1, __uint_as_float(v[e + 1]), (float)(e + 1), nz, pz);
        top2_update(a2, s2, i2, __uint_as_float(v[e + 2]), (float)(e + 2), nz, pz);
        top2_update(a3, s3, i3, __uint_as_float(v[e + 3]), (float)(e + 3), nz, pz);
      }
    } else if (MODE == 9 || MODE == 23) {  // raw min3 only (no pack): 16 ALU per chunk
#pragma unroll
      for (int e = 0; e < 32; e += 8) {
        a0 = fmin3(a0, __uint_as_float(v[e]), __uint_as_float(v[e + 1])); a1 = fmin3(a1, __uint_as_float(v[e + 2]), __uint_as_float(v[e + 3]));
        a2 = fmin3(a2, __uint_as_float(v[e + 4]), __uint_as_float(v[e + 5])); a3 = fmin3(a3, __uint_as_float(v[e + 6]), __uint_as_float(v[e + 7]));
      }
    } else if (MODE == 10 || MODE == 24) {  // raw min3 + 3 FMA-pipe ops per score (sat, count, index)
#pragma unroll
      for (int e = 0; e < 32; e += 8) {
        a0 = fmin3(a0, __uint_as_float(v[e]), __uint_as_float(v[e + 1])); a1 = fmin3(a1, __uint_as_float(v[e + 2]), __uint_as_float(v[e + 3]));
        a2 = fmin3(a2, __uint_as_float(v[e + 4]), __uint_as_float(v[e + 5])); a3 = fmin3(a3, __uint_as_float(v[e + 6]), __uint_as_float(v[e + 7]));
      }
#pragma unroll
      for (int e = 0; e < 32; e += 2) {
        const float u0 = ffma_sat(__uint_as_float(v[e]), big, thr), u1 = ffma_sat(__uint_as_float(v[e + 1]), big, thr);
        c0 = fadd(c0, u0); c1 = fadd(c1, u1);
        i0 = ffma(u0, (float)e, i0); i1 = ffma(u1, (float)(e + 1), i1);
      }
    } else if (MODE == 25) {  // tcgen05.ld only
    }
  }
  const long long t1 = clock64();
  float acc = a0 + a1 + a2 + a3 + c0 + c1 + c2 + c3 + i0 + i1 + i2 + i3 + s0 + s1 + s2 + s3;
#pragma unroll
  for (int e = 0; e < 32; ++e) acc += __uint_as_float(v[e]);
  out[threadIdx.x] = acc;
  if (lane == 0) cyc[warp] = t1 - t0;
  if (MODE >= 20) {
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

template <int MODE>
void run(const char* name, const float* in, float* out, long long* cyc) {
  for (int warps : {4, 8, 16}) {
    bench<MODE><<<1, warps * 32>>>(in, out, cyc, 0xFFFFFE00u, -0.0f, 0.0f);
    bench<MODE><<<1, warps * 32>>>(in, out, cyc, 0xFFFFFE00u, -0.0f, 0.0f);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
      printf("%-46s warps=%2d  CUDA error %s\n", name, warps, cudaGetErrorString(e));
      return;
    }
    long long h[16];
    cudaMemcpy(h, cyc, sizeof(long long) * warps, cudaMemcpyDeviceToHost);
    long long mx = 0;
    for (int i = 0; i < warps; ++i) mx = h[i] > mx ? h[i] : mx;
    const double per_chunk_warp = (double)mx / REP;
    const double per_chunk_sched = per_chunk_warp / (warps / 4);
    printf("%-46s warps/sched=%d  cyc/chunk/warp=%7.1f  cyc/chunk/sched=%7.1f  cyc/score/sched=%5.2f\n", name, warps / 4,
           per_chunk_warp, per_chunk_sched, per_chunk_sched / 32.0);
  }
}

int main() {
  float *in, *out;
  long long* cyc;
  cudaMalloc(&in, 8192 * 4);
  cudaMalloc(&out, 1024 * 4);
  cudaMalloc(&cyc, 64 * 8);
  float h[8192];
  for (int i = 0; i < 8192; ++i) h[i] = 1.0f + (float)((i * 2654435761u) % 1000) * 1e-3f;
  cudaMemcpy(in, h, sizeof(h), cudaMemcpyHostToDevice);
  run<0>("FMNMX 2-input x32", in, out, cyc);
  run<1>("FMNMX3 x16", in, out, cyc);
  run<2>("LOP3 x32", in, out, cyc);
  run<3>("FADD x32", in, out, cyc);
  run<4>("FFMA 3-reg x32", in, out, cyc);
  run<5>("FFMA.SAT+FADD x32", in, out, cyc);
  run<6>("pack+min3 (32 LOP3 + 16 FMNMX3)", in, out, cyc);
  run<7>("pack+min3 + count (+64 FMA-pipe)", in, out, cyc);
  run<8>("round-1 top2_update x32", in, out, cyc);
  run<9>("raw min3 x16", in, out, cyc);
  run<10>("raw min3 + sat/count/index (16 ALU + 96 FMA)", in, out, cyc);
  run<25>("tcgen05.ld x32 only", in, out, cyc);
  run<23>("tcgen05.ld + raw min3", in, out, cyc);
  run<20>("tcgen05.ld + pack+min3", in, out, cyc);
  run<21>("tcgen05.ld + pack+min3 + count", in, out, cyc);
  run<22>("tcgen05.ld + round-1 top2", in, out, cyc);
  run<24>("tcgen05.ld + raw min3 + sat/count/index", in, out, cyc);
  return 0;
}
