// Micro-benchmarks that decide where the k-mer table lives on B200 (shared memory vs L2-resident global memory).
// Prints lane-ops per second for: shared atomics (u32 add, u64 add, u64 CAS), shared random loads, global RED/ATOM
// to an L2-resident region, global random loads. Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench ubench_atomics.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

typedef unsigned long long u64;
typedef unsigned int u32;

__device__ __forceinline__ u32 lcg(u32 &s) {
  s = s * 1664525u + 1013904223u;
  return s ^ (s >> 15);
}

constexpr int SM_WORDS = 48 * 1024;  // 192 KB of u32

template <int MODE>
__global__ void __launch_bounds__(1024) smem_kernel(u64 *sink, int iters) {
  extern __shared__ u32 sm[];
  for (int i = threadIdx.x; i < SM_WORDS; i += blockDim.x) sm[i] = 0;
  __syncthreads();
  u32 s = threadIdx.x * 2654435761u + blockIdx.x * 97u + 1u;
  u64 acc = 0;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int u = 0; u < 8; u++) {
      const u32 r = lcg(s);
      if (MODE == 0) {  // u32 atomic add, result unused (RED-like)
        atomicAdd(&sm[r % SM_WORDS], 1u);
      } else if (MODE == 1) {  // u32 atomic add, result used
        acc += atomicAdd(&sm[r % SM_WORDS], 1u);
      } else if (MODE == 2) {  // u64 atomic add, result unused
        atomicAdd((u64 *)sm + (r % (SM_WORDS / 2)), 1ull);
      } else if (MODE == 3) {  // u64 CAS
        acc += atomicCAS((u64 *)sm + (r % (SM_WORDS / 2)), 0ull, (u64)r);
      } else if (MODE == 4) {  // random 8-byte load
        acc += ((u64 *)sm)[r % (SM_WORDS / 2)];
      } else if (MODE == 5) {  // random 4-byte load
        acc += sm[r % SM_WORDS];
      } else if (MODE == 6) {  // 2 x u32 atomic add to adjacent words (same slot)
        const u32 a = (r % (SM_WORDS / 8)) * 8;
        atomicAdd(&sm[a + (r >> 28 & 3)], 1u);
        atomicAdd(&sm[a + 4 + (r >> 30)], 1u);
      } else if (MODE == 7) {  // random 16-byte load
        const uint4 v = ((uint4 *)sm)[r % (SM_WORDS / 4)];
        acc += v.x + v.w;
      }
    }
  }
  if (acc == 0x1234567812345678ull) sink[0] = acc;
}

template <int MODE>
__global__ void __launch_bounds__(1024) gmem_kernel(u32 *buf, u64 nwords, u64 *sink, int iters) {
  u32 s = (threadIdx.x + blockIdx.x * blockDim.x) * 2654435761u + 1u;
  u64 acc = 0;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int u = 0; u < 8; u++) {
      const u64 r = ((u64)lcg(s) << 8) ^ lcg(s);
      if (MODE == 0) {  // RED u32
        atomicAdd(&buf[r % nwords], 1u);
      } else if (MODE == 1) {  // ATOM u32 (result used)
        acc += atomicAdd(&buf[r % nwords], 1u);
      } else if (MODE == 2) {  // RED u64
        atomicAdd((u64 *)buf + (r % (nwords / 2)), 1ull);
      } else if (MODE == 3) {  // CAS u64
        acc += atomicCAS((u64 *)buf + (r % (nwords / 2)), 0ull, r);
      } else if (MODE == 4) {  // random 8-byte load, L2 (ld.cg)
        acc += __ldcg((u64 *)buf + (r % (nwords / 2)));
      } else if (MODE == 5) {  // load key (8 B) then 2 REDs in the same 32-byte slot
        const u64 a = (r % (nwords / 8)) * 8;
        acc += __ldcg((u64 *)(buf + a));
        atomicAdd(&buf[a + 2 + (r >> 40 & 1)], 1u);
        atomicAdd(&buf[a + 4 + (r >> 41 & 3)], 1u);
      } else if (MODE == 6) {  // one 64-bit RED per 32-byte slot
        const u64 a = (r % (nwords / 8)) * 8;
        atomicAdd((u64 *)(buf + a + 2), 1ull << (16 * (r >> 40 & 3)));
      }
    }
  }
  if (acc == 0x1234567812345678ull) sink[0] = acc;
}

template <typename F>
float time_ms(F f, int reps = 3) {
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  f();
  cudaDeviceSynchronize();
  float best = 1e30f;
  for (int i = 0; i < reps; i++) {
    cudaEventRecord(a);
    f();
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    if (ms < best) best = ms;
  }
  return best;
}

int main() {
  int sms = 0, clk = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  printf("SMs %d clock %d kHz\n", sms, clk);
  u64 *sink;
  cudaMalloc(&sink, 64);
  const int iters = 2000;
  const size_t smem = SM_WORDS * 4;
  const char *sn[] = {"smem atomicAdd u32 (no result)", "smem atomicAdd u32 (result)", "smem atomicAdd u64", "smem CAS u64",
                      "smem load 8B random", "smem load 4B random", "smem 2x atomicAdd same slot", "smem load 16B random"};
#define RUN_S(M, T)                                                                                        \
  {                                                                                                        \
    cudaFuncSetAttribute(smem_kernel<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);          \
    float ms = time_ms([&] { smem_kernel<M><<<sms, T, smem>>>(sink, iters); });                            \
    double ops = (double)sms * T * iters * 8 * (M == 6 ? 2 : 1);                                           \
    printf("%-36s threads %4d: %8.3f ms  %8.1f Gop/s  %.3f cyc/lane-op/SM\n", sn[M], T, ms, ops / ms / 1e6, \
           ms * 1e-3 * clk * 1e3 / (ops / sms));                                                           \
  }
  for (int T : {256, 512, 1024}) {
    RUN_S(0, T) RUN_S(1, T) RUN_S(2, T) RUN_S(3, T) RUN_S(4, T) RUN_S(5, T) RUN_S(6, T) RUN_S(7, T)
  }
  const char *gn[] = {"gmem RED u32", "gmem ATOM u32 (result)", "gmem RED u64", "gmem CAS u64", "gmem ld.cg 8B random",
                      "gmem key load + 2 RED same slot", "gmem one RED.64 per slot"};
  for (size_t mb : {8, 32, 64, 96, 512}) {
    const u64 nwords = mb * 1024 * 1024 / 4;
    u32 *buf;
    cudaMalloc(&buf, nwords * 4);
    cudaMemset(buf, 0, nwords * 4);
    const int git = 200;
#define RUN_G(M)                                                                                                   \
  {                                                                                                                \
    float ms = time_ms([&] { gmem_kernel<M><<<sms * 2, 1024>>>(buf, nwords, sink, git); });                        \
    double ops = (double)sms * 2 * 1024 * git * 8;                                                                 \
    printf("%4zu MB %-34s: %8.3f ms  %8.1f G/s  %.3f cyc/lane-op/SM\n", mb, gn[M], ms, ops / ms / 1e6,             \
           ms * 1e-3 * clk * 1e3 / (ops / sms));                                                                   \
  }
    RUN_G(0) RUN_G(1) RUN_G(2) RUN_G(3) RUN_G(4) RUN_G(5) RUN_G(6)
    cudaFree(buf);
  }
  cudaError_t e = cudaDeviceSynchronize();
  printf("status: %s\n", cudaGetErrorString(e));
  return e != cudaSuccess;
}
