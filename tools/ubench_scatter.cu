// Microbenchmark behind the fill kernel's design: N 16-byte records appended to B buckets through a 32-bit cursor atomic
// (the fill kernel's memory pattern), for several B and the three ingredients on their own.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_scatter tools/ubench_scatter.cu && tools/ubench_scatter
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long u64;
typedef unsigned int u32;

__device__ __forceinline__ u64 mix(u64 x) {
  x ^= x >> 32; x *= 0xD6E8FEB86659FD93ULL; x ^= x >> 32; x *= 0xD6E8FEB86659FD93ULL; x ^= x >> 32;
  return x;
}

// mode 0: atomic + store, 1: store at a random slot (no atomic), 2: atomic only, 3: atomic + store, 4 records in flight,
// 4: 32-byte records, both halves stored by the owning thread, 5: 32-byte records, lane pairs store whole sectors
__global__ void scatter(u32 *cursor, uint4 *rec, u64 n, u32 B, u32 cap, int mode) {
  const u64 stride = (u64)gridDim.x * blockDim.x;
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (mode == 3) {
    for (; i + 3 * stride < n; i += 4 * stride) {
      u32 b[4], pos[4];
#pragma unroll
      for (int u = 0; u < 4; u++) { b[u] = (u32)(((mix(i + u * stride) >> 32) * B) >> 32); pos[u] = atomicAdd(&cursor[b[u]], 1u); }
#pragma unroll
      for (int u = 0; u < 4; u++) rec[pos[u]] = make_uint4((u32)i, b[u], pos[u], u);
    }
    return;
  }
  if (mode == 4 || mode == 5) {  // n/2 records of 32 bytes (rec indexed in 16-byte units: slot 2 * pos)
    const u64 n2 = (n / 2) & ~31ULL;
    for (; i < n2; i += stride) {
      const u32 b = (u32)(((mix(i) >> 32) * B) >> 32);
      const u32 pos = atomicAdd(&cursor[b], 2u);
      uint4 *dst = rec + pos;
      if (mode == 4) {
        dst[0] = make_uint4((u32)i, b, pos, 0);
        dst[1] = make_uint4((u32)i, b, pos, 1);
      } else {
        const bool odd = threadIdx.x & 1;
        const u64 od = __shfl_xor_sync(0xffffffffu, (u64)dst, 1);
        const u32 r = __shfl_xor_sync(0xffffffffu, odd ? (u32)i : pos, 1);
        *(uint4 *)(odd ? od + 16 : (u64)dst) = make_uint4((u32)i, b, r, 0);
        *(uint4 *)(odd ? (u64)dst + 16 : od) = make_uint4((u32)i, b, r, 1);
      }
    }
    return;
  }
  for (; i < n; i += stride) {
    const u64 h = mix(i);
    const u32 b = (u32)(((h >> 32) * B) >> 32);
    if (mode == 0) {
      const u32 pos = atomicAdd(&cursor[b], 1u);
      rec[pos] = make_uint4((u32)i, b, pos, 0);
    } else if (mode == 1) {
      const u64 pos = (u64)b * cap + (u32)(((h & 0xFFFFFFFFu) * cap) >> 32);
      rec[pos] = make_uint4((u32)i, b, (u32)pos, 0);
    } else {
      atomicAdd(&cursor[b], 1u);
    }
  }
}

__global__ void init_cursor(u32 *cursor, u32 B, u32 cap) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < B) cursor[i] = i * cap;
}

int main() {
  const u64 n = 230000000ULL;  // one 8 M-read batch at k=21
  uint4 *rec;
  u32 *cursor;
  const u64 rec_cap = (u64)(n * 1.5) + (1u << 22);
  cudaMalloc(&rec, rec_cap * 16);
  cudaMalloc(&cursor, 4u << 20);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const u32 Bs[] = {32768, 376000};
  for (u32 B : Bs) {
    const u32 cap = (u32)(n / B * 1.3) + 64;
    if ((u64)cap * B > rec_cap) { printf("B=%u skipped\n", B); continue; }
    for (int mode = 0; mode < 6; mode++) {
      float best = 1e9f;
      for (int rep = 0; rep < 3; rep++) {
        init_cursor<<<(B + 255) / 256, 256>>>(cursor, B, cap);
        cudaEventRecord(e0);
        scatter<<<148 * 16, 128>>>(cursor, rec, n, B, cap, mode);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
      }
      printf("B=%8u footprint=%6.2f GB mode=%d (%s): %7.3f ms  %6.1f G records/s\n", B, (double)cap * B * 16 / 1e9, mode,
             mode == 0 ? "atomic+store" : mode == 1 ? "store only" : mode == 2 ? "atomic only" : mode == 3 ? "atomic+store x4" : mode == 4 ? "32 B records, one thread" : "32 B records, lane pairs", best, n / best / 1e6);
    }
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
