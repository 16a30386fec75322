// Micro-benchmark: cost of finding, for every lane of a warp, the lanes that hold the same 8-bit digit
// (the ranking step of k3_sort_scatter).  Variants: 8 ballots (4 instructions per bit, the shipped form),
// match.any on the whole digit, match.any on 4-bit / 2-bit pieces, shared-memory atomicOr masks.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o rank_bench rank_bench.cu && ./rank_bench
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>

constexpr unsigned kFull = 0xffffffffu;
constexpr int kItems = 16;
constexpr int kThreads = 512;

__device__ __forceinline__ unsigned match_bit(unsigned peers, uint32_t v, uint32_t bit) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b32 t, s, m;\n\tand.b32 t, %1, %2;\n\tsetp.ne.u32 p, t, 0;\n\t"
      "vote.sync.ballot.b32 s, p, 0xffffffff;\n\tselp.b32 m, 0, 0xffffffff, p;\n\tlop3.b32 %0, %0, s, m, 0x60;\n\t}"
      : "+r"(peers) : "r"(v), "r"(bit));
  return peers;
}

template <int V>
__device__ __forceinline__ unsigned peers_of(uint32_t dk, uint32_t* masks, int lane) {
  if constexpr (V == 0) {
    unsigned p = kFull;
#pragma unroll
    for (int b = 0; b < 8; ++b) p = match_bit(p, dk, 1u << b);
    return p;
  } else if constexpr (V == 1) {
    return __match_any_sync(kFull, dk);
  } else if constexpr (V == 2) {
    return __match_any_sync(kFull, dk & 15u) & __match_any_sync(kFull, dk >> 4);
  } else if constexpr (V == 3) {
    return __match_any_sync(kFull, dk & 3u) & __match_any_sync(kFull, (dk >> 2) & 3u) &
           __match_any_sync(kFull, (dk >> 4) & 3u) & __match_any_sync(kFull, dk >> 6);
  } else {  // V == 4: atomicOr into a per-warp mask table, read back, clear
    atomicOr(&masks[dk], 1u << lane);
    __syncwarp();
    const unsigned p = masks[dk];
    __syncwarp();
    masks[dk] = 0;
    __syncwarp();
    return p;
  }
}

template <int V>
__global__ void __launch_bounds__(kThreads, 2) bench(const uint8_t* __restrict__ digits, int64_t n, int iters,
                                                     unsigned long long* out) {
  __shared__ uint32_t masks[kThreads / 32][256];
  __shared__ uint32_t wh[kThreads / 32][256];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < (kThreads / 32) * 256; i += kThreads) { (&masks[0][0])[i] = 0; (&wh[0][0])[i] = 0; }
  __syncthreads();
  unsigned long long acc = 0;
  const unsigned lt = (1u << lane) - 1u;
  for (int it = 0; it < iters; ++it) {
    const int64_t base = ((int64_t)(blockIdx.x * iters + it) * kThreads + w * 32) * kItems % (n - kThreads * kItems);
    uint32_t dk[kItems];
#pragma unroll
    for (int r = 0; r < kItems; ++r) dk[r] = digits[base + r * 32 + lane];
    if constexpr (V == 5) {
#pragma unroll
      for (int r = 0; r < kItems; ++r) {
        const unsigned p = peers_of<0>(dk[r], masks[w], lane);
        const uint32_t cnt = wh[w][dk[r]];          // every lane of the group reads the running count
        const uint32_t rank = __popc(p & lt);
        __syncwarp();
        if ((p >> lane) == 1u) wh[w][dk[r]] = cnt + rank + 1u;  // the group's highest lane knows the group size
        __syncwarp();
        acc += cnt + rank;
      }
    } else {
#pragma unroll
    for (int r = 0; r < kItems; ++r) {
      const unsigned p = peers_of<V>(dk[r], masks[w], lane);
      const int leader = __ffs(p) - 1;
      uint32_t old = 0;
      if (lane == leader) old = atomicAdd(&wh[w][dk[r]], (uint32_t)__popc(p));
      acc += __shfl_sync(kFull, old, leader) + __popc(p & lt);
    }
    }
  }
  atomicAdd(out, acc);
}

template <int V>
float run(const uint8_t* d, int64_t n, int iters, unsigned long long* out, unsigned long long* h) {
  cudaMemset(out, 0, 8);
  bench<V><<<296, kThreads>>>(d, n, 2, out);
  cudaDeviceSynchronize();
  cudaMemset(out, 0, 8);
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  cudaEventRecord(a);
  bench<V><<<296, kThreads>>>(d, n, iters, out);
  cudaEventRecord(b);
  cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  cudaMemcpy(h, out, 8, cudaMemcpyDeviceToHost);
  return ms;
}

int main() {
  const int64_t n = 1 << 24;
  uint8_t* hd = (uint8_t*)malloc(n);
  uint8_t* d; unsigned long long* out; unsigned long long h;
  cudaMalloc(&d, n); cudaMalloc(&out, 8);
  const int iters = 200;
  const double keys = 296.0 * iters * kThreads * kItems;
  for (int mode = 0; mode < 3; ++mode) {
    srand(1);
    for (int64_t i = 0; i < n; ++i) hd[i] = mode == 0 ? rand() & 255 : (mode == 1 ? (rand() % 100 < 90 ? 0x7e : rand() & 15) : rand() & 15);
    cudaMemcpy(d, hd, n, cudaMemcpyHostToDevice);
    const char* name = mode == 0 ? "uniform 256" : (mode == 1 ? "skewed (90% one digit)" : "16 values");
    float t0 = run<0>(d, n, iters, out, &h); unsigned long long c0 = h;
    float t1 = run<1>(d, n, iters, out, &h); unsigned long long c1 = h;
    float t2 = run<2>(d, n, iters, out, &h); unsigned long long c2 = h;
    float t3 = run<3>(d, n, iters, out, &h); unsigned long long c3 = h;
    float t4 = run<4>(d, n, iters, out, &h); unsigned long long c4 = h;
    float t5 = run<5>(d, n, iters, out, &h); unsigned long long c5 = h;
    printf("%-24s Gkeys/s: ballot8 %.1f | match8 %.1f | match4x2 %.1f | match2x4 %.1f | atomicOr %.1f | ballot8+LDS/STS %.1f  checks %d%d%d%d%d\n", name,
           keys / t0 / 1e6, keys / t1 / 1e6, keys / t2 / 1e6, keys / t3 / 1e6, keys / t4 / 1e6, keys / t5 / 1e6, c1 == c0, c2 == c0, c3 == c0, c4 == c0, c5 == c0);
  }
  return 0;
}
