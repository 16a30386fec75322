// Micro-benchmark: stable counting sort of an 8192-key tile by one 8-bit digit, entirely in shared memory.
//   A  (shipped in round 1)  warp ranking with eight ballots per round + per-warp digit counters
//   B  (candidate)           two stable 4-bit splits: every thread owns 16 CONSECUTIVE keys, counts its nibbles in a
//                            thread-private column of a [16][512] u16 matrix (bank = f(thread) only: conflict-free),
//                            one CTA-wide scan of the matrix, scatter; no warp votes at all
// Both write the tile in digit order to global memory; the host checks B against std::stable_sort.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tile_sort_bench tile_sort_bench.cu && ./tile_sort_bench
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

constexpr unsigned kFull = 0xffffffffu;
constexpr int kThreads = 512, kItems = 16, kTile = kThreads * kItems, kWarps = kThreads / 32;

__device__ __forceinline__ int pad(int i) { return i + ((i >> 4) << 2); }  // 16-key strips 20 words apart
constexpr int kTileWords = kTile + (kTile >> 4) * 4;

// ------------------------------------------------------------------ B: two 4-bit splits
__device__ __forceinline__ void split4(uint32_t (&key)[kItems], int nib_shift, uint16_t* cnt, uint32_t* tile,
                                       uint32_t* warp_tot) {
  const int t = threadIdx.x, lane = t & 31, w = t >> 5;
  // zero the [16][512] count matrix: 16 KB = 2 x 16 bytes per thread
  reinterpret_cast<uint4*>(cnt)[t] = make_uint4(0, 0, 0, 0);
  reinterpret_cast<uint4*>(cnt)[t + kThreads] = make_uint4(0, 0, 0, 0);
  __syncthreads();
  uint32_t rank[kItems];
#pragma unroll
  for (int j = 0; j < kItems; ++j) {
    const int nib = (key[j] >> nib_shift) & 15;
    uint16_t* c = cnt + nib * kThreads + t;
    const uint32_t old = *c;
    *c = (uint16_t)(old + 1);
    rank[j] = old;
  }
  __syncthreads();
  // exclusive scan of the matrix in (bin, thread) order: thread t owns entries [16 t, 16 t + 16)
  uint4 a = reinterpret_cast<uint4*>(cnt)[2 * t], b = reinterpret_cast<uint4*>(cnt)[2 * t + 1];
  uint32_t v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
  uint32_t e[16];
  uint32_t run = 0;
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    e[2 * q] = run;
    run += v[q] & 0xffffu;
    e[2 * q + 1] = run;
    run += v[q] >> 16;
  }
  uint32_t incl = run;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t o = __shfl_up_sync(kFull, incl, d);
    if (lane >= d) incl += o;
  }
  if (lane == 31) warp_tot[w] = incl;
  __syncthreads();
  uint32_t before = 0;
#pragma unroll
  for (int ww = 0; ww < kWarps; ++ww) before += (ww < w) ? warp_tot[ww] : 0u;
  const uint32_t base = before + incl - run;
  uint32_t o[8];
#pragma unroll
  for (int q = 0; q < 8; ++q) o[q] = (base + e[2 * q]) | ((base + e[2 * q + 1]) << 16);
  reinterpret_cast<uint4*>(cnt)[2 * t] = make_uint4(o[0], o[1], o[2], o[3]);
  reinterpret_cast<uint4*>(cnt)[2 * t + 1] = make_uint4(o[4], o[5], o[6], o[7]);
  __syncthreads();
#pragma unroll
  for (int j = 0; j < kItems; ++j) {
    const int nib = (key[j] >> nib_shift) & 15;
    const int pos = cnt[nib * kThreads + t] + rank[j];
    tile[pad(pos)] = key[j];
  }
  __syncthreads();
}

__global__ void __launch_bounds__(kThreads, 2) sort_b(const uint32_t* __restrict__ in, uint32_t* __restrict__ out,
                                                      int shift, int tiles_per_cta) {
  extern __shared__ uint32_t sm[];
  uint32_t* tile = sm;                                    // kTileWords
  uint16_t* cnt = (uint16_t*)(sm + kTileWords);           // [16][512]
  uint32_t* warp_tot = sm + kTileWords + 16 * kThreads / 2;
  const int t = threadIdx.x;
  for (int it = 0; it < tiles_per_cta; ++it) {
    const int64_t tile0 = ((int64_t)blockIdx.x * tiles_per_cta + it) * kTile;
    uint32_t key[kItems];
    const uint4* src = reinterpret_cast<const uint4*>(in + tile0 + t * kItems);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint4 r = __ldg(src + q);
      key[4 * q] = r.x; key[4 * q + 1] = r.y; key[4 * q + 2] = r.z; key[4 * q + 3] = r.w;
    }
    split4(key, shift, cnt, tile, warp_tot);
    const uint4* strip = reinterpret_cast<const uint4*>(tile + t * 20);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint4 r = strip[q];
      key[4 * q] = r.x; key[4 * q + 1] = r.y; key[4 * q + 2] = r.z; key[4 * q + 3] = r.w;
    }
    __syncthreads();
    split4(key, shift + 4, cnt, tile, warp_tot);
    for (int i = t; i < kTile; i += kThreads) out[tile0 + i] = tile[pad(i)];
    __syncthreads();
  }
}

// ------------------------------------------------------------------ A: ballots (round-1 kernel body, no global offsets)
__device__ __forceinline__ unsigned match_bit(unsigned peers, uint32_t v, uint32_t bit) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b32 t, s, m;\n\tand.b32 t, %1, %2;\n\tsetp.ne.u32 p, t, 0;\n\t"
      "vote.sync.ballot.b32 s, p, 0xffffffff;\n\tselp.b32 m, 0, 0xffffffff, p;\n\tlop3.b32 %0, %0, s, m, 0x60;\n\t}"
      : "+r"(peers) : "r"(v), "r"(bit));
  return peers;
}

__global__ void __launch_bounds__(kThreads, 2) sort_a(const uint32_t* __restrict__ in, uint32_t* __restrict__ out,
                                                      int shift, int tiles_per_cta) {
  extern __shared__ uint32_t sm[];
  uint32_t(*wh)[256] = reinterpret_cast<uint32_t(*)[256]>(sm);
  uint32_t* sorted = sm + kWarps * 256;
  __shared__ uint32_t warp_tot[8];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const unsigned lt = (1u << lane) - 1u;
  for (int it = 0; it < tiles_per_cta; ++it) {
    const int64_t tile0 = ((int64_t)blockIdx.x * tiles_per_cta + it) * kTile;
    for (int i = threadIdx.x; i < kWarps * 256; i += kThreads) sm[i] = 0;
    __syncthreads();
    uint32_t key[kItems], rank[kItems];
#pragma unroll
    for (int r = 0; r < kItems; ++r) key[r] = __ldg(in + tile0 + w * 512 + r * 32 + lane);
#pragma unroll
    for (int r = 0; r < kItems; ++r) {
      const uint32_t dk = (key[r] >> shift) & 255u;
      unsigned peers = kFull;
#pragma unroll
      for (int bit = 0; bit < 8; ++bit) peers = match_bit(peers, dk, 1u << bit);
      const int leader = __ffs(peers) - 1;
      uint32_t old = 0;
      if (lane == leader) old = atomicAdd(&wh[w][dk], (uint32_t)__popc(peers));
      rank[r] = __shfl_sync(kFull, old, leader) + __popc(peers & lt);
    }
    __syncthreads();
    uint32_t total = 0;
    if (threadIdx.x < 256)
#pragma unroll
      for (int ww = 0; ww < kWarps; ++ww) total += wh[ww][threadIdx.x];
    uint32_t incl = total;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t o = __shfl_up_sync(kFull, incl, d);
      if (lane >= d) incl += o;
    }
    if (w < 8 && lane == 31) warp_tot[w] = incl;
    __syncthreads();
    if (threadIdx.x < 256) {
      uint32_t before = 0;
#pragma unroll
      for (int ww = 0; ww < 8; ++ww) before += (ww < w) ? warp_tot[ww] : 0u;
      uint32_t run = before + incl - total;
#pragma unroll
      for (int ww = 0; ww < kWarps; ++ww) {
        const uint32_t c = wh[ww][threadIdx.x];
        wh[ww][threadIdx.x] = run;
        run += c;
      }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kItems; ++r) sorted[wh[w][(key[r] >> shift) & 255u] + rank[r]] = key[r];
    __syncthreads();
    for (int i = threadIdx.x; i < kTile; i += kThreads) out[tile0 + i] = sorted[i];
    __syncthreads();
  }
}

int main() {
  const int ctas = 296, per = 24;
  const int64_t n = (int64_t)ctas * per * kTile;
  std::vector<uint32_t> h(n), got(n);
  uint32_t *din, *dout;
  cudaMalloc(&din, n * 4);
  cudaMalloc(&dout, n * 4);
  const int smem_a = (kWarps * 256 + kTile) * 4, smem_b = (kTileWords + 16 * kThreads / 2 + 32) * 4;
  cudaFuncSetAttribute(sort_a, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_a);
  cudaFuncSetAttribute(sort_b, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_b);
  for (int mode = 0; mode < 2; ++mode) {
    srand(7);
    for (int64_t i = 0; i < n; ++i) {
      uint32_t r = ((uint32_t)rand() << 16) ^ (uint32_t)rand();
      if (mode == 1 && rand() % 10) r = (r & 0xffff00ffu) | 0x7e00u;  // 90 % of the keys share the digit at shift 8
      h[i] = r;
    }
    cudaMemcpy(din, h.data(), n * 4, cudaMemcpyHostToDevice);
    const int shift = 8;
    float ms[2];
    for (int v = 0; v < 2; ++v) {
      cudaEvent_t e0, e1;
      cudaEventCreate(&e0); cudaEventCreate(&e1);
      for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(e0);
        if (v == 0) sort_a<<<ctas, kThreads, smem_a>>>(din, dout, shift, per);
        else sort_b<<<ctas, kThreads, smem_b>>>(din, dout, shift, per);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms[v], e0, e1);
      }
      cudaError_t err = cudaGetLastError();
      if (err != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(err)); return 1; }
      cudaMemcpy(got.data(), dout, n * 4, cudaMemcpyDeviceToHost);
      int64_t bad = 0;
      for (int64_t t0 = 0; t0 < n; t0 += kTile) {
        if (t0 / kTile % 97) continue;  // check a sample of the tiles
        std::vector<uint32_t> ref(h.begin() + t0, h.begin() + t0 + kTile);
        std::stable_sort(ref.begin(), ref.end(), [&](uint32_t x, uint32_t y) { return ((x >> shift) & 255) < ((y >> shift) & 255); });
        for (int i = 0; i < kTile; ++i) bad += ref[i] != got[t0 + i];
      }
      printf("%s  %c: %.3f ms  %.1f Gkeys/s  mismatches %lld\n", mode ? "skewed " : "uniform", v ? 'B' : 'A', ms[v],
             n / ms[v] / 1e6, (long long)bad);
    }
  }
  return 0;
}
