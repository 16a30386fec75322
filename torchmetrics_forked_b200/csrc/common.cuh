// Shared device/host helpers for the tmb200 kernels (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/tmb200.h"

#ifndef __CUDA_ARCH__
#define TMB_HOST_ONLY 1
#endif

namespace tmb {

constexpr int kWarp = 32;
constexpr unsigned kFull = 0xffffffffu;

// ---- device properties, cached per device ------------------------------------------------------
struct DevInfo {
  int sm_count;
  int max_smem_optin;
};

inline const DevInfo& dev_info() {
  static DevInfo info[64];
  static bool have[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (!have[dev]) {
    int sms = 148, smem = 227 * 1024;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    info[dev].sm_count = sms > 0 ? sms : 148;
    info[dev].max_smem_optin = smem;
    have[dev] = true;
  }
  return info[dev];
}

// grid for a grid-stride kernel: whole waves of `per_sm` CTAs, never more CTAs than work items
inline int wave_grid(int64_t work_ctas, int per_sm) {
  int64_t cap = (int64_t)dev_info().sm_count * per_sm;
  if (work_ctas < 1) work_ctas = 1;
  return (int)(work_ctas < cap ? work_ctas : cap);
}

// grid for a grid-stride kernel whose CTAs must all be co-resident to share the work evenly: whole
// waves of exactly as many CTAs as the occupancy calculator says fit (registers / smem of THIS kernel)
template <typename Kernel>
inline int resident_grid(Kernel kernel, int block, size_t dyn_smem, int64_t work_ctas) {
  int per_sm = 1;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, block, dyn_smem) != cudaSuccess || per_sm < 1) {
    cudaGetLastError();
    per_sm = 1;
  }
  return wave_grid(work_ctas, per_sm);
}

inline int launch_status() {
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : (int)e;
}

// ---- streaming loads: read-only path, do not pollute L1 ------------------------------------------
__device__ __forceinline__ uint4 ldg_stream16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ uint2 ldg_stream8(const void* p) {
  uint2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ uint32_t ldg_stream4(const void* p) {
  uint32_t r;
  asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ uint16_t ldg_stream2(const void* p) {
  uint16_t r;
  asm volatile("ld.global.nc.L1::no_allocate.u16 %0, [%1];" : "=h"(r) : "l"(p));
  return r;
}

// pull the 128-byte line holding `p` into L2 without occupying a register or a scoreboard slot
__device__ __forceinline__ void prefetch_l2(const void* p) {
  asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}
__host__ __device__ constexpr int label_bytes_dev(int dtype) {
  return dtype == TMB_I64 ? 8 : (dtype == TMB_I32 ? 4 : (dtype == TMB_I16 ? 2 : 1));
}

// ---- element decoding ---------------------------------------------------------------------------
// Every float dtype is widened to fp32 exactly (bf16/f16 -> f32 is lossless), so comparisons give
// the same answer torch gets when it compares the narrow values.
template <int DT>
struct Elem;
template <>
struct Elem<TMB_F32> {
  static constexpr int kBytes = 4;
  __device__ static __forceinline__ float get(uint32_t word, int) { return __uint_as_float(word); }
};
template <>
struct Elem<TMB_BF16> {
  static constexpr int kBytes = 2;
  __device__ static __forceinline__ float get(uint32_t word, int half) {
    return __uint_as_float(half ? (word & 0xffff0000u) : (word << 16));
  }
};
template <>
struct Elem<TMB_F16> {
  static constexpr int kBytes = 2;
  __device__ static __forceinline__ float get(uint32_t word, int half) {
    return __half2float(__ushort_as_half((unsigned short)(half ? (word >> 16) : (word & 0xffffu))));
  }
};

template <int DT>
__device__ __forceinline__ float load_elem(const void* base, int64_t i) {
  if constexpr (DT == TMB_F32) {
    return __uint_as_float(ldg_stream4((const char*)base + i * 4));
  } else {
    uint32_t w = ldg_stream2((const char*)base + i * 2);
    return Elem<DT>::get(w, 0);
  }
}

// integer labels of any supported width -> int64
__device__ __forceinline__ int64_t load_label(const void* p, int dtype, int64_t i) {
  switch (dtype) {
    case TMB_I64: return __ldg((const long long*)p + i);
    case TMB_I32: return __ldg((const int*)p + i);
    case TMB_I16: return __ldg((const short*)p + i);
    case TMB_U8: return __ldg((const unsigned char*)p + i);
    default: return __ldg((const signed char*)p + i);
  }
}

inline int label_bytes(int dtype) {
  switch (dtype) {
    case TMB_I64: return 8;
    case TMB_I32: return 4;
    case TMB_I16: return 2;
    case TMB_U8: return 1;
    case TMB_I8: return 1;
    default: return 0;
  }
}
inline int float_bytes(int dtype) {
  switch (dtype) {
    case TMB_F32: return 4;
    case TMB_F16: return 2;
    case TMB_BF16: return 2;
    default: return 0;
  }
}

// ---- argmax bookkeeping: torch.argmax rule = first maximal index, NaN counts as maximal -----------
struct Best {
  float v;
  int i;
};
__device__ __forceinline__ Best best_init() { return Best{-INFINITY, 0x7fffffff}; }

// scan step, candidates visited in increasing index order inside one thread
__device__ __forceinline__ void best_step(Best& b, float v, int i) {
  // v replaces b when it is strictly greater, or when it is the first NaN
  // (bitwise operators and selects: with `||` / `&&` and an `if` the compiler emitted a branch per element)
  const bool take = (v > b.v) | ((v != v) & (b.v == b.v));
  b.v = take ? v : b.v;
  b.i = take ? i : b.i;
}
// merge of two partial results that cover disjoint index sets
__device__ __forceinline__ Best best_merge(Best a, Best o) {
  const bool an = a.v != a.v, on = o.v != o.v, lower = o.i < a.i;
  const bool take = on ? (!an | lower) : (!an & ((o.v > a.v) | ((o.v == a.v) & lower)));
  Best r;
  r.v = take ? o.v : a.v;
  r.i = take ? o.i : a.i;
  return r;
}
__device__ __forceinline__ Best best_warp(Best b) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    Best o;
    o.v = __shfl_xor_sync(kFull, b.v, d);
    o.i = __shfl_xor_sync(kFull, b.i, d);
    b = best_merge(b, o);
  }
  return b;
}
// a row/pixel whose every element is -inf never triggers best_step: its argmax is index 0
__device__ __forceinline__ int best_index(const Best& b) { return b.i == 0x7fffffff ? 0 : b.i; }

}  // namespace tmb
