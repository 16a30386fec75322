// K5 — multiclass curves: the row softmax of
//   _multiclass_precision_recall_curve_format   functional/classification/precision_recall_curve.py:423-458
// (`if not torch.all((preds >= 0) * (preds <= 1)): preds = preds.softmax(1)`, evaluated AFTER the rows with
// target == ignore_index were removed) as a conditional, bit-exact kernel.
//
// Threshold compares and sort order are decided by the low bits of the probabilities, so the softmax has to
// return what torch's CUDA kernel returns.  For a contiguous [M, C] matrix with C <= 1024 that kernel is the
// "persistent warp softmax" (aten/src/ATen/native/cuda/PersistentSoftmax.cuh, softmax_warp_forward); its result
// is fixed by four things, all reproduced here:
//   * P = next power of two >= C, W = min(32, P) lanes per row, element c lives in lane c % W, slot c / W;
//     slots past C hold -inf
//   * max: per lane over its slots in slot order with `m > e ? m : e`, then an xor butterfly over the W lanes
//     (offsets W/2 .. 1) with `a < b ? b : a`
//   * e = expf(x - max) (full-precision expf), per-lane sum in slot order starting from 0, same butterfly with +
//   * result e / sum (IEEE division), rounded to the input dtype for bf16 / f16
// Rows wider than 1024 use a different torch kernel (block reduction); the host side keeps torch's own softmax
// for those and only takes the in/out-of-range decision from here.
#include "softmax_rows.cuh"

namespace tmb {

__device__ __forceinline__ unsigned long long ld_gen(const unsigned long long* p) {  // device-scope relaxed load
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// pass A: atomicMax(flag, gen) as soon as a value outside [0,1] (NaN included) shows up in a row that is not
// ignored.  Same flag protocol as k3_outside_flag.
// With `out` non-null the rows are copied while they are tested (ignored rows too): for inputs that already are
// probabilities this one pass is the whole normalisation (pass B returns at once), so the data is read once.
template <int DT>
__global__ void __launch_bounds__(256) k5_rows_outside_flag(const void* __restrict__ x, void* __restrict__ out,
                                                            int64_t rows, int C, RowMap rm,
                                                            const void* __restrict__ target, int tdt,
                                                            int has_ignore, int64_t ignore, unsigned long long* flag,
                                                            unsigned long long gen) {
  __shared__ int go;
  if (threadIdx.x == 0) go = ld_gen(flag) != gen;
  __syncthreads();
  if (!go) return;
  const int lane = threadIdx.x & 31;
  const int sub = lane >> rm.log2W, j = lane & (rm.W - 1);
  const int rows_per_warp = 32 >> rm.log2W;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  bool found = false;
  int trip = 0;
  for (int64_t r0 = warp * rows_per_warp; r0 < rows; r0 += nwarps * rows_per_warp, ++trip) {  // warp-uniform
    const int64_t r = r0 + sub;
    bool counted = r < rows;
    const bool in_range = counted;
    if (counted && has_ignore && target != nullptr) counted = load_label(target, tdt, r) != ignore;
    if (counted || (in_range && out != nullptr)) {
      for (int c = j; c < C; c += rm.W) {
        const float v = load_elem<DT>(x, r * C + c);
        if (out != nullptr) store_elem<DT>(out, r * C + c, v);
        found |= counted & !(v >= 0.0f && v <= 1.0f);
      }
    }
    if (__any_sync(kFull, found)) break;
    if ((trip & 63) == 63 && ld_gen(flag) == gen) break;
  }
  if (__syncthreads_or(found) && threadIdx.x == 0 && ld_gen(flag) != gen) atomicMax(flag, gen);  // once per CTA
}

// pass B: out = (flag == gen) ? softmax(x, dim=1) : x.  A lane takes 8 elements per trip — SLOTS of each of
// ROWS = 8 / SLOTS rows when a row gives it fewer — so that narrow rows still keep several loads in flight per lane
// (one 400-byte row per warp and trip left the kernel waiting on DRAM latency at 35 % of the HBM roofline), and the
// rows of the next trip are requested into L2 while this trip is computed.
template <int DT, int SLOTS>
__global__ void __launch_bounds__(256) k5_softmax_rows(const void* __restrict__ x, void* __restrict__ out,
                                                       int64_t rows, int C, RowMap rm,
                                                       const unsigned long long* __restrict__ flag,
                                                       unsigned long long gen) {
  constexpr int ROWS = SLOTS < 8 ? 8 / SLOTS : 1;
  if (*flag != gen) return;  // nothing outside [0,1]: the copy made by pass A is the result
  const bool soft = true;
  const int lane = threadIdx.x & 31;
  const int W = SLOTS > 1 ? 32 : rm.W, log2W = SLOTS > 1 ? 5 : rm.log2W;  // several slots: the row takes the warp
  const int sub = lane >> log2W, j = lane & (W - 1);
  const int rows_per_warp = 32 >> log2W;
  const int rows_per_trip = rows_per_warp * ROWS;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  // one 64-bit pointer pair per trip; what a lane adds is a 32-bit loop-invariant offset (see k2_mc_rows)
  const int trip_elems = rows_per_trip * C;
  const int64_t first = warp * trip_elems * Elem<DT>::kBytes, step = nwarps * trip_elems * Elem<DT>::kBytes;
  const char* tp = (const char*)x + first;
  char* op = (char*)out + first;
  const char* const x_end = (const char*)x + rows * C * Elem<DT>::kBytes;
  for (int64_t r0 = warp * rows_per_trip; r0 < rows; r0 += nwarps * rows_per_trip, tp += step, op += step) {
    {
      const char* nx = tp + step;
      const char* mine = (const char*)((uintptr_t)nx & ~(uintptr_t)127) + lane * 128;
      if (mine < nx + trip_elems * Elem<DT>::kBytes && mine < x_end) prefetch_l2(mine);
    }
    const int left = (int)min((int64_t)rows_per_trip, rows - r0);
    float e[ROWS][SLOTS];
#pragma unroll
    for (int q = 0; q < ROWS; ++q) {
      const int row = q * rows_per_warp + sub;
      const char* lp = tp + (row * C + j) * Elem<DT>::kBytes;  // slot offsets become immediates of the loads
#pragma unroll
      for (int it = 0; it < SLOTS; ++it) e[q][it] = (row < left && it * W + j < C) ? load_elem<DT>(lp, it * W) : -INFINITY;
    }
    if (soft) softmax_rows_inplace<DT, SLOTS, ROWS>(e, rm);
#pragma unroll
    for (int q = 0; q < ROWS; ++q) {
      const int row = q * rows_per_warp + sub;
      char* sp = op + (row * C + j) * Elem<DT>::kBytes;
#pragma unroll
      for (int it = 0; it < SLOTS; ++it)
        if (row < left && it * W + j < C) store_elem<DT>(sp, it * W, e[q][it]);
    }
  }
}

// C <= 32: one thread per row (softmax_row_thread), rows staged through shared memory in both directions so that
// global loads and stores stay coalesced.
constexpr int kSoftmaxSmallRows = 256;

template <int DT, int CX>
__global__ void __launch_bounds__(kSoftmaxSmallRows) k5_softmax_small(const void* __restrict__ x, void* __restrict__ out,
                                                                      int64_t rows,
                                                                      const unsigned long long* __restrict__ flag,
                                                                      unsigned long long gen) {
  constexpr int C = CX;  // compile-time class count: the per-class loops carry no bounds checks
  constexpr int W = CX <= 2 ? 2 : (CX <= 4 ? 4 : (CX <= 8 ? 8 : (CX <= 16 ? 16 : 32)));  // torch's lanes per row
  extern __shared__ float stage[];  // [kSoftmaxSmallRows][C | 1]
  if (*flag != gen) return;  // nothing outside [0,1]: the copy made by pass A is the result
  const bool soft = true;
  constexpr int stride = C | 1;
  constexpr uint32_t magic = (uint32_t)(0x100000000ull / (unsigned)C) + 1u;  // i / C for i * C < 2^32
  for (int64_t tile0 = (int64_t)blockIdx.x * kSoftmaxSmallRows; tile0 < rows;
       tile0 += (int64_t)gridDim.x * kSoftmaxSmallRows) {
    const int n_rows = (int)min((int64_t)kSoftmaxSmallRows, rows - tile0);
    const int total = n_rows * C;
    const char* src = (const char*)x + tile0 * C * Elem<DT>::kBytes;
    char* dst = (char*)out + tile0 * C * Elem<DT>::kBytes;
    __syncthreads();
    for (int i = threadIdx.x; i < total; i += kSoftmaxSmallRows) {
      const int r = (int)__umulhi((uint32_t)i, magic);
      stage[r * stride + (i - r * C)] = load_elem<DT>(src, i);
    }
    __syncthreads();
    if (soft && (int)threadIdx.x < n_rows) {
      float e[W];
#pragma unroll
      for (int c = 0; c < W; ++c) e[c] = c < C ? stage[threadIdx.x * stride + (c < C ? c : 0)] : 0.0f;
      softmax_row_thread<DT, W>(e, C);
#pragma unroll
      for (int c = 0; c < C; ++c) stage[threadIdx.x * stride + c] = e[c];
    }
    __syncthreads();
    for (int i = threadIdx.x; i < total; i += kSoftmaxSmallRows) {
      const int r = (int)__umulhi((uint32_t)i, magic);
      store_elem<DT>(dst, i, stage[r * stride + (i - r * C)]);
    }
  }
}

template <int DT>
static int launch_rows_flag(const void* x, void* out, int64_t rows, int C, const void* target, int tdt, int has_ignore,
                            int64_t ignore, unsigned long long* flag, unsigned long long gen, cudaStream_t st) {
  const RowMap rm = row_map(C < kSoftmaxMaxClasses ? C : kSoftmaxMaxClasses);  // wider rows: 32 lanes stride over C
  auto k = k5_rows_outside_flag<DT>;
  const int64_t work = (rows * rm.W + 255) / 256;
  k<<<resident_grid(k, 256, 0, work), 256, 0, st>>>(x, out, rows, C, rm, target, tdt, has_ignore, ignore, flag, gen);
  return launch_status();
}

template <int DT>
static int launch_softmax_rows(const void* x, void* out, int64_t rows, int C, const unsigned long long* flag,
                               unsigned long long gen, cudaStream_t st) {
  const RowMap rm = row_map(C);
  if (C <= 32 && C >= 2) {
    const int smem = kSoftmaxSmallRows * (C | 1) * 4;
    const int64_t work = (rows + kSoftmaxSmallRows - 1) / kSoftmaxSmallRows;
    switch (C) {
#define TMB_SM_CASE(N)                                                                                       \
  case N: {                                                                                                  \
    auto k = k5_softmax_small<DT, N>;                                                                        \
    k<<<resident_grid(k, kSoftmaxSmallRows, smem, work), kSoftmaxSmallRows, smem, st>>>(x, out, rows, flag, gen); \
  } break;
      TMB_SM_CASE(2) TMB_SM_CASE(3) TMB_SM_CASE(4) TMB_SM_CASE(5) TMB_SM_CASE(6) TMB_SM_CASE(7) TMB_SM_CASE(8)
      TMB_SM_CASE(9) TMB_SM_CASE(10) TMB_SM_CASE(11) TMB_SM_CASE(12) TMB_SM_CASE(13) TMB_SM_CASE(14) TMB_SM_CASE(15)
      TMB_SM_CASE(16) TMB_SM_CASE(17) TMB_SM_CASE(18) TMB_SM_CASE(19) TMB_SM_CASE(20) TMB_SM_CASE(21) TMB_SM_CASE(22)
      TMB_SM_CASE(23) TMB_SM_CASE(24) TMB_SM_CASE(25) TMB_SM_CASE(26) TMB_SM_CASE(27) TMB_SM_CASE(28) TMB_SM_CASE(29)
      TMB_SM_CASE(30) TMB_SM_CASE(31) TMB_SM_CASE(32)
#undef TMB_SM_CASE
    }
    return launch_status();
  }
  TMB_DISPATCH_SLOTS(rm.slots, {
    constexpr int kRowsPerLane = SLOTS < 8 ? 8 / SLOTS : 1;
    const int64_t work = (rows * rm.W / kRowsPerLane + 255) / 256 + 1;
    auto k = k5_softmax_rows<DT, SLOTS>;
    k<<<resident_grid(k, 256, 0, work), 256, 0, st>>>(x, out, rows, C, rm, flag, gen);
  })
  return launch_status();
}

}  // namespace tmb

using namespace tmb;

extern "C" int tmb_rows_outside_flag(const void* preds, int preds_dtype, int64_t n_rows, int64_t num_classes,
                                     const void* target, int target_dtype, int has_ignore, int64_t ignore_index,
                                     int64_t* flag, int64_t generation, void* stream) {
  if (float_bytes(preds_dtype) == 0) return TMB_E_DTYPE;
  if (target != nullptr && label_bytes(target_dtype) == 0) return TMB_E_DTYPE;
  if (n_rows < 0 || num_classes < 1 || num_classes > (1 << 30) || generation < 1) return TMB_E_SHAPE;
  if (!flag) return TMB_E_NULL;
  if (n_rows == 0) return 0;
  if (!preds) return TMB_E_NULL;
  cudaStream_t st = (cudaStream_t)stream;
  unsigned long long* f = (unsigned long long*)flag;
  const unsigned long long gen = (unsigned long long)generation;
  const int C = (int)num_classes;
  switch (preds_dtype) {
    case TMB_F32: return launch_rows_flag<TMB_F32>(preds, nullptr, n_rows, C, target, target_dtype, has_ignore, ignore_index, f, gen, st);
    case TMB_F16: return launch_rows_flag<TMB_F16>(preds, nullptr, n_rows, C, target, target_dtype, has_ignore, ignore_index, f, gen, st);
    default: return launch_rows_flag<TMB_BF16>(preds, nullptr, n_rows, C, target, target_dtype, has_ignore, ignore_index, f, gen, st);
  }
}

extern "C" int tmb_softmax_normalize(const void* preds, int preds_dtype, int64_t n_rows, int64_t num_classes,
                                     const void* target, int target_dtype, int has_ignore, int64_t ignore_index,
                                     void* out, int64_t* flag, int64_t generation, void* stream) {
  if (float_bytes(preds_dtype) == 0) return TMB_E_DTYPE;
  if (target != nullptr && label_bytes(target_dtype) == 0) return TMB_E_DTYPE;
  if (n_rows < 0 || num_classes < 1 || generation < 1) return TMB_E_SHAPE;
  if (num_classes > kSoftmaxMaxClasses) return TMB_E_LIMIT;
  if (!flag) return TMB_E_NULL;
  if (n_rows == 0) return 0;
  if (!preds || !out) return TMB_E_NULL;
  cudaStream_t st = (cudaStream_t)stream;
  unsigned long long* f = (unsigned long long*)flag;
  const unsigned long long gen = (unsigned long long)generation;
  const int C = (int)num_classes;
  int rc;
  switch (preds_dtype) {  // pass A: copy + test, ends at the first value outside [0,1]
    case TMB_F32: rc = launch_rows_flag<TMB_F32>(preds, out, n_rows, C, target, target_dtype, has_ignore, ignore_index, f, gen, st); break;
    case TMB_F16: rc = launch_rows_flag<TMB_F16>(preds, out, n_rows, C, target, target_dtype, has_ignore, ignore_index, f, gen, st); break;
    default: rc = launch_rows_flag<TMB_BF16>(preds, out, n_rows, C, target, target_dtype, has_ignore, ignore_index, f, gen, st); break;
  }
  if (rc) return rc;
  switch (preds_dtype) {  // pass B: softmax, only when pass A raised the flag
    case TMB_F32: return launch_softmax_rows<TMB_F32>(preds, out, n_rows, C, f, gen, st);
    case TMB_F16: return launch_softmax_rows<TMB_F16>(preds, out, n_rows, C, f, gen, st);
    default: return launch_softmax_rows<TMB_BF16>(preds, out, n_rows, C, f, gen, st);
  }
}

extern "C" int tmb_softmax_rows(const void* preds, int preds_dtype, int64_t n_rows, int64_t num_classes, void* out,
                                const int64_t* flag, int64_t generation, void* stream) {
  if (float_bytes(preds_dtype) == 0) return TMB_E_DTYPE;
  if (n_rows < 0 || num_classes < 1 || generation < 1) return TMB_E_SHAPE;
  if (num_classes > kSoftmaxMaxClasses) return TMB_E_LIMIT;
  if (!flag) return TMB_E_NULL;
  if (n_rows == 0) return 0;
  if (!preds || !out) return TMB_E_NULL;
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned long long* f = (const unsigned long long*)flag;
  const unsigned long long gen = (unsigned long long)generation;
  const int C = (int)num_classes;
  switch (preds_dtype) {
    case TMB_F32: return launch_softmax_rows<TMB_F32>(preds, out, n_rows, C, f, gen, st);
    case TMB_F16: return launch_softmax_rows<TMB_F16>(preds, out, n_rows, C, f, gen, st);
    default: return launch_softmax_rows<TMB_BF16>(preds, out, n_rows, C, f, gen, st);
  }
}
