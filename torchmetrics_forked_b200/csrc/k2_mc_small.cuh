// K2, multiclass rows with up to 32 classes: one thread per row.  The kernel is instantiated for every class count
// (31 of them, two passes each); one translation unit per score dtype (k2_mc_small_{f32,bf16,f16}.cu) includes this
// file, so that they compile side by side with the other sources instead of one after the other.
#pragma once
#include "k2_binning.cuh"

namespace tmb {

// ---- multiclass binned curves, C <= 32: one THREAD per row ------------------------------------------------------
// For up to 32 classes torch's softmax layout gives every lane one element (W = next_pow2(C) lanes per row), so its
// arithmetic is: max over the row, e_c = expf(x_c - max), and the xor-butterfly sum — which is the balanced tree
// "s[i] += s[i + W/2], s[i] += s[i + W/4], ..." over the W values (zeros past C; float addition is commutative, so
// every lane of the butterfly ends with this very tree).  A single thread holding the row reproduces that bit for
// bit with W - 1 adds and no shuffles; the sub-warp version spends ~8 shuffle steps per row on one element per
// lane.  Rows come through shared memory (coalesced loads, odd row stride: conflict-free reads).

template <int PDT, bool SOFT, int CX>
__global__ void __launch_bounds__(kSmallRows) k2_mc_small(TbArgs a, int64_t rows, int micro, int sh_bins,
                                                          int low_k) {
  constexpr int C = CX;  // the class count is a compile-time constant: every "c < C" below folds away
  constexpr int W = CX <= 2 ? 2 : (CX <= 4 ? 4 : (CX <= 8 ? 8 : (CX <= 16 ? 16 : 32)));  // torch's lanes per row
  extern __shared__ uint32_t sh_u32[];
  if (!pass_enabled(a, SOFT)) return;
  float* sh_pad = (float*)sh_u32;
  const int pad_words = (a.T + 2 + 31) / 32 * 32;
  const int kbins = a.T + 1;
  uint32_t* hist = sh_u32 + pad_words;
  const int hist_words = (sh_bins + 31) / 32 * 32;
  float* stage = (float*)(hist + hist_words);  // [kSmallRows][stride]
  const int stride = C | 1;
  for (int i = threadIdx.x; i < sh_bins; i += blockDim.x) hist[i] = 0;
  Binner bn = make_binner(a, sh_pad, 1 << 30, nullptr);  // contains __syncthreads
  bn.gt = false;  // tmb_mc_curve_update only knows `>=`: a compile-time constant for the compares below
  const bool on_grid = bn.grid_eps < 0.01f;
  unsigned long long* gh = hist_base(a, SOFT ? 1 : 0);
  const uint32_t magic = C > 1 ? (uint32_t)(0x100000000ull / (unsigned)C) + 1u : 0u;  // e / C for e * C < 2^32
  const int class_stride = micro ? 0 : 2 * low_k;  // [C][2][low_k] counters in shared memory
  bool bad_target = false;
  int stop = 0;
  constexpr int kGroups = (W + kPer - 1) / kPer;
  for (int64_t tile0 = (int64_t)blockIdx.x * kSmallRows; tile0 < rows && !stop; tile0 += (int64_t)gridDim.x * kSmallRows) {
    const int n_rows = (int)min((int64_t)kSmallRows, rows - tile0);
    const int total = n_rows * C;
    __syncthreads();  // the previous tile's rows have been read
    for (int e = threadIdx.x; e < total; e += kSmallRows) {
      const int r = C > 1 ? (int)__umulhi((uint32_t)e, magic) : e;
      stage[r * stride + (e - r * C)] = load_elem<PDT>(a.preds, tile0 * C + e);
    }
    __syncthreads();
    const int r = threadIdx.x;
    const bool active = r < n_rows;
    const int64_t t = active ? load_label(a.target, a.tdt, tile0 + r) : 0;
    const bool counted = active && !(a.has_ignore && t == a.ignore);
    if (counted && (t < 0 || t >= C)) bad_target = true;
    const int tc = (int)t;
    float e[W];
#pragma unroll
    for (int c = 0; c < W; ++c) e[c] = 0.0f;
    if (counted) {
#pragma unroll
      for (int c = 0; c < W; ++c) {
        if (c >= C) break;  // CTA-uniform
        e[c] = stage[r * stride + c];
      }
    }
    if (!SOFT) {
      bool mine = false;
      if (counted) {
#pragma unroll
        for (int c = 0; c < W; ++c) {
          if (c >= C) break;
          mine |= !(e[c] >= 0.0f && e[c] <= 1.0f);
        }
      }
      if (mine) raise_flag(a);
      stop = __syncthreads_or(mine || (threadIdx.x == 0 && flag_up(a)));
      if (stop) break;
    } else if (counted) {
      softmax_row_thread<PDT, W>(e, C);
    }
#pragma unroll
    for (int gi = 0; gi < kGroups; ++gi) {
      if (gi * kPer >= C) break;  // CTA-uniform
      Decoded d;
      d.bad = 0;
      d.outside = 0;
#pragma unroll
      for (int i = 0; i < kPer; ++i) {
        const int c = gi * kPer + i;
        const bool valid = c < W && counted && c < C;
        d.x[i] = valid ? e[c < W ? c : 0] : 0.5f;
        d.inc[i] = valid ? 1u : 0u;
        d.toff[i] = 0u;
      }
      int g[kPer];
      if (on_grid)
        batch_bins<PDT, false, true>(bn, d, g);
      else
        batch_bins<PDT, false, false>(bn, d, g);
#pragma unroll
      for (int i = 0; i < kPer; ++i) {
        const int c = gi * kPer + i;
        if (d.inc[i]) {
          const bool hot = c == tc;
          if (g[i] < low_k)
            atomicAdd(&hist[c * class_stride + (hot ? low_k : 0) + g[i]], 1u);
          else
            atomicAdd(&gh[((micro ? 0 : c) * 2 + (hot ? 1 : 0)) * kbins + g[i]], 1ULL);
        }
      }
    }
  }
  if (bad_target) atomicOr(a.err, TMB_ERR_TARGET_RANGE);
  __syncthreads();
  if (!SOFT && (stop || flag_up(a))) return;  // finalize discards the raw histogram once the flag is up
  for (int i = threadIdx.x; i < sh_bins; i += blockDim.x) {
    const uint32_t v = hist[i];
    if (v) {
      const int64_t dst = (int64_t)(i / low_k) * kbins + (i % low_k);
      atomicAdd(&gh[dst], (unsigned long long)v);
    }
  }
}


template <int PDT, int CX>
static int launch_small_c(const TbArgs& a, int64_t rows, int micro, int small_bins, int low_k, int smem,
                          cudaStream_t st) {
  auto ka = k2_mc_small<PDT, false, CX>;
  auto kb = k2_mc_small<PDT, true, CX>;
  cudaFuncSetAttribute(ka, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(kb, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int64_t work = (rows + kSmallRows - 1) / kSmallRows;
  const int grid = resident_grid(kb, kSmallRows, smem, work);
  ka<<<grid, kSmallRows, smem, st>>>(a, rows, micro, small_bins, low_k);
  if (int rc = launch_status()) return rc;
  kb<<<grid, kSmallRows, smem, st>>>(a, rows, micro, small_bins, low_k);
  return launch_status();
}

template <int PDT>
static int launch_small_p(const TbArgs& a, int64_t rows, int C, int micro, int small_bins, int low_k, int smem,
                          cudaStream_t st) {
  switch (C) {
#define TMB_SMALL_CASE(N) \
  case N: return launch_small_c<PDT, N>(a, rows, micro, small_bins, low_k, smem, st);
    TMB_SMALL_CASE(2) TMB_SMALL_CASE(3) TMB_SMALL_CASE(4) TMB_SMALL_CASE(5) TMB_SMALL_CASE(6) TMB_SMALL_CASE(7)
    TMB_SMALL_CASE(8) TMB_SMALL_CASE(9) TMB_SMALL_CASE(10) TMB_SMALL_CASE(11) TMB_SMALL_CASE(12) TMB_SMALL_CASE(13)
    TMB_SMALL_CASE(14) TMB_SMALL_CASE(15) TMB_SMALL_CASE(16) TMB_SMALL_CASE(17) TMB_SMALL_CASE(18) TMB_SMALL_CASE(19)
    TMB_SMALL_CASE(20) TMB_SMALL_CASE(21) TMB_SMALL_CASE(22) TMB_SMALL_CASE(23) TMB_SMALL_CASE(24) TMB_SMALL_CASE(25)
    TMB_SMALL_CASE(26) TMB_SMALL_CASE(27) TMB_SMALL_CASE(28) TMB_SMALL_CASE(29) TMB_SMALL_CASE(30) TMB_SMALL_CASE(31)
    TMB_SMALL_CASE(32)
#undef TMB_SMALL_CASE
    default: return TMB_E_SHAPE;
  }
}

}  // namespace tmb
