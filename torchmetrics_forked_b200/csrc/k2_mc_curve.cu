// K2, multiclass rows: binned precision-recall / ROC state update with the row softmax fused in.
// Split from k2_thresh_bin.cu so that the two translation units compile side by side.
#include "k2_binning.cuh"

namespace tmb {

// ---- multiclass binned curves: row softmax fused into the binning ------------------------------
// _multiclass_precision_recall_curve_format/_update (precision_recall_curve.py:423-527): softmax over the classes
// if any kept value is outside [0,1], then every (row, class) element is binned against the thresholds with the
// one-hot target `class == target[row]`.  The reference materialises the probabilities and an N x C x T
// comparison (or loops T times over N x C); here a sub-warp keeps a row in registers in the lane layout of
// torch's persistent softmax (softmax_rows.cuh: bit-identical probabilities), bins its elements and adds them to
// the CTA-wide [C][2][T+1] histogram in shared memory — the logits are read once, nothing is written but counts.
// `micro` folds all classes into label 0 (the reference flattens preds and one-hot targets into one binary task).
// Histograms that do not fit one SM go to global memory with one atomic per element.
// A lane takes kPer elements per trip through the binning code: SLOTS elements of each of ROWS = kPer / SLOTS rows
// when a row gives it fewer than kPer (C <= 128), so narrow rows (C = 10: one element per lane and row) still fill
// the batch and keep ROWS loads in flight.  When the [C][2][T+1] histogram does not fit one SM (C = 1000, T = 200
// is 1.6 MB) only the lowest `low_k` bins of every (class, target) pair live in shared memory — softmax mass of a
// wide row sits near 0 — and the rest goes to global memory with one atomic per element.
template <int PDT, bool SOFT, int SLOTS>
__global__ void __launch_bounds__(SLOTS <= 8 ? 1024 : 512, 1)
    k2_mc_rows(TbArgs a, RowMap rm, int64_t rows, int C, int micro, int sh_bins, int low_k) {
  constexpr int ROWS = SLOTS < kPer ? kPer / SLOTS : 1;
  constexpr int kGroups = (SLOTS + kPer - 1) / kPer;
  extern __shared__ uint32_t sh_u32[];
  if (!pass_enabled(a, SOFT)) return;
  float* sh_pad = (float*)sh_u32;
  const int pad_words = (a.T + 2 + 31) / 32 * 32;
  const int kbins = a.T + 1;
  uint32_t* hist = sh_u32 + pad_words;
  for (int i = threadIdx.x; i < sh_bins; i += blockDim.x) hist[i] = 0;
  Binner bn = make_binner(a, sh_pad, 1 << 30, nullptr);  // contains __syncthreads
  bn.gt = false;  // tmb_mc_curve_update only knows `>=`: a compile-time constant for the compares below
  const bool on_grid = bn.grid_eps < 0.01f;
  unsigned long long* gh = hist_base(a, SOFT ? 1 : 0);
  const int lane = threadIdx.x & 31;
  // more than one slot per lane means the row takes the whole warp: W is a compile-time 32 there, which turns the
  // slot offsets into immediates of the load instructions
  const int W = SLOTS > 1 ? 32 : rm.W, log2W = SLOTS > 1 ? 5 : rm.log2W;
  const int sub = lane >> log2W, j = lane & (W - 1);
  const int rows_per_warp = 32 >> log2W;
  const int rows_per_trip = rows_per_warp * ROWS;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  bool stop = false;
  unsigned bad_target = 0;
  int trip = 0;
  // Addresses: one 64-bit pointer per trip (advanced by a constant), everything a lane adds to it is a 32-bit,
  // loop-invariant offset.  (With `r * C + c` spelled out per element the compiler carried a dozen 64-bit induction
  // variables: a third of the kernel's instructions were address arithmetic.)
  const int trip_elems = rows_per_trip * C;  // <= 8 rows of <= 1024 classes
  const char* tp = (const char*)a.preds + warp * trip_elems * Elem<PDT>::kBytes;
  const int64_t tp_step = nwarps * trip_elems * Elem<PDT>::kBytes;
  const char* const preds_end = (const char*)a.preds + rows * C * Elem<PDT>::kBytes;
  const bool labels64 = a.tdt == TMB_I64;
  // shared-memory counters of one class: [2][low_k] and one word of padding — an odd stride, so that the 32 lanes of
  // a warp (32 consecutive classes, and for softmax scores mostly the same one or two bins) hit 32 different banks;
  // with the even stride 2 * low_k the atomics of a warp took 3 (C = 100) to 8 (C = 1000) passes
  const int class_words = 2 * low_k + 1;
  const int class_stride = micro ? 0 : class_words;
  auto count = [&](int c, bool hot, int k) {
    if (k < low_k) {  // always, when the whole histogram is in shared memory (low_k == kbins)
      atomicAdd(&hist[c * class_stride + (hot ? low_k : 0) + k], 1u);
    } else {
      atomicAdd(&gh[((micro ? 0 : c) * 2 + (hot ? 1 : 0)) * kbins + k], 1ULL);
    }
  };
  for (int64_t r0 = warp * rows_per_trip; r0 < rows && !stop; r0 += nwarps * rows_per_trip, tp += tp_step, ++trip) {
    {  // the rows of this warp's NEXT trip are one contiguous region of at most 4 KB: one L2 prefetch per 128-byte
       // line, so that the register loads of that trip wait for L2, not for DRAM (half of all stall samples)
      const char* nx = tp + tp_step;
      const char* mine = (const char*)((uintptr_t)nx & ~(uintptr_t)127) + lane * 128;
      if (mine < nx + trip_elems * Elem<PDT>::kBytes && mine < preds_end) prefetch_l2(mine);
    }
    const int left = (int)min((int64_t)rows_per_trip, rows - r0);  // rows of this trip that exist
    // Scores and labels are requested together and the labels are first looked at after the softmax: a row whose
    // loads waited for `target != ignore_index` paid two memory latencies back to back (13 % of the stall samples).
    // Ignored rows are normalised like the others and dropped when the batch is counted.
    float e[ROWS][SLOTS];
    int64_t label[ROWS];
#pragma unroll
    for (int q = 0; q < ROWS; ++q) {
      const int row = q * rows_per_warp + sub;  // within the trip
      const bool active = row < left;
      label[q] = a.ignore;
      if (active) label[q] = labels64 ? __ldg((const long long*)a.target + (r0 + row)) : load_label(a.target, a.tdt, r0 + row);
      const char* lp = tp + (row * C + j) * Elem<PDT>::kBytes;  // slot offsets become immediates of the loads
#pragma unroll
      for (int it = 0; it < SLOTS; ++it) e[q][it] = (active && it * W + j < C) ? load_elem<PDT>(lp, it * W) : -INFINITY;
    }
    if (SOFT) softmax_rows_inplace<PDT, SLOTS, ROWS>(e, rm);
    int t[ROWS];
    bool counted[ROWS];
#pragma unroll
    for (int q = 0; q < ROWS; ++q) {
      const bool ignored = (q * rows_per_warp + sub >= left) | ((a.has_ignore != 0) & (label[q] == a.ignore));
      counted[q] = !ignored;
      bad_target |= (unsigned)(!ignored & ((uint64_t)label[q] >= (uint64_t)C));  // label < 0 or >= C
      t[q] = (int)label[q];  // out-of-range labels raise the error bit; what they count as does not matter
    }
    if (!SOFT) {
      bool mine = false;
#pragma unroll
      for (int q = 0; q < ROWS; ++q)
#pragma unroll
        for (int it = 0; it < SLOTS; ++it)
          if (counted[q] && it * W + j < C) mine |= !(e[q][it] >= 0.0f && e[q][it] <= 1.0f);
      if (mine) raise_flag(a);
      stop = __any_sync(kFull, mine || ((trip & 7) == 7 && flag_up(a)));
      if (stop) break;
    }
    // One batch of kPer elements through the binning code.  Element i is (row i / SLOTS, slot i % SLOTS) when a
    // lane holds several rows (SLOTS < kPer), else slot slot0 + i of the one row, found in register e[0][i].
    auto bin_group = [&](int slot0) {
      Decoded d;
      d.bad = 0;
      d.outside = 0;
#pragma unroll
      for (int i = 0; i < kPer; ++i) {
        const int q = ROWS > 1 ? i / SLOTS : 0;
        const int reg = ROWS > 1 ? i % SLOTS : (i < SLOTS ? i : 0);
        const int it = ROWS > 1 ? reg : slot0 + i;
        const bool valid = i < ROWS * SLOTS && counted[q] && it * W + j < C;
        d.x[i] = valid ? e[q][reg] : 0.5f;
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
        const int q = ROWS > 1 ? i / SLOTS : 0;
        const int it = ROWS > 1 ? i % SLOTS : slot0 + i;
        if (d.inc[i]) {
          const int c = it * W + j;
          count(c, c == t[q], g[i]);
        }
      }
    };
    if constexpr (kGroups == 1) {
      bin_group(0);
    } else {
      // wide rows: a rolled loop over the groups, the row shifted down by one group after each (the unrolled form
      // was 5400 instructions, 86 KB of code, and a quarter of the stall samples were instruction fetches)
#pragma unroll 1
      for (int gi = 0; gi < kGroups; ++gi) {
        bin_group(gi * kPer);
#pragma unroll
        for (int i = 0; i + kPer < SLOTS; ++i) e[0][i] = e[0][i + kPer];
      }
    }
  }
  if (bad_target) atomicOr(a.err, TMB_ERR_TARGET_RANGE);
  const bool dead = !SOFT && __syncthreads_or(stop);
  if (SOFT) __syncthreads();
  if (dead || (!SOFT && flag_up(a))) return;  // finalize discards the raw histogram once the flag is up
  for (int i = threadIdx.x; i < sh_bins; i += blockDim.x) {
    const uint32_t v = hist[i];
    if (v) {  // never the padding word
      const int cls = i / class_words, rem = i - cls * class_words;
      const int hot = rem >= low_k ? 1 : 0;
      atomicAdd(&gh[(int64_t)(cls * 2 + hot) * kbins + (rem - hot * low_k)], (unsigned long long)v);
    }
  }
}

template <int PDT>
static int launch_mc_rows(const TbArgs& a, int64_t rows, int C, int micro, cudaStream_t st) {
  const RowMap rm = row_map(C);
  const int64_t pad_bytes = (int64_t)(a.T + 2 + 31) / 32 * 32 * 4;
  const int64_t pairs = (int64_t)a.L * 2, kbins = a.T + 1;
  const int64_t room = ((int64_t)dev_info().max_smem_optin - 1024 - pad_bytes) / 4;  // 32-bit counters that fit
  // k2_mc_rows keeps 2 * low_k + 1 words per class (see class_words there), k2_mc_small 2 * low_k
  int low_k = (int)((room / a.L - 1) / 2);  // <= 0 for an absurd class count: everything goes to global memory
  if (low_k > kbins) low_k = (int)kbins;    // every bin is a "low" bin
  if (low_k <= 0) low_k = -1;               // "nothing in shared memory": k < low_k never holds
  const int sh_bins = low_k > 0 ? (int)((int64_t)a.L * (2 * low_k + 1)) : 0;
  const int smem = (int)(pad_bytes + (int64_t)sh_bins * 4);
  int rc = 0;
  if (C <= 32 && low_k > 0) {  // one thread per row (k2_mc_small.cu)
    const int small_bins = (int)(pairs * low_k);
    const int small_smem = (int)pad_bytes + (small_bins + 31) / 32 * 32 * 4 + kSmallRows * (C | 1) * 4;
    if (small_smem <= dev_info().max_smem_optin - 1024) {
      if constexpr (PDT == TMB_F32) return launch_mc_small_f32(a, rows, C, micro, small_bins, low_k, small_smem, st);
      if constexpr (PDT == TMB_BF16) return launch_mc_small_bf16(a, rows, C, micro, small_bins, low_k, small_smem, st);
      if constexpr (PDT == TMB_F16) return launch_mc_small_f16(a, rows, C, micro, small_bins, low_k, small_smem, st);
    }
  }
  TMB_DISPATCH_SLOTS(rm.slots, {
    auto ka = k2_mc_rows<PDT, false, SLOTS>;
    auto kb = k2_mc_rows<PDT, true, SLOTS>;
    cudaFuncSetAttribute(ka, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaFuncSetAttribute(kb, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    const int block = SLOTS <= 8 ? 1024 : 512;
    constexpr int kRowsPerLane = SLOTS < kPer ? kPer / SLOTS : 1;
    const int64_t work = (rows * rm.W / kRowsPerLane + block - 1) / block + 1;
    const int grid = resident_grid(kb, block, smem, work);
    ka<<<grid, block, smem, st>>>(a, rm, rows, C, micro, sh_bins, low_k);
    rc = launch_status();
    if (rc == 0) {
      kb<<<grid, block, smem, st>>>(a, rm, rows, C, micro, sh_bins, low_k);
      rc = launch_status();
    }
  })
  return rc;
}

}  // namespace tmb

using namespace tmb;

extern "C" int tmb_mc_curve_update(const void* preds, int preds_dtype, const void* target, int target_dtype,
                                   int64_t n_rows, int64_t num_classes, const float* thr, int64_t num_thresholds,
                                   int has_ignore, int64_t ignore_index, int micro, int64_t* scratch,
                                   int64_t scratch_elems, int32_t* err, void* stream) {
  if (float_bytes(preds_dtype) == 0 || label_bytes(target_dtype) == 0) return TMB_E_DTYPE;
  if (n_rows < 0 || num_classes < 2) return TMB_E_SHAPE;
  if (num_classes > kSoftmaxMaxClasses) return TMB_E_LIMIT;
  if (num_thresholds < 1 || num_thresholds > (1 << 26)) return TMB_E_SHAPE;
  if (!scratch || !err || !thr) return TMB_E_NULL;
  const int64_t labels = micro ? 1 : num_classes;
  if (scratch_elems < scratch_elems_for(labels, num_thresholds)) return TMB_E_SCRATCH;
  if (n_rows == 0) return 0;
  if (!preds || !target) return TMB_E_NULL;
  cudaStream_t st = (cudaStream_t)stream;
  TbArgs a;
  a.preds = preds;
  a.target = target;
  a.thr = thr;
  a.scratch = (unsigned long long*)scratch;
  a.err = err;
  a.n_total = n_rows * num_classes;
  a.S = 1;
  a.ignore = ignore_index;
  a.L = (int)labels;
  a.T = (int)num_thresholds;
  a.pdt = preds_dtype;
  a.tdt = target_dtype;
  a.preds_is_label = 0;
  a.cmp_gt = 0;
  a.has_ignore = has_ignore;
  a.flag_counts_ignored = 0;
  switch (preds_dtype) {
    case TMB_F32: return launch_mc_rows<TMB_F32>(a, n_rows, (int)num_classes, micro, st);
    case TMB_F16: return launch_mc_rows<TMB_F16>(a, n_rows, (int)num_classes, micro, st);
    default: return launch_mc_rows<TMB_BF16>(a, n_rows, (int)num_classes, micro, st);
  }
}
