// K1 — multiclass argmax + histogram, one pass over the logits.
//
// Replaces the reference's argmax -> flatten -> (boolean-mask) -> target*C+pred -> bincount chain
// (functional/classification/confusion_matrix.py:306-337, stat_scores.py:325-341 and :394-418,
// utilities/data.py:169-199) with kernels that read every logit exactly once and accumulate straight
// into the int64 metric state (confusion matrix) or into a small int64 scratch (stat scores).
//
// Three input layouts, three kernels:
//   rows_warp   [N, C] row-major, long rows   : one warp per row, vector loads, shuffle merge
//   rows_thread [N, C] row-major, short rows  : one thread per row, rows staged through shared memory
//   strided     [N, C, S] class stride S > 1  : one thread per group of adjacent pixels, loop over C
// plus a label kernel for integer preds.  All are HBM-streaming: algorithmic bytes per sample are
// C*sizeof(elem) + sizeof(target), there is no reuse, so the only shared-memory use is the coalescing
// stage of rows_thread and the privatised histogram for small C.
#include "common.cuh"
#include "softmax_rows.cuh"

namespace tmb {

// ---- accumulation target -------------------------------------------------------------------------
struct HistArgs {
  int mode;        // TMB_MC_*
  int C;           // num_classes
  int sh_bins;     // > 0: privatise this many int32 bins in shared memory, flush at CTA exit
  int has_ignore;
  int64_t ignore;
  unsigned long long* out;
  int* err;
  int64_t per_sample;  // > 0 (TMB_MC_STATS_SAMPLEWISE): one [3C] histogram block per sample, `per_sample` bins apart
  int64_t sample_len;  // flattened positions per sample (label-valued preds need it to find their sample)
};

// per-thread state: the micro counters (mode MICRO), flushed once per CTA, and ONE pending (cell, count) pair of the
// run-length aggregation below
struct Micro {
  unsigned int match, n;
  int pbin;           // pending cell of this thread's register slot (valid while pcnt != 0)
  unsigned int pcnt;
};

// Histograms too large for shared memory (C = 1000: 10^6 cells) are updated with RED.64 straight on the int64 state.
// L2 serialises the atomics of one address: a batch whose rows all fall into ONE cell (a collapsed model, a
// single-class validation shard, the background of a segmentation map) ran 4.8x slower than the bench data
// (config 2, 2^20 rows: 1.49 ms against 0.31 ms).  So the atomics are aggregated where the repeats are, in a thread's
// own sequence of samples: it keeps a pending (cell, count) pair per histogram it feeds and only sends it when the
// cell changes.  Random cells cost a compare more per sample; a run of equal cells costs one atomic per thread
// instead of one per sample.
//   * One pair lives in registers (Micro::pbin / pcnt).  In the warp-per-row kernels lane s owns histogram s of the
//     mode (count_sample_warp), so all of a row's histograms are handled side by side in registers.
//   * Kernels where ONE thread classifies a sample keep the pair of the first histogram in registers and those of the
//     others in thread-private columns of shared memory (all four pairs in registers cost the row kernels 18
//     registers and a CTA per SM; all four in shared memory put four dependent LDS round trips behind every row:
//     2.6 % on config 2, 25 % on the four-histogram TMB_MC_BOTH mode).
constexpr int kPendThreads = 256;
__shared__ int g_pend_bin[3][kPendThreads];
__shared__ unsigned int g_pend_cnt[3][kPendThreads];

__device__ __forceinline__ void hist_add_reg(const HistArgs& h, int* sh, Micro& mi, int64_t bin) {
  if (h.sh_bins > 0) {
    atomicAdd(&sh[bin], 1);
    return;
  }
  if (bin > 0x7fffffffLL) {  // cells beyond 2^31 (per-sample blocks of many classes): not aggregated
    atomicAdd(&h.out[bin], 1ULL);
    return;
  }
  if (mi.pcnt != 0 && mi.pbin == (int)bin) {
    mi.pcnt += 1;
    return;
  }
  if (mi.pcnt != 0) atomicAdd(&h.out[mi.pbin], (unsigned long long)mi.pcnt);
  mi.pbin = (int)bin;
  mi.pcnt = 1;
}
// slot 0 = the register pair, slots 1 .. 3 = shared-memory columns
__device__ __forceinline__ void hist_add(const HistArgs& h, int* sh, Micro& mi, int slot, int64_t bin) {
  if (slot == 0) {
    hist_add_reg(h, sh, mi, bin);
    return;
  }
  if (h.sh_bins > 0) {
    atomicAdd(&sh[bin], 1);
    return;
  }
  if (bin > 0x7fffffffLL || threadIdx.x >= kPendThreads) {
    atomicAdd(&h.out[bin], 1ULL);
    return;
  }
  const unsigned cnt = g_pend_cnt[slot - 1][threadIdx.x];
  const int pending = g_pend_bin[slot - 1][threadIdx.x];
  if (cnt != 0 && pending == (int)bin) {
    g_pend_cnt[slot - 1][threadIdx.x] = cnt + 1;
    return;
  }
  if (cnt != 0) atomicAdd(&h.out[pending], (unsigned long long)cnt);
  g_pend_bin[slot - 1][threadIdx.x] = (int)bin;
  g_pend_cnt[slot - 1][threadIdx.x] = 1;
}
__device__ __forceinline__ void hist_pending_init() {
  if (threadIdx.x < kPendThreads) {
#pragma unroll
    for (int s = 0; s < 3; ++s) g_pend_cnt[s][threadIdx.x] = 0;
  }
}
__device__ __forceinline__ void hist_flush_pending(const HistArgs& h, const Micro& mi) {
  if (h.sh_bins > 0) return;
  if (mi.pcnt) atomicAdd(&h.out[mi.pbin], (unsigned long long)mi.pcnt);
  if (threadIdx.x >= kPendThreads) return;
#pragma unroll
  for (int s = 0; s < 3; ++s) {
    const unsigned cnt = g_pend_cnt[s][threadIdx.x];
    if (cnt) atomicAdd(&h.out[g_pend_bin[s][threadIdx.x]], (unsigned long long)cnt);
  }
}

// A warp that classified ONE sample together (every lane holds t and p): lane s feeds histogram s of the mode from
// its own register pair, so the histograms of a row are updated side by side.  Must be called by the whole warp.
__device__ __forceinline__ void count_sample_warp(const HistArgs& h, int* sh, Micro& mi, int64_t t, int64_t p, int lane) {
  if (h.has_ignore && t == h.ignore) return;
  if (t < 0 || t >= h.C) {
    if (lane == 0) atomicOr(h.err, TMB_ERR_TARGET_RANGE);
    return;
  }
  if (h.mode == TMB_MC_MICRO) {
    if (lane == 0) {
      mi.n += 1;
      mi.match += (p == t);
    }
    return;
  }
  int64_t bin = -1;
  if (h.mode == TMB_MC_CONFMAT) {
    if (lane == 0) bin = t * h.C + p;
  } else {
    const int64_t base = h.mode == TMB_MC_BOTH ? (int64_t)h.C * h.C : 0;
    const int s = h.mode == TMB_MC_BOTH ? lane - 1 : lane;  // BOTH: lane 0 takes the confusion matrix
    if (h.mode == TMB_MC_BOTH && lane == 0) bin = t * h.C + p;
    if (s == 0) bin = base + t;
    if (s == 1) bin = base + h.C + p;
    if (s == 2 && p == t) bin = base + 2 * (int64_t)h.C + t;
  }
  if (bin >= 0) hist_add_reg(h, sh, mi, bin);
}

// one classified sample: t = target, p = predicted class
__device__ __forceinline__ void count_sample(const HistArgs& h, int* sh, Micro& mi, int64_t t, int64_t p,
                                             bool p_is_label, int64_t sample = 0) {
  if (h.has_ignore && t == h.ignore) return;
  if (t < 0 || t >= h.C) {
    atomicOr(h.err, TMB_ERR_TARGET_RANGE);
    return;
  }
  if (p_is_label && (p < 0 || p >= h.C)) {
    atomicOr(h.err, TMB_ERR_PREDS_RANGE);
    return;
  }
  if (h.mode == TMB_MC_CONFMAT) {
    hist_add(h, sh, mi, 0, t * h.C + p);
  } else if (h.mode == TMB_MC_BOTH) {  // batch confusion matrix followed by the three stat-score histograms
    const int64_t cc = (int64_t)h.C * h.C;
    hist_add(h, sh, mi, 0, t * h.C + p);
    hist_add(h, sh, mi, 1, cc + t);
    hist_add(h, sh, mi, 2, cc + h.C + p);
    if (p == t) hist_add(h, sh, mi, 3, cc + 2 * (int64_t)h.C + t);
  } else if (h.mode == TMB_MC_STATS) {
    const int64_t base = sample * h.per_sample;  // 0 unless samplewise (then sh_bins == 0: global atomics)
    hist_add(h, sh, mi, 0, base + t);
    hist_add(h, sh, mi, 1, base + h.C + p);
    if (p == t) hist_add(h, sh, mi, 2, base + 2 * (int64_t)h.C + t);
  } else {
    mi.n += 1;
    mi.match += (p == t);
  }
}

__device__ __forceinline__ void hist_prologue(const HistArgs& h, int* sh) {
  hist_pending_init();
  if (h.sh_bins > 0) {
    for (int i = threadIdx.x; i < h.sh_bins; i += blockDim.x) sh[i] = 0;
    __syncthreads();
  }
}

__device__ __forceinline__ void hist_epilogue(const HistArgs& h, int* sh, Micro mi) {
  hist_flush_pending(h, mi);
  if (h.sh_bins > 0) {
    __syncthreads();
    for (int i = threadIdx.x; i < h.sh_bins; i += blockDim.x) {
      int v = sh[i];
      if (v) atomicAdd(&h.out[i], (unsigned long long)v);
    }
  }
  if (h.mode == TMB_MC_MICRO) {
    unsigned m = __reduce_add_sync(kFull, mi.match);
    unsigned n = __reduce_add_sync(kFull, mi.n);
    if ((threadIdx.x & 31) == 0 && n) {
      atomicAdd(&h.out[0], (unsigned long long)m);
      atomicAdd(&h.out[1], (unsigned long long)n);
    }
  }
}

// ---- vector helpers ------------------------------------------------------------------------------
template <int VB>
struct Vec;
template <>
struct Vec<16> {
  uint32_t w[4];
  __device__ __forceinline__ void load(const void* p) {
    uint4 r = ldg_stream16(p);
    w[0] = r.x; w[1] = r.y; w[2] = r.z; w[3] = r.w;
  }
};
template <>
struct Vec<8> {
  uint32_t w[2];
  __device__ __forceinline__ void load(const void* p) {
    uint2 r = ldg_stream8(p);
    w[0] = r.x; w[1] = r.y;
  }
};
template <>
struct Vec<4> {
  uint32_t w[1];
  __device__ __forceinline__ void load(const void* p) { w[0] = ldg_stream4(p); }
};
template <>
struct Vec<2> {
  uint32_t w[1];
  __device__ __forceinline__ void load(const void* p) { w[0] = ldg_stream2(p); }
};

// element e of a vector of VB bytes holding elements of dtype DT
template <int DT, int VB>
__device__ __forceinline__ float vec_elem(const Vec<VB>& v, int e) {
  if constexpr (Elem<DT>::kBytes == 4)
    return Elem<DT>::get(v.w[e], 0);
  else
    return Elem<DT>::get(v.w[e >> 1], e & 1);
}

// ---- rows_warp: one warp per row -----------------------------------------------------------------
// Lane L reads vectors L, L+32, ... of the row (U of them in flight), scans its elements in increasing
// index order, then the 32 partial results are merged with the (value, index) rule of torch.argmax.
template <int DT, int VB>
__global__ void __launch_bounds__(256) k1_rows_warp(const char* __restrict__ preds, const void* __restrict__ target,
                                                    int target_dtype, int64_t n_rows, HistArgs h) {
  extern __shared__ int sh[];
  constexpr int ES = Elem<DT>::kBytes;
  constexpr int EPV = VB / ES;
  constexpr int U = 4;
  hist_prologue(h, sh);
  Micro mi{0, 0, 0, 0};
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int64_t row_bytes = (int64_t)h.C * ES;
  const int nvec = (int)(row_bytes / VB);

  for (int64_t row = warp; row < n_rows; row += nwarps) {
    const char* rp = preds + row * row_bytes;
    Best b = best_init();
    for (int v0 = 0; v0 < nvec; v0 += 32 * U) {
      Vec<VB> r[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        int vi = v0 + u * 32 + lane;
        if (vi < nvec) r[u].load(rp + (int64_t)vi * VB);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        int vi = v0 + u * 32 + lane;
        if (vi < nvec) {
#pragma unroll
          for (int e = 0; e < EPV; ++e) best_step(b, vec_elem<DT, VB>(r[u], e), vi * EPV + e);
        }
      }
    }
    b = best_warp(b);
    count_sample_warp(h, sh, mi, load_label(target, target_dtype, row), best_index(b), lane);
  }
  hist_epilogue(h, sh, mi);
}

// ---- rows_pipe: warp per row, the whole row in registers, next row in flight -----------------------
// The config-2 shape (2000-byte bf16 rows) fits one register tile: U vectors of 16 B per lane.  While a
// row is reduced, the NEXT row's vectors and target are already in flight (software pipelining), so each
// warp keeps 2 x 2 KB outstanding and the target load is off the critical path.
//
// 16-bit dtypes use a two-phase packed argmax that costs ~2 instructions per element instead of ~4:
//   phase 1  row maximum with NaN-propagating packed max (HMNMX2.NAN) + warp shuffle reduce;
//   phase 2  first position that equals the maximum (or the first NaN when the row has one), found with
//            packed equality masks; lanes report their lowest matching index, the warp takes the minimum.
// "first index equal to the maximum, a NaN counts as the maximum" is exactly torch.argmax's rule.
template <int DT>
struct Packed;
template <>
struct Packed<TMB_BF16> {
  using V2 = __nv_bfloat162;
  static constexpr uint32_t kNegInf2 = 0xff80ff80u;
  __device__ static __forceinline__ V2 as_v2(uint32_t w) { return *reinterpret_cast<V2*>(&w); }
  __device__ static __forceinline__ uint32_t as_u32(V2 v) { return *reinterpret_cast<uint32_t*>(&v); }
  __device__ static __forceinline__ uint32_t max_nan(uint32_t a, uint32_t b) {
    return as_u32(__hmax2_nan(as_v2(a), as_v2(b)));
  }
  __device__ static __forceinline__ uint32_t eq_mask(uint32_t a, uint32_t b) { return __heq2_mask(as_v2(a), as_v2(b)); }
  __device__ static __forceinline__ uint32_t nan_mask(uint32_t a) { return __hneu2_mask(as_v2(a), as_v2(a)); }
  __device__ static __forceinline__ uint32_t gt_mask(uint32_t a, uint32_t b) { return __hgt2_mask(as_v2(a), as_v2(b)); }
  __device__ static __forceinline__ uint32_t splat(float m) {
    return as_u32(__float2bfloat162_rn(m));  // m is one of the row's values: the conversion is exact
  }
};
template <>
struct Packed<TMB_F16> {
  using V2 = __half2;
  static constexpr uint32_t kNegInf2 = 0xfc00fc00u;
  __device__ static __forceinline__ V2 as_v2(uint32_t w) { return *reinterpret_cast<V2*>(&w); }
  __device__ static __forceinline__ uint32_t as_u32(V2 v) { return *reinterpret_cast<uint32_t*>(&v); }
  __device__ static __forceinline__ uint32_t max_nan(uint32_t a, uint32_t b) {
    return as_u32(__hmax2_nan(as_v2(a), as_v2(b)));
  }
  __device__ static __forceinline__ uint32_t eq_mask(uint32_t a, uint32_t b) { return __heq2_mask(as_v2(a), as_v2(b)); }
  __device__ static __forceinline__ uint32_t nan_mask(uint32_t a) { return __hneu2_mask(as_v2(a), as_v2(a)); }
  __device__ static __forceinline__ uint32_t gt_mask(uint32_t a, uint32_t b) { return __hgt2_mask(as_v2(a), as_v2(b)); }
  __device__ static __forceinline__ uint32_t splat(float m) { return as_u32(__float2half2_rn(m)); }
};

// argmax of one row held as U 16-byte vectors per lane (vector u of lane L is vector u*32+L of the row)
template <int DT, int U>
__device__ __forceinline__ int row_argmax_packed16(const Vec<16> (&r)[U], int nvec, int lane) {
  using P = Packed<DT>;
  // phase 1: maximum
  uint32_t m2 = P::kNegInf2;
#pragma unroll
  for (int u = 0; u < U; ++u) {
    if (u * 32 + lane < nvec) {
#pragma unroll
      for (int j = 0; j < 4; ++j) m2 = P::max_nan(m2, r[u].w[j]);
    }
  }
  float lo = Elem<DT>::get(m2, 0), hi = Elem<DT>::get(m2, 1);
  const bool lane_nan = (lo != lo) || (hi != hi);
  float m = fmaxf(lo, hi);  // NaN-free when it matters: rows with a NaN take the other branch below
  const bool row_nan = __any_sync(kFull, lane_nan);
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) m = fmaxf(m, __shfl_xor_sync(kFull, m, d));
  // phase 2: first matching position
  const uint32_t M2 = P::splat(m);
  int mine = 0x7fffffff;
#pragma unroll
  for (int u = U - 1; u >= 0; --u) {
    const int vi = u * 32 + lane;
    if (vi < nvec) {
      uint32_t k[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) k[j] = row_nan ? P::nan_mask(r[u].w[j]) : P::eq_mask(r[u].w[j], M2);
      if (k[0] | k[1] | k[2] | k[3]) {
        int pos = 0;
#pragma unroll
        for (int j = 3; j >= 0; --j)
          if (k[j]) pos = 2 * j + ((k[j] & 0xffffu) ? 0 : 1);
        mine = vi * 8 + pos;  // u descends, so the lowest vector index wins
      }
    }
  }
  return __reduce_min_sync(kFull, mine);
}

template <int DT, int VB>
__global__ void __launch_bounds__(256) k1_rows_pipe(const char* __restrict__ preds, const void* __restrict__ target,
                                                    int target_dtype, int64_t n_rows, HistArgs h) {
  extern __shared__ int sh[];
  constexpr int ES = Elem<DT>::kBytes;
  constexpr int EPV = VB / ES;
  constexpr int U = 4;
  hist_prologue(h, sh);
  Micro mi{0, 0, 0, 0};
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int64_t row_bytes = (int64_t)h.C * ES;
  const int nvec = (int)(row_bytes / VB);  // <= 32 * U, checked by the launcher

  Vec<VB> cur[U], nxt[U];
  int64_t t_cur = 0, t_nxt = 0;
  auto issue = [&](int64_t row, Vec<VB>(&r)[U], int64_t& t) {
    const char* rp = preds + row * row_bytes;
#pragma unroll
    for (int u = 0; u < U; ++u) {
      int vi = u * 32 + lane;
      if (vi < nvec) r[u].load(rp + (int64_t)vi * VB);
    }
    t = load_label(target, target_dtype, row);  // same address in every lane: one broadcast transaction
  };

  int64_t row = warp;
  if (row < n_rows) issue(row, cur, t_cur);
  for (; row < n_rows; row += nwarps) {
    const int64_t next = row + nwarps;
    if (next < n_rows) issue(next, nxt, t_nxt);
    int p;
    if constexpr (ES == 2 && VB == 16) {
      p = row_argmax_packed16<DT, U>(cur, nvec, lane);
    } else {
      Best b = best_init();
#pragma unroll
      for (int u = 0; u < U; ++u) {
        int vi = u * 32 + lane;
        if (vi < nvec) {
#pragma unroll
          for (int e = 0; e < EPV; ++e) best_step(b, vec_elem<DT, VB>(cur[u], e), vi * EPV + e);
        }
      }
      p = best_index(best_warp(b));
    }
    count_sample_warp(h, sh, mi, t_cur, p, lane);
#pragma unroll
    for (int u = 0; u < U; ++u) cur[u] = nxt[u];
    t_cur = t_nxt;
  }
  hist_epilogue(h, sh, mi);
}

// ---- rows_thread: one thread per row, rows staged through shared memory ---------------------------
// Rows shorter than a few vectors cannot keep a warp busy, so a CTA copies TILE consecutive rows
// (a contiguous byte range) into shared memory with full-width coalesced loads and each thread then
// scans its own row from there.
constexpr int kTileRows = 128;

template <int DT, int VB>
__global__ void __launch_bounds__(kTileRows) k1_rows_thread(const char* __restrict__ preds,
                                                            const void* __restrict__ target, int target_dtype,
                                                            int64_t n_rows, HistArgs h, int pad) {
  extern __shared__ int sh_all[];
  constexpr int ES = Elem<DT>::kBytes;
  const int64_t row_bytes = (int64_t)h.C * ES;
  // `pad`: rows of a multiple of four words are stored one word apart (odd word stride), word by word — read at
  // their natural stride the 32 rows of a warp fell into 32 / gcd(words, 32) banks (C = 64 bf16: 3.9 ms against
  // 0.53 ms at C = 100)
  const int row_words = (int)(row_bytes / 4);
  const int stride_bytes = (int)row_bytes + (pad ? 4 : 0);
  const uint32_t word_magic = pad ? (uint32_t)(0x100000000ull / (unsigned)row_words) + 1u : 0u;  // g / row_words
  // layout: [tile bytes, rounded up to 16][histogram bins]
  const int tile_cap = (int)((kTileRows * (int64_t)stride_bytes + 15) / 16 * 16);
  char* tile = (char*)sh_all;
  int* sh = (int*)(tile + tile_cap);
  hist_prologue(h, sh);
  Micro mi{0, 0, 0, 0};
  const int64_t n_tiles = (n_rows + kTileRows - 1) / kTileRows;

  for (int64_t tIdx = blockIdx.x; tIdx < n_tiles; tIdx += gridDim.x) {
    const int64_t row0 = tIdx * kTileRows;
    const int rows = (int)min((int64_t)kTileRows, n_rows - row0);
    const int64_t bytes = rows * row_bytes;
    const char* src = preds + row0 * row_bytes;
    const int nv = (int)(bytes / VB);
    // the row's label is requested before the tile is staged: it arrives while the scores are copied and scanned
    int64_t t = 0;
    if ((int)threadIdx.x < rows) t = load_label(target, target_dtype, row0 + threadIdx.x);
    for (int i = threadIdx.x; i < nv; i += kTileRows) {
      Vec<VB> v;
      v.load(src + (int64_t)i * VB);
      if constexpr (VB >= 4) {
        if (pad) {
#pragma unroll
          for (int e = 0; e < VB / 4; ++e) {
            const int g = i * (VB / 4) + e;  // word of the tile
            const int r = (int)__umulhi((uint32_t)g, word_magic);
            ((uint32_t*)tile)[g + r] = v.w[e];
          }
          continue;
        }
      }
      if constexpr (VB == 16) {
        *(uint4*)(tile + i * 16) = make_uint4(v.w[0], v.w[1], v.w[2], v.w[3]);
      } else if constexpr (VB == 8) {
        *(uint2*)(tile + i * 8) = make_uint2(v.w[0], v.w[1]);
      } else if constexpr (VB == 4) {
        *(uint32_t*)(tile + i * 4) = v.w[0];
      } else {
        *(uint16_t*)(tile + i * 2) = (uint16_t)v.w[0];
      }
    }
    // remainder (< VB bytes), element granular
    for (int64_t off = (int64_t)nv * VB + threadIdx.x * ES; off < bytes; off += kTileRows * ES) {
      const int shift = pad ? 4 * (int)__umulhi((uint32_t)(off >> 2), word_magic) : 0;
      if constexpr (ES == 4)
        *(uint32_t*)(tile + off + shift) = ldg_stream4(src + off);
      else
        *(uint16_t*)(tile + off + shift) = ldg_stream2(src + off);
    }
    __syncthreads();
    if ((int)threadIdx.x < rows) {
      Best b = best_init();
      const char* rp = tile + threadIdx.x * stride_bytes;
      for (int c = 0; c < h.C; ++c) {
        float v;
        if constexpr (ES == 4)
          v = *(const float*)(rp + c * 4);
        else
          v = Elem<DT>::get(*(const uint16_t*)(rp + c * 2), 0);
        best_step(b, v, c);
      }
      count_sample(h, sh, mi, t, best_index(b), false);
    }
    __syncthreads();
  }
  hist_epilogue(h, sh, mi);
}

// ---- strided: [N, C, S], class stride S ------------------------------------------------------------
// One thread owns PPT = VB/ES adjacent pixels of one outer index and walks the C class planes with U
// loads in flight; a warp therefore reads 32*VB contiguous bytes of every plane.
template <int DT, int VB>
__global__ void __launch_bounds__(256, 3) k1_strided(const char* __restrict__ preds, const void* __restrict__ target,
                                                  int target_dtype, int64_t n_outer, int64_t S, HistArgs h) {
  extern __shared__ int sh[];
  constexpr int ES = Elem<DT>::kBytes;
  constexpr int PPT = VB / ES;
  constexpr int U = 8;
  hist_prologue(h, sh);
  Micro mi{0, 0, 0, 0};
  const int64_t groups_per_outer = S / PPT;
  const int64_t n_groups = n_outer * groups_per_outer;
  const int64_t plane_bytes = S * ES;

  for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < n_groups;
       g += (int64_t)gridDim.x * blockDim.x) {
    const int64_t n = g / groups_per_outer;
    const int64_t s0 = (g - n * groups_per_outer) * PPT;
    const char* p0 = preds + (n * h.C * S + s0) * ES;
    if constexpr (ES == 2 && VB == 16) {
      // 16-bit scores: keep the running maximum and its class for two pixels per register (the launcher only
      // picks this variant for C <= 65535).  A plane replaces the incumbent where it is strictly greater, or
      // where it is the first NaN — torch.argmax's rule — with 3 packed compares and 3 bit-selects per pair.
      using P = Packed<DT>;
      uint32_t best2[4], idx2[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        best2[j] = P::kNegInf2;
        idx2[j] = 0u;
      }
      for (int c0 = 0; c0 < h.C; c0 += U) {
        Vec<VB> r[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
          if (c0 + u < h.C) r[u].load(p0 + (int64_t)(c0 + u) * plane_bytes);
#pragma unroll
        for (int u = 0; u < U; ++u)
          if (c0 + u < h.C) {
            const uint32_t c2 = (uint32_t)(c0 + u) * 0x10001u;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint32_t v2 = r[u].w[j];
              const uint32_t take = P::gt_mask(v2, best2[j]) | (P::nan_mask(v2) & P::eq_mask(best2[j], best2[j]));
              idx2[j] = (take & c2) | (~take & idx2[j]);
              best2[j] = (take & v2) | (~take & best2[j]);
            }
          }
      }
#pragma unroll
      for (int q = 0; q < PPT; ++q) {
        int64_t t = load_label(target, target_dtype, n * S + s0 + q);
        count_sample(h, sh, mi, t, (idx2[q >> 1] >> (16 * (q & 1))) & 0xffffu, false, n);
      }
      continue;
    }
    Best b[PPT];
#pragma unroll
    for (int q = 0; q < PPT; ++q) b[q] = best_init();
    for (int c0 = 0; c0 < h.C; c0 += U) {
      Vec<VB> r[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (c0 + u < h.C) r[u].load(p0 + (int64_t)(c0 + u) * plane_bytes);
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (c0 + u < h.C) {
#pragma unroll
          for (int q = 0; q < PPT; ++q) best_step(b[q], vec_elem<DT, VB>(r[u], q), c0 + u);
        }
    }
#pragma unroll
    for (int q = 0; q < PPT; ++q) {
      int64_t t = load_label(target, target_dtype, n * S + s0 + q);
      count_sample(h, sh, mi, t, best_index(b[q]), false, n);
    }
  }
  hist_epilogue(h, sh, mi);
}

// ---- labels: integer preds -------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k1_labels(const void* __restrict__ preds, int preds_dtype,
                                                 const void* __restrict__ target, int target_dtype, int64_t n,
                                                 HistArgs h) {
  extern __shared__ int sh[];
  hist_prologue(h, sh);
  Micro mi{0, 0, 0, 0};
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t t = load_label(target, target_dtype, i);
    int64_t p = load_label(preds, preds_dtype, i);
    count_sample(h, sh, mi, t, p, true, h.per_sample > 0 ? i / h.sample_len : 0);
  }
  hist_epilogue(h, sh, mi);
}

// ---- stat-score finalize ---------------------------------------------------------------------------
// tp/fp/tn/fn += f(hist_target, hist_pred, match), scratch re-zeroed for the next update.
__global__ void __launch_bounds__(1024) k1_stats_finalize(unsigned long long* scratch, int C, int mode,
                                                          long long* tp, long long* fp, long long* tn,
                                                          long long* fn) {
  if (mode == TMB_MC_MICRO) {
    if (threadIdx.x == 0) {
      long long m = (long long)scratch[0], n = (long long)scratch[1];
      tp[0] += m;
      fp[0] += n - m;
      fn[0] += n - m;
      tn[0] += (long long)C * n - 2 * n + m;  // C*n - (tp + fp + fn)
      scratch[0] = 0;
      scratch[1] = 0;
    }
    return;
  }
  __shared__ long long red[32];
  __shared__ long long total;
  long long part = 0;
  for (int c = threadIdx.x; c < C; c += blockDim.x) part += (long long)scratch[c];
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) part += __shfl_xor_sync(kFull, part, d);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = part;
  __syncthreads();
  if (threadIdx.x == 0) {
    long long s = 0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += red[w];
    total = s;
  }
  __syncthreads();
  const long long n = total;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    long long ht = (long long)scratch[c], hp = (long long)scratch[C + c], m = (long long)scratch[2 * C + c];
    tp[c] += m;
    fp[c] += hp - m;
    fn[c] += ht - m;
    tn[c] += n - ht - hp + m;
    scratch[c] = 0;
    scratch[C + c] = 0;
    scratch[2 * C + c] = 0;
  }
}

// ---- host side --------------------------------------------------------------------------------------
constexpr int kShBinsMax = 4096;  // privatise in shared memory up to this many bins (16 KB)

// ---- shared update of a MetricCollection -------------------------------------------------------
// TMB_MC_BOTH fills one scratch [C*C | 3C] per batch; these two kernels hand it out: the batch confusion matrix is
// added into the confusion-matrix state, the histograms into as many stat-score states as subscribe (per class, or
// summed for `average="micro"`), and the last consumer clears.
__global__ void __launch_bounds__(256) k1_add_clear(unsigned long long* __restrict__ scratch, long long* __restrict__ dst,
                                                    int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const unsigned long long v = scratch[i];
    if (v) {
      dst[i] += (long long)v;
      scratch[i] = 0;
    }
  }
}

__global__ void __launch_bounds__(1024) k1_stats_share(unsigned long long* scratch, int C, int micro, int clear,
                                                       long long* tp, long long* fp, long long* tn, long long* fn) {
  __shared__ long long red[2][32];
  __shared__ long long total[2];
  long long pt = 0, pm = 0;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    pt += (long long)scratch[c];
    pm += (long long)scratch[2 * C + c];
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    pt += __shfl_xor_sync(kFull, pt, d);
    pm += __shfl_xor_sync(kFull, pm, d);
  }
  if ((threadIdx.x & 31) == 0) {
    red[0][threadIdx.x >> 5] = pt;
    red[1][threadIdx.x >> 5] = pm;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    long long a = 0, b = 0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) {
      a += red[0][w];
      b += red[1][w];
    }
    total[0] = a;
    total[1] = b;
  }
  __syncthreads();
  const long long n = total[0], m = total[1];
  if (micro) {
    if (threadIdx.x == 0) {  // same expressions as the TMB_MC_MICRO finalize
      tp[0] += m;
      fp[0] += n - m;
      fn[0] += n - m;
      tn[0] += (long long)C * n - 2 * n + m;
    }
  } else {
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      const long long ht = (long long)scratch[c], hp = (long long)scratch[C + c], mm = (long long)scratch[2 * C + c];
      tp[c] += mm;
      fp[c] += hp - mm;
      fn[c] += ht - mm;
      tn[c] += n - ht - hp + mm;
    }
  }
  if (clear) {
    __syncthreads();  // every thread has read what it needs (the totals were taken before the barrier above)
    for (int c = threadIdx.x; c < 3 * C; c += blockDim.x) scratch[c] = 0;
  }
}

// samplewise (multidim_average="samplewise", stat_scores.py:363-393 with sum_dim = [1]): one block per sample turns
// its [3C] histograms into rows of the [N, C] outputs and clears them
__global__ void __launch_bounds__(256) k1_stats_finalize_rows(unsigned long long* scratch, int C, long long* tp,
                                                              long long* fp, long long* tn, long long* fn) {
  __shared__ long long red[8];
  __shared__ long long total;
  unsigned long long* h = scratch + (int64_t)blockIdx.x * 3 * C;
  long long part = 0;
  for (int c = threadIdx.x; c < C; c += blockDim.x) part += (long long)h[c];
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) part += __shfl_xor_sync(kFull, part, d);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = part;
  __syncthreads();
  if (threadIdx.x == 0) {
    long long s = 0;
    for (int w = 0; w < 8; ++w) s += red[w];
    total = s;
  }
  __syncthreads();
  const long long n = total;
  const int64_t row = (int64_t)blockIdx.x * C;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const long long ht = (long long)h[c], hp = (long long)h[C + c], m = (long long)h[2 * C + c];
    tp[row + c] = m;
    fp[row + c] = hp - m;
    fn[row + c] = ht - m;
    tn[row + c] = n - ht - hp + m;
    h[c] = 0;
    h[C + c] = 0;
    h[2 * C + c] = 0;
  }
}

static int hist_bins(int mode, int64_t C) {
  if (mode == TMB_MC_CONFMAT) return (int)(C * C);
  if (mode == TMB_MC_STATS) return (int)(3 * C);
  return 0;
}

static int make_hist(HistArgs& h, int mode, int64_t C, int has_ignore, int64_t ignore, int64_t* out,
                     int32_t* err) {
  if (mode != TMB_MC_CONFMAT && mode != TMB_MC_STATS && mode != TMB_MC_MICRO && mode != TMB_MC_BOTH) return TMB_E_MODE;
  if (C < 1 || C > (1 << 24)) return TMB_E_SHAPE;
  if (!out || !err) return TMB_E_NULL;
  h.mode = mode;
  h.C = (int)C;
  int64_t bins = mode == TMB_MC_CONFMAT ? C * C : (mode == TMB_MC_STATS ? 3 * C : (mode == TMB_MC_BOTH ? C * C + 3 * C : 0));
  h.sh_bins = (bins > 0 && bins <= kShBinsMax) ? (int)bins : 0;
  h.has_ignore = has_ignore;
  h.ignore = ignore;
  h.out = (unsigned long long*)out;
  h.err = err;
  h.per_sample = 0;
  h.sample_len = 1;
  (void)hist_bins;
  return 0;
}

// widest power-of-two vector (<= 16 B) that divides both an address and a byte pitch
static int vec_bytes(const void* p, int64_t pitch_bytes, int elem_bytes) {
  uint64_t a = (uint64_t)(uintptr_t)p | (uint64_t)pitch_bytes;
  int v = 16;
  while (v > elem_bytes && (a & (uint64_t)(v - 1))) v >>= 1;
  return v;
}

// row_argmax_packed16 for a row held by LANES lanes (vector v of the row in lane v % LANES of the group): the same
// two phases, the reductions as butterflies over the xor offsets below LANES (they stay inside the group)
template <int DT, int U, int LANES>
__device__ __forceinline__ int row_argmax_packed16_group(const Vec<16> (&r)[U], int nvec, int j) {
  using P = Packed<DT>;
  uint32_t m2 = P::kNegInf2;
#pragma unroll
  for (int u = 0; u < U; ++u) {
    if (u * LANES + j < nvec) {
#pragma unroll
      for (int q = 0; q < 4; ++q) m2 = P::max_nan(m2, r[u].w[q]);
    }
  }
  float lo = Elem<DT>::get(m2, 0), hi = Elem<DT>::get(m2, 1);
  unsigned row_nan = (unsigned)((lo != lo) || (hi != hi));
  float m = fmaxf(lo, hi);  // NaN-free when it matters: rows with a NaN take the other branch below
#pragma unroll
  for (int d = LANES / 2; d > 0; d >>= 1) {
    m = fmaxf(m, __shfl_xor_sync(kFull, m, d));
    row_nan |= __shfl_xor_sync(kFull, row_nan, d);
  }
  const uint32_t M2 = P::splat(m);
  int mine = 0x7fffffff;
#pragma unroll
  for (int u = U - 1; u >= 0; --u) {
    const int vi = u * LANES + j;
    if (vi < nvec) {
      uint32_t k[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) k[q] = row_nan ? P::nan_mask(r[u].w[q]) : P::eq_mask(r[u].w[q], M2);
      if (k[0] | k[1] | k[2] | k[3]) {
        int pos = 0;
#pragma unroll
        for (int q = 3; q >= 0; --q)
          if (k[q]) pos = 2 * q + ((k[q] & 0xffffu) ? 0 : 1);
        mine = vi * 8 + pos;  // u descends, so the lowest vector index wins
      }
    }
  }
#pragma unroll
  for (int d = LANES / 2; d > 0; d >>= 1) mine = min(mine, __shfl_xor_sync(kFull, mine, d));
  return mine;
}

// ---- rows_sub: several rows per warp -----------------------------------------------------------------------
// Rows of 17 .. 64 vectors (e.g. C = 136 bf16, C = 100 fp32) left half of a warp idle in the kernels above and paid
// the per-row costs (label load, 5-step merge, atomics, loop) once per 272 .. 1024 bytes: 2 - 6 times the HBM time.
// Here 2^k lanes (the smallest power of two with 4 vectors per lane covering the row) hold a row, a warp takes
// 32 >> k rows per trip, and the merge runs over the k butterfly steps that stay inside a group.
template <int DT, int VB, int LANES>
__global__ void __launch_bounds__(256) k1_rows_sub(const char* __restrict__ preds, const void* __restrict__ target,
                                                   int target_dtype, int64_t n_rows, HistArgs h) {
  extern __shared__ int sh[];
  constexpr int ES = Elem<DT>::kBytes;
  constexpr int EPV = VB / ES;
  constexpr int U = 4;
  hist_prologue(h, sh);
  Micro mi{0, 0, 0, 0};
  const int lane = threadIdx.x & 31;
  constexpr int lanes = LANES, rows_per_warp = 32 / LANES;  // 8 or 16 lanes per row
  const int sub = lane / LANES, j = lane % LANES;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int64_t row_bytes = (int64_t)h.C * ES;
  const int nvec = (int)(row_bytes / VB);  // <= lanes * U, checked by the launcher

  Vec<VB> cur[U], nxt[U];
#pragma unroll
  for (int u = 0; u < U; ++u) cur[u] = Vec<VB>{}, nxt[u] = Vec<VB>{};
  int64_t t_cur = 0, t_nxt = 0;
  auto issue = [&](int64_t row, Vec<VB>(&r)[U], int64_t& t) {
    if (row < n_rows) {
      const char* rp = preds + row * row_bytes;
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int vi = u * lanes + j;
        if (vi < nvec) r[u].load(rp + (int64_t)vi * VB);
      }
      t = load_label(target, target_dtype, row);
    }
  };
  int64_t row0 = warp * rows_per_warp;  // warp-uniform
  if (row0 < n_rows) issue(row0 + sub, cur, t_cur);
  const int64_t trip_bytes = (int64_t)rows_per_warp * row_bytes;  // a warp's rows of one trip are contiguous
  const char* const preds_end = preds + n_rows * row_bytes;
  for (; row0 < n_rows; row0 += nwarps * rows_per_warp) {
    const int64_t next = row0 + nwarps * rows_per_warp;
    if (next < n_rows) issue(next + sub, nxt, t_nxt);
    // The register loads reach one trip ahead — 1 - 2 KB per warp, too little in flight to cover DRAM latency (fp32
    // C = 100 ran at 0.45 of the roofline with 60 % of the issue slots idle).  The trips after that are pulled into L2
    // with one prefetch per 128-byte line; the register loads then only wait for L2.
    // (fp32 rows: -8 %; the packed 16-bit argmax is issue bound and does not gain: not compiled in there)
#pragma unroll
    for (int ahead = 2; ahead <= (ES == 4 ? 3 : 1); ++ahead) {
      const char* first = preds + (row0 + ahead * nwarps * rows_per_warp) * row_bytes;
      const char* mine = (const char*)((uintptr_t)first & ~(uintptr_t)127) + lane * 128;
      if (mine < first + trip_bytes && mine < preds_end) prefetch_l2(mine);
    }
    int p;
    if constexpr (ES == 2 && VB == 16) {
      p = row_argmax_packed16_group<DT, U, LANES>(cur, nvec, j);
      if (p == 0x7fffffff) p = 0;  // rows past the end (never counted)
    } else {
      Best b = best_init();
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int vi = u * lanes + j;
        if (vi < nvec) {
#pragma unroll
          for (int e = 0; e < EPV; ++e) best_step(b, vec_elem<DT, VB>(cur[u], e), vi * EPV + e);
        }
      }
#pragma unroll
      for (int d = LANES / 2; d > 0; d >>= 1) {  // xor offsets below LANES stay inside the group
        Best o;
        o.v = __shfl_xor_sync(kFull, b.v, d);
        o.i = __shfl_xor_sync(kFull, b.i, d);
        b = best_merge(b, o);
      }
      p = best_index(b);
    }
    if (j == 0 && row0 + sub < n_rows) count_sample(h, sh, mi, t_cur, p, false);
#pragma unroll
    for (int u = 0; u < U; ++u) cur[u] = nxt[u];
    t_cur = t_nxt;
  }
  hist_epilogue(h, sh, mi);
}

template <int DT>
static int launch_rows(const void* preds, const void* target, int tdt, int64_t n_rows, const HistArgs& h,
                       cudaStream_t st) {
  constexpr int ES = Elem<DT>::kBytes;
  const int64_t row_bytes = (int64_t)h.C * ES;
  const int hist_smem = h.sh_bins * (int)sizeof(int);
  // (rows up to 256 bytes; at 272 - 1100 bytes the same kernel measured 0.99 - 4.3 ms against 0.57 - 1.19 ms for the
  // lanes-per-row kernel below on 2 GB of scores: the staged tile no longer leaves room for enough CTAs per SM)
  if (row_bytes <= 256) {
    // short rows: thread per row through shared memory
    const int vb = vec_bytes(preds, kTileRows * row_bytes, ES);
    const int pad = (row_bytes % 16 == 0 && vb >= 4) ? 1 : 0;  // a multiple of four words per row: see the kernel
    const int tile_cap = (int)((kTileRows * (row_bytes + (pad ? 4 : 0)) + 15) / 16 * 16);
    const int smem = tile_cap + hist_smem;
    const int64_t tiles = (n_rows + kTileRows - 1) / kTileRows;
    const char* p = (const char*)preds;
#define TMB_LAUNCH_THREAD(VB_)                                                                          \
  {                                                                                                     \
    auto k = k1_rows_thread<DT, VB_>;                                                                   \
    if (smem > 48 * 1024) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);   \
    k<<<resident_grid(k, kTileRows, smem, tiles), kTileRows, smem, st>>>(p, target, tdt, n_rows, h, pad); \
  }
    switch (vb) {
      case 16: TMB_LAUNCH_THREAD(16) break;
      case 8: TMB_LAUNCH_THREAD(8) break;
      case 4: TMB_LAUNCH_THREAD(4) break;
      default:
        if constexpr (ES == 2) {
          TMB_LAUNCH_THREAD(2)
          break;
        } else {
          return TMB_E_SHAPE;
        }
    }
#undef TMB_LAUNCH_THREAD
    return launch_status();
  }
  const int vb = vec_bytes(preds, row_bytes, ES);
  const int block = 256;
  const int64_t ctas = (n_rows + 7) / 8;
  const char* p = (const char*)preds;
  const bool one_tile = row_bytes / vb <= 128;  // the row fits U = 4 vectors per lane: pipelined kernel
  if (row_bytes / vb <= 64) {  // at most half a warp's worth of vectors: several rows per warp (k1_rows_sub)
    const int per_lane4 = (int)((row_bytes / vb + 3) / 4);
    const int lanes = per_lane4 <= 8 ? 8 : 16;  // rows of 17 .. 64 vectors: 5 .. 16 lanes' worth
    const int64_t sub_ctas = (n_rows + 8 * (32 / lanes) - 1) / (8 * (32 / lanes));
#define TMB_LAUNCH_SUB(VB_)                                                                                     \
  if (lanes == 8) {                                                                                             \
    auto k = k1_rows_sub<DT, VB_, 8>;                                                                           \
    k<<<resident_grid(k, block, hist_smem, sub_ctas), block, hist_smem, st>>>(p, target, tdt, n_rows, h);       \
  } else {                                                                                                      \
    auto k = k1_rows_sub<DT, VB_, 16>;                                                                          \
    k<<<resident_grid(k, block, hist_smem, sub_ctas), block, hist_smem, st>>>(p, target, tdt, n_rows, h);       \
  }
    switch (vb) {
      case 16: TMB_LAUNCH_SUB(16) break;
      case 8: TMB_LAUNCH_SUB(8) break;
      case 4: TMB_LAUNCH_SUB(4) break;
      default:
        if constexpr (ES == 2) {
          TMB_LAUNCH_SUB(2)
          break;
        } else {
          return TMB_E_SHAPE;
        }
    }
#undef TMB_LAUNCH_SUB
    return launch_status();
  }
#define TMB_LAUNCH_ROWS(VB_)                                                                              \
  if (one_tile) {                                                                                         \
    auto k = k1_rows_pipe<DT, VB_>;                                                                       \
    k<<<resident_grid(k, block, hist_smem, ctas), block, hist_smem, st>>>(p, target, tdt, n_rows, h);     \
  } else {                                                                                                \
    auto k = k1_rows_warp<DT, VB_>;                                                                       \
    k<<<resident_grid(k, block, hist_smem, ctas), block, hist_smem, st>>>(p, target, tdt, n_rows, h);     \
  }
  switch (vb) {
    case 16: TMB_LAUNCH_ROWS(16) break;
    case 8: TMB_LAUNCH_ROWS(8) break;
    case 4: TMB_LAUNCH_ROWS(4) break;
    default:
      if constexpr (ES == 2) {
        TMB_LAUNCH_ROWS(2)
        break;
      } else {
        return TMB_E_SHAPE;
      }
  }
#undef TMB_LAUNCH_ROWS
  return launch_status();
}

template <int DT>
static int launch_strided(const void* preds, const void* target, int tdt, int64_t n_outer, int64_t S,
                          const HistArgs& h, cudaStream_t st) {
  constexpr int ES = Elem<DT>::kBytes;
  const int hist_smem = h.sh_bins * (int)sizeof(int);
  int vb = vec_bytes(preds, S * ES, ES);
  if (ES == 2 && vb == 16 && h.C > 65535) vb = 8;  // the packed 16-byte variant keeps class indices in 16 bits
  const int64_t groups = n_outer * (S / (vb / ES));
  const int block = 256;
  const int64_t ctas = (groups + block - 1) / block;
  const char* p = (const char*)preds;
#define TMB_LAUNCH_STRIDED(VB_)                                                                            \
  {                                                                                                        \
    auto k = k1_strided<DT, VB_>;                                                                          \
    k<<<resident_grid(k, block, hist_smem, ctas), block, hist_smem, st>>>(p, target, tdt, n_outer, S, h);  \
  }
  switch (vb) {
    case 16: TMB_LAUNCH_STRIDED(16) break;
    case 8: TMB_LAUNCH_STRIDED(8) break;
    case 4: TMB_LAUNCH_STRIDED(4) break;
    default:
      if constexpr (ES == 2) {
        TMB_LAUNCH_STRIDED(2)
        break;
      } else {
        return TMB_E_SHAPE;
      }
  }
#undef TMB_LAUNCH_STRIDED
  return launch_status();
}

// ---- top-k rows: stat scores with top_k > 1 -------------------------------------------------------
// _multiclass_stat_scores_update with top_k != 1 (functional/classification/stat_scores.py:363-393): select_topk
// (utilities/data.py:115-139, torch.topk) turns a row into a k-hot prediction, the target into a one-hot, and
// per class  tp = target & pred,  fn = target & !pred,  fp = !target & pred,  tn = the rest; rows whose target is
// ignored drop out of all four.  That is three histograms per batch — T[c] rows with target c, P[c] rows with c
// among the top k, M[c] rows with target c among the top k — in the layout k1_stats_finalize already reduces
// (tn = n - T - P + M holds for any k).  A warp keeps the row in registers (element c in lane c % 32, slot c / 32)
// as the order-preserving unsigned keys of torch's radix select (NaN largest, -0.0 below +0.0) and extracts the maximum k times:
// lane-local best in slot order, REDUX.MAX over the warp, REDUX.MIN over the indices of the lanes that hold it —
// i.e. ties go to the lowest class index, the order torch's radix-select gather produces.
__device__ __forceinline__ uint32_t topk_key(float v) {
  if (v != v) return 0xffffffffu;
  const uint32_t b = __float_as_uint(v);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);  // >= 0x007fffff for every non-NaN: 0 marks "taken / padding"
}

template <int DT, int SLOTS>
__global__ void __launch_bounds__(256) k1_rows_topk(const void* __restrict__ preds, const void* __restrict__ target,
                                                    int tdt, int64_t rows, int C, int top_k, int has_ignore,
                                                    int64_t ignore, unsigned long long* __restrict__ out, int* err) {
  extern __shared__ int sh_topk[];  // [3 * C]: T | P | M
  for (int i = threadIdx.x; i < 3 * C; i += blockDim.x) sh_topk[i] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < rows; r += nwarps) {  // warp-uniform
    {  // this warp's next row -> L2 (one request per 128-byte line): its register loads then wait for L2, not DRAM
      const int64_t rn = r + nwarps;
      const int64_t first = rn * C * Elem<DT>::kBytes, end = first + (int64_t)C * Elem<DT>::kBytes;
      const int64_t mine = (first & ~(int64_t)127) + (int64_t)lane * 128;
      if (rn < rows && mine < end) prefetch_l2((const char*)preds + mine);
    }
    const int64_t t = load_label(target, tdt, r);
    if (has_ignore && t == ignore) continue;
    if (t < 0 || t >= C) {
      if (lane == 0) atomicOr(err, TMB_ERR_TARGET_RANGE);
      continue;
    }
    uint32_t key[SLOTS];
#pragma unroll
    for (int it = 0; it < SLOTS; ++it) {
      const int c = it * 32 + lane;
      key[it] = c < C ? topk_key(load_elem<DT>(preds, r * C + c)) : 0u;
    }
    bool hit = false;
    for (int round = 0; round < top_k; ++round) {
      uint32_t best = 0;
      int best_it = 0;
#pragma unroll
      for (int it = 0; it < SLOTS; ++it) {
        if (key[it] > best) {  // strictly greater: the lowest slot (= lowest class of this lane) wins a tie
          best = key[it];
          best_it = it;
        }
      }
      const uint32_t top = __reduce_max_sync(kFull, best);
      const uint32_t cand = (best == top) ? (uint32_t)(best_it * 32 + lane) : 0xffffffffu;
      const uint32_t cls = __reduce_min_sync(kFull, cand);
#pragma unroll
      for (int it = 0; it < SLOTS; ++it)
        if ((uint32_t)(it * 32 + lane) == cls) key[it] = 0u;  // taken
      if (lane == 0) atomicAdd(&sh_topk[C + (int)cls], 1);
      hit |= (int64_t)cls == t;
    }
    if (lane == 0) {
      atomicAdd(&sh_topk[(int)t], 1);
      if (hit) atomicAdd(&sh_topk[2 * C + (int)t], 1);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 3 * C; i += blockDim.x) {
    const int v = sh_topk[i];
    if (v) atomicAdd(&out[i], (unsigned long long)v);
  }
}

// 16-bit scores leave 16 spare bits in a 32-bit key: value key in the high half, 0xffff - class in the low half.
// One integer max per slot then finds value AND lowest class at once, one REDUX.MAX per round finishes the warp,
// and the row arrives as 16-byte vectors (8 scores each, vector v of the row in lane v % 32) instead of 32 two-byte
// loads per lane.  Needs C % 8 == 0 (16-byte aligned rows) and C <= 1024; other shapes take k1_rows_topk.
template <int DT>
__device__ __forceinline__ uint32_t topk_key16(uint32_t b) {  // b: the raw 16-bit pattern
  constexpr uint32_t kExpAll = DT == TMB_BF16 ? 0x7f80u : 0x7c00u;
  if ((b & 0x7fffu) > kExpAll) return 0xffffu;            // NaN ranks highest
  return (b & 0x8000u) ? (~b & 0xffffu) : (b | 0x8000u);  // torch's TopKTypeConfig<half / bfloat16>::convert
}

// ---- top-k through the lane maxima ------------------------------------------------------------------------------
// Ranking every score costs ~8 instructions each (key, class, per-vector maximum) plus ~60 per selection round: ~600
// instructions for a row of 1000.  Most of that is spent on scores that are nowhere near the top.  Instead:
//   1. a lane reduces its 32 scores to their maximum with packed max instructions on the raw 16-bit patterns
//      (two scores per operand, three operands per instruction);
//   2. k REDUX rounds over the 32 lane maxima give tau, the k-th largest of them.  The k largest scores of the row
//      are all >= tau (the k lane maxima found so far are k different scores >= tau), so
//   3. ONE packed compare pass marks, per lane, the scores >= tau: the candidates, m >= k of them;
//   4. m == k (the k largest sit in k different lanes and nothing ties with tau): the candidates are the answer;
//      otherwise the m candidates get their full keys (value key << 16 | 0xffff - class, as below), one per lane, and
//      k REDUX rounds pick the k largest — the same order as the full ranking, ties included.
// ~170 instructions for a row that ends in step 4's first case, ~280 otherwise.  The full ranking below remains for
// rows that hold a NaN (packed compares do not order them) and for m > 32 (rows of mostly equal scores).
template <int DT>
struct Pair16 {  // operations on two raw 16-bit scores in one register
  static constexpr uint32_t kNegInf2 = DT == TMB_BF16 ? 0xff80ff80u : 0xfc00fc00u;
  static __device__ __forceinline__ uint32_t max_nan(uint32_t a, uint32_t b) {  // NaN if either is NaN
    uint32_t d;
    if constexpr (DT == TMB_BF16) {
      asm("max.NaN.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
    } else {
      asm("max.NaN.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
    }
    return d;
  }
  static __device__ __forceinline__ uint32_t ge_mask(uint32_t a, uint32_t b) {  // 0xffff in each half where a >= b
    if constexpr (DT == TMB_BF16) {
      return __hge2_mask(*reinterpret_cast<const __nv_bfloat162*>(&a), *reinterpret_cast<const __nv_bfloat162*>(&b));
    } else {
      return __hge2_mask(*reinterpret_cast<const __half2*>(&a), *reinterpret_cast<const __half2*>(&b));
    }
  }
  static __device__ __forceinline__ bool is_nan(uint32_t bits16) {
    constexpr uint32_t kExpAll = DT == TMB_BF16 ? 0x7f80u : 0x7c00u;
    return (bits16 & 0x7fffu) > kExpAll;
  }
  static __device__ __forceinline__ uint32_t key_of(uint32_t bits16) {  // topk_key16 for a score that is no NaN
    return (bits16 & 0x8000u) ? (~bits16 & 0xffffu) : (bits16 | 0x8000u);
  }
  static __device__ __forceinline__ uint32_t bits_of(uint32_t key) {  // its inverse
    return (key & 0x8000u) ? (key & 0x7fffu) : (~key & 0xffffu);
  }
};

// One row held in the lanes' `raw` vectors (vector q of this lane is vector lane + 32 q of the row, 8 scores each).
// Returns false when the row needs the full ranking; otherwise the row's k classes have been added to the
// "predicted" histogram `pred_hist` and `hit` says whether `t` is among them.  `list` is this warp's 32-word buffer.
template <int DT, int VECS>
__device__ __forceinline__ bool topk16_by_lane_maxima(const uint4 (&raw)[VECS], const void* __restrict__ preds,
                                                      int64_t row_base, int lane, int n_vec, int top_k, int64_t t,
                                                      int* pred_hist, uint32_t* list, bool& hit) {
  using P = Pair16<DT>;
  uint32_t m2 = P::kNegInf2;
#pragma unroll
  for (int q = 0; q < VECS; ++q) {
    const uint32_t v2 = P::max_nan(P::max_nan(raw[q].x, raw[q].y), P::max_nan(raw[q].z, raw[q].w));
    m2 = P::max_nan(m2, lane + 32 * q < n_vec ? v2 : P::kNegInf2);
  }
  const uint32_t lo = m2 & 0xffffu, hi = m2 >> 16;
  const bool nan = P::is_nan(lo) | P::is_nan(hi);
  const uint32_t kb = max(P::key_of(lo), P::key_of(hi));  // this lane's best score as a key
  if (__any_sync(kFull, nan)) return false;
  // tau: the lane number makes the words unique; which lane of a tie goes first does not matter for tau
  uint32_t mine = (kb << 16) | (uint32_t)lane;
  uint32_t kth = 0;
  for (int round = 0; round < top_k; ++round) {
    const uint32_t top = __reduce_max_sync(kFull, mine);
    kth = top >> 16;
    if (top == mine) mine = 0u;
  }
  // the compare pass.  Word i of the lane (i = 4 q + h) reports its low score in bit i and its high score in bit
  // 16 + i: one AND-OR per word with an immediate.
  const uint32_t tau = P::bits_of(kth);
  const uint32_t tau2 = tau | (tau << 16);
  uint32_t marks = 0;
#pragma unroll
  for (int q = 0; q < VECS; ++q) {
    const uint32_t w[4] = {raw[q].x, raw[q].y, raw[q].z, raw[q].w};
    if (lane + 32 * q < n_vec) {
#pragma unroll
      for (int h = 0; h < 4; ++h) marks |= P::ge_mask(w[h], tau2) & (0x00010001u << (q * 4 + h));
    }
  }
  const int count = __popc(marks);
  const int m = __reduce_add_sync(kFull, count);
  if (m > 32) return false;
  auto class_of = [&](int bit) {  // bit i or 16 + i of `marks` -> class
    const int word = bit & 15;
    return (lane + 32 * (word >> 2)) * 8 + 2 * (word & 3) + (bit >> 4);
  };
  hit = false;
  if (m == top_k) {
    bool mine_hit = false;
    for (uint32_t left = marks; left; left &= left - 1u) {
      const int cls = class_of(__ffs(left) - 1);
      atomicAdd(&pred_hist[cls], 1);
      mine_hit |= (int64_t)cls == t;
    }
    hit = __any_sync(kFull, mine_hit);
    return true;
  }
  // m > k: every candidate gets its full key and a lane of its own
  int at = count;  // exclusive prefix sum of the counts over the lanes
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int up = __shfl_up_sync(kFull, at, d);
    if (lane >= d) at += up;
  }
  at -= count;
  for (uint32_t left = marks; left; left &= left - 1u) {
    const int cls = class_of(__ffs(left) - 1);
    // a lane's only candidate is its maximum; a lane with several reads each score again (the row is in L2)
    const uint32_t key = count == 1 ? kb : P::key_of((uint32_t)__ldg((const unsigned short*)preds + row_base + cls));
    list[at++] = (key << 16) | (0xffffu - (uint32_t)cls);
  }
  __syncwarp();
  uint32_t cand = lane < m ? list[lane] : 0u;
  __syncwarp();
  for (int round = 0; round < top_k; ++round) {
    const uint32_t top = __reduce_max_sync(kFull, cand);
    const int cls = (int)(0xffffu - (top & 0xffffu));
    if (cand == top) {
      atomicAdd(&pred_hist[cls], 1);
      cand = 0u;
    }
    hit |= (int64_t)cls == t;
  }
  return true;
}

template <int DT, int VECS>
__global__ void __launch_bounds__(256, 4) k1_rows_topk16(const void* __restrict__ preds, const void* __restrict__ target,
                                                      int tdt, int64_t rows, int C, int top_k, int has_ignore,
                                                      int64_t ignore, unsigned long long* __restrict__ out, int* err) {
  extern __shared__ int sh_topk[];  // [3 * C]: T | P | M
  __shared__ uint32_t cand_list[8][32];  // per warp: the candidates of topk16_by_lane_maxima
  for (int i = threadIdx.x; i < 3 * C; i += blockDim.x) sh_topk[i] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int n_vec = C >> 3;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  // vector q of a lane sits 512 q bytes behind its first one, and rows follow each other at a fixed distance: one
  // pointer, advanced by a constant, and immediate offsets.  Vectors past the row end are never loaded and stay zero.
  uint4 raw[VECS];
#pragma unroll
  for (int q = 0; q < VECS; ++q) raw[q] = make_uint4(0, 0, 0, 0);
  int64_t t_ahead = 0;  // the row's label travels with the row: asked for a trip ahead, not when it is needed
  const char* ahead = (const char*)preds + (warp * C + (int64_t)lane * 8) * 2;  // the row `fetch` loads next
  const int64_t row_step = nwarps * C * 2;
  auto fetch = [&](int64_t r) {
    t_ahead = load_label(target, tdt, r);
#pragma unroll
    for (int q = 0; q < VECS; ++q)
      if (lane + 32 * q < n_vec) raw[q] = ldg_stream16(ahead + q * 512);
    ahead += row_step;
  };
  int64_t r = warp;
  if (r < rows) fetch(r);
  for (; r < rows; r += nwarps) {  // warp-uniform
    const int64_t t = t_ahead;
    if ((has_ignore && t == ignore) || t < 0 || t >= C) {  // nothing of this row is counted
      if (!(has_ignore && t == ignore) && lane == 0) atomicOr(err, TMB_ERR_TARGET_RANGE);
      if (r + nwarps < rows) fetch(r + nwarps);
      continue;
    }
    if (top_k <= 32) {
      bool hit;
      if (topk16_by_lane_maxima<DT, VECS>(raw, preds, r * C, lane, n_vec, top_k, t, sh_topk + C,
                                          cand_list[threadIdx.x >> 5], hit)) {
        if (r + nwarps < rows) fetch(r + nwarps);
        if (lane == 0) {
          atomicAdd(&sh_topk[(int)t], 1);
          if (hit) atomicAdd(&sh_topk[2 * C + (int)t], 1);
        }
        continue;
      }
    }
    // Keys, four instructions each: the score moves to the high half (it is there already for the odd elements),
    // torch's order-preserving map is "flip everything if negative, else the sign" on the whole word, and one LOP3
    // puts 0xffff - class into the low half.  NaN (-> 0xffff) needs a compare per element; instead one test per
    // 32-bit word looks for an all-ones exponent (inf or NaN) in either half, and a lane that saw one redoes its
    // keys with the exact rule.
    constexpr uint32_t kExp2 = DT == TMB_BF16 ? 0x7f807f80u : 0x7c007c00u;  // exponent field of both halves
    constexpr uint32_t kCarry2 = DT == TMB_BF16 ? 0x00800080u : 0x04000400u;  // + lowest exponent bit: carry into bit 15
    uint32_t key[VECS * 8], group[VECS], special = 0;
#pragma unroll
    for (int q = 0; q < VECS; ++q) {
      const int v = lane + 32 * q;
      const uint32_t w[4] = {raw[q].x, raw[q].y, raw[q].z, raw[q].w};
#pragma unroll
      for (int h = 0; h < 4; ++h) special |= ((w[h] & kExp2) + kCarry2) & 0x80008000u;
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        uint32_t k = (e & 1) ? (w[e >> 1] & 0xffff0000u) : (w[e >> 1] << 16);
        k ^= (uint32_t)((int32_t)k >> 31) | 0x80000000u;
        key[q * 8 + e] = (k & 0xffff0000u) | (0xffffu - ((uint32_t)v * 8u + e));
      }
    }
    if (special) {
#pragma unroll
      for (int q = 0; q < VECS; ++q) {
        const uint32_t w[4] = {raw[q].x, raw[q].y, raw[q].z, raw[q].w};
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const uint32_t bits = (e & 1) ? (w[e >> 1] >> 16) : (w[e >> 1] & 0xffffu);
          key[q * 8 + e] = (topk_key16<DT>(bits) << 16) | (key[q * 8 + e] & 0xffffu);
        }
      }
    }
    // A lane keeps the maximum of each of its vectors (8 keys) and the maximum of those.  Keys are unique, so a round
    // is one REDUX.MAX, and only the lane that owned the winner has anything to update: it clears the key in the one
    // vector whose maximum it was and redoes that maximum (8 of its 32 keys; clearing and re-scanning all of them in
    // every lane was 80 instructions a round, two fifths of the kernel).  Vectors past the row end never win.
    auto group_max = [&](int q) {
      uint32_t m = 0;
#pragma unroll
      for (int e = 0; e < 8; ++e) m = max(m, key[q * 8 + e]);
      return m;
    };
    uint32_t best = 0;
#pragma unroll
    for (int q = 0; q < VECS; ++q) {
      group[q] = lane + 32 * q < n_vec ? group_max(q) : 0u;
      best = max(best, group[q]);
    }
    bool hit = false;
    for (int round = 0; round < top_k; ++round) {
      const uint32_t top = __reduce_max_sync(kFull, best);
      const int cls = (int)(0xffffu - (top & 0xffffu));
      if (best == top && round + 1 < top_k) {  // the owner
        best = 0;
#pragma unroll
        for (int q = 0; q < VECS; ++q) {
          if (group[q] == top) {
#pragma unroll
            for (int e = 0; e < 8; ++e) key[q * 8 + e] = key[q * 8 + e] == top ? 0u : key[q * 8 + e];
            group[q] = group_max(q);
          }
          best = max(best, group[q]);
        }
      }
      if (lane == 0) atomicAdd(&sh_topk[C + cls], 1);
      hit |= (int64_t)cls == t;
    }
    if (lane == 0) {
      atomicAdd(&sh_topk[(int)t], 1);
      if (hit) atomicAdd(&sh_topk[2 * C + (int)t], 1);
    }
    // the full ranking is the rare route (see topk16_by_lane_maxima): its keys and the next row's vectors do not have
    // to be live together, which keeps the kernel at four CTAs per SM
    if (r + nwarps < rows) fetch(r + nwarps);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 3 * C; i += blockDim.x) {
    const int v = sh_topk[i];
    if (v) atomicAdd(&out[i], (unsigned long long)v);
  }
}

// fp32 scores: no spare key bits, so the (value key, class) pair stays split as in k1_rows_topk, but the row arrives
// as 16-byte vectors (vector v of the row in lane v % 32, 4 scores each) with the next row already in flight.
// Needs C % 4 == 0 (16-byte aligned rows) and C <= 1024.
// topk16_by_lane_maxima for fp32 scores: the lane maximum is a chain of three-operand float max instructions
// (NaN-propagating), the 32-bit value key has no room for the class, so equal keys are told apart by a second REDUX
// (lowest class first), and a candidate travels as two words.
__device__ __forceinline__ float fmax3_nan(float a, float b, float c) {
  float d;
  asm("max.NaN.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}
__device__ __forceinline__ uint32_t key32_of(float v) {  // topk_key for a score that is no NaN
  const uint32_t b = __float_as_uint(v);
  return b ^ ((uint32_t)((int32_t)b >> 31) | 0x80000000u);
}

template <int VECS>
__device__ __forceinline__ bool topk32_by_lane_maxima(const uint4 (&raw)[VECS], const void* __restrict__ preds,
                                                      int64_t row_base, int lane, int n_vec, int top_k, int64_t t,
                                                      int* pred_hist, uint32_t* list_key, uint32_t* list_cls,
                                                      bool& hit) {
  float best = -INFINITY;
#pragma unroll
  for (int q = 0; q < VECS; ++q) {
    const float x = __uint_as_float(raw[q].x), y = __uint_as_float(raw[q].y);
    const float z = __uint_as_float(raw[q].z), w = __uint_as_float(raw[q].w);
    const float v = fmax3_nan(fmax3_nan(x, y, z), w, -INFINITY);
    best = fmax3_nan(best, lane + 32 * q < n_vec ? v : -INFINITY, -INFINITY);
  }
  if (__any_sync(kFull, best != best)) return false;
  const uint32_t kb = key32_of(best);
  // tau = the k-th largest lane maximum; equal maxima leave one lane per round (the lowest)
  uint32_t mine = kb, kth = 0;
  for (int round = 0; round < top_k; ++round) {
    kth = __reduce_max_sync(kFull, mine);
    const unsigned holders = __ballot_sync(kFull, mine == kth);
    if (lane == __ffs(holders) - 1) mine = 0u;  // 0 is below the key of every score, -inf included
  }
  const uint32_t tb = (kth & 0x80000000u) ? (kth & 0x7fffffffu) : ~kth;
  const float tau = __uint_as_float(tb);
  uint32_t marks = 0;
#pragma unroll
  for (int q = 0; q < VECS; ++q) {
    const float x[4] = {__uint_as_float(raw[q].x), __uint_as_float(raw[q].y), __uint_as_float(raw[q].z),
                        __uint_as_float(raw[q].w)};
    if (lane + 32 * q < n_vec) {
#pragma unroll
      for (int e = 0; e < 4; ++e) marks |= (x[e] >= tau ? 1u : 0u) << (q * 4 + e);
    }
  }
  const int count = __popc(marks);
  const int m = __reduce_add_sync(kFull, count);
  if (m > 32) return false;
  auto class_of = [&](int bit) { return (lane + 32 * (bit >> 2)) * 4 + (bit & 3); };
  hit = false;
  if (m == top_k) {
    bool mine_hit = false;
    for (uint32_t left = marks; left; left &= left - 1u) {
      const int cls = class_of(__ffs(left) - 1);
      atomicAdd(&pred_hist[cls], 1);
      mine_hit |= (int64_t)cls == t;
    }
    hit = __any_sync(kFull, mine_hit);
    return true;
  }
  int at = count;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int up = __shfl_up_sync(kFull, at, d);
    if (lane >= d) at += up;
  }
  at -= count;
  for (uint32_t left = marks; left; left &= left - 1u) {
    const int cls = class_of(__ffs(left) - 1);
    list_key[at] = count == 1 ? kb : key32_of(__ldg((const float*)preds + row_base + cls));
    list_cls[at] = (uint32_t)cls;
    ++at;
  }
  __syncwarp();
  uint32_t ck = lane < m ? list_key[lane] : 0u;
  const uint32_t cc = lane < m ? list_cls[lane] : 0xffffffffu;
  __syncwarp();
  for (int round = 0; round < top_k; ++round) {
    const uint32_t top = __reduce_max_sync(kFull, ck);
    const uint32_t cls = __reduce_min_sync(kFull, ck == top ? cc : 0xffffffffu);
    if (ck == top && cc == cls) {
      atomicAdd(&pred_hist[cls], 1);
      ck = 0u;
    }
    hit |= (int64_t)cls == t;
  }
  return true;
}

template <int VECS>
__global__ void __launch_bounds__(256, 3) k1_rows_topk32(const void* __restrict__ preds, const void* __restrict__ target,
                                                      int tdt, int64_t rows, int C, int top_k, int has_ignore,
                                                      int64_t ignore, unsigned long long* __restrict__ out, int* err) {
  extern __shared__ int sh_topk[];  // [3 * C]: T | P | M
  __shared__ uint32_t cand_key[8][32], cand_cls[8][32];  // per warp: the candidates of topk32_by_lane_maxima
  for (int i = threadIdx.x; i < 3 * C; i += blockDim.x) sh_topk[i] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int n_vec = C >> 2;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  // vector q of a lane sits 512 q bytes behind its first one, and rows follow each other at a fixed distance: one
  // pointer, advanced by a constant, and immediate offsets.  Vectors past the row end are never loaded and stay zero.
  uint4 raw[VECS];
#pragma unroll
  for (int q = 0; q < VECS; ++q) raw[q] = make_uint4(0, 0, 0, 0);
  int64_t t_ahead = 0;  // the row's label travels with the row: asked for a trip ahead, not when it is needed
  const char* ahead = (const char*)preds + (warp * C + (int64_t)lane * 4) * 4;  // the row `fetch` loads next
  const int64_t row_step = nwarps * C * 4;
  auto fetch = [&](int64_t r) {
    t_ahead = load_label(target, tdt, r);
#pragma unroll
    for (int q = 0; q < VECS; ++q)
      if (lane + 32 * q < n_vec) raw[q] = ldg_stream16(ahead + q * 512);
    ahead += row_step;
  };
  int64_t r = warp;
  if (r < rows) fetch(r);
  for (; r < rows; r += nwarps) {  // warp-uniform
    const int64_t t = t_ahead;
    if ((has_ignore && t == ignore) || t < 0 || t >= C) {  // nothing of this row is counted
      if (!(has_ignore && t == ignore) && lane == 0) atomicOr(err, TMB_ERR_TARGET_RANGE);
      if (r + nwarps < rows) fetch(r + nwarps);
      continue;
    }
    if (top_k <= 32) {
      bool hit;
      if (topk32_by_lane_maxima<VECS>(raw, preds, r * C, lane, n_vec, top_k, t, sh_topk + C,
                                      cand_key[threadIdx.x >> 5], cand_cls[threadIdx.x >> 5], hit)) {
        if (r + nwarps < rows) fetch(r + nwarps);
        if (lane == 0) {
          atomicAdd(&sh_topk[(int)t], 1);
          if (hit) atomicAdd(&sh_topk[2 * C + (int)t], 1);
        }
        continue;
      }
    }
    // keys: "flip everything if negative, else the sign" is two instructions; NaN (-> all ones) is looked for once per
    // lane (an all-ones exponent, i.e. inf or NaN, in any of its scores) and fixed up in the rare lane that has one
    uint32_t key[VECS * 4], group[VECS];
    bool special = false;
#pragma unroll
    for (int q = 0; q < VECS; ++q) {
      const uint32_t w[4] = {raw[q].x, raw[q].y, raw[q].z, raw[q].w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        special |= (w[e] & 0x7f800000u) == 0x7f800000u;
        key[q * 4 + e] = w[e] ^ ((uint32_t)((int32_t)w[e] >> 31) | 0x80000000u);
      }
    }
    if (special) {
#pragma unroll
      for (int q = 0; q < VECS; ++q) {
        const uint32_t w[4] = {raw[q].x, raw[q].y, raw[q].z, raw[q].w};
#pragma unroll
        for (int e = 0; e < 4; ++e) key[q * 4 + e] = topk_key(__uint_as_float(w[e]));
      }
    }
    // Per-vector maxima as in k1_rows_topk16.  Equal scores are possible here: the lanes holding the warp maximum
    // name the lowest class they hold it for (vectors and elements ascend with the class inside a lane), REDUX.MIN
    // picks the lowest of those, and the lane that named it clears that one key and redoes that one vector.
    auto group_max = [&](int q) { return max(max(key[q * 4], key[q * 4 + 1]), max(key[q * 4 + 2], key[q * 4 + 3])); };
    uint32_t best = 0;
#pragma unroll
    for (int q = 0; q < VECS; ++q) {
      group[q] = lane + 32 * q < n_vec ? group_max(q) : 0u;  // 0: padding, below every real key
      best = max(best, group[q]);
    }
    bool hit = false;
    for (int round = 0; round < top_k; ++round) {
      const uint32_t top = __reduce_max_sync(kFull, best);
      uint32_t mine = 0xffffffffu;
      if (best == top) {
#pragma unroll
        for (int q = VECS - 1; q >= 0; --q) {
          if (group[q] == top) {
            const int e = key[q * 4] == top ? 0 : (key[q * 4 + 1] == top ? 1 : (key[q * 4 + 2] == top ? 2 : 3));
            mine = (uint32_t)((lane + 32 * q) * 4 + e);
          }
        }
      }
      const uint32_t cls = __reduce_min_sync(kFull, mine);
      if (mine == cls && round + 1 < top_k) {  // the owner
        best = 0;
#pragma unroll
        for (int q = 0; q < VECS; ++q) {
          if ((uint32_t)(lane + 32 * q) == (cls >> 2)) {
#pragma unroll
            for (int e = 0; e < 4; ++e) key[q * 4 + e] = (uint32_t)e == (cls & 3u) ? 0u : key[q * 4 + e];
            group[q] = group_max(q);
          }
          best = max(best, group[q]);
        }
      }
      if (lane == 0) atomicAdd(&sh_topk[C + (int)cls], 1);
      hit |= (int64_t)cls == t;
    }
    if (lane == 0) {
      atomicAdd(&sh_topk[(int)t], 1);
      if (hit) atomicAdd(&sh_topk[2 * C + (int)t], 1);
    }
    // the full ranking is the rare route (see topk16_by_lane_maxima): its keys and the next row's vectors do not have
    // to be live together, which keeps the kernel at four CTAs per SM
    if (r + nwarps < rows) fetch(r + nwarps);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 3 * C; i += blockDim.x) {
    const int v = sh_topk[i];
    if (v) atomicAdd(&out[i], (unsigned long long)v);
  }
}

// Up to 128 classes: one THREAD per row.  A warp per row pays ~300 instructions of fixed cost per row (k REDUX rounds,
// atomics, loop) however short the row is — at C = 10 that was 32 ms for 2^30 scores.  Here a CTA stages a tile of rows
// into shared memory as keys (coalesced loads, odd row stride: conflict-free reads), and every thread runs the k
// selection passes over its own row: pass j takes the largest key below the one pass j - 1 took, so nothing is marked.
// 16-bit scores: unique keys `value key << 16 | 0xfffe - class`; fp32 scores: the value key alone, equal keys are
// taken in class order (`key == previous && class > previous class` stays eligible).
constexpr int kTopkGroup = 8;  // keys per group of the two-level selection in k1_rows_topk_thread

template <int DT>
__global__ void __launch_bounds__(256) k1_rows_topk_thread(const void* __restrict__ preds,
                                                           const void* __restrict__ target, int tdt, int64_t rows,
                                                           int C, int top_k, int has_ignore, int64_t ignore,
                                                           unsigned long long* __restrict__ out, int* err) {
  // [3 * C] T | P | M, the key tile [blockDim.x][stride] (rows padded to whole groups of kTopkGroup keys; the padding
  // is zero, below every key), the group maxima [groups][blockDim.x] (fp32: and their classes)
  extern __shared__ int sh_topk[];
  uint32_t* keys = (uint32_t*)(sh_topk + (3 * C + 31) / 32 * 32);
  const int groups = (C + kTopkGroup - 1) / kTopkGroup;
  const int stride = (groups * kTopkGroup) | 1, tile_rows = blockDim.x;
  uint32_t* gmax = keys + tile_rows * stride;
  uint32_t* gcls = gmax + groups * tile_rows;  // fp32 only
  for (int i = threadIdx.x; i < tile_rows * stride; i += blockDim.x) keys[i] = 0u;
  const uint32_t magic = C > 1 ? (uint32_t)(0x100000000ull / (unsigned)C) + 1u : 0u;  // i / C for i * C < 2^32
  const bool vec_ok = (((uintptr_t)preds) & 15) == 0;  // tiles are 128 or 256 rows: every tile then starts 16-byte aligned
  for (int i = threadIdx.x; i < 3 * C; i += blockDim.x) sh_topk[i] = 0;
  for (int64_t tile0 = (int64_t)blockIdx.x * tile_rows; tile0 < rows; tile0 += (int64_t)gridDim.x * tile_rows) {
    const int n_rows = (int)min((int64_t)tile_rows, rows - tile0);
    const int total = n_rows * C;
    const char* src = (const char*)preds + tile0 * C * Elem<DT>::kBytes;
    __syncthreads();  // the previous tile has been ranked (and, first time round, the counters are zero)
    // staging: 16-byte vectors, four per thread in flight (with one 2-byte load per trip the CTA waited a DRAM
    // latency per element: 7.6 ms at C = 100); what is left of a ragged last tile, or everything when the tensor is
    // not 16-byte aligned, goes element by element
    constexpr int kPerVec = 16 / Elem<DT>::kBytes;
    const int n_vec = vec_ok ? total / kPerVec : 0;
    auto put = [&](int i, uint32_t bits) {  // element i of the tile; bits: fp32 pattern or the 16-bit pattern
      const int r = C > 1 ? (int)__umulhi((uint32_t)i, magic) : i;
      const int c = i - r * C;
      uint32_t key;
      if constexpr (DT == TMB_F32) {
        key = topk_key(__uint_as_float(bits));
      } else {
        key = (topk_key16<DT>(bits) << 16) | (0xfffeu - (uint32_t)c);
      }
      keys[r * stride + c] = key;
    };
    for (int v0 = threadIdx.x; v0 < n_vec; v0 += 4 * blockDim.x) {
      uint4 w[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int v = v0 + u * blockDim.x;
        w[u] = v < n_vec ? ldg_stream16(src + (int64_t)v * 16) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int v = v0 + u * blockDim.x;
        if (v < n_vec) {
          const uint32_t x[4] = {w[u].x, w[u].y, w[u].z, w[u].w};
#pragma unroll
          for (int e = 0; e < kPerVec; ++e) {
            if constexpr (DT == TMB_F32) {
              put(v * kPerVec + e, x[e]);
            } else {
              put(v * kPerVec + e, (e & 1) ? (x[e >> 1] >> 16) : (x[e >> 1] & 0xffffu));
            }
          }
        }
      }
    }
    for (int i = n_vec * kPerVec + threadIdx.x; i < total; i += blockDim.x) {
      if constexpr (DT == TMB_F32) {
        put(i, ldg_stream4(src + (int64_t)i * 4));
      } else {
        put(i, ldg_stream2(src + (int64_t)i * 2));
      }
    }
    __syncthreads();
    if ((int)threadIdx.x >= n_rows) continue;
    const int64_t t = load_label(target, tdt, tile0 + threadIdx.x);
    if (has_ignore && t == ignore) continue;
    if (t < 0 || t >= C) {
      atomicOr(err, TMB_ERR_TARGET_RANGE);
      continue;
    }
    // Two levels: the maximum of every group of kTopkGroup keys is found once (one walk of the row) and kept in this
    // thread's column of `gmax`; a selection round then looks at the group maxima, takes the largest, and walks only
    // that one group again for its largest key below the one just taken.  k walks of the whole row had cost
    // 4 C k instructions a row (2000 at C = 100, k = 5); this is about 2 C + k (2 C / 8 + 30).
    const uint32_t* row = keys + threadIdx.x * stride;
    uint32_t* mine = gmax + threadIdx.x;
    uint32_t* mine_cls = gcls + threadIdx.x;
    for (int g = 0; g < groups; ++g) {
      uint32_t m = 0;
      int mc = 0;
#pragma unroll
      for (int e = 0; e < kTopkGroup; ++e) {
        const uint32_t k = row[g * kTopkGroup + e];
        if constexpr (DT == TMB_F32) {
          if (k > m) {  // strictly greater: the lowest class of equal scores; padding (0) never wins
            m = k;
            mc = g * kTopkGroup + e;
          }
        } else {
          m = max(m, k);
        }
      }
      mine[g * tile_rows] = m;
      if constexpr (DT == TMB_F32) mine_cls[g * tile_rows] = (uint32_t)mc;
    }
    bool hit = false;
    for (int round = 0; round < top_k; ++round) {
      uint32_t best = 0;
      int bg = 0;
      for (int g = 0; g < groups; ++g) {
        const uint32_t v = mine[g * tile_rows];
        if constexpr (DT == TMB_F32) {
          if (v > best) {  // groups ascend with the class: strictly greater keeps the lowest class of equal scores
            best = v;
            bg = g;
          }
        } else {
          best = max(best, v);
        }
      }
      int cls;
      if constexpr (DT == TMB_F32) {
        cls = (int)mine_cls[bg * tile_rows];
      } else {
        cls = (int)(0xfffeu - (best & 0xffffu));  // 16-bit keys are unique and carry their class
        bg = cls / kTopkGroup;
      }
      atomicAdd(&sh_topk[C + cls], 1);
      hit |= (int64_t)cls == t;
      if (round + 1 == top_k) break;
      // the group's next maximum: the largest (key, lowest class) after the pair just taken.  Pairs taken in earlier
      // rounds come before it in that order, so they are out as well.
      uint32_t m = 0;
      int mc = 0;
#pragma unroll
      for (int e = 0; e < kTopkGroup; ++e) {
        const int c = bg * kTopkGroup + e;
        const uint32_t k = row[c];
        if constexpr (DT == TMB_F32) {
          const bool after = (k < best) | ((k == best) & (c > cls));
          if (after & (k > m)) {
            m = k;
            mc = c;
          }
        } else {
          m = max(m, k < best ? k : 0u);
        }
      }
      mine[bg * tile_rows] = m;
      if constexpr (DT == TMB_F32) mine_cls[bg * tile_rows] = (uint32_t)mc;
    }
    atomicAdd(&sh_topk[(int)t], 1);
    if (hit) atomicAdd(&sh_topk[2 * C + (int)t], 1);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 3 * C; i += blockDim.x) {
    const int v = sh_topk[i];
    if (v) atomicAdd(&out[i], (unsigned long long)v);
  }
}

// Rows that do not fit the register-resident kernels (C > 1024), and the per-sample histograms of
// multidim_average="samplewise": a warp walks the row once per selection round (element c in lane c % 32; after the
// first round the row comes from L1 / L2) and takes the largest (key, lowest class) pair strictly below the pair the
// previous round took — nothing is marked, so the row needs no storage.  Ties at the k-th score go to the lower class
// index, NaN ranks highest: the order of the other top-k kernels.  `rows_per_sample` > 0: the three histograms of row
// r go to the block of sample r / rows_per_sample (global atomics); 0: one [3 C] histogram, privatised in shared
// memory when `use_shared`.
template <int DT>
__global__ void __launch_bounds__(256) k1_rows_topk_wide(const void* __restrict__ preds, const void* __restrict__ target,
                                                         int tdt, int64_t rows, int C, int top_k, int has_ignore,
                                                         int64_t ignore, int64_t rows_per_sample, int use_shared,
                                                         unsigned long long* __restrict__ out, int* err) {
  extern __shared__ int sh_topk[];  // [3 * C] when use_shared
  if (use_shared) {
    for (int i = threadIdx.x; i < 3 * C; i += blockDim.x) sh_topk[i] = 0;
    __syncthreads();
  }
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < rows; r += nwarps) {  // warp-uniform
    const int64_t t = load_label(target, tdt, r);
    if (has_ignore && t == ignore) continue;
    if (t < 0 || t >= C) {
      if (lane == 0) atomicOr(err, TMB_ERR_TARGET_RANGE);
      continue;
    }
    unsigned long long* blk = out + (rows_per_sample > 0 ? (r / rows_per_sample) * 3 * (int64_t)C : 0);
    auto add = [&](int64_t bin) {
      if (use_shared)
        atomicAdd(&sh_topk[bin], 1);
      else
        atomicAdd(&blk[bin], 1ULL);
    };
    // the pair taken by the previous round; (key, class) pairs are ordered by key descending, then class ascending
    unsigned long long prev = ~0ULL;  // packed = key << 32 | (0xffffffff - class); a NaN in class 0 packs to ~0 itself,
                                      // so the first round takes every pair instead of comparing with `prev`
    bool hit = false;
    for (int round = 0; round < top_k; ++round) {
      unsigned long long best = 0ULL;
      for (int c = lane; c < C; c += 32) {
        const uint32_t key = topk_key(load_elem<DT>(preds, r * C + c));
        const unsigned long long packed = ((unsigned long long)key << 32) | (0xffffffffu - (uint32_t)c);
        if ((round == 0 || packed < prev) && packed > best) best = packed;
      }
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) {
        const unsigned long long o = __shfl_xor_sync(kFull, best, d);
        best = o > best ? o : best;
      }
      prev = best;
      const int64_t cls = (int64_t)(0xffffffffu - (uint32_t)(best & 0xffffffffu));
      if (lane == 0) add(C + cls);
      hit |= cls == t;
    }
    if (lane == 0) {
      add(t);
      if (hit) add(2 * (int64_t)C + t);
    }
  }
  if (use_shared) {
    __syncthreads();
    for (int i = threadIdx.x; i < 3 * C; i += blockDim.x) {
      const int v = sh_topk[i];
      if (v) atomicAdd(&out[i], (unsigned long long)v);
    }
  }
}

template <int DT>
static int launch_rows_topk_wide(const void* preds, const void* target, int tdt, int64_t rows, int C, int top_k,
                                 int has_ignore, int64_t ignore, int64_t rows_per_sample, unsigned long long* out,
                                 int* err, cudaStream_t st) {
  auto k = k1_rows_topk_wide<DT>;
  const int64_t want = (int64_t)3 * C * (int)sizeof(int);
  const int use_shared = rows_per_sample == 0 && want <= (int64_t)dev_info().max_smem_optin - 1024;
  const int smem = use_shared ? (int)want : 0;
  if (smem > 48 * 1024) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int64_t work = (rows * 32 + 255) / 256;
  k<<<resident_grid(k, 256, smem, work), 256, smem, st>>>(preds, target, tdt, rows, C, top_k, has_ignore, ignore,
                                                         rows_per_sample, use_shared, out, err);
  return launch_status();
}

template <int DT>
static int launch_rows_topk(const void* preds, const void* target, int tdt, int64_t rows, int C, int top_k,
                            int has_ignore, int64_t ignore, unsigned long long* out, int* err, cudaStream_t st) {
  if (C <= 128) {  // one thread per row
    const int tile_rows = C <= 64 ? 256 : 128;
    const int groups = (C + kTopkGroup - 1) / kTopkGroup;
    const int smem = ((3 * C + 31) / 32 * 32 + tile_rows * ((groups * kTopkGroup) | 1) +
                      tile_rows * groups * (DT == TMB_F32 ? 2 : 1)) * (int)sizeof(int);
    auto k = k1_rows_topk_thread<DT>;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    const int64_t work = (rows + tile_rows - 1) / tile_rows;
    k<<<resident_grid(k, tile_rows, smem, work), tile_rows, smem, st>>>(preds, target, tdt, rows, C, top_k, has_ignore,
                                                                       ignore, out, err);
    return launch_status();
  }
  const RowMap rm = row_map(C < 32 ? 32 : C);  // always 32 lanes per row here; slots = next_pow2(C) / 32
  const int smem = 3 * C * (int)sizeof(int);
  const int64_t work = (rows * 32 + 255) / 256;
  if constexpr (DT == TMB_F32) {
    if (C % 4 == 0 && (((uintptr_t)preds) & 15) == 0) {  // 16-byte aligned rows: vector loads, next row in flight
      const int vecs = (C / 4 + 31) / 32;  // 16-byte vectors per lane: 1 .. 8
#define TMB_LAUNCH_TOPK32(V)                                                                                   \
  {                                                                                                            \
    auto k = k1_rows_topk32<V>;                                                                                \
    k<<<resident_grid(k, 256, smem, work), 256, smem, st>>>(preds, target, tdt, rows, C, top_k, has_ignore,   \
                                                           ignore, out, err);                                 \
  }
      switch (vecs) {
        case 1: TMB_LAUNCH_TOPK32(1) break;
        case 2: TMB_LAUNCH_TOPK32(2) break;
        case 3: TMB_LAUNCH_TOPK32(3) break;
        case 4: TMB_LAUNCH_TOPK32(4) break;
        case 5: TMB_LAUNCH_TOPK32(5) break;
        case 6: TMB_LAUNCH_TOPK32(6) break;
        case 7: TMB_LAUNCH_TOPK32(7) break;
        default: TMB_LAUNCH_TOPK32(8) break;
      }
#undef TMB_LAUNCH_TOPK32
      return launch_status();
    }
  }
  if constexpr (DT != TMB_F32) {
    if (C % 8 == 0 && (((uintptr_t)preds) & 15) == 0) {  // 16-byte aligned rows: packed keys, vector loads
      const int vecs = (C / 8 + 31) / 32;  // 16-byte vectors per lane: 1 .. 4
#define TMB_LAUNCH_TOPK16(V)                                                                                   \
  {                                                                                                            \
    auto k = k1_rows_topk16<DT, V>;                                                                            \
    k<<<resident_grid(k, 256, smem, work), 256, smem, st>>>(preds, target, tdt, rows, C, top_k, has_ignore,   \
                                                           ignore, out, err);                                 \
  }
      switch (vecs) {
        case 1: TMB_LAUNCH_TOPK16(1) break;
        case 2: TMB_LAUNCH_TOPK16(2) break;
        case 3: TMB_LAUNCH_TOPK16(3) break;
        default: TMB_LAUNCH_TOPK16(4) break;
      }
#undef TMB_LAUNCH_TOPK16
      return launch_status();
    }
  }
  TMB_DISPATCH_SLOTS(rm.slots, {
    auto k = k1_rows_topk<DT, SLOTS>;
    k<<<resident_grid(k, 256, smem, work), 256, smem, st>>>(preds, target, tdt, rows, C, top_k, has_ignore, ignore,
                                                           out, err);
  })
  return launch_status();
}

}  // namespace tmb

using namespace tmb;

extern "C" int tmb_build_arch(void) { return 100; }

extern "C" int64_t tmb_mc_scratch_bytes(int64_t num_classes, int mode) {
  if (mode == TMB_MC_STATS) return 3 * num_classes * (int64_t)sizeof(int64_t);
  if (mode == TMB_MC_MICRO) return 2 * (int64_t)sizeof(int64_t);
  if (mode == TMB_MC_BOTH) return (num_classes * num_classes + 3 * num_classes) * (int64_t)sizeof(int64_t);
  return 0;
}

extern "C" int tmb_mc_argmax_hist(const void* preds, int preds_dtype, const void* target, int target_dtype,
                                  int64_t n_outer, int64_t num_classes, int64_t n_inner, int has_ignore,
                                  int64_t ignore_index, int mode, int64_t* out, int32_t* err, void* stream) {
  if (float_bytes(preds_dtype) == 0 || label_bytes(target_dtype) == 0) return TMB_E_DTYPE;
  if (n_outer < 0 || n_inner < 1) return TMB_E_SHAPE;
  HistArgs h;
  if (int rc = make_hist(h, mode, num_classes, has_ignore, ignore_index, out, err)) return rc;
  if (n_outer == 0) return 0;
  if (!preds || !target) return TMB_E_NULL;
  cudaStream_t st = (cudaStream_t)stream;
  if (n_inner == 1) {
    switch (preds_dtype) {
      case TMB_F32: return launch_rows<TMB_F32>(preds, target, target_dtype, n_outer, h, st);
      case TMB_F16: return launch_rows<TMB_F16>(preds, target, target_dtype, n_outer, h, st);
      default: return launch_rows<TMB_BF16>(preds, target, target_dtype, n_outer, h, st);
    }
  }
  switch (preds_dtype) {
    case TMB_F32: return launch_strided<TMB_F32>(preds, target, target_dtype, n_outer, n_inner, h, st);
    case TMB_F16: return launch_strided<TMB_F16>(preds, target, target_dtype, n_outer, n_inner, h, st);
    default: return launch_strided<TMB_BF16>(preds, target, target_dtype, n_outer, n_inner, h, st);
  }
}

extern "C" int tmb_mc_hist_samplewise(const void* preds, int preds_dtype, int preds_is_label, const void* target,
                                      int target_dtype, int64_t n_samples, int64_t num_classes, int64_t n_inner,
                                      int has_ignore, int64_t ignore_index, int64_t* scratch, int32_t* err,
                                      void* stream) {
  if (label_bytes(target_dtype) == 0) return TMB_E_DTYPE;
  if (preds_is_label ? label_bytes(preds_dtype) == 0 : float_bytes(preds_dtype) == 0) return TMB_E_DTYPE;
  if (n_samples < 0 || n_inner < 1) return TMB_E_SHAPE;
  HistArgs h;
  if (int rc = make_hist(h, TMB_MC_STATS, num_classes, has_ignore, ignore_index, scratch, err)) return rc;
  h.sh_bins = 0;  // every sample has its own histograms: straight to global memory
  h.per_sample = 3 * num_classes;
  h.sample_len = n_inner;
  if (n_samples == 0) return 0;
  if (!preds || !target) return TMB_E_NULL;
  cudaStream_t st = (cudaStream_t)stream;
  if (preds_is_label) {
    const int64_t n = n_samples * n_inner;
    k1_labels<<<wave_grid((n + 1023) / 1024, 8), 256, 0, st>>>(preds, preds_dtype, target, target_dtype, n, h);
    return launch_status();
  }
  switch (preds_dtype) {  // [N, C, S] with S >= 1: the strided kernel (a thread per pixel group) knows its sample
    case TMB_F32: return launch_strided<TMB_F32>(preds, target, target_dtype, n_samples, n_inner, h, st);
    case TMB_F16: return launch_strided<TMB_F16>(preds, target, target_dtype, n_samples, n_inner, h, st);
    default: return launch_strided<TMB_BF16>(preds, target, target_dtype, n_samples, n_inner, h, st);
  }
}

extern "C" int tmb_mc_stats_finalize_samplewise(int64_t* scratch, int64_t n_samples, int64_t num_classes,
                                                int64_t* tp, int64_t* fp, int64_t* tn, int64_t* fn, void* stream) {
  if (!scratch || !tp || !fp || !tn || !fn) return TMB_E_NULL;
  if (n_samples < 0 || num_classes < 1 || num_classes > (1 << 24) || n_samples > 0x7fffffff) return TMB_E_SHAPE;
  if (n_samples == 0) return 0;
  k1_stats_finalize_rows<<<(unsigned)n_samples, 256, 0, (cudaStream_t)stream>>>(
      (unsigned long long*)scratch, (int)num_classes, (long long*)tp, (long long*)fp, (long long*)tn, (long long*)fn);
  return launch_status();
}

extern "C" int tmb_mc_shared_confmat(int64_t* scratch, int64_t num_classes, int64_t* confmat, void* stream) {
  if (!scratch || !confmat) return TMB_E_NULL;
  if (num_classes < 1 || num_classes > (1 << 15)) return TMB_E_SHAPE;
  const int64_t n = num_classes * num_classes;
  k1_add_clear<<<wave_grid((n + 255) / 256, 4), 256, 0, (cudaStream_t)stream>>>((unsigned long long*)scratch,
                                                                                (long long*)confmat, n);
  return launch_status();
}

extern "C" int tmb_mc_shared_stats(int64_t* scratch, int64_t num_classes, int micro, int clear, int64_t* tp,
                                   int64_t* fp, int64_t* tn, int64_t* fn, void* stream) {
  if (!scratch || !tp || !fp || !tn || !fn) return TMB_E_NULL;
  if (num_classes < 1 || num_classes > (1 << 15)) return TMB_E_SHAPE;
  unsigned long long* stats = (unsigned long long*)scratch + num_classes * num_classes;
  k1_stats_share<<<1, 1024, 0, (cudaStream_t)stream>>>(stats, (int)num_classes, micro, clear, (long long*)tp,
                                                       (long long*)fp, (long long*)tn, (long long*)fn);
  return launch_status();
}

extern "C" int tmb_mc_topk_hist(const void* preds, int preds_dtype, const void* target, int target_dtype,
                                int64_t n_rows, int64_t num_classes, int64_t top_k, int has_ignore,
                                int64_t ignore_index, int64_t* scratch, int32_t* err, void* stream) {
  if (float_bytes(preds_dtype) == 0 || label_bytes(target_dtype) == 0) return TMB_E_DTYPE;
  if (n_rows < 0 || num_classes < 2 || top_k < 1 || top_k > num_classes) return TMB_E_SHAPE;
  if (num_classes > (1 << 24)) return TMB_E_LIMIT;
  if (!scratch || !err) return TMB_E_NULL;
  if (n_rows == 0) return 0;
  if (!preds || !target) return TMB_E_NULL;
  cudaStream_t st = (cudaStream_t)stream;
  unsigned long long* out = (unsigned long long*)scratch;
  const int C = (int)num_classes, k = (int)top_k;
  if (num_classes > kSoftmaxMaxClasses) {  // the row no longer fits 32 registers per lane: one walk per round
    switch (preds_dtype) {
      case TMB_F32: return launch_rows_topk_wide<TMB_F32>(preds, target, target_dtype, n_rows, C, k, has_ignore, ignore_index, 0, out, err, st);
      case TMB_F16: return launch_rows_topk_wide<TMB_F16>(preds, target, target_dtype, n_rows, C, k, has_ignore, ignore_index, 0, out, err, st);
      default: return launch_rows_topk_wide<TMB_BF16>(preds, target, target_dtype, n_rows, C, k, has_ignore, ignore_index, 0, out, err, st);
    }
  }
  switch (preds_dtype) {
    case TMB_F32: return launch_rows_topk<TMB_F32>(preds, target, target_dtype, n_rows, C, k, has_ignore, ignore_index, out, err, st);
    case TMB_F16: return launch_rows_topk<TMB_F16>(preds, target, target_dtype, n_rows, C, k, has_ignore, ignore_index, out, err, st);
    default: return launch_rows_topk<TMB_BF16>(preds, target, target_dtype, n_rows, C, k, has_ignore, ignore_index, out, err, st);
  }
}

extern "C" int tmb_mc_topk_hist_samplewise(const void* preds, int preds_dtype, const void* target, int target_dtype,
                                           int64_t n_samples, int64_t rows_per_sample, int64_t num_classes,
                                           int64_t top_k, int has_ignore, int64_t ignore_index, int64_t* scratch,
                                           int32_t* err, void* stream) {
  if (float_bytes(preds_dtype) == 0 || label_bytes(target_dtype) == 0) return TMB_E_DTYPE;
  if (n_samples < 0 || rows_per_sample < 1 || num_classes < 2 || top_k < 1 || top_k > num_classes) return TMB_E_SHAPE;
  if (num_classes > (1 << 24)) return TMB_E_LIMIT;
  if (!scratch || !err) return TMB_E_NULL;
  if (n_samples == 0) return 0;
  if (!preds || !target) return TMB_E_NULL;
  cudaStream_t st = (cudaStream_t)stream;
  unsigned long long* out = (unsigned long long*)scratch;
  const int C = (int)num_classes, k = (int)top_k;
  const int64_t rows = n_samples * rows_per_sample;
  switch (preds_dtype) {
    case TMB_F32: return launch_rows_topk_wide<TMB_F32>(preds, target, target_dtype, rows, C, k, has_ignore, ignore_index, rows_per_sample, out, err, st);
    case TMB_F16: return launch_rows_topk_wide<TMB_F16>(preds, target, target_dtype, rows, C, k, has_ignore, ignore_index, rows_per_sample, out, err, st);
    default: return launch_rows_topk_wide<TMB_BF16>(preds, target, target_dtype, rows, C, k, has_ignore, ignore_index, rows_per_sample, out, err, st);
  }
}

extern "C" int tmb_mc_label_hist(const void* preds, int preds_dtype, const void* target, int target_dtype,
                                 int64_t n, int64_t num_classes, int has_ignore, int64_t ignore_index, int mode,
                                 int64_t* out, int32_t* err, void* stream) {
  if (label_bytes(preds_dtype) == 0 || label_bytes(target_dtype) == 0) return TMB_E_DTYPE;
  if (n < 0) return TMB_E_SHAPE;
  HistArgs h;
  if (int rc = make_hist(h, mode, num_classes, has_ignore, ignore_index, out, err)) return rc;
  if (n == 0) return 0;
  if (!preds || !target) return TMB_E_NULL;
  const int block = 256;
  const int grid = wave_grid((n + block * 4 - 1) / (block * 4), 8);
  k1_labels<<<grid, block, h.sh_bins * sizeof(int), (cudaStream_t)stream>>>(preds, preds_dtype, target,
                                                                            target_dtype, n, h);
  return launch_status();
}

extern "C" int tmb_mc_stats_finalize(int64_t* scratch, int64_t num_classes, int mode, int64_t* tp, int64_t* fp,
                                     int64_t* tn, int64_t* fn, void* stream) {
  if (mode != TMB_MC_STATS && mode != TMB_MC_MICRO) return TMB_E_MODE;
  if (!scratch || !tp || !fp || !tn || !fn) return TMB_E_NULL;
  if (num_classes < 1 || num_classes > (1 << 24)) return TMB_E_SHAPE;
  k1_stats_finalize<<<1, 1024, 0, (cudaStream_t)stream>>>((unsigned long long*)scratch, (int)num_classes, mode,
                                                          (long long*)tp, (long long*)fp, (long long*)tn,
                                                          (long long*)fn);
  return launch_status();
}
