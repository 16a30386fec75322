// K2 / K1b — one-pass threshold binning.
//
// The reference evaluates `preds >= thresholds[i]` once per threshold (200 full passes and ~1800
// launches at T = 200, functional/classification/precision_recall_curve.py:228-250) or materialises an
// N x T comparison (:210-225, :785-793).  Here every sample is read once: its bin index
// k = #{j : thr[j] <= q} is found with a uniform-grid guess verified by exact compares against the
// metric's own threshold buffer (never a recomputed linspace, H4), and hist[label][target][k] is bumped.
// The per-threshold counts are suffix sums of that histogram, formed by a tiny finalize kernel.
//
// Conditional sigmoid (H1): the reference applies sigmoid iff ANY value of the batch lies outside [0,1].
// Two launches share one read of the data: pass A bins the raw values and raises a device flag at the
// first outside value (every CTA polls the flag and quits, so for logits this pass ends after the first
// wave); pass B bins sigmoid(x) and returns immediately when the flag is down.  Finalize picks.
//
// Hot case (binary, config 3): a shared-memory atomic per sample would cap the SM at <1 sample/clk
// (B300_MICROARCH: ATOMS ~2 clk/lane), an order of magnitude under the HBM rate.  Instead every LANE owns
// private counters — one byte per (bin, target class, lane), plain LDS.U8 / IADD / STS.U8, no conflicts, no
// atomics — that are folded into a CTA-wide 32-bit histogram every 31 batches (a byte holds 255 hits).  The
// kernel is instruction-issue bound, not HBM bound (12 B per sample against ~35 instructions), so the rest of
// the design is about issue slots: a 4-instruction approximate sigmoid with a rigorous error bound decides the
// bin unless the score is within that bound of a threshold (then the exact expression is evaluated), evenly
// spaced thresholds are recognised on the device and tested arithmetically instead of by loads, each thread
// works on 8 samples stage by stage so their dependency chains interleave, and 16 warps per SM (bytes, not
// 16-bit pairs, made that fit) cover the rest.  Loads: one register batch ahead plus L2 prefetches four
// batches ahead.
#include "common.cuh"
#include "softmax_rows.cuh"

namespace tmb {

constexpr int kHdr = 8;  // scratch[0] = "saw a value outside [0,1]" flag, [1..7] reserved

struct TbArgs {
  const void* preds;
  const void* target;
  const float* thr;
  unsigned long long* scratch;
  int* err;
  int64_t n_total;  // n_outer * L * S
  int64_t S;
  int64_t ignore;
  int L;
  int T;
  int pdt;  // float code, or label code when preds_is_label
  int tdt;
  int preds_is_label;
  int cmp_gt;
  int has_ignore;
  int flag_counts_ignored;
};

__device__ __forceinline__ unsigned long long* hist_base(const TbArgs& a, int variant) {
  return a.scratch + kHdr + (int64_t)variant * a.L * 2 * (a.T + 1);
}

// torch's CUDA sigmoid for fp32 opmath: 1 / (1 + exp(-x)), IEEE division, full-precision expf
// (aten/src/ATen/native/cuda/UnarySpecialOpsKernel.cu; the reference calls it at
// precision_recall_curve.py:184).  Narrow dtypes round the result back, like the reference tensor would.
template <int DT>
__device__ __forceinline__ float sigmoid_as(float x) {
  float y = 1.0f / (1.0f + expf(-x));
  if constexpr (DT == TMB_BF16) y = __bfloat162float(__float2bfloat16_rn(y));
  if constexpr (DT == TMB_F16) y = __half2float(__float2half_rn(y));
  return y;
}

// Number of thresholds j with thr[j] <= q (GE mode) or thr[j] < q (GT mode); thr ascending; NaN -> 0.
// `pad` is the threshold array with sentinels: pad[0] = -inf, pad[1 + j] = thr[j], pad[T + 1] = +inf, so
// "k is the right bin" reads  pad[k] <= q  and not  pad[k + 1] <= q  without any index special-casing.
struct Binner {
  const float* pad;  // shared or global, T + 2 entries
  int T;
  float lo, scale;  // guess = (q - lo) * scale
  // u(q) = fmaf(q, scale, 1 - lo*scale) maps threshold j to (nearly) the integer j + 1.  grid_eps is the largest
  // distance of any u(thr[j]) from j + 1: because u() is monotone in q, a sample whose u lies further than grid_eps
  // from every integer is on a known side of every threshold — no compare against the buffer needed.  A small
  // grid_eps means the thresholds are (numerically) evenly spaced, e.g. torch.linspace.
  float grid_eps;
  bool gt;
  __device__ __forceinline__ bool below(float t, float q) const { return gt ? (t < q) : (t <= q); }
  __device__ __forceinline__ int guess(float q) const {
    float gf = (q - lo) * scale;
    gf = fminf(fmaxf(gf, -1.0f), (float)T);  // NaN -> -1
    int g = (int)gf + 1;
    return g > T ? T : g;
  }
  __device__ __forceinline__ bool verified(int g, float q) const {
    return (below(pad[g], q) && !below(pad[g + 1], q)) || (q != q && g == 0);
  }
  __device__ __forceinline__ int bisect(float q) const {  // cold path: guess was off
    int lo_i = 0, hi_i = T;
    while (lo_i < hi_i) {
      int mid = (lo_i + hi_i) >> 1;
      if (below(pad[mid + 1], q)) lo_i = mid + 1; else hi_i = mid;
    }
    return lo_i;
  }
  __device__ __forceinline__ int bin(float q) const {
    int g = guess(q);
    return verified(g, q) ? g : bisect(q);
  }
};

// `sh_pad` must hold T + 2 floats when T + 2 <= sh_cap; otherwise the padded copy in the scratch tail is
// not available and the caller passes a global padded array instead.
__device__ __forceinline__ Binner make_binner(const TbArgs& a, float* sh_pad, int sh_cap, const float* global_pad) {
  Binner b;
  b.T = a.T;
  b.gt = a.cmp_gt != 0;
  if (a.T + 2 <= sh_cap) {
    for (int i = threadIdx.x; i < a.T + 2; i += blockDim.x)
      sh_pad[i] = (i == 0) ? -INFINITY : (i == a.T + 1 ? INFINITY : a.thr[i - 1]);
    __syncthreads();
    b.pad = sh_pad;
  } else {
    b.pad = global_pad;
  }
  float lo = a.thr[0], hi = a.thr[a.T - 1];
  float span = hi - lo;
  b.lo = lo;
  b.scale = (span > 0.0f && a.T > 1) ? (float)(a.T - 1) / span : 0.0f;
  // how far the thresholds are from the evenly spaced grid, in bins (CTA-wide maximum)
  __shared__ unsigned grid_eps_bits;
  if (threadIdx.x == 0) grid_eps_bits = 0u;
  __syncthreads();
  const float bias = 1.0f - lo * b.scale;
  float worst = 0.0f;
  for (int j = threadIdx.x; j < a.T; j += blockDim.x) {
    float d = fabsf(fmaf(a.thr[j], b.scale, bias) - (float)(j + 1));
    worst = (d == d) ? fmaxf(worst, d) : INFINITY;
  }
  atomicMax(&grid_eps_bits, __float_as_uint(worst));  // non-negative floats order like their bit patterns
  __syncthreads();
  b.grid_eps = (b.scale > 0.0f) ? __uint_as_float(grid_eps_bits) : INFINITY;
  return b;
}

__device__ __forceinline__ bool flag_up(const TbArgs& a) {
  return *((volatile unsigned long long*)a.scratch) != 0ULL;
}
__device__ __forceinline__ void raise_flag(const TbArgs& a) { *((volatile unsigned long long*)a.scratch) = 1ULL; }

// CTA-uniform decision whether this pass has anything to do (the flag may flip while pass A runs, so
// one thread samples it for everybody).
__device__ __forceinline__ bool pass_enabled(const TbArgs& a, bool sig_pass) {
  __shared__ int go;
  if (threadIdx.x == 0) go = sig_pass ? flag_up(a) : !flag_up(a);
  __syncthreads();
  return go != 0;
}

// ---- binary fast path: lane-private packed histograms ------------------------------------------------
// Targets are decoded from their raw 32-bit halves (no 64-bit arithmetic in the hot loop): four
// consecutive targets starting at element i (i % 4 == 0, base pointer 16-byte aligned).
template <int TDT>
__device__ __forceinline__ void load_targets4(const void* base, int64_t i, uint32_t lo[4], uint32_t hi[4]) {
  if constexpr (TDT == TMB_I64) {
    uint4 a = ldg_stream16((const char*)base + i * 8);
    uint4 b = ldg_stream16((const char*)base + i * 8 + 16);
    lo[0] = a.x; hi[0] = a.y; lo[1] = a.z; hi[1] = a.w;
    lo[2] = b.x; hi[2] = b.y; lo[3] = b.z; hi[3] = b.w;
  } else if constexpr (TDT == TMB_I32) {
    uint4 a = ldg_stream16((const char*)base + i * 4);
    lo[0] = a.x; lo[1] = a.y; lo[2] = a.z; lo[3] = a.w;
#pragma unroll
    for (int e = 0; e < 4; ++e) hi[e] = (uint32_t)((int)lo[e] >> 31);
  } else {
    uint32_t w = ldg_stream4((const char*)base + i);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      uint32_t b = (w >> (8 * e)) & 0xffu;
      if (TDT == TMB_I8) {
        lo[e] = (uint32_t)(int)(signed char)b;
        hi[e] = (uint32_t)((int)lo[e] >> 31);
      } else {
        lo[e] = b;
        hi[e] = 0u;
      }
    }
  }
}

template <int PDT>
__device__ __forceinline__ void load_preds4(const void* base, int64_t i, float x[4]) {
  if constexpr (PDT == TMB_F32) {
    uint4 a = ldg_stream16((const char*)base + i * 4);
    x[0] = __uint_as_float(a.x); x[1] = __uint_as_float(a.y);
    x[2] = __uint_as_float(a.z); x[3] = __uint_as_float(a.w);
  } else {
    uint2 a = ldg_stream8((const char*)base + i * 2);
    x[0] = Elem<PDT>::get(a.x, 0); x[1] = Elem<PDT>::get(a.x, 1);
    x[2] = Elem<PDT>::get(a.y, 0); x[3] = Elem<PDT>::get(a.y, 1);
  }
}

// Lane-private counters, one BYTE per (bin, target class, lane): byte (k*2 + t)*32 + lane of the warp's block.
// Four neighbouring lanes share a 32-bit word and therefore a bank, but they touch different bytes of it, which
// the shared-memory pipe merges like any sub-word access: no conflicts, no atomics.  Bytes instead of 16-bit
// pairs halve the footprint — (T+1)*64 B per warp, 16 warps per SM at T = 200 instead of 8 — and occupancy is
// what this kernel is short of (with 2 warps per scheduler every dependent instruction pair left issue slots
// empty: "stall_wait" was the top stall reason at 49 % issue utilisation).  A byte overflows after 255 hits, so
// each warp folds its block into the CTA-wide 32-bit histogram every kFoldEvery batches: one lane per row,
// 32 bytes summed with eight dp4a, about 1 extra instruction per sample.
struct LaneHist {
  uint8_t* col;        // this warp's [(T+1)*2][32] bytes, already offset by the lane
  uint8_t* warp_base;  // the same block without the lane offset (16-byte aligned)
  uint32_t* cta_hist;  // [2*(T+1)] words shared by the CTA, row r = k*2 + t
  int lane;
  int batches;         // batches binned since the last fold
};

constexpr int kBatch = 2;         // vectors of 4 samples per thread per iteration
constexpr int kPer = kBatch * 4;  // samples per thread per iteration
constexpr int kFoldEvery = 255 / kPer;  // a lane adds at most kPer to one byte per batch

__device__ __forceinline__ void lane_fold(const TbArgs& a, LaneHist& h) {
  __syncwarp();
  const int rows = 2 * (a.T + 1);
  for (int r = h.lane; r < rows; r += 32) {
    uint4* row = reinterpret_cast<uint4*>(h.warp_base + r * 32);
    const uint4 lo = row[0], hi = row[1];
    row[0] = make_uint4(0u, 0u, 0u, 0u);
    row[1] = make_uint4(0u, 0u, 0u, 0u);
    unsigned s = 0;
    s = __dp4a(lo.x, 0x01010101u, s); s = __dp4a(lo.y, 0x01010101u, s);
    s = __dp4a(lo.z, 0x01010101u, s); s = __dp4a(lo.w, 0x01010101u, s);
    s = __dp4a(hi.x, 0x01010101u, s); s = __dp4a(hi.y, 0x01010101u, s);
    s = __dp4a(hi.z, 0x01010101u, s); s = __dp4a(hi.w, 0x01010101u, s);
    if (s) atomicAdd(&h.cta_hist[r], s);
  }
  h.batches = 0;
  __syncwarp();
}

struct Batch {  // what the loads deliver
  float x[kPer];
  uint32_t lo[kPer], hi[kPer];
};

struct Decoded {  // what is left of a batch once its registers have been released to the next loads
  float x[kPer];
  uint32_t inc[kPer];   // 1 = counted, 0 = not
  uint32_t toff[kPer];  // byte offset of the target class inside a bin: t * 32
  unsigned bad;         // non-zero: a target that is neither 0, 1 nor ignored
  unsigned outside;     // non-zero: a value outside [0,1] that takes part in the sigmoid test (pass A)
};

// Filtered exact binning of sigmoid(x) (fp32 only).  The bin only changes when q crosses a threshold, so the
// exact value of q is needed only near one.  A 4-instruction approximation qa = 1/(1 + 2^(-x*log2e))
// (FMUL, MUFU.EX2, FADD, MUFU.RCP) comes with a rigorous bound on its distance from q, the value torch's
// kernel (full-precision expf + IEEE divide) returns:
//   2^(..) vs e^-x : ex2.approx is good to 2 ulp and the rounding of x*log2e costs |x*log2e| * 2^-24 in the
//                    exponent, i.e. at most (2 + 1.173|x|) ulp in e  (the bound CUDA documents for __expf)
//   1 + e          : 0.5 ulp;   approximate reciprocal: 1 ulp
//   q itself       : torch's expression is within 3 ulp of the true value
// so |qa - q| <= (6.5 + 1.173|x|) ulp * q.  The margin used is twice that, rel(x) = 1.6e-6 + 2.9e-7|x|, plus
// 1e-30 absolute for the flushed-to-zero ends (|x| > 87, where both expressions saturate at 0 / 1).  If the
// approximate score clears every threshold by more than the margin the bin is certain; otherwise (about 0.2 %
// of the samples at T = 200) the sample is re-done with the exact expression.  The result is bit-identical to
// binning torch's sigmoid at a third of the instructions.
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Cold path.  Arguments by value so that the caller's Binner stays in registers (a by-reference call put it in
// local memory and the reloads showed up as long-scoreboard stalls): exact score, then guess / verify / bisect
// against the threshold buffer.
template <int PDT, bool SIG>
__device__ __noinline__ int exact_bin(const float* pad, int T, float lo, float scale, bool gt, float x) {
  Binner b;
  b.pad = pad;
  b.T = T;
  b.lo = lo;
  b.scale = scale;
  b.gt = gt;
  b.grid_eps = 0.0f;
  if (SIG && PDT == TMB_F32) {
    // The grid test loses its resolution right at the first / last threshold, where saturated sigmoids pile up
    // (q ~ 1e-9 is certainly above a threshold of 0, yet u(q) rounds onto the integer).  Before paying for the
    // exact expression, bracket the score: the bin is monotone in q, so if q - margin and q + margin fall into
    // the same bin of the real threshold buffer, that is the bin.
    const float q = rcp_approx(1.0f + ex2_approx(x * -1.4426950408889634f));
    const float margin = fmaf(q, fmaf(fabsf(x), 2.9e-7f, 1.6e-6f), 1e-30f);
    if (margin < 1.0f) {  // false for NaN / infinite inputs: those go straight to the exact expression
      const int k_low = b.bin(q - margin), k_high = b.bin(q + margin);
      if (k_low == k_high) return k_low;
    }
  }
  return b.bin(SIG ? sigmoid_as<PDT>(x) : x);
}

// First touch of a freshly loaded batch: everything that depends on the targets, and a copy of the scores.
// After this the load registers are dead, so the NEXT batch can be requested into them right away; the wait
// that precedes this function therefore only ever covers loads that are really needed now.  (With two register
// buffers the wait for batch i also waited for the just-issued loads of batch i+1 whenever both shared one of
// the six scoreboard slots — 21 % of the kernel's stall samples sat on that one instruction.)
template <bool SIG, bool FULL, bool IGN>
__device__ __forceinline__ void decode_batch(const TbArgs& a, const Batch& s, unsigned live, Decoded& d) {
  const uint32_t ign_lo = (uint32_t)a.ignore, ign_hi = (uint32_t)((uint64_t)a.ignore >> 32);
  d.bad = 0;
  d.outside = 0;
#pragma unroll
  for (int i = 0; i < kPer; ++i) {
    const unsigned is_live = FULL ? 1u : ((live >> i) & 1u);
    const unsigned kept = IGN ? (unsigned)!((s.lo[i] == ign_lo) & (s.hi[i] == ign_hi)) : 1u;
    const unsigned not01 = s.hi[i] | (s.lo[i] >> 1);  // zero exactly for targets 0 and 1
    const unsigned counted = is_live & kept & (unsigned)(not01 == 0u);
    d.bad |= (is_live & kept) ? not01 : 0u;
    if (!SIG) {
      const unsigned outside = (unsigned)!(s.x[i] >= 0.0f && s.x[i] <= 1.0f);
      d.outside |= is_live & outside & (kept | (unsigned)(a.flag_counts_ignored != 0));
    }
    d.inc[i] = counted;
    d.toff[i] = (s.lo[i] & 1u) * 32u;
    d.x[i] = s.x[i];
  }
}

// Bins of kPer decoded samples.  Written stage by stage over all kPer samples (not sample by sample) so that the
// dependency chains of different samples interleave.  GRID = thresholds on a (numerically) even grid: certainty
// comes from arithmetic alone; otherwise the two neighbouring thresholds of the guess are fetched and compared.
template <int PDT, bool SIG, bool GRID>
__device__ __forceinline__ void batch_bins(const Binner& bn, const Decoded& d, int (&g)[kPer]) {
  constexpr bool kFiltered = SIG && PDT == TMB_F32;
  float q[kPer], lowq[kPer], highq[kPer], slack[kPer];
  const float bias = 1.0f - bn.lo * bn.scale;
  const float top = (float)bn.T;

  // stage 1: the score each sample is binned by
  if (kFiltered) {
    float e[kPer];
#pragma unroll
    for (int i = 0; i < kPer; ++i) e[i] = ex2_approx(d.x[i] * -1.4426950408889634f);
#pragma unroll
    for (int i = 0; i < kPer; ++i) q[i] = rcp_approx(1.0f + e[i]);
    if (GRID) {
      // uncertainty of u(q) in bins: the score margin scaled, the grid's own unevenness, 2 ulp of u.  The margin
      // rel(x) * q + 1e-30 is bounded with q <= 1, which makes the slack ONE fma of |x| (constants folded per
      // kernel) instead of three dependent ones; the bound is looser only where q is small, i.e. below the first
      // thresholds, and the fraction of samples sent to the exact path stays ~0.2 %.
      const float c1 = 2.9e-7f * bn.scale;
      const float c0 = fmaf(1.6e-6f, bn.scale, bn.grid_eps + top * 2.4e-7f + 1e-30f * bn.scale);
#pragma unroll
      for (int i = 0; i < kPer; ++i) slack[i] = fmaf(fabsf(d.x[i]), c1, c0);
    } else {
#pragma unroll
      for (int i = 0; i < kPer; ++i) {
        const float rel = fmaf(fabsf(d.x[i]), 2.9e-7f, 1.6e-6f);
        const float margin = fmaf(q[i], rel, 1e-30f);
        lowq[i] = q[i] - margin;
        highq[i] = q[i] + margin;
      }
    }
  } else {
#pragma unroll
    for (int i = 0; i < kPer; ++i) q[i] = SIG ? sigmoid_as<PDT>(d.x[i]) : d.x[i];
  }
  // stage 2: guessed bin
  float below[kPer], above[kPer], off_centre[kPer];
  if (GRID) {
#pragma unroll
    for (int i = 0; i < kPer; ++i) {
      // below the first / above the last threshold (and NaN) clamp to the middle of the end bins: certain
      const float u = fminf(fmaxf(fmaf(q[i], bn.scale, bias), 0.5f), top + 0.5f);
      const float fl = floorf(u);
      g[i] = (int)fl;
      off_centre[i] = fabsf((u - fl) - 0.5f);  // 0 in the middle of a bin, 0.5 on a threshold
    }
  } else {
#pragma unroll
    for (int i = 0; i < kPer; ++i) g[i] = (int)fminf(fmaxf(fmaf(q[i], bn.scale, bias), 0.0f), top);  // NaN -> 0
#pragma unroll
    for (int i = 0; i < kPer; ++i) {
      below[i] = bn.pad[g[i]];
      above[i] = bn.pad[g[i] + 1];
    }
  }
  // stage 3: is the guess certain?
  unsigned recheck = 0;
#pragma unroll
  for (int i = 0; i < kPer; ++i) {
    unsigned sure;
    if (GRID) {
      // exact scores only carry the grid's unevenness; approximate ones add their own margin (see stage 1)
      const float allowed = kFiltered ? slack[i] : bn.grid_eps;
      sure = (unsigned)(off_centre[i] + allowed < 0.5f);
    } else if (kFiltered) {
      // certain for `>=` thresholds and for `>` thresholds alike: q lies strictly between the neighbours
      sure = (unsigned)((below[i] < lowq[i]) & (highq[i] < above[i]));
    } else {
      const unsigned lo_ok = bn.gt ? (unsigned)(below[i] < q[i]) : (unsigned)(below[i] <= q[i]);
      const unsigned hi_ok = bn.gt ? (unsigned)!(above[i] < q[i]) : (unsigned)!(above[i] <= q[i]);
      sure = (lo_ok & hi_ok) | (unsigned)((q[i] != q[i]) & (g[i] == 0));
    }
    recheck |= (sure ^ 1u) << i;
  }
  // cold path: too close to a threshold for the approximation, or a guess that missed
  if (recheck) {
#pragma unroll
    for (int i = 0; i < kPer; ++i)
      if ((recheck >> i) & 1u) g[i] = exact_bin<PDT, SIG>(bn.pad, bn.T, bn.lo, bn.scale, bn.gt, d.x[i]);
  }
}

template <int PDT, bool SIG, bool GRID>
__device__ __forceinline__ void bin_batch(const Binner& bn, LaneHist& h, const Decoded& d) {
  int g[kPer];
  batch_bins<PDT, SIG, GRID>(bn, d, g);
  // stage 4: the read-modify-writes on this lane's own bytes.  Done four at a time: the four loads are issued
  // together, a sample whose counter repeats an earlier one of the group takes that one's updated value instead
  // of its own (stale) load, and the stores go out in order so the last writer holds the complete count.  This
  // cuts the LDS->IADD->STS dependency chain from kPer links to kPer/4.
#pragma unroll
  for (int i = 0; i < kPer; i += 4) {
    const uint32_t o0 = g[i] * 64 + d.toff[i], o1 = g[i + 1] * 64 + d.toff[i + 1];
    const uint32_t o2 = g[i + 2] * 64 + d.toff[i + 2], o3 = g[i + 3] * 64 + d.toff[i + 3];
    const uint32_t v0 = h.col[o0], v1 = h.col[o1], v2 = h.col[o2], v3 = h.col[o3];
    const uint32_t u0 = v0 + d.inc[i];
    const uint32_t u1 = (o1 == o0 ? u0 : v1) + d.inc[i + 1];
    const uint32_t u2 = (o2 == o1 ? u1 : (o2 == o0 ? u0 : v2)) + d.inc[i + 2];
    const uint32_t u3 = (o3 == o2 ? u2 : (o3 == o1 ? u1 : (o3 == o0 ? u0 : v3))) + d.inc[i + 3];
    h.col[o0] = (uint8_t)u0;
    h.col[o1] = (uint8_t)u1;
    h.col[o2] = (uint8_t)u2;
    h.col[o3] = (uint8_t)u3;
  }
  h.batches += 1;
}

template <int PDT, int TDT, bool SIG>
__global__ void __launch_bounds__(512, 1) k2_binary_lanes(TbArgs a, int warps_per_cta) {
  extern __shared__ uint32_t sh_u32[];
  if (!pass_enabled(a, SIG)) return;

  // shared memory: [T + 2 padded thresholds][2*(T+1) CTA histogram words][per-warp byte counters]
  float* sh_pad = (float*)sh_u32;
  const int pad_words = (a.T + 2 + 31) / 32 * 32;
  const int rows = 2 * (a.T + 1);
  const int cta_words = (rows + 31) / 32 * 32;
  uint32_t* cta_hist = sh_u32 + pad_words;
  for (int i = threadIdx.x; i < cta_words; i += blockDim.x) cta_hist[i] = 0;
  Binner bn = make_binner(a, sh_pad, 1 << 30, nullptr);  // contains __syncthreads
  const bool on_grid = bn.grid_eps < 0.01f;  // CTA-uniform: thresholds within 1 % of a bin of an even grid
  const int warp_in_cta = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  uint8_t* warp_base = (uint8_t*)(cta_hist + cta_words) + (int64_t)warp_in_cta * rows * 32;
  LaneHist h{warp_base + lane, warp_base, cta_hist, lane, 0};
  for (int r = 0; r < rows; ++r) h.col[r * 32] = 0;
  __syncwarp();

  const int64_t n = a.n_total;
  const int64_t n_vec = n / 4;  // groups of 4 consecutive samples
  const int64_t gthreads = (int64_t)gridDim.x * blockDim.x;
  const int64_t gtid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t step = gthreads * kBatch;
  bool stop = false, bad_target = false;

  auto fetch_full = [&](int64_t b, Batch& s) {
#pragma unroll
    for (int u = 0; u < kBatch; ++u) {
      const int64_t v = b + lane + (int64_t)u * gthreads;
      load_preds4<PDT>(a.preds, v * 4, &s.x[u * 4]);
      load_targets4<TDT>(a.target, v * 4, &s.lo[u * 4], &s.hi[u * 4]);
    }
  };
  // With few warps per SM the register loads alone do not cover HBM latency x bandwidth (~90 KB per SM), and a
  // second register buffer does not help (see decode_batch).  Batches further ahead are pulled into L2 with
  // prefetch instructions, which occupy neither registers nor scoreboard slots; the register loads then only
  // have to hide L2 latency.
  auto prefetch_batch = [&](int64_t bb) {
#pragma unroll
    for (int u = 0; u < kBatch; ++u) {
      const int64_t v = bb + lane + (int64_t)u * gthreads;
      const char* pp = (const char*)a.preds + v * 4 * Elem<PDT>::kBytes;
      const char* tp = (const char*)a.target + v * 4 * label_bytes_dev(TDT);
      if ((lane & 7) == 0) prefetch_l2(pp);  // one request per 128-byte line is enough
      if (label_bytes_dev(TDT) == 8 ? (lane & 3) == 0 : (lane & 7) == 0) prefetch_l2(tp);
    }
  };
  // after a batch is binned: the pass-A flag protocol, and keeping the byte counters from overflowing
  auto after_batch = [&](const Decoded& d) {
    bad_target |= d.bad != 0u;
    if (!SIG) {
      const bool mine = d.outside != 0u;
      if (mine) raise_flag(a);
      // once anybody saw a logit the raw histogram is moot; other CTAs' flags are polled only now and then (the
      // poll is a dependent L2 round trip in every iteration otherwise)
      const bool poll = (h.batches & 7) == 0;
      stop = __any_sync(kFull, mine || (poll && flag_up(a)));
    }
    if (h.batches >= kFoldEvery) lane_fold(a, h);  // `batches` is identical in all lanes
  };
  auto process = [&](const Batch& s, unsigned live, bool full, Decoded& d) {
    if (full) {
      if (a.has_ignore) decode_batch<SIG, true, true>(a, s, 0u, d);
      else decode_batch<SIG, true, false>(a, s, 0u, d);
    } else {
      if (a.has_ignore) decode_batch<SIG, false, true>(a, s, live, d);
      else decode_batch<SIG, false, false>(a, s, live, d);
    }
  };

  // Full batches: `b` (the warp's first vector) is warp-uniform and b + 32 + (kBatch-1)*gthreads <= n_vec holds
  // for every lane, so no masking.  One load buffer: a batch is decoded (first touch), the next one is requested
  // into the same registers, then the decoded one is binned while those loads are in flight.
  Batch S;
  Decoded D;
  constexpr int kAhead = 4;  // batches ahead of the register loads that are requested into L2
  int64_t b = gtid - lane;
  const int64_t last_full = n_vec - ((int64_t)(kBatch - 1) * gthreads + 32);  // b <= last_full  <=>  batch is full
  bool have = b <= last_full;
  if (have) {
    fetch_full(b, S);
#pragma unroll
    for (int p = 1; p <= kAhead; ++p)
      if (b + p * step <= last_full) prefetch_batch(b + p * step);
  }
  while (have && !stop) {
    process(S, 0u, true, D);
    b += step;
    have = b <= last_full;
    if (have) fetch_full(b, S);
    if (b + kAhead * step <= last_full) prefetch_batch(b + kAhead * step);
    if (on_grid)
      bin_batch<PDT, SIG, true>(bn, h, D);
    else
      bin_batch<PDT, SIG, false>(bn, h, D);
    after_batch(D);
  }
  // ragged remainder: at most one masked batch per warp
  if (!stop && b < n_vec) {
    unsigned live = 0;
#pragma unroll
    for (int u = 0; u < kBatch; ++u) {
      const int64_t v = b + lane + (int64_t)u * gthreads;
      if (v < n_vec) {
        load_preds4<PDT>(a.preds, v * 4, &S.x[u * 4]);
        load_targets4<TDT>(a.target, v * 4, &S.lo[u * 4], &S.hi[u * 4]);
        live |= 0xfu << (u * 4);
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          S.x[u * 4 + e] = 0.0f;
          S.lo[u * 4 + e] = 0u;
          S.hi[u * 4 + e] = 0u;
        }
      }
    }
    process(S, live, false, D);
    bin_batch<PDT, SIG, false>(bn, h, D);
    after_batch(D);
  }
  // tail samples (n % 4) — handled by the first threads of CTA 0
  bool tail_stop = false;
  if (!stop && blockIdx.x == 0) {
    int64_t i = n_vec * 4 + threadIdx.x;
    if (i < n) {
      float x = load_elem<PDT>(a.preds, i);
      int64_t t = load_label(a.target, a.tdt, i);
      const bool ignored = a.has_ignore && t == a.ignore;
      if (!SIG && !(x >= 0.0f && x <= 1.0f) && (!ignored || a.flag_counts_ignored)) {
        tail_stop = true;
        raise_flag(a);
      }
      if (!ignored) {
        if (t != 0 && t != 1) {
          bad_target = true;
        } else {
          int k = exact_bin<PDT, SIG>(bn.pad, bn.T, bn.lo, bn.scale, bn.gt, x);
          h.col[k * 64 + (int)t * 32] += 1;
        }
      }
    }
  }
  if (bad_target) atomicOr(a.err, TMB_ERR_TARGET_BINARY);
  // finalize discards the raw histogram once the flag is up: a stopped pass A has nothing worth flushing
  const bool dead = !SIG && (stop || __any_sync(kFull, tail_stop));
  if (!dead) lane_fold(a, h);
  __syncthreads();
  if (!SIG && flag_up(a)) return;
  unsigned long long* gh = hist_base(a, SIG ? 1 : 0);
  for (int r = threadIdx.x; r < rows; r += blockDim.x) {
    const uint32_t v = cta_hist[r];
    if (v) atomicAdd(&gh[(r & 1) * (a.T + 1) + (r >> 1)], (unsigned long long)v);  // row r = k*2 + t
  }
}

// ---- multilabel binned curves: [N, L] row-major, CTA-wide shared histogram ---------------------------------------
// Same load / decode / filtered-sigmoid pipeline as k2_binary_lanes; what differs is where the counts go.  With L
// labels the lane-private byte counters would need L times the space, so the CTA keeps ONE 32-bit histogram
// [L][2][T+1] in shared memory (161 KB at L = 100, T = 200: one CTA per SM) and every counted sample is one
// shared-memory atomic.  The earlier generic kernel did the same atomics but one sample per loop trip — load,
// exact sigmoid, compare, atomic, next load — and was bound by that latency chain (27 % of the HBM roofline),
// not by the atomics.  A thread's vector of 4 consecutive elements covers 4 consecutive labels (wrapping at L);
// the label of its first element advances by a constant from batch to batch, so no division in the loop.
template <int PDT, int TDT, bool SIG>
__global__ void __launch_bounds__(512, 1) k2_multi(TbArgs a) {
  extern __shared__ uint32_t sh_u32[];
  if (!pass_enabled(a, SIG)) return;
  float* sh_pad = (float*)sh_u32;
  const int pad_words = (a.T + 2 + 31) / 32 * 32;
  const int kbins = a.T + 1;
  const int nbins = a.L * 2 * kbins;
  uint32_t* hist = sh_u32 + pad_words;
  for (int i = threadIdx.x; i < nbins; i += blockDim.x) hist[i] = 0;
  Binner bn = make_binner(a, sh_pad, 1 << 30, nullptr);  // contains __syncthreads
  const bool on_grid = bn.grid_eps < 0.01f;
  const int lane = threadIdx.x & 31;
  const int L = a.L;

  const int64_t n = a.n_total;
  const int64_t n_vec = n / 4;
  const int64_t gthreads = (int64_t)gridDim.x * blockDim.x;
  const int64_t gtid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t step = gthreads * kBatch;
  bool stop = false, bad_target = false;
  int batches = 0;
  // label of the first element of this thread's vector u of the current batch, and its advance per batch
  int lab[kBatch];
#pragma unroll
  for (int u = 0; u < kBatch; ++u) lab[u] = (int)(((gtid + (int64_t)u * gthreads) * 4) % L);
  const int dlab = (int)((step * 4) % L);

  auto fetch_full = [&](int64_t b, Batch& s) {
#pragma unroll
    for (int u = 0; u < kBatch; ++u) {
      const int64_t v = b + lane + (int64_t)u * gthreads;
      load_preds4<PDT>(a.preds, v * 4, &s.x[u * 4]);
      load_targets4<TDT>(a.target, v * 4, &s.lo[u * 4], &s.hi[u * 4]);
    }
  };
  auto prefetch_batch = [&](int64_t bb) {
#pragma unroll
    for (int u = 0; u < kBatch; ++u) {
      const int64_t v = bb + lane + (int64_t)u * gthreads;
      const char* pp = (const char*)a.preds + v * 4 * Elem<PDT>::kBytes;
      const char* tp = (const char*)a.target + v * 4 * label_bytes_dev(TDT);
      if ((lane & 7) == 0) prefetch_l2(pp);
      if (label_bytes_dev(TDT) == 8 ? (lane & 3) == 0 : (lane & 7) == 0) prefetch_l2(tp);
    }
  };
  auto process = [&](const Batch& s, unsigned live, bool full, Decoded& d) {
    if (full) {
      if (a.has_ignore) decode_batch<SIG, true, true>(a, s, 0u, d);
      else decode_batch<SIG, true, false>(a, s, 0u, d);
    } else {
      if (a.has_ignore) decode_batch<SIG, false, true>(a, s, live, d);
      else decode_batch<SIG, false, false>(a, s, live, d);
    }
  };
  // count the batch (labels of the CURRENT batch), then move the labels on to the next one
  auto count = [&](const Decoded& d, const int (&g)[kPer]) {
#pragma unroll
    for (int u = 0; u < kBatch; ++u) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int i = u * 4 + e;
        int l = lab[u] + e;
        l -= (l >= L) ? L : 0;
        if (d.inc[i]) atomicAdd(&hist[(l * 2 + (int)(d.toff[i] >> 5)) * kbins + g[i]], 1u);
      }
      lab[u] += dlab;
      lab[u] -= (lab[u] >= L) ? L : 0;
    }
  };
  auto after_batch = [&](const Decoded& d) {
    bad_target |= d.bad != 0u;
    batches += 1;
    if (!SIG) {
      const bool mine = d.outside != 0u;
      if (mine) raise_flag(a);
      const bool poll = (batches & 7) == 0;
      stop = __any_sync(kFull, mine || (poll && flag_up(a)));
    }
  };

  Batch S;
  Decoded D;
  int g[kPer];
  constexpr int kAhead = 4;
  int64_t b = gtid - lane;
  const int64_t last_full = n_vec - ((int64_t)(kBatch - 1) * gthreads + 32);
  bool have = b <= last_full;
  if (have) {
    fetch_full(b, S);
#pragma unroll
    for (int p = 1; p <= kAhead; ++p)
      if (b + p * step <= last_full) prefetch_batch(b + p * step);
  }
  while (have && !stop) {
    process(S, 0u, true, D);
    b += step;
    have = b <= last_full;
    if (have) fetch_full(b, S);
    if (b + kAhead * step <= last_full) prefetch_batch(b + kAhead * step);
    if (on_grid)
      batch_bins<PDT, SIG, true>(bn, D, g);
    else
      batch_bins<PDT, SIG, false>(bn, D, g);
    count(D, g);
    after_batch(D);
  }
  // ragged remainder: at most one masked batch per warp
  if (!stop && b < n_vec) {
    unsigned live = 0;
#pragma unroll
    for (int u = 0; u < kBatch; ++u) {
      const int64_t v = b + lane + (int64_t)u * gthreads;
      if (v < n_vec) {
        load_preds4<PDT>(a.preds, v * 4, &S.x[u * 4]);
        load_targets4<TDT>(a.target, v * 4, &S.lo[u * 4], &S.hi[u * 4]);
        live |= 0xfu << (u * 4);
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          S.x[u * 4 + e] = 0.0f;
          S.lo[u * 4 + e] = 0u;
          S.hi[u * 4 + e] = 0u;
        }
      }
    }
    process(S, live, false, D);
    batch_bins<PDT, SIG, false>(bn, D, g);
    count(D, g);
    after_batch(D);
  }
  // tail elements (n % 4) — handled by the first threads of CTA 0
  bool tail_stop = false;
  if (!stop && blockIdx.x == 0) {
    const int64_t i = n_vec * 4 + threadIdx.x;
    if (i < n) {
      const float x = load_elem<PDT>(a.preds, i);
      const int64_t t = load_label(a.target, a.tdt, i);
      const bool ignored = a.has_ignore && t == a.ignore;
      if (!SIG && !(x >= 0.0f && x <= 1.0f) && (!ignored || a.flag_counts_ignored)) {
        tail_stop = true;
        raise_flag(a);
      }
      if (!ignored) {
        if (t != 0 && t != 1) {
          bad_target = true;
        } else {
          const int k = exact_bin<PDT, SIG>(bn.pad, bn.T, bn.lo, bn.scale, bn.gt, x);
          atomicAdd(&hist[((int)(i % L) * 2 + (int)t) * kbins + k], 1u);
        }
      }
    }
  }
  if (bad_target) atomicOr(a.err, TMB_ERR_TARGET_BINARY);
  const bool dead = !SIG && __syncthreads_or(stop || tail_stop);
  if (SIG) __syncthreads();
  if (dead || (!SIG && flag_up(a))) return;  // finalize discards the raw histogram once the flag is up
  unsigned long long* gh = hist_base(a, SIG ? 1 : 0);
  for (int i = threadIdx.x; i < nbins; i += blockDim.x) {
    const uint32_t v = hist[i];
    if (v) atomicAdd(&gh[i], (unsigned long long)v);
  }
}

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

// ---- multiclass binned curves, C <= 32: one THREAD per row ------------------------------------------------------
// For up to 32 classes torch's softmax layout gives every lane one element (W = next_pow2(C) lanes per row), so its
// arithmetic is: max over the row, e_c = expf(x_c - max), and the xor-butterfly sum — which is the balanced tree
// "s[i] += s[i + W/2], s[i] += s[i + W/4], ..." over the W values (zeros past C; float addition is commutative, so
// every lane of the butterfly ends with this very tree).  A single thread holding the row reproduces that bit for
// bit with W - 1 adds and no shuffles; the sub-warp version spends ~8 shuffle steps per row on one element per
// lane.  Rows come through shared memory (coalesced loads, odd row stride: conflict-free reads).
constexpr int kSmallRows = 256;  // rows per CTA tile = threads per CTA

template <int PDT, bool SOFT, int W>
__global__ void __launch_bounds__(kSmallRows) k2_mc_small(TbArgs a, int64_t rows, int C, int micro, int sh_bins,
                                                          int low_k) {
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

// ---- K1b fast path: ONE threshold, [N, L] row-major (binary = L 1) ---------------------------------------------
// Stat scores / confusion matrices / accuracy / F1 / Jaccard of binary and multilabel tasks compare against a
// single threshold, so a sample only needs "above or not" — four counters per label.  The grid-stride is a
// multiple of L vectors of 4 samples, which pins each of a thread's 4 vector slots to ONE label for the whole
// kernel: the counters live in registers (packed bytes, widened every 255 iterations) and reach memory through
// 16 atomics per thread at the very end.  No shared memory, no per-sample atomics; HBM-bound.
template <int PDT, int TDT, bool SIG>
__global__ void __launch_bounds__(256) k2_single(TbArgs a, int64_t G, int sh_bins) {
  extern __shared__ uint32_t sh_cnt[];  // [L * 4] when sh_bins > 0
  if (!pass_enabled(a, SIG)) return;
  for (int i = threadIdx.x; i < sh_bins; i += blockDim.x) sh_cnt[i] = 0;
  __syncthreads();
  const int64_t gtid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const float thr = a.thr[0];
  const bool gt = a.cmp_gt != 0;
  const int64_t n_vec = a.n_total / 4;
  const uint32_t ign_lo = (uint32_t)a.ignore, ign_hi = (uint32_t)((uint64_t)a.ignore >> 32);
  constexpr bool kFiltered = SIG && PDT == TMB_F32;
  unsigned long long* gh = hist_base(a, SIG ? 1 : 0);
  bool bad_target = false, stop = false;

  if (gtid < G) {
    uint32_t wide[4][4];  // [slot][target * 2 + above]
    uint32_t packed[4];   // the same four counters as bytes, widened before they can overflow
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      packed[e] = 0;
#pragma unroll
      for (int f = 0; f < 4; ++f) wide[e][f] = 0;
    }
    auto widen = [&]() {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
#pragma unroll
        for (int f = 0; f < 4; ++f) wide[e][f] += (packed[e] >> (8 * f)) & 0xffu;
        packed[e] = 0;
      }
    };
    auto one = [&](float x, uint32_t lo, uint32_t hi, int e) -> bool {  // returns "outside [0,1]" for pass A
      const bool kept = !(a.has_ignore && lo == ign_lo && hi == ign_hi);
      const bool zero_or_one = (hi | (lo >> 1)) == 0u;
      bad_target |= kept && !zero_or_one;
      bool above;
      if (kFiltered) {
        const float q = rcp_approx(1.0f + ex2_approx(x * -1.4426950408889634f));
        const float margin = fmaf(q, fmaf(fabsf(x), 2.9e-7f, 1.6e-6f), 1e-30f);  // see the filtered sigmoid above
        above = thr < q;
        if (!(fabsf(q - thr) > margin)) {  // too close to call (or NaN): exact expression
          const float qe = sigmoid_as<PDT>(x);
          above = gt ? (thr < qe) : (thr <= qe);
        }
      } else {
        const float q = SIG ? sigmoid_as<PDT>(x) : x;
        above = gt ? (thr < q) : (thr <= q);
      }
      const uint32_t field = (lo & 1u) * 2u + (above ? 1u : 0u);
      packed[e] += (kept && zero_or_one) ? (1u << (8 * field)) : 0u;
      return !SIG && !(x >= 0.0f && x <= 1.0f) && (kept || a.flag_counts_ignored);
    };
    int since_widen = 0;
    constexpr int U = 2;
    for (int64_t v0 = gtid; v0 < n_vec && !stop; v0 += G * U) {
      float x[U][4];
      uint32_t lo[U][4], hi[U][4];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t v = v0 + (int64_t)u * G;
        if (v < n_vec) {
          load_preds4<PDT>(a.preds, v * 4, x[u]);
          load_targets4<TDT>(a.target, v * 4, lo[u], hi[u]);
        }
      }
      bool outside = false;
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (v0 + (int64_t)u * G < n_vec) {
#pragma unroll
          for (int e = 0; e < 4; ++e) outside |= one(x[u][e], lo[u][e], hi[u][e], e);
        }
      }
      since_widen += U;
      if (since_widen > 255 - U) {
        widen();
        since_widen = 0;
        if (!SIG && flag_up(a)) stop = true;  // another CTA saw a logit: the raw counts are moot
      }
      if (outside) {
        raise_flag(a);
        stop = true;
      }
    }
    widen();
    if (!(!SIG && stop)) {
      // slot e of this thread always saw label (4 * gtid + e) mod L  (G is a multiple of L).  Counters meet in a
      // CTA-wide shared histogram first when it fits: with L = 1 every thread of the grid would otherwise send
      // its 16 atomics to the same four addresses.
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int l = (int)((4 * gtid + e) % a.L);
#pragma unroll
        for (int f = 0; f < 4; ++f) {
          if (wide[e][f]) {
            if (sh_bins > 0)
              atomicAdd(&sh_cnt[l * 4 + f], wide[e][f]);
            else
              atomicAdd(&gh[(int64_t)l * 4 + f], (unsigned long long)wide[e][f]);  // [l][t][k], T+1 = 2
          }
        }
      }
    }
  }
  if (sh_bins > 0) {
    __syncthreads();
    for (int i = threadIdx.x; i < sh_bins; i += blockDim.x) {
      const uint32_t v = sh_cnt[i];
      if (v) atomicAdd(&gh[i], (unsigned long long)v);
    }
  }
  // tail samples (n % 4): a handful, straight to the global counters
  if (blockIdx.x == 0 && !stop) {
    const int64_t i = n_vec * 4 + threadIdx.x;
    if (i < a.n_total) {
      const float x = load_elem<PDT>(a.preds, i);
      const int64_t t = load_label(a.target, a.tdt, i);
      const bool kept = !(a.has_ignore && t == a.ignore);
      if (!SIG && !(x >= 0.0f && x <= 1.0f) && (kept || a.flag_counts_ignored)) raise_flag(a);
      if (kept) {
        if (t != 0 && t != 1) {
          bad_target = true;
        } else {
          const float q = SIG ? sigmoid_as<PDT>(x) : x;
          const int above = gt ? (thr < q) : (thr <= q);
          atomicAdd(&gh[(i % a.L) * 4 + t * 2 + above], 1ULL);
        }
      }
    }
  }
  if (bad_target) atomicOr(a.err, TMB_ERR_TARGET_BINARY);
}

// ---- generic path: any L / S / T, atomics into shared (small) or global (large) histograms ------------
template <int PDT, bool SIG, bool LABELS>
__global__ void __launch_bounds__(1024) k2_generic(TbArgs a, int sh_bins, int sh_pad_cap, const float* global_pad) {
  extern __shared__ uint32_t sh_u32[];
  if (!LABELS) {
    if (!pass_enabled(a, SIG)) return;
  }
  float* sh_pad = (float*)sh_u32;
  int* sh_hist = (int*)(sh_u32 + sh_pad_cap);
  Binner bn{};
  if (!LABELS) bn = make_binner(a, sh_pad, sh_pad_cap, global_pad);
  for (int i = threadIdx.x; i < sh_bins; i += blockDim.x) sh_hist[i] = 0;
  __syncthreads();

  unsigned long long* gh = hist_base(a, SIG ? 1 : 0);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  // label / inner position of element i0, then advanced incrementally (no per-element division)
  int64_t s = i0 % a.S;
  int l = (int)((i0 / a.S) % a.L);
  const int64_t ds = stride % a.S;
  const int dl = (int)((stride / a.S) % a.L);
  const int kbins = a.T + 1;
  bool stop = false;

  for (int64_t i = i0; i < a.n_total; i += stride) {
    int64_t t = load_label(a.target, a.tdt, i);
    bool ignored = a.has_ignore && t == a.ignore;
    int k = -1;
    if (LABELS) {
      int64_t p = load_label(a.preds, a.pdt, i);
      if (!ignored) {
        if (p != 0 && p != 1) atomicOr(a.err, TMB_ERR_PREDS_RANGE);
        k = (p == 1) ? 1 : 0;
        // reference semantics for stray labels (stat_scores.py:127-130): only p == target counts as a match
        if (p != 0 && p != 1) k = (t == 1) ? 0 : 1;
      }
    } else {
      float x = load_elem<PDT>(a.preds, i);
      if (!SIG) {
        bool outside = !(x >= 0.0f && x <= 1.0f);
        if (outside && (!ignored || a.flag_counts_ignored)) {
          raise_flag(a);
          stop = true;
        }
      }
      if (!ignored && !stop) k = bn.bin(SIG ? sigmoid_as<PDT>(x) : x);
    }
    if (!SIG && !LABELS) {
      if (__any_sync(__activemask(), stop)) break;
    }
    if (k >= 0) {
      if (t != 0 && t != 1) {
        atomicOr(a.err, TMB_ERR_TARGET_BINARY);
      } else {
        int64_t bin = ((int64_t)l * 2 + t) * kbins + k;
        if (sh_bins > 0)
          atomicAdd(&sh_hist[bin], 1);
        else
          atomicAdd(&gh[bin], 1ULL);
      }
    }
    // advance (l, s) by `stride` elements
    s += ds;
    int carry = s >= a.S;
    s -= carry ? a.S : 0;
    l += dl + carry;
    if (l >= a.L) l -= a.L;
    if (l >= a.L) l -= a.L;
  }
  if (!SIG && !LABELS) {
    if (__syncthreads_or(stop) || flag_up(a)) return;
  } else {
    __syncthreads();
  }
  for (int i = threadIdx.x; i < sh_bins; i += blockDim.x) {
    int v = sh_hist[i];
    if (v) atomicAdd(&gh[i], (unsigned long long)v);
  }
}

// padded threshold copy in global memory for threshold counts that do not fit shared memory
__global__ void k2_pad_thresholds(const float* __restrict__ thr, int T, float* __restrict__ pad) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < T + 2; i += gridDim.x * blockDim.x)
    pad[i] = (i == 0) ? -INFINITY : (i == T + 1 ? INFINITY : thr[i - 1]);
}

// ---- finalize -------------------------------------------------------------------------------------------
// One warp per (label, target class).  The per-threshold counts are suffix sums of the bin histogram:
// above(j) = #{samples with q >= thr[j]} = sum_{k > j} hist[k].  Lanes take 32 bins at a time from the
// top, a shuffle scan forms the running suffix sum, and the counts are ADDED into the state.
__global__ void __launch_bounds__(256) k2_finalize(unsigned long long* scratch, int L, int T,
                                                   const int* __restrict__ perm, int layout, long long* out0,
                                                   long long* out1, long long* out2, long long* out3) {
  const int variant = scratch[0] != 0ULL ? 1 : 0;
  const int kbins = T + 1;
  const unsigned long long* h = scratch + kHdr + (int64_t)variant * L * 2 * kbins;
  const int lane = threadIdx.x & 31;
  const int job = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (job >= L * 2) return;
  const int l = job >> 1, t = job & 1;
  const unsigned long long* hc = h + ((int64_t)l * 2 + t) * kbins;
  if (layout == TMB_TB_CURVE) {
    long long total = 0;
    for (int k = lane; k <= T; k += 32) total += (long long)hc[k];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) total += __shfl_xor_sync(kFull, total, d);
    long long carry = 0;  // sum of the bins above the current chunk
    for (int top = T; top >= 1; top -= 32) {
      const int k = top - lane;  // lane 0 owns the highest bin of the chunk
      long long v = (k >= 1) ? (long long)hc[k] : 0;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        long long o = __shfl_up_sync(kFull, v, d);
        if (lane >= d) v += o;
      }
      const long long above = carry + v;  // sum_{k' >= k} hist[k']  ==  count for threshold j = k - 1
      if (k >= 1) {
        const int j = k - 1;
        const int row = perm ? perm[j] : j;
        long long* cell = out0 + ((int64_t)row * L + l) * 4 + t * 2;  // [row][l][t][pred]
        cell[1] += above;
        cell[0] += total - above;
      }
      carry += __shfl_sync(kFull, v, 31);
    }
  } else if (lane == 0) {
    if (layout == TMB_TB_CONFMAT) {
      long long* cell = out0 + (int64_t)l * 4 + t * 2;
      cell[0] += (long long)hc[0];
      cell[1] += (long long)hc[1];
    } else if (t == 1) {
      out0[l] += (long long)hc[1];  // tp
      out3[l] += (long long)hc[0];  // fn
    } else {
      out1[l] += (long long)hc[1];  // fp
      out2[l] += (long long)hc[0];  // tn
    }
  }
}

__global__ void __launch_bounds__(256) k2_clear(unsigned long long* scratch, int64_t scratch_elems) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < scratch_elems;
       i += (int64_t)gridDim.x * blockDim.x)
    scratch[i] = 0ULL;
}

// ---- utilities ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_bincount(const void* __restrict__ x, int dt, int64_t n, int64_t length,
                                                  unsigned long long* out, int* err) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t v = load_label(x, dt, i);
    if (v < 0 || v >= length)
      atomicOr(err, TMB_ERR_PREDS_RANGE);
    else
      atomicAdd(&out[v], 1ULL);
  }
}

__global__ void __launch_bounds__(256) k_label_max(const void* __restrict__ x, int dt, int64_t n, long long* out) {
  long long m = LLONG_MIN;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    long long v = load_label(x, dt, i);
    m = v > m ? v : m;
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    long long o = __shfl_xor_sync(kFull, m, d);
    m = o > m ? o : m;
  }
  if ((threadIdx.x & 31) == 0) atomicMax(out, m);
}

__global__ void k_set_i64(long long* p, long long v) { *p = v; }

// ---- host ------------------------------------------------------------------------------------------------
// scratch layout (int64 elements): [0, kHdr) header | 2 histogram variants | padded thresholds (as floats)
static int64_t hist_elems(int64_t L, int64_t T) { return 2 * L * 2 * (T + 1); }
static int64_t pad_elems(int64_t T) { return (T + 2 + 1) / 2; }
static int64_t scratch_elems_for(int64_t L, int64_t T) { return kHdr + hist_elems(L, T) + pad_elems(T); }

template <int PDT, int TDT>
static int launch_binary_lanes(const TbArgs& a, int warps, int smem, cudaStream_t st) {
  auto ka = k2_binary_lanes<PDT, TDT, false>;
  auto kb = k2_binary_lanes<PDT, TDT, true>;
  cudaFuncSetAttribute(ka, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(kb, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int block = warps * 32;
  // persistent grid: as many CTAs as stay resident (1 per SM at T = 200, more when the histograms are small)
  const int64_t work = (a.n_total / 4 + (int64_t)block * kBatch - 1) / ((int64_t)block * kBatch);
  const int grid = resident_grid(kb, block, smem, work);
  ka<<<grid, block, smem, st>>>(a, warps);
  if (int rc = launch_status()) return rc;
  kb<<<grid, block, smem, st>>>(a, warps);
  return launch_status();
}

template <int PDT>
static int launch_binary_lanes_t(const TbArgs& a, int warps, int smem, cudaStream_t st) {
  switch (a.tdt) {
    case TMB_I64: return launch_binary_lanes<PDT, TMB_I64>(a, warps, smem, st);
    case TMB_I32: return launch_binary_lanes<PDT, TMB_I32>(a, warps, smem, st);
    case TMB_U8: return launch_binary_lanes<PDT, TMB_U8>(a, warps, smem, st);
    default: return launch_binary_lanes<PDT, TMB_I8>(a, warps, smem, st);
  }
}

template <int PDT, int TDT>
static int launch_multi_pt(const TbArgs& a, int smem, cudaStream_t st) {
  auto ka = k2_multi<PDT, TDT, false>;
  auto kb = k2_multi<PDT, TDT, true>;
  cudaFuncSetAttribute(ka, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(kb, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int block = 512;
  const int64_t work = (a.n_total / 4 + (int64_t)block * kBatch - 1) / ((int64_t)block * kBatch);
  const int grid = resident_grid(kb, block, smem, work);
  ka<<<grid, block, smem, st>>>(a);
  if (int rc = launch_status()) return rc;
  kb<<<grid, block, smem, st>>>(a);
  return launch_status();
}

template <int PDT>
static int launch_multi_p(const TbArgs& a, int smem, cudaStream_t st) {
  switch (a.tdt) {
    case TMB_I64: return launch_multi_pt<PDT, TMB_I64>(a, smem, st);
    case TMB_I32: return launch_multi_pt<PDT, TMB_I32>(a, smem, st);
    case TMB_U8: return launch_multi_pt<PDT, TMB_U8>(a, smem, st);
    default: return launch_multi_pt<PDT, TMB_I8>(a, smem, st);
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
  if (C <= 32 && low_k > 0) {  // one thread per row (see k2_mc_small)
    const int small_bins = (int)(pairs * low_k);
    const int hist_words = (small_bins + 31) / 32 * 32;
    const int small_smem = (int)pad_bytes + hist_words * 4 + kSmallRows * (C | 1) * 4;
    if (small_smem <= dev_info().max_smem_optin - 1024) {
      const int64_t work = (rows + kSmallRows - 1) / kSmallRows;
#define TMB_LAUNCH_SMALL(W_)                                                                    \
  {                                                                                             \
    auto ka = k2_mc_small<PDT, false, W_>;                                                      \
    auto kb = k2_mc_small<PDT, true, W_>;                                                       \
    cudaFuncSetAttribute(ka, cudaFuncAttributeMaxDynamicSharedMemorySize, small_smem);          \
    cudaFuncSetAttribute(kb, cudaFuncAttributeMaxDynamicSharedMemorySize, small_smem);          \
    const int grid = resident_grid(kb, kSmallRows, small_smem, work);                           \
    ka<<<grid, kSmallRows, small_smem, st>>>(a, rows, C, micro, small_bins, low_k);                \
    rc = launch_status();                                                                       \
    if (rc == 0) {                                                                              \
      kb<<<grid, kSmallRows, small_smem, st>>>(a, rows, C, micro, small_bins, low_k);              \
      rc = launch_status();                                                                     \
    }                                                                                           \
  }
      switch (rm.W) {
        case 2: TMB_LAUNCH_SMALL(2) break;
        case 4: TMB_LAUNCH_SMALL(4) break;
        case 8: TMB_LAUNCH_SMALL(8) break;
        case 16: TMB_LAUNCH_SMALL(16) break;
        default: TMB_LAUNCH_SMALL(32) break;
      }
#undef TMB_LAUNCH_SMALL
      return rc;
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

template <int PDT, int TDT>
static int launch_single_pt(const TbArgs& a, cudaStream_t st) {
  auto ka = k2_single<PDT, TDT, false>;
  auto kb = k2_single<PDT, TDT, true>;
  const int block = 256;
  const int64_t n_vec = a.n_total / 4;
  const int sh_bins = (int64_t)a.L * 4 <= 8192 ? a.L * 4 : 0;
  const int smem = sh_bins * 4;
  const int grid = resident_grid(kb, block, smem, (n_vec + block * 2 - 1) / (block * 2));
  const int64_t threads = (int64_t)grid * block;
  const int64_t G = threads / a.L * a.L;  // threads taking part: a multiple of L, so vector slots keep their label
  if (G < 1) return TMB_E_LIMIT;          // more labels than threads: let the generic path do it
  ka<<<grid, block, smem, st>>>(a, G, sh_bins);
  if (int rc = launch_status()) return rc;
  kb<<<grid, block, smem, st>>>(a, G, sh_bins);
  return launch_status();
}

template <int PDT>
static int launch_single_p(const TbArgs& a, cudaStream_t st) {
  switch (a.tdt) {
    case TMB_I64: return launch_single_pt<PDT, TMB_I64>(a, st);
    case TMB_I32: return launch_single_pt<PDT, TMB_I32>(a, st);
    case TMB_U8: return launch_single_pt<PDT, TMB_U8>(a, st);
    default: return launch_single_pt<PDT, TMB_I8>(a, st);
  }
}

static int launch_single(const TbArgs& a, cudaStream_t st) {
  switch (a.pdt) {
    case TMB_F32: return launch_single_p<TMB_F32>(a, st);
    case TMB_F16: return launch_single_p<TMB_F16>(a, st);
    default: return launch_single_p<TMB_BF16>(a, st);
  }
}

template <int PDT>
static int launch_generic(const TbArgs& a, cudaStream_t st) {
  const int64_t bins = (int64_t)a.L * 2 * (a.T + 1);
  const int sh_pad_cap = a.T + 2 <= 4096 ? (a.T + 2 + 31) / 32 * 32 : 0;
  // privatise the histogram in shared memory whenever it fits the opt-in limit (multilabel L = 100, T = 200 is
  // 40 200 bins = 157 KB: one big CTA per SM beats 2e8 global atomics)
  const int64_t sh_room = ((int64_t)dev_info().max_smem_optin - 2048) / 4 - sh_pad_cap;
  const int sh_bins = bins <= sh_room ? (int)bins : 0;
  const int smem = (sh_pad_cap + sh_bins) * 4;
  const int block = smem > 48 * 1024 ? 1024 : 256;
  const int64_t work = (a.n_total + block - 1) / block;
  float* global_pad = (float*)(a.scratch + kHdr + hist_elems(a.L, a.T));
  if (a.preds_is_label) {
    // labels are never sigmoided: they go to histogram variant 0, the one finalize reads while the flag is down
    auto k = k2_generic<TMB_F32, false, true>;
    if (smem > 48 * 1024) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    k<<<resident_grid(k, block, smem, work), block, smem, st>>>(a, sh_bins, sh_pad_cap, global_pad);
    return launch_status();
  }
  if (sh_pad_cap == 0) {
    k2_pad_thresholds<<<wave_grid((a.T + 2 + 255) / 256, 2), 256, 0, st>>>(a.thr, a.T, global_pad);
    if (int rc = launch_status()) return rc;
  }
  auto ka = k2_generic<PDT, false, false>;
  auto kb = k2_generic<PDT, true, false>;
  if (smem > 48 * 1024) {
    cudaFuncSetAttribute(ka, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaFuncSetAttribute(kb, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  }
  ka<<<resident_grid(ka, block, smem, work), block, smem, st>>>(a, sh_bins, sh_pad_cap, global_pad);
  if (int rc = launch_status()) return rc;
  kb<<<resident_grid(kb, block, smem, work), block, smem, st>>>(a, sh_bins, sh_pad_cap, global_pad);
  return launch_status();
}

}  // namespace tmb

using namespace tmb;

extern "C" int64_t tmb_thresh_bin_scratch_elems(int64_t num_labels, int64_t num_thresholds) {
  if (num_labels < 1 || num_thresholds < 1) return 0;
  return scratch_elems_for(num_labels, num_thresholds);
}

extern "C" int tmb_thresh_bin_update(const void* preds, int preds_dtype, int preds_is_label, const void* target,
                                     int target_dtype, int64_t n_outer, int64_t num_labels, int64_t n_inner,
                                     const float* thr, int64_t num_thresholds, int cmp_gt, int has_ignore,
                                     int64_t ignore_index, int flag_counts_ignored, int64_t* scratch,
                                     int64_t scratch_elems, int32_t* err, void* stream) {
  if (label_bytes(target_dtype) == 0) return TMB_E_DTYPE;
  if (preds_is_label ? label_bytes(preds_dtype) == 0 : float_bytes(preds_dtype) == 0) return TMB_E_DTYPE;
  if (n_outer < 0 || num_labels < 1 || n_inner < 1 || num_labels > (1 << 24)) return TMB_E_SHAPE;
  if (num_thresholds < 1 || num_thresholds > (1 << 26)) return TMB_E_SHAPE;
  if (!scratch || !err || !thr) return TMB_E_NULL;
  if (scratch_elems < scratch_elems_for(num_labels, num_thresholds)) return TMB_E_SCRATCH;
  const int64_t n_total = n_outer * num_labels * n_inner;
  if (n_total == 0) return 0;
  if (!preds || !target) return TMB_E_NULL;
  cudaStream_t st = (cudaStream_t)stream;

  TbArgs a;
  a.preds = preds;
  a.target = target;
  a.thr = thr;
  a.scratch = (unsigned long long*)scratch;
  a.err = err;
  a.n_total = n_total;
  a.S = n_inner;
  a.ignore = ignore_index;
  a.L = (int)num_labels;
  a.T = (int)num_thresholds;
  a.pdt = preds_dtype;
  a.tdt = target_dtype;
  a.preds_is_label = preds_is_label;
  a.cmp_gt = cmp_gt;
  a.has_ignore = has_ignore;
  a.flag_counts_ignored = flag_counts_ignored;

  // K1b fast path: one threshold, labels along the fastest axis (binary, or multilabel [N, L])
  if (num_thresholds == 1 && n_inner == 1 && !preds_is_label && target_dtype != TMB_I16 &&
      ((((uintptr_t)preds | (uintptr_t)target) & 15) == 0)) {
    if (int rc = launch_single(a, st); rc != TMB_E_LIMIT) return rc;
  }
  // binary fast path: lane-private histograms need (T+1)*128 B per warp plus the padded thresholds
  const bool aligned = (((uintptr_t)preds | (uintptr_t)target) & 15) == 0;
  const bool tdt_ok = target_dtype != TMB_I16;
  if (num_labels == 1 && !preds_is_label && aligned && tdt_ok) {
    // shared memory: padded thresholds + CTA histogram words + (T+1)*64 bytes of lane-private counters per warp
    const int pad_bytes = (a.T + 2 + 31) / 32 * 32 * 4;
    const int cta_bytes = (2 * (a.T + 1) + 31) / 32 * 32 * 4;
    const int64_t per_warp = (int64_t)(a.T + 1) * 64;
    const int64_t avail = (int64_t)dev_info().max_smem_optin - pad_bytes - cta_bytes - 1024;
    int warps = (int)(avail / per_warp);
    if (warps >= 1) {
      if (warps > 16) warps = 16;
      const int smem = pad_bytes + cta_bytes + (int)(warps * per_warp);
      switch (preds_dtype) {
        case TMB_F32: return launch_binary_lanes_t<TMB_F32>(a, warps, smem, st);
        case TMB_F16: return launch_binary_lanes_t<TMB_F16>(a, warps, smem, st);
        default: return launch_binary_lanes_t<TMB_BF16>(a, warps, smem, st);
      }
    }
  }
  // multilabel [N, L] curves: CTA-wide shared histogram, batched loads (needs the histogram to fit one SM)
  if (num_labels >= 4 && n_inner == 1 && !preds_is_label && aligned && tdt_ok) {
    const int64_t smem = ((int64_t)(a.T + 2 + 31) / 32 * 32 + (int64_t)a.L * 2 * (a.T + 1)) * 4;
    if (smem <= (int64_t)dev_info().max_smem_optin - 1024) {
      switch (preds_dtype) {
        case TMB_F32: return launch_multi_p<TMB_F32>(a, (int)smem, st);
        case TMB_F16: return launch_multi_p<TMB_F16>(a, (int)smem, st);
        default: return launch_multi_p<TMB_BF16>(a, (int)smem, st);
      }
    }
  }
  if (preds_is_label) return launch_generic<TMB_F32>(a, st);
  switch (preds_dtype) {
    case TMB_F32: return launch_generic<TMB_F32>(a, st);
    case TMB_F16: return launch_generic<TMB_F16>(a, st);
    default: return launch_generic<TMB_BF16>(a, st);
  }
}

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

extern "C" int tmb_thresh_bin_finalize(int64_t* scratch, int64_t scratch_elems, int64_t num_labels,
                                       int64_t num_thresholds, const int32_t* perm, int layout, int64_t* out0,
                                       int64_t* out1, int64_t* out2, int64_t* out3, void* stream) {
  if (!scratch || !out0) return TMB_E_NULL;
  if (layout != TMB_TB_CURVE && layout != TMB_TB_CONFMAT && layout != TMB_TB_STATS) return TMB_E_MODE;
  if (layout != TMB_TB_CURVE && num_thresholds != 1) return TMB_E_SHAPE;
  if (layout == TMB_TB_STATS && (!out1 || !out2 || !out3)) return TMB_E_NULL;
  if (num_labels < 1 || num_thresholds < 1) return TMB_E_SHAPE;
  if (scratch_elems < scratch_elems_for(num_labels, num_thresholds)) return TMB_E_SCRATCH;
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t threads = num_labels * 2 * 32;
  k2_finalize<<<(int)((threads + 255) / 256), 256, 0, st>>>((unsigned long long*)scratch, (int)num_labels,
                                                            (int)num_thresholds, perm, layout, (long long*)out0,
                                                            (long long*)out1, (long long*)out2, (long long*)out3);
  if (int rc = launch_status()) return rc;
  const int64_t used = scratch_elems_for(num_labels, num_thresholds);
  k2_clear<<<wave_grid((used + 255) / 256, 4), 256, 0, st>>>((unsigned long long*)scratch, used);
  return launch_status();
}

extern "C" int tmb_bincount(const void* x, int x_dtype, int64_t n, int64_t length, int64_t* out, int32_t* err,
                            void* stream) {
  if (label_bytes(x_dtype) == 0) return TMB_E_DTYPE;
  if (n < 0 || length < 0) return TMB_E_SHAPE;
  if (n == 0) return 0;
  if (!x || !out || !err) return TMB_E_NULL;
  const int grid = wave_grid((n + 255) / 256, 8);
  k_bincount<<<grid, 256, 0, (cudaStream_t)stream>>>(x, x_dtype, n, length, (unsigned long long*)out, err);
  return launch_status();
}

extern "C" int tmb_label_max(const void* x, int x_dtype, int64_t n, int64_t* out_max, void* stream) {
  if (label_bytes(x_dtype) == 0) return TMB_E_DTYPE;
  if (n < 0) return TMB_E_SHAPE;
  if (!out_max) return TMB_E_NULL;
  cudaStream_t st = (cudaStream_t)stream;
  k_set_i64<<<1, 1, 0, st>>>((long long*)out_max, LLONG_MIN);
  if (int rc = launch_status()) return rc;
  if (n == 0) return 0;
  if (!x) return TMB_E_NULL;
  k_label_max<<<wave_grid((n + 255) / 256, 8), 256, 0, st>>>(x, x_dtype, n, (long long*)out_max);
  return launch_status();
}

// ---- binned binary curve update for callers that cannot allocate device memory -------------------------------
//
// Everything above takes its histogram scratch from the caller.  Some callers have nowhere to get it from:
//   * a host that binds this library over a bare FFI (ctypes without torch, cgo, JNI) and owns no device
//     allocator — it holds the input and state pointers somebody else gave it and nothing more;
//   * a service that hands every byte of device memory to one pool at start-up, so that a later cudaMalloc
//     either fails or fragments what the pool thinks it owns;
//   * an evaluation loop with a latency budget: cudaMalloc / cudaFree synchronise the whole device.
// For them the scratch lives in the library image itself: one statically sized __device__ array per GPU that
// loads the library (about 36 KB, held until the process exits).  The price is a hard cap on the number of
// thresholds and the loss of re-entrancy — see the notes on the entry point.
constexpr int kNoAllocMaxThresholds = 1024;
constexpr int64_t kNoAllocScratchElems = kHdr + 2 * 2 * ((int64_t)kNoAllocMaxThresholds + 1) +
                                         ((int64_t)kNoAllocMaxThresholds + 3) / 2;  // scratch_elems_for(1, 1024)
static __device__ __align__(256) unsigned long long g_noalloc_scratch[kNoAllocScratchElems];

// `g_noalloc_scratch` in HOST code is only a handle for the runtime, not a device address; the address a
// kernel can use has to be asked for, once per device, while that device is current.
static int noalloc_scratch_address(unsigned long long** out) {
  static unsigned long long* address[64] = {};
  static bool known[64] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= 64) return TMB_E_LIMIT;
  if (!known[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess) return (int)e;
    if (p == nullptr) return (int)cudaErrorInvalidSymbol;
    address[dev] = (unsigned long long*)p;
    known[dev] = true;
  }
  *out = address[dev];
  return 0;
}

extern "C" int tmb_binary_curve_update_noalloc(const void* preds, int preds_dtype, const void* target,
                                               int target_dtype, int64_t n, const float* thr,
                                               int64_t num_thresholds, int has_ignore, int64_t ignore_index,
                                               int64_t* confmat, int32_t* err, void* stream) {
  if (num_thresholds < 1) return TMB_E_SHAPE;
  if (num_thresholds > kNoAllocMaxThresholds) return TMB_E_LIMIT;  // the static array is sized for 1024
  if (!confmat || !err || !thr) return TMB_E_NULL;
  if (scratch_elems_for(1, num_thresholds) > kNoAllocScratchElems) return TMB_E_LIMIT;
  unsigned long long* scratch = nullptr;
  if (int rc = noalloc_scratch_address(&scratch)) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  // The array still holds whatever the previous call (or a failed one) left behind: the outside-[0,1] flag and
  // both histograms must start at zero, so clear before use rather than trusting the last finalize.
  k2_clear<<<wave_grid((kNoAllocScratchElems + 255) / 256, 1), 256, 0, st>>>(scratch, kNoAllocScratchElems);
  if (int rc = launch_status()) return rc;
  if (int rc = tmb_thresh_bin_update(preds, preds_dtype, 0, target, target_dtype, n, 1, 1, thr, num_thresholds,
                                     /*cmp_gt=*/0, has_ignore, ignore_index, /*flag_counts_ignored=*/0,
                                     (int64_t*)scratch, kNoAllocScratchElems, err, stream))
    return rc;
  return tmb_thresh_bin_finalize((int64_t*)scratch, kNoAllocScratchElems, 1, num_thresholds, nullptr, TMB_TB_CURVE,
                                 confmat, nullptr, nullptr, nullptr, stream);
}
