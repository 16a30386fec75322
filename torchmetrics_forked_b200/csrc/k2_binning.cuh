// Shared by the threshold-binning translation units (k2_thresh_bin.cu: binary / multilabel / generic layouts,
// k2_mc_curve.cu: multiclass rows with the fused softmax): arguments, the threshold binner, the batch binning code.
#pragma once
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

// The "saw a value outside [0,1]" flag lives in scratch[0].  It is read and written with device-scope relaxed
// accesses, and a CTA touches it as rarely as it can: `volatile` accesses compile to system-scope ones
// (LDG/STG.STRONG.SYS), and when every warp of the grid stored to and then loaded from that one word — which is what
// logits make pass A do: each warp finds an outside value in its first batch — the word's L2 slice serialised ~5000
// system-scope requests and pass A took 65-90 us instead of ~10 (ncu: 69 % of its stall samples on the compare after
// that load).  Now a CTA stores at most once (tb_cta_flag remembers), and once it has raised or seen the flag it
// never loads it again.
__shared__ int tb_cta_flag;  // this CTA has raised the global flag or seen it up; reset by pass_enabled()

__device__ __forceinline__ unsigned long long ld_flag_word(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_flag_word(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ bool flag_up(const TbArgs& a) {
  if (*((volatile int*)&tb_cta_flag)) return true;
  const bool up = ld_flag_word(a.scratch) != 0ULL;
  if (up) *((volatile int*)&tb_cta_flag) = 1;
  return up;
}
// may be called from divergent code by any subset of a warp
__device__ __forceinline__ void raise_flag(const TbArgs& a) {
  const unsigned m = __activemask();
  if ((threadIdx.x & 31) == (unsigned)(__ffs(m) - 1) && atomicExch(&tb_cta_flag, 1) == 0) st_flag_word(a.scratch, 1ULL);
}

// CTA-uniform decision whether this pass has anything to do (the flag may flip while pass A runs, so
// one thread samples it for everybody).
__device__ __forceinline__ bool pass_enabled(const TbArgs& a, bool sig_pass) {
  __shared__ int go;
  if (threadIdx.x == 0) {
    const bool up = ld_flag_word(a.scratch) != 0ULL;
    tb_cta_flag = up ? 1 : 0;
    go = sig_pass ? up : !up;
  }
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
  if (FULL && !IGN) {
    // The usual batch: every target is 0 or 1.  One OR over the batch's words settles that (instead of a range test,
    // a select and an error-word update per sample); then every sample counts and its class is the low bit.
    unsigned any_hi = 0, any_lo = 0;
#pragma unroll
    for (int i = 0; i < kPer; ++i) {
      any_hi |= s.hi[i];
      any_lo |= s.lo[i];
    }
    if ((any_hi | (any_lo >> 1)) == 0u) {
#pragma unroll
      for (int i = 0; i < kPer; ++i) {
        if (!SIG) d.outside |= (unsigned)!(s.x[i] >= 0.0f && s.x[i] <= 1.0f);
        d.inc[i] = 1u;
        d.toff[i] = s.lo[i] << 5;
        d.x[i] = s.x[i];
      }
      return;
    }
  }
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
      // `x < q` or `x <= q` by the run-time mode, as predicate logic (two compares and a select per side otherwise)
      const bool lo_ok = (below[i] < q[i]) | (!bn.gt & (below[i] == q[i]));
      const bool hi_bad = (above[i] < q[i]) | (!bn.gt & (above[i] == q[i]));
      sure = (unsigned)((lo_ok & !hi_bad) | ((q[i] != q[i]) & (g[i] == 0)));
    }
    if (!sure) recheck |= 1u << i;
  }
  // A guess that missed, exact scores, thresholds off the even grid (e.g. a linspace rounded to bf16, which 16-bit
  // scores then hit exactly): the right bin is almost always the neighbour on the side the failed compare points to.
  // One more threshold settles that inline; without this step such inputs sent a third of all elements, and so
  // nearly every warp, through the bisection below (C = 10, bf16: 1.64 ms against 0.50 ms for fp32).
  if (!GRID && !kFiltered && recheck) {
#pragma unroll
    for (int i = 0; i < kPer; ++i) {
      if ((recheck >> i) & 1u) {
        const bool up = bn.gt ? (above[i] < q[i]) : (above[i] <= q[i]);  // else the lower neighbour failed (or NaN)
        const int g2 = up ? min(g[i] + 1, bn.T) : max(g[i] - 1, 0);
        const float t_lo = bn.pad[g2], t_hi = bn.pad[g2 + 1];
        const bool ok = (bn.gt ? (t_lo < q[i]) : (t_lo <= q[i])) && !(bn.gt ? (t_hi < q[i]) : (t_hi <= q[i]));
        if (ok) {
          g[i] = g2;
          recheck &= ~(1u << i);
        }
      }
    }
  }
  // cold path: too close to a threshold for the approximation, or a guess that missed by more than one bin
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

// scratch layout (int64 elements): [0, kHdr) header | 2 histogram variants | padded thresholds (as floats)
inline int64_t hist_elems(int64_t L, int64_t T) { return 2 * L * 2 * (T + 1); }
inline int64_t pad_elems(int64_t T) { return (T + 2 + 1) / 2; }
inline int64_t scratch_elems_for(int64_t L, int64_t T) { return kHdr + hist_elems(L, T) + pad_elems(T); }

// multiclass rows with C <= 32 (k2_mc_small.cu); shared-memory layout [pad][2 * C * low_k counters][256 rows]
constexpr int kSmallRows = 256;  // rows per CTA tile = threads per CTA
int launch_mc_small_f32(const TbArgs& a, int64_t rows, int C, int micro, int small_bins, int low_k, int smem,
                        cudaStream_t st);
int launch_mc_small_bf16(const TbArgs& a, int64_t rows, int C, int micro, int small_bins, int low_k, int smem,
                         cudaStream_t st);
int launch_mc_small_f16(const TbArgs& a, int64_t rows, int C, int micro, int small_bins, int low_k, int smem,
                        cudaStream_t st);

}  // namespace tmb
