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
#include <atomic>

#include "k2_binning.cuh"

namespace tmb {
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
      // the reference's check of label-valued preds looks at every element, ignored positions included
      // (stat_scores.py:80-88, confusion_matrix.py:95-103)
      if (p != 0 && p != 1) atomicOr(a.err, TMB_ERR_PREDS_RANGE);
      if (!ignored) {
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
__device__ __forceinline__ void k2_finalize_body(unsigned long long* scratch, int L, int T,
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

__global__ void __launch_bounds__(256) k2_finalize(unsigned long long* scratch, int L, int T,
                                                   const int* __restrict__ perm, int layout, long long* out0,
                                                   long long* out1, long long* out2, long long* out3) {
  k2_finalize_body(scratch, L, T, perm, layout, out0, out1, out2, out3);
}

__global__ void __launch_bounds__(256) k2_clear(unsigned long long* scratch, int64_t scratch_elems) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < scratch_elems;
       i += (int64_t)gridDim.x * blockDim.x)
    scratch[i] = 0ULL;
}

// finalize + clear in one launch: the last CTA of k2_finalize to finish (ticket in scratch[1]) zeroes the scratch —
// flag, ticket and both histograms — for the next update.  One launch less per update (K2 updates are short enough
// for a launch to show: config 3 sharded over 8 GPUs is ~0.1 ms of binning).
__global__ void __launch_bounds__(256) k2_finalize_clear(unsigned long long* scratch, int L, int T,
                                                         const int* __restrict__ perm, int layout, long long* out0,
                                                         long long* out1, long long* out2, long long* out3,
                                                         int64_t scratch_elems) {
  k2_finalize_body(scratch, L, T, perm, layout, out0, out1, out2, out3);
  __shared__ int last;
  __syncthreads();  // every warp of this CTA has read what it needs from the scratch
  if (threadIdx.x == 0) {
    __threadfence();
    last = atomicAdd(&scratch[1], 1ULL) == (unsigned long long)(gridDim.x - 1);
  }
  __syncthreads();
  if (!last) return;
  for (int64_t i = threadIdx.x; i < scratch_elems; i += blockDim.x) scratch[i] = 0ULL;
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

// float32 sum of x[0..n) in index order, one rounding per addition (what Python's `sum(tensor)` does): the CTA stages
// tiles in shared memory, thread 0 adds them left to right.  For the handful of class-length vectors of MCC.
__global__ void __launch_bounds__(256) k_sum_sequential_f32(const float* __restrict__ x, int64_t n, float* out) {
  __shared__ float tile[2048];
  float acc = 0.0f;
  for (int64_t base = 0; base < n; base += 2048) {
    const int m = (int)(n - base < 2048 ? n - base : 2048);
    for (int i = threadIdx.x; i < m; i += 256) tile[i] = x[base + i];
    __syncthreads();
    if (threadIdx.x == 0)
      for (int i = 0; i < m; ++i) acc = __fadd_rn(acc, tile[i]);
    __syncthreads();
  }
  if (threadIdx.x == 0) *out = acc;
}

// ---- host ------------------------------------------------------------------------------------------------

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
  const int64_t used = scratch_elems_for(num_labels, num_thresholds);
  const int grid = (int)((threads + 255) / 256);
  if (used <= (int64_t)1 << 16) {  // the last CTA clears (a few hundred stores per thread at most)
    k2_finalize_clear<<<grid, 256, 0, st>>>((unsigned long long*)scratch, (int)num_labels, (int)num_thresholds, perm,
                                            layout, (long long*)out0, (long long*)out1, (long long*)out2,
                                            (long long*)out3, used);
    return launch_status();
  }
  k2_finalize<<<grid, 256, 0, st>>>((unsigned long long*)scratch, (int)num_labels, (int)num_thresholds, perm, layout,
                                    (long long*)out0, (long long*)out1, (long long*)out2, (long long*)out3);
  if (int rc = launch_status()) return rc;
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

// ---- K4 helpers: pack / unpack the sum-states of a metric for ONE all-reduce -----------------------------------------
// table (device, int64): [0 .. k] byte offset of state i inside the packed buffer (table[k] = total bytes, all
// multiples of 4) | [k+1 .. 2k] address of state i.  One launch moves every state, whatever their number.
__global__ void __launch_bounds__(256) k4_pack_states(const long long* __restrict__ table, int k, uint32_t* packed,
                                                      int unpack) {
  const long long total_words = table[k] >> 2;
  for (long long w = (long long)blockIdx.x * blockDim.x + threadIdx.x; w < total_words;
       w += (long long)gridDim.x * blockDim.x) {
    int lo = 0, hi = k - 1;  // last state that starts at or before byte 4 w
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if ((table[mid] >> 2) <= w) lo = mid; else hi = mid - 1;
    }
    uint32_t* state = reinterpret_cast<uint32_t*>(table[k + 1 + lo]) + (w - (table[lo] >> 2));
    if (unpack) *state = packed[w]; else packed[w] = *state;
  }
}

static int pack_states_impl(const int64_t* table, int64_t num_states, int64_t total_bytes, void* packed, int unpack,
                            void* stream) {
  if (num_states < 1 || num_states > (1 << 20) || total_bytes < 0 || (total_bytes & 3)) return TMB_E_SHAPE;
  if (!table) return TMB_E_NULL;
  if (total_bytes == 0) return 0;
  if (!packed) return TMB_E_NULL;
  const int64_t words = total_bytes >> 2;
  k4_pack_states<<<wave_grid((words + 255) / 256, 8), 256, 0, (cudaStream_t)stream>>>(
      (const long long*)table, (int)num_states, (uint32_t*)packed, unpack);
  return launch_status();
}

extern "C" int tmb_pack_states(const int64_t* table, int64_t num_states, int64_t total_bytes, void* packed,
                               void* stream) {
  return pack_states_impl(table, num_states, total_bytes, packed, 0, stream);
}
extern "C" int tmb_unpack_states(const int64_t* table, int64_t num_states, int64_t total_bytes, const void* packed,
                                 void* stream) {
  return pack_states_impl(table, num_states, total_bytes, const_cast<void*>(packed), 1, stream);
}

extern "C" int tmb_sum_sequential_f32(const float* x, int64_t n, float* out, void* stream) {
  if (n < 0) return TMB_E_SHAPE;
  if (!out || (n > 0 && !x)) return TMB_E_NULL;
  k_sum_sequential_f32<<<1, 256, 0, (cudaStream_t)stream>>>(x, n, out);
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
  // one slot per device; two host threads making the first call together both look the address up and store the
  // same value, so a plain atomic pointer is all the synchronisation this needs
  static std::atomic<unsigned long long*> address[64];
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= 64) return TMB_E_LIMIT;
  unsigned long long* a = address[dev].load(std::memory_order_acquire);
  if (a == nullptr) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess) return (int)e;
    if (p == nullptr) return (int)cudaErrorInvalidSymbol;
    a = (unsigned long long*)p;
    address[dev].store(a, std::memory_order_release);
  }
  *out = a;
  return 0;
}

extern "C" int tmb_binary_curve_update_noalloc(const void* preds, int preds_dtype, const void* target,
                                               int target_dtype, int64_t n, const float* thr,
                                               int64_t num_thresholds, int has_ignore, int64_t ignore_index,
                                               int64_t* confmat, int32_t* err, void* stream) {
  // every argument check of the calls below comes first: a negative return means that nothing at all was enqueued
  // (not even the clear of the shared array)
  if (label_bytes(target_dtype) == 0 || float_bytes(preds_dtype) == 0) return TMB_E_DTYPE;
  if (n < 0 || num_thresholds < 1) return TMB_E_SHAPE;
  if (num_thresholds > kNoAllocMaxThresholds) return TMB_E_LIMIT;  // the static array is sized for 1024
  if (!confmat || !err || !thr) return TMB_E_NULL;
  if (n > 0 && (!preds || !target)) return TMB_E_NULL;
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
