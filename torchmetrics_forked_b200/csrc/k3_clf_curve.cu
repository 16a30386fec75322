// K3 — exact curves: conditional-sigmoid normalisation at update time, and at compute time the batched
// replacement of `_binary_clf_curve` (functional/classification/precision_recall_curve.py:28-80) and of the
// Python loop over labels that calls it (:819-829, roc.py:345-355).
//
// Reference per label: argsort(descending) -> gather preds/target -> boundaries where consecutive scores
// differ -> float32 cumsum of the positives at the boundaries -> fps = 1 + idx - tps.  Here, for ALL labels
// in one set of launches:
//   pack    [N, L] scores + targets  ->  label-major 32-bit sort keys.  Scores on this path are sigmoid /
//           probability values, i.e. non-negative or NaN, so their sign bit is free: the key is
//           ~(((bits & 0x7fffffff) << 1 | is_positive) + 1) — ascending key order == descending score order,
//           the target rides along as the lowest bit (order inside a tie is irrelevant to the outputs, so no
//           stable key/value sort is needed), ignored entries get the largest key and sink to the end.
//   sort    LSD radix sort, 4 passes x 8 bits, one segment per label: per-tile digit histogram ->
//           exclusive scan over (digit, tile) -> stable scatter with match-based ranking.  Three plain
//           kernels per pass and no inter-CTA spinning (decoupled look-back would save one read of the keys
//           per pass but a bug there hangs the GPU).
//   curve   tile partial sums (positives, boundaries) -> scan over tiles -> emit (fps, tps, threshold) at
//           every boundary.  tps is an exact integer prefix count converted to float32, which is what the
//           reference's float32 cumsum yields while counts stay below 2^24 (H6).
// HBM traffic per label-sample: pack 4+8 read / 4 write, each sort pass 4+4 read / 4 write (+ counters),
// curve 4+4 read / up to 12 write.
#include "common.cuh"

namespace tmb {

constexpr uint32_t kSentinel = 0xffffffffu;  // key of an ignored entry: sorts last

// ---- sort keys --------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t make_key(float v, bool positive, bool& negative_score) {
  uint32_t u = __float_as_uint(v);
  if (v != v) {
    u = 0x7fc00000u;  // every NaN is one canonical NaN; each still forms its own threshold (NaN - NaN != 0)
  } else {
    negative_score |= (u > 0x80000000u);  // strictly negative: not representable in a 31-bit key
    u &= 0x7fffffffu;                     // -0.0 joins +0.0 (their difference is zero in the reference too)
  }
  return ~(((u << 1) | (positive ? 1u : 0u)) + 1u);
}
__device__ __forceinline__ bool key_valid(uint32_t key) { return key != kSentinel; }
__device__ __forceinline__ uint32_t key_positive(uint32_t key) { return (~key - 1u) & 1u; }
__device__ __forceinline__ float key_score(uint32_t key) { return __uint_as_float((~key - 1u) >> 1); }

// ---- conditional sigmoid (exact-mode update) ----------------------------------------------------------------
// pass A: atomicMax(flag, gen) as soon as a value outside [0,1] (or NaN) is seen among the elements that take
// part in the test; pass B: out = (flag == gen) ? sigmoid(x) : x.  `gen` grows with every call, so the flag
// never needs clearing.
template <int DT>
__global__ void __launch_bounds__(256) k3_outside_flag(const void* __restrict__ x, int64_t n,
                                                       const void* __restrict__ target, int tdt, int has_ignore,
                                                       int64_t ignore, unsigned long long* flag,
                                                       unsigned long long gen) {
  __shared__ int go;
  if (threadIdx.x == 0) go = *((volatile unsigned long long*)flag) != gen;
  __syncthreads();
  if (!go) return;
  bool found = false;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float v = load_elem<DT>(x, i);
    bool outside = !(v >= 0.0f && v <= 1.0f);
    if (outside && target != nullptr && has_ignore) outside = load_label(target, tdt, i) != ignore;
    found |= outside;
    if (__any_sync(__activemask(), found)) break;  // somebody in the warp found one: the answer is known
    if ((i & 0xfffff) < (int64_t)gridDim.x * blockDim.x && *((volatile unsigned long long*)flag) == gen) break;
  }
  // one atomic per warp, and none once the flag is known to be up: with logits EVERY thread finds an outside
  // value at once, and 300 k atomics on one address took longer than the whole sigmoid pass
  const unsigned voters = __activemask();
  const bool warp_found = __any_sync(voters, found);
  if (warp_found && (threadIdx.x & 31) == (__ffs(voters) - 1) && *((volatile unsigned long long*)flag) != gen)
    atomicMax(flag, gen);
}

template <int DT>
__device__ __forceinline__ float sigmoid_round(float x) {
  float y = 1.0f / (1.0f + expf(-x));  // same expression as torch's CUDA kernel, IEEE division
  if constexpr (DT == TMB_BF16) y = __bfloat162float(__float2bfloat16_rn(y));
  if constexpr (DT == TMB_F16) y = __half2float(__float2half_rn(y));
  return y;
}

template <int DT>
__global__ void __launch_bounds__(256) k3_normalize(const void* __restrict__ x, void* __restrict__ out, int64_t n,
                                                    const unsigned long long* flag, unsigned long long gen) {
  const bool sig = *flag == gen;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float v = load_elem<DT>(x, i);
    if (sig) v = sigmoid_round<DT>(v);
    if constexpr (DT == TMB_F32) {
      ((float*)out)[i] = v;
    } else if constexpr (DT == TMB_BF16) {
      ((__nv_bfloat16*)out)[i] = __float2bfloat16_rn(v);
    } else {
      ((__half*)out)[i] = __float2half_rn(v);
    }
  }
}

// 16-byte vector versions (aligned pointers, every element takes part in the test): U vectors in flight per thread
template <int DT>
struct VecIO {
  static constexpr int kElems = 16 / Elem<DT>::kBytes;
  __device__ static __forceinline__ void load(const void* base, int64_t v, float (&x)[kElems]) {
    uint4 r = ldg_stream16((const char*)base + v * 16);
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int e = 0; e < kElems; ++e)
      x[e] = (kElems == 4) ? Elem<DT>::get(w[e], 0) : Elem<DT>::get(w[e >> 1], e & 1);
  }
  __device__ static __forceinline__ void store(void* base, int64_t v, const float (&x)[kElems]) {
    uint32_t w[4];
    if constexpr (DT == TMB_F32) {
#pragma unroll
      for (int e = 0; e < 4; ++e) w[e] = __float_as_uint(x[e]);
    } else if constexpr (DT == TMB_BF16) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        __nv_bfloat162 p = __floats2bfloat162_rn(x[2 * e], x[2 * e + 1]);
        w[e] = *reinterpret_cast<uint32_t*>(&p);
      }
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        __half2 p = __floats2half2_rn(x[2 * e], x[2 * e + 1]);
        w[e] = *reinterpret_cast<uint32_t*>(&p);
      }
    }
    *reinterpret_cast<uint4*>((char*)base + v * 16) = make_uint4(w[0], w[1], w[2], w[3]);
  }
};

template <int DT>
__global__ void __launch_bounds__(256) k3_outside_flag_vec(const void* __restrict__ x, int64_t n_vec,
                                                           unsigned long long* flag, unsigned long long gen) {
  constexpr int E = VecIO<DT>::kElems;
  constexpr int U = 4;
  __shared__ int go;
  if (threadIdx.x == 0) go = *((volatile unsigned long long*)flag) != gen;
  __syncthreads();
  if (!go) return;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int lane = threadIdx.x & 31;
  bool found = false;
  // the loop variable is warp-uniform so that the vote below is reached by the whole warp
  for (int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x - lane; b < n_vec; b += stride * U) {
    float v[U][E];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = b + lane + (int64_t)u * stride;
      if (i < n_vec) {
        VecIO<DT>::load(x, i, v[u]);
      } else {
#pragma unroll
        for (int e = 0; e < E; ++e) v[u][e] = 0.0f;
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int e = 0; e < E; ++e) found |= !(v[u][e] >= 0.0f && v[u][e] <= 1.0f);
    if (__any_sync(kFull, found)) break;
    if (*((volatile unsigned long long*)flag) == gen) break;  // another CTA already settled it
  }
  // one atomic per warp, and none once the flag is known to be up: with logits EVERY thread finds an outside
  // value at once, and 300 k atomics on one address took longer than the whole sigmoid pass
  const unsigned voters = __activemask();
  const bool warp_found = __any_sync(voters, found);
  if (warp_found && (threadIdx.x & 31) == (__ffs(voters) - 1) && *((volatile unsigned long long*)flag) != gen)
    atomicMax(flag, gen);
}

template <int DT>
__global__ void __launch_bounds__(256) k3_normalize_vec(const void* __restrict__ x, void* __restrict__ out,
                                                        int64_t n_vec, const unsigned long long* flag,
                                                        unsigned long long gen) {
  constexpr int E = VecIO<DT>::kElems;
  constexpr int U = 4;
  const bool sig = *flag == gen;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; b < n_vec; b += stride * U) {
    float v[U][E];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = b + (int64_t)u * stride;
      if (i < n_vec) VecIO<DT>::load(x, i, v[u]);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = b + (int64_t)u * stride;
      if (i < n_vec) {
        if (sig) {
#pragma unroll
          for (int e = 0; e < E; ++e) v[u][e] = sigmoid_round<DT>(v[u][e]);
        }
        VecIO<DT>::store(out, i, v[u]);
      }
    }
  }
}

// ---- pack: [N, L] row-major -> label-major keys ----------------------------------------------------------------
template <int DT>
__global__ void __launch_bounds__(256) k3_pack(const void* __restrict__ preds, const void* __restrict__ target, int tdt,
                                               int64_t N, int L, int has_ignore, int64_t ignore, int64_t pos_label,
                                               uint32_t* __restrict__ keys, unsigned long long* n_valid, int* err) {
  __shared__ uint32_t tile[32][33];
  __shared__ int cnt[32];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t n0 = (int64_t)blockIdx.x * 32;
  const int l0 = blockIdx.y * 32;
  if (threadIdx.x < 32) cnt[threadIdx.x] = 0;
  bool negative = false;
  for (int j = ty; j < 32; j += 8) {
    const int64_t n = n0 + j;
    const int l = l0 + tx;
    uint32_t key = kSentinel;
    if (n < N && l < L) {
      const int64_t idx = n * L + l;
      const int64_t t = load_label(target, tdt, idx);
      if (!(has_ignore && t == ignore)) key = make_key(load_elem<DT>(preds, idx), t == pos_label, negative);
    }
    tile[j][tx] = key;
  }
  if (negative) atomicOr(err, TMB_ERR_SCORE_RANGE);
  __syncthreads();
  for (int j = ty; j < 32; j += 8) {  // j is warp-uniform: every lane reaches the ballot
    const int l = l0 + j;
    const int64_t n = n0 + tx;
    const bool in_range = (l < L) && (n < N);
    const uint32_t key = in_range ? tile[tx][j] : kSentinel;
    if (in_range) keys[(int64_t)l * N + n] = key;
    const unsigned b = __ballot_sync(kFull, in_range && key_valid(key));
    if (tx == 0 && b) atomicAdd(&cnt[j], __popc(b));
  }
  __syncthreads();
  if (threadIdx.x < 32 && l0 + (int)threadIdx.x < L && cnt[threadIdx.x])
    atomicAdd(&n_valid[l0 + threadIdx.x], (unsigned long long)cnt[threadIdx.x]);
}

// ---- radix sort ----------------------------------------------------------------------------------------------
constexpr int kSortThreads = 512;
constexpr int kSortItems = 16;
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kSortTile = kSortThreads * kSortItems;  // 8192 keys per CTA
constexpr int kWarpSpan = 32 * kSortItems;            // 512 consecutive keys per warp

// counts[(seg * 256 + digit) * nb + tile] = number of keys of `tile` whose digit is `digit`
__global__ void __launch_bounds__(kSortThreads) k3_sort_hist(const uint32_t* __restrict__ keys, int64_t N, int shift,
                                                             uint32_t* __restrict__ counts, int nb) {
  __shared__ uint32_t wh[kSortWarps][256];
  const int seg = blockIdx.y, blk = blockIdx.x;
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < kSortWarps * 256; i += kSortThreads) (&wh[0][0])[i] = 0;
  __syncthreads();
  const uint32_t* k = keys + (int64_t)seg * N;
  const int64_t start = (int64_t)blk * kSortTile + w * kWarpSpan;
#pragma unroll
  for (int r = 0; r < kSortItems; ++r) {
    const int64_t idx = start + r * 32 + lane;
    if (idx < N) atomicAdd(&wh[w][(k[idx] >> shift) & 255u], 1u);
  }
  __syncthreads();
  if (threadIdx.x < 256) {
    uint32_t s = 0;
#pragma unroll
    for (int ww = 0; ww < kSortWarps; ++ww) s += wh[ww][threadIdx.x];
    counts[((int64_t)seg * 256 + threadIdx.x) * nb + blk] = s;
  }
}

// in-place exclusive scan of E consecutive uint32 per segment (one CTA per segment, chunks of 4096);
// optionally writes the segment total
__global__ void __launch_bounds__(1024) k3_scan_u32(uint32_t* __restrict__ data, int64_t E, long long* totals) {
  __shared__ uint32_t warp_sums[32];
  uint32_t* c = data + (int64_t)blockIdx.x * E;
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t carry = 0;
  for (int64_t base = 0; base < E; base += 4096) {
    const int64_t idx = base + (int64_t)threadIdx.x * 4;
    uint32_t v[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) v[e] = (idx + e < E) ? c[idx + e] : 0u;
    const uint32_t s = v[0] + v[1] + v[2] + v[3];
    uint32_t incl = s;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      uint32_t o = __shfl_up_sync(kFull, incl, d);
      if (lane >= d) incl += o;
    }
    if (lane == 31) warp_sums[w] = incl;
    __syncthreads();
    if (w == 0) {
      uint32_t ws = warp_sums[lane];
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        uint32_t o = __shfl_up_sync(kFull, ws, d);
        if (lane >= d) ws += o;
      }
      warp_sums[lane] = ws;  // inclusive over warps
    }
    __syncthreads();
    uint32_t run = carry + (w > 0 ? warp_sums[w - 1] : 0u) + (incl - s);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      if (idx + e < E) c[idx + e] = run;
      run += v[e];
    }
    carry += warp_sums[31];
    __syncthreads();
  }
  if (totals != nullptr && threadIdx.x == 0) totals[blockIdx.x] = (long long)carry;
}

// stable scatter of one tile.  Ranking inside a warp: lanes holding the same digit find each other with
// match_any, the lowest such lane bumps the warp's digit counter for the whole group, each lane adds its
// position inside the group.  Order = (warp, round, lane) = memory order, so the pass is stable.
// The tile is then put in digit order in shared memory and written out from there: neighbouring threads hold
// neighbouring keys of one digit run (32 keys = 128 B on average), so a warp store touches a few sectors instead
// of 32 (writing straight from the ranking registers cost 2.6 ms per pass at 2^21 x 100 keys — one 32-byte
// sector transaction per 4-byte key).
constexpr int kScatterSmem = (kSortWarps * 256 + 256 + 256 + kSortTile) * 4;  // 51 200 B, needs the opt-in

__global__ void __launch_bounds__(kSortThreads, 2) k3_sort_scatter(const uint32_t* __restrict__ in,
                                                                uint32_t* __restrict__ out, int64_t N, int shift,
                                                                const uint32_t* __restrict__ scanned, int nb) {
  extern __shared__ uint32_t sm[];
  uint32_t(*wh)[256] = reinterpret_cast<uint32_t(*)[256]>(sm);  // [kSortWarps][256]: counts, then warp prefixes
  uint32_t* tile_off = sm + kSortWarps * 256;                    // [256] first local position of each digit
  uint32_t* gdelta = tile_off + 256;                             // [256] global position minus local position
  uint32_t* sorted = gdelta + 256;                               // [kSortTile] the tile in digit order
  const int seg = blockIdx.y, blk = blockIdx.x;
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < kSortWarps * 256; i += kSortThreads) sm[i] = 0;
  __syncthreads();
  const uint32_t* k = in + (int64_t)seg * N;
  const int64_t tile0 = (int64_t)blk * kSortTile;
  const int64_t start = tile0 + w * kWarpSpan;
  const int tile_n = (int)min((int64_t)kSortTile, N - tile0);
  const unsigned lt_mask = (1u << lane) - 1u;
  uint32_t key[kSortItems];
  uint32_t rank[kSortItems];
#pragma unroll
  for (int r = 0; r < kSortItems; ++r) {
    const int64_t idx = start + r * 32 + lane;
    key[r] = (idx < N) ? k[idx] : 0u;
  }
#pragma unroll
  for (int r = 0; r < kSortItems; ++r) {
    const bool valid = (start + r * 32 + lane) < N;
    const uint32_t d = (key[r] >> shift) & 255u;
    // lanes holding the same digit, from eight ballots (one per digit bit).  match.any gives the same mask in one
    // instruction, but its latency grows with the number of distinct values in the warp (~28 for random digits)
    // and made the ranking, not the memory traffic, the cost of this kernel.
    unsigned peers = __ballot_sync(kFull, valid);
#pragma unroll
    for (int bit = 0; bit < 8; ++bit) {
      const unsigned set = __ballot_sync(kFull, (d >> bit) & 1u);
      peers &= ((d >> bit) & 1u) ? set : ~set;
    }
    const int leader = __ffs(peers) - 1;
    uint32_t old = 0;
    if (lane == leader && valid) {
      old = wh[w][d];
      wh[w][d] = old + __popc(peers);
    }
    old = __shfl_sync(kFull, old, leader);
    rank[r] = old + __popc(peers & lt_mask);
    __syncwarp();
  }
  __syncthreads();
  // per digit: prefix over the warps, tile total; then an exclusive scan of the totals over the 256 digits
  uint32_t total = 0;
  if (threadIdx.x < 256) {
#pragma unroll
    for (int ww = 0; ww < kSortWarps; ++ww) {
      uint32_t t = wh[ww][threadIdx.x];
      wh[ww][threadIdx.x] = total;
      total += t;
    }
  }
  {
    uint32_t incl = total;  // threads >= 256 carry 0 and only take part in the shuffles
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      uint32_t o = __shfl_up_sync(kFull, incl, d);
      if (lane >= d) incl += o;
    }
    __shared__ uint32_t warp_tot[8];
    if (w < 8 && lane == 31) warp_tot[w] = incl;
    __syncthreads();
    if (threadIdx.x < 256) {
      uint32_t before = 0;
#pragma unroll
      for (int ww = 0; ww < 8; ++ww) before += (ww < w) ? warp_tot[ww] : 0u;
      const uint32_t first = before + incl - total;  // local position of the first key with this digit
      tile_off[threadIdx.x] = first;
      gdelta[threadIdx.x] = scanned[((int64_t)seg * 256 + threadIdx.x) * nb + blk] - first;
    }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < kSortItems; ++r) {
    if ((start + r * 32 + lane) < N) {
      const uint32_t d = (key[r] >> shift) & 255u;
      sorted[tile_off[d] + wh[w][d] + rank[r]] = key[r];
    }
  }
  __syncthreads();
  uint32_t* o = out + (int64_t)seg * N;
  for (int i = threadIdx.x; i < tile_n; i += kSortThreads) {
    const uint32_t kk = sorted[i];
    o[gdelta[(kk >> shift) & 255u] + (uint32_t)i] = kk;  // unsigned wrap-around makes "global - local" exact
  }
}

// ---- curve: boundaries + prefix counts ---------------------------------------------------------------------------
constexpr int kCurveThreads = 256;
constexpr int kCurveItems = 16;
constexpr int kCurveTile = kCurveThreads * kCurveItems;  // 4096

// Shared staging of one tile of sorted keys.  A thread owns 16 CONSECUTIVE positions (the scan needs that), but
// loading them straight from global memory makes every warp load touch 32 different 64-byte pieces; so the tile
// (plus the one key after it, for the last boundary test) is brought in with full-width coalesced loads and
// re-read from shared memory.  Word i lives at i + i/32: with that skew the 32 threads of a warp reading "their
// j-th key" (stride 16 words) hit 32 different banks.
constexpr int kKeyWords = kCurveTile + 1 + (kCurveTile + 1) / 32 + 1;
__device__ __forceinline__ int skew(int i) { return i + (i >> 5); }

struct CurveItems {
  uint32_t pos[kCurveItems];  // 1 for a counted positive
  uint32_t bnd[kCurveItems];  // 1 where a threshold ends
  float score[kCurveItems];
};

// stage tile `blk` of the segment starting at `k`; then thread t decodes positions tile0 + 16 t ... + 15
__device__ __forceinline__ void curve_load(const uint32_t* __restrict__ k, int64_t N, int64_t nv, int64_t tile0,
                                           uint32_t* sh_keys, CurveItems& it, uint32_t& pos_sum, uint32_t& bnd_sum) {
  for (int i = threadIdx.x; i <= kCurveTile; i += kCurveThreads)
    sh_keys[skew(i)] = (tile0 + i < N) ? k[tile0 + i] : kSentinel;
  __syncthreads();
  const int base = threadIdx.x * kCurveItems;
  uint32_t key[kCurveItems + 1];
#pragma unroll
  for (int j = 0; j <= kCurveItems; ++j) key[j] = sh_keys[skew(base + j)];
  const int64_t i0 = tile0 + base;
  pos_sum = 0;
  bnd_sum = 0;
#pragma unroll
  for (int j = 0; j < kCurveItems; ++j) {
    const bool valid = (i0 + j) < nv;  // the nv counted entries are exactly the first nv sorted positions
    const float v = key_score(key[j]);
    bool boundary = false;
    if (valid) {
      if (i0 + j == nv - 1) {
        boundary = true;  // the reference always appends the last index
      } else {
        const float diff = key_score(key[j + 1]) - v;  // reference: torch.where(preds[1:] - preds[:-1])
        boundary = !(diff == 0.0f);                     // NaN / inf-inf differences are "non-zero"
      }
    }
    it.pos[j] = valid ? key_positive(key[j]) : 0u;
    it.bnd[j] = boundary ? 1u : 0u;
    it.score[j] = v;
    pos_sum += it.pos[j];
    bnd_sum += it.bnd[j];
  }
}

// block-wide exclusive prefix of one uint32 per thread (256 threads), also returns the block total
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* warp_buf, uint32_t& total) {
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t incl = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    uint32_t o = __shfl_up_sync(kFull, incl, d);
    if (lane >= d) incl += o;
  }
  if (lane == 31) warp_buf[w] = incl;
  __syncthreads();
  uint32_t before = 0, tot = 0;
#pragma unroll
  for (int ww = 0; ww < kCurveThreads / 32; ++ww) {
    uint32_t s = warp_buf[ww];
    if (ww < w) before += s;
    tot += s;
  }
  __syncthreads();
  total = tot;
  return before + incl - v;
}

// pos_part / bnd_part[seg * ntile + tile] = positives / boundaries inside the tile
__global__ void __launch_bounds__(kCurveThreads) k3_curve_partials(const uint32_t* __restrict__ keys, int64_t N,
                                                                   const unsigned long long* __restrict__ n_valid,
                                                                   uint32_t* __restrict__ pos_part,
                                                                   uint32_t* __restrict__ bnd_part, int ntile) {
  __shared__ uint32_t buf[kCurveThreads / 32];
  __shared__ uint32_t sh_keys[kKeyWords];
  const int seg = blockIdx.y, blk = blockIdx.x;
  const int64_t nv = (int64_t)n_valid[seg];
  const int64_t tile0 = (int64_t)blk * kCurveTile;
  uint32_t ps = 0, bs = 0;
  if (tile0 < nv) {  // CTA-uniform
    CurveItems it;
    curve_load(keys + (int64_t)seg * N, N, nv, tile0, sh_keys, it, ps, bs);
  }
  uint32_t tot_p, tot_b;
  block_excl_scan(ps, buf, tot_p);
  block_excl_scan(bs, buf, tot_b);
  if (threadIdx.x == 0) {
    pos_part[(int64_t)seg * ntile + blk] = tot_p;
    bnd_part[(int64_t)seg * ntile + blk] = tot_b;
  }
}

// The curve points of a tile are compacted into shared memory first and written out from there, so that the
// three output streams leave as full-width coalesced stores.
constexpr int kEmitSmem = (kKeyWords + 3 * kCurveTile + 16) * 4;  // ~66 KB, needs the opt-in

__global__ void __launch_bounds__(kCurveThreads) k3_curve_emit(const uint32_t* __restrict__ keys, int64_t N,
                                                               const unsigned long long* __restrict__ n_valid,
                                                               const uint32_t* __restrict__ pos_excl,
                                                               const uint32_t* __restrict__ bnd_excl, int ntile,
                                                               float* __restrict__ fps, float* __restrict__ tps,
                                                               float* __restrict__ thr) {
  extern __shared__ uint32_t sm[];
  uint32_t* sh_keys = sm;
  float* sh_f = reinterpret_cast<float*>(sm + kKeyWords);
  float* sh_t = sh_f + kCurveTile;
  float* sh_h = sh_t + kCurveTile;
  uint32_t* buf = reinterpret_cast<uint32_t*>(sh_h + kCurveTile);
  const int seg = blockIdx.y, blk = blockIdx.x;
  const int64_t nv = (int64_t)n_valid[seg];
  const int64_t tile0 = (int64_t)blk * kCurveTile;
  if (tile0 >= nv) return;  // CTA-uniform
  CurveItems it;
  uint32_t ps, bs;
  curve_load(keys + (int64_t)seg * N, N, nv, tile0, sh_keys, it, ps, bs);
  uint32_t dummy, tile_points;
  uint32_t p_run = pos_excl[(int64_t)seg * ntile + blk] + block_excl_scan(ps, buf, dummy);
  uint32_t b_loc = block_excl_scan(bs, buf, tile_points);  // position inside this tile's output
  const int64_t i0 = tile0 + (int64_t)threadIdx.x * kCurveItems;
#pragma unroll
  for (int j = 0; j < kCurveItems; ++j) {
    p_run += it.pos[j];  // inclusive count of positives up to this position
    if (it.bnd[j]) {
      const float tp = __uint2float_rn(p_run);
      sh_t[b_loc] = tp;
      sh_f[b_loc] = __ll2float_rn(1 + i0 + j) - tp;  // reference: 1 + threshold_idxs - tps, evaluated in float32
      sh_h[b_loc] = it.score[j];
      b_loc += 1;
    }
  }
  __syncthreads();
  const int64_t out0 = (int64_t)seg * N + bnd_excl[(int64_t)seg * ntile + blk];
  for (int i = threadIdx.x; i < (int)tile_points; i += kCurveThreads) {
    fps[out0 + i] = sh_f[i];
    tps[out0 + i] = sh_t[i];
    thr[out0 + i] = sh_h[i];
  }
}

// ---- curve -> area reductions ----------------------------------------------------------------------------------------
// AUROC and average precision straight from the emitted (fps, tps) points, one read of the two arrays, instead
// of the reference's per-label chain of cat / divide / flip / trapz kernels over 2^24-element tensors
// (roc.py:53-78, utilities/compute.py:88-96, average_precision.py:43-53).  Per label, accumulated in float64:
//   out[4l+0] = sum_b (fps_b - fps_{b-1}) * (tps_b + tps_{b-1}) / 2      (trapezoid area before normalisation)
//   out[4l+1] = sum_b (tps_b - tps_{b-1}) * tps_b / (tps_b + fps_b)      (step-wise precision-recall sum)
//   out[4l+2] = fps of the last point, out[4l+3] = tps of the last point (the normalisers)
// with fps_{-1} = tps_{-1} = 0 (the (0,0) point the reference prepends).  `out` must be zero on entry.
__global__ void __launch_bounds__(256) k3_curve_reduce(const float* __restrict__ fps, const float* __restrict__ tps,
                                                       int64_t N, const long long* __restrict__ counts,
                                                       double* __restrict__ out) {
  const int seg = blockIdx.y;
  const int64_t D = counts[seg];
  const float* f = fps + (int64_t)seg * N;
  const float* t = tps + (int64_t)seg * N;
  double area = 0.0, ap = 0.0;
  for (int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; b < D; b += (int64_t)gridDim.x * blockDim.x) {
    const double fb = f[b], tb = t[b];
    const double fp = b ? (double)f[b - 1] : 0.0, tp = b ? (double)t[b - 1] : 0.0;
    area += (fb - fp) * (tb + tp) * 0.5;
    ap += (tb - tp) * (tb / (tb + fb));
    if (b == D - 1) {
      out[4 * seg + 2] = fb;
      out[4 * seg + 3] = tb;
    }
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    area += __shfl_xor_sync(kFull, area, d);
    ap += __shfl_xor_sync(kFull, ap, d);
  }
  __shared__ double red[2][8];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
    red[0][w] = area;
    red[1][w] = ap;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0.0, p = 0.0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      a += red[0][i];
      p += red[1][i];
    }
    if (a != 0.0) atomicAdd(&out[4 * seg + 0], a);
    if (p != 0.0) atomicAdd(&out[4 * seg + 1], p);
  }
}

// ---- ordered compaction (ignore_index filtering with data-dependent output size) -----------------------------------
__device__ __forceinline__ void copy_elem(void* dst, int64_t o, const void* src, int64_t i, int size) {
  switch (size) {
    case 8: ((uint64_t*)dst)[o] = ((const uint64_t*)src)[i]; break;
    case 4: ((uint32_t*)dst)[o] = ((const uint32_t*)src)[i]; break;
    case 2: ((uint16_t*)dst)[o] = ((const uint16_t*)src)[i]; break;
    default: ((uint8_t*)dst)[o] = ((const uint8_t*)src)[i]; break;
  }
}

__global__ void __launch_bounds__(kCurveThreads) k3_compact_count(const void* __restrict__ target, int tdt, int64_t n,
                                                                  int64_t ignore, uint32_t* __restrict__ tile_counts) {
  __shared__ uint32_t buf[kCurveThreads / 32];
  const int64_t i0 = (int64_t)blockIdx.x * kCurveTile + (int64_t)threadIdx.x * kCurveItems;
  uint32_t c = 0;
#pragma unroll
  for (int j = 0; j < kCurveItems; ++j)
    if (i0 + j < n) c += load_label(target, tdt, i0 + j) != ignore;
  uint32_t total;
  block_excl_scan(c, buf, total);
  if (threadIdx.x == 0) tile_counts[blockIdx.x] = total;
}

__global__ void __launch_bounds__(kCurveThreads) k3_compact_scatter(const void* __restrict__ preds, int psize,
                                                                    const void* __restrict__ target, int tdt,
                                                                    int tsize, int64_t n, int64_t ignore,
                                                                    const uint32_t* __restrict__ tile_excl,
                                                                    void* __restrict__ out_preds,
                                                                    void* __restrict__ out_target) {
  __shared__ uint32_t buf[kCurveThreads / 32];
  const int64_t i0 = (int64_t)blockIdx.x * kCurveTile + (int64_t)threadIdx.x * kCurveItems;
  bool keep[kCurveItems];
  uint32_t c = 0;
#pragma unroll
  for (int j = 0; j < kCurveItems; ++j) {
    keep[j] = (i0 + j < n) && load_label(target, tdt, i0 + j) != ignore;
    c += keep[j];
  }
  uint32_t total;
  int64_t o = (int64_t)tile_excl[blockIdx.x] + block_excl_scan(c, buf, total);
#pragma unroll
  for (int j = 0; j < kCurveItems; ++j) {
    if (keep[j]) {
      copy_elem(out_preds, o, preds, i0 + j, psize);
      copy_elem(out_target, o, target, i0 + j, tsize);
      ++o;
    }
  }
}

// ---- host -----------------------------------------------------------------------------------------------------------
static int64_t align_up(int64_t x, int64_t a) { return (x + a - 1) / a * a; }

struct CurvePlan {
  int64_t N, L;
  int nb, ntile;
  int64_t off_keys_a, off_keys_b, off_counts, off_pos, off_bnd, off_nvalid, total;
};

static CurvePlan curve_plan(int64_t N, int64_t L) {
  CurvePlan p;
  p.N = N;
  p.L = L;
  p.nb = (int)((N + kSortTile - 1) / kSortTile);
  p.ntile = (int)((N + kCurveTile - 1) / kCurveTile);
  int64_t off = 0;
  p.off_keys_a = off; off = align_up(off + L * N * 4, 256);
  p.off_keys_b = off; off = align_up(off + L * N * 4, 256);
  p.off_counts = off; off = align_up(off + L * 256 * (int64_t)p.nb * 4, 256);
  p.off_pos = off; off = align_up(off + L * (int64_t)p.ntile * 4, 256);
  p.off_bnd = off; off = align_up(off + L * (int64_t)p.ntile * 4, 256);
  p.off_nvalid = off; off = align_up(off + L * 8, 256);
  p.total = off;
  return p;
}

template <int DT>
static int run_normalize(const void* x, void* out, int64_t n, const void* target, int tdt, int has_ignore,
                         int64_t ignore, unsigned long long* flag, unsigned long long gen, cudaStream_t st) {
  constexpr int E = VecIO<DT>::kElems;
  constexpr int ES = Elem<DT>::kBytes;
  const bool aligned = (((uintptr_t)x | (uintptr_t)out) & 15) == 0;
  const bool test_all = !(target != nullptr && has_ignore);
  const int64_t n_vec = aligned ? n / E : 0;          // elements [0, n_vec*E) go through the vector kernels
  const int64_t tail0 = n_vec * E, n_tail = n - tail0;  // the rest through the scalar ones
  const char* x_tail = (const char*)x + tail0 * ES;
  char* out_tail = (char*)out + tail0 * ES;
  const void* t_tail = target ? (const char*)target + tail0 * label_bytes(tdt) : nullptr;
  // pass A: is any tested element outside [0,1]?
  if (test_all && n_vec > 0) {
    auto k = k3_outside_flag_vec<DT>;
    k<<<resident_grid(k, 256, 0, (n_vec + 256 * 4 - 1) / (256 * 4)), 256, 0, st>>>(x, n_vec, flag, gen);
    if (int rc = launch_status()) return rc;
    if (n_tail > 0) {
      k3_outside_flag<DT><<<1, 256, 0, st>>>(x_tail, n_tail, t_tail, tdt, has_ignore, ignore, flag, gen);
      if (int rc = launch_status()) return rc;
    }
  } else {
    k3_outside_flag<DT><<<wave_grid((n + 255) / 256, 8), 256, 0, st>>>(x, n, target, tdt, has_ignore, ignore, flag, gen);
    if (int rc = launch_status()) return rc;
  }
  // pass B: copy or sigmoid
  if (n_vec > 0) {
    auto k = k3_normalize_vec<DT>;
    k<<<resident_grid(k, 256, 0, (n_vec + 256 * 4 - 1) / (256 * 4)), 256, 0, st>>>(x, out, n_vec, flag, gen);
    if (int rc = launch_status()) return rc;
  }
  if (n_tail > 0) {
    k3_normalize<DT><<<wave_grid((n_tail + 255) / 256, 8), 256, 0, st>>>(x_tail, out_tail, n_tail, flag, gen);
    if (int rc = launch_status()) return rc;
  }
  return 0;
}

}  // namespace tmb

using namespace tmb;

extern "C" int tmb_sigmoid_normalize(const void* preds, int preds_dtype, int64_t n, const void* target,
                                     int target_dtype, int has_ignore, int64_t ignore_index, void* out,
                                     int64_t* flag, int64_t generation, void* stream) {
  if (float_bytes(preds_dtype) == 0) return TMB_E_DTYPE;
  if (target != nullptr && label_bytes(target_dtype) == 0) return TMB_E_DTYPE;
  if (n < 0 || generation < 1) return TMB_E_SHAPE;
  if (!flag) return TMB_E_NULL;
  if (n == 0) return 0;
  if (!preds || !out) return TMB_E_NULL;
  cudaStream_t st = (cudaStream_t)stream;
  unsigned long long* f = (unsigned long long*)flag;
  const unsigned long long gen = (unsigned long long)generation;
  switch (preds_dtype) {
    case TMB_F32: return run_normalize<TMB_F32>(preds, out, n, target, target_dtype, has_ignore, ignore_index, f, gen, st);
    case TMB_F16: return run_normalize<TMB_F16>(preds, out, n, target, target_dtype, has_ignore, ignore_index, f, gen, st);
    default: return run_normalize<TMB_BF16>(preds, out, n, target, target_dtype, has_ignore, ignore_index, f, gen, st);
  }
}

extern "C" int64_t tmb_clf_curve_scratch_bytes(int64_t n, int64_t num_labels) {
  if (n < 1 || num_labels < 1) return 0;
  return curve_plan(n, num_labels).total;
}

extern "C" int tmb_clf_curves(const void* preds, int preds_dtype, const void* target, int target_dtype, int64_t n,
                              int64_t num_labels, int has_ignore, int64_t ignore_index, int64_t pos_label,
                              void* scratch, int64_t scratch_bytes, float* fps, float* tps, float* thresholds,
                              int64_t* counts, int32_t* err, void* stream) {
  if (float_bytes(preds_dtype) == 0 || label_bytes(target_dtype) == 0) return TMB_E_DTYPE;
  if (n < 0 || num_labels < 1 || num_labels > 65535 * 32) return TMB_E_SHAPE;
  if (n >= ((int64_t)1 << 31)) return TMB_E_LIMIT;  // 32-bit ranks / prefix counts per label
  if (!counts || !err) return TMB_E_NULL;
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) {
    cudaMemsetAsync(counts, 0, num_labels * 8, st);
    return launch_status();
  }
  if (!preds || !target || !scratch || !fps || !tps || !thresholds) return TMB_E_NULL;
  const CurvePlan p = curve_plan(n, num_labels);
  if (scratch_bytes < p.total) return TMB_E_SCRATCH;
  char* base = (char*)scratch;
  uint32_t* keys_a = (uint32_t*)(base + p.off_keys_a);
  uint32_t* keys_b = (uint32_t*)(base + p.off_keys_b);
  uint32_t* cnt = (uint32_t*)(base + p.off_counts);
  uint32_t* pos_part = (uint32_t*)(base + p.off_pos);
  uint32_t* bnd_part = (uint32_t*)(base + p.off_bnd);
  unsigned long long* n_valid = (unsigned long long*)(base + p.off_nvalid);
  const int L = (int)num_labels;

  cudaMemsetAsync(n_valid, 0, num_labels * 8, st);
  {
    dim3 grid((unsigned)((n + 31) / 32), (unsigned)((L + 31) / 32));
    switch (preds_dtype) {
      case TMB_F32:
        k3_pack<TMB_F32><<<grid, 256, 0, st>>>(preds, target, target_dtype, n, L, has_ignore, ignore_index, pos_label,
                                               keys_a, n_valid, err);
        break;
      case TMB_F16:
        k3_pack<TMB_F16><<<grid, 256, 0, st>>>(preds, target, target_dtype, n, L, has_ignore, ignore_index, pos_label,
                                               keys_a, n_valid, err);
        break;
      default:
        k3_pack<TMB_BF16><<<grid, 256, 0, st>>>(preds, target, target_dtype, n, L, has_ignore, ignore_index, pos_label,
                                                keys_a, n_valid, err);
        break;
    }
    if (int rc = launch_status()) return rc;
  }
  uint32_t* src = keys_a;
  uint32_t* dst = keys_b;
  for (int pass = 0; pass < 4; ++pass) {
    const int shift = 8 * pass;
    dim3 grid((unsigned)p.nb, (unsigned)L);
    k3_sort_hist<<<grid, kSortThreads, 0, st>>>(src, n, shift, cnt, p.nb);
    if (int rc = launch_status()) return rc;
    k3_scan_u32<<<L, 1024, 0, st>>>(cnt, (int64_t)256 * p.nb, nullptr);
    if (int rc = launch_status()) return rc;
    cudaFuncSetAttribute(k3_sort_scatter, cudaFuncAttributeMaxDynamicSharedMemorySize, kScatterSmem);
    k3_sort_scatter<<<grid, kSortThreads, kScatterSmem, st>>>(src, dst, n, shift, cnt, p.nb);
    if (int rc = launch_status()) return rc;
    uint32_t* tmp = src;
    src = dst;
    dst = tmp;
  }
  // after 4 passes the sorted keys are back in keys_a (== src)
  {
    dim3 grid((unsigned)p.ntile, (unsigned)L);
    k3_curve_partials<<<grid, kCurveThreads, 0, st>>>(src, n, n_valid, pos_part, bnd_part, p.ntile);
    if (int rc = launch_status()) return rc;
    k3_scan_u32<<<L, 1024, 0, st>>>(pos_part, p.ntile, nullptr);
    if (int rc = launch_status()) return rc;
    k3_scan_u32<<<L, 1024, 0, st>>>(bnd_part, p.ntile, (long long*)counts);
    if (int rc = launch_status()) return rc;
    cudaFuncSetAttribute(k3_curve_emit, cudaFuncAttributeMaxDynamicSharedMemorySize, kEmitSmem);
    k3_curve_emit<<<grid, kCurveThreads, kEmitSmem, st>>>(src, n, n_valid, pos_part, bnd_part, p.ntile, fps, tps, thresholds);
    if (int rc = launch_status()) return rc;
  }
  return 0;
}

extern "C" int tmb_curve_reduce(const float* fps, const float* tps, int64_t n, int64_t num_labels,
                                const int64_t* counts, double* out, void* stream) {
  if (n < 0 || num_labels < 1 || num_labels > 65535) return TMB_E_SHAPE;
  if (!counts || !out) return TMB_E_NULL;
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(out, 0, num_labels * 4 * sizeof(double), st);
  if (int rc = launch_status()) return rc;
  if (n == 0) return 0;
  if (!fps || !tps) return TMB_E_NULL;
  int64_t gx = (n + 256 * 8 - 1) / (256 * 8);
  if (gx > 1024) gx = 1024;
  dim3 grid((unsigned)gx, (unsigned)num_labels);
  k3_curve_reduce<<<grid, 256, 0, st>>>(fps, tps, n, (const long long*)counts, out);
  return launch_status();
}

extern "C" int64_t tmb_compact_scratch_bytes(int64_t n) {
  if (n < 1) return 0;
  return align_up(((n + kCurveTile - 1) / kCurveTile) * 4, 256);
}

extern "C" int tmb_compact_not_ignored(const void* preds, int preds_elem_bytes, const void* target, int target_dtype,
                                       int64_t n, int64_t ignore_index, void* out_preds, void* out_target,
                                       int64_t* count, void* scratch, int64_t scratch_bytes, void* stream) {
  const int tsize = label_bytes(target_dtype);
  if (tsize == 0) return TMB_E_DTYPE;
  if (preds_elem_bytes != 1 && preds_elem_bytes != 2 && preds_elem_bytes != 4 && preds_elem_bytes != 8)
    return TMB_E_DTYPE;
  if (n < 0) return TMB_E_SHAPE;
  if (n >= ((int64_t)1 << 32)) return TMB_E_LIMIT;
  if (!count) return TMB_E_NULL;
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) {
    cudaMemsetAsync(count, 0, 8, st);
    return launch_status();
  }
  if (!preds || !target || !out_preds || !out_target || !scratch) return TMB_E_NULL;
  if (scratch_bytes < tmb_compact_scratch_bytes(n)) return TMB_E_SCRATCH;
  const int tiles = (int)((n + kCurveTile - 1) / kCurveTile);
  uint32_t* tc = (uint32_t*)scratch;
  k3_compact_count<<<tiles, kCurveThreads, 0, st>>>(target, target_dtype, n, ignore_index, tc);
  if (int rc = launch_status()) return rc;
  k3_scan_u32<<<1, 1024, 0, st>>>(tc, tiles, (long long*)count);
  if (int rc = launch_status()) return rc;
  k3_compact_scatter<<<tiles, kCurveThreads, 0, st>>>(preds, preds_elem_bytes, target, target_dtype, tsize, n,
                                                      ignore_index, tc, out_preds, out_target);
  return launch_status();
}
