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
#include <cstdlib>

#include "common.cuh"

namespace tmb {

constexpr uint32_t kSentinel = 0xffffffffu;  // key of an ignored entry: sorts last

// the generation flag of the conditional normalisation, read with a device-scope relaxed load (a `volatile` read is a
// system-scope one; thousands of warps polling one word that way queue up at its L2 slice)
__device__ __forceinline__ unsigned long long ld_gen(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
// end of a testing pass: ONE atomicMax per CTA that found something, none once the flag is known to be up (with
// logits every thread finds an outside value at once)
__device__ __forceinline__ void publish_found(bool found, unsigned long long* flag, unsigned long long gen) {
  if (__syncthreads_or(found) && threadIdx.x == 0 && ld_gen(flag) != gen) atomicMax(flag, gen);
}

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
// pass A: copy x to out while looking for a value outside [0,1] (or NaN) among the elements that take part in the
// test; atomicMax(flag, gen) at the first one, after which every CTA stops (what it has copied so far is overwritten
// by pass B).  pass B: out = sigmoid(x) if flag == gen, else nothing — the copy pass A made stands.  Either way the
// data is read ONCE and written once: probabilities cost pass A alone, logits pass A's first wave plus pass B.
// `gen` grows with every call, so the flag never needs clearing.
template <int DT>
__device__ __forceinline__ void store_raw(void* out, int64_t i, float v) {
  if constexpr (DT == TMB_F32) {
    ((float*)out)[i] = v;
  } else if constexpr (DT == TMB_BF16) {
    ((__nv_bfloat16*)out)[i] = __float2bfloat16_rn(v);  // exact: v came from a bf16
  } else {
    ((__half*)out)[i] = __float2half_rn(v);
  }
}

template <int DT>
__global__ void __launch_bounds__(256) k3_outside_flag(const void* __restrict__ x, void* __restrict__ out, int64_t n,
                                                       const void* __restrict__ target, int tdt, int has_ignore,
                                                       int64_t ignore, unsigned long long* flag,
                                                       unsigned long long gen) {
  __shared__ int go;
  if (threadIdx.x == 0) go = ld_gen(flag) != gen;
  __syncthreads();
  if (!go) return;
  bool found = false;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float v = load_elem<DT>(x, i);
    store_raw<DT>(out, i, v);
    bool outside = !(v >= 0.0f && v <= 1.0f);
    if (outside && target != nullptr && has_ignore) outside = load_label(target, tdt, i) != ignore;
    found |= outside;
    if (__any_sync(__activemask(), found)) break;  // somebody in the warp found one: the answer is known
    if ((i & 0xfffff) < (int64_t)gridDim.x * blockDim.x && ld_gen(flag) == gen) break;
  }
  publish_found(found, flag, gen);
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
  if (*flag != gen) return;  // all values were inside [0,1]: pass A's copy is the result
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float v = sigmoid_round<DT>(load_elem<DT>(x, i));
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
__global__ void __launch_bounds__(256) k3_outside_flag_vec(const void* __restrict__ x, void* __restrict__ out,
                                                           int64_t n_vec, unsigned long long* flag,
                                                           unsigned long long gen) {
  constexpr int E = VecIO<DT>::kElems;
  constexpr int U = 8;
  __shared__ int go;
  if (threadIdx.x == 0) go = ld_gen(flag) != gen;
  __syncthreads();
  if (!go) return;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int lane = threadIdx.x & 31;
  bool found = false;
  int trip = 0;
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
    if (__any_sync(kFull, found)) break;  // before the stores: with logits pass A writes nothing at all
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = b + lane + (int64_t)u * stride;
      if (i < n_vec) VecIO<DT>::store(out, i, v[u]);
    }
    if ((++trip & 15) == 0 && ld_gen(flag) == gen) break;  // another CTA already settled it (polled now and then)
  }
  publish_found(found, flag, gen);
}

template <int DT>
__global__ void __launch_bounds__(256) k3_normalize_vec(const void* __restrict__ x, void* __restrict__ out,
                                                        int64_t n_vec, const unsigned long long* flag,
                                                        unsigned long long gen) {
  constexpr int E = VecIO<DT>::kElems;
  constexpr int U = 8;
  // Flag down (all values inside [0,1]): pass A's copy is the result and the loop below runs zero trips.  Written as
  // a loop bound the compiler cannot see through, not as an early return: once it knows `sig` inside the loop it drops
  // the copy branch, allocates 64 registers instead of 76 and no longer keeps all eight loads of a trip in flight
  // ahead of the arithmetic — 292 us at [2^21,100] instead of 272 us.
  const bool sig = *flag == gen;
  long long none = 0;
  asm volatile("" : "+l"(none));  // opaque zero: the compiler keeps the loop's copy branch and its register budget
  const int64_t n_run = sig ? n_vec : none;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; b < n_run; b += stride * U) {
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
// The scores / targets of an exact-mode metric live in one buffer per `update` call (the reference concatenates
// them in compute(): classification/precision_recall_curve.py:159 `dim_zero_cat`).  The pack kernel reads the
// chunks where they lie through a device table, so the concatenated copy (a second 12 bytes per label-sample of
// traffic and of HBM) never exists.
//   table[0 .. C]        first row of chunk c (table[C] = N)
//   table[C+1 .. 2C]     address of chunk c's scores   ([rows_c, L] row-major, dtype DT)
//   table[2C+1 .. 3C]    address of chunk c's targets  ([rows_c, L] row-major, dtype tdt)
struct PackSrc {
  const void* preds;       // used when table == nullptr: one chunk
  const void* target;
  const long long* table;  // device memory, 3 C + 1 entries
  int num_chunks;
  int class_targets;       // 1: targets hold ONE class index per row ([rows_c]); label l is positive where it equals l
};

// Where the keys of label `lab`, rows n0 ... go.  Local form: keys[lab * N + row].  Label-sharded form (state sync of
// the exact curves over several GPUs, K4): the labels are dealt out in blocks of `labels_per_rank` to the ranks and
// every rank's pack kernel stores the keys of a label straight into the OWNER's buffer — plain 512-byte runs of
// st.global over NVLink peer mappings, no staging copy and no separate all-to-all — at the rows
// [row_offset, row_offset + n_local) of that label's n_total-row segment.  Ignored-entry counts go to the owner's
// n_valid with remote atomics.  (TMA is not involved: these are ordinary stores; nothing here multicasts.)
struct PackDst {
  uint32_t* keys;                         // local form
  unsigned long long* n_valid;
  const unsigned long long* peer_keys;    // device array [world]: address of rank r's key buffer (nullptr = local form)
  const unsigned long long* peer_nvalid;  // device array [world]: address of rank r's n_valid
  int labels_per_rank;
  int64_t n_total, row_offset;
};

constexpr int kPackRows = 128;    // rows per CTA: every label leaves as one 512-byte run of keys
constexpr int kPackLabels = 128;  // labels per CTA
constexpr int kPackThreads = 256;
constexpr int kPackUnroll = 8;
__host__ __device__ constexpr int pack_stride(int labels) {
  // words between two label rows of the staging tile.  A warp stores 32 consecutive (row, label) elements: with
  // >= 32 labels they differ in the label, bank (label + row) % 32 is conflict-free with stride = 1 mod 32; with
  // few labels the warp spans 32 / labels rows and the stride has to be that many banks.
  return kPackRows + (labels >= 32 ? 1 : ((32 + labels - 1) / labels) % 32);
}

// A CTA reads rows [n0, n0 + 128) x labels [l0, l0 + lt): with lt == L (any L <= 128) that is ONE contiguous
// 128 L-element region per chunk, read with full-width coalesced loads whatever L is (the earlier 32 x 32
// transpose tile fetched 1.7x the algorithmic bytes at L = 100: 400-byte rows cut into 128-byte pieces that
// straddle sectors).  Element e of the region -> (row, label) by a multiply-high with ceil(2^32 / lt).
template <int DT>
__global__ void __launch_bounds__(kPackThreads) k3_pack(PackSrc src, int tdt, int64_t N, int L, int has_ignore,
                                                        int64_t ignore, int64_t pos_label, PackDst dst, int* err) {
  extern __shared__ uint32_t tile[];  // [lt][stride]
  __shared__ int sh_chunk;
  __shared__ int cnt[kPackLabels];
  const int lane = threadIdx.x & 31;
  const int64_t n0 = (int64_t)blockIdx.x * kPackRows;
  const int l0 = blockIdx.y * kPackLabels;
  const int lt = min(kPackLabels, L - l0);
  const int rows = (int)min((int64_t)kPackRows, N - n0);
  const int stride = pack_stride(lt);
  const uint32_t magic = lt > 1 ? (uint32_t)(0x100000000ull / (unsigned)lt) + 1u : 0u;  // e / lt for e * lt < 2^32
  const int C = src.num_chunks;
  if (threadIdx.x < kPackLabels) cnt[threadIdx.x] = 0;
  if (src.table != nullptr) {
    if (threadIdx.x == 0) {  // last chunk that starts at or before n0
      int lo = 0, hi = C - 1;
      while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (src.table[mid] <= n0) lo = mid; else hi = mid - 1;
      }
      sh_chunk = lo;
    }
  }
  __syncthreads();
  // The tile's rows are walked chunk by chunk (CTA-uniform outer loop, normally one trip): inside, the base
  // pointers are loop constants and the element loop contains no load but the data loads.  (A per-element
  // "did the row cross into the next chunk?" test with the table read in its rare branch made every element
  // wait for the previous element's data: the branch's load and the data loads shared a scoreboard.)
  int c = src.table != nullptr ? sh_chunk : 0;
  bool negative = false;
  int r_begin = 0;
  while (r_begin < rows) {
    int64_t c_begin = 0, c_end = N;
    const void* pb = src.preds;
    const void* tb = src.target;
    if (src.table != nullptr) {
      while (src.table[c + 1] <= n0 + r_begin) ++c;  // skips empty chunks too
      c_begin = src.table[c];
      c_end = src.table[c + 1];
      pb = (const void*)src.table[C + 1 + c];
      tb = (const void*)src.table[2 * C + 1 + c];
    }
    const int r_end = (int)min((int64_t)rows, c_end - n0);
    const int64_t row_off = n0 - c_begin;            // row n0 inside the chunk
    const int64_t base_off = row_off * L + l0;        // element offset of (row n0, label l0) inside the chunk
    const int e_end = r_end * lt;
    for (int e0 = r_begin * lt + threadIdx.x; e0 < e_end; e0 += kPackThreads * kPackUnroll) {
      float v[kPackUnroll];
      int64_t t[kPackUnroll];
      int slot[kPackUnroll];
#pragma unroll
      for (int q = 0; q < kPackUnroll; ++q) {
        const int e = e0 + q * kPackThreads;
        slot[q] = -1;
        if (e < e_end) {
          const int r = lt > 1 ? (int)__umulhi((uint32_t)e, magic) : e;
          const int col = e - r * lt;
          const int64_t idx = base_off + (r * L + col);
          v[q] = load_elem<DT>(pb, idx);
          t[q] = load_label(tb, tdt, src.class_targets ? row_off + r : idx);
          slot[q] = (col << 8) | r;  // r < 128
        }
      }
#pragma unroll
      for (int q = 0; q < kPackUnroll; ++q) {
        if (slot[q] >= 0) {
          const int col = slot[q] >> 8, r = slot[q] & 255;
          uint32_t key = kSentinel;
          if (!(has_ignore && t[q] == ignore)) {
            const int64_t positive_value = src.class_targets ? (int64_t)(l0 + col) : pos_label;
            key = make_key(v[q], t[q] == positive_value, negative);
          }
          tile[col * stride + r] = key;
        }
      }
    }
    r_begin = r_end;
  }
  __syncthreads();
  if (__any_sync(kFull, negative) && lane == 0) atomicOr(err, TMB_ERR_SCORE_RANGE);
  // write-out: 128 consecutive keys per label; a warp stays inside one label (128 % 32 == 0)
  for (int o = threadIdx.x; o < lt * kPackRows; o += kPackThreads) {
    const int col = o >> 7, r = o & (kPackRows - 1);
    const bool in_range = r < rows;
    uint32_t key = kSentinel;
    if (in_range) {
      key = tile[col * stride + r];
      const int lab = l0 + col;
      if (dst.peer_keys == nullptr) {
        dst.keys[(int64_t)lab * N + n0 + r] = key;
      } else {  // the owner's buffer (a warp stays inside one label, so `owner` is warp-uniform)
        const int owner = lab / dst.labels_per_rank;
        uint32_t* base = reinterpret_cast<uint32_t*>(dst.peer_keys[owner]);
        base[(int64_t)(lab - owner * dst.labels_per_rank) * dst.n_total + dst.row_offset + n0 + r] = key;
      }
    }
    if (has_ignore) {
      const unsigned b = __ballot_sync(kFull, in_range && key_valid(key));
      if (lane == 0 && b) atomicAdd(&cnt[col], __popc(b));
    }
  }
  if (has_ignore) {
    __syncthreads();
    if (threadIdx.x < lt && cnt[threadIdx.x]) {
      const int lab = l0 + threadIdx.x;
      unsigned long long* nv = dst.n_valid + lab;
      if (dst.peer_nvalid != nullptr) {
        const int owner = lab / dst.labels_per_rank;
        nv = reinterpret_cast<unsigned long long*>(dst.peer_nvalid[owner]) + (lab - owner * dst.labels_per_rank);
      }
      atomicAdd(nv, (unsigned long long)cnt[threadIdx.x]);
    }
  } else if (dst.peer_keys == nullptr && blockIdx.x == 0 && threadIdx.x < lt) {
    dst.n_valid[l0 + threadIdx.x] = (unsigned long long)N;  // nothing is ignored: every label counts all N rows
  }
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
  const uint32_t* k = keys + (int64_t)seg * N;
  const int64_t start = (int64_t)blk * kSortTile + w * kWarpSpan;
  const int span = (int)min((int64_t)kWarpSpan, N - start);  // <= 0 for a warp past the end
  // all 16 loads first: interleaved with the shared-memory atomics the compiler kept them in program order and
  // every warp paid 16 DRAM latencies back to back
  uint32_t key[kSortItems];
#pragma unroll
  for (int r = 0; r < kSortItems; ++r) key[r] = (r * 32 + lane < span) ? ldg_stream4(k + start + r * 32 + lane) : 0u;
  for (int i = threadIdx.x; i < kSortWarps * 256; i += kSortThreads) (&wh[0][0])[i] = 0;
  __syncthreads();
#pragma unroll
  for (int r = 0; r < kSortItems; ++r)
    if (r * 32 + lane < span) atomicAdd(&wh[w][(key[r] >> shift) & 255u], 1u);
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

// One-sweep front end: the digit histograms of ALL four passes of every label from one read of the keys:
// ghist[(seg * 4 + pass) * 256 + digit].  Per-warp private counters (no inter-warp contention); a round whose 32 keys
// share the digit — the exponent byte mostly does, quantised scores do in the low bytes — is one add of 32 instead of
// a 32-way serialised atomic (two REDUX over the keys tell which of the four bytes are uniform).
constexpr int kHistAllTiles = 4;  // tiles per CTA
constexpr int kHistAllSmem = 4 * kSortWarps * 256 * 4;

__global__ void __launch_bounds__(kSortThreads) k3_digit_hist_all(const uint32_t* __restrict__ keys, int64_t N,
                                                                  uint32_t* __restrict__ ghist, int nb) {
  extern __shared__ uint32_t hh[];  // [4][kSortWarps][256]
  const int seg = blockIdx.y;
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 4 * kSortWarps * 256; i += kSortThreads) hh[i] = 0;
  __syncthreads();
  const uint32_t* k = keys + (int64_t)seg * N;
  for (int tt = 0; tt < kHistAllTiles; ++tt) {
    const int blk = blockIdx.x * kHistAllTiles + tt;
    if (blk >= nb) break;
    const int64_t start = (int64_t)blk * kSortTile + w * kWarpSpan;
    const int span = (int)min((int64_t)kWarpSpan, N - start);
    uint32_t key[kSortItems];
#pragma unroll
    for (int r = 0; r < kSortItems; ++r) key[r] = (r * 32 + lane < span) ? ldg_stream4(k + start + r * 32 + lane) : 0u;
#pragma unroll
    for (int r = 0; r < kSortItems; ++r) {
      const bool valid = r * 32 + lane < span;
      const bool full = (r + 1) * 32 <= span;  // warp-uniform
      // bytes in which the 32 keys of the round agree: AND and OR over the warp differ in no bit of that byte
      const uint32_t differ = full ? (__reduce_and_sync(kFull, key[r]) ^ __reduce_or_sync(kFull, key[r])) : ~0u;
      uint32_t* rows = hh + w * 256;
      if (full && ((differ - 0x01010101u) & ~differ & 0x80808080u) == 0u) {
        // the usual round: a full warp and no byte uniform — four plain atomics, no per-byte tests (the tests were
        // three quarters of this kernel's 75 instructions per round)
#pragma unroll
        for (int p = 0; p < 4; ++p) atomicAdd(&rows[p * kSortWarps * 256 + ((key[r] >> (8 * p)) & 255u)], 1u);
      } else {
#pragma unroll
        for (int p = 0; p < 4; ++p) {
          const uint32_t d = (key[r] >> (8 * p)) & 255u;
          uint32_t* row = rows + p * kSortWarps * 256;
          if (((differ >> (8 * p)) & 255u) == 0u) {
            if (lane == 0) atomicAdd(&row[d], 32u);
          } else if (valid) {
            atomicAdd(&row[d], 1u);
          }
        }
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 4 * 256; i += kSortThreads) {
    const int p = i >> 8, d = i & 255;
    uint32_t sum = 0;
#pragma unroll
    for (int ww = 0; ww < kSortWarps; ++ww) sum += hh[(p * kSortWarps + ww) * 256 + d];
    if (sum) atomicAdd(&ghist[((int64_t)seg * 4 + p) * 256 + d], sum);
  }
}

// in place: ghist[(seg, pass)][0 .. 255] -> exclusive prefix over the digits; one warp per (label, pass)
__global__ void __launch_bounds__(256) k3_digit_base(uint32_t* __restrict__ ghist, int n_rows) {
  const int row = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (row >= n_rows) return;
  uint32_t* g = ghist + (int64_t)row * 256 + lane * 8;
  uint32_t v[8], sum = 0;
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    v[e] = g[e];
    sum += v[e];
  }
  uint32_t incl = sum;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t o = __shfl_up_sync(kFull, incl, d);
    if (lane >= d) incl += o;
  }
  uint32_t run = incl - sum;
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    g[e] = run;
    run += v[e];
  }
}

// peers & (lanes whose `bit` of v equals mine): one step of an 8-ballot match over the digit
__device__ __forceinline__ unsigned match_bit(unsigned peers, uint32_t v, uint32_t bit) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      ".reg .b32 t, s, m;\n\t"
      "and.b32 t, %1, %2;\n\t"
      "setp.ne.u32 p, t, 0;\n\t"
      "vote.sync.ballot.b32 s, p, 0xffffffff;\n\t"
      "selp.b32 m, 0, 0xffffffff, p;\n\t"
      "lop3.b32 %0, %0, s, m, 0x60;\n\t"  // peers & (s ^ m): s where my bit is set, ~s where it is clear
      "}"
      : "+r"(peers)
      : "r"(v), "r"(bit));
  return peers;
}

__device__ __forceinline__ void put_rank(uint32_t* rank2, int r, uint32_t v) {  // r is a compile-time constant
  if (r & 1) rank2[r >> 1] |= v << 16; else rank2[r >> 1] = v;
}
__device__ __forceinline__ uint32_t get_rank(const uint32_t* rank2, int r) {
  return (r & 1) ? rank2[r >> 1] >> 16 : rank2[r >> 1] & 0xffffu;
}

// stable scatter of one tile.  Ranking inside a warp: lanes holding the same digit find each other with
// match_any, the lowest such lane bumps the warp's digit counter for the whole group, each lane adds its
// position inside the group.  Order = (warp, round, lane) = memory order, so the pass is stable.
// The tile is then put in digit order in shared memory and written out from there: neighbouring threads hold
// neighbouring keys of one digit run (32 keys = 128 B on average), so a warp store touches a few sectors instead
// of 32 (writing straight from the ranking registers cost 2.6 ms per pass at 2^21 x 100 keys — one 32-byte
// sector transaction per 4-byte key).
constexpr int kScatterSmem = (kSortWarps * 256 + 256 + kSortTile) * 4;  // 50 176 B, needs the opt-in

// One-sweep variant (LB = true): the per-label digit histograms of all four passes are taken up front
// (k3_digit_hist_all) and a tile learns how many keys of each digit the EARLIER tiles of its label hold by decoupled
// look-back instead of from a per-pass histogram + scan: tiles take their index from a ticket counter (so every tile
// with a lower index is already running or done: the wait below cannot deadlock), publish their 256 digit counts as
// "aggregate" words, walk back over their predecessors adding aggregates until they meet an "inclusive prefix", and
// publish their own prefix.  A status word carries its data (3-bit pass tag | prefix bit | 28-bit count), so plain
// device-scope relaxed loads and stores are all the protocol needs, and the tag makes one zeroing serve four passes.
constexpr uint32_t kLbPrefixBit = 1u << 28;
constexpr uint32_t kLbValueMask = kLbPrefixBit - 1u;
__device__ __forceinline__ uint32_t ld_relaxed_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

struct LookBack {          // only read when LB
  uint32_t* status;        // [L][nb][256], zero before pass 0
  const uint32_t* gbase;   // [L][4][256] exclusive digit offsets of every label and pass
  unsigned int* ticket;    // one counter per pass
  int pass, L;
  int group;               // labels whose tiles are interleaved in ticket order
};

template <bool LB>
__global__ void __launch_bounds__(kSortThreads, 2) k3_sort_scatter(const uint32_t* __restrict__ in,
                                                                uint32_t* __restrict__ out, int64_t N, int shift,
                                                                const uint32_t* __restrict__ scanned, int nb,
                                                                LookBack lb) {
  extern __shared__ uint32_t sm[];
  uint32_t(*wh)[256] = reinterpret_cast<uint32_t(*)[256]>(sm);  // [kSortWarps][256]: counts, then warp prefixes
  uint32_t* gdelta = sm + kSortWarps * 256;                      // [256] global position minus local position
  uint32_t* sorted = gdelta + 256;                               // [kSortTile] the tile in digit order
  __shared__ unsigned int sh_ticket;
  if (LB && threadIdx.x == 0) sh_ticket = atomicAdd(lb.ticket, 1u);
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < kSortWarps * 256; i += kSortThreads) sm[i] = 0;
  __syncthreads();
  // LB: consecutive tickets go round the labels, so the CTAs that start together work on ~grid / L tiles of each
  // label and the first look-backs are that short
  int seg = blockIdx.y, blk = blockIdx.x;
  if (LB) {  // tickets go round the labels of one group of lb.group labels, group after group
    const unsigned per_group = (unsigned)lb.group * (unsigned)nb;
    unsigned g = sh_ticket / per_group;
    const unsigned n_groups = ((unsigned)lb.L + lb.group - 1) / lb.group;
    if (g >= n_groups) g = n_groups - 1;
    const unsigned idx = sh_ticket - g * per_group;
    const unsigned in_group = min((unsigned)lb.group, (unsigned)lb.L - g * lb.group);
    seg = (int)(g * lb.group + idx % in_group);
    blk = (int)(idx / in_group);
  }
  // global position of this tile's first key of digit threadIdx.x (LB: of the label's first key of that digit; the
  // look-back adds the earlier tiles); fetched now, needed after the ranking
  uint32_t digit_base = 0u;
  if (threadIdx.x < 256)
    digit_base = LB ? lb.gbase[((int64_t)seg * 4 + lb.pass) * 256 + threadIdx.x]
                    : scanned[((int64_t)seg * 256 + threadIdx.x) * nb + blk];
  const uint32_t* k = in + (int64_t)seg * N;
  const int64_t tile0 = (int64_t)blk * kSortTile;
  const int64_t start = tile0 + w * kWarpSpan;
  const int tile_n = (int)min((int64_t)kSortTile, N - tile0);
  const int span = (int)min((int64_t)kWarpSpan, N - start);  // <= 0 for a warp past the end
  const unsigned lt_mask = (1u << lane) - 1u;
  // Positions past the end of the segment carry key 0xffffffff: digit 255 in every pass, and — being the last
  // positions of the tile — they rank behind every real key of that digit, so they land at the very end of the
  // digit-ordered tile (local index >= tile_n) and are never written.  No validity test in the loops below.
  uint32_t key[kSortItems];
  uint32_t rank2[kSortItems / 2];  // two 16-bit ranks per register (a warp rank is below 512 + 32)
#pragma unroll
  for (int r = 0; r < kSortItems; ++r)
    key[r] = (r * 32 + lane < span) ? ldg_stream4(k + start + r * 32 + lane) : 0xffffffffu;
  // The leader's atomic returns the warp's running count of the digit; its consumer (the shuffle that hands the
  // count to the group) is issued one round later, behind the next round's eight votes, so the warp does not
  // sit on the shared-memory round trip 16 times.  Same-warp atomics reach shared memory in issue order, which
  // is what keeps round r + 1 behind round r on a common digit.
  uint32_t prev_old = 0;
  unsigned prev_peers = 0;
  int prev_leader = 0;
#pragma unroll
  for (int r = 0; r < kSortItems; ++r) {
    const uint32_t dk = key[r] >> shift;
    // lanes holding the same digit, from eight ballots (one per digit bit).  match.any gives the same mask in one
    // instruction, but its latency grows with the number of distinct values in the warp (~28 for random digits)
    // and made the ranking, not the memory traffic, the cost of this kernel.
    // Four instructions per bit (test, vote, select, and-xor), pinned in PTX: from the C form the compiler
    // derived the predicate twice and spent six.
    unsigned peers = kFull;
#pragma unroll
    for (int bit = 0; bit < 8; ++bit) peers = match_bit(peers, dk, 1u << bit);
    const int leader = __ffs(peers) - 1;
    uint32_t old = 0;
    if (lane == leader) old = atomicAdd(&wh[w][dk & 255u], (uint32_t)__popc(peers));
    if (r > 0) put_rank(rank2, r - 1, __shfl_sync(kFull, prev_old, prev_leader) + __popc(prev_peers & lt_mask));
    prev_old = old;
    prev_peers = peers;
    prev_leader = leader;
  }
  put_rank(rank2, kSortItems - 1, __shfl_sync(kFull, prev_old, prev_leader) + __popc(prev_peers & lt_mask));
  __syncthreads();
  // per digit: tile total -> exclusive scan over the 256 digits -> wh[warp][digit] becomes the local position of
  // the warp's first key with that digit (digit start + keys of earlier warps)
  uint32_t total = 0;
  if (threadIdx.x < 256) {
#pragma unroll
    for (int ww = 0; ww < kSortWarps; ++ww) total += wh[ww][threadIdx.x];
  }
  uint32_t lb_mine = 0;
  if (LB && threadIdx.x < 256) {
    // publish what this tile really holds right away (the padding keys of a ragged last tile were ranked as digit
    // 255); the walk over the predecessors comes later, behind the shared-memory scatter, when they have had time
    // to publish too
    lb_mine = total - (threadIdx.x == 255 ? (uint32_t)(kSortTile - tile_n) : 0u);
    st_relaxed_u32(lb.status + (((int64_t)seg * nb + blk) * 256 + threadIdx.x),
                   ((uint32_t)(lb.pass + 1) << 29) | (blk == 0 ? kLbPrefixBit : 0u) | lb_mine);
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
      gdelta[threadIdx.x] = digit_base - first;    // LB: the earlier tiles' keys are added below
      uint32_t run = first;
#pragma unroll
      for (int ww = 0; ww < kSortWarps; ++ww) {
        const uint32_t t = wh[ww][threadIdx.x];
        wh[ww][threadIdx.x] = run;
        run += t;
      }
    }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < kSortItems; ++r) sorted[wh[w][(key[r] >> shift) & 255u] + get_rank(rank2, r)] = key[r];
  if (LB && threadIdx.x < 256) {
    const int d = threadIdx.x;
    uint32_t* st = lb.status + ((int64_t)seg * nb) * 256 + d;
    const uint32_t tag = (uint32_t)(lb.pass + 1) << 29;
    uint32_t excl = 0;
    int b = blk - 1;
    bool done = b < 0;
    while (!done) {
      // four predecessors per round trip (the loads are independent); they are consumed nearest first and the walk
      // stops at the first inclusive prefix, or retries from the first word that is not published yet (its tile is
      // running: ticket order)
      uint32_t v[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) v[q] = b - q >= 0 ? ld_relaxed_u32(st + (int64_t)(b - q) * 256) : 0u;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (done) break;
        if ((v[q] >> 29) != (uint32_t)(lb.pass + 1)) break;
        excl += v[q] & kLbValueMask;
        --b;
        done = (v[q] & kLbPrefixBit) != 0u || b < 0;
      }
    }
    if (blk > 0) st_relaxed_u32(st + (int64_t)blk * 256, tag | kLbPrefixBit | (excl + lb_mine));
    gdelta[d] += excl;
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

// all 16 loads of a thread are issued before the first store (a plain copy loop ran load -> store -> load ...
// and the curve kernels spent half their stall samples on that one shared-memory store)
__device__ __forceinline__ void stage_tile(const uint32_t* __restrict__ k, int64_t N, int64_t tile0, uint32_t* sh_keys) {
  const int n_load = (int)min((int64_t)kCurveTile + 1, N - tile0);
  uint32_t v[kCurveItems];
#pragma unroll
  for (int q = 0; q < kCurveItems; ++q) {
    const int i = threadIdx.x + q * kCurveThreads;
    v[q] = i < n_load ? ldg_stream4(k + tile0 + i) : kSentinel;
  }
#pragma unroll
  for (int q = 0; q < kCurveItems; ++q) sh_keys[skew(threadIdx.x + q * kCurveThreads)] = v[q];
  if (threadIdx.x == 0) sh_keys[skew(kCurveTile)] = kCurveTile < n_load ? ldg_stream4(k + tile0 + kCurveTile) : kSentinel;
  __syncthreads();
}

struct CurveItems {
  uint32_t pos[kCurveItems];  // 1 for a counted positive
  uint32_t bnd[kCurveItems];  // 1 where a threshold ends
  float score[kCurveItems];
};

// stage tile `blk` of the segment starting at `k`; then thread t decodes positions tile0 + 16 t ... + 15
__device__ __forceinline__ void curve_load(const uint32_t* __restrict__ k, int64_t N, int64_t nv, int64_t tile0,
                                           uint32_t* sh_keys, CurveItems& it, uint32_t& pos_sum, uint32_t& bnd_sum) {
  stage_tile(k, N, tile0, sh_keys);
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

// the same staging, but the thread keeps only two 16-bit masks (bit j: position base + j is a counted positive /
// ends a threshold) — all the partial sums and the area reductions need
__device__ __forceinline__ void curve_load_bits(const uint32_t* __restrict__ k, int64_t N, int64_t nv, int64_t tile0,
                                                uint32_t* sh_keys, uint32_t& pos_bits, uint32_t& bnd_bits) {
  stage_tile(k, N, tile0, sh_keys);
  const int base = threadIdx.x * kCurveItems;
  const int64_t left = nv - (tile0 + base);  // counted positions from this thread's first one on
  const int n_mine = left >= kCurveItems ? kCurveItems : (left > 0 ? (int)left : 0);
  const int last_j = left <= kCurveItems ? (int)left - 1 : -1;  // where the curve's final point sits, if here
  pos_bits = 0;
  bnd_bits = 0;
  uint32_t key = sh_keys[skew(base)];
#pragma unroll
  for (int j = 0; j < kCurveItems; ++j) {
    const uint32_t next = sh_keys[skew(base + j + 1)];
    const bool boundary = (j == last_j) || !(key_score(next) - key_score(key) == 0.0f);
    pos_bits |= key_positive(key) << j;
    bnd_bits |= (boundary ? 1u : 0u) << j;
    key = next;
  }
  const uint32_t mine = (1u << n_mine) - 1u;  // n_mine <= 16
  pos_bits &= mine;
  bnd_bits &= mine;
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

// pos_part / bnd_part[seg * ntile + tile] = positives / boundaries inside the tile;
// tile_last[...] = 0 if the tile holds no boundary, else ((local position of its last boundary + 1) << 16) |
// (positives of the tile up to and including that position) — what the next tile needs to know about "the
// previous curve point" in k3_curve_scores
__global__ void __launch_bounds__(kCurveThreads) k3_curve_partials(const uint32_t* __restrict__ keys, int64_t N,
                                                                   const unsigned long long* __restrict__ n_valid,
                                                                   uint32_t* __restrict__ pos_part,
                                                                   uint32_t* __restrict__ bnd_part,
                                                                   uint32_t* __restrict__ tile_last, int ntile) {
  __shared__ uint32_t buf[kCurveThreads / 32];
  __shared__ uint32_t sh_keys[kKeyWords];
  const int seg = blockIdx.y, blk = blockIdx.x;
  const int64_t nv = (int64_t)n_valid[seg];
  const int64_t tile0 = (int64_t)blk * kCurveTile;
  uint32_t pb = 0, bb = 0;
  if (tile0 < nv) curve_load_bits(keys + (int64_t)seg * N, N, nv, tile0, sh_keys, pb, bb);  // CTA-uniform
  uint32_t tot_p, tot_b;
  const uint32_t pe = block_excl_scan(__popc(pb), buf, tot_p);
  block_excl_scan(__popc(bb), buf, tot_b);
  uint32_t cand = 0;
  if (bb) {
    const int j = 31 - __clz(bb);
    cand = ((threadIdx.x * kCurveItems + j + 1) << 16) | (pe + __popc(pb & ((2u << j) - 1u)));
  }
  cand = __reduce_max_sync(kFull, cand);
  if ((threadIdx.x & 31) == 0) buf[threadIdx.x >> 5] = cand;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t last = 0;
#pragma unroll
    for (int ww = 0; ww < kCurveThreads / 32; ++ww) last = max(last, buf[ww]);
    pos_part[(int64_t)seg * ntile + blk] = tot_p;
    bnd_part[(int64_t)seg * ntile + blk] = tot_b;
    tile_last[(int64_t)seg * ntile + blk] = last;
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

// ---- sorted keys -> area reductions, no curve in memory ----------------------------------------------------------
// AUROC / average precision need sums over consecutive curve points only, so the fused metrics never write the
// three [L, N] float32 curves (12 bytes per label-sample) just to read two of them back: every thread walks its
// 16 sorted positions, and at a boundary adds the term of that curve point.  The previous curve point of a
// thread's first boundary comes from an exclusive max-scan of (position + 1) << 32 | tps over the CTA, seeded
// with the last boundary of the nearest earlier tile that has one (tile_last).  The float32 roundings are the
// ones the emitted curve would carry (tps = float32(count), fps = float32(1 + idx) - tps), the accumulation is
// float64 — same values as k3_curve_emit + k3_curve_reduce; `out` as documented there, zero on entry.
// kAP = false leaves column 1 (the average-precision sum, one float64 division per curve point) at zero.
template <bool kExact, bool kAP>
__global__ void __launch_bounds__(kCurveThreads) k3_curve_scores(const uint32_t* __restrict__ keys, int64_t N,
                                                                 const unsigned long long* __restrict__ n_valid,
                                                                 const uint32_t* __restrict__ pos_excl,
                                                                 const uint32_t* __restrict__ tile_last, int ntile,
                                                                 double* __restrict__ out) {
  __shared__ uint32_t buf[kCurveThreads / 32];
  __shared__ uint32_t sh_keys[kKeyWords];
  __shared__ unsigned long long wmax[kCurveThreads / 32];
  __shared__ unsigned long long sh_carry;
  __shared__ double red[2][kCurveThreads / 32];
  const int seg = blockIdx.y, blk = blockIdx.x;
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t nv = (int64_t)n_valid[seg];
  const int64_t tile0 = (int64_t)blk * kCurveTile;
  if (tile0 >= nv) return;  // CTA-uniform
  uint32_t pb, bb;
  curve_load_bits(keys + (int64_t)seg * N, N, nv, tile0, sh_keys, pb, bb);
  if (w == 0) {  // nearest earlier tile with a boundary, 32 tiles per step (ties spanning whole tiles are rare)
    unsigned long long carry = 0;
    for (int base = blk - 1; base >= 0; base -= 32) {
      const int j = base - lane;
      const uint32_t tl = j >= 0 ? tile_last[(int64_t)seg * ntile + j] : 0u;
      const unsigned found = __ballot_sync(kFull, tl != 0u);
      if (found) {  // warp-uniform
        const int src = __ffs(found) - 1;
        const uint32_t v = __shfl_sync(kFull, tl, src);
        const int jj = base - src;
        const unsigned long long idx1 = (unsigned long long)jj * kCurveTile + (v >> 16);
        carry = (idx1 << 32) | (unsigned long long)(pos_excl[(int64_t)seg * ntile + jj] + (v & 0xffffu));
        break;
      }
    }
    if (lane == 0) sh_carry = carry;
  }
  uint32_t dummy;
  const uint32_t pe = pos_excl[(int64_t)seg * ntile + blk] + block_excl_scan(__popc(pb), buf, dummy);
  const int64_t i0 = tile0 + (int64_t)threadIdx.x * kCurveItems;
  unsigned long long mine = 0;
  if (bb) {
    const int j = 31 - __clz(bb);
    mine = ((unsigned long long)(i0 + j + 1) << 32) | (unsigned long long)(pe + __popc(pb & ((2u << j) - 1u)));
  }
  unsigned long long incl = mine;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const unsigned long long o = __shfl_up_sync(kFull, incl, d);
    if (lane >= d) incl = max(incl, o);
  }
  if (lane == 31) wmax[w] = incl;
  __syncthreads();  // also orders sh_carry (written before the scan's barriers) — kept explicit for wmax
  unsigned long long prev = sh_carry;
#pragma unroll
  for (int ww = 0; ww < kCurveThreads / 32; ++ww)
    if (ww < w) prev = max(prev, wmax[ww]);
  {
    const unsigned long long up = __shfl_up_sync(kFull, incl, 1);
    if (lane > 0) prev = max(prev, up);
  }
  uint32_t p_run = pe;
  double area = 0.0, ap = 0.0;
  if constexpr (kExact) {
    // N < 2^24: every float32 curve value is an exact integer, so the trapezoid terms are integers too —
    // 2 * area accumulates in 64-bit integer arithmetic (< 2^49) and equals the float64 sum bit for bit
    uint32_t p_tps = (uint32_t)prev, p_idx1 = (uint32_t)(prev >> 32);
    unsigned long long area2 = 0;
    const uint32_t i1 = (uint32_t)i0 + 1u;
#pragma unroll
    for (int j = 0; j < kCurveItems; ++j) {
      p_run += (pb >> j) & 1u;
      if ((bb >> j) & 1u) {
        const uint32_t idx1 = i1 + j;
        const uint32_t df = (idx1 - p_run) - (p_idx1 - p_tps);
        area2 += (unsigned long long)df * (p_run + p_tps);
        if constexpr (kAP) ap += (double)(p_run - p_tps) * ((double)p_run / (double)idx1);
        p_tps = p_run;
        p_idx1 = idx1;
      }
    }
    area = (double)area2 * 0.5;
    if (p_idx1 == (uint32_t)nv && bb) {  // this thread holds the curve's last point
      out[4 * seg + 2] = (double)(p_idx1 - p_tps);
      out[4 * seg + 3] = (double)p_tps;
    }
  } else {
    float ptf = __uint2float_rn((uint32_t)prev);
    float pff = __ll2float_rn((long long)(prev >> 32)) - ptf;
#pragma unroll
    for (int j = 0; j < kCurveItems; ++j) {
      p_run += (pb >> j) & 1u;
      if ((bb >> j) & 1u) {
        const float tf = __uint2float_rn(p_run);
        const float ff = __ll2float_rn(1 + i0 + j) - tf;
        const double fb = ff, tb = tf, fp = pff, tp = ptf;
        area += (fb - fp) * (tb + tp) * 0.5;
        if constexpr (kAP) ap += (tb - tp) * (tb / (tb + fb));
        if (i0 + j == nv - 1) {
          out[4 * seg + 2] = fb;
          out[4 * seg + 3] = tb;
        }
        ptf = tf;
        pff = ff;
      }
    }
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    area += __shfl_xor_sync(kFull, area, d);
    ap += __shfl_xor_sync(kFull, ap, d);
  }
  if (lane == 0) {
    red[0][w] = area;
    red[1][w] = ap;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0.0, p = 0.0;
#pragma unroll
    for (int i = 0; i < kCurveThreads / 32; ++i) {
      a += red[0][i];
      p += red[1][i];
    }
    if (a != 0.0) atomicAdd(&out[4 * seg + 0], a);
    if (p != 0.0) atomicAdd(&out[4 * seg + 1], p);
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
  int64_t off_keys_a, off_keys_b, off_counts, off_pos, off_bnd, off_last, off_nvalid, off_ghist, off_ticket, total;
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
  p.off_last = off; off = align_up(off + L * (int64_t)p.ntile * 4, 256);
  p.off_nvalid = off; off = align_up(off + L * 8, 256);
  p.off_ghist = off; off = align_up(off + L * 4 * 256 * 4, 256);
  p.off_ticket = off; off = align_up(off + 4 * 4, 256);
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
    k<<<resident_grid(k, 256, 0, (n_vec + 256 * 8 - 1) / (256 * 8)), 256, 0, st>>>(x, out, n_vec, flag, gen);
    if (int rc = launch_status()) return rc;
    if (n_tail > 0) {
      k3_outside_flag<DT><<<1, 256, 0, st>>>(x_tail, out_tail, n_tail, t_tail, tdt, has_ignore, ignore, flag, gen);
      if (int rc = launch_status()) return rc;
    }
  } else {
    k3_outside_flag<DT><<<wave_grid((n + 255) / 256, 8), 256, 0, st>>>(x, out, n, target, tdt, has_ignore, ignore, flag,
                                                                       gen);
    if (int rc = launch_status()) return rc;
    // the scalar pass covered everything, including what the vector pass B will overwrite if the flag is up
  }
  // pass B: sigmoid, only when pass A raised the flag
  if (n_vec > 0) {
    auto k = k3_normalize_vec<DT>;
    k<<<resident_grid(k, 256, 0, (n_vec + 256 * 8 - 1) / (256 * 8)), 256, 0, st>>>(x, out, n_vec, flag, gen);
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

namespace tmb {

struct CurveBufs {
  uint32_t *keys_a, *keys_b, *cnt, *pos_part, *bnd_part, *tile_last, *ghist;
  unsigned int* ticket;
  unsigned long long* n_valid;
};

static CurveBufs curve_bufs(const CurvePlan& p, void* scratch) {
  char* base = (char*)scratch;
  CurveBufs b;
  b.keys_a = (uint32_t*)(base + p.off_keys_a);
  b.keys_b = (uint32_t*)(base + p.off_keys_b);
  b.cnt = (uint32_t*)(base + p.off_counts);
  b.pos_part = (uint32_t*)(base + p.off_pos);
  b.bnd_part = (uint32_t*)(base + p.off_bnd);
  b.tile_last = (uint32_t*)(base + p.off_last);
  b.n_valid = (unsigned long long*)(base + p.off_nvalid);
  b.ghist = (uint32_t*)(base + p.off_ghist);
  b.ticket = (unsigned int*)(base + p.off_ticket);
  return b;
}

template <int DT>
static int launch_pack(const PackSrc& src, int tdt, int64_t n, int L, int has_ignore, int64_t ignore,
                       int64_t pos_label, const PackDst& dst, int32_t* err, cudaStream_t st) {
  const int lt_full = L < kPackLabels ? L : kPackLabels;
  const int lt_last = L % kPackLabels ? L % kPackLabels : lt_full;
  int words = lt_full * pack_stride(lt_full);
  if (lt_last * pack_stride(lt_last) > words) words = lt_last * pack_stride(lt_last);
  auto k = k3_pack<DT>;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, kPackLabels * pack_stride(kPackLabels) * 4);
  dim3 grid((unsigned)((n + kPackRows - 1) / kPackRows), (unsigned)((L + kPackLabels - 1) / kPackLabels));
  k<<<grid, kPackThreads, words * 4, st>>>(src, tdt, n, L, has_ignore, ignore, pos_label, dst, err);
  return launch_status();
}

static int pack_keys(const PackSrc& src, int preds_dtype, int target_dtype, int64_t n, int L, int has_ignore,
                     int64_t ignore, int64_t pos_label, const PackDst& dst, int32_t* err, cudaStream_t st) {
  switch (preds_dtype) {
    case TMB_F32: return launch_pack<TMB_F32>(src, target_dtype, n, L, has_ignore, ignore, pos_label, dst, err, st);
    case TMB_F16: return launch_pack<TMB_F16>(src, target_dtype, n, L, has_ignore, ignore, pos_label, dst, err, st);
    default: return launch_pack<TMB_BF16>(src, target_dtype, n, L, has_ignore, ignore, pos_label, dst, err, st);
  }
}

// 4 LSD passes over [L, n] label-major keys; the sorted keys end up back in keys_a.
// One sweep (default, n < 2^28): one histogram kernel for all passes, then per pass ONE scatter launch whose tiles find
// their offsets by look-back.  Classic (TMB_K3_SORT=classic, or n >= 2^28): per pass tile histogram + scan + scatter.
static bool use_one_sweep(int64_t n) {
  static const bool classic = [] {
    const char* e = getenv("TMB_K3_SORT");
    return e != nullptr && e[0] == 'c';
  }();
  return !classic && n < ((int64_t)1 << 28);
}

static int sort_keys(const CurveBufs& b, int64_t n, int L, int nb, cudaStream_t st) {
  uint32_t* src_keys = b.keys_a;
  uint32_t* dst_keys = b.keys_b;
  if (use_one_sweep(n)) {
    auto k = k3_sort_scatter<true>;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, kScatterSmem);
    cudaFuncSetAttribute(k3_digit_hist_all, cudaFuncAttributeMaxDynamicSharedMemorySize, kHistAllSmem);
    cudaMemsetAsync(b.cnt, 0, (size_t)L * 256 * nb * 4, st);  // the tile status words (pass tag 0 = nothing published)
    cudaMemsetAsync(b.ghist, 0, (size_t)L * 4 * 256 * 4, st);
    cudaMemsetAsync(b.ticket, 0, 4 * 4, st);
    dim3 hgrid((unsigned)((nb + kHistAllTiles - 1) / kHistAllTiles), (unsigned)L);
    k3_digit_hist_all<<<hgrid, kSortThreads, kHistAllSmem, st>>>(src_keys, n, b.ghist, nb);
    if (int rc = launch_status()) return rc;
    k3_digit_base<<<(L * 4 * 32 + 255) / 256, 256, 0, st>>>(b.ghist, L * 4);
    if (int rc = launch_status()) return rc;
    for (int pass = 0; pass < 4; ++pass) {
      // all labels in one group (group sizes 8 .. 100 measured the same at L = 100; label-major order — one chain of
      // ~300 co-resident tiles — was 25 % slower)
      const LookBack lb{b.cnt, b.ghist, b.ticket + pass, pass, L, L};
      k<<<(unsigned)((int64_t)nb * L), kSortThreads, kScatterSmem, st>>>(src_keys, dst_keys, n, 8 * pass, nullptr, nb, lb);
      if (int rc = launch_status()) return rc;
      uint32_t* tmp = src_keys;
      src_keys = dst_keys;
      dst_keys = tmp;
    }
    return 0;
  }
  auto k = k3_sort_scatter<false>;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, kScatterSmem);
  for (int pass = 0; pass < 4; ++pass) {
    const int shift = 8 * pass;
    dim3 grid((unsigned)nb, (unsigned)L);
    k3_sort_hist<<<grid, kSortThreads, 0, st>>>(src_keys, n, shift, b.cnt, nb);
    if (int rc = launch_status()) return rc;
    k3_scan_u32<<<L, 1024, 0, st>>>(b.cnt, (int64_t)256 * nb, nullptr);
    if (int rc = launch_status()) return rc;
    k<<<grid, kSortThreads, kScatterSmem, st>>>(src_keys, dst_keys, n, shift, b.cnt, nb, LookBack{});
    if (int rc = launch_status()) return rc;
    uint32_t* tmp = src_keys;
    src_keys = dst_keys;
    dst_keys = tmp;
  }
  return 0;  // an even number of passes: the result is back in keys_a
}

// pack + 4 sort passes; the sorted keys end up in b.keys_a
static int pack_and_sort(const PackSrc& src, int preds_dtype, int target_dtype, int64_t n, int L, int has_ignore,
                         int64_t ignore, int64_t pos_label, const CurvePlan& p, const CurveBufs& b, int32_t* err,
                         cudaStream_t st) {
  cudaMemsetAsync(b.n_valid, 0, (size_t)L * 8, st);
  const PackDst dst{b.keys_a, b.n_valid, nullptr, nullptr, L, n, 0};
  if (int rc = pack_keys(src, preds_dtype, target_dtype, n, L, has_ignore, ignore, pos_label, dst, err, st)) return rc;
  return sort_keys(b, n, L, p.nb, st);
}

static int curve_tile_sums(int64_t n, int L, const CurvePlan& p, const CurveBufs& b, long long* counts,
                           cudaStream_t st) {
  dim3 grid((unsigned)p.ntile, (unsigned)L);
  k3_curve_partials<<<grid, kCurveThreads, 0, st>>>(b.keys_a, n, b.n_valid, b.pos_part, b.bnd_part, b.tile_last,
                                                    p.ntile);
  if (int rc = launch_status()) return rc;
  k3_scan_u32<<<L, 1024, 0, st>>>(b.pos_part, p.ntile, nullptr);
  if (int rc = launch_status()) return rc;
  if (counts != nullptr) {  // the curve lengths / output offsets are only needed when the curve is written
    k3_scan_u32<<<L, 1024, 0, st>>>(b.bnd_part, p.ntile, counts);
    if (int rc = launch_status()) return rc;
  }
  return 0;
}

static int curves_impl(const PackSrc& src, int preds_dtype, int target_dtype, int64_t n, int64_t num_labels,
                       int has_ignore, int64_t ignore_index, int64_t pos_label, void* scratch, int64_t scratch_bytes,
                       float* fps, float* tps, float* thresholds, int64_t* counts, int32_t* err, void* stream) {
  if (float_bytes(preds_dtype) == 0 || label_bytes(target_dtype) == 0) return TMB_E_DTYPE;
  if (n < 0 || num_labels < 1 || num_labels > 65535) return TMB_E_SHAPE;
  if (n >= ((int64_t)1 << 31)) return TMB_E_LIMIT;  // 32-bit ranks / prefix counts per label
  if (!counts || !err) return TMB_E_NULL;
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) {
    cudaMemsetAsync(counts, 0, num_labels * 8, st);
    return launch_status();
  }
  if (!scratch || !fps || !tps || !thresholds) return TMB_E_NULL;
  const CurvePlan p = curve_plan(n, num_labels);
  if (scratch_bytes < p.total) return TMB_E_SCRATCH;
  const CurveBufs b = curve_bufs(p, scratch);
  const int L = (int)num_labels;
  if (int rc = pack_and_sort(src, preds_dtype, target_dtype, n, L, has_ignore, ignore_index, pos_label, p, b, err, st))
    return rc;
  if (int rc = curve_tile_sums(n, L, p, b, (long long*)counts, st)) return rc;
  dim3 grid((unsigned)p.ntile, (unsigned)L);
  cudaFuncSetAttribute(k3_curve_emit, cudaFuncAttributeMaxDynamicSharedMemorySize, kEmitSmem);
  k3_curve_emit<<<grid, kCurveThreads, kEmitSmem, st>>>(b.keys_a, n, b.n_valid, b.pos_part, b.bnd_part, p.ntile, fps,
                                                       tps, thresholds);
  return launch_status();
}

__global__ void k3_fill_u64(unsigned long long* p, unsigned long long v, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// tile partial sums -> scan -> area reductions straight from the sorted keys (b.keys_a, b.n_valid)
static int scores_from_sorted(int64_t n, int L, const CurvePlan& p, const CurveBufs& b, int want_ap, double* scores,
                              cudaStream_t st) {
  if (int rc = curve_tile_sums(n, L, p, b, nullptr, st)) return rc;
  dim3 grid((unsigned)p.ntile, (unsigned)L);
  const bool exact = n < ((int64_t)1 << 24), ap = want_ap != 0;
  auto k = exact ? (ap ? k3_curve_scores<true, true> : k3_curve_scores<true, false>)
                 : (ap ? k3_curve_scores<false, true> : k3_curve_scores<false, false>);
  k<<<grid, kCurveThreads, 0, st>>>(b.keys_a, n, b.n_valid, b.pos_part, b.tile_last, p.ntile, scores);
  return launch_status();
}

static bool chunk_table_ok(const int64_t* table, int64_t num_chunks) {
  return table != nullptr && num_chunks >= 1 && num_chunks < ((int64_t)1 << 24);
}

}  // namespace tmb

extern "C" int tmb_clf_curves(const void* preds, int preds_dtype, const void* target, int target_dtype, int64_t n,
                              int64_t num_labels, int has_ignore, int64_t ignore_index, int64_t pos_label,
                              void* scratch, int64_t scratch_bytes, float* fps, float* tps, float* thresholds,
                              int64_t* counts, int32_t* err, void* stream) {
  if (n > 0 && (!preds || !target)) return TMB_E_NULL;
  const PackSrc src{preds, target, nullptr, 1, 0};
  return curves_impl(src, preds_dtype, target_dtype, n, num_labels, has_ignore, ignore_index, pos_label, scratch,
                     scratch_bytes, fps, tps, thresholds, counts, err, stream);
}

extern "C" int tmb_clf_curves_chunked(const int64_t* chunk_table, int64_t num_chunks, int preds_dtype,
                                      int target_dtype, int64_t n, int64_t num_labels, int has_ignore,
                                      int64_t ignore_index, int64_t pos_label, int class_targets, void* scratch,
                                      int64_t scratch_bytes, float* fps, float* tps, float* thresholds,
                                      int64_t* counts, int32_t* err, void* stream) {
  if (n > 0 && !chunk_table_ok(chunk_table, num_chunks)) return chunk_table ? TMB_E_SHAPE : TMB_E_NULL;
  const PackSrc src{nullptr, nullptr, (const long long*)chunk_table, (int)num_chunks, class_targets ? 1 : 0};
  return curves_impl(src, preds_dtype, target_dtype, n, num_labels, has_ignore, ignore_index, pos_label, scratch,
                     scratch_bytes, fps, tps, thresholds, counts, err, stream);
}

extern "C" int tmb_clf_curve_scores_chunked(const int64_t* chunk_table, int64_t num_chunks, int preds_dtype,
                                            int target_dtype, int64_t n, int64_t num_labels, int has_ignore,
                                            int64_t ignore_index, int64_t pos_label, int class_targets,
                                            int want_ap, void* scratch, int64_t scratch_bytes, double* scores,
                                            int32_t* err, void* stream) {
  if (float_bytes(preds_dtype) == 0 || label_bytes(target_dtype) == 0) return TMB_E_DTYPE;
  if (n < 0 || num_labels < 1 || num_labels > 65535) return TMB_E_SHAPE;
  if (n >= ((int64_t)1 << 31)) return TMB_E_LIMIT;
  if (!scores || !err) return TMB_E_NULL;
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(scores, 0, num_labels * 4 * sizeof(double), st);
  if (int rc = launch_status()) return rc;
  if (n == 0) return 0;
  if (!chunk_table_ok(chunk_table, num_chunks)) return chunk_table ? TMB_E_SHAPE : TMB_E_NULL;
  if (!scratch) return TMB_E_NULL;
  const CurvePlan p = curve_plan(n, num_labels);
  if (scratch_bytes < p.total) return TMB_E_SCRATCH;
  const CurveBufs b = curve_bufs(p, scratch);
  const int L = (int)num_labels;
  const PackSrc src{nullptr, nullptr, (const long long*)chunk_table, (int)num_chunks, class_targets ? 1 : 0};
  if (int rc = pack_and_sort(src, preds_dtype, target_dtype, n, L, has_ignore, ignore_index, pos_label, p, b, err, st))
    return rc;
  return scores_from_sorted(n, L, p, b, want_ap, scores, st);
}

// ---- label-sharded exact compute over several GPUs (K4 for cat states) ------------------------
// Rank s packs ITS rows of every label straight into the key buffer of the label's owner (tmb_clf_pack_keys_sharded,
// remote stores over NVLink); after a barrier every rank sorts and reduces only the labels it owns
// (tmb_clf_sorted_scores) — 1/world of the sort instead of all of it on every rank.
extern "C" int tmb_clf_pack_keys_sharded(const int64_t* chunk_table, int64_t num_chunks, int preds_dtype,
                                         int target_dtype, int64_t n_local, int64_t num_labels, int has_ignore,
                                         int64_t ignore_index, int64_t pos_label, int class_targets,
                                         const uint64_t* peer_keys, const uint64_t* peer_nvalid, int64_t world,
                                         int64_t labels_per_rank, int64_t n_total, int64_t row_offset, int32_t* err,
                                         void* stream) {
  if (float_bytes(preds_dtype) == 0 || label_bytes(target_dtype) == 0) return TMB_E_DTYPE;
  if (n_local < 0 || num_labels < 1 || num_labels > 65535 || world < 1 || labels_per_rank < 1) return TMB_E_SHAPE;
  if (labels_per_rank * world < num_labels || row_offset < 0 || row_offset + n_local > n_total) return TMB_E_SHAPE;
  if (n_total >= ((int64_t)1 << 31)) return TMB_E_LIMIT;
  if (!peer_keys || !peer_nvalid || !err) return TMB_E_NULL;
  if (n_local == 0) return 0;
  if (!chunk_table_ok(chunk_table, num_chunks)) return chunk_table ? TMB_E_SHAPE : TMB_E_NULL;
  const PackSrc src{nullptr, nullptr, (const long long*)chunk_table, (int)num_chunks, class_targets ? 1 : 0};
  const PackDst dst{nullptr, nullptr, (const unsigned long long*)peer_keys, (const unsigned long long*)peer_nvalid,
                    (int)labels_per_rank, n_total, row_offset};
  return pack_keys(src, preds_dtype, target_dtype, n_local, (int)num_labels, has_ignore, ignore_index, pos_label, dst,
                   err, (cudaStream_t)stream);
}

extern "C" int64_t tmb_clf_sorted_scores_scratch_bytes(int64_t n, int64_t num_labels) {
  if (n < 1 || num_labels < 1) return 0;
  const CurvePlan p = curve_plan(n, num_labels);
  return p.total - p.off_keys_b;  // everything but the first key buffer, which the caller owns
}

extern "C" int tmb_clf_sorted_scores(uint32_t* keys, uint64_t* n_valid, int64_t n, int64_t num_labels,
                                     int all_rows_valid, int want_ap, void* scratch, int64_t scratch_bytes,
                                     double* scores, void* stream) {
  if (n < 0 || num_labels < 1 || num_labels > 65535) return TMB_E_SHAPE;
  if (n >= ((int64_t)1 << 31)) return TMB_E_LIMIT;
  if (!scores) return TMB_E_NULL;
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(scores, 0, num_labels * 4 * sizeof(double), st);
  if (int rc = launch_status()) return rc;
  if (n == 0) return 0;
  if (!keys || !n_valid || !scratch) return TMB_E_NULL;
  const CurvePlan p = curve_plan(n, num_labels);
  if (scratch_bytes < p.total - p.off_keys_b) return TMB_E_SCRATCH;
  CurveBufs b = curve_bufs(p, (char*)scratch - p.off_keys_b);  // offsets of the plan are relative to keys_a
  b.keys_a = keys;
  b.n_valid = (unsigned long long*)n_valid;
  const int L = (int)num_labels;
  if (all_rows_valid) {
    k3_fill_u64<<<(L + 255) / 256, 256, 0, st>>>(b.n_valid, (unsigned long long)n, L);
    if (int rc = launch_status()) return rc;
  }
  if (int rc = sort_keys(b, n, L, p.nb, st)) return rc;
  return scores_from_sorted(n, L, p, b, want_ap, scores, st);
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
