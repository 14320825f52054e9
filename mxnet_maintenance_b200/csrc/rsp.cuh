// Row-sparse kernels for sm_100a, CUB-free:
//   * LSD radix sort of (row id, source position) pairs, 8-bit digits, only ceil(log2(num_rows)/8)
//     passes (the reference sorts full 64-bit keys with CUB, src/kvstore/kvstore_utils.cu:44-98)
//   * head-flag scan -> sorted unique ids + segment starts (replaces cub::DeviceSelect::Unique and
//     the O(num_rows) flag array + cub::DeviceScan of src/ndarray/ndarray_function.cu:104-190)
//   * segmented row reduce in source order (= the reference's device-order accumulation,
//     src/ndarray/ndarray_function.cc:81-113), optionally fused with the lazy row update
//     (SGDMomDnsRspDnsKernel / AdamDnsRspDnsKernel, src/operator/optimizer_op-inl.h:747-802,1342-1409)
//   * row gather for row_sparse_pull (SparseRetainRspThreadKernel, sparse_retain-inl.h:121-150)
// Work is O(nnz), never O(num_rows).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "optim.cuh"

namespace b200kv {

constexpr int kSortTile = 4096;     // keys per CTA: 8 warps x 16 rounds x 32 lanes
constexpr int kSortThreads = 256;
constexpr int kScanTile = 1024;     // keys per CTA in the head-flag scan

// ---- ids -> sort keys ------------------------------------------------------------------------------
struct RspInputs {
  const void* idx[kMaxCopies];   // row ids of each copy (int64 for pushes)
  const float* val[kMaxCopies];  // row values of each copy
  int64_t start[kMaxCopies + 1]; // position of each copy's first row in the concatenation
  int n;                         // copies
};

// dtype: kI64 | kI32 | kF32 (row_ids given as float arrays hold integers, Appendix A.9)
template <typename I>
__device__ __forceinline__ uint32_t id_to_key(I v) { return static_cast<uint32_t>(static_cast<int64_t>(v)); }

template <typename I>
__global__ void make_keys_kernel(const RspInputs in, int64_t total, uint32_t* __restrict__ keys,
                                 uint32_t* __restrict__ pay) {
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    int k = 0;
    while (k + 1 < in.n && i >= in.start[k + 1]) ++k;
    keys[i] = id_to_key(static_cast<const I*>(in.idx[k])[i - in.start[k]]);
    pay[i] = static_cast<uint32_t>(i);
  }
}

// ---- radix sort ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kSortThreads)
radix_hist_kernel(const uint32_t* __restrict__ keys, int n, int shift, uint32_t* __restrict__ counts,
                  int nblocks) {
  __shared__ uint32_t hist[256];
  hist[threadIdx.x] = 0;
  __syncthreads();
  const int base = blockIdx.x * kSortTile;
  for (int i = threadIdx.x; i < kSortTile; i += kSortThreads) {
    const int g = base + i;
    if (g < n) atomicAdd(&hist[(keys[g] >> shift) & 255u], 1u);
  }
  __syncthreads();
  counts[threadIdx.x * nblocks + blockIdx.x] = hist[threadIdx.x];
}

// exclusive scan of m values in place, one CTA of 1024 threads (m = 256 * nblocks, a few 10^4)
__global__ void __launch_bounds__(1024) scan_small_kernel(uint32_t* __restrict__ v, int m) {
  __shared__ uint32_t wsum[32];
  const int tid = threadIdx.x;
  const int per = (m + 1023) / 1024;
  const int b = tid * per, e = min(b + per, m);
  uint32_t s = 0;
  for (int i = b; i < e; ++i) s += v[i];
  // block-wide exclusive scan of s
  uint32_t x = s;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if ((tid & 31) >= o) x += y;
  }
  if ((tid & 31) == 31) wsum[tid >> 5] = x;
  __syncthreads();
  if (tid < 32) {
    uint32_t w = wsum[tid];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(0xffffffffu, w, o);
      if (tid >= o) w += y;
    }
    wsum[tid] = w;
  }
  __syncthreads();
  uint32_t run = x - s + ((tid >> 5) ? wsum[(tid >> 5) - 1] : 0u);
  for (int i = b; i < e; ++i) {
    const uint32_t t = v[i];
    v[i] = run;
    run += t;
  }
}

// stable scatter: rank of a key = #equal digits before it (global offset of its CTA + earlier
// warps of the CTA + earlier positions inside its warp segment, by __match_any_sync)
// SCANNED: `offsets` already holds the exclusive scan of the digit-major count matrix (large inputs);
// otherwise it holds the raw counts and each CTA derives its own offsets (one launch fewer — the
// matrix is a few thousand words for the row-id lists this path sees).
template <bool SCANNED>
__global__ void __launch_bounds__(kSortThreads)
radix_scatter_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ pay_in,
                     uint32_t* __restrict__ keys_out, uint32_t* __restrict__ pay_out, int n, int shift,
                     const uint32_t* __restrict__ offsets, int nblocks) {
  __shared__ uint32_t wcount[kSortThreads / 32][256];
  __shared__ uint32_t dscan[8];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  uint32_t my_offset;
  if (SCANNED) {
    my_offset = offsets[tid * nblocks + blockIdx.x];
  } else {
    // digit `tid`: keys with this digit in earlier CTAs, and in all CTAs
    uint32_t before = 0, total = 0;
    const uint32_t* row = offsets + tid * nblocks;
    for (int b = 0; b < nblocks; ++b) {
      const uint32_t c = row[b];
      total += c;
      if (b < static_cast<int>(blockIdx.x)) before += c;
    }
    // exclusive scan of `total` over the 256 digits
    uint32_t x = total;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) dscan[warp] = x;
    __syncthreads();
    uint32_t base = x - total;
    for (int w = 0; w < warp; ++w) base += dscan[w];
    my_offset = base + before;
  }
  for (int i = tid; i < (kSortThreads / 32) * 256; i += kSortThreads) (&wcount[0][0])[i] = 0;
  __syncthreads();
  constexpr int kRounds = kSortTile / kSortThreads;   // 16
  const int base = blockIdx.x * kSortTile + warp * (kRounds * 32);
  uint32_t key[kRounds];
  uint16_t rank[kRounds];
  const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
  for (int r = 0; r < kRounds; ++r) {
    const int i = base + r * 32 + lane;
    const bool valid = i < n;
    key[r] = valid ? keys_in[i] : 0u;
    const uint32_t d = valid ? ((key[r] >> shift) & 255u) : 256u;   // invalid lanes match each other only
    const uint32_t mask = __match_any_sync(0xffffffffu, d);
    const uint32_t before = __popc(mask & lt);
    uint32_t prev = 0;
    if (valid) prev = wcount[warp][d];
    __syncwarp();
    if (valid && before == 0) wcount[warp][d] = prev + __popc(mask);
    __syncwarp();
    rank[r] = static_cast<uint16_t>(prev + before);
  }
  __syncthreads();
  {  // digit `tid`: exclusive prefix over the warps, seeded with this CTA's global offset
    uint32_t run = my_offset;
#pragma unroll
    for (int w = 0; w < kSortThreads / 32; ++w) {
      const uint32_t t = wcount[w][tid];
      wcount[w][tid] = run;
      run += t;
    }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < kRounds; ++r) {
    const int i = base + r * 32 + lane;
    if (i < n) {
      const uint32_t pos = wcount[warp][(key[r] >> shift) & 255u] + rank[r];
      keys_out[pos] = key[r];
      pay_out[pos] = pay_in[i];
    }
  }
}

// ---- unique: head-flag scan ------------------------------------------------------------------------
__device__ __forceinline__ bool is_head(const uint32_t* keys, int i) {
  return i == 0 || keys[i] != keys[i - 1];
}

__global__ void __launch_bounds__(256)
head_count_kernel(const uint32_t* __restrict__ keys, int n, uint32_t* __restrict__ bsum) {
  __shared__ uint32_t ws[8];
  const int base = blockIdx.x * kScanTile;
  uint32_t c = 0;
  for (int j = threadIdx.x; j < kScanTile; j += 256) {
    const int i = base + j;
    if (i < n && is_head(keys, i)) ++c;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_down_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (int w = 0; w < 8; ++w) t += ws[w];
    bsum[blockIdx.x] = t;
  }
}

// after scan_small_kernel(bsum): bsum[b] = heads before tile b; count[0] = total number of heads
template <bool SCANNED>
__global__ void __launch_bounds__(256)
head_write_kernel(const uint32_t* __restrict__ keys, int n, const uint32_t* __restrict__ bsum,
                  int64_t* __restrict__ uniq, int32_t* __restrict__ seg_start, int64_t* __restrict__ count,
                  int nb) {
  __shared__ uint32_t ws[8];
  __shared__ uint32_t tile_base;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (SCANNED) {
    if (tid == 0) tile_base = bsum[blockIdx.x];
  } else {
    uint32_t part = 0;
    for (int b = tid; b < static_cast<int>(blockIdx.x); b += 256) part += bsum[b];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) part += __shfl_down_sync(0xffffffffu, part, o);
    if (lane == 0) ws[warp] = part;
    __syncthreads();
    if (tid == 0) {
      uint32_t t = 0;
      for (int w = 0; w < 8; ++w) t += ws[w];
      tile_base = t;
    }
  }
  __syncthreads();
  const int i0 = blockIdx.x * kScanTile + tid * 4;   // 4 consecutive keys per thread
  bool h[4];
  uint32_t c = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    h[j] = (i0 + j < n) && is_head(keys, i0 + j);
    c += h[j];
  }
  uint32_t x = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  if (lane == 31) ws[warp] = x;
  __syncthreads();
  uint32_t pos = tile_base + x - c;
  for (int w = 0; w < warp; ++w) pos += ws[w];
#pragma unroll
  for (int j = 0; j < 4; ++j)
    if (h[j]) {
      uniq[pos] = static_cast<int64_t>(keys[i0 + j]);
      seg_start[pos] = i0 + j;
      ++pos;
    }
  if (blockIdx.x == nb - 1 && tid == 255) {   // last thread of the last tile knows the total
    seg_start[pos] = n;
    count[0] = pos;
  }
}

// ---- segmented row reduce (+ optional lazy update) ---------------------------------------------------
enum RspMode { kRspStore = 0, kRspSgd = 1, kRspAdam = 2, kRspAdagrad = 3 };

struct RspUpdate {
  int mode;
  float lr, wd, momentum, beta1, beta2, eps, rescale, clip;
  int has_mom;
  int std_update;   // standard (non-lazy) update: every row is updated, Adam squares the whole gradient term
  float* table;  // [num_rows, row_len]
  float* s0;
  float* s1;
};

// the row update of one element held in registers: (w, s0, s1) <- gradient g
template <int MODE>
__device__ __forceinline__ void rsp_update_val(const RspUpdate& up, float& w, float& s0, float& s1, float g) {
  if (MODE == kRspSgd) {
    if (up.has_mom) {
      // SGDMomDnsRspDnsKernel (optimizer_op-inl.h:747-774): rate = lr*wd;
      // mom = momentum*mom - rate*w - lr*clip(rescale*g) | lr*rescale*g;  w += mom
      float mom = s0;
      const float a = __fmul_rn(up.momentum, mom);
      const float bb = __fmul_rn(__fmul_rn(up.lr, up.wd), w);
      const float cc = up.clip >= 0.f ? __fmul_rn(up.lr, clipf(__fmul_rn(up.rescale, g), up.clip))
                                      : __fmul_rn(__fmul_rn(up.lr, up.rescale), g);
      mom = __fsub_rn(__fsub_rn(a, bb), cc);
      s0 = mom;
      w = __fadd_rn(w, mom);
    } else {
      // SGDDnsRspKernel (optimizer_op-inl.h:452-475): (1 - lr*wd)*w - lr*clip(rescale*g) | (lr*rescale)*g
      // — NOT the multi-tensor op's mom = 0 form, which rounds differently
      const float keep = __fmul_rn(__fsub_rn(1.f, __fmul_rn(up.lr, up.wd)), w);
      const float step = up.clip >= 0.f ? __fmul_rn(up.lr, clipf(__fmul_rn(up.rescale, g), up.clip))
                                        : __fmul_rn(__fmul_rn(up.lr, up.rescale), g);
      w = __fsub_rn(keep, step);
    }
  } else if (MODE == kRspAdagrad) {
    // AdagradDnsRspDnsKernel (optimizer_op-inl.h:2514-2539): h += g'^2; w -= g'/sqrt(h+eps) * lr
    float gr = __fmul_rn(g, up.rescale);
    if (up.clip >= 0.f) gr = clipf(gr, up.clip);
    const float h = __fadd_rn(s0, __fmul_rn(gr, gr));
    s0 = h;
    const float div = __fdiv_rn(gr, __fsqrt_rn(__fadd_rn(h, up.eps)));
    w = __fsub_rn(w, __fmul_rn(div, up.lr));
  } else if (MODE == kRspAdam) {
    // AdamDnsRspDnsKernel<cpu> (optimizer_op-inl.h:1352-1382): note (1-beta2)*square(clip(g)) with
    // clipping but (1-beta2)*g*g (left to right) without
    float m = s0, v = s1;
    float gr = __fadd_rn(__fmul_rn(g, up.rescale), __fmul_rn(w, up.wd));
    const float omb1 = __fsub_rn(1.f, up.beta1), omb2 = __fsub_rn(1.f, up.beta2);
    if (up.clip >= 0.f) {
      gr = clipf(gr, up.clip);
      m = __fadd_rn(__fmul_rn(up.beta1, m), __fmul_rn(omb1, gr));
      v = __fadd_rn(__fmul_rn(up.beta2, v), __fmul_rn(omb2, __fmul_rn(gr, gr)));
    } else if (up.std_update) {   // AdamStdDnsRspDnsKernel (optimizer_op.cc:196-229): (1-beta2)*square(g')
      m = __fadd_rn(__fmul_rn(up.beta1, m), __fmul_rn(omb1, gr));
      v = __fadd_rn(__fmul_rn(up.beta2, v), __fmul_rn(omb2, __fmul_rn(gr, gr)));
    } else {
      m = __fadd_rn(__fmul_rn(up.beta1, m), __fmul_rn(omb1, gr));
      v = __fadd_rn(__fmul_rn(up.beta2, v), __fmul_rn(__fmul_rn(omb2, gr), gr));
    }
    s0 = m;
    s1 = v;
    w = __fsub_rn(w, __fdiv_rn(__fmul_rn(up.lr, m), __fadd_rn(__fsqrt_rn(v), up.eps)));
  }
}

// W consecutive elements (W = 4: one 128-bit load / store per array) of a row at table offset o
template <int MODE, int W>
__device__ __forceinline__ void rsp_update_vec(const RspUpdate& up, int64_t o, const float (&g)[W]) {
  constexpr bool kS0 = MODE == kRspSgd || MODE == kRspAdam || MODE == kRspAdagrad;
  constexpr bool kS1 = MODE == kRspAdam;
  const bool use_s0 = kS0 && (MODE != kRspSgd || up.has_mom);
  float w[W], s0[W], s1[W];
  if constexpr (W == 4) {
    const float4 a = *reinterpret_cast<const float4*>(up.table + o);
    w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w;
    if (use_s0) { const float4 b = *reinterpret_cast<const float4*>(up.s0 + o); s0[0] = b.x; s0[1] = b.y; s0[2] = b.z; s0[3] = b.w; }
    if (kS1) { const float4 c = *reinterpret_cast<const float4*>(up.s1 + o); s1[0] = c.x; s1[1] = c.y; s1[2] = c.z; s1[3] = c.w; }
  } else {
#pragma unroll
    for (int j = 0; j < W; ++j) {
      w[j] = up.table[o + j];
      if (use_s0) s0[j] = up.s0[o + j];
      if (kS1) s1[j] = up.s1[o + j];
    }
  }
#pragma unroll
  for (int j = 0; j < W; ++j) {
    if (!use_s0) s0[j] = 0.f;
    if (!kS1) s1[j] = 0.f;
    rsp_update_val<MODE>(up, w[j], s0[j], s1[j], g[j]);
  }
  if constexpr (W == 4) {
    *reinterpret_cast<float4*>(up.table + o) = make_float4(w[0], w[1], w[2], w[3]);
    if (use_s0) *reinterpret_cast<float4*>(up.s0 + o) = make_float4(s0[0], s0[1], s0[2], s0[3]);
    if (kS1) *reinterpret_cast<float4*>(up.s1 + o) = make_float4(s1[0], s1[1], s1[2], s1[3]);
  } else {
#pragma unroll
    for (int j = 0; j < W; ++j) {
      up.table[o + j] = w[j];
      if (use_s0) up.s0[o + j] = s0[j];
      if (kS1) up.s1[o + j] = s1[j];
    }
  }
}

__device__ __forceinline__ const float* rsp_row(const RspInputs& in, uint32_t p, int64_t row_len) {
  int k = 0;
  while (k + 1 < in.n && static_cast<int64_t>(p) >= in.start[k + 1]) ++k;
  return in.val[k] + (static_cast<int64_t>(p) - in.start[k]) * row_len;
}

// one CTA per output row; threads stride over the row in float4 (row_len % 4 == 0) or scalars
template <int MODE, bool VEC>
__global__ void __launch_bounds__(128)
rsp_reduce_kernel(const RspInputs in, const uint32_t* __restrict__ pay, const int32_t* __restrict__ seg_start,
                  const int64_t* __restrict__ uniq, const int64_t* __restrict__ count, int64_t row_len,
                  float* __restrict__ out_val, const RspUpdate up) {
  const int64_t nrows = count[0];
  for (int64_t r = blockIdx.x; r < nrows; r += gridDim.x) {
    const int b = seg_start[r], e = seg_start[r + 1];
    const int64_t row = uniq[r];
    constexpr int W = VEC ? 4 : 1;
    for (int64_t c = static_cast<int64_t>(threadIdx.x) * W; c < row_len; c += 128 * W) {
      float acc[W];
#pragma unroll
      for (int j = 0; j < W; ++j) acc[j] = 0.f;   // out = 0, then += every source row in order
      for (int s = b; s < e; ++s) {
        const float* src = rsp_row(in, pay[s], row_len) + c;
        float x[W];
        if constexpr (VEC) {
          const float4 v = *reinterpret_cast<const float4*>(src);
          x[0] = v.x; x[1] = v.y; x[2] = v.z; x[3] = v.w;
        } else {
          x[0] = src[0];
        }
#pragma unroll
        for (int j = 0; j < W; ++j) acc[j] = __fadd_rn(acc[j], x[j]);
      }
      if (MODE == kRspStore) {
#pragma unroll
        for (int j = 0; j < W; ++j) out_val[r * row_len + c + j] = acc[j];
      } else {
        const int64_t o = row * row_len + c;
        rsp_update_vec<MODE, W>(up, o, acc);
      }
    }
  }
}

// =====================================================================================================
// Bitmap path (tables up to 2^28 rows): the sorted unique set of row ids is a BITMAP over the row range
// — 1M rows are 128 KB, resident in L2 — so "sort + unique" becomes  mark (atomicOr)  ->  one chained
// popcount scan that also emits the ids in ascending order and clears the bitmap behind itself.
// Pushed copies are sorted-unique already (row-sparse invariant), so the rows of an output id are
// found by one binary search per copy inside the reduce kernel: no sort, no payload array.
//   push, one copy   : 1 kernel  (the copy IS the merged index set)
//   push, C copies   : mark + compact + reduce  = 3 kernels
//   row_sparse_pull  : mark + compact + gather  = 3 kernels
// (the radix-sort path above needs 10 / 20 launches and remains for tables beyond 2^28 rows)
// =====================================================================================================
constexpr int kMaxSources = 64;       // copies x ranks that one owner may reduce
constexpr int kCompactWords = 1024;   // bitmap words per CTA of the compaction kernel (256 threads x 4)
enum RspErr { kRspErrRange = 1, kRspErrOrder = 2 };

struct RspSources {
  const void* idx[kMaxSources];    // int64 row ids, sorted strictly ascending
  const float* val[kMaxSources];   // [len, row_len]
  int64_t len[kMaxSources];
  // when set, the length is read from device memory instead (a peer rank's header word, written by
  // that rank before the cross-GPU barrier that precedes the kernel)
  const int64_t* len_ptr[kMaxSources];
  int n;
};
__device__ __forceinline__ int64_t rsp_len(const RspSources& in, int k) {
  return in.len_ptr[k] != nullptr ? __ldcv(in.len_ptr[k]) : in.len[k];
}

// ids -> bitmap bits for the rows this rank owns: ids row_lo + j*stride, j in [0, nlocal) (one GPU: all
// rows; several ranks: the rows with id % W == rank — a CYCLIC partition, so that skewed (Zipf) ids
// spread evenly over the owners).  I: element type of the id arrays.
template <typename I, bool CHECK_SORTED>
__global__ void rsp_mark_kernel(const RspSources in, int64_t num_rows, int64_t row_lo, int64_t stride,
                                uint32_t* __restrict__ bitmap, uint32_t* __restrict__ err) {
  const int64_t gstride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int k = 0; k < in.n; ++k) {
    const I* ids = static_cast<const I*>(in.idx[k]);
    const int64_t len = rsp_len(in, k);
    // the trip count is the same for every lane of a warp (i0 is the warp's lane-0 index), so the
    // full-mask warp primitives below are legal
    for (int64_t i0 = blockIdx.x * static_cast<int64_t>(blockDim.x) + (threadIdx.x & ~31); i0 < len;
         i0 += gstride) {
      const int64_t i = i0 + (threadIdx.x & 31);
      bool mine = i < len;
      int64_t b = 0;
      if (mine) {
        const int64_t id = static_cast<int64_t>(ids[i]);
        if (id < 0 || id >= num_rows) {
          atomicOr(err, static_cast<uint32_t>(kRspErrRange));
          mine = false;
        } else {
          if (CHECK_SORTED && i > 0 && static_cast<int64_t>(ids[i - 1]) >= id)
            atomicOr(err, static_cast<uint32_t>(kRspErrOrder));
          mine = id >= row_lo && (id - row_lo) % stride == 0;
          b = (id - row_lo) / stride;
        }
      }
      // duplicate-heavy id lists (Zipf) would hammer a few words with atomics: lanes that hit the same
      // word merge their bits, and a word that already holds them is left alone
      const uint32_t word = mine ? static_cast<uint32_t>(b >> 5) : 0xffffffffu;
      const uint32_t bit = mine ? (1u << (b & 31)) : 0u;
      const uint32_t peers = __match_any_sync(0xffffffffu, word);
      const uint32_t bits = __reduce_or_sync(peers, bit);
      if (mine && (__ffs(peers) - 1) == static_cast<int>(threadIdx.x & 31)) {
        const uint32_t have = *reinterpret_cast<volatile uint32_t*>(bitmap + word);
        if ((have & bits) != bits) atomicOr(&bitmap[word], bits);
      }
    }
  }
}

// One-pass scan over the bitmap.  chain: 2 words per CTA {its count, epoch}; `epoch` differs on every
// launch, so nothing is ever reset.
// Emits the ids in ascending order, writes the total to count[0] and zeroes the bitmap.
__global__ void __launch_bounds__(256)
rsp_compact_kernel(uint32_t* __restrict__ bitmap, int nwords, int64_t row_lo, int64_t stride,
                   int64_t* __restrict__ uniq,
                   int64_t* __restrict__ count, volatile uint32_t* chain, uint32_t epoch) {
  __shared__ uint32_t ws[8];
  __shared__ uint32_t s_prefix;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int w0 = blockIdx.x * kCompactWords + tid * 4;
  uint32_t w[4] = {0u, 0u, 0u, 0u};
  if (w0 < nwords) {   // the bitmap allocation is padded to whole CTAs
    const uint4 v = *reinterpret_cast<const uint4*>(bitmap + w0);
    w[0] = v.x; w[1] = v.y; w[2] = v.z; w[3] = v.w;
    if (v.x | v.y | v.z | v.w) *reinterpret_cast<uint4*>(bitmap + w0) = make_uint4(0u, 0u, 0u, 0u);
  }
  const uint32_t c = __popc(w[0]) + __popc(w[1]) + __popc(w[2]) + __popc(w[3]);
  uint32_t x = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  if (lane == 31) ws[warp] = x;
  __syncthreads();
  uint32_t before = x - c, total = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    if (k < warp) before += ws[k];
    total += ws[k];
  }
  // publish this CTA's own count at once, then add up the counts of all CTAs before it: nobody waits
  // for a prefix, only for counts that every CTA publishes right after loading its words — no serial
  // chain (lower-numbered CTAs are scheduled first, so the waits always resolve)
  if (tid == 0) {
    chain[2 * blockIdx.x] = total;
    __threadfence();
    chain[2 * blockIdx.x + 1] = epoch;
  }
  uint32_t part = 0;
  for (int j = tid; j < static_cast<int>(blockIdx.x); j += 256) {
    while (chain[2 * j + 1] != epoch) {}
    __threadfence();
    part += chain[2 * j];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) part += __shfl_down_sync(0xffffffffu, part, o);
  __syncthreads();            // everyone has read ws[] of the first scan
  if (lane == 0) ws[warp] = part;
  __syncthreads();
  if (tid == 0) {
    uint32_t p = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) p += ws[k];
    s_prefix = p;
    if (blockIdx.x == gridDim.x - 1) count[0] = static_cast<int64_t>(p) + total;
  }
  __syncthreads();
  int64_t pos = static_cast<int64_t>(s_prefix) + before;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint32_t bits = w[j];
    const int64_t base = (static_cast<int64_t>(w0) + j) * 32;
    while (bits) {
      const int b = __ffs(bits) - 1;
      uniq[pos++] = row_lo + (base + b) * stride;
      bits &= bits - 1;
    }
  }
}

// One CTA per output row.  DIRECT: a single pushed copy — its index list is the merged set and source
// row r is output row r.  Otherwise the row's id is looked up in every source by binary search
// (thread k searches source k); rows are added in source order = the reference's device order
// (ndarray_function.cc:81-113).  `up.table` / states hold this rank's rows (id = row_lo + j*stride) densely.
// MODE kRspStore: merged rows go to out_val[r], ids to out_idx[r], the row count to out_nnr.
template <int MODE, bool VEC, bool DIRECT>
__global__ void __launch_bounds__(128)
rsp_reduce2_kernel(const RspSources in, const int64_t* __restrict__ uniq, const int64_t* __restrict__ count,
                   int64_t row_len, int64_t num_rows, int64_t row_lo, int64_t stride, float* __restrict__ out_val,
                   int64_t* __restrict__ out_idx, int64_t* __restrict__ out_nnr, const RspUpdate up,
                   uint32_t* __restrict__ err) {
  __shared__ int64_t s_pos[kMaxSources];
  const int64_t nrows = DIRECT ? rsp_len(in, 0) : count[0];
  if (MODE == kRspStore && blockIdx.x == 0 && threadIdx.x == 0 && out_nnr != nullptr) out_nnr[0] = nrows;
  constexpr int W = VEC ? 4 : 1;
  for (int64_t r = blockIdx.x; r < nrows; r += gridDim.x) {
    int64_t row;
    if (DIRECT) {
      const int64_t* ids = static_cast<const int64_t*>(in.idx[0]);
      row = ids[r];
      bool bad = row < 0 || row >= num_rows;
      if (bad && threadIdx.x == 0) atomicOr(err, static_cast<uint32_t>(kRspErrRange));
      if (r > 0 && ids[r - 1] >= row) {   // duplicates would race on the table row
        if (threadIdx.x == 0) atomicOr(err, static_cast<uint32_t>(kRspErrOrder));
        bad = true;
      }
      if (bad) continue;   // (the direct form is only used when this rank owns every row)
    } else {
      row = uniq[r];
      __syncthreads();   // the previous row's readers are done with s_pos
      for (int k = threadIdx.x; k < in.n; k += 128) {
        const int64_t* ids = static_cast<const int64_t*>(in.idx[k]);
        int64_t lo = 0, hi = rsp_len(in, k) - 1, found = -1;
        while (lo <= hi) {
          const int64_t m = lo + (hi - lo) / 2;
          const int64_t v = ids[m];
          if (v == row) { found = m; break; }
          if (v < row) lo = m + 1; else hi = m - 1;
        }
        s_pos[k] = found;
      }
      __syncthreads();
    }
    if (MODE == kRspStore && threadIdx.x == 0) out_idx[r] = row;
    for (int64_t c = static_cast<int64_t>(threadIdx.x) * W; c < row_len; c += 128 * W) {
      float acc[W];
#pragma unroll
      for (int j = 0; j < W; ++j) acc[j] = 0.f;   // out = 0, then += every source row in order
      for (int k = 0; k < in.n; ++k) {
        const int64_t p = DIRECT ? r : s_pos[k];
        if (p < 0) continue;
        const float* src = in.val[k] + p * row_len + c;
        if constexpr (VEC) {
          const float4 v = *reinterpret_cast<const float4*>(src);
          acc[0] = __fadd_rn(acc[0], v.x); acc[1] = __fadd_rn(acc[1], v.y);
          acc[2] = __fadd_rn(acc[2], v.z); acc[3] = __fadd_rn(acc[3], v.w);
        } else {
          acc[0] = __fadd_rn(acc[0], src[0]);
        }
      }
      if (MODE == kRspStore) {
#pragma unroll
        for (int j = 0; j < W; ++j) out_val[r * row_len + c + j] = acc[j];
      } else {
        const int64_t o = ((row - row_lo) / stride) * row_len + c;
        rsp_update_vec<MODE, W>(up, o, acc);
      }
    }
  }
}

// Standard (non-lazy) update, SGDMomStdDnsRspDnsKernel / AdamStdDnsRspDnsKernel (optimizer_op.cc:108-286,
// optimizer_op.cu:32-223): EVERY row of this rank's shard is updated; a row that is absent from the
// merged gradient (midx sorted unique, nnr rows, values mval) takes part with a zero gradient.
template <int MODE, bool VEC>
__global__ void __launch_bounds__(128)
rsp_std_update_kernel(const int64_t* __restrict__ midx, const int64_t* __restrict__ nnr_ptr,
                      const float* __restrict__ mval, int64_t row_len, int64_t row_lo, int64_t stride, int64_t nlocal,
                      const RspUpdate up) {
  __shared__ int64_t s_pos;
  const int64_t nnr = nnr_ptr[0];
  constexpr int W = VEC ? 4 : 1;
  for (int64_t jrow = blockIdx.x; jrow < nlocal; jrow += gridDim.x) {
    const int64_t row = row_lo + jrow * stride;
    __syncthreads();
    if (threadIdx.x == 0) {
      int64_t lo = 0, hi = nnr - 1, found = -1;
      while (lo <= hi) {
        const int64_t m = lo + (hi - lo) / 2;
        const int64_t v = midx[m];
        if (v == row) { found = m; break; }
        if (v < row) lo = m + 1; else hi = m - 1;
      }
      s_pos = found;
    }
    __syncthreads();
    const int64_t p = s_pos;
    for (int64_t c = static_cast<int64_t>(threadIdx.x) * W; c < row_len; c += 128 * W) {
      float g[W];
#pragma unroll
      for (int j = 0; j < W; ++j) g[j] = 0.f;
      if (p >= 0) {
        if constexpr (VEC) {
          const float4 v = *reinterpret_cast<const float4*>(mval + p * row_len + c);
          g[0] = v.x; g[1] = v.y; g[2] = v.z; g[3] = v.w;
        } else {
          g[0] = mval[p * row_len + c];
        }
      }
      const int64_t o = jrow * row_len + c;
      rsp_update_vec<MODE, W>(up, o, g);
    }
  }
}

// ---- row gather for row_sparse_pull ------------------------------------------------------------------
// full != 0: the store holds every row (dense table); otherwise (src_idx, src_val) is a compact
// row-sparse value and absent rows come back as zeros.
// the table may be partitioned cyclically over W ranks: row id lives in t[id % W] (peer memory for the
// other ranks) at local row id / W; one GPU: W = 1, t[0] holds everything
struct RspTables {
  const float* t[kMaxCopies];
  int64_t nranks;
};

template <bool VEC>
__global__ void __launch_bounds__(128)
rsp_gather_kernel(const int64_t* __restrict__ uniq, const int64_t* __restrict__ count, int64_t row_len,
                  int full, const RspTables tabs, int64_t num_rows,
                  const int64_t* __restrict__ src_idx, const float* __restrict__ src_val,
                  const int64_t* __restrict__ src_nnr, int64_t* __restrict__ out_idx,
                  float* __restrict__ out_val) {
  const int64_t nrows = count[0];
  for (int64_t r = blockIdx.x; r < nrows; r += gridDim.x) {
    const int64_t id = uniq[r];
    const float* src = nullptr;
    if (full) {
      if (id >= 0 && id < num_rows) {
        const int64_t o = id % tabs.nranks;
        src = tabs.t[o] + (id / tabs.nranks) * row_len;
      }
    } else {
      int64_t lo = 0, hi = src_nnr[0] - 1;
      while (lo <= hi) {
        const int64_t m = lo + (hi - lo) / 2;
        const int64_t v = src_idx[m];
        if (v == id) { src = src_val + m * row_len; break; }
        if (v < id) lo = m + 1; else hi = m - 1;
      }
    }
    if (threadIdx.x == 0) out_idx[r] = id;
    float* dst = out_val + r * row_len;
    if constexpr (VEC) {
      for (int64_t c = static_cast<int64_t>(threadIdx.x) * 4; c < row_len; c += 512)
        *reinterpret_cast<float4*>(dst + c) =
            src ? *reinterpret_cast<const float4*>(src + c) : make_float4(0.f, 0.f, 0.f, 0.f);
    } else {
      for (int64_t c = threadIdx.x; c < row_len; c += 128) dst[c] = src ? src[c] : 0.f;
    }
  }
}

__global__ void copy_i64_kernel(const int64_t* __restrict__ src, int64_t* __restrict__ dst,
                                const int64_t* __restrict__ count) {
  const int64_t n = count[0];
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x)
    dst[i] = src[i];
}

}  // namespace b200kv
