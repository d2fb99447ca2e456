// Bit-exact jagged / sparse integer ops for sm_100a.
//
// Replaces the torch.ops.fbgemm.* integer kernels torchrec calls
// (fbgemm_gpu/src/sparse_ops.cu): asynchronous_{complete,inclusive,exclusive}_cumsum :189-300,
// permute_2D_sparse_data :309-486, permute_1D_sparse_data :490-665,
// expand_into_jagged_permute :671-730, block_bucketize_sparse_features :736-1228,
// offsets_range :62-187, and bounds_check_indices (codegen/embedding_bounds_check.cu).
//
// Design: every op that needs a prefix sum is ONE cooperative launch sized to the machine
// (<= 148 SMs x resident CTAs) with grid-wide syncs between phases, instead of the
// reference's kernel + CUB-scan + kernel chains; scans are reduce-then-scan over contiguous
// per-CTA chunks, so no spin-waiting between CTAs is ever needed.  Copy kernels pick the
// number of lanes per jagged segment from the average segment length so that L = 1 inputs
// (Criteo) are coalesced thread-per-segment copies and long bags get a full warp.
#include <cooperative_groups.h>

#include <type_traits>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace b200 {

constexpr int kScanThreads = 512;
constexpr int kScanItems = 4;
constexpr int kMaxCoopBlocks = 2048;  // size of the block-sum scratch in `ws`

// ws layout (int64): [0, kMaxCoopBlocks) block sums A, [kMaxCoopBlocks, 2k) block sums B,
// then a few flags.
constexpr size_t kWsFlagOffset = 2 * kMaxCoopBlocks;
constexpr size_t kWsInt64s = kWsFlagOffset + 16;

template <typename T>
__device__ __forceinline__ T ld_as_i64(const void* p, int64_t i);

struct I32 {
  using type = int32_t;
};
struct I64 {
  using type = int64_t;
};

// ---- block-wide helpers -------------------------------------------------------------
__device__ __forceinline__ int64_t block_reduce_sum(int64_t v, int64_t* s_warp) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) s_warp[warp] = v;
  __syncthreads();
  int64_t r = 0;
  for (int w = 0; w < (int)(blockDim.x >> 5); ++w) r += s_warp[w];
  return r;
}

// Exclusive scan of one value per thread across the block; returns the exclusive prefix of
// this thread and the block total through `total`.  The warp totals are scanned once by warp 0
// (not summed by every thread): the scan kernels are instruction-bound, not HBM-bound.
__device__ __forceinline__ int64_t block_exclusive_scan(int64_t v, int64_t* s_warp, int64_t& total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nwarps = (int)(blockDim.x >> 5);
  int64_t inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int64_t n = __shfl_up_sync(kFull, inc, o);
    if (lane >= o) inc += n;
  }
  __syncthreads();  // s_warp may still be read by the previous call
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    const int64_t x = lane < nwarps ? s_warp[lane] : 0;
    int64_t winc = x;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int64_t n = __shfl_up_sync(kFull, winc, o);
      if (lane >= o) winc += n;
    }
    if (lane < nwarps) s_warp[lane] = winc - x;  // exclusive prefix of warp `lane`
    if (lane == 31) s_warp[nwarps] = winc;       // block total
  }
  __syncthreads();
  total = s_warp[nwarps];
  return s_warp[warp] + inc - v;
}

struct Chunk {
  int64_t begin, end;
};
__device__ __forceinline__ Chunk block_chunk(int64_t n) {
  // contiguous chunk of this CTA, a multiple of the scan tile so tiles never straddle CTAs
  const int64_t tile = (int64_t)kScanThreads * kScanItems;
  const int64_t tiles = (n + tile - 1) / tile;
  const int64_t per = (tiles + gridDim.x - 1) / gridDim.x;
  Chunk c;
  c.begin = min(n, (int64_t)blockIdx.x * per * tile);
  c.end = min(n, c.begin + per * tile);
  return c;
}

// Phase 1 of a grid scan: sum of this CTA's chunk -> sums[blockIdx.x].
template <typename LoadF>
__device__ __forceinline__ void scan_reduce_phase(LoadF load, int64_t n, int64_t* sums,
                                                  int64_t* s_warp) {
  const Chunk c = block_chunk(n);
  int64_t v = 0;
  for (int64_t i = c.begin + threadIdx.x; i < c.end; i += kScanThreads) v += load(i);
  const int64_t r = block_reduce_sum(v, s_warp);
  if (threadIdx.x == 0) sums[blockIdx.x] = r;
}

// Phase 2 (after a grid sync): exclusive scan of the chunk, store(i, exclusive_prefix, value).
// Returns the grand total (valid in every thread).
template <typename LoadF, typename StoreF>
__device__ __forceinline__ int64_t scan_finish_phase(LoadF load, StoreF store, int64_t n,
                                                     const int64_t* sums, int64_t* s_warp) {
  int64_t before = 0, all = 0;
  for (int bidx = threadIdx.x; bidx < (int)gridDim.x; bidx += kScanThreads) {
    const int64_t x = sums[bidx];
    if (bidx < (int)blockIdx.x) before += x;
    all += x;
  }
  const int64_t base0 = block_reduce_sum(before, s_warp);
  const int64_t grand = block_reduce_sum(all, s_warp);
  const Chunk c = block_chunk(n);
  int64_t running = base0;
  // Warp-striped arrangement: warp w owns the contiguous 32*kScanItems elements
  // [tile0 + w*32*kScanItems, ...), lane l holds elements seg + k*32 + l.  Loads and stores are
  // fully coalesced and the scan costs kScanItems warp scans + one scan of the warp totals.
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int64_t tile0 = c.begin; tile0 < c.end; tile0 += (int64_t)kScanThreads * kScanItems) {
    const int64_t seg = tile0 + (int64_t)warp * (32 * kScanItems) + lane;
    int64_t x[kScanItems], pre[kScanItems];
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) x[k] = (seg + 32 * k < c.end) ? load(seg + 32 * k) : 0;
    int64_t wrun = 0;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) {
      int64_t inc = x[k];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int64_t nb = __shfl_up_sync(kFull, inc, o);
        if (lane >= o) inc += nb;
      }
      pre[k] = wrun + inc - x[k];
      wrun += __shfl_sync(kFull, inc, 31);
    }
    // exclusive scan of the warp totals (one value per warp, held by every lane)
    __syncthreads();  // s_warp may still be read by the previous round
    if (lane == 0) s_warp[warp] = wrun;
    __syncthreads();
    if (warp == 0) {
      const int nwarps = kScanThreads / 32;
      const int64_t t = lane < nwarps ? s_warp[lane] : 0;
      int64_t winc = t;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int64_t nb = __shfl_up_sync(kFull, winc, o);
        if (lane >= o) winc += nb;
      }
      if (lane < nwarps) s_warp[lane] = winc - t;
      if (lane == 31) s_warp[nwarps] = winc;
    }
    __syncthreads();
    const int64_t wbase = running + s_warp[warp];
    const int64_t tile_total = s_warp[kScanThreads / 32];
#pragma unroll
    for (int k = 0; k < kScanItems; ++k)
      if (seg + 32 * k < c.end) store(seg + 32 * k, wbase + pre[k], x[k]);
    running += tile_total;
  }
  return grand;
}

static int coop_grid_for(const void* kernel, int64_t n) {
  static int sms = 0;
  int dev = 0;
  cudaGetDevice(&dev);
  if (sms == 0) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  int per_sm = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kScanThreads, 0);
  if (per_sm < 1) per_sm = 1;
  int64_t cap = (int64_t)sms * per_sm;
  if (cap > kMaxCoopBlocks) cap = kMaxCoopBlocks;
  int64_t want = ceil_div(n > 0 ? n : 1, (int64_t)kScanThreads * kScanItems);
  if (want < 1) want = 1;
  return (int)(want < cap ? want : cap);
}

// ---------------------------------------------------------------- cumsum
template <typename T>
__global__ void __launch_bounds__(kScanThreads)
cumsum_kernel(const T* __restrict__ in, T* __restrict__ out, int64_t n, int mode,
              int64_t* __restrict__ ws) {
  __shared__ int64_t s_warp[kScanThreads / 32 + 1];
  cg::grid_group grid = cg::this_grid();
  auto load = [&](int64_t i) -> int64_t { return (int64_t)in[i]; };
  scan_reduce_phase(load, n, ws, s_warp);
  grid.sync();
  const int64_t total = scan_finish_phase(
      load,
      [&](int64_t i, int64_t pre, int64_t v) {
        out[i] = (T)(mode == B200_CUMSUM_INCLUSIVE ? pre + v : pre);
      },
      n, ws, s_warp);
  if (mode == B200_CUMSUM_COMPLETE && blockIdx.x == 0 && threadIdx.x == 0) out[n] = (T)total;
}

// ---------------------------------------------------------------- permute (lengths + scans)
template <typename T>
__global__ void __launch_bounds__(kScanThreads)
permute_lengths_kernel(const int32_t* __restrict__ permute, const T* __restrict__ lengths,
                       int32_t T_in, int32_t Tp, int32_t B, T* __restrict__ out_lengths,
                       int64_t* __restrict__ in_offsets, int64_t* __restrict__ out_offsets,
                       int64_t* __restrict__ ws) {
  __shared__ int64_t s_warp[kScanThreads / 32 + 1];
  cg::grid_group grid = cg::this_grid();
  const int64_t n_in = (int64_t)T_in * B, n_out = (int64_t)Tp * B;
  auto load_in = [&](int64_t i) -> int64_t { return (int64_t)lengths[i]; };
  auto load_out = [&](int64_t i) -> int64_t {
    const int64_t tp = i / B;
    return (int64_t)lengths[(int64_t)permute[tp] * B + (i - tp * B)];
  };
  scan_reduce_phase(load_in, n_in, ws, s_warp);
  scan_reduce_phase(load_out, n_out, ws + kMaxCoopBlocks, s_warp);
  grid.sync();
  const int64_t tot_in = scan_finish_phase(
      load_in, [&](int64_t i, int64_t pre, int64_t) { in_offsets[i] = pre; }, n_in, ws, s_warp);
  const int64_t tot_out = scan_finish_phase(
      load_out,
      [&](int64_t i, int64_t pre, int64_t v) {
        out_offsets[i] = pre;
        out_lengths[i] = (T)v;
      },
      n_out, ws + kMaxCoopBlocks, s_warp);
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    in_offsets[n_in] = tot_in;
    out_offsets[n_out] = tot_out;
  }
}

template <int BYTES>
struct RawT;
template <>
struct RawT<1> {
  using type = uint8_t;
};
template <>
struct RawT<2> {
  using type = uint16_t;
};
template <>
struct RawT<4> {
  using type = uint32_t;
};
template <>
struct RawT<8> {
  using type = uint64_t;
};

// G lanes cooperate on one (tp, b) segment; G is a power of two <= 32 chosen by the host.
template <int VB, int WB>
__global__ void __launch_bounds__(256)
permute_data_kernel(const int32_t* __restrict__ permute, int32_t Tp, int32_t B, int G,
                    const int64_t* __restrict__ in_offsets, const int64_t* __restrict__ out_offsets,
                    const void* __restrict__ values, void* __restrict__ out_values,
                    const void* __restrict__ weights, void* __restrict__ out_weights,
                    int64_t out_capacity) {
  using V = typename RawT<VB>::type;
  using Wt = typename RawT<WB>::type;
  const int64_t n_seg = (int64_t)Tp * B;
  const int64_t gid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
  const int sub = threadIdx.x & (G - 1);
  const int64_t stride = ((int64_t)gridDim.x * blockDim.x) / G;
  for (int64_t seg = gid; seg < n_seg; seg += stride) {
    const int64_t tp = seg / B;
    const int64_t b = seg - tp * B;
    const int64_t src_seg = (int64_t)permute[tp] * B + b;
    const int64_t src = in_offsets[src_seg];
    const int64_t dst = out_offsets[seg];
    int64_t len = out_offsets[seg + 1] - dst;
    if (dst + len > out_capacity) len = out_capacity - dst;  // caller-supplied total too small
    for (int64_t k = sub; k < len; k += G) {
      reinterpret_cast<V*>(out_values)[dst + k] = reinterpret_cast<const V*>(values)[src + k];
      if (weights)
        reinterpret_cast<Wt*>(out_weights)[dst + k] = reinterpret_cast<const Wt*>(weights)[src + k];
    }
  }
}

static int pick_group(int64_t total, int64_t segments) {
  if (segments <= 0) return 1;
  const int64_t avg = total / segments;
  int g = 1;
  while (g < 32 && g < avg) g <<= 1;
  return g;
}

static unsigned grid_for_groups(int64_t segments, int G, int threads) {
  int64_t blocks = ceil_div(segments * G, threads);
  const int64_t cap = (int64_t)kNumSMs * 16;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (unsigned)blocks;
}

// ---------------------------------------------------------------- permute_2D, fused
// permute_2D_sparse_data moves whole feature rows: the B bags of input feature permute[tp] are ONE
// contiguous range of `values`, so only T per-feature totals are needed - not the T*B per-bag
// offsets the reference (and permute_lengths_kernel above) builds with two device-wide scans.
// One cooperative launch:  zero T sums | sync | per-feature length sums (warp per 1024-bag piece,
// one atomic each) | sync | every CTA scans the <= 1024 totals in shared memory, then copies
// out_lengths (row gather) and the values / weights (flat, binary search over Tp+1 offsets).
constexpr int kPermMaxFeatures = 1024;
constexpr int kPermPiece = 1024;

template <typename LT, int VB, int WB>
__global__ void __launch_bounds__(kScanThreads)
permute_2d_fused_kernel(const int32_t* __restrict__ permute, const LT* __restrict__ lengths, int32_t T,
                        int32_t Tp, int32_t B, LT* __restrict__ out_lengths,
                        const void* __restrict__ values, void* __restrict__ out_values,
                        const void* __restrict__ weights, void* __restrict__ out_weights,
                        int64_t out_capacity, int64_t* __restrict__ ws) {
  using V = typename RawT<VB>::type;
  using Wt = typename RawT<WB>::type;
  __shared__ int64_t s_in[kPermMaxFeatures + 1];
  __shared__ int64_t s_out[kPermMaxFeatures + 1];
  __shared__ int64_t s_warp[kScanThreads / 32 + 1];
  cg::grid_group grid = cg::this_grid();
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
  const int lane = threadIdx.x & 31;

  for (int64_t i = tid; i < T; i += nthreads) ws[i] = 0;
  grid.sync();

  // per-feature totals
  const int pieces_per_row = (int)ceil_div(B, kPermPiece);
  const int64_t n_pieces = (int64_t)T * pieces_per_row;
  for (int64_t pc = tid >> 5; pc < n_pieces; pc += nthreads >> 5) {
    const int t = (int)(pc / pieces_per_row);
    const int b_lo = (int)(pc - (int64_t)t * pieces_per_row) * kPermPiece;
    const int b_hi = min(B, b_lo + kPermPiece);
    const LT* __restrict__ row = lengths + (int64_t)t * B;
    int64_t v = 0;
    for (int b = b_lo + lane; b < b_hi; b += 32 * 8) {
      LT x[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) x[k] = (b + 32 * k < b_hi) ? row[b + 32 * k] : (LT)0;
#pragma unroll
      for (int k = 0; k < 8; ++k) v += (int64_t)x[k];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
    if (lane == 0 && v != 0) atomicAdd((unsigned long long*)&ws[t], (unsigned long long)v);
  }
  grid.sync();

  // feature offsets of the input and of the permuted output (every CTA computes its own copy)
  {
    const int i0 = 2 * threadIdx.x, i1 = i0 + 1;
    const int64_t a0 = i0 < T ? ws[i0] : 0, a1 = i1 < T ? ws[i1] : 0;
    int64_t tot;
    const int64_t pre = block_exclusive_scan(a0 + a1, s_warp, tot);
    if (i0 < T) s_in[i0] = pre;
    if (i1 < T) s_in[i1] = pre + a0;
    if (threadIdx.x == 0) s_in[T] = tot;
    const int64_t c0 = i0 < Tp ? ws[permute[i0]] : 0, c1 = i1 < Tp ? ws[permute[i1]] : 0;
    const int64_t pre2 = block_exclusive_scan(c0 + c1, s_warp, tot);
    if (i0 < Tp) s_out[i0] = pre2;
    if (i1 < Tp) s_out[i1] = pre2 + c0;
    if (threadIdx.x == 0) s_out[Tp] = tot;
  }
  __syncthreads();

  const int64_t n_out_len = (int64_t)Tp * B;
  for (int64_t i = tid; i < n_out_len; i += nthreads) {
    const int64_t tp = i / B;
    out_lengths[i] = lengths[(int64_t)permute[tp] * B + (i - tp * B)];
  }
  const int64_t total = min(s_out[Tp], out_capacity);
  for (int64_t j = tid; j < total; j += nthreads) {
    int lo = 0, hi = Tp;  // largest tp with s_out[tp] <= j  (empty features are skipped)
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (s_out[mid] <= j) lo = mid; else hi = mid;
    }
    const int64_t src = s_in[permute[lo]] + (j - s_out[lo]);
    reinterpret_cast<V*>(out_values)[j] = reinterpret_cast<const V*>(values)[src];
    if (weights) reinterpret_cast<Wt*>(out_weights)[j] = reinterpret_cast<const Wt*>(weights)[src];
  }
}

// ---------------------------------------------------------------- expand_into_jagged_permute
template <typename T>
__global__ void __launch_bounds__(256)
expand_into_jagged_permute_kernel(const int32_t* __restrict__ permute,
                                  const T* __restrict__ input_offsets,
                                  const T* __restrict__ output_offsets, int64_t permute_size, int G,
                                  T* __restrict__ out, int64_t output_size) {
  const int64_t gid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
  const int sub = threadIdx.x & (G - 1);
  const int64_t stride = ((int64_t)gridDim.x * blockDim.x) / G;
  for (int64_t i = gid; i < permute_size; i += stride) {
    const T dst = output_offsets[i];
    const T len = output_offsets[i + 1] - dst;
    const T src = input_offsets[permute[i]];
    for (T k = sub; k < len; k += G)
      if (dst + k < output_size) out[dst + k] = src + k;
  }
}

// ---------------------------------------------------------------- offsets_range
template <typename T>
__global__ void __launch_bounds__(256)
offsets_range_kernel(const T* __restrict__ offsets, int64_t n, int G, T* __restrict__ out,
                     int64_t range_size) {
  const int64_t gid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
  const int sub = threadIdx.x & (G - 1);
  const int64_t stride = ((int64_t)gridDim.x * blockDim.x) / G;
  for (int64_t i = gid; i < n; i += stride) {
    const int64_t start = offsets[i];
    const int64_t end = (i + 1 < n) ? (int64_t)offsets[i + 1] : range_size;
    for (int64_t k = sub; k < end - start; k += G) out[start + k] = (T)k;
  }
}

// ---------------------------------------------------------------- block bucketize
template <typename index_t>
__device__ __forceinline__ void bucket_of(index_t raw, index_t blk, int32_t W, uint64_t& p,
                                          uint64_t& new_idx) {
  using u = typename std::make_unsigned<index_t>::type;
  const u idx = (u)raw;
  const u ublk = (u)blk;
  const u bound = (u)(blk * (index_t)W);  // same wrap-around as the reference's index_t product
  if constexpr (sizeof(index_t) == 8) {
    // ids and block sizes almost always fit 32 bits: a 32-bit divide is ~5x cheaper than the
    // emulated 64-bit one
    if (((idx | ublk) >> 32) == 0) {
      const uint32_t i32 = (uint32_t)idx, b32 = (uint32_t)ublk;
      if (idx < bound) {
        const uint32_t q = i32 / b32;
        p = q;
        new_idx = i32 - q * b32;
      } else {
        const uint32_t q = i32 / (uint32_t)W;
        p = i32 - q * (uint32_t)W;
        new_idx = q;
      }
      return;
    }
  }
  if (idx < bound) {
    p = (uint64_t)(idx / ublk);
    new_idx = (uint64_t)(idx % ublk);
  } else {
    p = (uint64_t)(idx % (u)W);
    new_idx = (uint64_t)(idx / (u)W);
  }
}

// i / d for non-negative int64 operands, through a 32-bit divide when both fit
__device__ __forceinline__ int64_t fast_div(int64_t i, int64_t d) {
  if (((uint64_t)(i | d) >> 32) == 0) return (int64_t)((uint32_t)i / (uint32_t)d);
  return i / d;
}

template <typename offset_t, typename index_t>
__global__ void __launch_bounds__(kScanThreads)
block_bucketize_kernel(const offset_t* __restrict__ lengths, const index_t* __restrict__ indices,
                       const index_t* __restrict__ block_sizes, int32_t T, int32_t B, int32_t W,
                       const float* __restrict__ weights, offset_t* __restrict__ new_lengths,
                       index_t* __restrict__ new_indices, float* __restrict__ new_weights,
                       index_t* __restrict__ new_pos, index_t* __restrict__ unbucketize_permute,
                       int64_t* __restrict__ offsets, int64_t* __restrict__ new_offsets,
                       bool warp_per_row, int64_t* __restrict__ ws) {
  __shared__ int64_t s_warp[kScanThreads / 32 + 1];
  cg::grid_group grid = cg::this_grid();
  const int64_t TB = (int64_t)T * B;
  const int64_t WTB = TB * W;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
  const int lane = threadIdx.x & 31;

  // phase 1: offsets = exclusive scan(lengths); clear the bucket counters
  auto load_len = [&](int64_t i) -> int64_t { return (int64_t)lengths[i]; };
  scan_reduce_phase(load_len, TB, ws, s_warp);
  for (int64_t i = tid; i < WTB; i += nthreads) new_lengths[i] = 0;
  grid.sync();
  const int64_t total = scan_finish_phase(
      load_len, [&](int64_t i, int64_t pre, int64_t) { offsets[i] = pre; }, TB, ws, s_warp);
  if (blockIdx.x == 0 && threadIdx.x == 0) offsets[TB] = total;
  grid.sync();

  // phase 2: count ids per (bucket, row)
  if (!warp_per_row) {
    for (int64_t bt = tid; bt < TB; bt += nthreads) {
      const index_t blk = block_sizes[bt / B];
      const int64_t s = offsets[bt], e = offsets[bt + 1];
      for (int64_t i = s; i < e; ++i) {
        uint64_t p, ni;
        bucket_of<index_t>(indices[i], blk, W, p, ni);
        new_lengths[p * TB + bt] += 1;
      }
    }
  } else {
    for (int64_t bt = tid >> 5; bt < TB; bt += nthreads >> 5) {
      const index_t blk = block_sizes[bt / B];
      const int64_t s = offsets[bt], e = offsets[bt + 1];
      for (int64_t i0 = s; i0 < e; i0 += 32) {
        const int64_t i = i0 + lane;
        uint64_t p = 0, ni;
        const bool live = i < e;
        if (live) bucket_of<index_t>(indices[i], blk, W, p, ni);
        const unsigned act = __ballot_sync(kFull, live);
        if (live) {
          const unsigned peers = __match_any_sync(act, p);
          if (lane == (31 - __clz(peers))) new_lengths[p * TB + bt] += (offset_t)__popc(peers);
        }
        __syncwarp();
      }
    }
  }
  grid.sync();

  // phase 3: new_offsets = exclusive scan(new_lengths)
  auto load_nl = [&](int64_t i) -> int64_t { return (int64_t)new_lengths[i]; };
  scan_reduce_phase(load_nl, WTB, ws + kMaxCoopBlocks, s_warp);
  grid.sync();
  scan_finish_phase(
      load_nl, [&](int64_t i, int64_t pre, int64_t) { new_offsets[i] = pre; }, WTB,
      ws + kMaxCoopBlocks, s_warp);
  grid.sync();

  // phase 4: scatter, keeping the in-bag order inside each bucket
  if (!warp_per_row) {
    for (int64_t bt = tid; bt < TB; bt += nthreads) {
      const index_t blk = block_sizes[bt / B];
      const int64_t s = offsets[bt], e = offsets[bt + 1];
      for (int64_t i = s; i < e; ++i) {
        uint64_t p, ni;
        bucket_of<index_t>(indices[i], blk, W, p, ni);
        const int64_t pos = new_offsets[p * TB + bt];
        new_offsets[p * TB + bt] = pos + 1;
        new_indices[pos] = (index_t)ni;
        if (unbucketize_permute) unbucketize_permute[i] = (index_t)pos;
        if (new_weights) new_weights[pos] = weights[i];
        if (new_pos) new_pos[pos] = (index_t)(i - s);
      }
    }
  } else {
    for (int64_t bt = tid >> 5; bt < TB; bt += nthreads >> 5) {
      const index_t blk = block_sizes[bt / B];
      const int64_t s = offsets[bt], e = offsets[bt + 1];
      for (int64_t i0 = s; i0 < e; i0 += 32) {
        const int64_t i = i0 + lane;
        uint64_t p = 0, ni = 0;
        const bool live = i < e;
        if (live) bucket_of<index_t>(indices[i], blk, W, p, ni);
        const unsigned act = __ballot_sync(kFull, live);
        if (live) {
          const unsigned peers = __match_any_sync(act, p);
          const int rank = __popc(peers & ((1u << lane) - 1u));
          volatile int64_t* cur = new_offsets + p * TB + bt;
          const int64_t base = *cur;
          const int64_t pos = base + rank;
          new_indices[pos] = (index_t)ni;
          if (unbucketize_permute) unbucketize_permute[i] = (index_t)pos;
          if (new_weights) new_weights[pos] = weights[i];
          if (new_pos) new_pos[pos] = (index_t)(i - s);
          __syncwarp(act);
          if (lane == (31 - __clz(peers))) *cur = base + __popc(peers);
        }
        __syncwarp();
      }
    }
  }
}

// Short-bag variant (average bag <= 8 ids, e.g. Criteo's 1 id per bag, and N < 2^31): thread per
// bag; the scan of the W*T*B bucket counters is kept as int32 positions (the generic kernel writes
// and re-reads them as int64) and the per-element work of that scan is a plain coalesced store:
//   R1 | sync | F1: offsets[bt] = scan(lengths) and, in the same pass, the W bucket counts of bag
//   bt -> new_lengths (coalesced over bt) | sync | R2 | sync | F2: pos32 = exclusive scan of
//   new_lengths | sync | P3: thread per bag scatters its ids to pos32[bucket][bag] (+ rank in bag).
template <typename offset_t, typename index_t>
__global__ void __launch_bounds__(kScanThreads)
block_bucketize_short_kernel(const offset_t* __restrict__ lengths, const index_t* __restrict__ indices,
                             const index_t* __restrict__ block_sizes, int32_t T, int32_t B, int32_t W,
                             const float* __restrict__ weights, offset_t* __restrict__ new_lengths,
                             index_t* __restrict__ new_indices, float* __restrict__ new_weights,
                             index_t* __restrict__ new_pos, index_t* __restrict__ unbucketize_permute,
                             int64_t* __restrict__ offsets, int32_t* __restrict__ pos32,
                             int64_t* __restrict__ ws) {
  __shared__ int64_t s_warp[kScanThreads / 32 + 1];
  cg::grid_group grid = cg::this_grid();
  const int64_t TB = (int64_t)T * B;
  const int64_t WTB = TB * W;

  auto load_len = [&](int64_t i) -> int64_t { return (int64_t)lengths[i]; };
  scan_reduce_phase(load_len, TB, ws, s_warp);
  grid.sync();
  const int64_t total = scan_finish_phase(
      load_len,
      [&](int64_t bt, int64_t pre, int64_t len) {
        offsets[bt] = pre;
        const index_t blk = block_sizes[fast_div(bt, B)];
        if (len == 1) {  // the common case: one coalesced pass, no read-modify-write
          uint64_t p, ni;
          bucket_of<index_t>(indices[pre], blk, W, p, ni);
          for (int32_t q = 0; q < W; ++q) new_lengths[(int64_t)q * TB + bt] = (offset_t)(q == (int32_t)p);
        } else {
          for (int32_t q = 0; q < W; ++q) new_lengths[(int64_t)q * TB + bt] = 0;
          for (int64_t i = pre; i < pre + len; ++i) {
            uint64_t p, ni;
            bucket_of<index_t>(indices[i], blk, W, p, ni);
            new_lengths[(int64_t)p * TB + bt] += 1;  // own cells only
          }
        }
      },
      TB, ws, s_warp);
  if (blockIdx.x == 0 && threadIdx.x == 0) offsets[TB] = total;
  grid.sync();

  auto load_nl = [&](int64_t i) -> int64_t { return (int64_t)new_lengths[i]; };
  scan_reduce_phase(load_nl, WTB, ws + kMaxCoopBlocks, s_warp);
  grid.sync();
  scan_finish_phase(
      load_nl, [&](int64_t i, int64_t pre, int64_t) { pos32[i] = (int32_t)pre; }, WTB,
      ws + kMaxCoopBlocks, s_warp);
  grid.sync();

  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
  for (int64_t bt = tid; bt < TB; bt += nthreads) {
    const int64_t s = offsets[bt], e = offsets[bt + 1];
    if (e <= s) continue;
    const index_t blk = block_sizes[fast_div(bt, B)];
    if (e - s == 1) {
      uint64_t p, ni;
      bucket_of<index_t>(indices[s], blk, W, p, ni);
      const int64_t pos = pos32[(int64_t)p * TB + bt];
      new_indices[pos] = (index_t)ni;
      if (unbucketize_permute) unbucketize_permute[s] = (index_t)pos;
      if (new_weights) new_weights[pos] = weights[s];
      if (new_pos) new_pos[pos] = 0;
    } else {
      for (int64_t k = s; k < e; ++k) {
        uint64_t p, ni;
        bucket_of<index_t>(indices[k], blk, W, p, ni);
        int32_t* cur = pos32 + (int64_t)p * TB + bt;  // own cells only
        const int64_t pos = *cur;
        *cur = (int32_t)(pos + 1);
        new_indices[pos] = (index_t)ni;
        if (unbucketize_permute) unbucketize_permute[k] = (index_t)pos;
        if (new_weights) new_weights[pos] = weights[k];
        if (new_pos) new_pos[pos] = (index_t)(k - s);
      }
    }
  }
}

// ---------------------------------------------------------------- bounds check
// Pass A: validate offsets; if any bag range is malformed, the last CTA to finish repairs
// the whole offsets array with the reference's sequential rule
//   off'[0] = clamp(off[0], 0, N);  off'[i+1] = max(off'[i], min(off[i+1], N))
// (embedding_bounds_check.cu:57-71 applied bag after bag).
template <typename T>
__global__ void __launch_bounds__(256)
bounds_check_offsets_kernel(T* __restrict__ offsets, int64_t n_bags, int64_t num_indices, int mode,
                            int64_t* __restrict__ warning, int64_t* __restrict__ flags) {
  __shared__ int s_bad;
  __shared__ bool s_last;
  __shared__ int64_t s_warp[8];
  if (threadIdx.x == 0) s_bad = 0;
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int bad = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_bags; i += stride) {
    const int64_t s = offsets[i], e = offsets[i + 1];
    if (s < 0 || s > e || e > num_indices) {
      bad = 1;
      if (mode == B200_BOUNDS_FATAL) {
        printf("EmbeddingBoundsCheck: bad offsets for bag %lld: start %lld end %lld num_indices %lld\n",
               (long long)i, (long long)s, (long long)e, (long long)num_indices);
        __trap();
      }
      if (mode == B200_BOUNDS_WARNING) {
        if (atomicAdd((unsigned long long*)warning, 1ull) == 0)
          printf(
              "EmbeddingBoundsCheck: (at least one) Out of bounds access for bag: %lld, "
              "indices_start: %lld, indices_end: %lld, num_indices: %lld. Setting indices_start "
              "and indices_end within the range.\n",
              (long long)i, (long long)s, (long long)e, (long long)num_indices);
      }
    }
  }
  if (bad) atomicOr(&s_bad, 1);
  __syncthreads();
  if (threadIdx.x == 0) {
    if (s_bad) atomicExch((unsigned long long*)&flags[0], 1ull);
    __threadfence();
    const unsigned long long ticket = atomicAdd((unsigned long long*)&flags[1], 1ull);
    s_last = (ticket == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const bool any_bad = *((volatile int64_t*)&flags[0]) != 0;
  if (threadIdx.x == 0) {  // reset for the next call
    flags[0] = 0;
    flags[1] = 0;
  }
  if (!any_bad) return;
  // rare path: running max-scan over all offsets by this one CTA
  int64_t carry = 0;
  for (int64_t base = 0; base <= n_bags; base += 256) {
    const int64_t i = base + threadIdx.x;
    int64_t v = INT64_MIN;
    if (i <= n_bags) {
      v = (int64_t)offsets[i];
      if (v > num_indices) v = num_indices;
      if (i == 0 && v < 0) v = 0;
    }
    // inclusive max-scan within the block
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int64_t m = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int64_t nb = __shfl_up_sync(kFull, m, o);
      if (lane >= o && nb > m) m = nb;
    }
    __syncthreads();
    if (lane == 31) s_warp[warp] = m;
    __syncthreads();
    int64_t pre = (base == 0) ? INT64_MIN : carry;
    for (int w = 0; w < warp; ++w) pre = s_warp[w] > pre ? s_warp[w] : pre;
    if (pre > m) m = pre;
    if (i <= n_bags) offsets[i] = (T)m;
    int64_t blockmax = (base == 0) ? INT64_MIN : carry;
    for (int w = 0; w < 8; ++w) blockmax = s_warp[w] > blockmax ? s_warp[w] : blockmax;
    carry = blockmax;
    __syncthreads();
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
bounds_check_indices_kernel(const int64_t* __restrict__ rows_per_table, T* __restrict__ indices,
                            const T* __restrict__ offsets, int32_t Tn, int32_t B, int G, int mode,
                            int64_t* __restrict__ warning) {
  const int64_t n_bags = (int64_t)Tn * B;
  const int64_t gid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
  const int sub = threadIdx.x & (G - 1);
  const int64_t stride = ((int64_t)gridDim.x * blockDim.x) / G;
  for (int64_t bt = gid; bt < n_bags; bt += stride) {
    const int64_t t = bt / B;
    const int64_t rows = rows_per_table[t];
    const int64_t s = offsets[bt], e = offsets[bt + 1];
    for (int64_t i = s + sub; i < e; i += G) {
      const int64_t idx = (int64_t)indices[i];
      if (idx == -1) continue;  // pruned-row marker (embedding_bounds_check.cu:77-80)
      if (idx < 0 || idx >= rows) {
        if (mode == B200_BOUNDS_FATAL) {
          printf("EmbeddingBoundsCheck: idx %lld out of range [0, %lld) for table %lld bag %lld\n",
                 (long long)idx, (long long)rows, (long long)t, (long long)(bt - t * B));
          __trap();
        }
        if (mode == B200_BOUNDS_WARNING) {
          if (atomicAdd((unsigned long long*)warning, 1ull) == 0)
            printf(
                "EmbeddingBoundsCheck: (at least one) Out of bounds access for batch: %lld, "
                "table: %lld, bag element: %lld, idx: %lld, num_rows: %lld. Setting idx to zero.\n",
                (long long)(bt - t * B), (long long)t, (long long)(i - s), (long long)idx,
                (long long)rows);
        }
        indices[i] = 0;
      }
    }
  }
}

template <typename K, typename... Args>
static int coop_launch(K kernel, int64_t n, cudaStream_t s, Args... args) {
  const int grid = coop_grid_for((const void*)kernel, n);
  void* argv[] = {(void*)&args...};
  B200_CUDA_TRY(cudaLaunchCooperativeKernel((const void*)kernel, dim3(grid), dim3(kScanThreads),
                                            argv, 0, s));
  return B200_OK;
}

}  // namespace b200

using namespace b200;

extern "C" size_t b200_scan_workspace_bytes(void) { return kWsInt64s * sizeof(int64_t); }

extern "C" int b200_cumsum(const void* in, void* out, int64_t n, int dtype, int mode, void* ws,
                           void* stream) {
  if (n < 0 || mode < 0 || mode > 2) return B200_EINVAL;
  cudaStream_t s = (cudaStream_t)stream;
  int64_t* w = (int64_t*)ws;
  if (dtype == B200_I32)
    return coop_launch(cumsum_kernel<int32_t>, n, s, (const int32_t*)in, (int32_t*)out, n, mode, w);
  if (dtype == B200_I64)
    return coop_launch(cumsum_kernel<int64_t>, n, s, (const int64_t*)in, (int64_t*)out, n, mode, w);
  return B200_EINVAL;
}

extern "C" int b200_permute_lengths(const int32_t* permute, const void* lengths, int lengths_dtype,
                                    int32_t T, int32_t Tp, int32_t B, void* out_lengths,
                                    int64_t* in_offsets, int64_t* out_offsets, void* ws,
                                    void* stream) {
  if (T < 0 || Tp < 0 || B < 0) return B200_EINVAL;
  cudaStream_t s = (cudaStream_t)stream;
  int64_t* w = (int64_t*)ws;
  const int64_t n = (int64_t)(T > Tp ? T : Tp) * B;
  if (lengths_dtype == B200_I32)
    return coop_launch(permute_lengths_kernel<int32_t>, n, s, permute, (const int32_t*)lengths, T,
                       Tp, B, (int32_t*)out_lengths, in_offsets, out_offsets, w);
  if (lengths_dtype == B200_I64)
    return coop_launch(permute_lengths_kernel<int64_t>, n, s, permute, (const int64_t*)lengths, T,
                       Tp, B, (int64_t*)out_lengths, in_offsets, out_offsets, w);
  return B200_EINVAL;
}

template <int VB>
static int permute_data_w(int weight_bytes, unsigned grid, cudaStream_t s, const int32_t* permute,
                          int32_t Tp, int32_t B, int G, const int64_t* in_offsets,
                          const int64_t* out_offsets, const void* values, void* out_values,
                          const void* weights, void* out_weights, int64_t cap) {
#define B200_PD(WB)                                                                            \
  permute_data_kernel<VB, WB><<<grid, 256, 0, s>>>(permute, Tp, B, G, in_offsets, out_offsets, \
                                                   values, out_values, weights, out_weights, cap)
  switch (weights ? weight_bytes : 4) {
    case 2:
      B200_PD(2);
      break;
    case 4:
      B200_PD(4);
      break;
    case 8:
      B200_PD(8);
      break;
    default:
      return B200_EINVAL;
  }
#undef B200_PD
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_permute_data(const int32_t* permute, int32_t Tp, int32_t B,
                                 const int64_t* in_offsets, const int64_t* out_offsets,
                                 const void* values, void* out_values, int value_bytes,
                                 const void* weights, void* out_weights, int weight_bytes,
                                 int64_t out_capacity, void* stream) {
  const int64_t segs = (int64_t)Tp * B;
  if (segs == 0 || out_capacity == 0) return B200_OK;
  cudaStream_t s = (cudaStream_t)stream;
  const int G = pick_group(out_capacity, segs);
  const unsigned grid = grid_for_groups(segs, G, 256);
  switch (value_bytes) {
    case 1:
      return permute_data_w<1>(weight_bytes, grid, s, permute, Tp, B, G, in_offsets, out_offsets,
                               values, out_values, weights, out_weights, out_capacity);
    case 2:
      return permute_data_w<2>(weight_bytes, grid, s, permute, Tp, B, G, in_offsets, out_offsets,
                               values, out_values, weights, out_weights, out_capacity);
    case 4:
      return permute_data_w<4>(weight_bytes, grid, s, permute, Tp, B, G, in_offsets, out_offsets,
                               values, out_values, weights, out_weights, out_capacity);
    case 8:
      return permute_data_w<8>(weight_bytes, grid, s, permute, Tp, B, G, in_offsets, out_offsets,
                               values, out_values, weights, out_weights, out_capacity);
    default:
      return B200_EINVAL;
  }
}

template <typename LT, int VB>
static int permute_2d_fused_w(int weight_bytes, cudaStream_t s, int64_t work, const int32_t* permute,
                              const LT* lengths, int32_t T, int32_t Tp, int32_t B, LT* out_lengths,
                              const void* values, void* out_values, const void* weights,
                              void* out_weights, int64_t cap, int64_t* ws) {
  switch (weights ? weight_bytes : 4) {
    case 2:
      return coop_launch(permute_2d_fused_kernel<LT, VB, 2>, work, s, permute, lengths, T, Tp, B,
                         out_lengths, values, out_values, weights, out_weights, cap, ws);
    case 4:
      return coop_launch(permute_2d_fused_kernel<LT, VB, 4>, work, s, permute, lengths, T, Tp, B,
                         out_lengths, values, out_values, weights, out_weights, cap, ws);
    case 8:
      return coop_launch(permute_2d_fused_kernel<LT, VB, 8>, work, s, permute, lengths, T, Tp, B,
                         out_lengths, values, out_values, weights, out_weights, cap, ws);
    default:
      return B200_EINVAL;
  }
}

template <typename LT>
static int permute_2d_fused_v(int value_bytes, int weight_bytes, cudaStream_t s, int64_t work,
                              const int32_t* permute, const LT* lengths, int32_t T, int32_t Tp,
                              int32_t B, LT* out_lengths, const void* values, void* out_values,
                              const void* weights, void* out_weights, int64_t cap, int64_t* ws) {
#define B200_PF(VB)                                                                                 \
  return permute_2d_fused_w<LT, VB>(weight_bytes, s, work, permute, lengths, T, Tp, B, out_lengths, \
                                    values, out_values, weights, out_weights, cap, ws)
  switch (value_bytes) {
    case 1:
      B200_PF(1);
    case 2:
      B200_PF(2);
    case 4:
      B200_PF(4);
    case 8:
      B200_PF(8);
    default:
      return B200_EINVAL;
  }
#undef B200_PF
}

extern "C" int b200_permute_2d_fused(const int32_t* permute, const void* lengths, int lengths_dtype,
                                     int32_t T, int32_t Tp, int32_t B, void* out_lengths,
                                     const void* values, void* out_values, int value_bytes,
                                     const void* weights, void* out_weights, int weight_bytes,
                                     int64_t out_capacity, void* ws, void* stream) {
  if (T < 0 || Tp < 0 || B < 0 || out_capacity < 0) return B200_EINVAL;
  if (T > kPermMaxFeatures || Tp > kPermMaxFeatures) return B200_EUNSUPPORTED;
  if (Tp == 0 || B == 0) return B200_OK;
  cudaStream_t s = (cudaStream_t)stream;
  const int64_t nlen = (int64_t)(T > Tp ? T : Tp) * B;
  const int64_t work = nlen > out_capacity ? nlen : out_capacity;
  if (lengths_dtype == B200_I32)
    return permute_2d_fused_v<int32_t>(value_bytes, weight_bytes, s, work, permute,
                                       (const int32_t*)lengths, T, Tp, B, (int32_t*)out_lengths, values,
                                       out_values, weights, out_weights, out_capacity, (int64_t*)ws);
  if (lengths_dtype == B200_I64)
    return permute_2d_fused_v<int64_t>(value_bytes, weight_bytes, s, work, permute,
                                       (const int64_t*)lengths, T, Tp, B, (int64_t*)out_lengths, values,
                                       out_values, weights, out_weights, out_capacity, (int64_t*)ws);
  return B200_EINVAL;
}

extern "C" int b200_expand_into_jagged_permute(const int32_t* permute, const void* input_offsets,
                                               const void* output_offsets, int dtype,
                                               int64_t permute_size, void* output_permute,
                                               int64_t output_size, void* stream) {
  if (permute_size == 0 || output_size == 0) return B200_OK;
  cudaStream_t s = (cudaStream_t)stream;
  const int G = pick_group(output_size, permute_size);
  const unsigned grid = grid_for_groups(permute_size, G, 256);
  if (dtype == B200_I32)
    expand_into_jagged_permute_kernel<int32_t><<<grid, 256, 0, s>>>(
        permute, (const int32_t*)input_offsets, (const int32_t*)output_offsets, permute_size, G,
        (int32_t*)output_permute, output_size);
  else if (dtype == B200_I64)
    expand_into_jagged_permute_kernel<int64_t><<<grid, 256, 0, s>>>(
        permute, (const int64_t*)input_offsets, (const int64_t*)output_offsets, permute_size, G,
        (int64_t*)output_permute, output_size);
  else
    return B200_EINVAL;
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_offsets_range(const void* offsets, int dtype, int64_t n_offsets, void* out,
                                  int64_t range_size, void* stream) {
  if (n_offsets == 0 || range_size == 0) return B200_OK;
  cudaStream_t s = (cudaStream_t)stream;
  const int G = pick_group(range_size, n_offsets);
  const unsigned grid = grid_for_groups(n_offsets, G, 256);
  if (dtype == B200_I32)
    offsets_range_kernel<int32_t><<<grid, 256, 0, s>>>((const int32_t*)offsets, n_offsets, G,
                                                      (int32_t*)out, range_size);
  else if (dtype == B200_I64)
    offsets_range_kernel<int64_t><<<grid, 256, 0, s>>>((const int64_t*)offsets, n_offsets, G,
                                                      (int64_t*)out, range_size);
  else
    return B200_EINVAL;
  B200_LAUNCH_CHECK();
  return B200_OK;
}

template <typename offset_t, typename index_t>
static int bucketize_launch(const void* lengths, const void* indices, const void* block_sizes,
                            int32_t T, int32_t B, int32_t W, int64_t N, const float* weights,
                            void* new_lengths, void* new_indices, float* new_weights, void* new_pos,
                            void* unbucketize_permute, int64_t* offsets, int64_t* new_offsets,
                            int64_t* ws, cudaStream_t s) {
  const int64_t TB = (int64_t)T * B;
  if (TB > 0 && N <= 8 * TB && N < 0x7fffffffLL)  // short bags: thread per bag, int32 positions
    return coop_launch(block_bucketize_short_kernel<offset_t, index_t>, TB * W, s,
                       (const offset_t*)lengths, (const index_t*)indices, (const index_t*)block_sizes,
                       T, B, W, weights, (offset_t*)new_lengths, (index_t*)new_indices, new_weights,
                       (index_t*)new_pos, (index_t*)unbucketize_permute, offsets,
                       (int32_t*)new_offsets, ws);
  const bool warp_per_row = TB > 0 && (N / TB) >= 16;
  const int64_t work = warp_per_row ? TB * 32 : (TB * W > N ? TB * W : N);
  return coop_launch(block_bucketize_kernel<offset_t, index_t>, work, s, (const offset_t*)lengths,
                     (const index_t*)indices, (const index_t*)block_sizes, T, B, W, weights,
                     (offset_t*)new_lengths, (index_t*)new_indices, new_weights, (index_t*)new_pos,
                     (index_t*)unbucketize_permute, offsets, new_offsets, warp_per_row, ws);
}

extern "C" int b200_block_bucketize(const void* lengths, int offset_dtype, const void* indices,
                                    int index_dtype, const void* block_sizes, int32_t T, int32_t B,
                                    int32_t W, int64_t N, const float* weights, void* new_lengths,
                                    void* new_indices, float* new_weights, void* new_pos,
                                    void* unbucketize_permute, int64_t* offsets_scratch,
                                    int64_t* new_offsets_scratch, void* ws, void* stream) {
  if (T <= 0 || B < 0 || W <= 0) return B200_EINVAL;
  if ((weights == nullptr) != (new_weights == nullptr)) return B200_EINVAL;
  if (B == 0) return B200_OK;
  cudaStream_t s = (cudaStream_t)stream;
  int64_t* w = (int64_t*)ws;
#define B200_BB(OT, IT)                                                                       \
  return bucketize_launch<OT, IT>(lengths, indices, block_sizes, T, B, W, N, weights,         \
                                  new_lengths, new_indices, new_weights, new_pos,             \
                                  unbucketize_permute, offsets_scratch, new_offsets_scratch, w, s)
  if (offset_dtype == B200_I32 && index_dtype == B200_I32) B200_BB(int32_t, int32_t);
  if (offset_dtype == B200_I32 && index_dtype == B200_I64) B200_BB(int32_t, int64_t);
  if (offset_dtype == B200_I64 && index_dtype == B200_I32) B200_BB(int64_t, int32_t);
  if (offset_dtype == B200_I64 && index_dtype == B200_I64) B200_BB(int64_t, int64_t);
#undef B200_BB
  return B200_EINVAL;
}

extern "C" int b200_bounds_check(const int64_t* rows_per_table, void* indices, void* offsets,
                                 int dtype, int32_t T, int32_t B, int64_t num_indices, int mode,
                                 int64_t* warning, void* ws, void* stream) {
  if (T == 0 || B == 0) return B200_OK;
  if (mode < 0 || mode > 2) return B200_EINVAL;
  cudaStream_t s = (cudaStream_t)stream;
  int64_t* flags = (int64_t*)ws + kWsFlagOffset;
  const int64_t n_bags = (int64_t)T * B;
  if (mode == B200_BOUNDS_WARNING) B200_CUDA_TRY(cudaMemsetAsync(warning, 0, sizeof(int64_t), s));
  unsigned gridA = (unsigned)ceil_div(n_bags, 256 * 4);
  if (gridA > (unsigned)kNumSMs * 8) gridA = kNumSMs * 8;
  const int G = pick_group(num_indices, n_bags);
  const unsigned gridB = grid_for_groups(n_bags, G, 256);
  if (dtype == B200_I32) {
    bounds_check_offsets_kernel<int32_t><<<gridA, 256, 0, s>>>((int32_t*)offsets, n_bags,
                                                              num_indices, mode, warning, flags);
    B200_LAUNCH_CHECK();
    bounds_check_indices_kernel<int32_t><<<gridB, 256, 0, s>>>(
        rows_per_table, (int32_t*)indices, (const int32_t*)offsets, T, B, G, mode, warning);
  } else if (dtype == B200_I64) {
    bounds_check_offsets_kernel<int64_t><<<gridA, 256, 0, s>>>((int64_t*)offsets, n_bags,
                                                              num_indices, mode, warning, flags);
    B200_LAUNCH_CHECK();
    bounds_check_indices_kernel<int64_t><<<gridB, 256, 0, s>>>(
        rows_per_table, (int64_t*)indices, (const int64_t*)offsets, T, B, G, mode, warning);
  } else {
    return B200_EINVAL;
  }
  B200_LAUNCH_CHECK();
  return B200_OK;
}
