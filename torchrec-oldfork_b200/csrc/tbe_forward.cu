// Pooled embedding-bag lookup (table-batched) for sm_100a.
//
// Replaces split_embedding_codegen_forward_{un,}weighted_kernel
// (fbgemm_gpu/codegen/embedding_forward_split_template.cu:43-274).  Same contract:
// out[b, D_offsets[t] : D_offsets[t+1]] = sum_{p in bag(t,b)} (alpha_p *) W_t[idx_p, :],
// times 1/L for MEAN, fp32 accumulation, bag (t,b) at offsets[t*B + b].
//
// Design (differs from the reference's one-warp-per-bag):
//  * one warp owns a TILE of 32 consecutive bags of one table.  Their ids form one
//    contiguous range of `indices`, which the warp walks as a flat stream; bag boundaries
//    are detected on the fly from the 33 offsets held in registers.  The number of row
//    gathers in flight is therefore independent of the pooling factor: at L = 1 the warp
//    still keeps 8 x 16-byte loads per lane outstanding (the reference has one).
//  * ids of the next batch are fetched while the rows of the current batch are in flight
//    (two dependent DRAM latencies overlapped by software pipelining).
//  * rows are read with ld.global.nc.L1::no_allocate 128-bit loads and the output row is
//    written once with streaming 128-bit stores.
#include "common.cuh"

namespace b200 {

constexpr int kFwdThreads = 128;

template <typename W, typename O, int EPV, int NV>
__global__ void __launch_bounds__(kFwdThreads)
tbe_forward_kernel(const W* __restrict__ weights, const int64_t* __restrict__ weights_offsets,
                   const int32_t* __restrict__ D_offsets, int32_t T, int32_t B, int32_t total_D,
                   int32_t tiles_per_table, const void* __restrict__ indices,
                   const void* __restrict__ offsets, bool idx64,
                   const float* __restrict__ psw, bool mean, O* __restrict__ out) {
  constexpr int U = (32 / (NV * EPV)) > 0 ? (32 / (NV * EPV)) : 1;  // rows in flight per warp
  const int lane = threadIdx.x & 31;
  const int64_t tile = (int64_t)blockIdx.x * (kFwdThreads / 32) + (threadIdx.x >> 5);
  if (tile >= (int64_t)T * tiles_per_table) return;
  const int t = (int)(tile / tiles_per_table);
  const int b0 = (int)(tile - (int64_t)t * tiles_per_table) * 32;
  const int nb = min(32, B - b0);
  const int D_begin = D_offsets[t];
  const int D = D_offsets[t + 1] - D_begin;
  const W* __restrict__ wtab = weights + weights_offsets[t];

  // 33 offsets of the tile, relative to the first one.
  const int64_t obase = (int64_t)t * B + b0;
  const int64_t my_start = load_index(offsets, obase + min(lane, nb), idx64);
  const int64_t S0 = __shfl_sync(kFull, my_start, 0);
  const int32_t rel_start = (int32_t)(my_start - S0);
  int32_t rel_end = __shfl_down_sync(kFull, rel_start, 1);
  if (lane == 31) rel_end = (int32_t)(load_index(offsets, obase + nb, idx64) - S0);
  const int32_t len = __shfl_sync(kFull, rel_end, 31);
  const bool in_tile = lane < nb;
  unsigned remaining = __ballot_sync(kFull, in_tile && rel_end > rel_start);
  const unsigned empty = __ballot_sync(kFull, in_tile && rel_end <= rel_start);

  O* __restrict__ out_tile = out + (int64_t)b0 * total_D + D_begin;

  // Empty bags pool to zeros.
  for (unsigned m = empty; m; m &= m - 1) {
    const int j = __ffs(m) - 1;
    float z[EPV];
#pragma unroll
    for (int e = 0; e < EPV; ++e) z[e] = 0.f;
#pragma unroll
    for (int v = 0; v < NV; ++v) {
      const int d = EPV * (lane + 32 * v);
      if (d < D) store_vec<O, EPV, true>(out_tile + (int64_t)j * total_D + d, z);
    }
  }
  if (len <= 0) return;

  float acc[NV][EPV];
#pragma unroll
  for (int v = 0; v < NV; ++v)
#pragma unroll
    for (int e = 0; e < EPV; ++e) acc[v][e] = 0.f;

  int cur_bag = remaining ? (__ffs(remaining) - 1) : 0;
  int32_t cur_end = __shfl_sync(kFull, rel_end, cur_bag);
  int32_t prev_end = 0;
  const bool weighted = psw != nullptr;

  int64_t idx_cur[U];
  float w_cur[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    idx_cur[u] = 0;
    w_cur[u] = 1.f;
    if (u < len) {
      idx_cur[u] = load_index(indices, S0 + u, idx64);
      if (weighted) w_cur[u] = psw[S0 + u];
    }
  }

  for (int32_t q = 0; q < len; q += U) {
    // ids of the next batch (independent of the row loads below)
    int64_t idx_nxt[U];
    float w_nxt[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int32_t p = q + U + u;
      idx_nxt[u] = 0;
      w_nxt[u] = 1.f;
      if (p < len) {
        idx_nxt[u] = load_index(indices, S0 + p, idx64);
        if (weighted) w_nxt[u] = psw[S0 + p];
      }
    }
    // gather U rows
    float rows[U][NV][EPV];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const W* __restrict__ row = wtab + idx_cur[u] * D;
      const bool live = q + u < len;
#pragma unroll
      for (int v = 0; v < NV; ++v) {
        const int d = EPV * (lane + 32 * v);
        if (live && d < D) {
          load_vec<W, EPV, true>(row + d, rows[u][v]);
        } else {
#pragma unroll
          for (int e = 0; e < EPV; ++e) rows[u][v][e] = 0.f;
        }
      }
    }
    // accumulate in stream order, flushing at bag boundaries
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (q + u < len) {
        if (weighted) {
#pragma unroll
          for (int v = 0; v < NV; ++v)
#pragma unroll
            for (int e = 0; e < EPV; ++e) acc[v][e] = fmaf(w_cur[u], rows[u][v][e], acc[v][e]);
        } else {
#pragma unroll
          for (int v = 0; v < NV; ++v)
#pragma unroll
            for (int e = 0; e < EPV; ++e) acc[v][e] += rows[u][v][e];
        }
        if (q + u + 1 == cur_end) {
          const float scale = mean ? 1.0f / (float)(cur_end - prev_end) : 1.0f;
#pragma unroll
          for (int v = 0; v < NV; ++v) {
            const int d = EPV * (lane + 32 * v);
            if (mean) {
#pragma unroll
              for (int e = 0; e < EPV; ++e) acc[v][e] *= scale;
            }
            if (d < D) store_vec<O, EPV, true>(out_tile + (int64_t)cur_bag * total_D + d, acc[v]);
#pragma unroll
            for (int e = 0; e < EPV; ++e) acc[v][e] = 0.f;
          }
          prev_end = cur_end;
          remaining &= remaining - 1;
          cur_bag = remaining ? (__ffs(remaining) - 1) : 0;
          cur_end = remaining ? __shfl_sync(kFull, rel_end, cur_bag) : 0x7fffffff;
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      idx_cur[u] = idx_nxt[u];
      w_cur[u] = w_nxt[u];
    }
  }
}

template <typename W, typename O, int EPV>
static int launch_forward_nv(const void* weights, const int64_t* weights_offsets,
                             const int32_t* D_offsets, int32_t T, int32_t B, int32_t total_D,
                             int32_t max_D, const void* indices, const void* offsets, bool idx64,
                             const float* psw, bool mean, void* output, cudaStream_t s) {
  const int tiles_per_table = (B + 31) / 32;
  const int64_t tiles = (int64_t)T * tiles_per_table;
  const int64_t grid = ceil_div(tiles, kFwdThreads / 32);
  if (grid == 0) return B200_OK;
  if (grid > 0x7fffffffLL) return B200_EINVAL;
  const int vecs = (int)ceil_div(max_D, EPV * 32);
#define B200_FWD_LAUNCH(NV)                                                                     \
  tbe_forward_kernel<W, O, EPV, NV><<<(unsigned)grid, kFwdThreads, 0, s>>>(                      \
      (const W*)weights, weights_offsets, D_offsets, T, B, total_D, tiles_per_table, indices,    \
      offsets, idx64, psw, mean, (O*)output)
  if (vecs <= 1)
    B200_FWD_LAUNCH(1);
  else if (vecs <= 2)
    B200_FWD_LAUNCH(2);
  else if (vecs <= 4)
    B200_FWD_LAUNCH(4);
  else if (vecs <= 8)
    B200_FWD_LAUNCH(8);
  else
    return B200_EUNSUPPORTED;  // reference limit is max_D <= 1024 too
#undef B200_FWD_LAUNCH
  B200_LAUNCH_CHECK();
  return B200_OK;
}

template <typename W, int EPV>
static int launch_forward_out(int output_dtype, const void* weights, const int64_t* wo,
                              const int32_t* Do, int32_t T, int32_t B, int32_t total_D,
                              int32_t max_D, const void* indices, const void* offsets, bool idx64,
                              const float* psw, bool mean, void* output, cudaStream_t s) {
  switch (output_dtype) {
    case B200_F32:
      return launch_forward_nv<W, float, EPV>(weights, wo, Do, T, B, total_D, max_D, indices,
                                              offsets, idx64, psw, mean, output, s);
    case B200_F16:
      return launch_forward_nv<W, __half, EPV>(weights, wo, Do, T, B, total_D, max_D, indices,
                                               offsets, idx64, psw, mean, output, s);
    default:
      return B200_EUNSUPPORTED;
  }
}

}  // namespace b200

extern "C" int b200_tbe_forward(const void* weights, int weight_dtype,
                                const int64_t* weights_offsets, const int32_t* D_offsets,
                                int32_t T, int32_t B, int32_t total_D, int32_t max_D,
                                int32_t all_D_mult8, const void* indices, const void* offsets,
                                int index_dtype, const float* per_sample_weights,
                                int pooling_mode, void* output, int output_dtype, void* stream) {
  using namespace b200;
  if (T < 0 || B < 0 || total_D < 0 || (total_D % 4) != 0) return B200_EINVAL;
  if (index_dtype != B200_I32 && index_dtype != B200_I64) return B200_EINVAL;
  if (T == 0 || B == 0) return B200_OK;
  const bool idx64 = index_dtype == B200_I64;
  const bool mean = pooling_mode == B200_POOL_MEAN;
  cudaStream_t s = (cudaStream_t)stream;
  switch (weight_dtype) {
    case B200_F32:
      return launch_forward_out<float, 4>(output_dtype, weights, weights_offsets, D_offsets, T, B,
                                          total_D, max_D, indices, offsets, idx64,
                                          per_sample_weights, mean, output, s);
    case B200_F16:
      if (all_D_mult8)
        return launch_forward_out<__half, 8>(output_dtype, weights, weights_offsets, D_offsets, T,
                                             B, total_D, max_D, indices, offsets, idx64,
                                             per_sample_weights, mean, output, s);
      return launch_forward_out<__half, 4>(output_dtype, weights, weights_offsets, D_offsets, T, B,
                                           total_D, max_D, indices, offsets, idx64,
                                           per_sample_weights, mean, output, s);
    case B200_BF16:
      if (all_D_mult8)
        return launch_forward_out<__nv_bfloat16, 8>(output_dtype, weights, weights_offsets,
                                                    D_offsets, T, B, total_D, max_D, indices,
                                                    offsets, idx64, per_sample_weights, mean,
                                                    output, s);
      return launch_forward_out<__nv_bfloat16, 4>(output_dtype, weights, weights_offsets,
                                                  D_offsets, T, B, total_D, max_D, indices,
                                                  offsets, idx64, per_sample_weights, mean, output,
                                                  s);
    default:
      return B200_EUNSUPPORTED;
  }
}
