// Small entry points: error text, ABI version, per-sample-weight gradient, and the
// receive-side layout change of the pooled all-to-all.
#include "common.cuh"

namespace b200 {

// grad wrt per-sample weights: grad_w[p] = sum_d grad_out[b, D_begin + d] * W_t[idx_p, d]
// (fbgemm_gpu/codegen/embedding_backward_split_indice_weights_template.cu).  One warp per
// 32-bag tile of one table walks the tile's ids as a stream (same layout as the forward).
template <typename W, int EPV>
__global__ void __launch_bounds__(128)
grad_indice_weights_kernel(const float* __restrict__ grad, int64_t grad_stride,
                           const W* __restrict__ weights,
                           const int64_t* __restrict__ weights_offsets,
                           const int32_t* __restrict__ D_offsets, int32_t T, int32_t B,
                           int32_t tiles_per_table, const void* __restrict__ indices,
                           const void* __restrict__ offsets, bool idx64, float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int64_t tile = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
  if (tile >= (int64_t)T * tiles_per_table) return;
  const int t = (int)(tile / tiles_per_table);
  const int b0 = (int)(tile - (int64_t)t * tiles_per_table) * 32;
  const int nb = min(32, B - b0);
  const int D_begin = D_offsets[t];
  const int D = D_offsets[t + 1] - D_begin;
  const W* __restrict__ wtab = weights + weights_offsets[t];
  const int64_t obase = (int64_t)t * B + b0;
  for (int j = 0; j < nb; ++j) {
    const int64_t s = load_index(offsets, obase + j, idx64);
    const int64_t e = load_index(offsets, obase + j + 1, idx64);
    const float* __restrict__ g = grad + (int64_t)(b0 + j) * grad_stride + D_begin;
    for (int64_t p = s; p < e; ++p) {
      const int64_t idx = load_index(indices, p, idx64);
      const W* __restrict__ row = wtab + idx * D;
      float acc = 0.f;
      for (int d = EPV * lane; d < D; d += EPV * 32) {
        float w[EPV];
        load_vec<W, EPV, true>(row + d, w);
#pragma unroll
        for (int k = 0; k < EPV; ++k) acc = fmaf(g[d + k], w[k], acc);
      }
      acc = warp_sum(acc);
      if (lane == 0) out[p] = acc;
    }
  }
}

// dim_cumsum[r] .. dim_cumsum[r+1] are the columns rank r contributes.
// forward: src block r is [B_local, dims[r]] (contiguous), dst is [B_local, sumD].
// backward: src is [B_local, sumD], dst block r is [B_local, dims[r]], scaled.
__global__ void __launch_bounds__(256)
pooled_a2a_recat_kernel(const float* __restrict__ src, float* __restrict__ dst,
                        const int32_t* __restrict__ dim_cumsum, int32_t W, int32_t B_local,
                        bool forward, float scale) {
  const int sumD4 = dim_cumsum[W] >> 2;
  const int64_t total4 = (int64_t)B_local * sumD4;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total4;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = i / sumD4;
    const int c4 = (int)(i - b * sumD4);
    const int col = c4 << 2;
    int r = 0;
    while (dim_cumsum[r + 1] <= col) ++r;
    const int dr = dim_cumsum[r + 1] - dim_cumsum[r];
    const int64_t blk = (int64_t)B_local * dim_cumsum[r] + b * dr + (col - dim_cumsum[r]);
    const int64_t flat = b * ((int64_t)sumD4 << 2) + col;
    if (forward) {
      const uint4 v = ldg_stream_16(src + blk);
      st_stream_16(dst + flat, v);
    } else {
      uint4 v = ldg_stream_16(src + flat);
      float4 f = *reinterpret_cast<float4*>(&v);
      f.x *= scale;
      f.y *= scale;
      f.z *= scale;
      f.w *= scale;
      *reinterpret_cast<float4*>(dst + blk) = f;
    }
  }
}

// Sequence (non-pooled) embeddings, PoolingMode.NONE: every id is its own output row.  All tables
// of such a module share one dim D and sit back to back in the flat weight buffer, so table t's row
// r is row hash_size_cumsum[t] + r of ONE virtual table: after this linearisation the sequence
// lookup / backward are the pooled kernels with T = 1, B = N, one id per bag (no second kernel zoo).
// Position p belongs to the feature t with offsets[t*B] <= p < offsets[(t+1)*B] (table-major ids).
__global__ void __launch_bounds__(256)
linearize_seq_kernel(const void* __restrict__ indices, const void* __restrict__ offsets, bool idx64,
                     const int64_t* __restrict__ hash_size_cumsum, int32_t T, int32_t B, int64_t N,
                     int64_t* __restrict__ out) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < N;
       p += (int64_t)gridDim.x * blockDim.x) {
    int lo = 0, hi = T;  // largest t with offsets[t*B] <= p
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (load_index(offsets, (int64_t)mid * B, idx64) <= p) lo = mid; else hi = mid;
    }
    out[p] = load_index(indices, p, idx64) + hash_size_cumsum[lo];
  }
}

}  // namespace b200

extern "C" int b200_linearize_seq_indices(const void* indices, const void* offsets, int index_dtype,
                                          const int64_t* hash_size_cumsum, int32_t T, int32_t B,
                                          int64_t N, int64_t* out, void* stream) {
  using namespace b200;
  if (index_dtype != B200_I32 && index_dtype != B200_I64) return B200_EINVAL;
  if (T <= 0 || B < 0 || N < 0) return B200_EINVAL;
  if (N == 0) return B200_OK;
  int64_t grid = ceil_div(N, 256);
  if (grid > (int64_t)kNumSMs * 16) grid = (int64_t)kNumSMs * 16;
  linearize_seq_kernel<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>(
      indices, offsets, index_dtype == B200_I64, hash_size_cumsum, T, B, N, out);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" const char* b200_error_string(int code) {
  switch (code) {
    case B200_OK:
      return "ok";
    case B200_EINVAL:
      return "b200: invalid argument";
    case B200_EUNSUPPORTED:
      return "b200: configuration not supported by this build (see DESIGN.md)";
    case B200_ETOOBIG:
      return "b200: problem needs more scratch than this fixed-scratch entry point owns (use the variant "
             "that takes a scratch pointer)";
    default:
      return code > 0 ? cudaGetErrorString((cudaError_t)code) : "b200: unknown error";
  }
}

extern "C" int b200_abi_version(void) { return 5; }

extern "C" int b200_tbe_grad_indice_weights(const float* grad_output, int64_t grad_stride,
                                            const void* weights, int weight_dtype,
                                            const int64_t* weights_offsets,
                                            const int32_t* D_offsets, int32_t T, int32_t B,
                                            const void* indices, const void* offsets,
                                            int index_dtype, float* grad_indice_weights,
                                            void* stream) {
  using namespace b200;
  if (T == 0 || B == 0) return B200_OK;
  if (index_dtype != B200_I32 && index_dtype != B200_I64) return B200_EINVAL;
  const int tiles = (B + 31) / 32;
  const unsigned grid = (unsigned)ceil_div((int64_t)T * tiles, 4);
  const bool idx64 = index_dtype == B200_I64;
  cudaStream_t s = (cudaStream_t)stream;
  switch (weight_dtype) {
    case B200_F32:
      grad_indice_weights_kernel<float, 4><<<grid, 128, 0, s>>>(
          grad_output, grad_stride, (const float*)weights, weights_offsets, D_offsets, T, B, tiles,
          indices, offsets, idx64, grad_indice_weights);
      break;
    case B200_F16:
      grad_indice_weights_kernel<__half, 4><<<grid, 128, 0, s>>>(
          grad_output, grad_stride, (const __half*)weights, weights_offsets, D_offsets, T, B,
          tiles, indices, offsets, idx64, grad_indice_weights);
      break;
    case B200_BF16:
      grad_indice_weights_kernel<__nv_bfloat16, 4><<<grid, 128, 0, s>>>(
          grad_output, grad_stride, (const __nv_bfloat16*)weights, weights_offsets, D_offsets, T,
          B, tiles, indices, offsets, idx64, grad_indice_weights);
      break;
    default:
      return B200_EUNSUPPORTED;
  }
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_pooled_a2a_recat(const float* src, float* dst, const int32_t* dim_cumsum,
                                     int32_t W, int32_t B_local, int forward, float scale,
                                     void* stream) {
  using namespace b200;
  if (W <= 0 || B_local <= 0) return B200_OK;
  pooled_a2a_recat_kernel<<<kNumSMs * 8, 256, 0, (cudaStream_t)stream>>>(
      src, dst, dim_cumsum, W, B_local, forward != 0, scale);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// ---------------------------------------------------------------------------------------------
// Row-wise sharding, receive side of the fused reduce-scatter: every rank r wrote its partial sums
// for my B_local samples into slot r of `stage` ([W][B_local][Dw], peer stores from
// b200_tbe_forward_peer); the pooled rows are their sum in rank order (deterministic), written into
// columns [out_col, out_col + Dw) of the [B_local, pitch] output.
namespace b200 {
__global__ void __launch_bounds__(256)
peer_stage_reduce_kernel(float* __restrict__ stage, int32_t W, int64_t rows, int32_t Dw4,
                         float* __restrict__ out, int64_t pitch, const int32_t* __restrict__ out_col4,
                         bool clear) {
  const int64_t total = rows * Dw4;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = i / Dw4;
    const int c = (int)(i - b * Dw4);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r = 0; r < W; ++r) {
      float4* slot = reinterpret_cast<float4*>(stage) + (int64_t)r * total + i;
      const float4 f = __ldcg(slot);  // written by peers, re-zeroed below: no non-coherent path
      const uint4 u = make_uint4(__float_as_uint(f.x), __float_as_uint(f.y), __float_as_uint(f.z),
                                 __float_as_uint(f.w));
      // consume: the next step's peers only store the non-empty bags, the rest must read as zero
      if (clear && (u.x | u.y | u.z | u.w)) *slot = make_float4(0.f, 0.f, 0.f, 0.f);
      acc.x += __uint_as_float(u.x);
      acc.y += __uint_as_float(u.y);
      acc.z += __uint_as_float(u.z);
      acc.w += __uint_as_float(u.w);
    }
    *reinterpret_cast<float4*>(out + b * pitch + out_col4[c]) = acc;
  }
}
}  // namespace b200

// Cross-rank barrier over NVLink peer memory: one small kernel, no host round trip.  pads[r] is rank r's
// signal pad (uint32[W], symmetric memory); rank `rank` writes the barrier's epoch into slot `rank` of every
// peer's pad with a system-scope release store and waits until every slot of its own pad has reached the epoch.
// The epoch is a DEVICE-resident counter bumped by the kernel itself (every rank runs the same sequence of
// barriers, so the counters agree), which keeps the launch valid inside a replayed CUDA graph.  Everything the
// stream did before the barrier is visible to the peers after it, and vice versa.
namespace b200 {
__global__ void __launch_bounds__(32)
peer_barrier_kernel(uint32_t* const* __restrict__ pads, int rank, int W, uint32_t* __restrict__ epoch_ctr) {
  __shared__ uint32_t s_epoch;
  if (threadIdx.x == 0) s_epoch = ++(*epoch_ctr);
  __syncwarp();
  const uint32_t epoch = s_epoch;
  __threadfence_system();
  for (int t = threadIdx.x; t < W; t += 32) {
    uint32_t* remote = pads[t] + rank;
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(remote), "r"(epoch) : "memory");
  }
  for (int t = threadIdx.x; t < W; t += 32) {
    const uint32_t* mine = pads[rank] + t;
    uint32_t v;
    do {
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
    } while ((int32_t)(v - epoch) < 0);
  }
  __threadfence_system();
}
}  // namespace b200

extern "C" int b200_peer_barrier(uint32_t* const* pads, int32_t rank, int32_t W, uint32_t* epoch_ctr,
                                 void* stream) {
  if (!pads || !epoch_ctr || W <= 0 || rank < 0 || rank >= W) return B200_EINVAL;
  b200::peer_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(pads, rank, W, epoch_ctr);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// Flag-driven variant: flags[b][f][r] != 0 says that rank r stored a partial sum for feature f of sample b.
// One warp per (sample, feature): it reads the W flag bytes, adds only the flagged slots in rank order
// (deterministic), clears the flags it consumed and writes the pooled row.  The staging data itself is never
// cleared - unflagged slots are simply not read - so the pass moves ~1/W of the staging buffer when bags
// have one id.
namespace b200 {
__global__ void __launch_bounds__(256)
peer_stage_reduce_flagged_kernel(const float* __restrict__ stage, uint8_t* __restrict__ flags, int32_t W,
                                 int64_t rows, int32_t F, int32_t Dw4,
                                 const int32_t* __restrict__ feat_col4, float* __restrict__ out,
                                 int64_t pitch, const int32_t* __restrict__ out_col4) {
  const int lane = threadIdx.x & 31;
  const int64_t pair = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (pair >= rows * F) return;
  const int64_t b = pair / F;
  const int f = (int)(pair - b * F);
  uint8_t* fl = flags + pair * W;
  unsigned mask = 0;
  if (lane < W) mask = fl[lane] ? 1u : 0u;
  mask = __ballot_sync(kFull, mask != 0);
  if (lane < W && ((mask >> lane) & 1u)) fl[lane] = 0;
  const int c0 = feat_col4[f], c1 = feat_col4[f + 1];
  const int64_t slot = rows * Dw4;  // float4s per rank slot
  for (int c = c0 + lane; c < c1; c += 32) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (unsigned m = mask; m; m &= m - 1) {
      const int r = __ffs(m) - 1;
      const float4 v = __ldcg(reinterpret_cast<const float4*>(stage) + (int64_t)r * slot + b * Dw4 + c);
      acc.x += v.x;
      acc.y += v.y;
      acc.z += v.z;
      acc.w += v.w;
    }
    *reinterpret_cast<float4*>(out + b * pitch + out_col4[c]) = acc;
  }
}
}  // namespace b200

extern "C" int b200_peer_stage_reduce_flagged(const float* stage, uint8_t* flags, int32_t W,
                                              int32_t B_local, int32_t Dw, int32_t F,
                                              const int32_t* feat_col4, float* out, int64_t out_pitch,
                                              const int32_t* out_col4, void* stream) {
  using namespace b200;
  if (W <= 0 || W > 32 || B_local < 0 || Dw < 0 || (Dw & 3) || (out_pitch & 3) || F <= 0 || !out_col4 ||
      !feat_col4 || !flags)
    return B200_EINVAL;
  if (B_local == 0 || Dw == 0) return B200_OK;
  const int64_t warps = (int64_t)B_local * F;
  peer_stage_reduce_flagged_kernel<<<(unsigned)ceil_div(warps, 8), 256, 0, (cudaStream_t)stream>>>(
      stage, flags, W, B_local, F, Dw >> 2, feat_col4, out, out_pitch, out_col4);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_peer_stage_reduce(float* stage, int32_t W, int32_t B_local, int32_t Dw,
                                      float* out, int64_t out_pitch, const int32_t* out_col4,
                                      int32_t clear_stage, void* stream) {
  using namespace b200;
  if (W <= 0 || B_local < 0 || Dw < 0 || (Dw & 3) || (out_pitch & 3) || !out_col4) return B200_EINVAL;
  if (B_local == 0 || Dw == 0) return B200_OK;
  const int64_t total = (int64_t)B_local * (Dw >> 2);
  int64_t grid = ceil_div(total, 256);
  if (grid > (int64_t)kNumSMs * 16) grid = (int64_t)kNumSMs * 16;
  peer_stage_reduce_kernel<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>(
      stage, W, B_local, Dw >> 2, out, out_pitch, out_col4, clear_stage != 0);
  B200_LAUNCH_CHECK();
  return B200_OK;
}
