// int8 rowwise-quantized inference path for sm_100a.
//
// Replaces INT8_split_embedding_codegen_forward_{un,}weighted_kernel_small_L
// (fbgemm_gpu/codegen/embedding_forward_quantized_split_template.cu:154-450) and
// FloatToFusedNBitRowwiseQuantizedSBHalf for bit_rate 8 (fbgemm_gpu/src/quantize_ops.cu).
//
// Row layout (split_table_batched_embeddings_ops.py:1469-1485,2244-2249):
//   [fp16 scale][fp16 bias][D x uint8][pad to row_alignment bytes].
// out[b, :] = sum_p (scale_r * q_r[:] + bias_r); idx == -1 marks a pruned row (zeros).
//
// Same tile design as the training forward (tbe_forward.cu): a warp owns 32 consecutive
// bags of one table and walks their ids as one stream with several 16-byte row chunks in
// flight, so small-L inference batches still keep many loads outstanding.
#include "common.cuh"

namespace b200 {

constexpr int kI8Threads = 128;

__device__ __forceinline__ void unpack16(const uint4& r, float (&q)[16]) {
  const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    q[4 * i + 0] = (float)(w[i] & 0xffu);
    q[4 * i + 1] = (float)((w[i] >> 8) & 0xffu);
    q[4 * i + 2] = (float)((w[i] >> 16) & 0xffu);
    q[4 * i + 3] = (float)(w[i] >> 24);
  }
}

template <typename O, int NV>
__global__ void __launch_bounds__(kI8Threads)
tbe_int8_forward_kernel(const uint8_t* __restrict__ weights,
                        const int64_t* __restrict__ weights_offsets,
                        const int32_t* __restrict__ D_offsets, int32_t T, int32_t B,
                        int32_t total_D, int32_t tiles_per_table, int32_t bpt, int32_t row_alignment,
                        const void* __restrict__ indices, const void* __restrict__ offsets,
                        bool idx64, const float* __restrict__ psw, bool mean,
                        O* __restrict__ out) {
  constexpr int U = NV == 1 ? 4 : (NV == 2 ? 2 : 1);
  const int lane = threadIdx.x & 31;
  const int64_t tile = (int64_t)blockIdx.x * (kI8Threads / 32) + (threadIdx.x >> 5);
  if (tile >= (int64_t)T * tiles_per_table) return;
  const int t = (int)(tile / tiles_per_table);
  const int b0 = (int)(tile - (int64_t)t * tiles_per_table) * bpt;
  const int nb = min(bpt, B - b0);  // bags of this tile (lane k <-> bag k; lanes >= nb are idle)
  const int D_begin = D_offsets[t];
  const int D = D_offsets[t + 1] - D_begin;
  const int64_t row_bytes = ((int64_t)(D + 4) + row_alignment - 1) / row_alignment * row_alignment;
  const uint8_t* __restrict__ wtab = weights + weights_offsets[t];

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

  // chunk c = lane + 32*v covers row bytes [16c, 16c+16) = elements [16c-4, 16c+12)
  auto flush = [&](int bag, float (&a)[NV][16], float scale) {
#pragma unroll
    for (int v = 0; v < NV; ++v) {
      const int e0 = 16 * (lane + 32 * v) - 4;
#pragma unroll
      for (int k = 0; k < 16; k += 4) {
        const int e = e0 + k;  // multiple of 4; D % 4 == 0 so groups of 4 are all-in or all-out
        if (e >= 0 && e < D) {
          float f[4] = {a[v][k] * scale, a[v][k + 1] * scale, a[v][k + 2] * scale,
                        a[v][k + 3] * scale};
          store_vec<O, 4, true>(out_tile + (int64_t)bag * total_D + e, f);
        }
      }
    }
  };

  float acc[NV][16];
#pragma unroll
  for (int v = 0; v < NV; ++v)
#pragma unroll
    for (int k = 0; k < 16; ++k) acc[v][k] = 0.f;
  for (unsigned m = empty; m; m &= m - 1) flush(__ffs(m) - 1, acc, 1.f);
  if (len <= 0) return;

  int cur_bag = remaining ? (__ffs(remaining) - 1) : 0;
  int32_t cur_end = __shfl_sync(kFull, rel_end, cur_bag);
  int32_t prev_end = 0;
  const bool weighted = psw != nullptr;

  for (int32_t q = 0; q < len; q += U) {
    int64_t idx[U];
    float pw[U];
    uint4 raw[U][NV];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      idx[u] = -1;
      pw[u] = 1.f;
      if (q + u < len) {
        idx[u] = load_index(indices, S0 + q + u, idx64);
        if (weighted) pw[u] = psw[S0 + q + u];
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
#pragma unroll
      for (int v = 0; v < NV; ++v) {
        const int c = lane + 32 * v;
        raw[u][v] = make_uint4(0, 0, 0, 0);
        if (idx[u] >= 0 && 16 * c < D + 4)
          raw[u][v] = ldg_stream_16(wtab + idx[u] * row_bytes + 16 * c);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (q + u < len) {
        // header lives in the first 4 bytes of chunk 0 (lane 0, v = 0)
        const uint32_t hdr = __shfl_sync(kFull, raw[u][0].x, 0);
        float sc, bi;
        ElemCvt<__half>::unpack2(hdr, sc, bi);
        sc *= pw[u];
        bi *= pw[u];
        if (idx[u] < 0) {  // pruned row contributes zeros
          sc = 0.f;
          bi = 0.f;
        }
#pragma unroll
        for (int v = 0; v < NV; ++v) {
          float qv[16];
          unpack16(raw[u][v], qv);
#pragma unroll
          for (int k = 0; k < 16; ++k) acc[v][k] = fmaf(sc, qv[k], acc[v][k] + bi);
        }
        if (q + u + 1 == cur_end) {
          const float scale = mean ? 1.0f / (float)(cur_end - prev_end) : 1.0f;
          flush(cur_bag, acc, scale);
#pragma unroll
          for (int v = 0; v < NV; ++v)
#pragma unroll
            for (int k = 0; k < 16; ++k) acc[v][k] = 0.f;
          prev_end = cur_end;
          remaining &= remaining - 1;
          cur_bag = remaining ? (__ffs(remaining) - 1) : 0;
          cur_end = remaining ? __shfl_sync(kFull, rel_end, cur_bag) : 0x7fffffff;
        }
      }
    }
  }
}

// One warp per row: min/max, then q = round((x - min) / scale) with fp16-rounded scale/bias.
__global__ void __launch_bounds__(256)
quantize_rowwise_8bit_sbhalf_kernel(const float* __restrict__ in, int64_t nrows, int32_t ncols,
                                    uint8_t* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (row >= nrows) return;
  const float* x = in + row * ncols;
  uint8_t* o = out + row * ((int64_t)ncols + 4);
  float mn = INFINITY, mx = -INFINITY;
  for (int c = lane; c < ncols; c += 32) {
    const float v = x[c];
    mn = fminf(mn, v);
    mx = fmaxf(mx, v);
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) {
    mn = fminf(mn, __shfl_xor_sync(kFull, mn, s));
    mx = fmaxf(mx, __shfl_xor_sync(kFull, mx, s));
  }
  // quantize_ops_cpu / fbgemm FloatToFusedNBitRowwiseQuantizedSBHalf: bias = fp16(min),
  // scale = fp16((max - bias) / 255), falling back to 1 when the range is empty.
  const __half hmin = __float2half_rn(mn);
  const float minimum = __half2float(hmin);
  const float range = mx - minimum;
  __half hscale = __float2half_rn(range == 0.f ? 1.0f : range / 255.0f);
  float scale = __half2float(hscale);
  if (scale == 0.f) {
    // range too small for fp16: avoid a division by zero below
    scale = 1.0f;
    hscale = __float2half_rn(1.0f);
  }
  float inv = 1.0f / scale;
  if (isinf(inv)) {
    scale = 1.0f;
    hscale = __float2half_rn(1.0f);
    inv = 1.0f;
  }
  for (int c = lane; c < ncols; c += 32) {
    const float q = rintf((x[c] - minimum) * inv);
    o[c] = (uint8_t)fminf(fmaxf(q, 0.f), 255.f);
  }
  if (lane == 0) {
    *reinterpret_cast<__half*>(o + ncols) = hscale;
    *reinterpret_cast<__half*>(o + ncols + 2) = hmin;
  }
}

template <typename O>
static int int8_launch(int vecs, unsigned grid, cudaStream_t s, const uint8_t* weights,
                       const int64_t* wo, const int32_t* Do, int32_t T, int32_t B, int32_t total_D,
                       int32_t tiles, int32_t bpt, int32_t align, const void* indices, const void* offsets,
                       bool idx64, const float* psw, bool mean, void* out) {
#define B200_I8(NV)                                                                           \
  tbe_int8_forward_kernel<O, NV><<<grid, kI8Threads, 0, s>>>(weights, wo, Do, T, B, total_D,  \
                                                            tiles, bpt, align, indices, offsets, \
                                                            idx64, psw, mean, (O*)out)
  if (vecs <= 1)
    B200_I8(1);
  else if (vecs <= 2)
    B200_I8(2);
  else if (vecs <= 4)
    B200_I8(4);
  else
    return B200_EUNSUPPORTED;
#undef B200_I8
  B200_LAUNCH_CHECK();
  return B200_OK;
}

}  // namespace b200

extern "C" int b200_quantize_rowwise_8bit_sbhalf(const float* in, int64_t nrows, int32_t ncols,
                                                 uint8_t* out, void* stream) {
  using namespace b200;
  if (nrows == 0) return B200_OK;
  if (ncols <= 0 || (ncols & 1)) return B200_EINVAL;  // fp16 trailer must stay 2-byte aligned
  quantize_rowwise_8bit_sbhalf_kernel<<<(unsigned)ceil_div(nrows * 32, 256), 256, 0,
                                        (cudaStream_t)stream>>>(in, nrows, ncols, out);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_tbe_int8_forward(const uint8_t* weights, const int64_t* weights_offsets,
                                     const int32_t* D_offsets, int32_t T, int32_t B,
                                     int32_t total_D, int32_t max_D, int32_t row_alignment,
                                     const void* indices, const void* offsets, int index_dtype,
                                     const float* per_sample_weights, int pooling_mode,
                                     void* output, int output_dtype, void* stream) {
  using namespace b200;
  if (index_dtype != B200_I32 && index_dtype != B200_I64) return B200_EINVAL;
  if (row_alignment < 16 || (row_alignment % 16) != 0 || (total_D % 4) != 0) return B200_EINVAL;
  if (T == 0 || B == 0) return B200_OK;
  // bags per warp tile: 32 for big batches; small inference batches (cfg5: 26 x 4096 bags) get
  // smaller tiles so that the grid still fills 148 SMs x 64 warps
  int bpt = 32;
  while (bpt > 4 && (int64_t)T * ceil_div(B, bpt) < (int64_t)kNumSMs * 64) bpt >>= 1;
  const int tiles = (int)ceil_div(B, bpt);
  const unsigned grid = (unsigned)ceil_div((int64_t)T * tiles, kI8Threads / 32);
  const int vecs = (int)ceil_div(max_D + 4, 16 * 32);
  cudaStream_t s = (cudaStream_t)stream;
  const bool idx64 = index_dtype == B200_I64;
  const bool mean = pooling_mode == B200_POOL_MEAN;
  if (output_dtype == B200_F16)
    return int8_launch<__half>(vecs, grid, s, weights, weights_offsets, D_offsets, T, B, total_D,
                               tiles, bpt, row_alignment, indices, offsets, idx64,
                               per_sample_weights, mean, output);
  if (output_dtype == B200_F32)
    return int8_launch<float>(vecs, grid, s, weights, weights_offsets, D_offsets, T, B, total_D,
                              tiles, bpt, row_alignment, indices, offsets, idx64,
                              per_sample_weights, mean, output);
  return B200_EUNSUPPORTED;
}
