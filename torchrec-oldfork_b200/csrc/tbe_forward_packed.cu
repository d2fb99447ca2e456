// Pooled lookup for NARROW rows (row bytes <= 256): several rows per warp instruction.
//
// Same contract as tbe_forward.cu (embedding_forward_split_template.cu:43-274).  With one row per
// warp-wide load the main kernel spends the same ~40 issue slots on a 64-byte row as on a
// 512-byte one, so at dim 32..128 fp16 (SURVEY cfg4) it is issue-bound at ~14 G rows/s while HBM
// idles.  Here a row is owned by a GROUP of LPR = 4 / 8 / 16 lanes (16 bytes per lane) and the
// R = 32 / LPR groups of a warp gather R rows with each load instruction:
//  * the warp still owns a tile of 32 consecutive bags of one table (lane k holds the offsets of
//    bag k); group g owns the LPR consecutive bags [g*LPR, (g+1)*LPR) and walks THEIR ids, which
//    are one contiguous range of `indices`, as a flat stream - independent of the pooling factor.
//  * ids: lane j of a group loads id[c*LPR + j] (coalesced), the group shuffles them out one per
//    row slot; kU = 4 row slices per lane are in flight as plain 128-bit streaming loads.
//  * bag ends are detected per group; all shuffles run warp-convergent (uniform trip counts), the
//    stores are predicated.  A bag is summed by one group in id order: results are bit-identical
//    to the one-row-per-warp kernel and to the sequential oracle.
#include "common.cuh"

namespace b200 {

constexpr int kPkThreads = 128;
constexpr int kPkU = 4;  // row slots in flight per lane

// One lane's 16-byte (or 8-byte) slice of a row, kept raw until it is accumulated so that kPkU
// slices in flight cost 4 (2) registers each instead of EPV.
template <typename W, int EPV>
struct RawVec {
  static constexpr int kWords = (int)(EPV * sizeof(W) / 4);
  uint32_t r[kWords];
  __device__ __forceinline__ void zero() {
#pragma unroll
    for (int i = 0; i < kWords; ++i) r[i] = 0u;
  }
  __device__ __forceinline__ void load(const W* p) {
    if constexpr (kWords == 4) {
      const uint4 v = ldg_stream_16(p);
      r[0] = v.x, r[1] = v.y, r[2] = v.z, r[3] = v.w;
    } else {
      const uint2 v = ldg_stream_8(p);
      r[0] = v.x, r[1] = v.y;
    }
  }
  __device__ __forceinline__ void unpack(float (&f)[EPV]) const {
    if constexpr (sizeof(W) == 4) {
#pragma unroll
      for (int i = 0; i < EPV; ++i) f[i] = __uint_as_float(r[i]);
    } else {
#pragma unroll
      for (int i = 0; i < kWords; ++i) ElemCvt<W>::unpack2(r[i], f[2 * i], f[2 * i + 1]);
    }
  }
};

template <typename W, typename O, int EPV, int LPR>
__global__ void __launch_bounds__(kPkThreads)
tbe_forward_packed_kernel(const W* __restrict__ weights, const int64_t* __restrict__ weights_offsets,
                          const int32_t* __restrict__ D_offsets, int32_t T, int32_t B, int32_t total_D,
                          int32_t tiles_per_table, const void* __restrict__ indices,
                          const void* __restrict__ offsets, bool idx64, const float* __restrict__ psw,
                          bool mean, O* __restrict__ out) {
  static_assert(LPR >= kPkU && LPR <= 16, "group width");
  const int lane = threadIdx.x & 31;
  const int64_t tile = (int64_t)blockIdx.x * (kPkThreads / 32) + (threadIdx.x >> 5);
  if (tile >= (int64_t)T * tiles_per_table) return;
  const int t = (int)(tile / tiles_per_table);
  const int b0 = (int)(tile - (int64_t)t * tiles_per_table) * 32;
  const int nb = min(32, B - b0);
  const int D_begin = D_offsets[t];
  const int D = D_offsets[t + 1] - D_begin;
  const int gl = lane & (LPR - 1);   // lane within its group
  const int g0 = lane & ~(LPR - 1);  // first lane (= first bag) of the group
  const unsigned gmask = (LPR == 32 ? kFull : ((1u << LPR) - 1u)) << g0;
  const bool act = EPV * gl < D;
  const W* __restrict__ lrow = weights + weights_offsets[t] + EPV * gl;

  // lane k <-> bag k of the tile: offsets relative to the tile's first id
  const int64_t obase = (int64_t)t * B + b0;
  const int64_t my_start = load_index(offsets, obase + min(lane, nb), idx64);
  const int64_t S0 = __shfl_sync(kFull, my_start, 0);
  const int32_t rel_start = (int32_t)(my_start - S0);
  int32_t rel_end = __shfl_down_sync(kFull, rel_start, 1);
  if (lane == 31) rel_end = (int32_t)(load_index(offsets, obase + nb, idx64) - S0);
  const bool in_tile = lane < nb;
  unsigned remaining = __ballot_sync(kFull, in_tile && rel_end > rel_start) & gmask;
  const unsigned empty = __ballot_sync(kFull, in_tile && rel_end <= rel_start) & gmask;

  O* __restrict__ out_tile = out + (int64_t)b0 * total_D + D_begin + EPV * gl;

  // empty bags pool to zeros (each group writes its own)
  if (act) {
    float z[EPV];
#pragma unroll
    for (int e = 0; e < EPV; ++e) z[e] = 0.f;
    for (unsigned m = empty; m; m &= m - 1)
      store_vec<O, EPV, true>(out_tile + (int64_t)(__ffs(m) - 1) * total_D, z);
  }

  // the group's id range [gs, ge) and the warp-uniform trip count
  const int32_t gs = __shfl_sync(kFull, rel_start, g0);
  const int32_t ge = __shfl_sync(kFull, rel_end, g0 + LPR - 1);
  const int32_t glen = ge - gs;
  const int32_t maxlen = __reduce_max_sync(kFull, glen);
  if (maxlen <= 0) return;

  int cur_bag = remaining ? (__ffs(remaining) - 1) : g0;
  int32_t cur_end = __shfl_sync(kFull, rel_end, cur_bag);
  int32_t cur_begin = __shfl_sync(kFull, rel_start, cur_bag);
  if (!remaining) cur_end = 0x7fffffff;
  const bool weighted = psw != nullptr;

  float acc[EPV];
#pragma unroll
  for (int e = 0; e < EPV; ++e) acc[e] = 0.f;

  for (int32_t c = 0; c < maxlen; c += LPR) {
    // one coalesced id (and weight) per lane for the group's next LPR positions
    int64_t id = 0;
    float pw = 1.f;
    if (c + gl < glen) {
      id = load_index(indices, S0 + gs + c + gl, idx64);
      if (weighted) pw = psw[S0 + gs + c + gl];
    }
#pragma unroll
    for (int q0 = 0; q0 < LPR; q0 += kPkU) {
      RawVec<W, EPV> raw[kPkU];
      float wq[kPkU];
#pragma unroll
      for (int u = 0; u < kPkU; ++u) {
        const int64_t idq = __shfl_sync(kFull, id, g0 + q0 + u);
        wq[u] = weighted ? __shfl_sync(kFull, pw, g0 + q0 + u) : 1.f;
        raw[u].zero();  // an all-zero bit pattern is 0.0 in fp32 / fp16 / bf16
        if (act && c + q0 + u < glen) raw[u].load(lrow + idq * D);
      }
#pragma unroll
      for (int u = 0; u < kPkU; ++u) {
        float x[EPV];
        raw[u].unpack(x);
        if (weighted) {
#pragma unroll
          for (int e = 0; e < EPV; ++e) acc[e] = fmaf(wq[u], x[e], acc[e]);
        } else {
#pragma unroll
          for (int e = 0; e < EPV; ++e) acc[e] += x[e];
        }
        const bool hit = (c + q0 + u < glen) && (gs + c + q0 + u + 1 == cur_end);
        if (__any_sync(kFull, hit)) {  // warp-uniform: some group finished a bag
          if (hit) {
            if (mean) {
              const float scale = 1.0f / (float)(cur_end - cur_begin);
#pragma unroll
              for (int e = 0; e < EPV; ++e) acc[e] *= scale;
            }
            if (act) store_vec<O, EPV, true>(out_tile + (int64_t)cur_bag * total_D, acc);
#pragma unroll
            for (int e = 0; e < EPV; ++e) acc[e] = 0.f;
            remaining &= remaining - 1;
            cur_bag = remaining ? (__ffs(remaining) - 1) : g0;
          }
          const int32_t ne = __shfl_sync(kFull, rel_end, cur_bag);
          const int32_t nbg = __shfl_sync(kFull, rel_start, cur_bag);
          if (hit) {
            cur_end = remaining ? ne : 0x7fffffff;
            cur_begin = nbg;
          }
        }
      }
    }
  }
}

template <typename W, typename O, int EPV>
static int launch_packed_lpr(int lpr, const void* weights, const int64_t* wo, const int32_t* Do, int32_t T,
                             int32_t B, int32_t total_D, const void* indices, const void* offsets,
                             bool idx64, const float* psw, bool mean, void* out, cudaStream_t s) {
  const int tiles_per_table = (B + 31) / 32;
  const int64_t grid = ceil_div((int64_t)T * tiles_per_table, kPkThreads / 32);
  if (grid == 0) return B200_OK;
  if (grid > 0x7fffffffLL) return B200_EINVAL;
#define B200_PK(LPR)                                                                              \
  tbe_forward_packed_kernel<W, O, EPV, LPR><<<(unsigned)grid, kPkThreads, 0, s>>>(                  \
      (const W*)weights, wo, Do, T, B, total_D, tiles_per_table, indices, offsets, idx64, psw, mean, \
      (O*)out)
  if (lpr <= 4)
    B200_PK(4);
  else if (lpr <= 8)
    B200_PK(8);
  else
    B200_PK(16);
#undef B200_PK
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// Returns B200_EUNSUPPORTED when the rows are too wide for a packed layout (caller falls back).
int launch_forward_packed(int weight_dtype, int output_dtype, bool all_D_mult8, const void* weights,
                          const int64_t* wo, const int32_t* Do, int32_t T, int32_t B, int32_t total_D,
                          int32_t max_D, const void* indices, const void* offsets, bool idx64,
                          const float* psw, bool mean, void* out, cudaStream_t s) {
  const int epv = (weight_dtype == B200_F32) ? 4 : (all_D_mult8 ? 8 : 4);
  const int lpr = (int)ceil_div(max_D, epv);
  if (lpr > 16) return B200_EUNSUPPORTED;
#define B200_PK_OUT(WT, EPV)                                                                        \
  (output_dtype == B200_F32                                                                         \
       ? launch_packed_lpr<WT, float, EPV>(lpr, weights, wo, Do, T, B, total_D, indices, offsets,   \
                                           idx64, psw, mean, out, s)                                \
       : (output_dtype == B200_F16                                                                  \
              ? launch_packed_lpr<WT, __half, EPV>(lpr, weights, wo, Do, T, B, total_D, indices,    \
                                                   offsets, idx64, psw, mean, out, s)               \
              : (int)B200_EUNSUPPORTED))
  switch (weight_dtype) {
    case B200_F32:
      return B200_PK_OUT(float, 4);
    case B200_F16:
      return all_D_mult8 ? B200_PK_OUT(__half, 8) : B200_PK_OUT(__half, 4);
    case B200_BF16:
      return all_D_mult8 ? B200_PK_OUT(__nv_bfloat16, 8) : B200_PK_OUT(__nv_bfloat16, 4);
    default:
      return B200_EUNSUPPORTED;
  }
#undef B200_PK_OUT
}

}  // namespace b200
