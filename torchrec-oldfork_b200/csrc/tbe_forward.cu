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
//  * ids travel as one coalesced load per 32 positions (next chunk in flight while the current
//    one is used) and are broadcast through a 64-entry shared-memory ring.
//  * rows stream through a warp-private cp.async ring (8 stages of one row): every lane copies
//    and later consumes its own 16-byte slot, so there are no barriers, the kernel needs ~34
//    registers and 12+ CTAs/SM keep >100 KB of row reads in flight per SM.
//  * the pooled row is written once with streaming 128-bit stores.
#include "common.cuh"

namespace b200 {

constexpr int kFwdThreads = 128;
constexpr int kFwdWarps = kFwdThreads / 32;

// Ring geometry: one stage = one table row laid out as 16-byte lane slots ([v][lane]).
template <int NV>
struct FwdRing {
  static constexpr int kStageBytes = NV * 512;
  static constexpr int kStages = NV >= 4 ? 2 : (NV == 2 ? 4 : 8);  // power of two
  static constexpr int kWarpBytes = kStages * kStageBytes + 64 * 8 + 64 * 4;  // + id ring + weight ring
};

__device__ __forceinline__ void fwd_cp16(void* dst_smem, const void* src) {
  const uint32_t sa = (uint32_t)__cvta_generic_to_shared(dst_smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(src) : "memory");
}
__device__ __forceinline__ void fwd_cp8(void* dst_smem, const void* src) {
  const uint32_t sa = (uint32_t)__cvta_generic_to_shared(dst_smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sa), "l"(src) : "memory");
}
__device__ __forceinline__ void fwd_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void fwd_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// one lane's staged piece of a row (EPV elements of W) -> fp32
template <typename W, int EPV>
__device__ __forceinline__ void fwd_load_slot(const unsigned char* p, float (&f)[EPV]) {
  if constexpr (sizeof(W) == 4) {
    const float4 x = *reinterpret_cast<const float4*>(p);
    f[0] = x.x;
    f[1] = x.y;
    f[2] = x.z;
    f[3] = x.w;
  } else if constexpr (EPV == 8) {
    const uint4 r = *reinterpret_cast<const uint4*>(p);
    ElemCvt<W>::unpack2(r.x, f[0], f[1]);
    ElemCvt<W>::unpack2(r.y, f[2], f[3]);
    ElemCvt<W>::unpack2(r.z, f[4], f[5]);
    ElemCvt<W>::unpack2(r.w, f[6], f[7]);
  } else {
    const uint2 r = *reinterpret_cast<const uint2*>(p);
    ElemCvt<W>::unpack2(r.x, f[0], f[1]);
    ElemCvt<W>::unpack2(r.y, f[2], f[3]);
  }
}

// Destination of the pooled rows when they are written straight into (peer) output buffers:
// bag b goes to bases[b / B_dst] at row b % B_dst, column col[t], rows `pitch` elements apart.
// With NVLink peer mappings in `bases` this IS the output all-to-all of table-wise sharding (and the
// scatter half of row-wise sharding's reduce-scatter): no pooled [B_global, D_local] tensor, no NCCL
// pass and no re-layout copy exist.
struct FwdPeerDst {
  void* const* bases;
  const int32_t* col;
  int64_t pitch;
  int32_t B_dst;
  int32_t skip_empty;  // empty bags store nothing (row-wise partial sums)
  // Optional presence flags (row-wise sharding): every NON-EMPTY bag also sets the byte
  // flag_bases[b / B_dst][(b % B_dst) * flag_pitch + flag_col[t]], so that the reduction on the destination reads
  // only the slots that were written (with one id per bag that is 1/W of the staging buffer).
  uint8_t* const* flag_bases;
  const int32_t* flag_col;
  int64_t flag_pitch;
};

template <typename W, typename O, int EPV, int NV, bool PEER>
__global__ void __launch_bounds__(kFwdThreads)
tbe_forward_kernel(const W* __restrict__ weights, const int64_t* __restrict__ weights_offsets,
                   const int32_t* __restrict__ D_offsets, int32_t T, int32_t B, int32_t total_D,
                   int32_t tiles_per_table, int32_t bpt, const void* __restrict__ indices,
                   const void* __restrict__ offsets, bool idx64,
                   const float* __restrict__ psw, bool mean, O* __restrict__ out, FwdPeerDst peer) {
  constexpr int kStages = FwdRing<NV>::kStages;
  constexpr int kStageBytes = FwdRing<NV>::kStageBytes;
  constexpr int kAhead = kStages - 1;
  extern __shared__ __align__(16) unsigned char fwd_smem[];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int64_t tile = (int64_t)blockIdx.x * kFwdWarps + warp;
  if (tile >= (int64_t)T * tiles_per_table) return;
  const int t = (int)(tile / tiles_per_table);
  const int b0 = (int)(tile - (int64_t)t * tiles_per_table) * bpt;
  const int nb = min(bpt, B - b0);  // bags of this tile (lane k <-> bag k; lanes >= nb are idle)
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

  if constexpr (PEER) {
    if (peer.flag_bases != nullptr && in_tile && rel_end > rel_start) {
      const int b = b0 + lane;
      const int r = b / peer.B_dst;
      peer.flag_bases[r][(int64_t)(b - r * peer.B_dst) * peer.flag_pitch + peer.flag_col[t]] = 1;
    }
  }
  O* __restrict__ out_tile = PEER ? nullptr : out + (int64_t)b0 * total_D + D_begin + EPV * lane;
  const int peer_col = PEER ? peer.col[t] + EPV * lane : 0;
  auto row_ptr = [&](int bag) -> O* {
    if constexpr (PEER) {
      const int b = b0 + bag;
      const int r = b / peer.B_dst;
      return reinterpret_cast<O*>(peer.bases[r]) + (int64_t)(b - r * peer.B_dst) * peer.pitch + peer_col;
    } else {
      return out_tile + (int64_t)bag * total_D;
    }
  };
  bool act[NV];
#pragma unroll
  for (int v = 0; v < NV; ++v) act[v] = EPV * (lane + 32 * v) < D;

  // Empty bags pool to zeros.
  for (unsigned m = (PEER && peer.skip_empty) ? 0u : empty; m; m &= m - 1) {
    const int j = __ffs(m) - 1;
    float z[EPV];
#pragma unroll
    for (int e = 0; e < EPV; ++e) z[e] = 0.f;
#pragma unroll
    O* zr = row_ptr(j);
#pragma unroll
    for (int v = 0; v < NV; ++v)
      if (act[v]) store_vec<O, EPV, true>(zr + EPV * 32 * v, z);
  }
  if (len <= 0) {
    if constexpr (PEER) __threadfence_system();
    return;
  }

  unsigned char* const wsm = fwd_smem + (size_t)warp * FwdRing<NV>::kWarpBytes;
  unsigned char* const lring = wsm + lane * 16;
  int64_t* const sid = reinterpret_cast<int64_t*>(wsm + kStages * kStageBytes);  // [64] ids, 2 chunks
  float* const sw = reinterpret_cast<float*>(sid + 64);                           // [64] weights
  const bool weighted = psw != nullptr;
  const W* const lrow = wtab + EPV * lane;

  // ids travel one coalesced load per 32 positions; chunk c+1 is in flight while chunk c is used
  const int nchunks = (len + 31) >> 5;
  auto load_chunk = [&](int c, int64_t& id, float& w) {
    const int32_t p = 32 * c + lane;
    id = 0;
    w = 1.f;
    if (p < len) {
      id = load_index(indices, S0 + p, idx64);
      if (weighted) w = psw[S0 + p];
    }
  };
  int64_t id_nxt;
  float w_nxt;
  load_chunk(0, id_nxt, w_nxt);
  sid[lane] = id_nxt;
  sw[lane] = w_nxt;
  __syncwarp();
  if (nchunks > 1) load_chunk(1, id_nxt, w_nxt);

  auto issue = [&](int32_t p, int stage) {
    const W* row = lrow + sid[p & 63] * D;
    unsigned char* st = lring + stage * kStageBytes;
#pragma unroll
    for (int v = 0; v < NV; ++v)
      if (act[v]) {
        if constexpr (EPV * sizeof(W) == 16) {
          fwd_cp16(st + v * 512, row + EPV * 32 * v);
        } else {
          fwd_cp8(st + v * 512, row + EPV * 32 * v);
        }
      }
  };

  float acc[NV][EPV];
#pragma unroll
  for (int v = 0; v < NV; ++v)
#pragma unroll
    for (int e = 0; e < EPV; ++e) acc[v][e] = 0.f;
  int cur_bag = __ffs(remaining) - 1;  // len > 0 => some bag is non-empty
  int32_t cur_end = __shfl_sync(kFull, rel_end, cur_bag);
  int32_t prev_end = 0;

#pragma unroll
  for (int q = 0; q < kAhead; ++q) {
    if (q < len) issue(q, q);
    fwd_commit();
  }
  for (int c = 0; c < nchunks; ++c) {
    __syncwarp();
    if (c + 1 < nchunks) {  // publish the ids of the next chunk (its slot was last read a chunk ago)
      sid[((c + 1) & 1) * 32 + lane] = id_nxt;
      sw[((c + 1) & 1) * 32 + lane] = w_nxt;
    }
    __syncwarp();
    if (c + 2 < nchunks) load_chunk(c + 2, id_nxt, w_nxt);
#pragma unroll 1
    for (int p0 = 32 * c; p0 < 32 * c + 32 && p0 < len; p0 += 8) {
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const int32_t p = p0 + q;
        if (p + kAhead < len) issue(p + kAhead, (q + kAhead) & (kStages - 1));
        fwd_commit();
        if (p < len) {
          fwd_wait<kAhead>();
          const unsigned char* st = lring + (q & (kStages - 1)) * kStageBytes;
          const float wp = weighted ? sw[p & 63] : 1.f;
#pragma unroll
          for (int v = 0; v < NV; ++v)
            if (act[v]) {
              float x[EPV];
              fwd_load_slot<W, EPV>(st + v * 512, x);
              if (weighted) {
#pragma unroll
                for (int e = 0; e < EPV; ++e) acc[v][e] = fmaf(wp, x[e], acc[v][e]);
              } else {
#pragma unroll
                for (int e = 0; e < EPV; ++e) acc[v][e] += x[e];
              }
            }
          if (p + 1 == cur_end) {  // bag boundary: write the pooled row once
            const float scale = mean ? 1.0f / (float)(cur_end - prev_end) : 1.0f;
            O* orow = row_ptr(cur_bag);
#pragma unroll
            for (int v = 0; v < NV; ++v) {
              if (mean) {
#pragma unroll
                for (int e = 0; e < EPV; ++e) acc[v][e] *= scale;
              }
              if (act[v]) store_vec<O, EPV, true>(orow + EPV * 32 * v, acc[v]);
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
    }
  }
  fwd_wait<0>();
  if constexpr (PEER) __threadfence_system();  // peer stores visible before the cross-rank barrier
}

template <typename W, typename O, int EPV>
static int launch_forward_nv(const void* weights, const int64_t* weights_offsets,
                             const int32_t* D_offsets, int32_t T, int32_t B, int32_t total_D,
                             int32_t max_D, const void* indices, const void* offsets, bool idx64,
                             const float* psw, bool mean, void* output, const FwdPeerDst* peer,
                             cudaStream_t s) {
  // bags per warp tile: 32, fewer when the batch is too small to fill 148 SMs x 48 warps
  int bpt = 32;
  while (bpt > 4 && (int64_t)T * ceil_div(B, bpt) < (int64_t)kNumSMs * 48) bpt >>= 1;
  const int tiles_per_table = (int)ceil_div(B, bpt);
  const int64_t tiles = (int64_t)T * tiles_per_table;
  const int64_t grid = ceil_div(tiles, kFwdThreads / 32);
  if (grid == 0) return B200_OK;
  if (grid > 0x7fffffffLL) return B200_EINVAL;
  const int vecs = (int)ceil_div(max_D, EPV * 32);
#define B200_FWD_LAUNCH(NV)                                                                     \
  if (peer)                                                                                     \
    tbe_forward_kernel<W, O, EPV, NV, true>                                                      \
        <<<(unsigned)grid, kFwdThreads, kFwdWarps * FwdRing<NV>::kWarpBytes, s>>>(               \
            (const W*)weights, weights_offsets, D_offsets, T, B, total_D, tiles_per_table, bpt,  \
            indices, offsets, idx64, psw, mean, nullptr, *peer);                                 \
  else                                                                                          \
    tbe_forward_kernel<W, O, EPV, NV, false>                                                     \
        <<<(unsigned)grid, kFwdThreads, kFwdWarps * FwdRing<NV>::kWarpBytes, s>>>(               \
            (const W*)weights, weights_offsets, D_offsets, T, B, total_D, tiles_per_table, bpt,  \
            indices, offsets, idx64, psw, mean, (O*)output, FwdPeerDst{})
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
                              const float* psw, bool mean, void* output, const FwdPeerDst* peer,
                              cudaStream_t s) {
  switch (output_dtype) {
    case B200_F32:
      return launch_forward_nv<W, float, EPV>(weights, wo, Do, T, B, total_D, max_D, indices,
                                              offsets, idx64, psw, mean, output, peer, s);
    case B200_F16:
      return launch_forward_nv<W, __half, EPV>(weights, wo, Do, T, B, total_D, max_D, indices,
                                               offsets, idx64, psw, mean, output, peer, s);
    default:
      return B200_EUNSUPPORTED;
  }
}

// tbe_forward_packed.cu: rows of <= 256 bytes, several rows per warp instruction
int launch_forward_packed(int weight_dtype, int output_dtype, bool all_D_mult8, const void* weights,
                          const int64_t* wo, const int32_t* Do, int32_t T, int32_t B, int32_t total_D,
                          int32_t max_D, const void* indices, const void* offsets, bool idx64,
                          const float* psw, bool mean, void* out, cudaStream_t s);

}  // namespace b200

static int tbe_forward_impl(const void* weights, int weight_dtype,
                                const int64_t* weights_offsets, const int32_t* D_offsets,
                                int32_t T, int32_t B, int32_t total_D, int32_t max_D,
                                int32_t all_D_mult8, const void* indices, const void* offsets,
                                int index_dtype, const float* per_sample_weights,
                                int pooling_mode, void* output, int output_dtype,
                            const b200::FwdPeerDst* peer, void* stream) {
  using namespace b200;
  if (T < 0 || B < 0 || total_D < 0 || (total_D % 4) != 0) return B200_EINVAL;
  if (index_dtype != B200_I32 && index_dtype != B200_I64) return B200_EINVAL;
  if (T == 0 || B == 0) return B200_OK;
  const bool idx64 = index_dtype == B200_I64;
  const bool mean = pooling_mode == B200_POOL_MEAN;
  cudaStream_t s = (cudaStream_t)stream;
  if (!peer) {
    const int r = launch_forward_packed(weight_dtype, output_dtype, all_D_mult8 != 0, weights,
                                        weights_offsets, D_offsets, T, B, total_D, max_D, indices,
                                        offsets, idx64, per_sample_weights, mean, output, s);
    if (r != B200_EUNSUPPORTED) return r;
  }
  switch (weight_dtype) {
    case B200_F32:
      return launch_forward_out<float, 4>(output_dtype, weights, weights_offsets, D_offsets, T, B,
                                          total_D, max_D, indices, offsets, idx64,
                                          per_sample_weights, mean, output, peer, s);
    case B200_F16:
      if (all_D_mult8)
        return launch_forward_out<__half, 8>(output_dtype, weights, weights_offsets, D_offsets, T,
                                             B, total_D, max_D, indices, offsets, idx64,
                                             per_sample_weights, mean, output, peer, s);
      return launch_forward_out<__half, 4>(output_dtype, weights, weights_offsets, D_offsets, T, B,
                                           total_D, max_D, indices, offsets, idx64,
                                           per_sample_weights, mean, output, peer, s);
    case B200_BF16:
      if (all_D_mult8)
        return launch_forward_out<__nv_bfloat16, 8>(output_dtype, weights, weights_offsets,
                                                    D_offsets, T, B, total_D, max_D, indices,
                                                    offsets, idx64, per_sample_weights, mean,
                                                    output, peer, s);
      return launch_forward_out<__nv_bfloat16, 4>(output_dtype, weights, weights_offsets,
                                                  D_offsets, T, B, total_D, max_D, indices,
                                                  offsets, idx64, per_sample_weights, mean, output,
                                                  peer, s);
    default:
      return B200_EUNSUPPORTED;
  }
}

extern "C" int b200_tbe_forward(const void* weights, int weight_dtype,
                                const int64_t* weights_offsets, const int32_t* D_offsets,
                                int32_t T, int32_t B, int32_t total_D, int32_t max_D,
                                int32_t all_D_mult8, const void* indices, const void* offsets,
                                int index_dtype, const float* per_sample_weights,
                                int pooling_mode, void* output, int output_dtype, void* stream) {
  return tbe_forward_impl(weights, weight_dtype, weights_offsets, D_offsets, T, B, total_D, max_D,
                          all_D_mult8, indices, offsets, index_dtype, per_sample_weights,
                          pooling_mode, output, output_dtype, nullptr, stream);
}

extern "C" int b200_tbe_forward_peer(const void* weights, int weight_dtype,
                                     const int64_t* weights_offsets, const int32_t* D_offsets,
                                     int32_t T, int32_t B, int32_t total_D, int32_t max_D,
                                     int32_t all_D_mult8, const void* indices, const void* offsets,
                                     int index_dtype, const float* per_sample_weights,
                                     int pooling_mode, void* const* dst_bases, int32_t n_dst,
                                     int32_t B_dst, int64_t dst_pitch, const int32_t* dst_col,
                                     int output_dtype, int32_t skip_empty,
                                     uint8_t* const* flag_bases, int64_t flag_pitch,
                                     const int32_t* flag_col, void* stream) {
  if (!dst_bases || !dst_col || n_dst <= 0 || B_dst <= 0 || (int64_t)n_dst * B_dst < B ||
      (dst_pitch & 3))
    return B200_EINVAL;
  if (flag_bases && (!flag_col || flag_pitch <= 0)) return B200_EINVAL;
  b200::FwdPeerDst peer{dst_bases, dst_col, dst_pitch, B_dst, skip_empty ? 1 : 0,
                        flag_bases, flag_col, flag_pitch};
  return tbe_forward_impl(weights, weight_dtype, weights_offsets, D_offsets, T, B, total_D, max_D,
                          all_D_mult8, indices, offsets, index_dtype, per_sample_weights,
                          pooling_mode, nullptr, output_dtype, &peer, stream);
}
