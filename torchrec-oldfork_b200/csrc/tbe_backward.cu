// Fused backward + optimizer of the pooled lookup for sm_100a.
//
// Replaces split_embedding_backward_codegen_<opt>_{un,}weighted_exact_cuda
// (fbgemm_gpu/codegen/embedding_backward_split_template.cu:81-1138) with the optimizer bodies
// of embedding_backward_code_generator.py:376-401 (rowwise Adagrad) and :629-632 (SGD).
//
// Contract kept: duplicate ids are coalesced BEFORE the optimizer, one update per unique
// (table,row) per step, fp32 accumulation, weights updated in place, run-to-run
// deterministic (no float atomics; the per-row sum is taken in stable sorted order).
//
// Design (not the reference's warp-per-row / CTA-per-row split):
//  1. linearise keys (table row-offset + id) and sort (key, b_t) pairs once (csrc/tbe_sort.cu: a hand-written
//     segmented LSD radix sort) — the sort only depends on the ids, so the host layer issues it at forward
//     time on a side stream.
//  2. ONE balanced pass over the sorted positions: each warp owns 32 consecutive sorted
//     positions, keeps 4 gradient rows and the weight rows of the runs that end inside the
//     batch in flight, sums runs in registers and applies the update as soon as a run ends.
//     Work per warp is fixed (32 gradient rows) whatever the run-length distribution, so a
//     3-row Criteo table (runs of ~21k) costs the same per position as a 40M-row one.
//  3. runs that straddle warps are stitched through shared memory in fixed order; runs that
//     straddle CTAs leave one partial row per CTA in scratch and a small second kernel
//     finishes them, again in fixed order.
#include <mutex>

#include <cstdlib>

#include "common.cuh"

namespace b200 {

static inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

// ------------------------------------------------------------------ update
constexpr int kBwdWarps = 8;
constexpr int kBwdThreads = kBwdWarps * 32;
constexpr int kBwdTile = kBwdThreads;  // sorted positions per CTA

struct BwdMeta {  // per-CTA records in scratch, written unconditionally by every CTA
  unsigned long long t_key;
  int32_t t_feat;
  int32_t t_flag;    // a run starts in this CTA and continues into the next
  int32_t has_head;  // the run entering this CTA ends inside it (a head, or end of run/data)
  int32_t pad;
};

__device__ __forceinline__ uint4 ld_na_16(const void* p) {
  uint4 r;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

// fp32 gradient vector: EPV consecutive floats
template <int EPV>
__device__ __forceinline__ void load_grad(const float* p, float (&f)[EPV]) {
#pragma unroll
  for (int i = 0; i < EPV; i += 4) {
    uint4 r = ldg_stream_16(p + i);
    f[i] = __uint_as_float(r.x);
    f[i + 1] = __uint_as_float(r.y);
    f[i + 2] = __uint_as_float(r.z);
    f[i + 3] = __uint_as_float(r.w);
  }
}

template <typename W, int EPV>
__device__ __forceinline__ void load_wrow(const W* p, float (&f)[EPV]) {
  load_vec<W, EPV, false>(p, f);
}

// Optimizer step for one (table,row): acc holds the summed gradient, distributed over the
// warp as element EPV*(lane+32*v)+e.  wv (optional) holds the prefetched weight row.
template <typename W, int EPV, int NV, int OPT, bool PREFETCHED>
__device__ __forceinline__ void apply_update(float (&acc)[NV][EPV], float (&wv)[NV][EPV],
                                             float m_prefetched, W* __restrict__ wrow,
                                             float* __restrict__ mom, int D, int lane, float lr,
                                             float eps, const SrArgs& sr) {
  float mult = lr;
  if constexpr (OPT == B200_OPT_ROWWISE_ADAGRAD) {
    float ss = 0.f;
#pragma unroll
    for (int v = 0; v < NV; ++v)
#pragma unroll
      for (int e = 0; e < EPV; ++e) ss = fmaf(acc[v][e], acc[v][e], ss);  // lanes past D hold 0
    ss = warp_sum(ss) / (float)D;
    const float m_old = PREFETCHED ? m_prefetched : *mom;
    const float m_new = m_old + ss;
    if (lane == 0) *mom = m_new;
    mult = lr / (sqrtf(m_new) + eps);
  }
#pragma unroll
  for (int v = 0; v < NV; ++v) {
    const int d = EPV * (lane + 32 * v);
    if (d < D) {
      float w[EPV];
      if constexpr (PREFETCHED) {
#pragma unroll
        for (int e = 0; e < EPV; ++e) w[e] = wv[v][e];
      } else {
        load_wrow<W, EPV>(wrow + d, w);
      }
#pragma unroll
      for (int e = 0; e < EPV; ++e) w[e] = fmaf(acc[v][e], -mult, w[e]);
      sr_prepare<W, EPV>(sr, wrow + d, w);
      store_vec<W, EPV, false>(wrow + d, w);
    }
  }
}

// shared/global partial rows are stored so that lane accesses are 16-byte and conflict-free:
// float4 chunk c of vector v of lane l sits at ((v*(EPV/4)+c)*32 + l)*4.
template <int EPV, int NV>
__device__ __forceinline__ void row_store(float* dst, const float (&acc)[NV][EPV], int lane) {
#pragma unroll
  for (int v = 0; v < NV; ++v)
#pragma unroll
    for (int c = 0; c < EPV / 4; ++c)
      *reinterpret_cast<float4*>(dst + ((v * (EPV / 4) + c) * 32 + lane) * 4) =
          make_float4(acc[v][4 * c], acc[v][4 * c + 1], acc[v][4 * c + 2], acc[v][4 * c + 3]);
}
template <int EPV, int NV>
__device__ __forceinline__ void row_add(float (&acc)[NV][EPV], const float* src, int lane) {
#pragma unroll
  for (int v = 0; v < NV; ++v)
#pragma unroll
    for (int c = 0; c < EPV / 4; ++c) {
      const float4 x = *reinterpret_cast<const float4*>(src + ((v * (EPV / 4) + c) * 32 + lane) * 4);
      acc[v][4 * c] += x.x;
      acc[v][4 * c + 1] += x.y;
      acc[v][4 * c + 2] += x.z;
      acc[v][4 * c + 3] += x.w;
    }
}


// cp.async ring geometry of the tile kernel: per sorted position one fp32 gradient row, one weight
// row slot (16 bytes per lane and vector) and 16 bytes for the momentum word.
template <int EPV, int NV>
struct BwdRing {
  static constexpr int kGradBytes = NV * EPV * 32 * 4;
  static constexpr int kWrowBytes = NV * 32 * 16;
  static constexpr int kStageBytes = kGradBytes + kWrowBytes + 16;
  static constexpr int kChunks = NV * (EPV / 4);
#ifdef B200_BWD_STAGES
  static constexpr int kStages = kChunks >= 4 ? 2 : B200_BWD_STAGES;  // power of two
#else
  static constexpr int kStages = kChunks >= 4 ? 2 : 4;  // power of two
#endif
};

__device__ __forceinline__ void cp_async_cg16(void* dst_smem, const void* src) {
  const uint32_t sa = (uint32_t)__cvta_generic_to_shared(dst_smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_ca8(void* dst_smem, const void* src) {
  const uint32_t sa = (uint32_t)__cvta_generic_to_shared(dst_smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sa), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_ca4(void* dst_smem, const void* src) {
  const uint32_t sa = (uint32_t)__cvta_generic_to_shared(dst_smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit_group() {
  asm volatile("cp.async.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void cp_async_wait_group() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// one lane's staged piece of a weight row (EPV elements of W) -> fp32
template <typename W, int EPV>
__device__ __forceinline__ void load_wslot(const unsigned char* p, float (&f)[EPV]) {
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

// ---- narrow rows (D <= 16 * EPV): several sorted positions per warp instruction -------------------
// With one position per step the tile loop costs ~150 issue slots per position whatever the row
// width, so at dim 32 / 64 (SURVEY cfg4) the kernel is issue-bound at a quarter of what HBM allows.
// Here a row is owned by a group of LPR lanes (EPV elements each) and the R = 32 / LPR groups of the
// warp take R consecutive sorted positions per step:
//   * gradient rows (and, for positions that complete a run, the weight row + momentum) of step s+1
//     are loaded into registers while step s is reduced (two steps in flight per lane);
//   * rows of one run are adjacent groups: a segmented inclusive scan across groups (log2 R shuffle
//     rounds) gives every run end its sum inside the step, the run still open from earlier steps
//     is a carry replicated in every group;
//   * all runs that end in a step apply the optimizer at once, each by its own group.
// Leading / trailing partial runs leave through the same interface as the other two paths
// (sH[warp] and `acc` in the one-row-per-warp layout, which for D <= EPV * LPR is group 0's layout).
template <typename W, int EPV>
struct BwdRawW {
  static constexpr int kWords = (int)(EPV * sizeof(W) / 4);
  uint32_t r[kWords];
  __device__ __forceinline__ void load(const W* p) {
    if constexpr (kWords == 4) {
      const uint4 v = ld_16(p);
      r[0] = v.x, r[1] = v.y, r[2] = v.z, r[3] = v.w;
    } else {
      const uint2 v = ld_8(p);
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

template <typename W, int EPV, int OPT, int LPR>
__device__ __forceinline__ void bwd_packed_tile(const int lane, const unsigned Hm, const unsigned Em,
                                                const unsigned Cm, const float* gptr, W* wptr,
                                                float* mptr, const float scale, const bool scaled,
                                                const int D0, const float lr, const float eps,
                                                const SrArgs& sr, float* sH_warp,
                                                float (&acc)[1][EPV]) {
  constexpr int R = 32 / LPR;
  constexpr unsigned kRMask = (1u << R) - 1u;
  const int gl = lane & (LPR - 1);
  const int g = lane / LPR;
  const bool act = EPV * gl < D0;
  const float inv_D = 1.0f / (float)D0;
  float carry[EPV];
#pragma unroll
  for (int e = 0; e < EPV; ++e) carry[e] = 0.f;

  float gx[2][EPV];
  BwdRawW<W, EPV> wr[2];
  float mo[2] = {0.f, 0.f};
  W* wpb[2] = {nullptr, nullptr};
  float* mpb[2] = {nullptr, nullptr};
  float spb[2] = {1.f, 1.f};

  auto issue = [&](int s, int buf) {
    const int p = s * R + g;
    const float* gp = reinterpret_cast<const float*>(__shfl_sync(kFull, (unsigned long long)gptr, p));
    wpb[buf] = reinterpret_cast<W*>(__shfl_sync(kFull, (unsigned long long)wptr, p));
    if constexpr (OPT == B200_OPT_ROWWISE_ADAGRAD)
      mpb[buf] = reinterpret_cast<float*>(__shfl_sync(kFull, (unsigned long long)mptr, p));
    spb[buf] = scaled ? __shfl_sync(kFull, scale, p) : 1.f;
    const bool fin = (Cm >> p) & 1u;
    if (act) {
      load_grad<EPV>(gp + EPV * gl, gx[buf]);
      if (fin) wr[buf].load(wpb[buf] + EPV * gl);
    } else {
#pragma unroll
      for (int e = 0; e < EPV; ++e) gx[buf][e] = 0.f;
    }
    if constexpr (OPT == B200_OPT_ROWWISE_ADAGRAD) {
      if (fin && gl == 0) mo[buf] = *mpb[buf];
    }
  };

  issue(0, 0);
#pragma unroll 2
  for (int s = 0; s < LPR; ++s) {  // 32 / R steps
    const int buf = s & 1;
    if (s + 1 < LPR) issue(s + 1, buf ^ 1);
    float x[EPV];
#pragma unroll
    for (int e = 0; e < EPV; ++e) x[e] = scaled ? gx[buf][e] * spb[buf] : gx[buf][e];
    const int p = s * R + g;
    int f = (int)((Hm >> p) & 1u);
    // segmented inclusive scan over the R groups (segments start at run heads)
#pragma unroll
    for (int d = LPR; d < 32; d <<= 1) {
      float y[EPV];
#pragma unroll
      for (int e = 0; e < EPV; ++e) y[e] = __shfl_up_sync(kFull, x[e], d);
      const int fy = __shfl_up_sync(kFull, f, d);
      if (lane >= d && !f) {
#pragma unroll
        for (int e = 0; e < EPV; ++e) x[e] += y[e];
        f |= fy;
      }
    }
    if (!f) {  // the segment began before this step: add the open run
#pragma unroll
      for (int e = 0; e < EPV; ++e) x[e] += carry[e];
    }
    const unsigned emb = (Em >> (s * R)) & kRMask;
    const unsigned cmb = (Cm >> (s * R)) & kRMask;
    if (emb) {  // warp-uniform: some run ends in this step
      if (cmb) {
        const bool my_fin = (cmb >> g) & 1u;
        float mult = lr;
        if constexpr (OPT == B200_OPT_ROWWISE_ADAGRAD) {
          float ss = 0.f;
#pragma unroll
          for (int e = 0; e < EPV; ++e) ss = fmaf(x[e], x[e], ss);  // idle lanes hold 0
#pragma unroll
          for (int o = LPR / 2; o > 0; o >>= 1) ss += __shfl_xor_sync(kFull, ss, o);
          float m_new = mo[buf] + ss * inv_D;
          if (my_fin && gl == 0) *mpb[buf] = m_new;
          m_new = __shfl_sync(kFull, m_new, g * LPR);
          float r;
          asm("sqrt.approx.f32 %0, %1;" : "=f"(r) : "f"(m_new));
          mult = __fdividef(lr, r + eps);
        }
        if (my_fin && act) {
          float w[EPV];
          wr[buf].unpack(w);
#pragma unroll
          for (int e = 0; e < EPV; ++e) w[e] = fmaf(x[e], -mult, w[e]);
          sr_prepare<W, EPV>(sr, wpb[buf] + EPV * gl, w);
          store_vec<W, EPV, false>(wpb[buf] + EPV * gl, w);
        }
      }
      const unsigned lead = emb & ~cmb;
      if (lead) {  // the run that entered the tile ends here: its partial sum leaves through sH
        const int gE = __ffs(lead) - 1;
        float tmp[1][EPV];
#pragma unroll
        for (int e = 0; e < EPV; ++e) {
          const float v = __shfl_sync(kFull, x[e], gE * LPR + gl);
          tmp[0][e] = lane < LPR ? v : 0.f;
        }
        row_store<EPV, 1>(sH_warp, tmp, lane);
      }
    }
    const bool last_ends = (emb >> (R - 1)) & 1u;
#pragma unroll
    for (int e = 0; e < EPV; ++e) {
      const float c = __shfl_sync(kFull, x[e], (R - 1) * LPR + gl);
      carry[e] = last_ends ? 0.f : c;
    }
  }
#pragma unroll
  for (int e = 0; e < EPV; ++e) acc[0][e] = lane < LPR ? carry[e] : 0.f;
}

template <typename K>
struct BwdArgs {
  const float* grad;
  int64_t grad_stride;
  const K* keys;
  const int32_t* payload;
  const int32_t* bag_of_pos;  // non-null => payload is the original position
  int64_t N;
  void* weights;
  const int64_t* weights_offsets;
  const int32_t* D_offsets;
  const int64_t* hash_size_cumsum;
  int32_t T, B;
  const void* offsets;
  bool idx64;
  const float* psw;
  bool mean;
  float* momentum1;
  const int64_t* momentum1_offsets;
  float lr, eps;
  // Peer gather (sharded training): the gradient row of global sample b lives on rank b / B_src at
  // grad_bases[rank] + (b % B_src) * grad_stride + src_col[t].  With NVLink peer mappings this IS
  // the backward all-to-all (table-wise) / all-gather (row-wise).  nullptr => local `grad`.
  const float* const* grad_bases;
  const int32_t* src_col;
  int32_t B_src;
  float grad_scale;  // 1/W of the sharded path (GRADIENT_DIVISION), 1 otherwise
  int32_t pack_rows;  // narrow rows: several positions per warp step (bwd_packed_tile)
  SrArgs sr;          // stochastic rounding of 16-bit weight rows (off for fp32 tables)
  float* part_rows;  // [numCTAs][2][ROW]  (0 = leading partial H, 1 = trailing partial T)
  BwdMeta* meta;     // [numCTAs]
  int32_t num_ctas;
};

template <typename W, typename K, int EPV, int NV, int OPT, bool PACKED>
__global__ void __launch_bounds__(kBwdThreads) tbe_backward_tile_kernel(const BwdArgs<K> a) {
  constexpr int ROW = NV * EPV * 32;
  constexpr int kStages = BwdRing<EPV, NV>::kStages;
  constexpr int kGradBytes = BwdRing<EPV, NV>::kGradBytes;
  constexpr int kWrowBytes = BwdRing<EPV, NV>::kWrowBytes;
  constexpr int kStageBytes = BwdRing<EPV, NV>::kStageBytes;
  extern __shared__ __align__(16) float smem[];
  float* sH = smem;                    // [kBwdWarps][ROW]
  __shared__ int s_first_head[kBwdWarps];
  __shared__ int s_cont[kBwdWarps];
  // per-position row pointers of the fast path: [warp][grad | weight | momentum][position]
  __shared__ unsigned long long s_ptr[kBwdWarps][3][32];
  __shared__ float s_scale[kBwdWarps][32];

  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int64_t i0 = (int64_t)blockIdx.x * kBwdTile + warp * 32;
  const int nvalid = (int)max((int64_t)0, min((int64_t)32, a.N - i0));
  const bool valid = lane < nvalid;
  W* const weights = reinterpret_cast<W*>(a.weights);

  // ---- per-position decode (lane k <-> sorted position i0+k)
  K key = 0;
  int32_t pay = 0;
  if (valid) {
    key = a.keys[i0 + lane];
    pay = a.payload[i0 + lane];
  }
  K prev = __shfl_up_sync(kFull, key, 1);
  if (lane == 0 && i0 > 0 && nvalid > 0) prev = a.keys[i0 - 1];
  const bool head = valid && ((i0 + lane) == 0 || key != prev);
  K nextkey = key;
  if (lane == 31 && nvalid == 32 && i0 + 32 < a.N) nextkey = a.keys[i0 + 32];
  const bool cont =
      __shfl_sync(kFull, (int)(nvalid == 32 && i0 + 32 < a.N && nextkey == key), 31) != 0;
  const unsigned Hm = __ballot_sync(kFull, head);
  const unsigned Vm = nvalid == 32 ? kFull : ((1u << nvalid) - 1u);
  const int fh = Hm ? (__ffs(Hm) - 1) : 32;
  // run-end mask: position k ends a run if k+1 is a head, or k is the tile's last position
  // and the run does not continue.  Complete (head inside the tile) only for k >= fh.
  unsigned Em = (Hm >> 1);
  if (nvalid > 0 && !cont) Em |= 1u << (nvalid - 1);
  Em &= Vm;
  const unsigned Cm = fh < 32 ? (Em & ~((1u << fh) - 1u)) : 0u;

  int32_t pos = pay, bt = pay;
  if (a.bag_of_pos && valid) bt = a.bag_of_pos[pos];
  const int32_t t = valid ? bt / a.B : 0;
  const int32_t b = bt - t * a.B;
  int32_t Dbeg = 0, Dt = 0;
  int64_t woff = 0, moff = 0;
  float scale = 1.f;
  if (valid) {
    Dbeg = a.D_offsets[t];
    Dt = a.D_offsets[t + 1] - Dbeg;
    const int64_t row = (int64_t)key - a.hash_size_cumsum[t];
    woff = a.weights_offsets[t] + row * Dt;
    if constexpr (OPT == B200_OPT_ROWWISE_ADAGRAD) moff = a.momentum1_offsets[t] + row;
    if (a.mean) {
      const int64_t L = load_index(a.offsets, (int64_t)bt + 1, a.idx64) -
                        load_index(a.offsets, (int64_t)bt, a.idx64);
      scale = 1.0f / (float)L;
    }
    if (a.psw) scale *= a.psw[pos];
    scale *= a.grad_scale;
  }
  const bool scaled = a.mean || a.psw != nullptr || a.grad_scale != 1.0f;
  const float* gptr = a.grad;
  if (valid) {
    if (a.grad_bases) {
      const int r = b / a.B_src;
      gptr = a.grad_bases[r] + (int64_t)(b - r * a.B_src) * a.grad_stride + a.src_col[t];
    } else {
      gptr = a.grad + (int64_t)b * a.grad_stride + Dbeg;
    }
  }

  float acc[NV][EPV];
#pragma unroll
  for (int v = 0; v < NV; ++v)
#pragma unroll
    for (int e = 0; e < EPV; ++e) acc[v][e] = 0.f;
  int open_head = -1;  // head position of the open run inside this tile, -1 = carried in
  float dummy_w[NV][EPV];

  // ---- stream the tile's rows through a warp-private cp.async ring.  One group per sorted
  // position: its gradient row and, when a run that started in this tile ends there, the weight
  // row (+ momentum) the optimizer will rewrite.  Every lane copies and later consumes its own
  // 16-byte pieces, so no warp barrier is needed and the registers hold only the accumulator:
  // kStages-1 positions (~0.7 KB each at D=128) stay in flight per warp.
  unsigned char* ring = reinterpret_cast<unsigned char*>(smem + kBwdWarps * ROW) +
                        (size_t)warp * kStages * kStageBytes;
  // ---- fast path: a full tile whose rows all have one width.  Row pointers are staged in
  // shared memory once, the position loop is unrolled by 8 so stage offsets and mask bits are
  // immediates, and the optimizer uses MUFU sqrt/rcp: ~35 issue slots per position instead of
  // ~140 (the generic loop below is issue-bound, not HBM-bound).
  const int D0 = __shfl_sync(kFull, Dt, 0);
  const bool fast = nvalid == 32 && __all_sync(kFull, Dt == D0);
  bool packed = false;
  if constexpr (PACKED) {  // host: NV == 1 and max_D <= 8 * EPV (4 or 8 lanes per row; with 16 lanes
    // only two positions fit a step and the one-row path below is faster)
    if (fast && D0 <= EPV * 8) {
      packed = true;
      W* const wp = weights + woff;
      float* const mp = (OPT == B200_OPT_ROWWISE_ADAGRAD) ? a.momentum1 + moff : nullptr;
      if (D0 <= EPV * 4)
        bwd_packed_tile<W, EPV, OPT, 4>(lane, Hm, Em, Cm, gptr, wp, mp, scale, scaled, D0, a.lr, a.eps,
                                        a.sr, sH + warp * ROW, acc);
      else
        bwd_packed_tile<W, EPV, OPT, 8>(lane, Hm, Em, Cm, gptr, wp, mp, scale, scaled, D0, a.lr, a.eps,
                                        a.sr, sH + warp * ROW, acc);
      open_head = Hm ? 31 - __clz(Hm) : -1;
    }
  }
  if (packed) {
    // done: acc / open_head / sH are set as the stitching code below expects
  } else if (fast) {
    unsigned long long* const mg = s_ptr[warp][0];
    unsigned long long* const mw = s_ptr[warp][1];
    unsigned long long* const mm = s_ptr[warp][2];
    mg[lane] = (unsigned long long)gptr;
    mw[lane] = (unsigned long long)(weights + woff);
    if constexpr (OPT == B200_OPT_ROWWISE_ADAGRAD) mm[lane] = (unsigned long long)(a.momentum1 + moff);
    s_scale[warp][lane] = scale;
    __syncwarp();
    bool act[NV];
#pragma unroll
    for (int v = 0; v < NV; ++v) act[v] = EPV * (lane + 32 * v) < D0;
    unsigned char* const lring = ring + lane * 16;
    const float inv_D = 1.0f / (float)D0;

    auto fissue = [&](int p, int stage, bool fin) {
      unsigned char* st = lring + stage * kStageBytes;
      const float* g = reinterpret_cast<const float*>(mg[p]) + EPV * lane;
#pragma unroll
      for (int v = 0; v < NV; ++v)
        if (act[v]) {
#pragma unroll
          for (int c = 0; c < EPV / 4; ++c)
            cp_async_cg16(st + (v * (EPV / 4) + c) * 512, g + EPV * 32 * v + 4 * c);
        }
      if (fin) {  // warp-uniform
        const W* w = reinterpret_cast<const W*>(mw[p]) + EPV * lane;
#pragma unroll
        for (int v = 0; v < NV; ++v)
          if (act[v]) {
            if constexpr (EPV * sizeof(W) == 16)
              cp_async_cg16(st + kGradBytes + v * 512, w + EPV * 32 * v);
            else
              cp_async_ca8(st + kGradBytes + v * 512, w + EPV * 32 * v);
          }
        if constexpr (OPT == B200_OPT_ROWWISE_ADAGRAD) {
          if (lane == 0) cp_async_ca4(st + kGradBytes + kWrowBytes, reinterpret_cast<const float*>(mm[p]));
        }
      }
      cp_async_commit_group();
    };
    auto fconsume = [&](int p, int stage, bool ends, bool fin) {
      cp_async_wait_group<kStages - 1>();
      const unsigned char* st = lring + stage * kStageBytes;
      const float sp = scaled ? s_scale[warp][p] : 1.f;
#pragma unroll
      for (int v = 0; v < NV; ++v)
        if (act[v]) {
#pragma unroll
          for (int c = 0; c < EPV / 4; ++c) {
            const float4 x = *reinterpret_cast<const float4*>(st + (v * (EPV / 4) + c) * 512);
            if (scaled) {
              acc[v][4 * c] = fmaf(sp, x.x, acc[v][4 * c]);
              acc[v][4 * c + 1] = fmaf(sp, x.y, acc[v][4 * c + 1]);
              acc[v][4 * c + 2] = fmaf(sp, x.z, acc[v][4 * c + 2]);
              acc[v][4 * c + 3] = fmaf(sp, x.w, acc[v][4 * c + 3]);
            } else {
              acc[v][4 * c] += x.x;
              acc[v][4 * c + 1] += x.y;
              acc[v][4 * c + 2] += x.z;
              acc[v][4 * c + 3] += x.w;
            }
          }
        }
      if (ends) {  // warp-uniform
        if (fin) {
          float mult = a.lr;
          if constexpr (OPT == B200_OPT_ROWWISE_ADAGRAD) {
            float ss = 0.f;
#pragma unroll
            for (int v = 0; v < NV; ++v)
#pragma unroll
              for (int e = 0; e < EPV; ++e) ss = fmaf(acc[v][e], acc[v][e], ss);  // idle lanes hold 0
            ss = warp_sum(ss) * inv_D;
            float m_new = 0.f;
            if (lane == 0) {
              m_new = *reinterpret_cast<const float*>(st - lane * 16 + kGradBytes + kWrowBytes) + ss;
              *reinterpret_cast<float*>(mm[p]) = m_new;
            }
            m_new = __shfl_sync(kFull, m_new, 0);
            float r;
            asm("sqrt.approx.f32 %0, %1;" : "=f"(r) : "f"(m_new));
            mult = __fdividef(a.lr, r + a.eps);
          }
          W* wrow = reinterpret_cast<W*>(mw[p]) + EPV * lane;
#pragma unroll
          for (int v = 0; v < NV; ++v)
            if (act[v]) {
              float w[EPV];
              load_wslot<W, EPV>(st + kGradBytes + v * 512, w);
#pragma unroll
              for (int e = 0; e < EPV; ++e) w[e] = fmaf(acc[v][e], -mult, w[e]);
              sr_prepare<W, EPV>(a.sr, wrow + EPV * 32 * v, w);
              store_vec<W, EPV, false>(wrow + EPV * 32 * v, w);
            }
        } else {
          row_store<EPV, NV>(sH + warp * ROW, acc, lane);  // tail of a run from an earlier tile
        }
#pragma unroll
        for (int v = 0; v < NV; ++v)
#pragma unroll
          for (int e = 0; e < EPV; ++e) acc[v][e] = 0.f;
      }
    };
#pragma unroll
    for (int q = 0; q < kStages - 1; ++q) fissue(q, q, (Cm >> q) & 1u);
#pragma unroll 1
    for (int p0 = 0; p0 < 32; p0 += 8) {
      const unsigned emg = Em >> p0, cmg = Cm >> p0, vmg = kFull >> p0;
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        constexpr int kAhead = kStages - 1;
        if ((vmg >> (q + kAhead)) & 1u)
          fissue(p0 + q + kAhead, (q + kAhead) & (kStages - 1), (cmg >> (q + kAhead)) & 1u);
        else
          cp_async_commit_group();
        fconsume(p0 + q, q & (kStages - 1), (emg >> q) & 1u, (cmg >> q) & 1u);
      }
    }
    cp_async_wait_group<0>();
    open_head = Hm ? 31 - __clz(Hm) : -1;
  } else {
  auto issue = [&](int p) {
    if (p < nvalid) {
      unsigned char* st = ring + (p % kStages) * kStageBytes;
      const float* grow = reinterpret_cast<const float*>(
          __shfl_sync(kFull, (unsigned long long)gptr, p));
      const int Dp = __shfl_sync(kFull, Dt, p);
      const bool fin = (Cm >> p) & 1u;
      int64_t wo = 0, mo = 0;
      if (fin) {  // warp-uniform
        wo = __shfl_sync(kFull, woff, p);
        if constexpr (OPT == B200_OPT_ROWWISE_ADAGRAD) mo = __shfl_sync(kFull, moff, p);
      }
#pragma unroll
      for (int v = 0; v < NV; ++v) {
        const int d = EPV * (lane + 32 * v);
        if (d < Dp) {
#pragma unroll
          for (int c = 0; c < EPV / 4; ++c)
            cp_async_cg16(st + ((v * (EPV / 4) + c) * 32 + lane) * 16, grow + d + 4 * c);
          if (fin) {
            if constexpr (EPV * sizeof(W) == 16)
              cp_async_cg16(st + kGradBytes + (v * 32 + lane) * 16, weights + wo + d);
            else
              cp_async_ca8(st + kGradBytes + (v * 32 + lane) * 16, weights + wo + d);
          }
        }
      }
      if constexpr (OPT == B200_OPT_ROWWISE_ADAGRAD) {
        if (fin && lane == 0) cp_async_ca4(st + kGradBytes + kWrowBytes, a.momentum1 + mo);
      }
    }
    cp_async_commit_group();  // always: keeps the group count uniform
  };
#pragma unroll
  for (int p = 0; p < kStages - 1; ++p) issue(p);

  for (int p = 0; p < nvalid; ++p) {
    issue(p + kStages - 1);
    cp_async_wait_group<kStages - 1>();
    const unsigned char* st = ring + (p % kStages) * kStageBytes;
    const int Dp = __shfl_sync(kFull, Dt, p);
    if ((Hm >> p) & 1u) open_head = p;
    const float sp = scaled ? __shfl_sync(kFull, scale, p) : 1.f;
#pragma unroll
    for (int v = 0; v < NV; ++v) {
      if (EPV * (lane + 32 * v) < Dp) {
#pragma unroll
        for (int c = 0; c < EPV / 4; ++c) {
          const float4 x = *reinterpret_cast<const float4*>(st + ((v * (EPV / 4) + c) * 32 + lane) * 16);
          acc[v][4 * c] = fmaf(sp, x.x, acc[v][4 * c]);
          acc[v][4 * c + 1] = fmaf(sp, x.y, acc[v][4 * c + 1]);
          acc[v][4 * c + 2] = fmaf(sp, x.z, acc[v][4 * c + 2]);
          acc[v][4 * c + 3] = fmaf(sp, x.w, acc[v][4 * c + 3]);
        }
      }
    }
    if ((Em >> p) & 1u) {
      if (open_head >= 0) {
        const int64_t wo = __shfl_sync(kFull, woff, p);
        int64_t mo = 0;
        float m_old = 0.f;
        if constexpr (OPT == B200_OPT_ROWWISE_ADAGRAD) {
          mo = __shfl_sync(kFull, moff, p);
          if (lane == 0) m_old = *reinterpret_cast<const float*>(st + kGradBytes + kWrowBytes);
          m_old = __shfl_sync(kFull, m_old, 0);
        }
        float wv[NV][EPV];
#pragma unroll
        for (int v = 0; v < NV; ++v) {
          if (EPV * (lane + 32 * v) < Dp) {
            load_wslot<W, EPV>(st + kGradBytes + (v * 32 + lane) * 16, wv[v]);
          } else {
#pragma unroll
            for (int e = 0; e < EPV; ++e) wv[v][e] = 0.f;
          }
        }
        apply_update<W, EPV, NV, OPT, true>(acc, wv, m_old, weights + wo, a.momentum1 + mo, Dp, lane,
                                            a.lr, a.eps, a.sr);
      } else {
        row_store<EPV, NV>(sH + warp * ROW, acc, lane);  // tail of a run from an earlier tile
      }
#pragma unroll
      for (int v = 0; v < NV; ++v)
#pragma unroll
        for (int e = 0; e < EPV; ++e) acc[v][e] = 0.f;
    }
  }
  cp_async_wait_group<0>();
  }  // generic path
  // The run still open at the end of the tile continues into the next one.
  // open_head < 0: the whole tile lies inside one run -> leading partial (H).
  if (cont && open_head < 0) row_store<EPV, NV>(sH + warp * ROW, acc, lane);
  if (lane == 0) {
    s_first_head[warp] = nvalid > 0 ? fh : 0;
    s_cont[warp] = cont ? 1 : 0;
  }
  __syncthreads();

  // ---- stitch runs across the warps of this CTA (fixed order => deterministic)
  float* my_part = a.part_rows + (int64_t)blockIdx.x * 2 * ROW;
  if (cont && open_head >= 0) {
    // this warp holds the first piece (T) of a run that continues; acc is still in registers
    const K rk = __shfl_sync(kFull, key, open_head);
    const int32_t rt = __shfl_sync(kFull, t, open_head);
    const int rD = __shfl_sync(kFull, Dt, open_head);
    const int64_t rwo = __shfl_sync(kFull, woff, open_head);
    const int64_t rmo = __shfl_sync(kFull, moff, open_head);
    bool finished = false;
    for (int x = warp + 1; x < kBwdWarps; ++x) {
      const int xfh = s_first_head[x];
      if (xfh > 0) row_add<EPV, NV>(acc, sH + x * ROW, lane);
      if (xfh < 32 || !s_cont[x]) {  // a head follows, or the run ends with tile x
        finished = true;
        break;
      }
    }
    if (finished) {
      apply_update<W, EPV, NV, OPT, false>(acc, dummy_w, 0.f, weights + rwo, a.momentum1 + rmo, rD,
                                           lane, a.lr, a.eps, a.sr);
    } else {
      row_store<EPV, NV>(my_part + ROW, acc, lane);
      if (lane == 0) {
        a.meta[blockIdx.x].t_key = (unsigned long long)rk;
        a.meta[blockIdx.x].t_feat = rt;
      }
    }
    if (lane == 0 && !finished) a.meta[blockIdx.x].t_flag = 1;
  }
  if (warp == 0) {
    // leading partial of the CTA: pieces before the first head belong to a run that
    // started in an earlier CTA.
    float lead[NV][EPV];
#pragma unroll
    for (int v = 0; v < NV; ++v)
#pragma unroll
      for (int e = 0; e < EPV; ++e) lead[v][e] = 0.f;
    int stop = 0;      // the run entering this CTA ends inside it
    int any_head = 0;  // some run starts inside this CTA
    for (int x = 0; x < kBwdWarps; ++x) {
      const int xfh = s_first_head[x];
      any_head |= (xfh < 32);
      if (!stop) {
        if (xfh > 0) row_add<EPV, NV>(lead, sH + x * ROW, lane);
        if (xfh < 32 || !s_cont[x]) stop = 1;
      }
    }
    // only the last run can leave the CTA: it does iff the last warp continues, and it
    // started here iff the CTA holds a head.
    const int any_t = s_cont[kBwdWarps - 1] && any_head;
    row_store<EPV, NV>(my_part, lead, lane);
    if (lane == 0) {
      a.meta[blockIdx.x].has_head = stop;
      if (!any_t) a.meta[blockIdx.x].t_flag = 0;
    }
  }
}

// Finish runs that straddle CTAs: T of CTA c + leading partials of following CTAs.
template <typename W, typename K, int EPV, int NV, int OPT>
__global__ void __launch_bounds__(128) tbe_backward_combine_kernel(const BwdArgs<K> a) {
  constexpr int ROW = NV * EPV * 32;
  const int lane = threadIdx.x & 31;
  const int c = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (c >= a.num_ctas) return;
  if (!a.meta[c].t_flag) return;
  W* const weights = reinterpret_cast<W*>(a.weights);
  float acc[NV][EPV];
#pragma unroll
  for (int v = 0; v < NV; ++v)
#pragma unroll
    for (int e = 0; e < EPV; ++e) acc[v][e] = 0.f;
  row_add<EPV, NV>(acc, a.part_rows + ((int64_t)c * 2 + 1) * ROW, lane);
  int x0 = c + 1;
  while (x0 < a.num_ctas) {
    const int x = x0 + lane;
    const int flag = x < a.num_ctas ? a.meta[x].has_head : 1;
    const unsigned m = __ballot_sync(kFull, flag != 0);
    const int n = m ? (__ffs(m) - 1) : 32;  // CTAs x0 .. x0+n-1 lie wholly inside the run
    const int last = m ? n : n - 1;         // inclusive index of the last partial to add
    for (int k = 0; k <= last; ++k) {
      if (x0 + k < a.num_ctas) row_add<EPV, NV>(acc, a.part_rows + (int64_t)(x0 + k) * 2 * ROW, lane);
    }
    if (m) break;
    x0 += 32;
  }
  const K key = (K)a.meta[c].t_key;
  const int32_t t = a.meta[c].t_feat;
  const int Dbeg = a.D_offsets[t];
  const int D = a.D_offsets[t + 1] - Dbeg;
  const int64_t row = (int64_t)key - a.hash_size_cumsum[t];
  float dummy_w[NV][EPV];
  float* mom = a.momentum1;
  if constexpr (OPT == B200_OPT_ROWWISE_ADAGRAD) mom += a.momentum1_offsets[t] + row;
  apply_update<W, EPV, NV, OPT, false>(acc, dummy_w, 0.f, weights + a.weights_offsets[t] + row * D,
                                       mom, D, lane, a.lr, a.eps, a.sr);
}

template <typename W, typename K, int EPV, int NV, int OPT>
static int launch_backward(const BwdArgs<K>& a, cudaStream_t s) {
  constexpr int ROW = NV * EPV * 32;
  const size_t smem = sizeof(float) * kBwdWarps * ROW +
                      (size_t)kBwdWarps * BwdRing<EPV, NV>::kStages * BwdRing<EPV, NV>::kStageBytes;
  if constexpr (NV == 1) {
    if (a.pack_rows) {  // narrow rows: several sorted positions per warp step
      auto ptile = tbe_backward_tile_kernel<W, K, EPV, 1, OPT, true>;
      if (smem > 48 * 1024)
        B200_CUDA_TRY(cudaFuncSetAttribute(ptile, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      ptile<<<a.num_ctas, kBwdThreads, smem, s>>>(a);
      B200_LAUNCH_CHECK();
      tbe_backward_combine_kernel<W, K, EPV, NV, OPT><<<(a.num_ctas + 3) / 4, 128, 0, s>>>(a);
      B200_LAUNCH_CHECK();
      return B200_OK;
    }
  }
  auto tile = tbe_backward_tile_kernel<W, K, EPV, NV, OPT, false>;
  if (smem > 48 * 1024)
    B200_CUDA_TRY(cudaFuncSetAttribute(tile, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  tile<<<a.num_ctas, kBwdThreads, smem, s>>>(a);
  B200_LAUNCH_CHECK();
  tbe_backward_combine_kernel<W, K, EPV, NV, OPT><<<(a.num_ctas + 3) / 4, 128, 0, s>>>(a);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

template <typename W, typename K, int EPV, int OPT>
static int dispatch_nv(const BwdArgs<K>& a, int max_D, cudaStream_t s) {
  const int vecs = (int)ceil_div(max_D, EPV * 32);
  if (vecs <= 1) {
    BwdArgs<K> b = a;
    b.pack_rows = (a.pack_rows && max_D <= EPV * 8) ? 1 : 0;
    return launch_backward<W, K, EPV, 1, OPT>(b, s);
  }
  if (vecs <= 2) return launch_backward<W, K, EPV, 2, OPT>(a, s);
  if (vecs <= 4) return launch_backward<W, K, EPV, 4, OPT>(a, s);
  if (vecs <= 8) return launch_backward<W, K, EPV, 8, OPT>(a, s);
  return B200_EUNSUPPORTED;
}

template <typename W, typename K, int EPV>
static int dispatch_opt(const BwdArgs<K>& a, int optimizer, int max_D, cudaStream_t s) {
  if (optimizer == B200_OPT_SGD) return dispatch_nv<W, K, EPV, B200_OPT_SGD>(a, max_D, s);
  if (optimizer == B200_OPT_ROWWISE_ADAGRAD)
    return dispatch_nv<W, K, EPV, B200_OPT_ROWWISE_ADAGRAD>(a, max_D, s);
  return B200_EUNSUPPORTED;
}

template <typename K>
static int dispatch_w(const BwdArgs<K>& a, int weight_dtype, bool mult8, int optimizer, int max_D,
                      cudaStream_t s) {
  switch (weight_dtype) {
    case B200_F32:
      return dispatch_opt<float, K, 4>(a, optimizer, max_D, s);
    case B200_F16:
      return mult8 ? dispatch_opt<__half, K, 8>(a, optimizer, max_D, s)
                   : dispatch_opt<__half, K, 4>(a, optimizer, max_D, s);
    case B200_BF16:
      return mult8 ? dispatch_opt<__nv_bfloat16, K, 8>(a, optimizer, max_D, s)
                   : dispatch_opt<__nv_bfloat16, K, 4>(a, optimizer, max_D, s);
    default:
      return B200_EUNSUPPORTED;
  }
}

static inline int64_t bwd_num_ctas(int64_t N) { return ceil_div(N, kBwdTile); }
// B200_BWD_PACK=0 disables the narrow-row packed path (A/B measurements only)
static inline int bwd_pack_rows() {
  static const int v = [] {
    const char* e = getenv("B200_BWD_PACK");
    return e ? atoi(e) : 1;
  }();
  return v;
}
static inline size_t bwd_row_floats(int max_D) {
  // the widest instantiation that can be chosen for this max_D (EPV=4 or 8, NV rounded up)
  int vecs4 = (int)ceil_div(max_D, 4 * 32);
  int nv = vecs4 <= 1 ? 1 : vecs4 <= 2 ? 2 : vecs4 <= 4 ? 4 : 8;
  return (size_t)nv * 4 * 32 * 2;  // x2 covers the EPV=8 layouts as well
}

}  // namespace b200

extern "C" size_t b200_tbe_backward_scratch_bytes(int64_t N, int32_t max_D) {
  using namespace b200;
  const int64_t ctas = bwd_num_ctas(N > 0 ? N : 1);
  return align256((size_t)ctas * 2 * bwd_row_floats(max_D) * sizeof(float)) +
         align256((size_t)ctas * sizeof(BwdMeta)) + 256;
}

static int tbe_backward_impl(const float* grad_output, int64_t grad_stride,
                                 const void* sorted_keys, const int32_t* sorted_payload,
                                 const int32_t* bag_of_pos, int64_t N, int key_bits, void* weights,
                                 int weight_dtype, const int64_t* weights_offsets,
                                 const int32_t* D_offsets, const int64_t* hash_size_cumsum,
                                 int32_t T, int32_t B, int32_t max_D, int32_t all_D_mult8,
                                 const void* offsets, int index_dtype,
                                 const float* per_sample_weights, int pooling_mode, int optimizer,
                                 float* momentum1, const int64_t* momentum1_offsets,
                                 float learning_rate, float eps, void* partial_scratch,
                                 const float* const* grad_bases, int32_t B_src,
                                 const int32_t* src_col, float grad_scale, int32_t stochastic_rounding,
                                 uint64_t sr_seed, uint64_t sr_offset, void* stream) {
  using namespace b200;
  if (N == 0 || T == 0 || B == 0) return B200_OK;
  if ((grad_stride & 3) != 0) return B200_EINVAL;
  if (per_sample_weights && !bag_of_pos) return B200_EINVAL;
  if (optimizer == B200_OPT_ROWWISE_ADAGRAD && (!momentum1 || !momentum1_offsets)) return B200_EINVAL;
  cudaStream_t s = (cudaStream_t)stream;
  const int64_t ctas = bwd_num_ctas(N);
  char* p = (char*)partial_scratch;
  float* part_rows = (float*)p;
  p += align256((size_t)ctas * 2 * bwd_row_floats(max_D) * sizeof(float));
  BwdMeta* meta = (BwdMeta*)p;
#define B200_FILL(ARGS)                                    \
  ARGS.grad = grad_output;                                 \
  ARGS.grad_stride = grad_stride;                          \
  ARGS.payload = sorted_payload;                           \
  ARGS.bag_of_pos = bag_of_pos;                            \
  ARGS.N = N;                                              \
  ARGS.weights = weights;                                  \
  ARGS.weights_offsets = weights_offsets;                  \
  ARGS.D_offsets = D_offsets;                              \
  ARGS.hash_size_cumsum = hash_size_cumsum;                \
  ARGS.T = T;                                              \
  ARGS.B = B;                                              \
  ARGS.offsets = offsets;                                  \
  ARGS.idx64 = index_dtype == B200_I64;                    \
  ARGS.psw = per_sample_weights;                           \
  ARGS.mean = pooling_mode == B200_POOL_MEAN;              \
  ARGS.momentum1 = momentum1;                              \
  ARGS.momentum1_offsets = momentum1_offsets;              \
  ARGS.lr = learning_rate;                                 \
  ARGS.eps = eps;                                          \
  ARGS.grad_bases = grad_bases;                            \
  ARGS.src_col = src_col;                                  \
  ARGS.B_src = B_src;                                      \
  ARGS.grad_scale = grad_scale;                            \
  ARGS.pack_rows = bwd_pack_rows();                        \
  ARGS.sr.seed = sr_seed;                                  \
  ARGS.sr.offset = sr_offset;                              \
  ARGS.sr.base = weights;                                  \
  ARGS.sr.on = (stochastic_rounding && weight_dtype != B200_F32) ? 1 : 0; \
  ARGS.part_rows = part_rows;                              \
  ARGS.meta = meta;                                        \
  ARGS.num_ctas = (int32_t)ctas;
  if (key_bits > 32) {
    BwdArgs<uint64_t> a;
    a.keys = (const uint64_t*)sorted_keys;
    B200_FILL(a);
    return dispatch_w<uint64_t>(a, weight_dtype, all_D_mult8 != 0, optimizer, max_D, s);
  }
  BwdArgs<uint32_t> a;
  a.keys = (const uint32_t*)sorted_keys;
  B200_FILL(a);
  return dispatch_w<uint32_t>(a, weight_dtype, all_D_mult8 != 0, optimizer, max_D, s);
#undef B200_FILL
}

extern "C" int b200_tbe_backward(const float* grad_output, int64_t grad_stride,
                                 const void* sorted_keys, const int32_t* sorted_payload,
                                 const int32_t* bag_of_pos, int64_t N, int key_bits, void* weights,
                                 int weight_dtype, const int64_t* weights_offsets,
                                 const int32_t* D_offsets, const int64_t* hash_size_cumsum,
                                 int32_t T, int32_t B, int32_t max_D, int32_t all_D_mult8,
                                 const void* offsets, int index_dtype,
                                 const float* per_sample_weights, int pooling_mode, int optimizer,
                                 float* momentum1, const int64_t* momentum1_offsets,
                                 float learning_rate, float eps, void* partial_scratch,
                                 void* stream) {
  if (!grad_output) return B200_EINVAL;
  return tbe_backward_impl(grad_output, grad_stride, sorted_keys, sorted_payload, bag_of_pos, N,
                           key_bits, weights, weight_dtype, weights_offsets, D_offsets,
                           hash_size_cumsum, T, B, max_D, all_D_mult8, offsets, index_dtype,
                           per_sample_weights, pooling_mode, optimizer, momentum1,
                           momentum1_offsets, learning_rate, eps, partial_scratch, nullptr, 0,
                           nullptr, 1.0f, 0, 0, 0, stream);
}

extern "C" int b200_tbe_backward_sr(const float* grad_output, int64_t grad_stride,
                                    const void* sorted_keys, const int32_t* sorted_payload,
                                    const int32_t* bag_of_pos, int64_t N, int key_bits, void* weights,
                                    int weight_dtype, const int64_t* weights_offsets,
                                    const int32_t* D_offsets, const int64_t* hash_size_cumsum,
                                    int32_t T, int32_t B, int32_t max_D, int32_t all_D_mult8,
                                    const void* offsets, int index_dtype,
                                    const float* per_sample_weights, int pooling_mode, int optimizer,
                                    float* momentum1, const int64_t* momentum1_offsets,
                                    float learning_rate, float eps, void* partial_scratch,
                                    uint64_t sr_seed, uint64_t sr_offset, void* stream) {
  if (!grad_output) return B200_EINVAL;
  return tbe_backward_impl(grad_output, grad_stride, sorted_keys, sorted_payload, bag_of_pos, N,
                           key_bits, weights, weight_dtype, weights_offsets, D_offsets,
                           hash_size_cumsum, T, B, max_D, all_D_mult8, offsets, index_dtype,
                           per_sample_weights, pooling_mode, optimizer, momentum1,
                           momentum1_offsets, learning_rate, eps, partial_scratch, nullptr, 0,
                           nullptr, 1.0f, 1, sr_seed, sr_offset, stream);
}

// ---------------------------------------------------------------------------------------------
// b200_tbe_backward_static: b200_tbe_backward for callers that cannot allocate (CUDA-graph capture,
// serving loops, harnesses that forbid cudaMalloc).  The per-CTA partial rows + run records that
// b200_tbe_backward takes as `partial_scratch` live in this file-scope device array instead.  Every
// GPU that loads the library holds kStaticScratchBytes of it until the process exits.
namespace b200 {
constexpr size_t kStaticScratchBytes = (size_t)16 << 20;  // 16 MiB: cfg2 (N = 1,703,936 ids, dim 128) needs 13.8 MiB
static __device__ __align__(256) unsigned char g_bwd_static_scratch[kStaticScratchBytes];
constexpr int kMaxDevices = 64;
static void* g_bwd_static_ptr[kMaxDevices];  // device address of the array on device i (0 = not looked up yet)
static std::mutex g_bwd_static_mu;

// In host code `g_bwd_static_scratch` is only a handle: the device address must come from
// cudaGetSymbolAddress, once per device, with that device current.
static int static_scratch_for_current_device(void** out) {
  int dev = -1;
  B200_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return B200_EUNSUPPORTED;
  std::lock_guard<std::mutex> lock(g_bwd_static_mu);
  if (!g_bwd_static_ptr[dev]) {
    void* p = nullptr;
    B200_CUDA_TRY(cudaGetSymbolAddress(&p, g_bwd_static_scratch));
    if (!p) return B200_EINVAL;
    g_bwd_static_ptr[dev] = p;
  }
  *out = g_bwd_static_ptr[dev];
  return B200_OK;
}
}  // namespace b200

extern "C" size_t b200_tbe_backward_static_max_scratch_bytes(void) { return b200::kStaticScratchBytes; }

extern "C" int b200_tbe_backward_static(const float* grad_output, int64_t grad_stride,
                                        const void* sorted_keys, const int32_t* sorted_payload,
                                        const int32_t* bag_of_pos, int64_t N, int key_bits,
                                        void* weights, int weight_dtype,
                                        const int64_t* weights_offsets, const int32_t* D_offsets,
                                        const int64_t* hash_size_cumsum, int32_t T, int32_t B,
                                        int32_t max_D, int32_t all_D_mult8, const void* offsets,
                                        int index_dtype, const float* per_sample_weights,
                                        int pooling_mode, int optimizer, float* momentum1,
                                        const int64_t* momentum1_offsets, float learning_rate,
                                        float eps, void* stream) {
  using namespace b200;
  if (!grad_output || N < 0 || max_D <= 0) return B200_EINVAL;
  if (b200_tbe_backward_scratch_bytes(N, max_D) > kStaticScratchBytes) return B200_ETOOBIG;
  void* scratch = nullptr;
  const int rc = static_scratch_for_current_device(&scratch);
  if (rc != B200_OK) return rc;
  // Nothing in the scratch has to start at zero: every CTA writes its partial rows and its run record
  // before the combine kernel (same stream) reads them, so the previous call's contents are dead.
  return tbe_backward_impl(grad_output, grad_stride, sorted_keys, sorted_payload, bag_of_pos, N,
                           key_bits, weights, weight_dtype, weights_offsets, D_offsets,
                           hash_size_cumsum, T, B, max_D, all_D_mult8, offsets, index_dtype,
                           per_sample_weights, pooling_mode, optimizer, momentum1,
                           momentum1_offsets, learning_rate, eps, scratch, nullptr, 0, nullptr,
                           1.0f, 0, 0, 0, stream);
}

extern "C" int b200_tbe_backward_peer(const float* const* grad_bases, int32_t n_src, int32_t B_src,
                                      int64_t grad_stride, const int32_t* src_col,
                                      float grad_scale, const void* sorted_keys,
                                      const int32_t* sorted_payload, const int32_t* bag_of_pos,
                                      int64_t N, int key_bits, void* weights, int weight_dtype,
                                      const int64_t* weights_offsets, const int32_t* D_offsets,
                                      const int64_t* hash_size_cumsum, int32_t T, int32_t B,
                                      int32_t max_D, int32_t all_D_mult8, const void* offsets,
                                      int index_dtype, const float* per_sample_weights,
                                      int pooling_mode, int optimizer, float* momentum1,
                                      const int64_t* momentum1_offsets, float learning_rate,
                                      float eps, void* partial_scratch, int32_t stochastic_rounding,
                                      uint64_t sr_seed, uint64_t sr_offset, void* stream) {
  if (!grad_bases || !src_col || n_src <= 0 || B_src <= 0 || (int64_t)n_src * B_src < B)
    return B200_EINVAL;
  return tbe_backward_impl(nullptr, grad_stride, sorted_keys, sorted_payload, bag_of_pos, N,
                           key_bits, weights, weight_dtype, weights_offsets, D_offsets,
                           hash_size_cumsum, T, B, max_D, all_D_mult8, offsets, index_dtype,
                           per_sample_weights, pooling_mode, optimizer, momentum1,
                           momentum1_offsets, learning_rate, eps, partial_scratch, grad_bases,
                           B_src, src_col, grad_scale, stochastic_rounding ? 1 : 0, sr_seed, sr_offset, stream);
}
