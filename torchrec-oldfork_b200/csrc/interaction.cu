// DLRM dot interaction for sm_100a (fp32, exact FFMA arithmetic).
//
// Replaces the cat + torch.bmm + triu gather + cat sequence of InteractionArch.forward
// (torchrec/models/dlrm.py:193-219) with one kernel per direction:
//   X = [dense; sparse] in [B, n = F+1, D];  Z = X X^T;
//   out = [dense | Z[i][j] for i < j, row-major]  in [B, D + n(n-1)/2].
// Each sample is read from HBM once, staged in shared memory, and the n(n-1)/2 dot
// products are produced from 3x3 register blocks; the [B, n, n] Gram matrix, the
// concatenated input and the gather indices of the reference never exist in memory.
#include <cuda.h>

#include <cstdlib>
#include <mutex>

#include "common.cuh"

namespace b200 {

constexpr int kIxGroup = 64;             // threads cooperating on one sample
constexpr int kIxSamplesPerCta = 4;
constexpr int kIxThreads = kIxGroup * kIxSamplesPerCta;

__device__ __forceinline__ int tri_index(int i, int j, int n) {  // i < j
  return i * (2 * n - i - 1) / 2 + (j - i - 1);
}

// smem row stride D + 4 floats keeps 16-byte alignment and staggers banks between rows.
__global__ void __launch_bounds__(kIxThreads)
interaction_forward_kernel(const float* __restrict__ dense, const float* __restrict__ sparse,
                           int32_t B, int32_t F, int32_t D, float* __restrict__ out, int64_t out_stride) {
  extern __shared__ __align__(16) float smem[];
  const int n = F + 1;
  const int ld = D + 4;
  const int grp = threadIdx.x / kIxGroup;
  const int tid = threadIdx.x % kIxGroup;
  float* X = smem + (size_t)grp * n * ld;
  const int out_dim = D + n * (n - 1) / 2;
  const int nb = (n + 2) / 3;                 // 3-row blocks
  const int n_blocks = nb * (nb + 1) / 2;     // upper-triangular block pairs
  const int D4 = D >> 2;

  for (int64_t s0 = (int64_t)blockIdx.x * kIxSamplesPerCta; s0 < B;
       s0 += (int64_t)gridDim.x * kIxSamplesPerCta) {
    const int64_t s = s0 + grp;
    const bool live = s < B;
    __syncthreads();  // previous sample fully consumed
    if (live) {
      const float4* d4 = reinterpret_cast<const float4*>(dense + s * D);
      const float4* s4 = reinterpret_cast<const float4*>(sparse + s * (int64_t)F * D);
      for (int c = tid; c < D4; c += kIxGroup)
        *reinterpret_cast<uint4*>(X + 4 * c) = ldg_stream_16(d4 + c);
      for (int c = tid; c < F * D4; c += kIxGroup) {
        const int r = c / D4, cc = c - r * D4;
        *reinterpret_cast<uint4*>(X + (size_t)(r + 1) * ld + 4 * cc) = ldg_stream_16(s4 + c);
      }
    }
    __syncthreads();
    if (live) {
      // dense passthrough; output rows are only 4-byte aligned (out_dim = 479 at F=26, D=128)
      for (int c = tid; c < D; c += kIxGroup) out[s * out_stride + c] = X[c];
      for (int c = out_dim + tid; c < out_stride; c += kIxGroup) out[s * out_stride + c] = 0.f;  // row padding
      float* o = out + s * out_stride + D;
      for (int blk = tid; blk < n_blocks; blk += kIxGroup) {
        // decode block pair (bi <= bj) from the linear index
        int bi = 0, rem = blk;
        while (rem >= nb - bi) {
          rem -= nb - bi;
          ++bi;
        }
        const int bj = bi + rem;
        const int i0 = 3 * bi, j0 = 3 * bj;
        float acc[3][3];
#pragma unroll
        for (int a = 0; a < 3; ++a)
#pragma unroll
          for (int b = 0; b < 3; ++b) acc[a][b] = 0.f;
        const float* xa[3];
        const float* xb[3];
#pragma unroll
        for (int a = 0; a < 3; ++a) {
          xa[a] = X + (size_t)min(i0 + a, n - 1) * ld;
          xb[a] = X + (size_t)min(j0 + a, n - 1) * ld;
        }
        for (int k = 0; k < D; k += 4) {
          float4 va[3], vb[3];
#pragma unroll
          for (int a = 0; a < 3; ++a) {
            va[a] = *reinterpret_cast<const float4*>(xa[a] + k);
            vb[a] = *reinterpret_cast<const float4*>(xb[a] + k);
          }
#pragma unroll
          for (int a = 0; a < 3; ++a)
#pragma unroll
            for (int b = 0; b < 3; ++b) {
              acc[a][b] = fmaf(va[a].x, vb[b].x, acc[a][b]);
              acc[a][b] = fmaf(va[a].y, vb[b].y, acc[a][b]);
              acc[a][b] = fmaf(va[a].z, vb[b].z, acc[a][b]);
              acc[a][b] = fmaf(va[a].w, vb[b].w, acc[a][b]);
            }
        }
#pragma unroll
        for (int a = 0; a < 3; ++a)
#pragma unroll
          for (int b = 0; b < 3; ++b) {
            const int i = i0 + a, j = j0 + b;
            if (i < j && j < n) o[tri_index(i, j, n)] = acc[a][b];
          }
      }
    }
  }
}

// grad_X[i] = sum_j Gs[i][j] X[j], Gs symmetric with zero diagonal built from grad_out's
// triangle; grad_dense additionally receives the passthrough part grad_out[:, :D].
__global__ void __launch_bounds__(kIxThreads)
interaction_backward_kernel(const float* __restrict__ dense, const float* __restrict__ sparse,
                            const float* __restrict__ grad_out, int64_t grad_stride, int32_t B, int32_t F,
                            int32_t D, float* __restrict__ grad_dense, float* __restrict__ grad_sparse) {
  extern __shared__ __align__(16) float smem[];
  const int n = F + 1;
  const int ld = D + 4;
  const int ldg = n + 1;
  const int grp = threadIdx.x / kIxGroup;
  const int tid = threadIdx.x % kIxGroup;
  const int per_group = (n * ld + n * ldg + 3) & ~3;  // keep every group's X 16-byte aligned
  float* X = smem + (size_t)grp * per_group;
  float* G = X + (size_t)n * ld;
  const int out_dim = D + n * (n - 1) / 2;
  const int D4 = D >> 2;
  const int n_tri = n * (n - 1) / 2;

  for (int64_t s0 = (int64_t)blockIdx.x * kIxSamplesPerCta; s0 < B;
       s0 += (int64_t)gridDim.x * kIxSamplesPerCta) {
    const int64_t s = s0 + grp;
    const bool live = s < B;
    __syncthreads();
    if (live) {
      const float4* d4 = reinterpret_cast<const float4*>(dense + s * D);
      const float4* s4 = reinterpret_cast<const float4*>(sparse + s * (int64_t)F * D);
      for (int c = tid; c < D4; c += kIxGroup)
        *reinterpret_cast<uint4*>(X + 4 * c) = ldg_stream_16(d4 + c);
      for (int c = tid; c < F * D4; c += kIxGroup) {
        const int r = c / D4, cc = c - r * D4;
        *reinterpret_cast<uint4*>(X + (size_t)(r + 1) * ld + 4 * cc) = ldg_stream_16(s4 + c);
      }
      const float* g = grad_out + s * grad_stride + D;
      for (int i = tid; i < n; i += kIxGroup) G[i * ldg + i] = 0.f;
      // walk the triangle row by row
      for (int p = tid; p < n_tri; p += kIxGroup) {
        int i = 0, rem = p;
        while (rem >= n - 1 - i) {
          rem -= n - 1 - i;
          ++i;
        }
        const int j = i + 1 + rem;
        const float v = g[p];
        G[i * ldg + j] = v;
        G[j * ldg + i] = v;
      }
    }
    __syncthreads();
    if (live) {
      const float* gpass = grad_out + s * grad_stride;
      // work item = (row i, float4 column c)
      for (int w = tid; w < n * D4; w += kIxGroup) {
        const int i = w / D4, c = w - i * D4;
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int j = 0; j < n; ++j) {
          const float gij = G[i * ldg + j];
          const float4 x = *reinterpret_cast<const float4*>(X + (size_t)j * ld + 4 * c);
          a.x = fmaf(gij, x.x, a.x);
          a.y = fmaf(gij, x.y, a.y);
          a.z = fmaf(gij, x.z, a.z);
          a.w = fmaf(gij, x.w, a.w);
        }
        if (i == 0) {
          a.x += gpass[4 * c];
          a.y += gpass[4 * c + 1];
          a.z += gpass[4 * c + 2];
          a.w += gpass[4 * c + 3];
          *reinterpret_cast<float4*>(grad_dense + s * D + 4 * c) = a;
        } else {
          *reinterpret_cast<float4*>(grad_sparse + (s * F + (i - 1)) * (int64_t)D + 4 * c) = a;
        }
      }
    }
  }
}


// ---------------------------------------------------------------------------------------------
// Tensor-core path (n = F+1 <= 32, D % 64 == 0): one warp per sample, mma.sync m16n8k8 with the
// 3xTF32 split (x = hi + lo, products lo*hi + hi*lo + hi*hi accumulated in fp32), which keeps the
// result within a few ulp of the fp32 FFMA chain.  X X^T is symmetric, so the A fragment of row
// tile i IS the pair of B fragments of column tiles 2i, 2i+1: each k-step loads 8 floats per lane
// and issues 18 MMAs (6 upper-triangular 16x8 tiles x 3 passes).
// Samples stream through warp-private cp.async rings of 64-column half-sample stages (row stride
// 68 floats: fragment loads are bank-conflict free), so no CTA-wide barrier sits on the path and
// 8 warps x 2 stages (~115 KB) stay in flight per SM.  The [D + n(n-1)/2] output row is assembled
// in shared memory and written with coalesced stores (rows are only 4-byte aligned).
// ---------------------------------------------------------------------------------------------
// 12 warps per SM: the kernels are latency-bound per warp (LDS -> split -> dependent MMAs), the
// tensor pipe is ~1/3 busy at 8 warps (ncu), so residency is what shared memory is spent on.
constexpr int kFwdMmaWarps = 16;
constexpr int kFwdKC = 32;
constexpr int kFwdStages = 3;   // columns per pipeline stage, forward (2 stages per warp)
constexpr int kBwdMmaWarps = 12;
constexpr int kBwdKC = 32;   // backward stages are quarter samples

__device__ __forceinline__ void mma_tf32(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                         uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
      "{%0,%1,%2,%3};\n"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// x = hi + lo with hi, lo representable in TF32 (10-bit mantissa).  hi is x rounded to nearest
// (integer add of half an ulp, then mask), so x - hi is exact in fp32 and symmetric around 0; lo is
// that remainder with the low 13 bits cleared (error <= 2^-21 |x|).  4 integer/FP ops instead of
// the two cvt.rna.tf32.f32, which sm_100 expands to ~5 instructions each (measured with ncu).
__device__ __forceinline__ void split_tf32(float x, uint32_t& hi, uint32_t& lo) {
  hi = (__float_as_uint(x) + 0x1000u) & 0xffffe000u;
  lo = __float_as_uint(x - __uint_as_float(hi)) & 0xffffe000u;
}

__device__ __forceinline__ void cp_async_16(float* dst_smem, const float* src) {
  const uint32_t sa = (uint32_t)__cvta_generic_to_shared(dst_smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_4(float* dst_smem, const float* src) {
  const uint32_t sa = (uint32_t)__cvta_generic_to_shared(dst_smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

template <bool FULL>  // FULL: 25 <= n <= 32, all six 16x8 tiles are live (no per-tile guards)
__global__ void __launch_bounds__(kFwdMmaWarps * 32, 1)
interaction_forward_mma_kernel(const float* __restrict__ dense, const float* __restrict__ sparse,
                               int32_t B, int32_t F, int32_t D, float* __restrict__ out,
                               int64_t out_stride) {
  extern __shared__ __align__(16) float smem[];
  const int n = F + 1;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  constexpr int ld = kFwdKC + 4;
  const int stage_floats = n * ld;
  const int out_dim = D + n * (n - 1) / 2;
  const int outs_floats = (out_dim + 3) & ~3;
  float* wbase = smem + (size_t)warp * (kFwdStages * stage_floats + outs_floats);
  float* outs = wbase + kFwdStages * stage_floats;
  const int nch = D / kFwdKC;
  const int ntiles = (n + 7) >> 3;  // live 8-row tiles
  const int64_t gw = (int64_t)blockIdx.x * kFwdMmaWarps + warp, tw = (int64_t)gridDim.x * kFwdMmaWarps;
  const int64_t Q = gw < B ? ((B - gw + tw - 1) / tw) * nch : 0;
  // rows at a 16-byte pitch (out_stride == round_up(out_dim, 4)) leave with 128-bit stores; the pad
  // columns carry zeros so a consumer may treat the row as out_stride wide
  const bool vec_out = out_stride == outs_floats && (reinterpret_cast<uintptr_t>(out) & 15) == 0;
  for (int c = out_dim + lane; c < outs_floats; c += 32) outs[c] = 0.f;

  // copy pattern of a stage: lane l moves 16-byte chunk (l & 15) of rows (l >> 4) + 2k, k = 0, 1, ...
  // Row 0 is the dense vector, rows 1..F the sample's sparse block (contiguous, D floats apart).
  constexpr int kCH = kFwdKC / 4, kRP = 32 / kCH;  // 16-byte chunks per row, rows per pass
  const int r0 = lane / kCH, cc4 = 4 * (lane % kCH);
  int stage_of = 0;  // q % 2 of the next stage to issue
  int64_t s_i = gw;  // sample / column chunk of the next stage to issue
  int h_i = 0;
  auto issue = [&](int64_t q) {
    if (q < Q) {
      float* dst = wbase + stage_of * stage_floats + r0 * ld + cc4;
      const float* src = sparse + s_i * (int64_t)F * D + (int64_t)(r0 - 1) * D + h_i * kFwdKC + cc4;
      if (r0 == 0) cp_async_16(dst, dense + s_i * D + h_i * kFwdKC + cc4);
      else if (r0 < n) cp_async_16(dst, src);
      for (int r = r0 + kRP; r < n; r += kRP) {
        dst += kRP * ld;
        src += kRP * D;
        cp_async_16(dst, src);
      }
      if (++h_i == nch) {
        h_i = 0;
        s_i += tw;
      }
    }
    stage_of = stage_of == kFwdStages - 1 ? 0 : stage_of + 1;
    cp_async_commit();  // always: keeps the group count uniform
  };
#pragma unroll
  for (int q = 0; q < kFwdStages - 1; ++q) issue(q);

  constexpr int kTM[6] = {0, 0, 0, 0, 1, 1};
  constexpr int kTN[6] = {0, 1, 2, 3, 2, 3};
  float acc[6][4];
  int rowoff[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) rowoff[j] = min(8 * j + g, n - 1) * ld + t;

  // output slot of every accumulator element of this lane (-1: lower triangle / padding)
  int ooff[6][4];
#pragma unroll
  for (int x = 0; x < 6; ++x)
#pragma unroll
    for (int y = 0; y < 4; ++y) {
      const int i = 16 * kTM[x] + g + ((y & 2) ? 8 : 0);
      const int j = 8 * kTN[x] + 2 * t + (y & 1);
      ooff[x][y] = (i < j && j < n) ? D + tri_index(i, j, n) : -1;
    }
  int stage_cons = 0;
  int64_t it = 0;
  int h = 0;
  for (int64_t q = 0; q < Q; ++q) {
    cp_async_wait<kFwdStages - 2>();
    __syncwarp();
    issue(q + kFwdStages - 1);
    const float* st = wbase + stage_cons * stage_floats;
    stage_cons = stage_cons == kFwdStages - 1 ? 0 : stage_cons + 1;
    if (h == 0) {
#pragma unroll
      for (int x = 0; x < 6; ++x)
#pragma unroll
        for (int y = 0; y < 4; ++y) acc[x][y] = 0.f;
    }
#pragma unroll
    for (int c = 0; c < kFwdKC; c += 32) outs[h * kFwdKC + c + lane] = st[c + lane];  // dense passthrough
#pragma unroll 2
    for (int k0 = 0; k0 < kFwdKC; k0 += 8) {
      uint32_t bh[4][2], bl[4][2];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        split_tf32(st[rowoff[j] + k0], bh[j][0], bl[j][0]);
        split_tf32(st[rowoff[j] + k0 + 4], bh[j][1], bl[j][1]);
      }
#pragma unroll
      for (int x = 0; x < 6; ++x) {
        const int mi = kTM[x], nj = kTN[x];
        if (FULL || (nj < ntiles && 2 * mi < ntiles)) {
          mma_tf32(acc[x], bl[2 * mi][0], bl[2 * mi + 1][0], bl[2 * mi][1], bl[2 * mi + 1][1], bh[nj][0], bh[nj][1]);
          mma_tf32(acc[x], bh[2 * mi][0], bh[2 * mi + 1][0], bh[2 * mi][1], bh[2 * mi + 1][1], bl[nj][0], bl[nj][1]);
          mma_tf32(acc[x], bh[2 * mi][0], bh[2 * mi + 1][0], bh[2 * mi][1], bh[2 * mi + 1][1], bh[nj][0], bh[nj][1]);
        }
      }
    }
    if (h == nch - 1) {
#pragma unroll
      for (int x = 0; x < 6; ++x) {
#pragma unroll
        for (int y = 0; y < 4; ++y)
          if (ooff[x][y] >= 0) outs[ooff[x][y]] = acc[x][y];
      }
      __syncwarp();
      float* dst = out + (gw + it * tw) * out_stride;
      if (vec_out) {
        for (int c = lane; c < (outs_floats >> 2); c += 32)
          reinterpret_cast<float4*>(dst)[c] = reinterpret_cast<const float4*>(outs)[c];
      } else {
        for (int c = lane; c < out_dim; c += 32) dst[c] = outs[c];
        for (int c = out_dim + lane; c < out_stride; c += 32) dst[c] = 0.f;
      }
    }
    if (++h == nch) {
      h = 0;
      ++it;
    }
  }
}


// ---------------------------------------------------------------------------------------------
// TMA variant of the forward (default when the tensor maps can be encoded): the per-lane cp.async copies of
// a stage are replaced by TWO tensor-memory-accelerator bulk copies issued by one lane - the dense row
// (box 32 x 1 of the [B, D] tensor) and the sample's F sparse rows (box 32 x F of the [B*F, D] tensor) - into
// a 128-byte-swizzled tile (row = 128 bytes, 16-byte chunk c of row r lives at chunk c ^ (r & 7)), completing
// on a per-warp, per-stage mbarrier.  The swizzle makes the MMA fragment loads (8 rows x 4 consecutive
// floats per k-step) bank-conflict free without the row padding of the cp.async layout, the copy engine
// does the address generation, and no cp.async group bookkeeping sits in the warp's instruction stream.
// ---------------------------------------------------------------------------------------------
constexpr int kTmaTileBytes = 32 * 128;  // 32 rows x 32 floats, 1024-byte aligned (swizzle period)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(bar), "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const void* tmap, int32_t c0, int32_t c1, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
      ::"r"(dst), "l"(tmap), "r"(c0), "r"(c1), "r"(bar)
      : "memory");
}

template <bool FULL>
__global__ void __launch_bounds__(kFwdMmaWarps * 32, 1)
interaction_forward_tma_kernel(const __grid_constant__ CUtensorMap tm_dense,
                               const __grid_constant__ CUtensorMap tm_sparse, int32_t B, int32_t F, int32_t D,
                               float* __restrict__ out, int64_t out_stride) {
  extern __shared__ __align__(16) unsigned char smem_dyn[];
  // the swizzle pattern is a function of the shared-memory ADDRESS: tiles start on a 1024-byte boundary
  unsigned char* smem_raw = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
  const int n = F + 1;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int out_dim = D + n * (n - 1) / 2;
  const int outs_floats = (out_dim + 3) & ~3;
  // layout: [warps][stages] tiles (1024-aligned) | [warps] output rows | [warps][stages] mbarriers
  unsigned char* tiles = smem_raw + (size_t)warp * kFwdStages * kTmaTileBytes;
  float* outs = reinterpret_cast<float*>(smem_raw + (size_t)kFwdMmaWarps * kFwdStages * kTmaTileBytes) +
                (size_t)warp * outs_floats;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + (size_t)kFwdMmaWarps * kFwdStages * kTmaTileBytes +
                                               (size_t)kFwdMmaWarps * outs_floats * sizeof(float)) +
                   warp * kFwdStages;
  const int nch = D / kFwdKC;
  const int ntiles = (n + 7) >> 3;
  const int64_t gw = (int64_t)blockIdx.x * kFwdMmaWarps + warp, tw = (int64_t)gridDim.x * kFwdMmaWarps;
  const int64_t Q = gw < B ? ((B - gw + tw - 1) / tw) * nch : 0;
  const bool vec_out = out_stride == outs_floats && (reinterpret_cast<uintptr_t>(out) & 15) == 0;
  for (int c = out_dim + lane; c < outs_floats; c += 32) outs[c] = 0.f;
  if (lane == 0) {
#pragma unroll
    for (int st = 0; st < kFwdStages; ++st) mbar_init(smem_u32(&bars[st]), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  const uint32_t tx_bytes = (uint32_t)n * 128u;

  int stage_of = 0;
  int64_t s_i = gw;
  int h_i = 0;
  auto issue = [&](int64_t q) {
    if (q < Q && lane == 0) {
      const uint32_t bar = smem_u32(&bars[stage_of]);
      const uint32_t dst = smem_u32(tiles + (size_t)stage_of * kTmaTileBytes);
      mbar_expect_tx(bar, tx_bytes);
      tma_load_2d(dst, &tm_dense, h_i * kFwdKC, (int32_t)s_i, bar);
      tma_load_2d(dst + 128, &tm_sparse, h_i * kFwdKC, (int32_t)(s_i * F), bar);
    }
    if (q < Q) {
      if (++h_i == nch) {
        h_i = 0;
        s_i += tw;
      }
    }
    stage_of = stage_of == kFwdStages - 1 ? 0 : stage_of + 1;
  };
#pragma unroll
  for (int q = 0; q < kFwdStages - 1; ++q) issue(q);

  constexpr int kTM[6] = {0, 0, 0, 0, 1, 1};
  constexpr int kTN[6] = {0, 1, 2, 3, 2, 3};
  float acc[6][4];
  int rbase[4], rx[4];  // float offset of the lane's row in the tile, and its swizzle key (row & 7)
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int r = min(8 * j + g, n - 1);
    rbase[j] = r * 32 + t;
    rx[j] = r & 7;
  }
  int ooff[6][4];
#pragma unroll
  for (int x = 0; x < 6; ++x)
#pragma unroll
    for (int y = 0; y < 4; ++y) {
      const int i = 16 * kTM[x] + g + ((y & 2) ? 8 : 0);
      const int j = 8 * kTN[x] + 2 * t + (y & 1);
      ooff[x][y] = (i < j && j < n) ? D + tri_index(i, j, n) : -1;
    }
  int stage_cons = 0;
  uint32_t phase = 0;  // bit st = parity to wait for on stage st
  int64_t it = 0;
  int h = 0;
  for (int64_t q = 0; q < Q; ++q) {
    __syncwarp();  // every lane is done with the stage that the next issue overwrites
    issue(q + kFwdStages - 1);
    mbar_wait(smem_u32(&bars[stage_cons]), (phase >> stage_cons) & 1u);
    phase ^= 1u << stage_cons;
    const float* st = reinterpret_cast<const float*>(tiles + (size_t)stage_cons * kTmaTileBytes);
    stage_cons = stage_cons == kFwdStages - 1 ? 0 : stage_cons + 1;
    if (h == 0) {
#pragma unroll
      for (int x = 0; x < 6; ++x)
#pragma unroll
        for (int y = 0; y < 4; ++y) acc[x][y] = 0.f;
    }
    outs[h * kFwdKC + lane] = st[lane];  // dense passthrough: row 0 is not swizzled (r & 7 == 0)
#pragma unroll 2
    for (int k0 = 0; k0 < kFwdKC; k0 += 8) {
      uint32_t bh[4][2], bl[4][2];
      const int c0 = k0 >> 2;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        split_tf32(st[rbase[j] + ((c0 ^ rx[j]) << 2)], bh[j][0], bl[j][0]);
        split_tf32(st[rbase[j] + (((c0 + 1) ^ rx[j]) << 2)], bh[j][1], bl[j][1]);
      }
#pragma unroll
      for (int x = 0; x < 6; ++x) {
        const int mi = kTM[x], nj = kTN[x];
        if (FULL || (nj < ntiles && 2 * mi < ntiles)) {
          mma_tf32(acc[x], bl[2 * mi][0], bl[2 * mi + 1][0], bl[2 * mi][1], bl[2 * mi + 1][1], bh[nj][0], bh[nj][1]);
          mma_tf32(acc[x], bh[2 * mi][0], bh[2 * mi + 1][0], bh[2 * mi][1], bh[2 * mi + 1][1], bl[nj][0], bl[nj][1]);
          mma_tf32(acc[x], bh[2 * mi][0], bh[2 * mi + 1][0], bh[2 * mi][1], bh[2 * mi + 1][1], bh[nj][0], bh[nj][1]);
        }
      }
    }
    if (h == nch - 1) {
#pragma unroll
      for (int x = 0; x < 6; ++x) {
#pragma unroll
        for (int y = 0; y < 4; ++y)
          if (ooff[x][y] >= 0) outs[ooff[x][y]] = acc[x][y];
      }
      __syncwarp();
      float* dst = out + (gw + it * tw) * out_stride;
      if (vec_out) {
        for (int c = lane; c < (outs_floats >> 2); c += 32)
          reinterpret_cast<float4*>(dst)[c] = reinterpret_cast<const float4*>(outs)[c];
      } else {
        for (int c = lane; c < out_dim; c += 32) dst[c] = outs[c];
        for (int c = out_dim + lane; c < out_stride; c += 32) dst[c] = 0.f;
      }
    }
    if (++h == nch) {
      h = 0;
      ++it;
    }
  }
}


// Backward on the same machinery: grad_X = Gs X, Gs (n x n, symmetric, zero diagonal) scattered
// from grad_out's triangle into a warp-private 32x36 tile (A operand, split once per sample and
// kept in registers), X streamed in 64-column stages with row stride 72 (B-fragment loads are
// conflict free).  grad_out rows ride in the same cp.async groups (4-byte copies: rows are only
// 4-byte aligned).  Results leave as 8-byte stores that fill whole 32-byte sectors.
// SCATTER: the gradient row of feature f and sample s goes to (float*)feat_base[f] + s * feat_pitch[f]
// instead of grad_sparse[s][f][:] - with NVLink peer mappings in feat_base this kernel IS the
// backward all-to-all of table-wise sharded embeddings (every feature's gradient lands in its
// owner's buffer while the tile is being computed; torchrec comm_ops.py:585-640 runs that as a
// re-layout copy + NCCL all_to_all_single after the kernel).
template <bool FULL, bool SCATTER>  // FULL: 25 <= n <= 32 (four k-steps, both row tiles live)
__global__ void __launch_bounds__(kBwdMmaWarps * 32, 1)
interaction_backward_mma_kernel(const float* __restrict__ dense, const float* __restrict__ sparse,
                                const float* __restrict__ grad_out, int64_t grad_stride, int32_t B,
                                int32_t F, int32_t D, float* __restrict__ grad_dense,
                                float* __restrict__ grad_sparse,
                                const int64_t* __restrict__ feat_base,
                                const int64_t* __restrict__ feat_pitch) {
  __shared__ long long s_fbase[32];
  __shared__ long long s_fpitch[32];
  if constexpr (SCATTER) {
    if (threadIdx.x < F) {
      s_fbase[threadIdx.x] = feat_base[threadIdx.x];
      s_fpitch[threadIdx.x] = feat_pitch[threadIdx.x];
    }
  }
  extern __shared__ __align__(16) float smem[];
  const int n = F + 1;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  constexpr int ld = kBwdKC + 8;
  constexpr int ldG = 36;
  const int stage_floats = n * ld;
  const int n_tri = n * (n - 1) / 2;
  const int out_dim = D + n_tri;
  const int gs_floats = (out_dim + 3) & ~3;
  const int per_warp = 2 * stage_floats + 32 * ldG + 2 * gs_floats;
  uint16_t* ij = reinterpret_cast<uint16_t*>(smem + (size_t)kBwdMmaWarps * per_warp);  // [n_tri] (i | j << 8)
  float* wbase = smem + (size_t)warp * per_warp;
  float* Gs = wbase + 2 * stage_floats;
  float* gst = Gs + 32 * ldG;
  const int nch = D / kBwdKC;
  const int ksteps = (n + 7) >> 3;
  const int64_t gw = (int64_t)blockIdx.x * kBwdMmaWarps + warp, tw = (int64_t)gridDim.x * kBwdMmaWarps;
  const int64_t Q = gw < B ? ((B - gw + tw - 1) / tw) * nch : 0;
  // 16-byte copies of grad_out rows when their pitch allows it (the pad floats are read, never used)
  const bool vec_g = (grad_stride & 3) == 0 && grad_stride >= gs_floats &&
                     (reinterpret_cast<uintptr_t>(grad_out) & 15) == 0;

  for (int p = threadIdx.x; p < n_tri; p += blockDim.x) {
    int i = 0, rem = p;
    while (rem >= n - 1 - i) {
      rem -= n - 1 - i;
      ++i;
    }
    ij[p] = (uint16_t)(i | ((i + 1 + rem) << 8));
  }
  for (int c = lane; c < 32 * ldG; c += 32) Gs[c] = 0.f;  // diagonal and padding stay zero
  __syncthreads();

  // same copy pattern as the forward kernel; the grad_out row rides in the group of chunk 0
  constexpr int kCH = kBwdKC / 4, kRP = 32 / kCH;  // 16-byte chunks per row, rows per pass
  const int r0 = lane / kCH, cc4 = 4 * (lane % kCH);
  int64_t s_i = gw, it_i = 0;
  int h_i = 0, stage_of = 0;
  auto issue = [&](int64_t q) {
    if (q < Q) {
      float* dst = wbase + stage_of * stage_floats + r0 * ld + cc4;
      const float* src = sparse + s_i * (int64_t)F * D + (int64_t)(r0 - 1) * D + h_i * kBwdKC + cc4;
      if (r0 == 0) cp_async_16(dst, dense + s_i * D + h_i * kBwdKC + cc4);
      else if (r0 < n) cp_async_16(dst, src);
      for (int r = r0 + kRP; r < n; r += kRP) {
        dst += kRP * ld;
        src += kRP * D;
        cp_async_16(dst, src);
      }
      if (h_i == 0) {
        float* gd = gst + (it_i & 1) * gs_floats;
        const float* gsrc = grad_out + s_i * grad_stride;
        if (vec_g) {
          for (int c = lane; c < (gs_floats >> 2); c += 32) cp_async_16(gd + 4 * c, gsrc + 4 * c);
        } else {
          for (int c = lane; c < out_dim; c += 32) cp_async_4(gd + c, gsrc + c);
        }
      }
      if (++h_i == nch) {
        h_i = 0;
        s_i += tw;
        ++it_i;
      }
    }
    stage_of ^= 1;
    cp_async_commit();
  };
  issue(0);

  uint32_t ah[2][4][4], al[2][4][4];  // [row tile][k-step][fragment]
  int rowoff[4][2];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    rowoff[k][0] = min(8 * k + t, n - 1) * ld + g;
    rowoff[k][1] = min(8 * k + t + 4, n - 1) * ld + g;
  }

  int64_t it = 0, s = gw;
  int h = 0;
  for (int64_t q = 0; q < Q; ++q) {
    cp_async_wait<0>();
    __syncwarp();
    issue(q + 1);
    const float* st = wbase + (q & 1) * stage_floats;
    const float* gcur = gst + (it & 1) * gs_floats;
    if (h == 0) {
      for (int p = lane; p < n_tri; p += 32) {
        const int i = ij[p] & 0xff, j = ij[p] >> 8;
        const float v = gcur[D + p];
        Gs[i * ldG + j] = v;
        Gs[j * ldG + i] = v;
      }
      __syncwarp();
#pragma unroll
      for (int mi = 0; mi < 2; ++mi)
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float* a = Gs + (16 * mi + g) * ldG + 8 * k + t;
          split_tf32(a[0], ah[mi][k][0], al[mi][k][0]);
          split_tf32(a[8 * ldG], ah[mi][k][1], al[mi][k][1]);
          split_tf32(a[4], ah[mi][k][2], al[mi][k][2]);
          split_tf32(a[8 * ldG + 4], ah[mi][k][3], al[mi][k][3]);
        }
      __syncwarp();  // Gs is rewritten by the next sample only after every lane has its fragments
    }
#pragma unroll 1
    for (int nt = 0; nt < kBwdKC / 8; ++nt) {
      float acc[2][4];
#pragma unroll
      for (int mi = 0; mi < 2; ++mi)
#pragma unroll
        for (int y = 0; y < 4; ++y) acc[mi][y] = 0.f;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (FULL || k < ksteps) {
          uint32_t bh0, bl0, bh1, bl1;
          split_tf32(st[rowoff[k][0] + 8 * nt], bh0, bl0);
          split_tf32(st[rowoff[k][1] + 8 * nt], bh1, bl1);
#pragma unroll
          for (int mi = 0; mi < 2; ++mi) {
            if (FULL || 16 * mi < n) {
              mma_tf32(acc[mi], al[mi][k][0], al[mi][k][1], al[mi][k][2], al[mi][k][3], bh0, bh1);
              mma_tf32(acc[mi], ah[mi][k][0], ah[mi][k][1], ah[mi][k][2], ah[mi][k][3], bl0, bl1);
              mma_tf32(acc[mi], ah[mi][k][0], ah[mi][k][1], ah[mi][k][2], ah[mi][k][3], bh0, bh1);
            }
          }
        }
      }
      const int d = h * kBwdKC + 8 * nt + 2 * t;
      // SCATTER: rows may live on another GPU, where 8-byte stores waste the link (measured: 2.8x the
      // kernel time).  The finished 8-column block replaces the X block it was computed from (nobody
      // reads those columns again) and the quarter sample leaves below as 128-byte row segments.
      if constexpr (SCATTER) __syncwarp();  // every lane has loaded its B fragments of this block
#pragma unroll
      for (int mi = 0; mi < 2; ++mi)
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const int i = 16 * mi + g + 8 * half;
          if (i < n) {
            float2 v = make_float2(acc[mi][2 * half], acc[mi][2 * half + 1]);
            if (i == 0) {
              v.x += gcur[d];
              v.y += gcur[d + 1];
              *reinterpret_cast<float2*>(grad_dense + s * D + d) = v;
            } else if constexpr (SCATTER) {
              *reinterpret_cast<float2*>(const_cast<float*>(st) + i * ld + 8 * nt + 2 * t) = v;
            } else {
              *reinterpret_cast<float2*>(grad_sparse + (s * F + (i - 1)) * (int64_t)D + d) = v;
            }
          }
        }
    }
    if constexpr (SCATTER) {
      __syncwarp();
      constexpr int kChunks = kBwdKC / 4;  // 16-byte chunks per row segment
      for (int r = 1 + lane / kChunks; r < n; r += 32 / kChunks) {
        const float4 v = *reinterpret_cast<const float4*>(st + r * ld + 4 * (lane % kChunks));
        float* row = reinterpret_cast<float*>(s_fbase[r - 1]) + s * s_fpitch[r - 1];
        *reinterpret_cast<float4*>(row + h * kBwdKC + 4 * (lane % kChunks)) = v;
      }
      // the next iteration's __syncwarp (before it refills this stage) orders these loads
    }
    if (++h == nch) {
      h = 0;
      ++it;
      s += tw;
    }
  }
}

}  // namespace b200

namespace b200 {

// cuTensorMapEncodeTiled through the runtime's driver entry point (the library links cudart statically and does
// not link libcuda).
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return (EncodeTiledFn)p;
  }();
  return fn;
}

// B200_IX_TMA=1 makes the TMA kernel the default forward.  Measured at the Criteo shape (B 65536, F 26, D 128):
// 258.6 us against 249.4 us for the cp.async ring - the kernel is bound by its per-warp LDS -> split -> MMA chain,
// not by copy issue slots - so the cp.async kernel stays the default and the TMA kernel is reachable through
// b200_interaction_forward_tma (bit-identical results: same MMA sequence).
static bool ix_tma_enabled() {
  static const bool on = [] {
    const char* e = getenv("B200_IX_TMA");
    return e && e[0] == '1';
  }();
  return on && encode_tiled_fn() != nullptr;
}

// [rows, D] fp32 row-major -> tensor map with a 32-float x box_rows box, 128-byte swizzle
static bool encode_rows_map(CUtensorMap* tm, const float* base, uint64_t rows, uint32_t D, uint32_t box_rows) {
  const cuuint64_t dims[2] = {D, rows};
  const cuuint64_t strides[1] = {(cuuint64_t)D * sizeof(float)};
  const cuuint32_t box[2] = {(cuuint32_t)kFwdKC, box_rows};
  const cuuint32_t estr[2] = {1, 1};
  return encode_tiled_fn()(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box,
                           estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                           CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

static int interaction_forward_tma(const float* dense, const float* sparse, int32_t B, int32_t F, int32_t D,
                                   float* out, int64_t out_stride, cudaStream_t stream) {
  const int n = F + 1;
  if ((reinterpret_cast<uintptr_t>(dense) & 15) || (reinterpret_cast<uintptr_t>(sparse) & 15) || F > 256)
    return B200_EUNSUPPORTED;
  const int out_dim = D + n * (n - 1) / 2;
  const size_t smem = (size_t)kFwdMmaWarps * kFwdStages * kTmaTileBytes +
                      (size_t)kFwdMmaWarps * ((out_dim + 3) & ~3) * sizeof(float) +
                      (size_t)kFwdMmaWarps * kFwdStages * sizeof(uint64_t) + 1024;  // + alignment slack
  if (smem > 227 * 1024) return B200_EUNSUPPORTED;
  CUtensorMap tm_dense, tm_sparse;
  if (!encode_rows_map(&tm_dense, dense, (uint64_t)B, (uint32_t)D, 1) ||
      !encode_rows_map(&tm_sparse, sparse, (uint64_t)B * F, (uint32_t)D, (uint32_t)F))
    return B200_EUNSUPPORTED;
  auto kern = n > 24 ? interaction_forward_tma_kernel<true> : interaction_forward_tma_kernel<false>;
  B200_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int64_t grid = ceil_div(B, kFwdMmaWarps);
  if (grid > kNumSMs) grid = kNumSMs;
  kern<<<(unsigned)grid, kFwdMmaWarps * 32, smem, stream>>>(tm_dense, tm_sparse, B, F, D, out, out_stride);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

}  // namespace b200

extern "C" int b200_interaction_forward_tma(const float* dense, const float* sparse, int32_t B, int32_t F,
                                            int32_t D, float* out, int64_t out_stride, void* stream) {
  using namespace b200;
  if (B == 0) return B200_OK;
  if (D <= 0 || (D & 3) || F < 0) return B200_EINVAL;
  if (out_stride < D + (int64_t)(F + 1) * F / 2) return B200_EINVAL;
  if (F + 1 > 32 || F < 1 || D % kFwdKC != 0 || encode_tiled_fn() == nullptr) return B200_EUNSUPPORTED;
  return interaction_forward_tma(dense, sparse, B, F, D, out, out_stride, (cudaStream_t)stream);
}

extern "C" int b200_interaction_forward(const float* dense, const float* sparse, int32_t B,
                                        int32_t F, int32_t D, float* out, int64_t out_stride,
                                        void* stream) {
  using namespace b200;
  if (B == 0) return B200_OK;
  if (D <= 0 || (D & 3) || F < 0) return B200_EINVAL;
  if (out_stride < D + (int64_t)(F + 1) * F / 2) return B200_EINVAL;
  const int n = F + 1;
  if (n <= 32 && F >= 1 && D % kFwdKC == 0 && ix_tma_enabled()) {
    const int rc = interaction_forward_tma(dense, sparse, B, F, D, out, out_stride, (cudaStream_t)stream);
    if (rc != B200_EUNSUPPORTED) return rc;  // else: the cp.async kernel below
  }
  if (n <= 32 && D % kFwdKC == 0) {
    const int out_dim = D + n * (n - 1) / 2;
    const size_t msmem = sizeof(float) * kFwdMmaWarps * (kFwdStages * (size_t)n * (kFwdKC + 4) + ((out_dim + 3) & ~3));
    if (msmem <= 227 * 1024) {
      auto kern = n > 24 ? interaction_forward_mma_kernel<true> : interaction_forward_mma_kernel<false>;
      B200_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)msmem));
      int64_t mgrid = ceil_div(B, kFwdMmaWarps);
      if (mgrid > kNumSMs) mgrid = kNumSMs;
      kern<<<(unsigned)mgrid, kFwdMmaWarps * 32, msmem, (cudaStream_t)stream>>>(
          dense, sparse, B, F, D, out, out_stride);
      B200_LAUNCH_CHECK();
      return B200_OK;
    }
  }
  const size_t smem = sizeof(float) * kIxSamplesPerCta * (size_t)n * (D + 4);
  if (smem > 200 * 1024) return B200_EUNSUPPORTED;
  B200_CUDA_TRY(cudaFuncSetAttribute(interaction_forward_kernel,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int64_t grid = ceil_div(B, kIxSamplesPerCta);
  if (grid > (int64_t)kNumSMs * 8) grid = (int64_t)kNumSMs * 8;
  interaction_forward_kernel<<<(unsigned)grid, kIxThreads, smem, (cudaStream_t)stream>>>(
      dense, sparse, B, F, D, out, out_stride);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

static int interaction_backward_impl(const float* dense, const float* sparse, const float* grad_out,
                                     int64_t grad_stride, int32_t B, int32_t F, int32_t D,
                                     float* grad_dense, float* grad_sparse, const int64_t* feat_base,
                                     const int64_t* feat_pitch, void* stream) {
  using namespace b200;
  const bool scatter = feat_base != nullptr;
  if (B == 0) return B200_OK;
  if (D <= 0 || (D & 3) || F < 0) return B200_EINVAL;
  if (grad_stride < D + (int64_t)(F + 1) * F / 2) return B200_EINVAL;
  const int n = F + 1;
  if (n <= 32 && D % kBwdKC == 0) {
    const int n_tri = n * (n - 1) / 2;
    const size_t per_warp = 2 * (size_t)n * (kBwdKC + 8) + 32 * 36 + 2 * (size_t)((D + n_tri + 3) & ~3);
    const size_t msmem = sizeof(float) * kBwdMmaWarps * per_warp + ((sizeof(uint16_t) * n_tri + 15) & ~(size_t)15);
    if (msmem <= 226 * 1024) {
      auto kern = scatter ? (n > 24 ? interaction_backward_mma_kernel<true, true>
                                    : interaction_backward_mma_kernel<false, true>)
                          : (n > 24 ? interaction_backward_mma_kernel<true, false>
                                    : interaction_backward_mma_kernel<false, false>);
      B200_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)msmem));
      int64_t mgrid = ceil_div(B, kBwdMmaWarps);
      if (mgrid > kNumSMs) mgrid = kNumSMs;
      kern<<<(unsigned)mgrid, kBwdMmaWarps * 32, msmem, (cudaStream_t)stream>>>(
          dense, sparse, grad_out, grad_stride, B, F, D, grad_dense, grad_sparse, feat_base, feat_pitch);
      B200_LAUNCH_CHECK();
      return B200_OK;
    }
  }
  if (scatter) return B200_EUNSUPPORTED;  // the caller falls back to the dense layout
  const size_t smem = sizeof(float) * kIxSamplesPerCta * (((size_t)n * (D + 4) + (size_t)n * (n + 1) + 3) & ~(size_t)3);
  if (smem > 200 * 1024) return B200_EUNSUPPORTED;
  B200_CUDA_TRY(cudaFuncSetAttribute(interaction_backward_kernel,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int64_t grid = ceil_div(B, kIxSamplesPerCta);
  if (grid > (int64_t)kNumSMs * 8) grid = (int64_t)kNumSMs * 8;
  interaction_backward_kernel<<<(unsigned)grid, kIxThreads, smem, (cudaStream_t)stream>>>(
      dense, sparse, grad_out, grad_stride, B, F, D, grad_dense, grad_sparse);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_interaction_backward(const float* dense, const float* sparse,
                                         const float* grad_out, int64_t grad_stride, int32_t B,
                                         int32_t F, int32_t D, float* grad_dense,
                                         float* grad_sparse, void* stream) {
  return interaction_backward_impl(dense, sparse, grad_out, grad_stride, B, F, D, grad_dense, grad_sparse,
                                   nullptr, nullptr, stream);
}

extern "C" int b200_interaction_backward_scatter(const float* dense, const float* sparse,
                                                 const float* grad_out, int64_t grad_stride, int32_t B,
                                                 int32_t F, int32_t D, float* grad_dense,
                                                 const int64_t* feat_base, const int64_t* feat_pitch,
                                                 void* stream) {
  if (!feat_base || !feat_pitch) return B200_EINVAL;
  return interaction_backward_impl(dense, sparse, grad_out, grad_stride, B, F, D, grad_dense, nullptr,
                                   feat_base, feat_pitch, stream);
}
