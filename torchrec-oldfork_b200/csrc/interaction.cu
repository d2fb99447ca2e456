// DLRM dot interaction for sm_100a (fp32, exact FFMA arithmetic).
//
// Replaces the cat + torch.bmm + triu gather + cat sequence of InteractionArch.forward
// (torchrec/models/dlrm.py:193-219) with one kernel per direction:
//   X = [dense; sparse] in [B, n = F+1, D];  Z = X X^T;
//   out = [dense | Z[i][j] for i < j, row-major]  in [B, D + n(n-1)/2].
// Each sample is read from HBM once, staged in shared memory, and the n(n-1)/2 dot
// products are produced from 3x3 register blocks; the [B, n, n] Gram matrix, the
// concatenated input and the gather indices of the reference never exist in memory.
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
                           int32_t B, int32_t F, int32_t D, float* __restrict__ out) {
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
      for (int c = tid; c < D; c += kIxGroup) out[s * out_dim + c] = X[c];
      float* o = out + s * out_dim + D;
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
                            const float* __restrict__ grad_out, int32_t B, int32_t F, int32_t D,
                            float* __restrict__ grad_dense, float* __restrict__ grad_sparse) {
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
      const float* g = grad_out + s * out_dim + D;
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
      const float* gpass = grad_out + s * out_dim;
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

}  // namespace b200

extern "C" int b200_interaction_forward(const float* dense, const float* sparse, int32_t B,
                                        int32_t F, int32_t D, float* out, void* stream) {
  using namespace b200;
  if (B == 0) return B200_OK;
  if (D <= 0 || (D & 3) || F < 0) return B200_EINVAL;
  const int n = F + 1;
  const size_t smem = sizeof(float) * kIxSamplesPerCta * (size_t)n * (D + 4);
  if (smem > 200 * 1024) return B200_EUNSUPPORTED;
  B200_CUDA_TRY(cudaFuncSetAttribute(interaction_forward_kernel,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int64_t grid = ceil_div(B, kIxSamplesPerCta);
  if (grid > (int64_t)kNumSMs * 8) grid = (int64_t)kNumSMs * 8;
  interaction_forward_kernel<<<(unsigned)grid, kIxThreads, smem, (cudaStream_t)stream>>>(
      dense, sparse, B, F, D, out);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_interaction_backward(const float* dense, const float* sparse,
                                         const float* grad_out, int32_t B, int32_t F, int32_t D,
                                         float* grad_dense, float* grad_sparse, void* stream) {
  using namespace b200;
  if (B == 0) return B200_OK;
  if (D <= 0 || (D & 3) || F < 0) return B200_EINVAL;
  const int n = F + 1;
  const size_t smem = sizeof(float) * kIxSamplesPerCta * (((size_t)n * (D + 4) + (size_t)n * (n + 1) + 3) & ~(size_t)3);
  if (smem > 200 * 1024) return B200_EUNSUPPORTED;
  B200_CUDA_TRY(cudaFuncSetAttribute(interaction_backward_kernel,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int64_t grid = ceil_div(B, kIxSamplesPerCta);
  if (grid > (int64_t)kNumSMs * 8) grid = (int64_t)kNumSMs * 8;
  interaction_backward_kernel<<<(unsigned)grid, kIxThreads, smem, (cudaStream_t)stream>>>(
      dense, sparse, grad_out, B, F, D, grad_dense, grad_sparse);
  B200_LAUNCH_CHECK();
  return B200_OK;
}
