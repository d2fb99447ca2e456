// Backward glue of a Linear+ReLU layer (torchrec/modules/mlp.py Perceptron :19-79 of the reference):
//   grad_pre[b, n] = grad_out[b, n] * (out[b, n] > 0)          (ReLU backward)
//   grad_bias[n]   = sum_b grad_pre[b, n]                       (bias gradient)
// PyTorch runs these as threshold_backward + a separate column reduction, i.e. grad_pre makes a
// second trip through HBM (at the Criteo over-arch, 268 MB per 1024-wide layer).  Here one pass
// reads grad_out and out, writes grad_pre and keeps per-CTA column sums in registers; a second tiny
// kernel adds the per-CTA partials in a fixed order (deterministic, no atomics).  The GEMMs of the
// layer stay on cuBLAS.
#include "common.cuh"

namespace b200 {

constexpr int kMlpThreads = 256;
constexpr int kMlpMaxParts = 592;  // 148 SMs x 4 row chunks

// Thread t owns float4 column (t % C4) and rows r0 + (t / C4) + k * (256 / C4) of the CTA's chunk;
// N <= 1024 so C4 <= 256 (C4 is rounded up to a power of two that divides 256).
__global__ void __launch_bounds__(kMlpThreads)
relu_bgrad_kernel(const float* __restrict__ grad_out, int64_t go_stride, const float* __restrict__ out,
                  int64_t out_stride, int64_t B, int32_t N, int32_t C4, int64_t rows_per_cta,
                  float* __restrict__ grad_pre, float* __restrict__ partials) {
  __shared__ float4 s_acc[kMlpThreads];
  const int c4 = threadIdx.x % C4;
  const int rsub = threadIdx.x / C4;
  const int rstep = kMlpThreads / C4;
  const bool col_ok = 4 * c4 < N;
  const int64_t r_begin = (int64_t)blockIdx.x * rows_per_cta;
  const int64_t r_end = min(B, r_begin + rows_per_cta);
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (col_ok) {
#pragma unroll 4
    for (int64_t r = r_begin + rsub; r < r_end; r += rstep) {
      const uint4 gu = ldg_stream_16(grad_out + r * go_stride + 4 * c4);
      const float4 y = *reinterpret_cast<const float4*>(out + r * out_stride + 4 * c4);
      float4 g = *reinterpret_cast<const float4*>(&gu);
      g.x = y.x > 0.f ? g.x : 0.f;
      g.y = y.y > 0.f ? g.y : 0.f;
      g.z = y.z > 0.f ? g.z : 0.f;
      g.w = y.w > 0.f ? g.w : 0.f;
      *reinterpret_cast<float4*>(grad_pre + r * (int64_t)N + 4 * c4) = g;
      acc.x += g.x;
      acc.y += g.y;
      acc.z += g.z;
      acc.w += g.w;
    }
  }
  s_acc[threadIdx.x] = acc;
  __syncthreads();
  if (rsub == 0 && col_ok) {  // fixed order over the row sub-groups
    for (int k = 1; k < rstep; ++k) {
      const float4 o = s_acc[k * C4 + c4];
      acc.x += o.x;
      acc.y += o.y;
      acc.z += o.z;
      acc.w += o.w;
    }
    *reinterpret_cast<float4*>(partials + (int64_t)blockIdx.x * N + 4 * c4) = acc;
  }
}

// 32 columns per CTA, 8 slices of the partial rows per column; slices are added in a fixed order.
__global__ void __launch_bounds__(256)
bgrad_finish_kernel(const float* __restrict__ partials, int32_t parts, int32_t N,
                    float* __restrict__ grad_bias) {
  __shared__ float s[8][33];
  const int cx = threadIdx.x & 31, sy = threadIdx.x >> 5;
  const int n = blockIdx.x * 32 + cx;
  float a = 0.f;
  if (n < N) {
    const int per = (parts + 7) >> 3;
    const int p0 = sy * per, p1 = min(parts, p0 + per);
#pragma unroll 8
    for (int p = p0; p < p1; ++p) a += partials[(int64_t)p * N + n];
  }
  s[sy][cx] = a;
  __syncthreads();
  if (sy == 0 && n < N) {
#pragma unroll
    for (int k = 1; k < 8; ++k) a += s[k][cx];
    grad_bias[n] = a;
  }
}

}  // namespace b200

extern "C" size_t b200_relu_bias_grad_scratch_bytes(int32_t N) {
  return (size_t)b200::kMlpMaxParts * (size_t)(N > 0 ? N : 0) * sizeof(float);
}

extern "C" int b200_relu_bias_grad(const float* grad_out, int64_t grad_out_stride, const float* out,
                                   int64_t out_stride, int64_t B, int32_t N, float* grad_pre,
                                   float* grad_bias, float* scratch, void* stream) {
  using namespace b200;
  if (B < 0 || N <= 0 || (N & 3) || N > 1024 || (grad_out_stride & 3) || (out_stride & 3)) return B200_EINVAL;
  cudaStream_t s = (cudaStream_t)stream;
  if (B == 0) {
    B200_CUDA_TRY(cudaMemsetAsync(grad_bias, 0, sizeof(float) * N, s));
    return B200_OK;
  }
  int C4 = 1;
  while (C4 < N / 4) C4 <<= 1;  // power of two <= 256
  const int rstep = kMlpThreads / C4;
  int64_t parts = ceil_div(B, (int64_t)rstep * 16);  // >= 16 rows per thread
  if (parts > kMlpMaxParts) parts = kMlpMaxParts;
  if (parts < 1) parts = 1;
  const int64_t rows_per_cta = ceil_div(B, parts);
  parts = ceil_div(B, rows_per_cta);
  relu_bgrad_kernel<<<(unsigned)parts, kMlpThreads, 0, s>>>(grad_out, grad_out_stride, out, out_stride, B, N,
                                                             C4, rows_per_cta, grad_pre, scratch);
  B200_LAUNCH_CHECK();
  bgrad_finish_kernel<<<(unsigned)ceil_div(N, 32), 256, 0, s>>>(scratch, (int32_t)parts, N, grad_bias);
  B200_LAUNCH_CHECK();
  return B200_OK;
}
