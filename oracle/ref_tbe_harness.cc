// TEST INFRASTRUCTURE / CPU BASELINE: drives the reference's own CPU embedding kernels - FBGEMM's JIT-generated
// EmbeddingSpMDM and rowwise sparse Adagrad (third_party/fbgemm/src/EmbeddingSpMDM.cc, SparseAdagrad.cc,
// compiled UNMODIFIED where they lie, together with the vendored asmjit) - with exactly the arguments the
// reference's CPU table-batched embedding passes them:
//   forward   fbgemm_gpu/codegen/embedding_forward_split_cpu.cpp:71-116 (GenerateEmbeddingSpMDMWithStrides
//             <float, int64, int64>(D, weighted, mean, prefetch 16, false, use_offsets, output_stride),
//             one call per (table, batch slice) under at::parallel_for over the batch)
//   backward  fbgemm_gpu/codegen/embedding_backward_split_cpu_template.cpp:125-182 (per table: a CSR->CSC
//             transpose, then per 64-column block GenerateEmbeddingSpMDMWithStrides<float, int32, int32>
//             (D, weighted, false, 16, false, true, -1, grad_stride) to sum the gradient rows of each unique id
//             and GenerateSparseAdaGrad<int>(D, rowwise = true)(..., eps, -lr, 0, nullptr, 0)).
// The torch-op layer around those calls (embedding_forward_split_cpu.cpp itself, which also holds
// batched_csr2csc) cannot be compiled here: it includes folly and libdivide, which are not in this image, and
// the backward is jinja-generated.  So the loops and the transpose below are this harness's; every floating
// point operation is executed by the reference's kernels.
#include <omp.h>

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <vector>

#include "fbgemm/FbgemmEmbedding.h"

extern "C" {

int ref_tbe_num_threads(void) { return omp_get_max_threads(); }

// out[b, Do[t]:Do[t+1]] = sum of rows; returns 1 on success, 0 when a kernel reported a bad index.
int ref_tbe_forward_f32(const float* weights, const int64_t* weights_offsets, const int32_t* D_offsets,
                        const int64_t* rows_per_table, int32_t T, int32_t B, const int64_t* indices,
                        const int64_t* offsets, const float* indice_weights, int mean, float* out,
                        int64_t out_stride) {
  int ok = 1;
  const int nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(static) reduction(& : ok)
  for (int th = 0; th < nthreads; ++th) {  // at::parallel_for(0, B, 0, ...): one contiguous batch slice per thread
    const int64_t per = (B + nthreads - 1) / nthreads;
    const int64_t b_begin = std::min<int64_t>(B, th * per), b_end = std::min<int64_t>(B, b_begin + per);
    if (b_begin >= b_end) continue;
    for (int t = 0; t < T; ++t) {
      const int D = D_offsets[t + 1] - D_offsets[t];
      auto kernel = fbgemm::GenerateEmbeddingSpMDMWithStrides<float, int64_t, int64_t>(
          D, indice_weights != nullptr, mean != 0, /*prefetch=*/16, /*is_weight_positional=*/false,
          /*use_offsets=*/true, out_stride);
      const int64_t* offsets_begin = offsets + (int64_t)t * B + b_begin;
      const int64_t n_idx = offsets[(int64_t)t * B + b_end] - *offsets_begin;
      const bool success = kernel(b_end - b_begin, n_idx, rows_per_table[t], weights + weights_offsets[t],
                                  indices + *offsets_begin, offsets_begin,
                                  indice_weights ? indice_weights + *offsets_begin : nullptr,
                                  out + b_begin * out_stride + D_offsets[t]);
      ok &= success ? 1 : 0;
    }
  }
  return ok;
}

// Fused rowwise-Adagrad step, in place on weights / momentum1.  Returns 1 on success.
int ref_tbe_backward_rowwise_adagrad_f32(const float* grad, int64_t grad_stride, float* weights,
                                         const int64_t* weights_offsets, const int32_t* D_offsets,
                                         const int64_t* rows_per_table, int32_t T, int32_t B,
                                         const int64_t* indices, const int64_t* offsets, float* momentum1,
                                         const int64_t* momentum1_offsets, float lr, float eps) {
  int ok = 1;
  struct Csc {
    std::vector<int> row_indices, seg_ptr, seg_idx;
  };
  std::vector<Csc> cscs((size_t)T);
  // CSR -> CSC per table, tables in parallel (the reference's batched_csr2csc is a parallel radix sort):
  // (column id, row b) pairs sorted by id then b; one column segment per unique id
#pragma omp parallel for schedule(dynamic, 1)
  for (int t = 0; t < T; ++t) {
    const int64_t p0 = offsets[(int64_t)t * B], p1 = offsets[(int64_t)(t + 1) * B];
    const int64_t n = p1 - p0;
    Csc& c = cscs[(size_t)t];
    if (n == 0) continue;
    std::vector<uint64_t> keys((size_t)n);
    for (int32_t b = 0; b < B; ++b)
      for (int64_t p = offsets[(int64_t)t * B + b]; p < offsets[(int64_t)t * B + b + 1]; ++p)
        keys[(size_t)(p - p0)] = ((uint64_t)indices[p] << 32) | (uint32_t)b;
    std::sort(keys.begin(), keys.end());
    c.row_indices.resize((size_t)n);
    c.seg_ptr.reserve((size_t)n + 1);
    c.seg_idx.reserve((size_t)n);
    for (int64_t i = 0; i < n; ++i) {
      c.row_indices[(size_t)i] = (int)(keys[(size_t)i] & 0xffffffffu);
      if (i == 0 || (keys[(size_t)i] >> 32) != (keys[(size_t)i - 1] >> 32)) {
        c.seg_ptr.push_back((int)i);
        c.seg_idx.push_back((int)(keys[(size_t)i] >> 32));
      }
    }
    c.seg_ptr.push_back((int)n);
  }
  for (int t = 0; t < T; ++t) {
    const Csc& csc = cscs[(size_t)t];
    const int64_t c_end = (int64_t)csc.seg_idx.size();
    if (c_end == 0) continue;
    const int D = D_offsets[t + 1] - D_offsets[t];
    auto spmdm = fbgemm::GenerateEmbeddingSpMDMWithStrides<float, int32_t, int32_t>(
        D, /*has_weight=*/false, /*normalize_by_lengths=*/false, /*prefetch=*/16,
        /*is_weight_positional=*/false, /*use_offsets=*/true, /*output_stride=*/-1, /*input_stride=*/grad_stride);
    auto adagrad = fbgemm::GenerateSparseAdaGrad<int>(D, /*rowwise=*/true);
    constexpr int C_BLOCK = 64;
    const int64_t nblocks = (c_end + C_BLOCK - 1) / C_BLOCK;
#pragma omp parallel reduction(& : ok)
    {
      std::vector<float> buf((size_t)C_BLOCK * D);
#pragma omp for schedule(static)
      for (int64_t blk = 0; blk < nblocks; ++blk) {
        const int64_t c = blk * C_BLOCK, ce = std::min(c + C_BLOCK, c_end);
        const int* off_begin = csc.seg_ptr.data() + c;
        const bool success = spmdm(ce - c, csc.seg_ptr[(size_t)ce] - *off_begin, B, grad + D_offsets[t],
                                   csc.row_indices.data() + *off_begin, off_begin, nullptr, buf.data());
        const int done = adagrad((int)(ce - c), (uint64_t)rows_per_table[t] * D, weights + weights_offsets[t],
                                 buf.data(), momentum1 + momentum1_offsets[t], csc.seg_idx.data() + c, eps, -lr,
                                 /*weight_decay=*/0.f, /*counter=*/nullptr, /*counter_halflife=*/0);
        ok &= (success && done == ce - c) ? 1 : 0;
      }
    }
  }
  return ok;
}

}  // extern "C"
