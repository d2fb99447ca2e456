// TEST INFRASTRUCTURE: extern "C" doorway to the reference's own scalar implementations in
// third_party/fbgemm/src/RefImplementations.cc (EmbeddingSpMDM_ref :1092,
// rowwise_sparse_adagrad_ref :1569), which the Makefile compiles unmodified from the reference
// checkout.  Nothing here restates reference code; it only forwards arguments.
#include <cstdint>

#include "fbgemm/FbgemmEmbedding.h"
#include "src/RefImplementations.h"

extern "C" {

// fp32 rows, int64 indices and offsets, fp32 out — the instantiation the reference's CPU TBE uses.
int ref_spmdm_f32(int64_t block_size, int64_t output_size, int64_t index_size, int64_t data_size,
                  const float* input, const int64_t* indices, const int64_t* offsets,
                  const float* weights, int normalize_by_lengths, float* out, int64_t output_stride) {
  return fbgemm::EmbeddingSpMDM_ref<float, int64_t, int64_t, float>(
             block_size, output_size, index_size, data_size, input, indices, offsets, weights,
             normalize_by_lengths != 0, out, /*is_weight_positional=*/false, /*use_offsets=*/true,
             output_stride, /*input_stride=*/-1, /*scale_bias_last=*/true)
             ? 1
             : 0;
}

// int8 rows with fp16 scale/bias FIRST (scale_bias_last=false), the nbit TBE layout.
int ref_spmdm_u8_sbfront(int64_t block_size, int64_t output_size, int64_t index_size,
                         int64_t data_size, const uint8_t* input, const int64_t* indices,
                         const int64_t* offsets, const float* weights, int normalize_by_lengths,
                         float* out, int64_t output_stride, int64_t input_stride) {
  return fbgemm::EmbeddingSpMDM_ref<uint8_t, int64_t, int64_t, float>(
             block_size, output_size, index_size, data_size, input, indices, offsets, weights,
             normalize_by_lengths != 0, out, false, true, output_stride, input_stride,
             /*scale_bias_last=*/false)
             ? 1
             : 0;
}

// Rowwise Adagrad on already-coalesced gradient rows; the reference CPU TBE calls it with
// lr = -learning_rate (embedding_backward_split_cpu_template.cpp:171).
int ref_rowwise_adagrad(int num_rows, int block_size, uint64_t param_size, float* w, const float* g,
                        float* h, const int64_t* indices, float epsilon, float lr) {
  return fbgemm::rowwise_sparse_adagrad_ref<int64_t>(num_rows, block_size, param_size, w, g, h,
                                                     indices, epsilon, lr, 0.f, nullptr, 0);
}
}
